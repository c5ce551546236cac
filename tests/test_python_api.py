"""The reference-compatible python module on the host engine (CPU, multi-process)."""
import os
import sys

from conftest import ROOT, run_job, run_py


def _py(script, args, n, **kw):
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    return run_job([sys.executable, os.path.join(ROOT, script)] + list(args), n, env=env, **kw)


def test_guide_basic_py_world4():
    res = _py("guide/basic.py", [], 4)
    assert res.ok, res.outputs
    out = "".join(res.outputs.values())
    assert out.count("after-allreduce-max: a=[3. 4. 5.]") == 4
    assert out.count("after-allreduce-sum: a=[12. 16. 20.]") == 4


def test_guide_broadcast_py():
    res = _py("guide/broadcast.py", [], 3)
    assert res.ok, res.outputs
    assert "".join(res.outputs.values()).count("after-broadcast: s=\"{'hello world': 100, 2: 3}\"") == 3


def test_lazy_allreduce_py_prepare_skipped_after_restart():
    res = _py("guide/lazy_allreduce.py", ["mock=0,0,1,0"], 2)
    assert res.ok, res.outputs
    victim = [i for i, o in res.outputs.items() if "Hit Mock Error" in o]
    assert len(victim) == 1 and res.attempts[victim[0]] == 1
    second_life = res.outputs[victim[0]].split("Hit Mock Error")[-1]
    assert "run prepare function" not in second_life
    assert "after-allreduce-sum: a=[2. 4. 6.]" in second_life


def test_python_local_recover_with_real_kills():
    res = _py("test/local_recover.py", ["2000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=3,1,2,0", "mock=1,1,1,1"], 4,
              timeout=180)
    assert res.ok, {i: o[-800:] for i, o in res.outputs.items()}
    assert res.restarts == 4
    for o in res.outputs.values():
        assert "iter 2 checkpointed (version 3)" in o


def test_python_misc_api():
    code = r'''
import numpy as np
from rabit_b200 import rabit
rabit.init()
r, n = rabit.get_rank(), rabit.get_world_size()
assert rabit.is_distributed() and 0 <= rabit.get_ring_prev_rank() < n
assert len(rabit.get_processor_name()) > 0
for dt in rabit.DTYPE_ENUM__:
    x = (np.arange(10) + r).astype(dt)
    assert np.array_equal(rabit.allreduce(x, rabit.MAX), (np.arange(10) + n - 1).astype(dt))
    assert np.array_equal(rabit.allreduce(x, rabit.MIN), np.arange(10).astype(dt))
    if dt.kind in "iu":
        assert np.array_equal(rabit.allreduce(np.full(3, 1 << r, dtype=dt), rabit.BITOR), np.full(3, (1 << n) - 1, dtype=dt))
g = rabit.allgather(np.full(5, r, dtype=np.int32))
assert np.array_equal(g, np.repeat(np.arange(n, dtype=np.int32), 5))
big = np.frombuffer(bytes(range(256)) * 1024, dtype=np.uint8)      # > 64 KiB checkpoint blob (reference truncates the length)
v, m = rabit.load_checkpoint()
assert v == 0 and m is None
rabit.checkpoint(big)
assert rabit.version_number() == 1
rabit.tracker_print("rank %d fine\n" % r)
rabit.finalize()
'''
    res = run_py(code, 3)
    assert res.ok, res.outputs
    assert sum("fine" in m for m in res.messages) == 3


def test_lib_mpi_selects_the_library_backed_engine():
    """lib='mpi' (reference python/rabit.py:47-71 -> librabit_mpi) loads the library-backed engine of this build,
    librabit_nccl.so; a world of one needs neither NCCL nor a GPU"""
    code = ("from rabit_b200 import rabit\nimport numpy as np, os\n"
            "rabit.init(['prog'], lib='mpi')\n"
            "assert os.path.basename(rabit._LIB_PATH) == 'librabit_nccl.so', rabit._LIB_PATH\n"
            "a = rabit.allreduce(np.arange(4, dtype=np.float32), rabit.SUM)\n"
            "assert a.tolist() == [0, 1, 2, 3]\n"
            "assert rabit.get_world_size() == 1\n"
            "rabit.finalize()\nprint('mpi-lib-ok')\n")
    import subprocess
    import sys
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=120)
    assert r.returncode == 0 and b"mpi-lib-ok" in r.stdout, r.stdout.decode()


def test_native_torch_binding_on_cpu_tensors():
    """the pybind11 / libtorch extension (src/python/torch_ext.cc) carries the tensor calls; CPU tensors go through the
    host engine, also across a scripted kill (world 3), and the ctypes fallback gives the same answers"""
    code = ("import sys, os\nimport torch\nfrom rabit_b200 import rabit, torch_binding as rt\n"
            "want_native = os.environ.get('RABIT_TORCH_CTYPES') is None\n"
            "assert rt.native_binding() == want_native, rt.native_binding()\n"
            "rabit.init(sys.argv, lib='mock')\n"
            "me, n = rabit.get_rank(), rabit.get_world_size()\n"
            "v, m = rabit.load_checkpoint()\n"
            "for it in range(v, 3):\n"
            "    t = torch.arange(1000, dtype=torch.float64) * (me + 1) + it\n"
            "    rt.allreduce_(t)\n"
            "    assert torch.equal(t, torch.arange(1000, dtype=torch.float64) * (n * (n + 1) // 2) + it * n)\n"
            "    b = torch.full((77,), me, dtype=torch.int32)\n"
            "    rt.broadcast_(b, it % n)\n"
            "    assert torch.equal(b, torch.full((77,), it % n, dtype=torch.int32))\n"
            "    rabit.checkpoint(it)\n"
            "rabit.tracker_print('[%d] torch binding ok\\n' % me)\n"
            "rabit.finalize()\n")
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    for extra in ({}, {"RABIT_TORCH_CTYPES": "1"}):
        e = dict(env)
        e.update(extra)
        res = run_job([sys.executable, "-c", code, "mock=1,1,1,0"], 3, timeout=180, env=e)
        assert res.ok, res.outputs
        assert res.restarts == 1
        assert sum("torch binding ok" in m for m in res.messages) == 3
