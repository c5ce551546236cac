"""Multi-GPU tests: one process per GPU, real peer memory (needs >= 2 GPUs; skipped otherwise).

Covers BASELINE config 4 (checkpoint + result-cache replay after mock kills, on device buffers).
"""
import os
import sys

import pytest

from conftest import ROOT, run_job

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run(args, n, timeout=300, extra_env=None, **kw):
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), "RABIT_DEVICE_VERBOSE": "1"}
    env.update(extra_env or {})
    return run_job([sys.executable, os.path.join(ROOT, "tests", "gpu_worker.py")] + args, n, timeout=timeout,
                   env=env, per_rank_gpu=True, **kw)


@pytest.fixture(scope="module")
def world():
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    return min(n, 8)


@pytest.mark.parametrize("algo", ["auto", "oneshot", "twoshot", "tma"])
def test_device_collectives_no_fault(world, algo):
    res = _run(["200000", "rabit_device_algo=" + algo], world)
    assert res.ok, {i: o[-1500:] for i, o in res.outputs.items()}
    assert res.restarts == 0
    assert any("kernel launches" in m for m in res.messages)


def test_device_recovery_after_kills(world):
    """a rank dies in version 0 and another in version 1; both come back, load the checkpoint and
    replay the cached device results"""
    res = _run(["50000", "mock=0,0,1,0", "mock=1,1,2,0", "rabit_timeout=1", "rabit_timeout_sec=120"], world,
               timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 2
    for i, o in res.outputs.items():
        assert "iter 2 checkpointed" in o


def test_device_recovery_large_payload_over_nvlink(world):
    """12 MB results and checkpoint: recovery payloads of that size travel as ONE broadcast kernel over
    NVLink (RobustEngine::DeviceDeliver) instead of the TCP ring"""
    res = _run(["3000000", "iters=2", "mock=1,1,1,0", "rabit_debug=1", "rabit_timeout=1", "rabit_timeout_sec=120"],
               world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 1
    assert any("moved over NVLink" in o for o in res.outputs.values())


def test_native_cpp_device_workload_with_kills(world):
    """the C++ public API (rabit::Allreduce / Broadcast on device pointers) incl. recovery"""
    from conftest import bin_path
    res = run_job([bin_path("device_recover_cuda"), "100000", "mock=0,0,1,0", "mock=1,1,2,0"], world, timeout=300,
                  per_rank_gpu=True)
    assert res.ok, {i: o[-2000:] for i, o in res.outputs.items()}
    assert res.restarts == 2


def test_device_symmetric_zero_copy_buffers(world):
    """allreduce / broadcast on RabitDeviceSymAlloc buffers through the fault-tolerant API, also across a kill"""
    res = _run(["100000", "rabit_device_heap=64M", "mock=1,1,3,0"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 1


def test_device_symmetric_async_in_place(world):
    """heap buffers through the async API: reduced in place by the zero-copy kernels"""
    res = _run(["100000", "rabit_device_heap=64M", "iters=2"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 0


def test_host_buffers_over_nvlink(world):
    """large host buffers (numpy, pinned torch) are chunked through H2D | fused kernel | D2H, also across a kill"""
    res = _run(["30000", "hostpipe=1", "rabit_device_host_min=1M", "rabit_device_host_chunk=1M", "mock=1,1,8,0",
                "rabit_timeout=1", "rabit_timeout_sec=120"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 1
    assert any("kernel launches" in m for m in res.messages)


def test_local_models_replicated_over_nvlink(world):
    """per-rank (local) models: the replicas travel neighbour to neighbour by peer copies on the side stream, also
    when a restarted rank fetches its own model back; with rabit_debug the engine reports the bytes per transport"""
    from conftest import bin_path
    kills = ["mock=0,1,1,0"] + (["mock=%d,2,1,0" % (world - 1)] if world > 2 else [])
    res = run_job([bin_path("local_recover"), "60000", "rabit_debug=1", "rabit_device_blob_min=1"] + kills, world,
                  timeout=300, per_rank_gpu=True)
    assert res.ok, {i: o[-2000:] for i, o in res.outputs.items()}
    assert res.restarts == len(kills)
    lines = [l for o in res.outputs.values() for l in o.splitlines() if "local state complete" in l]
    assert lines, "no local-state report in the debug log"
    assert all(l.rstrip().endswith("over TCP 0") for l in lines), lines[:4]
    assert any("over NVLink 0," not in l for l in lines)
    # the global checkpoint of the restarted rank came over NVLink as well
    assert any("moved over NVLink" in o for o in res.outputs.values())


def test_lazy_checkpoint_large_model_fetched_over_nvlink(world):
    """BASELINE config 5 in small: LazyCheckPoint of a 96 MB model; a rank dies in the next version, the holder
    serialises the model on demand and the restarted rank pulls it over NVLink in several window pieces"""
    from conftest import bin_path
    res = run_job([bin_path("lazy_recover"), "24000000", "iters=2", "rabit_debug=1", "mock=1,1,0,0"], world, timeout=600,
                  per_rank_gpu=True)
    assert res.ok, {i: o[-2000:] for i, o in res.outputs.items()}
    assert res.restarts == 1
    moved = [l for o in res.outputs.values() for l in o.splitlines() if "moved over NVLink" in l]
    assert any("payload of 9600" in l for l in moved), moved[:4]


STREAM_SMALL = ["rabit_device_stream_min=1M", "rabit_device_stream_chunk=1M", "rabit_device_stream_first=256K",
                "rabit_device_host_min=1M", "rabit_device_host_chunk=1M"]


def test_streamed_allreduce_no_fault(world):
    """plain, symmetric and host tensors through the chunked allreduce (several chunk kernels per op)"""
    res = _run(["40000", "iters=2", "stream=1", "rabit_device_heap=64M"] + STREAM_SMALL, world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 0


def test_streamed_allreduce_mock_kill_at_entry(world):
    """a rank dies at the entry of a streamed op of version 1: the survivors' chunk kernels give up, the op is undone
    from the stage copy and run again once the rank is back"""
    res = _run(["40000", "iters=3", "stream=1", "rabit_device_heap=64M", "mock=1,1,6,0", "mock=0,1,9,0", "rabit_timeout=1",
                "rabit_timeout_sec=120"] + STREAM_SMALL, world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 2


@pytest.mark.parametrize("spec", ["1,2,2", "0,3,3", "1,1,5", "0,2,6", "1,2,0"])
def test_streamed_allreduce_death_inside_the_op(world, spec):
    """a rank dies INSIDE a streamed op (spec = rank, chunk, index of the streamed op in the worker's first iteration:
    0-1 symmetric, 2 plain sum, 3 max, 4 float64, 5 symmetric, 6 host tensor), after some chunks
    have already been written back into the survivors' buffers: their buffers must hold the input again when the op
    is retried, otherwise the self-checks of the worker fail"""
    env = {"RABIT_TEST_STREAM_KILL": spec}
    res = _run(["40000", "iters=2", "stream=1", "rabit_device_heap=64M", "rabit_timeout=1", "rabit_timeout_sec=120"] +
               STREAM_SMALL, world, timeout=600, extra_env=env)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 1
    assert any("RABIT_TEST_STREAM_KILL" in o for o in res.outputs.values())


def test_arena_grows_on_demand(world):
    """a message larger than the current stage makes every rank re-allocate and re-exchange its arena"""
    res = _run(["50000", "iters=2", "grow=1"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 0


def test_cuda_graph_replay_across_ranks(world):
    """collectives captured in a CUDA graph and replayed on every rank (device-side op epoch)"""
    res = _run(["20000", "iters=2", "graph=1"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}


def test_nccl_engine_as_comparator(world):
    """rabit_engine=nccl (python lib='mpi'): the library-backed, non-fault-tolerant engine, the counterpart of the
    reference's MPI engine; device tensors go to ncclAllReduce / ncclBroadcast, host arrays to the TCP collectives"""
    code = ("import os, sys\nimport numpy as np\nimport torch\n"
            "from rabit_b200 import rabit\nfrom rabit_b200 import torch_binding as rt\n"
            "torch.cuda.set_device(int(os.environ['RABIT_DEVICE_ORDINAL']) % torch.cuda.device_count())\n"
            "rabit.init(sys.argv, lib='mpi')\n"
            "me, n = rabit.get_rank(), rabit.get_world_size()\n"
            "x = torch.full((100003,), float(me + 1), device='cuda')\n"
            "rabit.allreduce(x, rabit.SUM)\n"
            "assert torch.equal(x, torch.full_like(x, n * (n + 1) / 2)), 'nccl allreduce'\n"
            "y = torch.arange(5000, device='cuda', dtype=torch.int32) * (me + 1)\n"
            "rabit.allreduce(y, rabit.MAX)\n"
            "assert torch.equal(y, torch.arange(5000, device='cuda', dtype=torch.int32) * n), 'nccl max'\n"
            "b = torch.full((70001,), me, dtype=torch.uint8, device='cuda')\n"
            "rt.broadcast_(b, n - 1)\n"
            "assert torch.equal(b, torch.full_like(b, n - 1)), 'nccl broadcast'\n"
            "h = rabit.allreduce(np.full(10, me, dtype=np.float64), rabit.SUM)\n"
            "assert h[0] == n * (n - 1) / 2\n"
            "assert not rt.device_active()\n"
            "rabit.tracker_print('[%d] nccl engine ok\\n' % me)\n"
            "rabit.finalize()\n")
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    res = run_job([sys.executable, "-c", code], world, timeout=300, env=env, per_rank_gpu=True)
    assert res.ok, {i: o[-2000:] for i, o in res.outputs.items()}
    assert sum("nccl engine ok" in m for m in res.messages) == world
