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


def _run(args, n, timeout=300, **kw):
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), "RABIT_DEVICE_VERBOSE": "1"}
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


def test_arena_grows_on_demand(world):
    """a message larger than the current stage makes every rank re-allocate and re-exchange its arena"""
    res = _run(["50000", "iters=2", "grow=1"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 0


def test_cuda_graph_replay_across_ranks(world):
    """collectives captured in a CUDA graph and replayed on every rank (device-side op epoch)"""
    res = _run(["20000", "iters=2", "graph=1"], world, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
