"""The engine <-> device-backend contract on the CPU tier: `rabit_device=fake` is a DeviceBackend made of host memory and
the TCP collectives (src/host/fake_device.cc). With `rabit_device_host_min=1` every typed allreduce of the test programs
takes the road of a device op: fused consensus in the backend, result cached in backend memory, recovery payloads and
local-model replicas moved by the backend's transport (`rabit_device_blob_min=1`). The backend checks on every op that
all ranks present the same op counter, so a lapse in the engine's re-alignment after an abandoned op fails loudly."""
import pytest

from conftest import bin_path, run_job

FAKE = ["rabit_device=fake", "rabit_device_host_min=1", "rabit_device_blob_min=1", "rabit_debug=1"]


def _out(res):
    return "\n".join(res.outputs.values())


@pytest.mark.parametrize("prog,args,world,restarts", [
    ("model_recover", ["20000"], 4, 0),
    ("model_recover", ["20000", "rabit_bootstrap_cache=1", "mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "mock=3,2,2,0"], 4, 4),
    ("local_recover", ["20000", "mock=0,1,1,0", "mock=3,2,1,0"], 5, 2),
    ("lazy_recover", ["20000", "mock=1,1,0,0", "mock=2,2,3,0"], 4, 2),
    ("reducer_multi", ["3000", "iters=3", "mock=1,1,1,0"], 4, 1),
])
def test_recovery_scenarios_through_the_backend(prog, args, world, restarts):
    res = run_job([bin_path(prog)] + args + FAKE, world, timeout=120)
    assert res.ok, res.outputs
    assert res.restarts == restarts
    if restarts:
        # cached results / checkpoints reached the restarted rank through the backend's transport
        assert "moved over NVLink" in _out(res)


def test_local_replicas_use_the_backend_transport_only():
    res = run_job([bin_path("local_recover"), "20000", "mock=0,1,1,0"] + FAKE, 5, timeout=120)
    assert res.ok, res.outputs
    lines = [l for l in _out(res).splitlines() if "local state complete" in l]
    assert lines and all(l.rstrip().endswith("over TCP 0") for l in lines), lines[:3]


def test_op_larger_than_the_arena_is_cut_into_recoverable_pieces():
    """64 KiB arena = 32 KiB per op: a 200 KB allreduce becomes 7 sequence-numbered ops; a rank that dies behind them
    replays every piece from its peers' caches"""
    res = run_job([bin_path("model_recover"), "50000", "rabit_device_arena=64K", "rabit_bootstrap_cache=1", "mock=2,1,7,0",
                   "mock=0,2,8,0"] + FAKE, 4, timeout=120)
    assert res.ok, res.outputs
    assert res.restarts >= 1
    assert "/6 " in _out(res) or "/6)" in _out(res)      # the op keys of the pieces carry their index


@pytest.mark.parametrize("spec", ["2,3", "0,0", "1,7"])
def test_consensus_mismatch_falls_back_and_realigns_op_counters(spec):
    """one rank presents a different consensus code once (as when ranks picked different kernels) and comes out with
    its op counter ahead (as after an abandoned multi-kernel op): the engine goes through the host protocol, retries,
    and the counters are equal again at the next op (checked by the backend on every op)"""
    res = run_job([bin_path("model_recover"), "20000", "rabit_bootstrap_cache=1", "mock=1,1,1,0"] + FAKE, 4, timeout=120,
                  env={"RABIT_FAKE_MISMATCH": spec})
    assert res.ok, res.outputs
    assert "op counters differ" not in _out(res)
