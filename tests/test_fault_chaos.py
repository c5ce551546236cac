"""Random SIGKILLs at random moments (tests/chaos_harness.py): the job must still finish everywhere.
Global-model workloads only: they tolerate any number of overlapping failures as long as one rank survives
(local models legitimately cannot, see doc/guide.md)."""
import pytest

from chaos_harness import run


@pytest.mark.parametrize("seed", [3, 11, 261, 393, 1007, 1042, 2043, 4219, 8559])
def test_random_sigkills(built, seed):
    ok, info = run(seed)
    assert ok, {k: info[k] for k in ("prog", "world", "args", "kills", "done", "tails")}


@pytest.mark.parametrize("seed", [2344, 2568, 2584])
def test_random_sigkills_device_side(built, seed):
    """The same with the engine's device side active on the host-memory stand-in of the backend (src/host/fake_device.cc):
    result cache in backend memory, host buffers through the backend, recovery payloads over the backend's transport.
    These seeds kill a rank during Init's arena bootstrap and used to end in a fatal error / a re-join livelock."""
    ok, info = run(seed, ("model_recover", "lazy_recover", "local_recover"),
                   extra_args=["rabit_device=fake", "rabit_device_host_min=1", "rabit_device_blob_min=1"])
    assert ok, {k: info[k] for k in ("prog", "world", "args", "kills", "done", "tails")}


@pytest.mark.parametrize("seed", [3, 1007, 4219])
def test_random_sigkills_sockets_only(built, seed):
    """the transport between hosts: no shared-memory rings, every byte on the sockets"""
    ok, info = run(seed, extra_args=["rabit_shm_ring=0"])
    assert ok, {k: info[k] for k in ("prog", "world", "args", "kills", "done", "tails")}


@pytest.mark.parametrize("seed", [11, 393, 2043, 8559])
def test_random_sigkills_mixed_media(built, seed):
    """several hosts with several workers each: some links are shared-memory rings, some are sockets (test hook:
    ranks of different parity pretend to be on different hosts)"""
    ok, info = run(seed, extra_args=["rabit_shm_test_split=2"])
    assert ok, {k: info[k] for k in ("prog", "world", "args", "kills", "done", "tails")}
