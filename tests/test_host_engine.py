"""Multi-process tests of the host engine over loopback TCP: guide programs (BASELINE config 1)
and the reference's fault-injection matrix (test/test.mk:15-37) with our tracker + launcher."""
import re

import pytest

from conftest import bin_path, run_job


def _lines(res):
    out = []
    for i in sorted(res.outputs):
        out.extend(res.outputs[i].splitlines())
    return out


def test_basic_world4_oracle():
    res = run_job([bin_path("basic")], 4)
    assert res.ok, res.outputs
    lines = _lines(res)
    assert sum("after-allreduce-max: a={3, 4, 5}" in l for l in lines) == 4
    assert sum("after-allreduce-sum: a={12, 16, 20}" in l for l in lines) == 4


def test_broadcast_world3():
    res = run_job([bin_path("broadcast")], 3)
    assert res.ok
    assert sum('after-broadcast: s="hello world"' in l for l in _lines(res)) == 3


def test_lazy_allreduce_prepare_skipped_on_recovery():
    """doc/guide.md:235-241 of the reference: rank 0 dies at the 2nd op; after the restart its first
    allreduce is recovered from the peer's cache, so the prepare lambda does NOT run again."""
    res = run_job([bin_path("lazy_allreduce"), "mock=0,0,1,0"], 2)
    assert res.ok, res.outputs
    # rank 0 is whichever task reached the tracker first
    victim = [i for i, o in res.outputs.items() if "Hit Mock Error" in o]
    assert len(victim) == 1 and res.restarts == 1 and res.attempts[victim[0]] == 1
    second_life = res.outputs[victim[0]].split("Hit Mock Error")[-1]
    assert "run prepare function" not in second_life
    assert "after-allreduce-max: a={1, 2, 3}" in second_life
    assert "after-allreduce-sum: a={2, 4, 6}" in second_life


@pytest.mark.parametrize("world", [1, 2, 3, 5, 8])
def test_model_recover_no_fault(world):
    res = run_job([bin_path("model_recover"), "1000", "rabit_bootstrap_cache=1"], world)
    assert res.ok, res.outputs
    assert res.restarts == 0


@pytest.mark.parametrize("extra", [["rabit_shm_ring=0"], ["rabit_shm_ring=64K", "rabit_shm_spin_us=0"],
                                   ["rabit_shm_test_split=2"], ["rabit_shm_test_split=3", "rabit_shm_ring=64K"]])
def test_transport_variants_no_fault(extra):
    """plain sockets, and tiny shared-memory rings (every message wraps them many times) without the look before sleeping"""
    res = run_job([bin_path("model_recover"), "300000"] + extra, 5)
    assert res.ok, res.outputs
    res = run_job([bin_path("model_recover"), "300000", "rabit_reduce_ring_mincount=1"] + extra, 4)
    assert res.ok, res.outputs


def test_ring_paths_forced():
    for mincount in ("1", "10"):
        res = run_job([bin_path("model_recover"), "10000", "rabit_reduce_ring_mincount=" + mincount], 10)
        assert res.ok, res.outputs


def test_tree_path_large_message_and_small_buffer():
    res = run_job([bin_path("model_recover"), "50000", "rabit_reduce_ring_mincount=100000000",
                   "rabit_reduce_buffer=4K"], 6)
    assert res.ok, res.outputs
    res = run_job([bin_path("model_recover"), "50000", "rabit_reduce_ring_mincount=1", "rabit_reduce_buffer=4K"], 6)
    assert res.ok, res.outputs


# expected deaths: a key fires only while the rank's restart count equals its 4th field, so e.g.
# mock=0,1,1,0 never fires once rank 0 has already died at mock=0,0,1,0
SCENARIOS = {
    "model_recover": ("model_recover", ["mock=0,0,1,0", "mock=1,1,1,0", "rabit_bootstrap_cache=true",
                                        "rabit_reduce_ring_mincount=1", "rabit_timeout=true",
                                        "rabit_timeout_sec=30"], 2),
    "model_die_same": ("model_recover", ["mock=0,0,1,0", "mock=1,1,1,0", "mock=0,1,1,0", "mock=4,1,1,0",
                                         "mock=9,1,1,0", "rabit_bootstrap_cache=1"], 4),
    "model_die_hard": ("model_recover", ["mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "mock=0,1,1,0",
                                         "mock=4,1,1,0", "mock=9,1,1,0", "mock=8,1,2,0", "mock=4,1,3,0",
                                         "rabit_bootstrap_cache=1"], 6),
    "local_recover": ("local_recover", ["mock=0,0,1,0", "mock=1,1,1,0", "mock=0,1,1,0", "mock=4,1,1,0",
                                        "mock=9,1,1,0", "mock=1,1,1,1"], 5),
    "lazy_die_hard": ("lazy_recover", ["mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "mock=0,1,1,0",
                                       "mock=4,1,1,0", "mock=9,1,1,0", "mock=8,1,2,0", "mock=4,1,3,0"], 6),
    "lazy_die_same": ("lazy_recover", ["mock=0,0,1,0", "mock=1,1,1,0", "mock=0,1,1,0", "mock=4,1,1,0",
                                       "mock=9,1,1,0"], 4),
}


@pytest.mark.parametrize("transport", ["shm", "tcp", "mixed"])
@pytest.mark.parametrize("name", sorted(SCENARIOS))
def test_fault_matrix_world10(name, transport):
    """the reference's test/test.mk matrix at world 10; once with the payload in shared-memory rings (what workers of one
    host do by default) and once on the sockets alone (what workers of different hosts do)"""
    prog, args, deaths = SCENARIOS[name]
    # "mixed": ranks of different parity pretend to sit on different hosts, so that one wait covers rings AND sockets
    # (several hosts with several workers each)
    more = {"shm": [], "tcp": ["rabit_shm_ring=0"], "mixed": ["rabit_shm_test_split=2"]}[transport]
    res = run_job([bin_path(prog), "10000"] + args + more, 10, timeout=180)
    assert res.ok, {i: o[-600:] for i, o in res.outputs.items()}
    assert res.restarts == deaths
    hits = sum(len(re.findall("terminate called", o)) for o in res.outputs.values())
    assert hits == deaths  # each death is an uncaught rabit::Error -> abort, like the reference
    # every rank finished all three versions
    for i, o in res.outputs.items():
        assert "iter 2" in o, (i, o[-400:])


def test_checkpoint_kill_points():
    """die exactly at CheckPoint / LazyCheckPoint entry (mock seq = ops issued in the version)"""
    # model_recover at world 4 issues 1 max + 4 broadcasts(2 ops each) + 1 sum + 1 allgather = 11 ops
    res = run_job([bin_path("model_recover"), "2000", "mock=2,0,11,0", "mock=0,1,11,0", "rabit_bootstrap_cache=1"], 4)
    assert res.ok and res.restarts == 2, res.outputs
    res = run_job([bin_path("local_recover"), "2000", "mock=1,0,10,0", "mock=3,2,10,0"], 4)
    assert res.ok and res.restarts == 2, res.outputs


def test_local_replica_survives_adjacent_deaths():
    """two ring-adjacent ranks die at the same op: with the default 2 replicas the local model of
    each is still held by a third rank"""
    res = run_job([bin_path("local_recover"), "3000", "mock=2,1,1,0", "mock=3,1,1,0"], 6)
    assert res.ok and res.restarts == 2, res.outputs


def test_watchdog_kills_stuck_survivors():
    """a rank dies and is never restarted: with rabit_timeout the survivors must exit (rc 10) instead
    of waiting forever (reference only prints)"""
    res = run_job([bin_path("model_recover"), "100", "mock=1,0,1,0", "rabit_timeout=1", "rabit_timeout_sec=2",
                   "rabit_bootstrap_cache=1"], 3, max_attempts=0, timeout=60)
    assert not res.ok
    # which task holds rank 1 depends on arrival order at the tracker: one abort, two watchdog exits
    assert sorted(res.returncodes.values()) == [-6, 10, 10]
    assert any("exit due to time out 2 s" in o for o in res.outputs.values())


def test_recovery_at_world_31(built):
    """the reference's sweep goes up to 31 machines (test/speed_runner.py:13-18): ring + tree + recovery at that size"""
    res = run_job([bin_path("model_recover"), "10000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=9,1,1,0", "mock=15,2,3,0",
                   "rabit_bootstrap_cache=1", "rabit_timeout=1", "rabit_timeout_sec=100"], 31, timeout=240)
    assert res.ok, {i: o[-400:] for i, o in res.outputs.items() if "iter 2" not in o}
    assert res.restarts == 4
    assert all("iter 2 checkpointed (version 3)" in o for o in res.outputs.values())


@pytest.mark.parametrize("extra", [[], ["rabit_reduce_ring_mincount=10"], ["rabit_reduce_ring_mincount=100000000"],
                                   ["rabit_reduce_buffer=8", "rabit_reduce_ring_mincount=100000000"],
                                   ["rabit_reduce_ring_mincount=10", "rabit_shm_ring=64K"],
                                   ["rabit_reduce_ring_mincount=100000000", "rabit_shm_ring=64K"],
                                   ["rabit_reduce_ring_mincount=10", "rabit_shm_ring=0"]])
def test_custom_reducers_across_ranks(extra):
    """Reducer<16-byte struct> (opaque merge, declared packed sum, mixed lanes), Reducer<12-byte struct> and
    SerializeReducer over tree and ring, world 5; the 4th case has a per-link buffer smaller than one element; then
    shared-memory rings small enough to wrap many times (12-byte items straddle the wrap), and plain TCP"""
    res = run_job([bin_path("reducer_multi"), "30000" if any("shm_ring" in e for e in extra) else "3000"] + extra, 5)
    assert res.ok, res.outputs
    assert res.restarts == 0


def test_custom_reducers_replayed_after_kill():
    res = run_job([bin_path("reducer_multi"), "2000", "iters=3", "mock=1,1,1,0", "mock=3,1,3,0"], 4)
    assert res.ok, res.outputs
    assert res.restarts == 2


def test_tree_reduce_minsize_small_and_large():
    for v in ("1", "4096", "100000000"):
        res = run_job([bin_path("model_recover"), "30000", "rabit_reduce_ring_mincount=100000000",
                       "rabit_tree_reduce_minsize=" + v], 6)
        assert res.ok, res.outputs


@pytest.mark.parametrize("ring", ["512K", "0"])
def test_waiting_for_a_peer_that_has_left_fails_instead_of_hanging(ring):
    """Rank 1 finishes and says bye; rank 0 then starts one more collective that needs rank 1. With the payload on the
    socket the reader hits the end of the stream; a shared-memory ring just stays empty, so the wait must notice that
    the peer's connection is closed (found by the kill campaigns: a re-run shutdown barrier waiting for ranks that
    had already passed it)."""
    import time
    from conftest import run_py
    code = (
        "import sys, time, numpy as np\n"
        "from rabit_b200 import rabit\n"
        "rabit.init(sys.argv[1:], lib='base')\n"
        "r = rabit.get_rank()\n"
        "rabit.allreduce(np.ones(4, dtype=np.float32), rabit.SUM)\n"
        "if r == 0:\n"
        "    time.sleep(1.0)\n"
        "    print('extra op starts', flush=True)\n"
        "    rabit.allreduce(np.ones(4, dtype=np.float32), rabit.SUM)\n"
        "    print('extra op returned', flush=True)\n"
        "rabit.finalize()\n")
    t0 = time.time()
    res = run_py(code, 2, args=["rabit_shm_ring=" + ring], timeout=60, max_attempts=1)
    assert time.time() - t0 < 30, res.outputs
    out = "".join(res.outputs.values())     # (rabit rank 0 is whichever task the tracker labelled 0)
    assert "extra op starts" in out and "extra op returned" not in out, res.outputs
    assert "peer-closed" in out, res.outputs


def test_rejoins_do_not_leak_file_descriptors():
    """every re-join opens sockets and shared-memory rings; after shutdown only stdin / stdout / stderr and the doorbell
    socket may be left, however many re-joins there were"""
    res = run_job([bin_path("model_recover"), "1000", "rabit_debug=1", "rabit_bootstrap_cache=1", "mock=0,0,1,0", "mock=1,1,1,0",
                   "mock=2,1,2,0", "mock=3,2,1,0", "mock=0,2,1,1"], 4)
    assert res.ok, res.outputs
    assert res.restarts == 5
    counts = re.findall(r"open file descriptors after shutdown: (\d+)", "".join(res.outputs.values()))
    assert len(counts) == 4 and all(int(c) <= 4 for c in counts), counts
