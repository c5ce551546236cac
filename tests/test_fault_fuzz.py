"""Randomised kill schedules for the recovery protocol (host engine, CPU).

The consensus state machine is the subtle part of the library: instead of only
replaying the reference's hand-written scenarios, draw random (rank, version, seq, trial) kill points
— including several ranks dying at the same op, re-deaths of a restarted rank, and kills at the
checkpoint call itself — and require that the self-checking workloads still finish everywhere.
Seeds are fixed, so failures are reproducible: the failing schedule is printed in the assertion.
"""
import random

import pytest

from conftest import bin_path, run_job

# ops per version (mock seq numbers): see the programs in test/
#   model_recover: max(1) + broadcasts(2 each) + sum(1) + allgather(1) ; then CheckPoint at that seq
#   local_recover / lazy_recover: max(1) + broadcasts(2 each) + sum(1)


def _schedule(rng, world, max_seq, kills):
    keys = set()
    trial = {}
    for _ in range(kills):
        r = rng.randrange(world)
        v = rng.randrange(3)
        s = rng.randrange(max_seq + 1)          # max_seq itself = the checkpoint call
        t = trial.get(r, 0) if rng.random() < 0.8 else 0
        keys.add((r, v, s, t))
        trial[r] = trial.get(r, 0) + 1
    return ["mock=%d,%d,%d,%d" % k for k in sorted(keys)]


@pytest.mark.parametrize("seed", range(8))
def test_random_kill_schedules_model_recover(seed):
    rng = random.Random(1000 + seed)
    world = rng.choice([3, 4, 5, 7, 8])
    nbcast = len(range(0, world, max(world // 3, 1)))
    max_seq = 1 + 2 * nbcast + 1 + 1
    kills = _schedule(rng, world, max_seq, rng.randrange(1, 7))
    args = ["500", "rabit_bootstrap_cache=1", "rabit_timeout=1", "rabit_timeout_sec=60"] + kills
    if rng.random() < 0.5:
        args.append("rabit_reduce_ring_mincount=1")
    res = run_job([bin_path("model_recover")] + args, world, timeout=150)
    assert res.ok, (world, kills, {i: o[-400:] for i, o in res.outputs.items() if "iter 2" not in o})
    for o in res.outputs.values():
        assert "iter 2 checkpointed (version 3)" in o, (world, kills)


@pytest.mark.parametrize("seed", range(6))
def test_random_kill_schedules_local_and_lazy(seed):
    rng = random.Random(2000 + seed)
    world = rng.choice([4, 5, 6, 8])
    prog = rng.choice(["local_recover", "lazy_recover"])
    nbcast = len(range(0, world, max(world // 3, 1)))
    max_seq = 1 + 2 * nbcast + 1
    nk = rng.randrange(1, 5)
    kills = _schedule(rng, world, max_seq, nk)
    if prog == "local_recover":
        # the local model of a rank survives as long as not all of {rank, rank+1, rank+2} die at once:
        # keep simultaneous deaths (same version and seq) to at most two ring-adjacent ranks
        seen = {}
        kept = []
        for k in kills:
            r, v, s, t = [int(x) for x in k[5:].split(",")]
            group = seen.setdefault((v, s), [])
            if len(group) < 2:
                group.append(r)
                kept.append(k)
        kills = kept
    res = run_job([bin_path(prog), "500", "rabit_timeout=1", "rabit_timeout_sec=60"] + kills, world, timeout=150)
    assert res.ok, (prog, world, kills, {i: o[-400:] for i, o in res.outputs.items() if "iter 2" not in o})
