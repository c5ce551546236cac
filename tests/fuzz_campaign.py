"""Long form of tests/test_fault_fuzz.py: many random `mock=` kill schedules over all three recovery workloads.

    python tests/fuzz_campaign.py <first_seed> <count>

A run passes when every worker ends with exit code 0 and either finished the last iteration or, having been killed at
the last CheckPoint call, came back to find the job complete ("start at iter=<iters>").
"""
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200.parallel.launcher import launch  # noqa: E402


def schedule(rng, world, max_seq, kills):
    keys, trial = set(), {}
    for _ in range(kills):
        r, v, s = rng.randrange(world), rng.randrange(3), rng.randrange(max_seq + 1)
        t = trial.get(r, 0) if rng.random() < 0.8 else 0
        keys.add((r, v, s, t))
        trial[r] = trial.get(r, 0) + 1
    return ["mock=%d,%d,%d,%d" % k for k in sorted(keys)]


def one(seed):
    rng = random.Random(50000 + seed)
    prog = rng.choice(["model_recover", "model_recover", "local_recover", "lazy_recover"])
    world = rng.choice([2, 3, 4, 5, 6, 7, 8, 10])
    nbcast = len(range(0, world, max(world // 3, 1)))
    max_seq = 1 + 2 * nbcast + 1 + (1 if prog == "model_recover" else 0)
    kills = schedule(rng, world, max_seq, rng.randrange(1, 8))
    if prog == "local_recover":
        # a local model survives as long as not all of {rank, rank+1, rank+2} die at once
        seen, kept = {}, []
        for k in kills:
            r, v, s, _ = [int(x) for x in k[5:].split(",")]
            group = seen.setdefault((v, s), [])
            if len(group) < 2 and world > 3:
                group.append(r)
                kept.append(k)
        kills = kept
    args = [str(rng.choice([500, 500, 20000])), "rabit_timeout=1", "rabit_timeout_sec=60"] + kills
    if prog == "model_recover":
        args.append("rabit_bootstrap_cache=1")
    if rng.random() < 0.5:
        args.append("rabit_reduce_ring_mincount=1")
    if rng.random() < 0.3:
        args.append("rabit_reduce_buffer=4K")
    res = launch([os.path.join(ROOT, "build", "bin", prog)] + args, world, capture=True, timeout=150)
    ok = res.ok and all(("iter 2" in o or "start at iter=3" in o) for o in res.outputs.values())
    return ok, dict(seed=seed, prog=prog, world=world, args=args, tails={i: o[-600:] for i, o in res.outputs.items()})


if __name__ == "__main__":
    first, count = int(sys.argv[1]), int(sys.argv[2])
    bad, t0 = 0, time.time()
    for s in range(first, first + count):
        good, info = one(s)
        if not good:
            bad += 1
            os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
            with open(os.path.join(ROOT, "build", "fuzz_fail_%d.json" % s), "w") as fo:
                json.dump(info, fo, indent=1)
            print("FAIL", json.dumps({k: info[k] for k in ("seed", "prog", "world", "args")}), flush=True)
    print("%d runs, %d failures, %.0f s" % (count, bad, time.time() - t0))
    sys.exit(1 if bad else 0)
