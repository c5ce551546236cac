"""SIGKILL random workers at random wall-clock moments (mid-transfer, mid-rendezvous, mid-commit) and restart them.

The scripted `mock=` kills only fire at the entry of a collective; this harness covers everything in between.
Used by tests/test_fault_chaos.py (a few fixed seeds) and runnable as a long campaign:
    python tests/chaos_harness.py <first_seed> <count> [--local]
"""
import json
import os
import random
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200.parallel.tracker import Tracker  # noqa: E402


def run(seed, programs=("model_recover", "model_recover", "lazy_recover"), timeout=120, extra_args=(), tail=700,
        on_stuck=None):
    rng = random.Random(seed)
    world = rng.choice([2, 3, 4, 6, 8])
    prog = rng.choice(list(programs))
    n = rng.choice([2000, 50000, 300000])
    iters = rng.choice([10, 20])
    nkill = rng.randrange(1, 9)
    gap = (0.005, 0.12)
    args = [str(n), "iters=%d" % iters, "rabit_timeout=1", "rabit_timeout_sec=40"]
    if prog == "model_recover":
        args.append("rabit_bootstrap_cache=1")
    if rng.random() < 0.5:
        args.append("rabit_reduce_ring_mincount=1")
    args += list(extra_args)
    tracker = Tracker(world).start()
    procs, outs, attempts, done = {}, {i: b"" for i in range(world)}, {i: 0 for i in range(world)}, {}

    def spawn(i):
        env = dict(os.environ)
        env.update(tracker.worker_env())
        env["DMLC_TASK_ID"] = str(i)
        env["DMLC_NUM_ATTEMPT"] = str(attempts[i])
        bindir = os.environ.get("RABIT_CHAOS_BIN", os.path.join(ROOT, "build", "bin"))   # e.g. build/san_address
        procs[i] = subprocess.Popen([os.path.join(bindir, prog)] + args, env=env,
                                    stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        os.set_blocking(procs[i].stdout.fileno(), False)

    def drain(i):
        # keep the pipe empty: a chatty worker (rabit_debug=1) must not block on a full pipe while it is alive
        try:
            while True:
                chunk = os.read(procs[i].stdout.fileno(), 1 << 16)
                if not chunk:
                    return
                outs[i] += chunk
        except (BlockingIOError, OSError, ValueError):
            return

    for i in range(world):
        spawn(i)
    kills = []
    next_kill = time.time() + rng.uniform(*gap)
    deadline = time.time() + timeout
    while len(done) < world and time.time() < deadline:
        time.sleep(0.002)
        for i in range(world):
            if i in done:
                continue
            drain(i)
            rc = procs[i].poll()
            if rc is None:
                continue
            drain(i)
            if rc == 0:
                done[i] = 0
            else:
                attempts[i] += 1
                if attempts[i] > 30:
                    done[i] = rc
                else:
                    spawn(i)
        if len(kills) < nkill and time.time() >= next_kill:
            live = [i for i in range(world) if i not in done and procs[i].poll() is None]
            if len(live) >= 2:              # somebody has to survive
                victim = rng.choice(live)
                procs[victim].kill()        # the exact process this harness started
                kills.append(victim)
            next_kill = time.time() + rng.uniform(*gap)
    ok = len(done) == world and all(rc == 0 for rc in done.values())
    if not ok and on_stuck is not None:
        on_stuck({i: procs[i] for i in range(world) if procs[i].poll() is None}, tracker)   # post-mortem before the kill
    for i in range(world):
        if procs[i].poll() is None:
            procs[i].kill()
        procs[i].wait()
        drain(i)
    tracker.close()
    return ok, dict(seed=seed, world=world, prog=prog, args=args, kills=kills, done=done,
                    tails={i: outs[i].decode(errors="replace")[-tail:] for i in range(world)})


if __name__ == "__main__":
    first, count = int(sys.argv[1]), int(sys.argv[2])
    progs = ("model_recover", "model_recover", "lazy_recover") + (("local_recover",) if "--local" in sys.argv else ())
    fails, delivered, t0 = [], 0, time.time()
    for s in range(first, first + count):
        good, info = run(s, progs)
        delivered += len(info["kills"])
        if not good:
            fails.append(info)
            os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
            with open(os.path.join(ROOT, "build", "chaos_fail_%d.json" % s), "w") as fo:
                json.dump(info, fo, indent=1)
            print("FAIL", json.dumps({k: info[k] for k in ("seed", "prog", "world", "args", "kills", "done")}), flush=True)
    print("%d runs, %d kills delivered, %d failures, %.0f s" % (count, delivered, len(fails), time.time() - t0))
    sys.exit(1 if fails else 0)
