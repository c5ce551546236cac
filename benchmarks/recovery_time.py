#!/usr/bin/env python
"""Recover-after-kill seconds (second half of the BASELINE metric).

Definition: (wall(run with k scripted kills) - wall(same run, no kills)) / k,
same launcher, same tracker, same kill script for both implementations:
  ours       build/bin/model_recover          (host engine over loopback TCP)
  reference  baseline/_ref/bin/ref_model_recover  (the unmodified reference, its own test program)
  device     tests/gpu_worker.py, one process per GPU (only with --gpus N on a GPU box)

    python benchmarks/recovery_time.py [--world 8] [--repeat 3] [--gpus 0] [--out profiles/recovery.json]
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200.parallel.launcher import launch  # noqa: E402

KILLS = {
    "2 kills (test.mk model_recover)": ["mock=0,0,1,0", "mock=1,1,1,0"],
    "4 kills, 3 at the same op (die_same)": ["mock=0,0,1,0", "mock=1,1,1,0", "mock=4,1,1,0", "mock=6,1,1,0"],
    "6 kills incl. a re-death (die_hard)": ["mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "mock=4,1,1,0",
                                            "mock=6,1,1,0", "mock=7,1,2,0"],
}


def wall(cmd, world, repeat, **kw):
    ts = []
    for _ in range(repeat):
        res = launch(cmd, world, capture=True, timeout=600, **kw)
        if not res.ok:
            raise RuntimeError("job failed: %r\n%s" % (res.returncodes, list(res.outputs.values())[0][-800:]))
        ts.append(res.wall)
    return statistics.median(ts), res.restarts


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--ndata", type=int, default=10000)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    impls = {
        "ours (host engine)": ([os.path.join(ROOT, "build", "bin", "model_recover"), str(args.ndata)], {}),
        "reference": ([os.path.join(ROOT, "baseline", "_ref", "bin", "ref_model_recover"), str(args.ndata)], {}),
    }
    if args.gpus > 0:
        env = {"PYTHONPATH": ROOT}
        impls = {"ours (device buffers, %d GPUs)" % args.gpus:
                 ([sys.executable, os.path.join(ROOT, "tests", "gpu_worker.py"), "200000"],
                  {"env": env, "per_rank_gpu": True}),
                 "ours (device buffers, native C++ worker, %d GPUs)" % args.gpus:
                 ([os.path.join(ROOT, "build", "bin", "device_recover_cuda"), "200000"], {"per_rank_gpu": True})}
        args.world = args.gpus
    common = ["rabit_bootstrap_cache=1", "rabit_enable_tcp_no_delay=true"]
    rows = []
    for name, (cmd, kw) in impls.items():
        if not os.path.exists(cmd[0] if cmd[0] != sys.executable else cmd[1]):
            continue
        base, _ = wall(cmd + common, args.world, args.repeat, **kw)
        rows.append({"impl": name, "scenario": "no kills", "world": args.world, "wall_s": base, "deaths": 0,
                     "sec_per_kill": None})
        print(json.dumps(rows[-1]), flush=True)
        for scen, kills in KILLS.items():
            kills = [k for k in kills if int(k.split("=")[1].split(",")[0]) < args.world]
            t, deaths = wall(cmd + common + kills, args.world, args.repeat, **kw)
            rows.append({"impl": name, "scenario": scen, "world": args.world, "wall_s": t, "deaths": deaths,
                         "sec_per_kill": (t - base) / max(deaths, 1)})
            print(json.dumps(rows[-1]), flush=True)
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
