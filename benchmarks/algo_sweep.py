#!/usr/bin/env python
"""Tuning tool: time every allreduce algorithm x block count x message size on N GPUs
(torchrun, one rank per GPU) and print block 0's phase breakdown.  Device-timed, max over ranks.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29720 \
        benchmarks/algo_sweep.py [--sizes 4194304,268435456] [--algos twoshot,tma,nvls] [--blocks 0,74,148,296]

    ... benchmarks/algo_sweep.py --emit-table profiles/dispatch_table.txt
        sweeps 1 KiB - 4 MiB over {ll, oneshot, twoshot, nvls} and appends the measured thresholds of this world
        size as lines "world key value" (the format the backend reads: rabit_device_tune=<file>)
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from rabit_b200 import rabit, torch_binding as rt  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1024,65536,1048576,16777216,268435456")
    ap.add_argument("--algos", default="oneshot,twoshot,tma,nvls")
    ap.add_argument("--blocks", default="0")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--sym", action="store_true", help="operate on a symmetric (zero-copy) buffer")
    ap.add_argument("--emit-table", default="", help="append the measured dispatch thresholds of this world size to this file")
    args = ap.parse_args()
    if args.emit_table:
        args.sizes = ",".join(str(1 << k) for k in range(10, 27))
        args.algos = "ll,oneshot,twoshot,nvls"
        args.blocks = "0"
        args.iters = max(args.iters, 20)
    rank, world, local = bench.dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    store = bench.make_store(rank, world)
    tracker, port = bench.start_tracker(rank, world, store)
    sizes = [int(s) for s in args.sizes.split(",")]
    init = ["rabit_device=on", "rabit_device_arena=%d" % (2 * max(sizes) + (4 << 20)), "rabit_device_stream_min=64G"]
    if args.sym:
        init += ["rabit_device_heap=%d" % (max(sizes) + (1 << 20))]
    if world > 1:
        init += ["rabit_tracker_uri=127.0.0.1", "rabit_tracker_port=%d" % port, "rabit_task_id=%d" % rank]
    else:
        init += ["rabit_device_force=1"]
    rabit.init(init)
    n = rabit.get_world_size()
    me = rabit.get_rank()
    if args.sym:
        x = rt.symmetric_empty(max(sizes) // 4, torch.float32)
    else:
        x = torch.zeros(max(sizes) // 4, dtype=torch.float32, device=dev)
    rows = []
    for sz in sizes:
        v = x[:sz // 4]
        for algo in args.algos.split(","):
            if algo == "oneshot" and sz > (8 << 20):
                continue
            if algo == "ll" and sz > 32768:
                continue
            for blocks in [int(b) for b in args.blocks.split(",")]:
                rt.set_tuning(algo, blocks, trace=False)
                # correctness first: small integers, exact in fp32
                v.copy_((torch.arange(v.numel(), device=dev, dtype=torch.int32) % 61).float() * (me + 1))
                rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
                torch.cuda.synchronize()
                rt.allreduce_async_(v)
                rt.synchronize()
                want = (torch.arange(v.numel(), device=dev, dtype=torch.int32) % 61).float() * float(n * (n + 1) // 2)
                ok = bool(torch.equal(v, want))
                del want
                v.zero_()
                for _ in range(3):
                    rt.allreduce_async_(v)
                rt.synchronize()
                rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.iters):
                    rt.allreduce_async_(v)
                e1.record()
                rt.synchronize()
                torch.cuda.synchronize()
                ms = bench.max_over_ranks(rabit, np, e0.elapsed_time(e1) / args.iters)
                rt.set_tuning(algo, blocks, trace=True)
                rt.allreduce_async_(v)
                rt.synchronize()
                tr = rt.get_trace()
                ph = [(tr[i + 1] - tr[i]) / 1e3 if tr[i + 1] >= tr[i] else -1 for i in range(5)]
                row = {"bytes": sz, "algo": ("sym-" if args.sym else "") + algo, "blocks": blocks, "correct": ok, "us": ms * 1e3,
                       "busbw_GBps": sz / (ms * 1e-3) / 1e9 * 2 * (n - 1) / max(n, 1),
                       "phases_us(stage,barrierA,rs,barrierB,ag)": [round(p, 1) for p in ph]}
                rows.append(row)
                if rank == 0:
                    print(json.dumps(row), flush=True)
    rt.set_tuning("auto", 0, False)
    desc = rt.describe()
    rabit.finalize()
    if rank == 0 and args.emit_table:
        best = {}
        for r in rows:
            if r["correct"] and (r["bytes"] not in best or r["us"] < best[r["bytes"]][1]):
                best[r["bytes"]] = (r["algo"], r["us"])
        by = sorted(best)
        ll_max = max([b for b in by if best[b][0] == "ll"] or [0])
        # nvls from the first size after which it never loses again
        nvls_min = None
        for b in reversed(by):
            if best[b][0] == "nvls":
                nvls_min = b
            else:
                break
        # one-shot vs two-shot among the sizes the switch does not take
        t = dict(((r["bytes"], r["algo"]), r["us"]) for r in rows if r["correct"])
        oneshot_max = max([b for b in by if (b, "oneshot") in t and t[(b, "oneshot")] <= t.get((b, "twoshot"), 1e30)] or [0])
        with open(args.emit_table, "a") as fo:
            fo.write("# world %d: winners %s\n" % (n, " ".join("%d:%s:%.1fus" % (b, best[b][0], best[b][1]) for b in by)))
            fo.write("%d ll_max %d\n" % (n, ll_max))
            fo.write("%d oneshot_max %d\n" % (n, oneshot_max))
            if nvls_min is not None:
                fo.write("%d nvls_min %d\n" % (n, nvls_min))
        print("TABLE world %d ll_max %d nvls_min %s" % (n, ll_max, nvls_min))
    if rank == 0:
        print(desc)
        if args.out:
            json.dump({"world": n, "backend": desc, "rows": rows}, open(args.out, "w"), indent=1)
    if tracker is not None:
        tracker.join(10)
        tracker.close()


if __name__ == "__main__":
    main()
