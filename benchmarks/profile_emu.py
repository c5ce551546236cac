#!/usr/bin/env python
"""Single-GPU workload for ncu: 8 emulated ranks allreduce 32 MiB of float32 each through the
two-shot (and TMA) kernel, one cooperative grid per call.  Prints CUDA-event timings so that the
plain run gives the number and the ncu run only gives the counters.

    python benchmarks/profile_emu.py [--algo 2] [--mib 32] [--iters 5]
    python benchmarks/profile_emu.py --algo 10 --chunk-mib 8     (the streamed allreduce: StreamP2P per chunk)
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200.ops import emulate as emu  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--algo", type=int, default=2)
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--mib", type=int, default=32)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--chunk-mib", type=int, default=8)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    count = args.mib * (1 << 20) // 4
    bufs = [torch.full((count,), float(r + 1), device="cuda") for r in range(args.ranks)]
    def run():
        if args.algo == 10:
            assert emu.stream_allreduce(bufs, op=1, chunk_bytes=args.chunk_mib << 20) == [0] * args.ranks
        else:
            emu.collective(emu.ALLREDUCE, bufs, algo=args.algo, op=1)   # Min keeps values bounded
    for _ in range(3):
        run()
    times = []
    for _ in range(args.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    assert all(float(b[0]) == 1.0 for b in bufs)
    ms = min(times)
    nbytes = count * 4
    # HBM traffic of the emulated job: stage copy (r+w), reduce-scatter (n reads + 1 write per tile), allgather (r+w)
    traffic = args.ranks * nbytes * (2 + 1 + 1.0 / args.ranks + 2)
    print(json.dumps({"algo": args.algo, "ranks": args.ranks, "bytes_per_rank": nbytes, "best_ms": ms,
                      "hbm_traffic_model_GB": traffic / 1e9, "hbm_GBps_model": traffic / (ms * 1e-3) / 1e9}))


if __name__ == "__main__":
    main()
