#!/usr/bin/env python
"""Chunk-plan sweep of the streamed allreduce (where does the time go, which plan wins at this world size?).

    python benchmarks/stream_diag.py                                              one GPU (kernels forced on)
    python -m rabit_b200.parallel.launcher -n 8 --gpus python benchmarks/stream_diag.py [plans=8M:64M,32M:512M] [mib=256]

For every plan "first:chunk" (and for the single-launch kernels, "legacy"): device time per op on the async API
and on the blocking fault-tolerant API, on an ordinary tensor and on a symmetric one, max over ranks; block 0's
phase stamps of the LAST chunk kernel (us: entry barrier, reduce, mid barrier, epilogue / gather).
One JSON line per plan on rank 0.
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rabit_b200 import rabit, torch_binding as rt  # noqa: E402


def main():
    mib = 256
    plans = [("8M", "64M"), ("32M", "512M"), ("512M", "512M")]
    for a in sys.argv[1:]:
        if a.startswith("plans="):
            plans = [tuple(p.split(":")) for p in a[6:].split(",")]
        if a.startswith("mib="):
            mib = int(a[4:])
    quick = any(a == "quick=1" for a in sys.argv[1:])
    S = mib << 20
    ordinal = int(os.environ.get("RABIT_DEVICE_ORDINAL", os.environ.get("RABIT_LOCAL_RANK", "0")))
    torch.cuda.set_device(ordinal % torch.cuda.device_count())
    passthrough = [a for a in sys.argv if a.startswith("rabit_") or a.startswith("DMLC_")]
    for first, chunk in plans + ([] if quick else [("legacy", "legacy")]):
        args = passthrough + ["rabit_device=on", "rabit_device_arena=%d" % (2 * S + (4 << 20)), "rabit_device_force=1",
                              "rabit_device_cache=32G", "rabit_device_heap=%d" % (S + (4 << 20))]
        if first == "legacy":
            args += ["rabit_device_algo=nostream"]
        else:
            args += ["rabit_device_stream_first=" + first, "rabit_device_stream_chunk=" + chunk, "rabit_device_stream_min=1M"]
        args += [a for a in sys.argv[1:] if a.startswith("rabit_device_host_chunk=")]
        rabit.init(args)
        me, n = rabit.get_rank(), rabit.get_world_size()
        x = torch.ones(S // 4, device="cuda")
        xs = rt.symmetric_empty(S // 4, torch.float32) if n > 1 else x
        rt.set_tuning(trace=True)

        def timed(fn, sync, steps=10, warm=3):
            for _ in range(warm):
                fn()
            sync()
            rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            sync()
            torch.cuda.synchronize()
            return float(rabit.allreduce(np.array([e0.elapsed_time(e1) / steps]), rabit.MAX)[0])
        row = {"world": n, "mib": mib, "first": first, "chunk": chunk}
        row["async_plain_ms"] = timed(lambda: rt.allreduce_async_(x), rt.synchronize)
        tr = rt.get_trace()
        row["last_kernel_phases_us"] = [round((tr[i] - tr[j]) / 1e3, 1) for i, j in ((2, 0), (3, 2), (4, 3), (5, 4))]
        for t, key in ((x, "robust_plain_ms"), (xs, "robust_sym_ms")):
            if key == "robust_sym_ms" and (n == 1 or quick):
                continue
            for _ in range(14):
                rabit.allreduce(t, rabit.SUM)
            rabit.checkpoint(0)
            t.fill_(1.0)
            row[key] = timed(lambda: rabit.allreduce(t, rabit.SUM), torch.cuda.synchronize)
            if key == "robust_plain_ms":
                tr = rt.get_trace()
                row["robust_last_kernel_phases_us"] = [round((tr[i] - tr[j]) / 1e3, 1) for i, j in ((2, 0), (3, 2), (4, 3), (5, 4))]
            rabit.checkpoint(0)
        if n > 1 and not quick:
            row["async_sym_ms"] = timed(lambda: rt.allreduce_async_(xs), rt.synchronize)
        if quick:
            if me == 0:
                print("DIAG " + json.dumps(row), flush=True)
            rabit.finalize()
            continue
        # pinned host tensor through the same call (H2D | kernels | D2H), wall clock
        import time
        h = torch.ones(S // 4).pin_memory()
        for _ in range(6):
            rt.allreduce_(h)
        rabit.checkpoint(0)
        h.fill_(1e-6)
        rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
        t0 = time.perf_counter()
        for _ in range(5):
            rt.allreduce_(h)
        row["host_tensor_ms"] = float(rabit.allreduce(np.array([(time.perf_counter() - t0) * 1e3 / 5]), rabit.MAX)[0])
        rabit.checkpoint(0)
        if me == 0:
            print("DIAG " + json.dumps(row), flush=True)
        rabit.finalize()


if __name__ == "__main__":
    main()
