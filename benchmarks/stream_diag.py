#!/usr/bin/env python
"""Where does the time of a streamed allreduce go?  One GPU (rabit_device_force), or one rank per GPU under the
launcher.  For a few chunk plans: device time per 256 MiB op on the async and the blocking API, and block 0's
phase stamps of the LAST chunk kernel (us: entry barrier, reduce, mid barrier, epilogue / gather)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rabit_b200 import rabit, torch_binding as rt  # noqa: E402


def main():
    S = 256 << 20
    plans = [("8M", "64M"), ("256M", "256M"), ("16M", "32M")]
    ordinal = int(os.environ.get("RABIT_DEVICE_ORDINAL", os.environ.get("RABIT_LOCAL_RANK", "0")))
    torch.cuda.set_device(ordinal % torch.cuda.device_count())
    out = []
    for first, chunk in plans + [("legacy", "legacy")]:
        args = list(sys.argv) + ["rabit_device=on", "rabit_device_arena=%d" % (2 * S + (4 << 20)), "rabit_device_force=1",
                                 "rabit_device_cache=32G"]
        if first == "legacy":
            args += ["rabit_device_algo=nostream"]
        else:
            args += ["rabit_device_stream_first=" + first, "rabit_device_stream_chunk=" + chunk]
        rabit.init(args)
        me = rabit.get_rank()
        x = torch.ones(S // 4, device="cuda")
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
        row = {"first": first, "chunk": chunk}
        row["async_ms"] = timed(lambda: rt.allreduce_async_(x), rt.synchronize)
        tr = rt.get_trace()
        row["last_kernel_phases_us"] = [round((tr[i] - tr[j]) / 1e3, 1) for i, j in ((2, 0), (3, 2), (4, 3), (5, 4))]
        for _ in range(14):
            rabit.allreduce(x, rabit.SUM)
        rabit.checkpoint(0)
        x.fill_(1.0)
        row["robust_ms"] = timed(lambda: rabit.allreduce(x, rabit.SUM), torch.cuda.synchronize)
        rabit.checkpoint(0)
        if me == 0:
            out.append(row)
            print(json.dumps(row), flush=True)
        rabit.finalize()


if __name__ == "__main__":
    main()
