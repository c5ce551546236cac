#!/usr/bin/env python
"""Host-side overhead of the blocking, fault-tolerant API on device tensors (1 GPU, kernels forced on)."""
import os, sys, time, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200 import rabit, torch_binding as rt
torch.cuda.set_device(0)
rabit.init(["rabit_device=on", "rabit_device_force=1"])
for nbytes in (4096, 4096000, 64 << 20):
    t = torch.zeros(nbytes // 8, dtype=torch.float64, device="cuda")
    for _ in range(5):
        rabit.allreduce(t, rabit.SUM)
    rabit.checkpoint(0)
    res = {}
    for name, every in (("no_checkpoint", 0), ("checkpoint_every_10", 10)):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for i in range(100):
            rabit.allreduce(t, rabit.SUM)
            if every and i % every == every - 1:
                rabit.checkpoint(0)
        torch.cuda.synchronize()
        res[name] = (time.perf_counter() - t0) * 1e6 / 100
        rabit.checkpoint(0)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(100):
        rt.allreduce_async_(t)
    rt.synchronize(); torch.cuda.synchronize()
    res["async_api"] = (time.perf_counter() - t0) * 1e6 / 100
    print(json.dumps({"bytes": nbytes, "us_per_op": res}))
rabit.finalize()
