#!/usr/bin/env python
"""Copy-engine probe: how fast are the cudaMemcpyAsync copies the streamed allreduce leans on?

    python benchmarks/ce_probe.py            (one GPU)

Prints one JSON line: D2D (same device), H2D and D2H from pinned memory, alone and both directions at once,
and D2D while an SM kernel of this library (a large emulated allreduce) is running.  Timed with CUDA events,
3 warm-up + 10 timed copies of 256 MiB (> the 126 MB L2).
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from cuda.bindings import runtime as rt  # noqa: E402


def memcpy_async(dst, src, nbytes, kind, stream):
    err, = rt.cudaMemcpyAsync(dst, src, nbytes, kind, stream.cuda_stream)
    assert err == rt.cudaError_t.cudaSuccess, err


def timed(fn, streams, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = [torch.cuda.Event(enable_timing=True) for _ in streams]
    e0.record(torch.cuda.current_stream())
    for s in streams:
        s.wait_event(e0)
    for _ in range(reps):
        fn()
    for s, e in zip(streams, e1):
        e.record(s)
    torch.cuda.synchronize()
    return max(e0.elapsed_time(e) for e in e1) / reps


def main():
    torch.cuda.set_device(0)
    S = 256 << 20
    a = torch.empty(S, dtype=torch.uint8, device="cuda")
    b = torch.empty(S, dtype=torch.uint8, device="cuda")
    h_in = torch.empty(S, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(S, dtype=torch.uint8).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    K = rt.cudaMemcpyKind
    out = {"bytes": S}
    ms = timed(lambda: memcpy_async(b.data_ptr(), a.data_ptr(), S, K.cudaMemcpyDeviceToDevice, s1), [s1])
    out["d2d_GBps"] = S / ms / 1e6
    ms = timed(lambda: b.copy_(a), [torch.cuda.current_stream()])
    out["d2d_sm_kernel_GBps"] = S / ms / 1e6
    ms = timed(lambda: memcpy_async(a.data_ptr(), h_in.data_ptr(), S, K.cudaMemcpyHostToDevice, s1), [s1])
    out["h2d_GBps"] = S / ms / 1e6
    ms = timed(lambda: memcpy_async(h_out.data_ptr(), b.data_ptr(), S, K.cudaMemcpyDeviceToHost, s2), [s2])
    out["d2h_GBps"] = S / ms / 1e6

    def both():
        memcpy_async(a.data_ptr(), h_in.data_ptr(), S, K.cudaMemcpyHostToDevice, s1)
        memcpy_async(h_out.data_ptr(), b.data_ptr(), S, K.cudaMemcpyDeviceToHost, s2)
    ms = timed(both, [s1, s2])
    out["h2d_plus_d2h_ms"] = ms
    out["h2d_plus_d2h_each_GBps"] = S / ms / 1e6

    def two_d2d():
        memcpy_async(b.data_ptr(), a.data_ptr(), S // 2, K.cudaMemcpyDeviceToDevice, s1)
        memcpy_async(b.data_ptr() + S // 2, a.data_ptr() + S // 2, S // 2, K.cudaMemcpyDeviceToDevice, s2)
    ms = timed(two_d2d, [s1, s2])
    out["d2d_two_streams_GBps"] = S / ms / 1e6
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
