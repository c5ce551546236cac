"""Self-checking device-buffer workload, one process per GPU (launched by test_gpu_multi.py).

The GPU twin of test/model_recover.cc: every collective runs on CUDA tensors through the
public API (robust engine -> fused kernels), results are compared with closed forms, the
model is checkpointed every iteration, and `mock=rank,version,seq,trial` kills ranks so
that LoadCheckPoint + result-cache replay are exercised on device buffers.
usage: gpu_worker.py <ndata> [iters=N] [mock=...] [rabit_*=...]
"""
import os
import sys

import numpy as np
import torch

from rabit_b200 import rabit
from rabit_b200 import torch_binding as rt


def contribution(n, r, it, mod, dev):
    i = torch.arange(n, device=dev, dtype=torch.int64)
    return ((i * (r + 1) + it) % mod).float()


def main():
    n = int(sys.argv[1])
    iters = 3
    for a in sys.argv[2:]:
        if a.startswith("iters="):
            iters = int(a[6:])
    ordinal = int(os.environ.get("RABIT_DEVICE_ORDINAL", os.environ.get("RABIT_LOCAL_RANK", "0")))
    torch.cuda.set_device(ordinal % torch.cuda.device_count())
    dev = torch.device("cuda", torch.cuda.current_device())
    rabit.init(sys.argv, lib="mock")
    me, world = rabit.get_rank(), rabit.get_world_size()
    if any(a == "staged=1" for a in sys.argv):
        # no arena (rabit_device=off / ranks on several hosts): device pointers are staged through host TCP
        assert not rt.device_active()
    else:
        assert rt.device_active(), rt.describe()
    grow = any(a == "grow=1" for a in sys.argv)
    graph = any(a == "graph=1" for a in sys.argv)
    sym = None
    if any(a.startswith("rabit_device_heap=") for a in sys.argv):
        # collective allocation, same order on every rank (also after a restart)
        sym = (rt.symmetric_empty(1 << 20, torch.float32), rt.symmetric_empty(2 << 20, torch.uint8))
    version, model = rabit.load_checkpoint()
    if version == 0:
        model = np.ones(n, dtype=np.float32)
    model = torch.from_numpy(model).to(dev)
    print("[%d] start at iter=%d on %s | %s" % (me, version, dev, rt.describe()), flush=True)
    for it in range(version, iters):
        # max with lazy prepare
        x = torch.empty(n, device=dev)

        def fill(t):
            t.copy_(contribution(n, me, it, 97 + it, dev) + model)
        rabit.allreduce(x, rabit.MAX, fill)
        want = torch.stack([contribution(n, r, it, 97 + it, dev) + model for r in range(world)]).max(0).values
        assert torch.equal(x, want), "[%d] max mismatch iter %d" % (me, it)
        # broadcast bytes from a rotating root
        root = it % world
        b = torch.full((n,), 255, dtype=torch.uint8, device=dev)
        truth = ((torch.arange(n, device=dev) * 7 + root + it) % 251).to(torch.uint8)
        if me == root:
            b.copy_(truth)
        rt.broadcast_(b, root)
        assert torch.equal(b, truth), "[%d] broadcast mismatch iter %d" % (me, it)
        # sum
        y = contribution(n, me, it, 113 + it, dev) + x
        rabit.allreduce(y, rabit.SUM)
        want = torch.stack([contribution(n, r, it, 113 + it, dev) + x for r in range(world)]).sum(0)
        assert torch.equal(y, want), "[%d] sum mismatch iter %d" % (me, it)
        # grad/hess pair histogram: float64 sum over 2x elements
        h = (contribution(2 * 256, me, it, 89, dev).double() + 0.5)
        rabit.allreduce(h, rabit.SUM)
        want = torch.stack([contribution(2 * 256, r, it, 89, dev).double() + 0.5 for r in range(world)]).sum(0)
        assert torch.equal(h, want), "[%d] pair-sum mismatch iter %d" % (me, it)
        # allgather
        per = 33
        g = torch.zeros(per * world, dtype=torch.int32, device=dev)
        g[me * per:(me + 1) * per] = me * 1000 + it
        rt.allgather_(g, me * per, per)
        want = torch.arange(world, device=dev, dtype=torch.int32).repeat_interleave(per) * 1000 + it
        assert torch.equal(g, want), "[%d] allgather mismatch iter %d" % (me, it)
        # large (>= 1 MiB) broadcast: scatter + multicast-allgather kernel where NVLS is up
        big = 3 * (1 << 20) + 4096
        bb = torch.zeros(big, dtype=torch.uint8, device=dev)
        btruth = ((torch.arange(big, device=dev) * 13 + root + it) % 253).to(torch.uint8)
        if me == root:
            bb.copy_(btruth)
        rt.broadcast_(bb, root)
        assert torch.equal(bb, btruth), "[%d] large broadcast mismatch iter %d" % (me, it)
        if sym is not None:
            # zero-copy path: the same collectives on buffers from the symmetric heap
            sx, sb = sym
            sx.copy_(contribution(sx.numel(), me, it, 101, dev))
            rabit.allreduce(sx, rabit.SUM)
            want = torch.stack([contribution(sx.numel(), r, it, 101, dev) for r in range(world)]).sum(0)
            assert torch.equal(sx, want), "[%d] symmetric allreduce mismatch iter %d" % (me, it)
            sx.copy_(contribution(sx.numel(), me, it, 103, dev))
            rabit.allreduce(sx, rabit.MAX)
            want = torch.stack([contribution(sx.numel(), r, it, 103, dev) for r in range(world)]).max(0).values
            assert torch.equal(sx, want), "[%d] symmetric max mismatch iter %d" % (me, it)
            if not any(a.startswith("mock=") for a in sys.argv) and "RABIT_TEST_STREAM_KILL" not in os.environ:
                # the async API reduces heap buffers in place (zero-copy kernels); it is not sequence-numbered,
                # so it cannot be part of a run that replays iterations after a kill
                sx.copy_(contribution(sx.numel(), me, it, 107, dev))
                rt.allreduce_async_(sx)
                rt.synchronize()
                want = torch.stack([contribution(sx.numel(), r, it, 107, dev) for r in range(world)]).sum(0)
                assert torch.equal(sx, want), "[%d] in-place async allreduce mismatch iter %d" % (me, it)
            if me == root:
                sb.copy_(btruth[:sb.numel()])
            else:
                sb.zero_()
            rt.broadcast_(sb, root)
            assert torch.equal(sb, btruth[:sb.numel()]), "[%d] symmetric broadcast mismatch iter %d" % (me, it)
        if any(a == "stream=1" for a in sys.argv):
            # the chunked (streamed) allreduce on an ordinary tensor, on a symmetric tensor and on a host tensor:
            # small rabit_device_stream_* values make a few MiB run as several chunk kernels
            sn = 3 * (1 << 20) + 1000
            t = contribution(sn, me, it, 127, dev)
            rabit.allreduce(t, rabit.SUM)
            want = torch.stack([contribution(sn, r, it, 127, dev) for r in range(world)]).sum(0)
            assert torch.equal(t, want), "[%d] streamed allreduce mismatch iter %d" % (me, it)
            t = contribution(sn, me, it, 131, dev)
            rabit.allreduce(t, rabit.MAX)
            want = torch.stack([contribution(sn, r, it, 131, dev) for r in range(world)]).max(0).values
            assert torch.equal(t, want), "[%d] streamed max mismatch iter %d" % (me, it)
            td = contribution(sn // 2, me, it, 59, dev).double()
            rabit.allreduce(td, rabit.SUM)
            want = torch.stack([contribution(sn // 2, r, it, 59, dev).double() for r in range(world)]).sum(0)
            assert torch.equal(td, want), "[%d] streamed f64 mismatch iter %d" % (me, it)
            if sym is not None:
                sx = sym[0]
                sx.copy_(contribution(sx.numel(), me, it, 137, dev))
                rabit.allreduce(sx, rabit.SUM)
                want = torch.stack([contribution(sx.numel(), r, it, 137, dev) for r in range(world)]).sum(0)
                assert torch.equal(sx, want), "[%d] streamed symmetric allreduce mismatch iter %d" % (me, it)
            th = contribution(sn, me, it, 139, dev).cpu().pin_memory()
            rt.allreduce_(th, rabit.SUM)
            want = torch.stack([contribution(sn, r, it, 139, dev) for r in range(world)]).sum(0).cpu()
            assert torch.equal(th, want), "[%d] streamed host allreduce mismatch iter %d" % (me, it)
            # async flavour of the same path (no undo, no cache; not sequence-numbered, so not in runs with kills)
            if any(a.startswith("mock=") for a in sys.argv) or "RABIT_TEST_STREAM_KILL" in os.environ:
                continue_async = False
            else:
                continue_async = True
            t = contribution(sn, me, it, 149, dev)
            if continue_async:
                rt.allreduce_async_(t)
                rt.synchronize()
                want = torch.stack([contribution(sn, r, it, 149, dev) for r in range(world)]).sum(0)
                assert torch.equal(t, want), "[%d] streamed async allreduce mismatch iter %d" % (me, it)
        if any(a == "hostpipe=1" for a in sys.argv):
            # large HOST buffers ride PCIe + NVLink in chunks (rabit_device_host_min): pageable numpy and pinned torch
            hn = 3 * (1 << 20) + 17
            base = (np.arange(hn, dtype=np.int64) % 61).astype(np.float32)
            out = rabit.allreduce(base * (me + 1) + it, rabit.SUM)
            want_h = base * (world * (world + 1) // 2) + it * world
            assert np.array_equal(out, want_h), "[%d] host-pipe numpy allreduce mismatch iter %d" % (me, it)
            pinned = torch.from_numpy(base * (me + 2)).pin_memory()
            rt.allreduce_(pinned, rabit.MAX)
            assert np.array_equal(pinned.numpy(), base * (world + 1)), "[%d] host-pipe pinned max mismatch" % me
            hd = (np.arange(2 * 300000, dtype=np.int64) % 53).astype(np.float64) + me
            out = rabit.allreduce(hd, rabit.SUM)
            want_d = (np.arange(2 * 300000, dtype=np.int64) % 53).astype(np.float64) * world + world * (world - 1) / 2
            assert np.array_equal(out, want_d), "[%d] host-pipe f64 allreduce mismatch iter %d" % (me, it)
        if graph:
            # capture three small collectives once, replay them several times on every rank
            gx = torch.zeros(2048, device=dev)
            gy = torch.zeros(64, device=dev, dtype=torch.float64)
            side = torch.cuda.Stream()
            with torch.cuda.stream(side):
                rt.allreduce_async_(gx)
                rt.allreduce_async_(gy)
            rt.synchronize()
            torch.cuda.synchronize()
            cg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(cg, stream=side):
                rt.allreduce_async_(gx)                 # LL kernel (<= 4 KiB? no: 8 KiB -> one-shot / NVLS)
                rt.allreduce_async_(gy, rabit.MAX)      # LL kernel
                rt.broadcast_async_(gx, root)
            for rep in range(4):
                gx.fill_(float(me + rep))
                gy.fill_(float(me * 10 + rep))
                torch.cuda.synchronize()
                rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)   # host barrier: every rank has refilled
                cg.replay()
                torch.cuda.synchronize()
                want = float(sum(r + rep for r in range(world)))
                assert torch.equal(gx, torch.full_like(gx, want)), "[%d] graph replay %d allreduce mismatch" % (me, rep)
                assert torch.equal(gy, torch.full_like(gy, float((world - 1) * 10 + rep))), "[%d] graph max mismatch" % me
            rt.allreduce_async_(gx)                     # eager op after the replays
            rt.synchronize()
        if grow:
            # 96 MB > the default 64 MB stage: the arena is re-allocated and re-exchanged on the fly
            gcount = 24 * (1 << 20)
            gx = (torch.arange(gcount, device=dev, dtype=torch.int32) % 7).float() + me
            rabit.allreduce(gx, rabit.SUM)
            want = (torch.arange(gcount, device=dev, dtype=torch.int32) % 7).float() * world + world * (world - 1) / 2
            assert torch.equal(gx, want), "[%d] allreduce after arena growth mismatch iter %d" % (me, it)
            del gx, want
        model = y % 1000.0
        rabit.checkpoint(model.cpu().numpy())
        print("[%d] iter %d checkpointed (version %d)" % (me, it, rabit.version_number()), flush=True)
    launches = rt.launch_count()
    rabit.tracker_print("[%d] done, %d kernel launches\n" % (me, launches))
    assert launches > 0 or any(a == "staged=1" for a in sys.argv)
    rabit.finalize()


if __name__ == "__main__":
    main()
