#!/usr/bin/env python
"""Headline benchmark: speed_test-style allreduce (float32, Sum) through rabit-b200.

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference --gpus N ...      (the unmodified reference, host TCP engine)

Metric (BASELINE.json: "allreduce/broadcast bus GB/s vs msg size"; definition from the
reference's test/speed_test.cc:52-69, bytes / mean time):
  value = N * S / t   aggregate bytes allreduced per second over all N GPUs, S = 256 MiB of
          float32 per rank (fixed as N grows: weak scaling).  Per-rank algorithm bandwidth is
          S / t and bus bandwidth is (S / t) * 2 (N-1) / N; both are in the JSON, and `sweep`
          has the bus-bandwidth curve from 1 KiB to 1 GiB for allreduce and broadcast.
  A "step" is one in-place allreduce of the S-byte device buffer.  Timed on the device with
  CUDA events after a barrier + synchronize, max over ranks.  The 256 MiB operands exceed the
  126 MB L2, so consecutive steps cannot be served from cache.
  e2e: the public blocking API (rabit.allreduce on a tensor, robust engine with fused consensus
  and result cache) including the H2D copy of the input from pinned memory and the D2H read of
  the whole result, every step.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HEADLINE_BYTES = 256 << 20
METRIC = "allreduce_f32_sum_aggregate_GBps"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bytes", type=int, default=HEADLINE_BYTES)
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--algo", default="auto")
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    return rank, world, local


class ClockSampler(object):
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.proc = None
        self.path = None
        self.gpu = gpu_index

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, reasons = [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    sm.append(float(f[1]))
                    out["sm_max_mhz"] = float(f[2])
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.remove(self.path)
        except OSError:
            pass
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
        out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def ensure_built(rank):
    """lib/librabit.so normally travels with the snapshot; build it (rank 0 only) if it is missing."""
    lib = os.path.join(ROOT, "lib", "librabit.so")
    if os.path.isfile(lib):
        return
    if rank == 0:
        import build as _build
        _build.build(cuda=True, verbose=False)
    else:
        for _ in range(1800):
            if os.path.isfile(lib):
                time.sleep(2.0)
                return
            time.sleep(1.0)


def make_store(rank, world):
    """torchrun's MASTER_ADDR/PORT rendezvous as a key-value store (also used to share the tracker port)."""
    import datetime
    import torch.distributed as dist
    if world == 1:
        return None
    return dist.TCPStore(os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ.get("MASTER_PORT", "29500")),
                         world, rank == 0, timeout=datetime.timedelta(seconds=300), wait_for_workers=True)


def start_tracker(rank, world, store):
    """rank 0 hosts the rabit tracker on a free port and publishes it through the store"""
    if world == 1:
        return None, None
    if rank == 0:
        from rabit_b200.parallel.tracker import Tracker
        tr = Tracker(world, host="127.0.0.1").start()
        store.set("rabit_tracker_port", str(tr.port))
        return tr, tr.port
    return None, int(store.get("rabit_tracker_port").decode())


def max_over_ranks(rabit, np, value):
    return float(rabit.allreduce(np.array([value], dtype=np.float64), rabit.MAX)[0])


# --------------------------------------------------------------------------- ours
def run_ours(args):
    import numpy as np
    import torch
    rank, world, local = dist_env()
    ensure_built(rank)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node == --gpus"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    store = make_store(rank, world)
    tracker, port = start_tracker(rank, world, store)
    from rabit_b200 import rabit, torch_binding as rt
    big_bytes = max(args.bytes, 1 << 30) if not args.no_sweep else args.bytes
    init_args = ["rabit_device=on", "rabit_device_algo=" + args.algo,
                 "rabit_device_arena=%d" % (2 * (args.bytes + 4096) + (1 << 20)),
                 "rabit_device_heap=%d" % (args.bytes + big_bytes + (4 << 20)),
                 "rabit_device_host_min=8M"]     # host buffers >= 8 MiB ride PCIe + NVLink in 16 MiB chunks
    if world > 1:
        init_args += ["rabit_tracker_uri=127.0.0.1", "rabit_tracker_port=%d" % port, "rabit_task_id=%d" % rank]
    else:
        init_args += ["rabit_device_force=1"]
    rabit.init(init_args)
    assert rt.device_active(), "device backend did not come up: " + rt.describe()
    n = rabit.get_world_size()
    me = rabit.get_rank()
    S = args.bytes
    count = S // 4
    # headline buffer: a symmetric (zero-copy) tensor, the library's documented fast path; the staged
    # path for arbitrary device pointers is measured as well and reported under "staged"
    x = rt.symmetric_empty(count, torch.float32) if n > 1 else torch.empty(count, dtype=torch.float32, device=dev)
    x_plain = torch.empty(count, dtype=torch.float32, device=dev)

    def fill():
        # small integers: the float sum is exact, so the check below is bit-exact
        x.copy_((torch.arange(count, device=dev, dtype=torch.int32) % 131).float() * (me + 1))

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()          # streams every 100 ms from here until the e2e section is over
    # correctness of the timed path, once
    fill()
    rt.allreduce_async_(x)
    rt.synchronize()
    want = (torch.arange(count, device=dev, dtype=torch.int32) % 131).float() * float(n * (n + 1) // 2)
    assert torch.equal(x, want), "allreduce result is wrong"
    del want

    def barrier():
        rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        rt.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        rt.synchronize()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        barrier()
        return max_over_ranks(rabit, np, ms)

    fill()
    launches0 = rt.launch_count()
    ms = timed(lambda: rt.allreduce_async_(x), args.steps, args.warmup)
    launches = rt.launch_count() - launches0
    # launches counted above include the warm-up; report the timed share
    launches = launches * args.steps // (args.steps + args.warmup)
    x_plain.copy_(x)
    staged_ms = timed(lambda: rt.allreduce_async_(x_plain), args.steps, args.warmup) if n > 1 else ms

    # ---- end to end through the public blocking API, host buffer in, host buffer out ----
    e2e_steps = max(3, min(args.steps, 10))
    h_in = torch.empty(count, dtype=torch.float32).pin_memory()
    h_in.copy_((torch.arange(count, dtype=torch.int32) % 131).float() * (me + 1))
    h_out = torch.empty(count, dtype=torch.float32).pin_memory()
    d = x  # the H2D copy lands directly in the symmetric buffer

    def e2e_step():
        d.copy_(h_in, non_blocking=True)
        rabit.allreduce(d, rabit.SUM)            # public API: robust engine, fused consensus, cached result
        h_out.copy_(d, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(2):
        e2e_step()
    rabit.checkpoint(0)                          # drop the result cache built so far
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks(rabit, np, (time.perf_counter() - t0) * 1e3 / e2e_steps)
    rabit.checkpoint(0)
    assert float(h_out[5]) == 5.0 * n * (n + 1) // 2
    e2e_manual_ms = e2e_ms
    e2e_path = "pinned host -> H2D -> rabit.allreduce (robust engine) -> D2H of the full result"
    # the same step as ONE call on the pinned host tensor: the engine chunks it through
    # H2D | fused kernel | D2H itself (RobustEngine::HostViaDevice), copies overlapped
    try:
        h_out.copy_(h_in)
        rt.allreduce_(h_out)
        assert float(h_out[5]) == 5.0 * n * (n + 1) // 2 and float(h_out[count - 1]) == float((count - 1) % 131) * n * (n + 1) // 2
        for _ in range(e2e_steps):               # warm-up: also fills the engine's block pools, which the
            rt.allreduce_(h_out)                 # checkpoint below hands back for reuse
        rabit.checkpoint(0)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            rt.allreduce_(h_out)                 # input: the pinned host tensor; output: the same tensor
        host_ms = max_over_ranks(rabit, np, (time.perf_counter() - t0) * 1e3 / e2e_steps)
        rabit.checkpoint(0)
        if host_ms < e2e_ms:
            e2e_ms = host_ms
            e2e_path = ("torch_binding.allreduce_(pinned host tensor): robust engine cuts it into 16 MiB chunks, "
                        "H2D | fused fault-tolerant kernel | D2H overlapped, result back in the same host tensor")
    except Exception as exc:                     # keep the manual number if the host path is unavailable
        e2e_path += " (host-buffer path failed: %s)" % exc
    clocks = sampler.stop() if rank == 0 else {}

    # ---- NCCL comparator + message-size sweep ----
    sweep = []
    nccl_ms = None
    if not args.no_sweep:
        import torch.distributed as dist
        have_nccl = False
        if world > 1:
            try:
                dist.init_process_group("nccl", store=store, rank=rank, world_size=world)
                have_nccl = True
            except Exception as e:  # pylint: disable=broad-except
                if rank == 0:
                    print("NCCL comparator unavailable: %r" % e, file=sys.stderr)
        if have_nccl:
            nccl_ms = timed_nccl(dist, torch, x, args.steps, args.warmup, rabit, np)
        sizes = [1 << 10, 16 << 10, 256 << 10, 4 << 20, 64 << 20, S]
        if S < (1 << 30):
            sizes.append(1 << 30)
        big = torch.empty((max(sizes) // 4,), dtype=torch.float32, device=dev)
        big_sym = rt.symmetric_empty(max(sizes) // 4, torch.float32) if n > 1 else big
        for sz in sizes:
            v = big[:sz // 4]
            vs = big_sym[:sz // 4]
            it = 20 if sz <= (4 << 20) else 5
            row = {"bytes": sz}
            t = timed(lambda: rt.allreduce_async_(v), it, 3)
            row["allreduce_us"] = t * 1e3
            row["allreduce_busbw_GBps"] = sz / (t * 1e-3) / 1e9 * (2.0 * (n - 1) / n)
            if n > 1:
                t = timed(lambda: rt.allreduce_async_(vs), it, 3)
                row["allreduce_sym_us"] = t * 1e3
                row["allreduce_sym_busbw_GBps"] = sz / (t * 1e-3) / 1e9 * (2.0 * (n - 1) / n)
            if sz <= (4 << 20):
                # launch-bound sizes: the same op replayed from a CUDA graph of 20 launches (no host work per op)
                side = torch.cuda.Stream()
                with torch.cuda.stream(side):
                    rt.allreduce_async_(vs)
                rt.synchronize()
                torch.cuda.synchronize()
                cg = torch.cuda.CUDAGraph()
                with torch.cuda.graph(cg, stream=side):
                    for _ in range(20):
                        rt.allreduce_async_(vs)
                t = timed(cg.replay, 5, 2) / 20.0
                row["allreduce_graph_us"] = t * 1e3
                del cg
            t = timed(lambda: rt.broadcast_async_(v, 0), it, 3)
            row["broadcast_us"] = t * 1e3
            row["broadcast_busbw_GBps"] = sz / (t * 1e-3) / 1e9
            if n > 1:
                t = timed(lambda: rt.broadcast_async_(vs, 0), it, 3)
                row["broadcast_sym_us"] = t * 1e3
                row["broadcast_sym_busbw_GBps"] = sz / (t * 1e-3) / 1e9
            if have_nccl:
                t = timed_nccl(dist, torch, v, it, 3, rabit, np)
                row["nccl_allreduce_us"] = t * 1e3
                row["nccl_allreduce_busbw_GBps"] = sz / (t * 1e-3) / 1e9 * (2.0 * (n - 1) / n)
                t = timed_nccl(dist, torch, v, it, 3, rabit, np, bcast=True)
                row["nccl_broadcast_us"] = t * 1e3
                row["nccl_broadcast_busbw_GBps"] = sz / (t * 1e-3) / 1e9
            sweep.append(row)
        # BASELINE config 3: XGBoost-style gradient histogram, 1000 features x 256 bins x {double grad, double hess}
        # = 4.096 MB, reduced with the pair-sum reducer = float64 Sum over 2x the elements
        hcount = 1000 * 256 * 2
        hs = rt.symmetric_empty(hcount, torch.float64) if n > 1 else torch.zeros(hcount, dtype=torch.float64, device=dev)
        hp = torch.zeros(hcount, dtype=torch.float64, device=dev)
        hist = {"bytes": hcount * 8}
        hist["sym_us"] = timed(lambda: rt.allreduce_async_(hs), 50, 5) * 1e3
        hist["staged_us"] = timed(lambda: rt.allreduce_async_(hp), 50, 5) * 1e3
        if have_nccl:
            hist["nccl_us"] = timed_nccl(dist, torch, hp, 50, 5, rabit, np) * 1e3
        for _ in range(3):
            rabit.allreduce(hp, rabit.SUM)          # robust public API (consensus + cache copy fused)
        rabit.checkpoint(0)
        barrier()
        t0 = time.perf_counter()
        for _ in range(50):
            rabit.allreduce(hp, rabit.SUM)
        torch.cuda.synchronize()
        hist["robust_api_us"] = max_over_ranks(rabit, np, (time.perf_counter() - t0) * 1e6 / 50)
        rabit.checkpoint(0)
        if have_nccl:
            dist.destroy_process_group()

    describe = rt.describe()
    rabit.finalize()
    if rank == 0:
        agg = n * S / (ms * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": agg, "unit": "GB/s", "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "impl": "ours",
            "algbw_GBps": S / (ms * 1e-3) / 1e9,
            "busbw_GBps": S / (ms * 1e-3) / 1e9 * (2.0 * (n - 1) / n),
            "nvlink_roofline_frac_of_770": (S / (ms * 1e-3) / 1e9 * (2.0 * (n - 1) / n)) / 770.0 if n > 1 else None,
            "config": {"model": "speed_test allreduce Sum float32 (reference test/speed_test.cc)",
                       "global_batch": n * S, "seq_len": S, "bytes_per_rank": S, "parallelism": "dp%d" % n,
                       "l2": "operands (256 MiB) larger than the 126 MB L2",
                       "timing": "CUDA events, barrier+sync both sides, max over ranks",
                       "algo": args.algo, "backend": describe,
                       "buffer": "symmetric heap tensor (RabitDeviceSymAlloc / torch_binding.symmetric_empty), zero-copy; "
                                 "'staged' is the same step on an ordinary torch tensor"},
            "staged": {"ms_per_step": staged_ms, "busbw_GBps": S / (staged_ms * 1e-3) / 1e9 * (2.0 * (n - 1) / n),
                       "value": n * S / (staged_ms * 1e-3) / 1e9},
            "e2e": {"value": n * S / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e2e_ms,
                    "steps": e2e_steps, "h2d_bytes_per_step": S, "d2h_bytes_per_step": S,
                    "path": e2e_path, "manual_copy_ms_per_step": e2e_manual_ms},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        if n == 1:
            out["config"]["single_rank_note"] = (
                "one rank: an allreduce is the identity and the reference returns without moving a byte; the timed kernel "
                "still runs its publish / reduce / write-back passes over local HBM (no NVLink traffic), so this is an "
                "HBM-bound figure of the kernel, not a communication figure; busbw is the metric to compare across N")
        if nccl_ms is not None:
            out["nccl"] = {"ms_per_step": nccl_ms, "busbw_GBps": S / (nccl_ms * 1e-3) / 1e9 * (2.0 * (n - 1) / n)}
        if sweep:
            out["sweep"] = sweep
            out["histogram_pair_sum_f64_4MB"] = hist
        print(json.dumps(out), flush=True)
    if tracker is not None:
        tracker.join(10)
        tracker.close()


def timed_nccl(dist, torch, t, steps, warmup, rabit, np, bcast=False):
    fn = (lambda: dist.broadcast(t, 0)) if bcast else (lambda: dist.all_reduce(t))
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return max_over_ranks(rabit, np, e0.elapsed_time(e1) / steps)


# ---------------------------------------------------------------------- reference
def run_reference(args):
    """The UNMODIFIED reference (sources of /root/reference copied to baseline/_ref by
    baseline/build_ref.py, compiled against a minimal dmlc header shim) running the same step:
    rabit::Allreduce<op::Sum>(float*, S/4) on host memory through its own socket engine."""
    rank, world, _ = dist_env()
    exe = os.path.join(ROOT, "baseline", "_ref", "bin", "ref_speed")
    if not os.path.isfile(exe):
        try:
            sys.path.insert(0, os.path.join(ROOT, "baseline"))
            import build_ref
            build_ref.build()
        except Exception as e:  # pylint: disable=broad-except
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "reference build failed: %s" % str(e)[:200]}))
            return
    if args.gpus == 1 and world == 1:
        # With one worker the reference's Allreduce returns before touching the buffer. For orientation the
        # same step is also run as TWO reference workers on this host's CPUs (its smallest real job).
        info = None
        try:
            from rabit_b200.parallel.launcher import launch
            res = launch([exe, str(args.bytes // 4), "2", "1", "rabit_enable_tcp_no_delay=true"], 2, capture=True,
                         timeout=600)
            line = [l for o in res.outputs.values() for l in o.splitlines() if l.startswith("REF_RESULT")]
            if res.ok and line:
                ms = float(line[0].split()[1])
                info = {"workers": 2, "ms_per_step": ms, "aggregate_GBps": 2 * args.bytes / (ms * 1e-3) / 1e9}
        except Exception:  # pylint: disable=broad-except
            info = None
        print(json.dumps({"impl": "reference", "unavailable":
                          "world_size=1: rabit::Allreduce returns immediately (allreduce_robust.cc:253-256), "
                          "there is no transfer to time", "two_host_workers_for_orientation": info}))
        return
    store = make_store(rank, world)
    tracker, port = start_tracker(rank, world, store)
    env = dict(os.environ)
    env.update({"DMLC_TRACKER_URI": "127.0.0.1", "DMLC_TRACKER_PORT": str(port), "DMLC_TASK_ID": str(rank),
                "DMLC_ROLE": "worker", "DMLC_NUM_ATTEMPT": "0"})
    steps = max(1, min(args.steps, 5))   # ~1 s per 256 MiB op on loopback TCP
    warm = max(1, min(args.warmup, 2))
    r = subprocess.run([exe, str(args.bytes // 4), str(steps), str(warm), "rabit_enable_tcp_no_delay=true"],
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=3000)
    if tracker is not None:
        tracker.join(20)
        tracker.close()
    # rabit rank 0 (the one that prints) is whatever process the tracker labelled 0: route its line to
    # torchrun rank 0 through the store
    mine = [l for l in r.stdout.decode().splitlines() if l.startswith("REF_RESULT")]
    if mine:
        store.set("ref_result", mine[0])
    store.set("ref_rc_%d" % rank, str(r.returncode))
    if rank == 0:
        line = []
        if all(store.get("ref_rc_%d" % q).decode() == "0" for q in range(world)):
            line = [store.get("ref_result").decode()]
        if r.returncode != 0 or not line:
            print(json.dumps({"impl": "reference", "unavailable": "reference run failed rc=%d: %s" %
                              (r.returncode, r.stderr.decode()[-200:])}))
            return
        ms = float(line[0].split()[1])
        n, S = world, args.bytes
        print(json.dumps({
            "metric": METRIC, "value": n * S / (ms * 1e-3) / 1e9, "unit": "GB/s", "n_gpus": n, "steps": steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic", "impl": "reference",
            "busbw_GBps": S / (ms * 1e-3) / 1e9 * (2.0 * (n - 1) / n),
            "config": {"model": "speed_test allreduce Sum float32 (reference test/speed_test.cc)",
                       "bytes_per_rank": S, "parallelism": "dp%d" % n,
                       "engine": "reference AllreduceRobust over loopback TCP, host memory, rabit_enable_tcp_no_delay=true",
                       "timing": "host wall clock inside the reference program (it has no device path), max over ranks"},
            "e2e": {"value": n * S / (ms * 1e-3) / 1e9, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "path": "host buffer in, host buffer out (the reference's only mode)"},
            "gpu_launches": 0}), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
