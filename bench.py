#!/usr/bin/env python
"""Headline benchmark: speed_test-style allreduce (float32, Sum) through rabit-b200.

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference --gpus N ...      (the unmodified reference, host TCP engine)

Metric (BASELINE.json: "allreduce/broadcast bus GB/s vs msg size ...; recover-after-kill secs"; throughput
definition from the reference's test/speed_test.cc:52-69, bytes / mean time):
  value = N * S / t   aggregate bytes allreduced per second over all N GPUs, S = 256 MiB of float32 per
          rank (fixed as N grows: weak scaling).  Per-rank algorithm bandwidth is S / t, bus bandwidth
          (S / t) * 2 (N-1) / N; both are in the JSON next to their fraction of 900 GB/s, and `sweep`
          has the curve from 1 KiB to 1 GiB for allreduce, broadcast and allgather against NCCL.
  A "step" is ONE CALL OF THE PUBLIC, BLOCKING, FAULT-TOLERANT API, rabit.allreduce(tensor, SUM), on an
  ordinary torch CUDA tensor: consensus fused into the kernels, result cached for recovery, the caller's
  buffer intact if a peer dies.  Timed on the device with CUDA events after a barrier + synchronize, max
  over ranks.  The 256 MiB operands exceed the 126 MB L2.  The stream-ordered, non-recoverable calls
  (allreduce_async_ on plain and symmetric tensors) are reported as extra keys.
  e2e: the same public call on a PINNED HOST tensor: the engine moves it H2D chunk by chunk, reduces on the
  GPUs and moves the result D2H (one sequence-numbered, recoverable op); wall clock, every step.
  recovery: seconds per scripted kill of the reference's model_recover scenario on device buffers.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HEADLINE_BYTES = 256 << 20
METRIC = "allreduce_f32_sum_aggregate_GBps"
NVLINK_NOMINAL = 900.0      # GB/s per direction per GPU (BASELINE.json)
NVLINK_MEASURED = 770.0     # peer copy on this pool (B200_PROFILING.md)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bytes", type=int, default=HEADLINE_BYTES)
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-recovery", action="store_true")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--extra", default="", help="extra name=value engine parameters, comma separated")
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    return rank, world, local


class ClockSampler(object):
    """nvidia-smi clocks/throttle reasons sampled every 100 ms while the timed region runs."""
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


def nvlink_bytes(gpu):
    """Cumulative NVLink (rx, tx) KiB of one GPU, summed over its links: NVML field values first, the nvidia-smi
    table as a fallback; None if the counters are not exposed."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu)
        vals = pynvml.nvmlDeviceGetFieldValues(h, [(pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF),
                                                   (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF)])
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                raise RuntimeError("field not available")
            out.append(int(v.value.ullVal))
        return tuple(out)
    except Exception:  # pylint: disable=broad-except
        pass
    try:
        r = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(gpu)], stdout=subprocess.PIPE,
                           stderr=subprocess.DEVNULL, timeout=20)
        rx = tx = 0
        seen = False
        for line in r.stdout.decode().splitlines():
            line = line.strip()
            if "Data Tx:" in line:
                tx += int(line.split("Data Tx:")[1].split()[0]); seen = True
            elif "Data Rx:" in line:
                rx += int(line.split("Data Rx:")[1].split()[0]); seen = True
        return (rx, tx) if seen else None
    except Exception:  # pylint: disable=broad-except
        return None


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
                         world, rank == 0, timeout=datetime.timedelta(seconds=600), wait_for_workers=True)


def start_tracker(rank, world, store, tag="rabit_tracker_port"):
    """rank 0 hosts the rabit tracker on a free port and publishes it through the store"""
    if world == 1:
        return None, None
    if rank == 0:
        from rabit_b200.parallel.tracker import Tracker
        tr = Tracker(world, host="127.0.0.1").start()
        store.set(tag, str(tr.port))
        return tr, tr.port
    return None, int(store.get(tag).decode())


def max_over_ranks(rabit, np, value):
    return float(rabit.allreduce(np.array([value], dtype=np.float64), rabit.MAX)[0])


def recovery_seconds(exe_args, world, per_rank_gpu, repeat=1, timeout=600, standby=False):
    """(wall(run with 2 scripted kills) - wall(same run, no kills)) / 2 under OUR launcher (SURVEY section 6;
    scenario test/test.mk:15-16 of the reference). Returns a dict, or {"error": ...}."""
    from rabit_b200.parallel.launcher import launch
    common = ["rabit_bootstrap_cache=1", "rabit_enable_tcp_no_delay=true"]
    kills = [k for k in ("mock=0,0,1,0", "mock=1,1,1,0") if int(k.split("=")[1].split(",")[0]) < world]
    out = {"world": world, "scenario": "model_recover, kills " + " ".join(kills), "standby_workers": standby}
    try:
        walls = {}
        for name, extra in (("no_kill", []), ("kills", kills)):
            best = None
            for _ in range(repeat):
                res = launch(exe_args + common + extra, world, capture=True, timeout=timeout, per_rank_gpu=per_rank_gpu,
                             standby=standby)
                if not res.ok:
                    return {"error": "job failed (%s): %s" % (name, str(res.returncodes)[:200])}
                best = res.wall if best is None else min(best, res.wall)
                out[name + "_restarts"] = res.restarts
            walls[name] = best
        out["wall_no_kill_s"] = walls["no_kill"]
        out["wall_kills_s"] = walls["kills"]
        out["recover_s_per_kill"] = (walls["kills"] - walls["no_kill"]) / max(len(kills), 1)
    except Exception as e:  # pylint: disable=broad-except
        return {"error": str(e)[:300]}
    return out


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
    S = args.bytes
    big_bytes = max(S, 1 << 30) if not args.no_sweep else S
    init_args = ["rabit_device=on", "rabit_device_algo=" + args.algo,
                 "rabit_device_arena=%d" % (2 * (big_bytes + 4096) + (1 << 20)),
                 "rabit_device_heap=%d" % (S + big_bytes + (768 << 20)),
                 "rabit_device_cache=64G"]
    init_args += [a for a in args.extra.split(",") if a]
    if world > 1:
        init_args += ["rabit_tracker_uri=127.0.0.1", "rabit_tracker_port=%d" % port, "rabit_task_id=%d" % rank]
    else:
        init_args += ["rabit_device_force=1"]
    rabit.init(init_args)
    assert rt.device_active(), "device backend did not come up: " + rt.describe()
    n = rabit.get_world_size()
    me = rabit.get_rank()
    count = S // 4
    tri = float(n * (n + 1) // 2)

    def pattern(cnt, scale):
        # small integers: float sums are exact, so every check below is bit-exact
        return (torch.arange(cnt, device=dev, dtype=torch.int32) % 131).float() * scale

    x = torch.empty(count, dtype=torch.float32, device=dev)                 # ordinary torch tensor (cudaMalloc)
    x_sym = rt.symmetric_empty(count, torch.float32) if n > 1 else x       # zero-copy heap tensor

    def barrier():
        rabit.allreduce(np.zeros(1, dtype=np.float32), rabit.SUM)
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, sync=None):
        for _ in range(warmup):
            fn()
        (sync or torch.cuda.synchronize)()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        (sync or torch.cuda.synchronize)()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        barrier()
        return max_over_ranks(rabit, np, ms)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()          # streams every 100 ms from here until the e2e section is over

    # ---- headline: the public blocking fault-tolerant call on an ordinary tensor ----
    x.copy_(pattern(count, me + 1))
    rabit.allreduce(x, rabit.SUM)
    assert torch.equal(x, pattern(count, tri)), "robust allreduce result is wrong"
    # fill the engine's cache-block pool (one block per op until the next checkpoint), then hand it back:
    # the timed steps below must not contain a cudaMalloc
    for _ in range(args.steps + args.warmup):
        rabit.allreduce(x, rabit.SUM)
    rabit.checkpoint(0)
    x.copy_(pattern(count, me + 1))
    nv0 = nvlink_bytes(local) if rank == 0 else None
    launches0 = rt.launch_count()
    ms = timed(lambda: rabit.allreduce(x, rabit.SUM), args.steps, args.warmup)
    launches = (rt.launch_count() - launches0) * args.steps // (args.steps + args.warmup)
    nv1 = nvlink_bytes(local) if rank == 0 else None
    rabit.checkpoint(0)
    # the same call on a symmetric-heap tensor (robust, in place behind a backup)
    robust_sym_ms = None
    if n > 1:
        x_sym.copy_(pattern(count, me + 1))
        rabit.allreduce(x_sym, rabit.SUM)
        assert torch.equal(x_sym, pattern(count, tri)), "robust allreduce on a symmetric tensor is wrong"
        robust_sym_ms = timed(lambda: rabit.allreduce(x_sym, rabit.SUM), args.steps, args.warmup)
        rabit.checkpoint(0)
    # stream-ordered, non-recoverable calls
    x_sym.copy_(pattern(count, me + 1))
    rt.allreduce_async_(x_sym)
    rt.synchronize()
    assert torch.equal(x_sym, pattern(count, tri)), "async allreduce result is wrong"
    async_sym_ms = timed(lambda: rt.allreduce_async_(x_sym), args.steps, args.warmup, rt.synchronize)
    async_plain_ms = timed(lambda: rt.allreduce_async_(x), args.steps, args.warmup, rt.synchronize) if n > 1 else async_sym_ms

    # ---- copy ceiling: what PCIe and host memory allow when every rank moves S in and S out at once ----
    h_in = torch.empty(count, dtype=torch.float32).pin_memory()
    h_out = torch.empty(count, dtype=torch.float32).pin_memory()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def copy_both():
        with torch.cuda.stream(s_in):
            x.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s_out):
            h_out.copy_(x_sym, non_blocking=True)
        s_in.synchronize()
        s_out.synchronize()
    ceil_steps = 5
    for _ in range(2):
        copy_both()
    barrier()
    t0 = time.perf_counter()
    for _ in range(ceil_steps):
        copy_both()
    copy_ceiling_ms = max_over_ranks(rabit, np, (time.perf_counter() - t0) * 1e3 / ceil_steps)

    # ---- end to end through the public blocking API, pinned host buffer in, host buffer out ----
    e2e_steps = max(3, min(args.steps, 10))
    base = ((torch.arange(count, dtype=torch.int32) % 131).float() * (me + 1))
    h_in.copy_(base)
    e2e = {}
    # (a) ONE call on the pinned host tensor: the engine runs H2D | kernels | D2H itself
    h_in.copy_(base)
    rt.allreduce_(h_in)
    ok_host = bool(float(h_in[5]) == 5.0 * tri and float(h_in[count - 1]) == float((count - 1) % 131) * tri)
    host_ms = None
    if ok_host:
        for _ in range(e2e_steps + 2):
            rt.allreduce_(h_in)
        rabit.checkpoint(0)
        h_in.copy_(base / 1e6)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            rt.allreduce_(h_in)                  # input: the pinned host tensor; output: the same tensor
        host_ms = max_over_ranks(rabit, np, (time.perf_counter() - t0) * 1e3 / e2e_steps)
        rabit.checkpoint(0)
    # (b) the copies made by the caller around the device call
    h_in.copy_(base)

    def manual_step():
        x.copy_(h_in, non_blocking=True)
        rabit.allreduce(x, rabit.SUM)
        h_out.copy_(x, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    for _ in range(2):
        manual_step()
    assert float(h_out[5]) == 5.0 * tri
    rabit.checkpoint(0)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        manual_step()
    manual_ms = max_over_ranks(rabit, np, (time.perf_counter() - t0) * 1e3 / e2e_steps)
    rabit.checkpoint(0)
    if host_ms is not None and host_ms <= manual_ms:
        e2e_ms = host_ms
        e2e_path = ("rabit.allreduce(pinned host tensor): one recoverable op; the engine streams it H2D -> stage | "
                    "reduce kernels | result -> D2H in chunks, result back in the same host tensor")
    else:
        e2e_ms = manual_ms
        e2e_path = "pinned host -> H2D -> rabit.allreduce(device tensor) -> D2H of the full result"
        if not ok_host:
            e2e_path += " (host-buffer path gave a wrong result and is not reported)"
    e2e = {"value": n * S / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e2e_ms, "steps": e2e_steps,
           "h2d_bytes_per_step": S, "d2h_bytes_per_step": S, "path": e2e_path,
           "host_tensor_call_ms_per_step": host_ms, "manual_copy_ms_per_step": manual_ms,
           "copy_ceiling_ms": copy_ceiling_ms,
           "copy_ceiling_note": "every rank copies S host->device and S device->host at once on two streams, no "
                                "collective: what PCIe and the host memory allow at this N",
           "frac_of_copy_ceiling": copy_ceiling_ms / e2e_ms}
    clocks = sampler.stop() if rank == 0 else {}

    # ---- bit-exact self-check of every algorithm on real peers (the driver's multi-GPU correctness run) ----
    selfcheck = run_selfcheck(rabit, rt, torch, np, dev, me, n)

    # ---- NCCL comparator + message-size sweep ----
    sweep = []
    nccl_ms = None
    hist = None
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
            bus = 2.0 * (n - 1) / n

            def put(key, t, factor=bus):
                row[key + "_us"] = t * 1e3
                row[key + "_busbw_GBps"] = sz / (t * 1e-3) / 1e9 * factor
            # the product path first: blocking fault-tolerant API on an ordinary tensor
            for _ in range(it + 3):
                rabit.allreduce(v, rabit.SUM)
            rabit.checkpoint(0)
            put("robust", timed(lambda: rabit.allreduce(v, rabit.SUM), it, 3))
            rabit.checkpoint(0)
            put("allreduce", timed(lambda: rt.allreduce_async_(v), it, 3, rt.synchronize))
            if n > 1:
                put("allreduce_sym", timed(lambda: rt.allreduce_async_(vs), it, 3, rt.synchronize))
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
                row["allreduce_graph_us"] = timed(cg.replay, 5, 2) / 20.0 * 1e3
                del cg
            put("broadcast", timed(lambda: rt.broadcast_async_(v, 0), it, 3, rt.synchronize), 1.0)
            if n > 1:
                put("broadcast_sym", timed(lambda: rt.broadcast_async_(vs, 0), it, 3, rt.synchronize), 1.0)
                # allgather of sz bytes in total (every rank owns sz / n)
                per = sz // 4 // n
                if per > 0:
                    put("allgather", timed(lambda: rt.allgather_async_(v[:per * n], me * per, per), it, 3, rt.synchronize),
                        (n - 1.0) / n)
                    put("allgather_sym", timed(lambda: rt.allgather_async_(vs[:per * n], me * per, per), it, 3, rt.synchronize),
                        (n - 1.0) / n)
            if have_nccl:
                put("nccl_allreduce", timed_nccl(dist, torch, v, it, 3, rabit, np))
                put("nccl_broadcast", timed_nccl(dist, torch, v, it, 3, rabit, np, kind="bcast"), 1.0)
                per = sz // 4 // n
                if per > 0:
                    put("nccl_allgather", timed_nccl(dist, torch, v[:per * n], it, 3, rabit, np, kind="gather", rank=me, n=n),
                        (n - 1.0) / n)
            sweep.append(row)
        # BASELINE config 3: XGBoost-style gradient histogram, 1000 features x 256 bins x {double grad, double hess}
        # = 4.096 MB, reduced with the pair-sum reducer = float64 Sum over 2x the elements
        hcount = 1000 * 256 * 2
        hs = rt.symmetric_empty(hcount, torch.float64) if n > 1 else torch.zeros(hcount, dtype=torch.float64, device=dev)
        hp = torch.zeros(hcount, dtype=torch.float64, device=dev)
        hist = {"bytes": hcount * 8}
        hist["sym_us"] = timed(lambda: rt.allreduce_async_(hs), 50, 5, rt.synchronize) * 1e3
        hist["staged_us"] = timed(lambda: rt.allreduce_async_(hp), 50, 5, rt.synchronize) * 1e3
        if have_nccl:
            hist["nccl_us"] = timed_nccl(dist, torch, hp, 50, 5, rabit, np) * 1e3
        for _ in range(55):
            rabit.allreduce(hp, rabit.SUM)          # robust public API (consensus + cache copy fused)
        rabit.checkpoint(0)
        hist["robust_api_us"] = timed(lambda: rabit.allreduce(hp, rabit.SUM), 50, 5) * 1e3
        rabit.checkpoint(0)
        if have_nccl:
            dist.destroy_process_group()

    describe = rt.describe()
    rabit.finalize()
    torch.cuda.synchronize()
    # ---- recover-after-kill on device buffers (native workers under our launcher), rank 0 drives it ----
    recovery = None
    if not args.no_recovery and n > 1:
        if store is not None:
            store.set("bench_done_%d" % rank, "1")
        if rank == 0:
            for q in range(world):
                store.get("bench_done_%d" % q)
            exe = os.path.join(ROOT, "build", "bin", "device_recover_cuda")
            if os.path.isfile(exe):
                # the launcher's standby workers (a pre-started process per task, CUDA context ready) are part of
                # this framework; the same scenario without them is reported next to it
                recovery = recovery_seconds([exe, "200000"], n, per_rank_gpu=True, standby=True)
                cold = recovery_seconds([exe, "200000"], n, per_rank_gpu=True, standby=False)
                recovery["without_standby_workers"] = cold
            else:
                recovery = {"error": "build/bin/device_recover_cuda missing"}
            store.set("bench_recovery_done", "1")
        else:
            store.get("bench_recovery_done")
    same_prog = None
    if not args.no_recovery and n > 1 and store is not None:
        try:
            same_prog = same_program_on_this_library(rank, world, local, store, S, max(1, min(args.steps, 5)))
        except Exception as e:  # pylint: disable=broad-except
            same_prog = {"error": str(e)[:200]}
    if rank == 0:
        bus = 2.0 * (n - 1) / n

        def rates(t_ms):
            alg = S / (t_ms * 1e-3) / 1e9
            return {"ms_per_step": t_ms, "value": n * alg, "algbw_GBps": alg, "busbw_GBps": alg * bus,
                    "busbw_frac_of_900": alg * bus / NVLINK_NOMINAL if n > 1 else None,
                    "busbw_frac_of_770_measured": alg * bus / NVLINK_MEASURED if n > 1 else None}
        head = rates(ms)
        out = {
            "metric": METRIC, "value": head["value"], "unit": "GB/s", "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "impl": "ours",
            "algbw_GBps": head["algbw_GBps"], "busbw_GBps": head["busbw_GBps"],
            "nvlink_roofline_frac": head["busbw_frac_of_900"],
            "busbw_frac_of_770_measured": head["busbw_frac_of_770_measured"],
            "scaling_note": "the N=1 value is an HBM figure (no NVLink traffic); compare busbw_GBps / 900 across N",
            "config": {"model": "speed_test allreduce Sum float32 (reference test/speed_test.cc)",
                       "global_batch": n * S, "seq_len": S, "bytes_per_rank": S, "parallelism": "dp%d" % n,
                       "l2": "operands (256 MiB) larger than the 126 MB L2",
                       "timing": "CUDA events, barrier+sync both sides, max over ranks",
                       "algo": args.algo, "backend": describe,
                       "api": "rabit.allreduce(torch CUDA tensor, SUM): public blocking fault-tolerant call "
                              "(RobustEngine: consensus fused into the kernels, result cached, undo on failure) on an "
                              "ordinary cudaMalloc'ed tensor"},
            "robust_symmetric": rates(robust_sym_ms) if robust_sym_ms else None,
            "async_plain": rates(async_plain_ms), "async_symmetric": rates(async_sym_ms),
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "selfcheck": selfcheck,
        }
        if nv0 is not None and nv1 is not None:
            tot = args.steps + args.warmup + 2       # + the barrier's tiny ops: negligible bytes
            out["nvlink_counters"] = {
                "rx_GB_per_step": (nv1[0] - nv0[0]) * 1024.0 / 1e9 / tot, "tx_GB_per_step": (nv1[1] - nv0[1]) * 1024.0 / 1e9 / tot,
                "algorithmic_GB_per_step_per_direction": S / 1e9 * (n - 1.0) / n * (1.0 if "NVLS) on" in describe and n >= 3 else 2.0),
                "source": "nvidia-smi nvlink -gt d, rank 0's GPU, around warm-up + timed steps"}
        if n == 1:
            out["config"]["single_rank_note"] = (
                "one rank: an allreduce is the identity and the reference returns without moving a byte; the timed op "
                "still runs its copy-in / reduce / write-back passes over local HBM (no NVLink traffic), so this is an "
                "HBM-bound figure, not a communication figure")
        if nccl_ms is not None:
            out["nccl"] = rates(nccl_ms)
        if sweep:
            out["sweep"] = sweep
            out["histogram_pair_sum_f64_4MB"] = hist
        if same_prog is not None:
            out["same_program"] = same_prog
        if recovery is not None:
            out["recovery"] = recovery
            if "recover_s_per_kill" in recovery:
                out["recover_s_per_kill"] = recovery["recover_s_per_kill"]
        print(json.dumps(out), flush=True)
    if tracker is not None:
        tracker.join(10)
        tracker.close()


def run_selfcheck(rabit, rt, torch, np, dev, me, n):
    """Every allreduce algorithm x {f32 Sum, f64 Sum, i32 Max, i64 BitOR} x {async plain, async symmetric, robust
    plain, robust symmetric, host buffer} x three sizes, plus broadcast and allgather, compared bit for bit with
    the closed form.  Returns {"checked": k, "failed": [...]}; never raises (the timed numbers stay reportable)."""
    failed, checked = [], 0
    sizes = [1024, 1 << 20, 40 << 20]
    algos = ["auto", "oneshot", "twoshot", "tma", "nvls", "ll"] if n > 1 else ["auto"]
    sym_cache = {}

    def sym(cnt, dt):
        key = (cnt, dt)
        if key not in sym_cache:
            sym_cache[key] = rt.symmetric_empty(cnt, dt) if n > 1 else torch.empty(cnt, dtype=dt, device=dev)
        return sym_cache[key]

    def contrib(cnt, dt, r, op):
        i = torch.arange(cnt, device=dev, dtype=torch.int64)
        if op == rabit.BITOR:
            return (((i * 7 + 3) % 251) | (1 << (r + 8))).to(dt)
        return ((i * (r + 3) + r) % 113).to(dt)

    def expect(cnt, dt, op):
        rows = torch.stack([contrib(cnt, dt, r, op) for r in range(n)])
        if op == rabit.SUM:
            return rows.sum(0).to(dt)
        if op == rabit.MAX:
            return rows.max(0).values
        out = rows[0].clone()
        for r in range(1, n):
            out |= rows[r]
        return out
    combos = [(torch.float32, rabit.SUM), (torch.float64, rabit.SUM), (torch.int32, rabit.MAX), (torch.int64, rabit.BITOR)]
    try:
        for algo in algos:
            rt.set_tuning(algo=algo)
            for dt, op in combos:
                for sz in sizes:
                    if algo == "ll" and sz > 32768:
                        continue
                    if algo != "auto" and sz > (1 << 20) and (dt, op) != (torch.float32, rabit.SUM):
                        continue
                    cnt = sz // torch.empty((), dtype=dt).element_size()
                    want = expect(cnt, dt, op)
                    paths = ["async_plain", "robust_plain"]
                    if n > 1:
                        paths += ["async_sym", "robust_sym"]
                    if algo == "auto" and sz >= (1 << 20):
                        paths.append("host")
                    for path in paths:
                        checked += 1
                        mine = contrib(cnt, dt, me, op)
                        if path == "host":
                            t = mine.cpu().pin_memory()
                            rt.allreduce_(t, op)
                            got = t.to(dev)
                        else:
                            t = sym(cnt, dt) if path.endswith("sym") else torch.empty(cnt, dtype=dt, device=dev)
                            t.copy_(mine)
                            if path.startswith("async"):
                                rt.allreduce_async_(t, op)
                                rt.synchronize()
                            else:
                                rabit.allreduce(t, op)
                            got = t
                        if not torch.equal(got, want):
                            failed.append("%s/%s/op%d/%d/%s" % (algo, str(dt).split(".")[1], op, sz, path))
            rabit.checkpoint(0)
        rt.set_tuning(algo="auto")
        # broadcast and allgather, plain and symmetric
        for sz in sizes:
            for root in sorted(set([0, n - 1])):
                truth = ((torch.arange(sz, device=dev) * 13 + root) % 253).to(torch.uint8)
                for which in (["plain", "sym"] if n > 1 else ["plain"]):
                    checked += 1
                    b = sym(sz, torch.uint8) if which == "sym" else torch.empty(sz, dtype=torch.uint8, device=dev)
                    b.zero_()
                    if me == root:
                        b.copy_(truth)
                    rt.broadcast_(b, root)
                    if not torch.equal(b, truth):
                        failed.append("broadcast/%d/root%d/%s" % (sz, root, which))
            per = sz // 4 // n
            if per > 0:
                checked += 1
                g = torch.zeros(per * n, dtype=torch.int32, device=dev)
                g[me * per:(me + 1) * per] = me + 1
                rt.allgather_(g, me * per, per)
                if not torch.equal(g, torch.arange(1, n + 1, device=dev, dtype=torch.int32).repeat_interleave(per)):
                    failed.append("allgather/%d" % sz)
        rabit.checkpoint(0)
    except Exception as e:  # pylint: disable=broad-except
        failed.append("exception: %s" % str(e)[:200])
        try:
            rt.set_tuning(algo="auto")
        except Exception:  # pylint: disable=broad-except
            pass
    return {"checked": checked, "failed": failed, "algorithms": algos, "bit_exact": True}


def timed_nccl(dist, torch, t, steps, warmup, rabit, np, kind="allreduce", rank=0, n=1):
    if kind == "bcast":
        fn = lambda: dist.broadcast(t, 0)  # noqa: E731
    elif kind == "gather":
        per = t.numel() // n
        fn = lambda: dist.all_gather_into_tensor(t, t[rank * per:(rank + 1) * per])  # noqa: E731
    else:
        fn = lambda: dist.all_reduce(t)  # noqa: E731
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
def ref_run(exe, argv, rank, world, store, tag, timeout=3000, env_extra=None):
    """all torchrun ranks start one worker of the program each under a fresh tracker; returns rank 0's REF_RESULT
    lines (on torchrun rank 0) or None. A worker that fails or overruns `timeout` makes the result None, never an
    exception or a wait without end: every rank always posts its return code."""
    tracker, port = start_tracker(rank, world, store, "ref_tracker_" + tag)
    env = dict(os.environ)
    env.update({"DMLC_TRACKER_URI": "127.0.0.1", "DMLC_TRACKER_PORT": str(port), "DMLC_TASK_ID": str(rank),
                "DMLC_ROLE": "worker", "DMLC_NUM_ATTEMPT": "0"})
    env.update(env_extra or {})
    rc, stdout = "failed", ""
    try:
        r = subprocess.run([exe] + argv, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=timeout)
        rc, stdout = str(r.returncode), r.stdout.decode(errors="replace")
        if r.returncode != 0:
            sys.stderr.write("[bench] %s worker %d rc=%d: %s\n" % (tag, rank, r.returncode, r.stderr.decode(errors="replace")[-400:]))
    except Exception as e:  # pylint: disable=broad-except
        sys.stderr.write("[bench] %s worker %d: %s\n" % (tag, rank, str(e)[:200]))
    if tracker is not None:
        tracker.join(20)
        tracker.close()
    # rabit rank 0 (the one that prints) is whatever process the tracker labelled 0: route its lines to
    # torchrun rank 0 through the store
    mine = [l for l in stdout.splitlines() if l.startswith("REF_RESULT")]
    if mine:
        store.set("ref_result_" + tag, "\n".join(mine))
    store.set("ref_rc_%s_%d" % (tag, rank), rc)
    if rank != 0:
        return None
    if not all(store.get("ref_rc_%s_%d" % (tag, q)).decode() == "0" for q in range(world)):
        return None
    return store.get("ref_result_" + tag).decode().splitlines()


def same_program_on_this_library(rank, world, local, store, nbytes, steps):
    """baseline/ref_speed.cc, the harness of the reference arm, compiled UNCHANGED against this library
    (build/bin/ours_speed): host buffers in, host buffers out (pageable std::vector), wall clock inside the program,
    max over ranks: what a user of the reference gets by re-linking. Once with the device backend off (host engine:
    sockets for rendezvous and liveness, payload through shared-memory rings between the workers of one box) and once
    with it on (the host buffers ride the GPUs: H2D | kernels | D2H)."""
    exe = os.path.join(ROOT, "build", "bin", "ours_speed")
    if not os.path.isfile(exe):
        return {"error": "build/bin/ours_speed missing"} if rank == 0 else None
    out = {"program": "baseline/ref_speed.cc (unchanged) linked with lib/librabit_mock.so, rabit_engine=robust",
           "steps": steps, "bytes_per_rank": nbytes}
    for tag, extra in (("host_engine", ["rabit_device=off"]), ("via_gpus", ["rabit_device=on"])):
        lines = ref_run(exe, [str(nbytes // 4), str(steps), "1", "rabit_engine=robust", "rabit_enable_tcp_no_delay=true"] + extra,
                        rank, world, store, "same_" + tag, timeout=240, env_extra={"RABIT_DEVICE_ORDINAL": str(local)})
        if rank == 0:
            if lines:
                ms = float(lines[0].split()[1])
                out[tag] = {"ms_per_step": ms, "value": world * nbytes / (ms * 1e-3) / 1e9, "unit": "GB/s"}
            else:
                out[tag] = {"error": "run failed (see stderr)"}
    return out if rank == 0 else None


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
    # the requested step count is honoured up to a budget of about two minutes of loopback TCP (one 256 MiB op
    # takes 0.4-0.5 s at 8 workers); whatever was used is in the JSON
    steps = max(1, min(args.steps, 100))
    warm = max(1, min(args.warmup, 20))
    n, S = world, args.bytes
    lines = ref_run(exe, [str(S // 4), str(steps), str(warm), "rabit_enable_tcp_no_delay=true"], rank, world, store, "main")
    # same step with the reference's DEFAULT parameters (Nagle on), fewer repetitions
    lines_def = ref_run(exe, [str(S // 4), str(max(1, min(steps, 5))), "1"], rank, world, store, "default")
    # message-size sweep of the reference on this box (BASELINE config 2)
    sweep = []
    for sz, reps in ((1 << 10, 200), (16 << 10, 200), (256 << 10, 100), (4 << 20, 20), (64 << 20, 5), (1 << 30, 2)):
        ls = ref_run(exe, [str(sz // 4), str(reps), "2", "rabit_enable_tcp_no_delay=true"], rank, world, store, "s%d" % sz)
        if rank == 0 and ls:
            t = float(ls[0].split()[1])
            sweep.append({"bytes": sz, "allreduce_us": t * 1e3, "allreduce_busbw_GBps": sz / (t * 1e-3) / 1e9 * 2.0 * (n - 1) / n})
    # recover-after-kill of the reference's own model_recover under the same launcher and kill script
    recovery = None
    if not args.no_recovery:
        store.set("refbench_done_%d" % rank, "1")
        if rank == 0:
            for q in range(world):
                store.get("refbench_done_%d" % q)
            rexe = os.path.join(ROOT, "baseline", "_ref", "bin", "ref_model_recover")
            recovery = recovery_seconds([rexe, "200000"], n, per_rank_gpu=False) if os.path.isfile(rexe) else \
                {"error": "ref_model_recover missing"}
            store.set("refbench_recovery_done", "1")
        else:
            store.get("refbench_recovery_done")
    if rank == 0:
        if not lines:
            print(json.dumps({"impl": "reference", "unavailable": "reference run failed (see stderr of the workers)"}))
            return
        ms = float(lines[0].split()[1])
        out = {
            "metric": METRIC, "value": n * S / (ms * 1e-3) / 1e9, "unit": "GB/s", "n_gpus": n, "steps": steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic", "impl": "reference",
            "busbw_GBps": S / (ms * 1e-3) / 1e9 * (2.0 * (n - 1) / n),
            "config": {"model": "speed_test allreduce Sum float32 (reference test/speed_test.cc)",
                       "bytes_per_rank": S, "parallelism": "dp%d" % n,
                       "engine": "reference AllreduceRobust over loopback TCP, host memory, rabit_enable_tcp_no_delay=true",
                       "steps_note": "requested steps=%d warmup=%d; used %d / %d" % (args.steps, args.warmup, steps, warm),
                       "timing": "host wall clock inside the reference program (it has no device path), max over ranks"},
            "e2e": {"value": n * S / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0, "path": "host buffer in, host buffer out (the reference's only mode)"},
            "gpu_launches": 0, "sweep": sweep}
        if lines_def:
            dms = float(lines_def[0].split()[1])
            out["default_params"] = {"ms_per_step": dms, "value": n * S / (dms * 1e-3) / 1e9,
                                     "note": "reference defaults (TCP_NODELAY off)"}
        if recovery is not None:
            out["recovery"] = recovery
            if "recover_s_per_kill" in recovery:
                out["recover_s_per_kill"] = recovery["recover_s_per_kill"]
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
