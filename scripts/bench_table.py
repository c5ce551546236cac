#!/usr/bin/env python
"""Markdown summary of bench.py JSON lines (ours + reference arm) for profiles/.

    python scripts/bench_table.py profiles/r2_bench2.json profiles/r2_bench4.json profiles/r2_bench8.json \
        [--ref profiles/r2_bench2_ref.json ...] > profiles/r2_bench_vs_nccl.md
"""
import json
import sys


def load(path):
    rows = [json.loads(l) for l in open(path) if l.startswith("{")]
    return rows[-1] if rows else None


def f(v, nd=1):
    return "-" if v is None else ("%." + str(nd) + "f") % v


def main():
    args = sys.argv[1:]
    refs = {}
    if "--ref" in args:
        i = args.index("--ref")
        for p in args[i + 1:]:
            d = load(p)
            if d and "n_gpus" in d:
                refs[d["n_gpus"]] = d
        args = args[:i]
    print("# bench.py results, round 2: ours vs NCCL (same processes, same buffers) vs the unmodified reference\n")
    print("Device-timed (CUDA events), barrier + synchronize on both sides, max over ranks; headline operand 256 MiB > 126 MB L2.")
    print("`robust` = `rabit.allreduce(tensor)`: the public, blocking, fault-tolerant call (consensus fused into the kernels, result")
    print("cached, caller's buffer intact on failure) on an ordinary torch tensor; `async` = stream-ordered, non-recoverable call;")
    print("`sym` = tensor from the symmetric heap (zero-copy kernels). NCCL = torch.distributed, NCCL 2.28.9, default settings, eager.\n")
    for p in args:
        d = load(p)
        if d is None:
            continue
        n = d["n_gpus"]
        print("## N = %d\n" % n)
        print("`%s`\n" % d["config"]["backend"])
        rs, asym, apl, nc = d.get("robust_symmetric"), d.get("async_symmetric"), d.get("async_plain"), d.get("nccl")
        print("* headline 256 MiB f32 Sum, **robust API, ordinary tensor: %s us** (busbw %s GB/s = %s of 900, %s of the "
              "measured 770)" % (f(d["ms_per_step"] * 1e3), f(d["busbw_GBps"]), f(d.get("nvlink_roofline_frac"), 2),
                                 f(d.get("busbw_frac_of_770_measured"), 2)))
        print("  - robust API, symmetric tensor: %s us; async plain: %s us; async symmetric: %s us (busbw %s = %s of 900); "
              "NCCL: %s us (busbw %s)"
              % (f(rs and rs["ms_per_step"] * 1e3), f(apl and apl["ms_per_step"] * 1e3), f(asym and asym["ms_per_step"] * 1e3),
                 f(asym and asym["busbw_GBps"]), f(asym and asym.get("busbw_frac_of_900"), 2), f(nc and nc["ms_per_step"] * 1e3),
                 f(nc and nc["busbw_GBps"])))
        e = d["e2e"]
        r = refs.get(n)
        print("* e2e (pinned host tensor -> `rabit.allreduce` -> same host tensor, H2D + D2H of 256 MiB each, every step): **%s ms/step**"
              " (%s GB/s aggregate); the caller doing the two copies around the device call: %s ms; copy ceiling of this box at this N"
              " (all ranks H2D + D2H at once, no collective): %s ms%s"
              % (f(e["ms_per_step"], 2), f(e["value"]), f(e["manual_copy_ms_per_step"], 2), f(e["copy_ceiling_ms"], 2),
                 ("; reference (host TCP engine): %s ms/step -> **%sx**"
                  % (f(r["ms_per_step"], 0), f(r["ms_per_step"] / e["ms_per_step"]))) if r else ""))
        rec = d.get("recovery") or {}
        cold = rec.get("without_standby_workers") or {}
        rr = (r or {}).get("recovery") or {}
        if rec:
            print("* recover-after-kill (model_recover scenario, 2 scripted kills, device buffers, native workers): %s s per kill with"
                  " standby workers, %s s without; reference (host only, no CUDA context): %s s"
                  % (f(rec.get("recover_s_per_kill"), 3), f(cold.get("recover_s_per_kill"), 3), f(rr.get("recover_s_per_kill"), 3)))
        sp = d.get("same_program") or {}
        if sp.get("host_engine") or sp.get("via_gpus"):
            print("* the reference arm's harness (`baseline/ref_speed.cc`, unchanged) linked with this library, host buffers in and out: "
                  "%s ms/step on the host engine (device backend off), %s ms/step with the host buffers riding the GPUs%s"
                  % (f((sp.get("host_engine") or {}).get("ms_per_step"), 1), f((sp.get("via_gpus") or {}).get("ms_per_step"), 1),
                     ("; the reference itself: %s ms/step" % f(r["ms_per_step"], 0)) if r else ""))
        sc = d.get("selfcheck") or {}
        print("* bit-exact self-check on these ranks: %s cases, failed: %s; kernels launched in the timed region: %s; clocks %s"
              % (sc.get("checked"), sc.get("failed"), d.get("gpu_launches"), json.dumps(d.get("clocks"))))
        h = d.get("histogram_pair_sum_f64_4MB")
        if h:
            print("* histogram pair-sum 4.096 MB f64: async sym %s us, async plain %s us, NCCL %s us, robust API %s us"
                  % (f(h.get("sym_us")), f(h.get("staged_us")), f(h.get("nccl_us")), f(h.get("robust_api_us"))))
        sw = d.get("sweep") or []
        if sw:
            print("\n| bytes | allreduce robust us | async plain | async sym | graph | NCCL | broadcast plain | sym | NCCL "
                  "| allgather plain | sym | NCCL |")
            print("|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
            for row in sw:
                g = lambda k: f(row.get(k))  # noqa: E731
                print("| %d | %s | %s | %s | %s | %s | %s | %s | %s | %s | %s | %s |" % (
                    row["bytes"], g("robust_us"), g("allreduce_us"), g("allreduce_sym_us"), g("allreduce_graph_us"),
                    g("nccl_allreduce_us"), g("broadcast_us"), g("broadcast_sym_us"), g("nccl_broadcast_us"), g("allgather_us"),
                    g("allgather_sym_us"), g("nccl_allgather_us")))
        if r and r.get("sweep"):
            print("\nreference (unmodified, host TCP, nodelay) allreduce us: " +
                  ", ".join("%d B: %s" % (x["bytes"], f(x["allreduce_us"], 0)) for x in r["sweep"]) +
                  "; 256 MiB with its default parameters: %s ms" % f((r.get("default_params") or {}).get("ms_per_step"), 0))
        print()


if __name__ == "__main__":
    main()
