#!/usr/bin/env python
"""Host-engine sweep (role of reference test/speed_runner.py): speed_test over sizes x worker counts for
our host engine and, when built, the unmodified reference (baseline/_ref/bin/ref_speed is a different
harness; here both run the SAME program text: ours = test/speed_test.cc, reference = its own
test/speed_test.cc built by baseline/build_ref.py if present).

    python benchmarks/speed_runner.py [--workers 2,4,8] [--sizes 256,262144,16777216] [--out profiles/host_speed.json]
"""
import argparse
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200.parallel.launcher import launch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", default="2,4,8")
    ap.add_argument("--sizes", default="256,262144,16777216")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    progs = {"ours": os.path.join(ROOT, "build", "bin", "speed_test"),
             "reference": os.path.join(ROOT, "baseline", "_ref", "bin", "ref_speed_test")}
    rows = []
    for nw in [int(x) for x in args.workers.split(",")]:
        for n in [int(x) for x in args.sizes.split(",")]:
            nrep = max(3, min(1000, (64 << 20) // max(n * 4, 1)))
            for impl, exe in progs.items():
                if not os.path.exists(exe):
                    continue
                res = launch([exe, str(n), str(nrep), "rabit_enable_tcp_no_delay=true"], nw, capture=True, timeout=900)
                if not res.ok:
                    continue
                row = {"impl": impl, "workers": nw, "floats": n, "nrep": nrep}
                for m in res.messages:
                    g = re.match(r"(\w+)-speed: ([0-9.e+-]+) MB/sec", m)
                    if g:
                        row[g.group(1) + "_MBps"] = float(g.group(2))
                rows.append(row)
                print(json.dumps(row), flush=True)
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
