#!/usr/bin/env python
"""Build the host engine and its test programs with a sanitizer and run the fault scenarios under it
(the reference has no sanitizer support at all).

    python scripts/sanitize.py address      # ASan + UBSan
    python scripts/sanitize.py thread       # TSan (engine thread + recovery watchdog thread)

Everything is linked statically into build/san_<kind>/<program> (host-only: device_stub.cc), then the
programs run under the normal launcher/tracker with scripted kills. Exit code 0 = no sanitizer report.
"""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rabit_b200.parallel.launcher import launch  # noqa: E402

HOST = ["comm", "engine_base", "engine_robust", "engine_recover", "engine_ops", "engine_mock", "result_store", "fake_device",
        "c_api", "engine"]
PROGRAMS = ["model_recover", "local_recover", "lazy_recover", "speed_test", "unit_tests"]


def build(kind):
    out = os.path.join(ROOT, "build", "san_" + kind)
    os.makedirs(out, exist_ok=True)
    san = "-fsanitize=address,undefined" if kind == "address" else "-fsanitize=thread"
    flags = ["-std=c++17", "-O1", "-g", "-fno-omit-frame-pointer", san, "-pthread", "-I" + os.path.join(ROOT, "include"),
             '-DRABIT_DEFAULT_ENGINE="mock"']
    objs = []
    procs = []
    for name in HOST:
        o = os.path.join(out, name + ".o")
        objs.append(o)
        procs.append(subprocess.Popen(["g++"] + flags + ["-c", os.path.join(ROOT, "src/host", name + ".cc"), "-o", o]))
    o = os.path.join(out, "device_stub.o")
    objs.append(o)
    procs.append(subprocess.Popen(["g++"] + flags + ["-c", os.path.join(ROOT, "src/device/device_stub.cc"), "-o", o]))
    for p in procs:
        assert p.wait() == 0, "compile failed"
    procs = [subprocess.Popen(["g++"] + flags + [os.path.join(ROOT, "test", prog + ".cc")] + objs +
                              ["-o", os.path.join(out, prog), "-ldl"]) for prog in PROGRAMS]
    for p in procs:
        assert p.wait() == 0, "link failed"
    return out


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "address"
    assert kind in ("address", "thread")
    out = build(kind)
    env = {"ASAN_OPTIONS": "detect_leaks=1:abort_on_error=0:exitcode=66", "UBSAN_OPTIONS": "print_stacktrace=1:halt_on_error=1:exitcode=66",
           "TSAN_OPTIONS": "exitcode=66:halt_on_error=1"}
    scenarios = [
        ("unit_tests", [], 1),
        ("speed_test", ["100000", "3"], 4),
        ("model_recover", ["10000", "rabit_bootstrap_cache=1"], 6),
        ("model_recover", ["10000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=2,1,0,0", "rabit_bootstrap_cache=1",
                           "rabit_reduce_ring_mincount=1"], 6),
        ("model_recover", ["10000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "mock=0,1,1,0", "mock=4,1,1,0",
                           "mock=5,1,1,0", "rabit_bootstrap_cache=1"], 6),
        ("local_recover", ["10000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=0,1,1,1", "mock=4,1,1,0", "mock=5,1,1,0"], 6),
        ("lazy_recover", ["10000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "mock=0,1,1,0", "mock=4,1,1,0"], 6),
        ("model_recover", ["2000", "mock=1,1,1,0", "rabit_timeout=1", "rabit_timeout_sec=60"], 4),
        # the device-backend side of the engines (result cache in backend memory, op pieces, backend transport for
        # recovery payloads and local replicas) on the host-memory stand-in of the backend
        ("model_recover", ["20000", "mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1", "rabit_bootstrap_cache=1", "rabit_device=fake",
                           "rabit_device_host_min=1", "rabit_device_blob_min=1", "rabit_device_arena=64K"], 4),
        ("local_recover", ["20000", "mock=0,1,1,0", "mock=3,2,1,0", "rabit_device=fake", "rabit_device_host_min=1",
                           "rabit_device_blob_min=1"], 5),
        ("lazy_recover", ["20000", "mock=1,1,0,0", "mock=2,2,3,0", "rabit_device=fake", "rabit_device_host_min=1",
                          "rabit_device_blob_min=1"], 4),
    ]
    bad = 0
    for prog, args, n in scenarios:
        exe = os.path.join(out, prog)
        if n == 1:
            r = subprocess.run([exe] + args, env=dict(os.environ, **env), stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
            text, ok = r.stdout.decode(errors="replace"), r.returncode == 0
        else:
            res = launch([exe] + args, n, capture=True, timeout=600, env=env)
            text, ok = "\n".join(res.outputs.values()), res.ok
        report = ("Sanitizer" in text and "ERROR" in text) or "runtime error:" in text or "WARNING: ThreadSanitizer" in text
        print("%-14s n=%d %-60s %s" % (prog, n, " ".join(args)[:60], "ok" if ok and not report else "FAILED"), flush=True)
        if not ok or report:
            bad += 1
            lines = text.splitlines()
            hits = [i for i, l in enumerate(lines) if "Sanitizer" in l or "runtime error" in l]
            start = hits[0] if hits else max(len(lines) - 40, 0)
            print("\n".join(lines[start:start + 60]))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
