#!/usr/bin/env python
"""Build the UNMODIFIED reference for the `bench.py --impl reference` arm.

The reference ships no setup.py / pyproject.toml, so
  python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
      --target baseline/_ref /root/reference
cannot work (recorded in DESIGN.md); it is a C++ library that expects dmlc-core beside it.
This script copies the reference's `include/` and `src/` byte-for-byte into
baseline/_ref/reference/ (git-ignored; travels to the GPU box), compiles them with g++ against the
three-header shim in baseline/dmlc_shim/ (which provides the dmlc API generation the snapshot was
written for), and links baseline/ref_speed.cc, a harness that only uses the reference's public API.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
DST = os.path.join(HERE, "_ref")


def build(force=False):
    copy = os.path.join(DST, "reference")
    if os.path.isdir(REF_SRC) and (force or not os.path.isdir(copy)):
        shutil.rmtree(copy, ignore_errors=True)
        os.makedirs(copy)
        for sub in ("include", "src", "test", "guide", "LICENSE"):
            s = os.path.join(REF_SRC, sub)
            if os.path.isdir(s):
                shutil.copytree(s, os.path.join(copy, sub))
            elif os.path.isfile(s):
                shutil.copy(s, os.path.join(copy, sub))
    if not os.path.isdir(copy):
        raise RuntimeError("no copy of the reference under %s and %s is not mounted" % (copy, REF_SRC))
    bin_dir = os.path.join(DST, "bin")
    os.makedirs(bin_dir, exist_ok=True)
    flags = ["-std=c++11", "-O3", "-msse2", "-pthread", "-w", "-I" + os.path.join(copy, "include"),
             "-I" + os.path.join(HERE, "dmlc_shim", "include"), "-I" + os.path.join(copy, "src")]
    srcs = [os.path.join(copy, "src", f) for f in ("allreduce_base.cc", "allreduce_robust.cc", "engine.cc", "c_api.cc")]
    for name, main in (("ref_speed", os.path.join(HERE, "ref_speed.cc")),
                       ("ref_model_recover", os.path.join(copy, "test", "model_recover.cc")),
                       ("ref_speed_test", os.path.join(copy, "test", "speed_test.cc")),
                       ("ref_basic", os.path.join(copy, "guide", "basic.cc"))):
        # the reference's test programs are linked with its mock engine (test/Makefile:67)
        eng = [os.path.join(copy, "src", "engine_mock.cc" if name == "ref_model_recover" else "engine.cc")]
        cmd = ["g++"] + flags + [main] + srcs[:2] + eng + [srcs[3], "-o", os.path.join(bin_dir, name)]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        if r.returncode != 0:
            raise RuntimeError("g++ failed for %s:\n%s" % (name, r.stdout.decode()[-2000:]))
    return bin_dir


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
