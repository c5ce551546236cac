#!/usr/bin/env python
"""Style gate of the repository (role of the reference's `make lint`, Makefile:95-97 + scripts/travis_script.sh:7-9,
which runs dmlc-core's cpplint wrapper). Self-contained: no external linter is needed.

C/C++/CUDA: no tabs, no trailing blanks, lines <= 140 columns, every header has an include guard that matches
its path, files end with a newline. Python: the same whitespace rules and every file must compile.
Exit code = number of findings (0 = clean).
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CXX_DIRS = ["include", "src", "test", "guide", "baseline/dmlc_shim", "baseline"]
PY_DIRS = ["rabit_b200", "tests", "benchmarks", "scripts", "guide", "test", "python", "baseline", "."]
CXX_EXT = (".h", ".cc", ".cu", ".cuh", ".inc")
MAX_COLS = 140


def files(dirs, exts, recurse=True):
    seen = set()
    for d in dirs:
        base = os.path.join(ROOT, d)
        if not os.path.isdir(base):
            continue
        for dirpath, dirnames, names in os.walk(base):
            dirnames[:] = [x for x in dirnames if x not in ("_ref", "build", "lib", "__pycache__", ".git", "gpurun_out")]
            for n in names:
                if n.endswith(exts):
                    seen.add(os.path.join(dirpath, n))
            if d == "." or not recurse:
                break
    return sorted(seen)


def guard_for(path):
    rel = os.path.relpath(path, ROOT)
    if rel.startswith("include/"):
        rel = rel[len("include/"):]
    elif rel.startswith("baseline/dmlc_shim/include/"):
        rel = rel[len("baseline/dmlc_shim/include/"):]
    else:
        rel = "rabit_" + rel
    return re.sub(r"[^A-Za-z0-9]", "_", rel).upper() + "_"


def check_text(path, findings, max_cols):
    data = open(path, "rb").read()
    if data and not data.endswith(b"\n"):
        findings.append("%s: no newline at end of file" % os.path.relpath(path, ROOT))
    for i, line in enumerate(data.decode("utf-8", errors="replace").split("\n"), 1):
        where = "%s:%d" % (os.path.relpath(path, ROOT), i)
        if "\t" in line:
            findings.append(where + ": tab character")
        if line != line.rstrip():
            findings.append(where + ": trailing whitespace")
        if len(line) > max_cols and "NOLINT" not in line and "http" not in line:
            findings.append(where + ": %d columns (max %d)" % (len(line), max_cols))


def main():
    findings = []
    for path in files(CXX_DIRS, CXX_EXT):
        check_text(path, findings, MAX_COLS)
        if path.endswith((".h", ".cuh")):
            text = open(path).read()
            m = re.search(r"#ifndef\s+(\w+)\s*\n\s*#define\s+(\w+)", text)
            if not m or m.group(1) != m.group(2):
                findings.append("%s: missing include guard" % os.path.relpath(path, ROOT))
    for path in files(PY_DIRS, (".py",)):
        check_text(path, findings, 140)
        try:
            compile(open(path).read(), path, "exec")
        except SyntaxError as e:
            findings.append("%s: %s" % (os.path.relpath(path, ROOT), e))
    for f in findings:
        print(f)
    print("%d finding(s)" % len(findings))
    return min(len(findings), 255)


if __name__ == "__main__":
    sys.exit(main())
