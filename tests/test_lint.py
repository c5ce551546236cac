"""The repository passes its own style gate (scripts/lint.py; role of the reference's `make lint`)."""
import os
import subprocess
import sys

from conftest import ROOT


def test_lint_clean():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "lint.py")], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT)
    assert r.returncode == 0, r.stdout.decode(errors="replace")[-4000:]
