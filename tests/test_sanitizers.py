"""ASan/UBSan and TSan runs of the host engine's fault scenarios (scripts/sanitize.py).
They rebuild everything with the sanitizer (about a minute each), so they only run when
RABIT_RUN_SANITIZERS=1 (scripts/ci.sh sets it)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.skipif(os.environ.get("RABIT_RUN_SANITIZERS") != "1",
                                reason="set RABIT_RUN_SANITIZERS=1 to run the sanitizer builds")


@pytest.mark.parametrize("kind", ["address", "thread"])
def test_fault_scenarios_under_sanitizer(kind):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize.py"), kind],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=1800)
    assert r.returncode == 0, r.stdout.decode(errors="replace")[-6000:]
