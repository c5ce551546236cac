"""pytest configuration: markers, in-tree build, helpers to run multi-process jobs."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200 box)")


def _have_nvcc():
    import shutil
    return shutil.which("nvcc") is not None or os.path.exists("/usr/local/cuda/bin/nvcc")


@pytest.fixture(scope="session", autouse=True)
def built():
    """Make sure lib/ and build/bin are up to date before any test runs."""
    import build as _build
    _build.build(cuda=_have_nvcc(), verbose=False)
    return True


def bin_path(name):
    return os.path.join(ROOT, "build", "bin", name)


def run_job(cmd, nworker, timeout=120, **kw):
    """Launch cmd under a private tracker; returns rabit_b200.parallel.launcher.JobResult."""
    from rabit_b200.parallel.launcher import launch
    return launch(cmd, nworker, capture=True, timeout=timeout, **kw)


def run_py(code, nworker, args=(), timeout=120, **kw):
    """Run a python snippet as nworker rabit workers."""
    env = dict(kw.pop("env", {}) or {})
    env["PYTHONPATH"] = ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")
    return run_job([sys.executable, "-c", code] + list(args), nworker, timeout=timeout, env=env, **kw)
