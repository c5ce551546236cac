"""Runs the C++ unit-test binary (test/unit_tests.cc): parameters, streams, mock, watchdog,
consensus decision table.  Mirrors the reference's gtest suite (test/cpp/*)."""
import subprocess

from conftest import bin_path


def test_cpp_unit_tests():
    r = subprocess.run([bin_path("unit_tests")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=120)
    out = r.stdout.decode()
    assert r.returncode == 0, out
    assert "0 failed" in out
