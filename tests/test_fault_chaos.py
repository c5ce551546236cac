"""Random SIGKILLs at random moments (tests/chaos_harness.py): the job must still finish everywhere.
Global-model workloads only: they tolerate any number of overlapping failures as long as one rank survives
(local models legitimately cannot, see doc/guide.md)."""
import pytest

from chaos_harness import run


@pytest.mark.parametrize("seed", [3, 11, 261, 393, 1007, 1042, 2043, 4219, 8559])
def test_random_sigkills(built, seed):
    ok, info = run(seed)
    assert ok, {k: info[k] for k in ("prog", "world", "args", "kills", "done", "tails")}
