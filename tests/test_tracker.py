"""Tracker topology invariants the workers rely on."""
import pytest

from rabit_b200.parallel.tracker import build_topology


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 7, 8, 10, 16, 31, 64])
def test_topology_invariants(n):
    tree, parent, ring = build_topology(n)
    assert sorted(tree) == list(range(n))
    if n == 1:
        assert tree[0] == [] and parent[0] == -1
        return
    # ring is r -> r+1 (mod n): required by the ring collectives
    for r in range(n):
        assert ring[r] == ((r - 1) % n, (r + 1) % n)
    # tree: symmetric adjacency, n-1 edges, connected, exactly one root, parent is a neighbour
    edges = set()
    for r, nb in tree.items():
        assert len(nb) <= 3
        for q in nb:
            assert r in tree[q]
            edges.add((min(r, q), max(r, q)))
    assert len(edges) == n - 1
    roots = [r for r in range(n) if parent[r] == -1]
    assert len(roots) == 1
    seen, todo = {roots[0]}, [roots[0]]
    while todo:
        for q in tree[todo.pop()]:
            if q not in seen:
                seen.add(q)
                todo.append(q)
    assert len(seen) == n
    for r in range(n):
        if parent[r] != -1:
            assert parent[r] in tree[r]
