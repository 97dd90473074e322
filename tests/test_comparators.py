"""GPU tests (-m gpu) of SURVEY.md §8f row N3: the reference's alternative strategies as comparator
engines - prefix-sum queue generation (runCudaScanBfs, bfs.cu:748-781) and the topology-driven level
relax (multiBfs, bfs.cu:101-130) - against the oracle: bit-exact depths, Graph500-valid trees."""
import numpy as np
import pytest

import oracle
from conftest import golden_names, load_golden

pytestmark = pytest.mark.gpu
INT_MAX = oracle.INT_MAX


def _check(b2, h, g, root):
    want = oracle.bfs_serial(g, root)
    for engine in (b2.ENGINE_SCAN, b2.ENGINE_PULL):
        h.set_option("engine", engine)
        depth, parent, st = h.run(root, want_parent=True)
        bad = oracle.check_output(depth, want)
        assert bad == -1, f"engine {engine} root {root}: vertex {bad} got {depth[bad]} want {want[bad]}"
        assert oracle.validate_tree(g, root, depth, parent) == (0, -1)
        assert st["levels"] == int(want[want != INT_MAX].max()) + 1
        vr, dr = oracle.reach_stats(g, want)
        assert (st["reached_vertices"], st["reached_entries"]) == (vr, dr)
    h.set_option("engine", b2.ENGINE_FUSED)


@pytest.mark.parametrize("kind,scale,ef,seed", [(0, 14, 16, 1), (1, 13, 16, 5), (2, 10, 0, 0), (0, 10, 1, 4)])
def test_comparator_engines_match_the_oracle(b2, kind, scale, ef, seed):
    n, s, d = oracle.generate_edges(kind, scale, ef, seed)
    g = oracle.build_csr(n, s, d)
    with b2.Bfs(0) as h:
        h.generate(kind, scale, ef, seed)
        rng = np.random.default_rng(seed)
        roots = list(rng.choice(np.flatnonzero(g.sizes > 0), 4, replace=False)) + [int(np.argmax(g.sizes))]
        if (g.sizes == 0).any():
            roots.append(int(np.flatnonzero(g.sizes == 0)[0]))
        for r in roots:
            _check(b2, h, g, int(r))


@pytest.mark.parametrize("name", golden_names())
def test_comparator_engines_on_the_reference_vectors(b2, name):
    path, gold = load_golden(name)
    with b2.Bfs(0, dedup=0) as h:  # the reference's multigraph: a neighbour listed twice is queued once
        h.load_edge_file(path)
        for engine in (b2.ENGINE_SCAN, b2.ENGINE_PULL):
            h.set_option("engine", engine)
            for s in gold["searches"]:
                depth, _ = h.run(s["root"])
                assert oracle.check_output(depth, np.array(s["depth"], np.int32)) == -1, (name, engine, s["root"])


def test_comparator_engines_hub_row_multigraph_and_directed(b2):
    n = 20000
    s = np.concatenate([np.zeros(n - 1, np.int32), [5, 5, 7, 0, 0]]).astype(np.int32)
    d = np.concatenate([np.arange(1, n, dtype=np.int32), [5, 6, 8, 9, 9]]).astype(np.int32)
    g = oracle.build_csr(n, s, d)
    for dedup in (0, 1):
        with b2.Bfs(0, dedup=dedup) as h:
            h.load_edges(n, s, d)
            for root in (0, 5, n - 1):
                _check(b2, h, g, root)
    rng = np.random.default_rng(3)
    n, m = 4000, 20000
    s = rng.integers(0, n, m).astype(np.int32)
    d = rng.integers(0, n, m).astype(np.int32)
    gd = oracle.build_csr_directed(n, s, d)
    with b2.Bfs(0, directed=1) as h:
        h.load_edges(n, s, d)
        for engine in (b2.ENGINE_SCAN, b2.ENGINE_PULL):
            h.set_option("engine", engine)
            for root in (0, 7, 3999):
                depth, _ = h.run(root)
                assert oracle.check_output(depth, oracle.bfs_serial(gd, root)) == -1
