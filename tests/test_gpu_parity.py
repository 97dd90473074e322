"""GPU parity tests (-m gpu): the CUDA path, called through the C-ABI (ctypes mirror), against the
oracle on the same inputs — bit-exact depths (the reference's own contract, checkOutput
bfs.cu:374-384), Graph500-valid parent trees, the golden vectors the reference's host code
produced, and the error behaviour of the boundary."""
import os
import subprocess

import numpy as np
import pytest

import oracle
from conftest import golden_names, load_golden

pytestmark = pytest.mark.gpu
INT_MAX = oracle.INT_MAX


@pytest.fixture(scope="module")
def gpu(b2):
    h = b2.Bfs(0)
    yield h
    h.close()


def _configs(b2):
    return [(b2.MODE_DIRECTION_OPT, b2.ENGINE_FUSED), (b2.MODE_TOP_DOWN, b2.ENGINE_FUSED),
            (b2.MODE_DIRECTION_OPT, b2.ENGINE_STEPWISE), (b2.MODE_TOP_DOWN, b2.ENGINE_STEPWISE)]


def _check_search(b2, h, g, root):
    want = oracle.bfs_serial(g, root)
    for mode, engine in _configs(b2):
        h.set_option("mode", mode)
        h.set_option("engine", engine)
        depth, parent, st = h.run(root, want_parent=True)
        bad = oracle.check_output(depth, want)
        assert bad == -1, f"root {root} mode {mode} engine {engine}: vertex {bad} got {depth[bad]} want {want[bad]}"
        assert oracle.validate_tree(g, root, depth, parent) == (0, -1)
        vr, dr = oracle.reach_stats(g, want)
        assert (st["reached_vertices"], st["reached_entries"]) == (vr, dr)
        assert st["levels"] == int(want[want != INT_MAX].max()) + 1
        assert st["kernel_launches"] >= 1
    h.set_option("mode", b2.MODE_DIRECTION_OPT)
    h.set_option("engine", b2.ENGINE_FUSED)


@pytest.mark.parametrize("kind,scale,ef,seed", [(0, 12, 16, 1), (1, 12, 16, 7), (2, 12, 0, 0), (0, 5, 3, 9)])
def test_device_generator_is_the_host_generator(b2, gpu, kind, scale, ef, seed):
    gpu.generate(kind, scale, ef, seed)
    n, s, d = oracle.generate_edges(kind, scale, ef, seed)
    gs, gd = gpu.edges()
    assert gpu.num_vertices == n and gpu.num_input_edges == len(s) and gpu.num_edges == 2 * len(s)
    assert np.array_equal(gs, s) and np.array_equal(gd, d)


@pytest.mark.parametrize("dedup", [0, 1])
@pytest.mark.parametrize("kind,scale", [(0, 11), (1, 10), (2, 8)])
def test_csr_build_matches_reference_rows(b2, kind, scale, dedup):
    """Row CONTENT equals the reference's rows (bfs.cu:851-875) as multisets (dedup=0) or as sets
    without self-loops (dedup=1); within-row order is ours (hubs first: descending floor(log2 degree),
    then ascending id), which no depth depends on."""
    n, s, d = oracle.generate_edges(kind, scale, 16, 3)
    g = oracle.build_csr(n, s, d)
    with b2.Bfs(0, dedup=dedup) as h:
        h.load_edges(n, s, d)
        rowptr, col = h.csr()
        assert np.array_equal(h.degrees(), g.sizes)  # the reference's edgesSize[]
        assert rowptr[0] == 0 and rowptr[-1] == len(col) == h.csr_entries
        assert np.all(np.diff(rowptr) >= 0)
        nl = np.bincount(s[s != d], minlength=n) + np.bincount(d[s != d], minlength=n)
        cls = np.floor(np.log2(np.maximum(nl, 1))).astype(np.int64).clip(0, 15)
        for v in range(n):
            ref_row = np.sort(g.adj[g.offsets[v]:g.offsets[v] + g.sizes[v]])
            if dedup:
                ref_row = np.unique(ref_row[ref_row != v])
            row = col[rowptr[v]:rowptr[v + 1]]
            assert np.array_equal(np.sort(row), ref_row), f"row {v}"
            key = (15 - cls[row]) * (1 << 32) + row  # hub-first order: class descending, then id ascending
            assert np.all(np.diff(key) >= 0), f"row {v} order"


@pytest.mark.parametrize("name", golden_names())
def test_golden_vectors_from_the_reference(b2, gpu, name):
    path, gold = load_golden(name)
    gpu.load_edge_file(path)
    assert gpu.num_vertices == gold["n"] and gpu.num_edges == gold["num_edges_printed"]
    for s in gold["searches"]:
        for mode, engine in _configs(b2):
            gpu.set_option("mode", mode)
            gpu.set_option("engine", engine)
            depth, _ = gpu.run(s["root"])
            assert oracle.check_output(depth, np.array(s["depth"], np.int32)) == -1
    gpu.set_option("mode", b2.MODE_DIRECTION_OPT)
    gpu.set_option("engine", b2.ENGINE_FUSED)


def test_config1_rmat16_src2_through_a_text_file(b2, gpu, tmp_path):
    """BASELINE.json configs[0]: RMAT scale-16 ef-16 edge-list file, src=2, depth bit-exact."""
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 16, 16, 1)
    path = str(tmp_path / "rmat16.txt")
    oracle.write_edge_file(path, n, s, d)
    gpu.load_edge_file(path)
    gs, gd = gpu.edges()
    assert np.array_equal(gs, s) and np.array_equal(gd, d)  # the threaded parser read every tuple
    g = oracle.build_csr(n, s, d)
    _check_search(b2, gpu, g, 2)
    if oracle.ref_available():  # build container only; the GPU box has the prebuilt library too
        oracle.ref_set_graph(g)
        ref_depth, _ = oracle.ref_bfs_cpu(2, n)
        depth, _ = gpu.run(2)
        assert oracle.check_output(depth, ref_depth) == -1


@pytest.mark.parametrize("kind,scale,ef,seed,nroots", [(0, 14, 16, 1, 6), (1, 14, 16, 5, 6), (2, 12, 0, 0, 4), (0, 10, 1, 4, 6)])
def test_random_roots_depth_and_parent(b2, gpu, kind, scale, ef, seed, nroots):
    n, s, d = oracle.generate_edges(kind, scale, ef, seed)
    g = oracle.build_csr(n, s, d)
    gpu.generate(kind, scale, ef, seed)
    rng = np.random.default_rng(seed)
    roots = list(rng.choice(np.flatnonzero(g.sizes > 0), nroots, replace=False)) + [int(np.flatnonzero(g.sizes == 0)[0])] \
        if (g.sizes == 0).any() else list(rng.choice(n, nroots, replace=False))
    for r in roots:
        _check_search(b2, gpu, g, int(r))


@pytest.mark.parametrize("gamma", [0, 1, 24, 1 << 20])
@pytest.mark.parametrize("kind,scale,ef,seed", [(0, 16, 16, 1), (1, 15, 16, 5), (2, 12, 0, 0), (0, 13, 2, 4)])
def test_claim_only_levels_change_no_depth(b2, kind, scale, ef, seed, gamma):
    """Option "gamma" decides when a top-down level only claims (no work lists for the next level)
    and forces the switch to bottom-up: gamma = 0 never, 2^20 at the first level >= 1 with any
    frontier at all.  Depths, tree validity and the reach statistics must not depend on it."""
    n, s, d = oracle.generate_edges(kind, scale, ef, seed)
    g = oracle.build_csr(n, s, d)
    with b2.Bfs(0, gamma=gamma) as h:
        h.generate(kind, scale, ef, seed)
        assert h.get_option("gamma") == gamma
        rng = np.random.default_rng(seed + gamma)
        for r in rng.choice(np.flatnonzero(g.sizes > 0), 4, replace=False):
            _check_search(b2, h, g, int(r))


@pytest.mark.parametrize("huge_min", [0, 4096])
@pytest.mark.parametrize("kind,scale,seed", [(0, 17, 1), (1, 16, 5)])
def test_hub_bin_changes_no_depth(b2, kind, scale, seed, huge_min):
    """Rows of at least "huge_min" entries are expanded by whole CTAs from shared memory (bulk
    asynchronous copies behind an mbarrier); 0 leaves them to the 128-entry warp items.  Same depths."""
    n, s, d = oracle.generate_edges(kind, scale, 16, seed)
    g = oracle.build_csr(n, s, d)
    with b2.Bfs(0, huge_min=huge_min) as h:
        h.generate(kind, scale, 16, seed)
        assert h.get_option("max_degree") >= 4096  # the graph does have hub rows
        hubs = np.argsort(g.sizes)[-3:]
        rng = np.random.default_rng(seed)
        roots = [int(hubs[-1]), int(hubs[0])] + [int(r) for r in rng.choice(np.flatnonzero(g.sizes > 0), 3, replace=False)]
        for r in roots:  # a hub as the source, and sources whose first levels meet the hubs
            _check_search(b2, h, g, r)


def test_hub_rows_and_no_dedup(b2):
    """A star whose centre row is cut into many wide work items, plus duplicates and self-loops kept."""
    n = 20000
    s = np.concatenate([np.zeros(n - 1, np.int32), [5, 5, 7]]).astype(np.int32)
    d = np.concatenate([np.arange(1, n, dtype=np.int32), [5, 6, 8]]).astype(np.int32)
    g = oracle.build_csr(n, s, d)
    for dedup in (0, 1):
        with b2.Bfs(0, dedup=dedup, ) as h:
            h.load_edges(n, s, d)
            assert h.get_option("wide_items") >= (n - 1 + 255) // 256
            for root in (0, 5, n - 1):
                _check_search(b2, h, g, root)


def test_odd_sizes_and_tiny_graphs(b2):
    rng = np.random.default_rng(3)
    for n in (1, 2, 31, 33, 127, 129, 1000):
        m = int(rng.integers(0, 3 * n + 1))
        s = rng.integers(0, n, m).astype(np.int32)
        d = rng.integers(0, n, m).astype(np.int32)
        g = oracle.build_csr(n, s, d)
        with b2.Bfs(0) as h:
            h.load_edges(n, s, d)
            for root in {0, n - 1, n // 2}:
                _check_search(b2, h, g, root)


def test_large_graph_properties(b2, gpu):
    """Scale-20 (1M vertices, 16.7M tuples): engines and modes agree bit for bit, and the oracle
    agrees with them and validates the tree."""
    gpu.generate(b2.KIND_RMAT, 20, 16, 1)
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 20, 16, 1)
    g = oracle.build_csr(n, s, d)
    deg = gpu.degrees()
    assert np.array_equal(deg, g.sizes)
    roots = np.random.default_rng(2).choice(np.flatnonzero(deg > 0), 3, replace=False)
    for r in roots:
        _check_search(b2, gpu, g, int(r))


def test_full_16m_vertex_grid_from_corner_and_centre(b2):
    """BASELINE configs[3]: the 4096 x 4096 grid (16.7 M vertices, thousands of levels).  From the corner
    (8191 levels, the longest search the graph has) and from the centre, every one of the n depths equals
    the reference's own bfsCPU where oracle/_ref is present (else the restatement); the cluster-mode
    kernel runs all of it, the plain kernel (option cluster = 0) must agree."""
    n, s, d = oracle.generate_edges(oracle.GEN_GRID, 24, 0, 0)
    g = oracle.build_csr_parallel(n, s, d)
    use_ref = oracle.ref_available()
    if use_ref:
        oracle.ref_set_graph(g)
    side = 4096
    with b2.Bfs(0) as h:
        h.generate(b2.KIND_GRID, 24, 0, 0)
        assert h.num_vertices == n == side * side and h.get_option("max_degree") == 4
        for root, levels in ((0, 2 * side - 1), (side // 2 * side + side // 2, side + 1)):
            want = oracle.ref_bfs_cpu(root, n)[0] if use_ref else oracle.bfs_serial(g, root)
            for cluster in (1, 0):
                h.set_option("cluster", cluster)
                depth, parent, st = h.run(root, want_parent=True)
                bad = oracle.check_output(depth, want)
                assert bad == -1, f"root {root} cluster {cluster}: vertex {bad} got {depth[bad]} want {want[bad]}"
                assert st["levels"] == levels and st["reached_vertices"] == n
                assert oracle.validate_tree(g, root, depth, parent) == (0, -1)


def test_boundary_errors(b2, tmp_path):
    with b2.Bfs(0) as h:
        with pytest.raises(b2.B2BfsError) as e:
            h.run(0)
        assert e.value.code == -7  # no graph yet
        p = tmp_path / "short.txt"
        p.write_text("4 3\n0 1\n1 2\n")
        with pytest.raises(b2.B2BfsError) as e:
            h.load_edge_file(str(p))
        assert e.value.code == -3
        p.write_text("4 1\n0 4\n")
        with pytest.raises(b2.B2BfsError) as e:
            h.load_edge_file(str(p))
        assert e.value.code == -4
        with pytest.raises(b2.B2BfsError) as e:
            h.load_edge_file(str(tmp_path / "missing.txt"))
        assert e.value.code == -2
        with pytest.raises(b2.B2BfsError) as e:
            h.load_edges(4, np.array([0, 9], np.int32), np.array([1, 2], np.int32))
        assert e.value.code == -4
        p.write_text("3 2\r\n0 1\r\n\r\n1   2\r\n")  # CRLF, blank lines, runs of blanks
        h.load_edge_file(str(p))
        depth, _ = h.run(0)
        assert depth.tolist() == [0, 1, 2]
        for bad in (-1, 3):
            with pytest.raises(b2.B2BfsError):
                h.run(bad)
        with pytest.raises(b2.B2BfsError):
            h.set_option("dedup", 0)  # graph-time option after load


def test_cli_prints_the_reference_lines(b2, tmp_path):
    """`b2bfs <src> <file>` — the reference's command line and stdout (bfs.cu:783-823)."""
    path, gold = load_golden("tiny")
    dump = str(tmp_path / "depth.txt")
    r = subprocess.run([b2.CLI_PATH, "2", path, "--dump-depth", dump], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    out = r.stdout.splitlines()
    assert out[0] == path
    for line in ("nodes num: 6", "edge num: 5", "finish load graph", "Number of vertices 6", "Number of edges 10",
                 "Starting sequential bfs.", "Starting queue parallel bfs.", "Output OK!"):
        assert line in out, line
    assert sum(l.startswith("Elapsed time in milliseconds : ") and l.endswith(" ms.") for l in out) == 2
    assert [int(x) for x in open(dump).read().split()] == gold["searches"][0]["depth"]
    r = subprocess.run([b2.CLI_PATH, "2", str(tmp_path / "nope.txt")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "not open" in r.stdout
