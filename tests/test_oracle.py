"""CPU tests: the oracle (oracle/bfs_oracle.c) against the golden vectors that the reference's own host
code produced (tests/golden/make_golden.py), against the reference library itself where it is
present (oracle/_ref, build container only), and against scipy as an independent third opinion."""
import os

import numpy as np
import pytest
import scipy.sparse as sp
from scipy.sparse.csgraph import shortest_path

import oracle
from conftest import golden_names, load_golden

INT_MAX = oracle.INT_MAX


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_golden(name):
    path, gold = load_golden(name)
    n, src, dst = oracle.read_edge_file(path)
    assert n == gold["n"] and len(src) == gold["m"]
    g = oracle.build_csr(n, src, dst)
    assert g.num_edges == gold["num_edges_printed"] == 2 * gold["m"]  # bfs.cu:875
    if gold["adj"] is not None:
        assert g.offsets.tolist() == gold["offsets"]
        assert g.sizes.tolist() == gold["sizes"]
        assert g.adj.tolist() == gold["adj"]
    chk = int((g.adj.astype(np.int64) * (np.arange(g.adj.size) % 1000003 + 1)).sum() % (2**61 - 1))
    assert chk == gold["csr_checksum"]
    for s in gold["searches"]:
        depth, parent = oracle.bfs_serial(g, s["root"], want_parent=True)
        assert oracle.check_output(depth, np.array(s["depth"], np.int32)) == -1
        assert parent.tolist() == s["parent_edge"]  # edge indices, bfs.cu:940


def test_survey_golden_vector():
    # the vector obtained from the reference's host code during the survey
    path, _ = load_golden("tiny")
    n, src, dst = oracle.read_edge_file(path)
    g = oracle.build_csr(n, src, dst)
    assert g.offsets.tolist() == [0, 2, 5, 7, 10, 10]
    assert g.sizes.tolist() == [2, 3, 2, 3, 0, 0]
    assert g.adj.tolist() == [1, 1, 0, 2, 0, 1, 3, 2, 3, 3]
    depth, parent = oracle.bfs_serial(g, 2, want_parent=True)
    assert depth.tolist() == [2, 1, 0, 1, INT_MAX, INT_MAX]
    assert parent.tolist() == [2, 5, 2, 6, INT_MAX, INT_MAX]


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("kind,scale,seed", [(oracle.GEN_RMAT, 10, 1), (oracle.GEN_KRON, 10, 7), (oracle.GEN_GRID, 10, 0)])
def test_oracle_matches_reference_library(tmp_path, kind, scale, seed):
    """Live comparison with the reference's loader and bfsCPU, through a real text file."""
    n, src, dst = oracle.generate_edges(kind, scale, 16, seed)
    path = str(tmp_path / "g.txt")
    oracle.write_edge_file(path, n, src, dst)
    gr = oracle.ref_load_file(path)
    n2, s2, d2 = oracle.read_edge_file(path)
    g = oracle.build_csr(n2, s2, d2)
    assert np.array_equal(g.offsets, gr.offsets) and np.array_equal(g.sizes, gr.sizes) and np.array_equal(g.adj, gr.adj)
    for root in (2, 0, n - 1, n // 2):
        dr, pr, _ = oracle.ref_bfs_cpu(root, n, want_parent=True)
        do, po = oracle.bfs_serial(g, root, want_parent=True)
        assert oracle.check_output(do, dr) == -1
        assert np.array_equal(po.astype(np.int32), pr)


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")
def test_ref_set_graph_equals_file_load(tmp_path):
    n, src, dst = oracle.generate_edges(oracle.GEN_RMAT, 9, 8, 3)
    g = oracle.build_csr(n, src, dst)
    oracle.ref_set_graph(g)
    d1, _ = oracle.ref_bfs_cpu(2, n)
    path = str(tmp_path / "g.txt")
    oracle.write_edge_file(path, n, src, dst)
    oracle.ref_load_file(path)
    d2, _ = oracle.ref_bfs_cpu(2, n)
    assert np.array_equal(d1, d2)


@pytest.mark.parametrize("seed", range(6))
def test_oracle_vs_scipy(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(2, 300))
    m = int(rng.integers(0, 3 * n))
    src = rng.integers(0, n, m).astype(np.int32)
    dst = rng.integers(0, n, m).astype(np.int32)
    g = oracle.build_csr(n, src, dst)
    a = sp.coo_matrix((np.ones(2 * m), (np.concatenate([src, dst]), np.concatenate([dst, src]))), shape=(n, n)).tocsr()
    root = int(rng.integers(0, n))
    ref = shortest_path(a, unweighted=True, indices=root)
    want = np.where(np.isinf(ref), INT_MAX, ref).astype(np.int64).astype(np.int32)
    assert np.array_equal(oracle.bfs_serial(g, root), want)


@pytest.mark.parametrize("k", [1, 2, 4, 8])
@pytest.mark.parametrize("seed", range(4))
def test_kdev_protocol_equals_serial(k, seed):
    """The reference's multi-device protocol (private distance arrays + owner buckets + min-merge,
    bfs.cu:569-619, 424-441 / bfs_mpi.cu:581-634) yields the serial depths."""
    rng = np.random.default_rng(100 + seed)
    n = 8 * int(rng.integers(1, 40))
    m = int(rng.integers(1, 4 * n))
    g = oracle.build_csr(n, rng.integers(0, n, m).astype(np.int32), rng.integers(0, n, m).astype(np.int32))
    for root in (0, n - 1, int(rng.integers(0, n))):
        depth, levels = oracle.bfs_kdev(g, root, k)
        want = oracle.bfs_serial(g, root)
        assert oracle.check_output(depth, want) == -1
        assert levels >= int(want[want != INT_MAX].max()) + 1  # re-discovered duplicates may add a level


def test_reader_rejects_what_the_reference_mishandles(tmp_path):
    p = tmp_path / "short.txt"
    p.write_text("4 3\n0 1\n1 2\n")  # fewer pairs than the header says (bfs.cu:851-861 reads garbage)
    with pytest.raises(ValueError):
        oracle.read_edge_file(str(p))
    p.write_text("4 1\n0 4\n")  # id >= n (bfs.cu:860 indexes out of range)
    with pytest.raises(ValueError):
        oracle.read_edge_file(str(p))
    with pytest.raises(ValueError):
        oracle.read_edge_file(str(tmp_path / "missing.txt"))
    p.write_text("3 2\r\n0 1\r\n\r\n1   2\r\n")  # CRLF / blank lines / runs of blanks: `>>` skips all
    n, s, d = oracle.read_edge_file(str(p))
    assert (n, s.tolist(), d.tolist()) == (3, [0, 1], [1, 2])


def test_generators_shape_and_determinism():
    for kind in (oracle.GEN_RMAT, oracle.GEN_KRON):
        n, s, d = oracle.generate_edges(kind, 12, 16, 1)
        assert n == 4096 and len(s) == 16 * 4096
        assert s.min() >= 0 and s.max() < n and d.min() >= 0 and d.max() < n
        _, s2, d2 = oracle.generate_edges(kind, 12, 16, 1, first=1000, count=500)
        assert np.array_equal(s2, s[1000:1500]) and np.array_equal(d2, d[1000:1500])
        _, s3, _ = oracle.generate_edges(kind, 12, 16, 2)
        assert not np.array_equal(s, s3)
    # RMAT without permutation: vertex 2 is a hub in the giant component
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 16, 16, 1)
    deg = np.bincount(np.concatenate([s, d]), minlength=n)
    assert deg[2] > 100
    # quadrant frequencies ~ (0.57, 0.19, 0.19, 0.05) on the top bit
    top = ((s >> 15) & 1) * 2 + ((d >> 15) & 1)
    f = np.bincount(top, minlength=4) / len(s)
    assert np.allclose(f, [0.57, 0.19, 0.19, 0.05], atol=0.01)
    # Kronecker = relabelled RMAT: the scramble is a bijection, so the degree multiset is the same
    nk, sk, dk = oracle.generate_edges(oracle.GEN_KRON, 16, 16, 1)
    degk = np.bincount(np.concatenate([sk, dk]), minlength=nk)
    assert np.array_equal(np.sort(deg), np.sort(degk))
    # grid: 2^5 x 2^5, 4-neighbour, each tuple once
    n, s, d = oracle.generate_edges(oracle.GEN_GRID, 10, 0, 0)
    assert n == 1024 and len(s) == 2 * 32 * 31
    assert set((d - s).tolist()) == {1, 32}
    g = oracle.build_csr(n, s, d)
    assert sorted(set(g.sizes.tolist())) == [2, 3, 4]
    depth = oracle.bfs_serial(g, 0)
    assert depth.max() == 62  # corner to corner: (32-1)+(32-1)


def test_validator_accepts_good_and_rejects_corrupted_trees():
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 9, 8, 5)
    g = oracle.build_csr(n, s, d)
    root = 2
    depth, pe = oracle.bfs_serial(g, root, want_parent=True)
    # edge index -> vertex id of the discovering endpoint: the row that contains entry i
    rows = np.repeat(np.arange(n, dtype=np.int32), g.sizes)
    parent = np.full(n, -1, np.int32)
    reached = depth != INT_MAX
    parent[reached] = rows[pe[reached].clip(0, g.adj.size - 1)]
    parent[root] = root
    assert oracle.validate_tree(g, root, depth, parent) == (0, -1)
    bad = parent.copy(); bad[root] = 3
    assert oracle.validate_tree(g, root, depth, bad)[0] == 1
    v = int(np.flatnonzero(reached & (depth >= 2))[0])
    bad = parent.copy(); bad[v] = root if root not in g.adj[g.offsets[v]:g.offsets[v] + g.sizes[v]] else v
    assert oracle.validate_tree(g, root, depth, bad)[0] in (2, 5)
    badd = depth.copy(); badd[v] += 1
    assert oracle.validate_tree(g, root, badd, parent)[0] in (2, 3)
    iso = int(np.flatnonzero(~reached)[0])
    bad = parent.copy(); bad[iso] = root
    assert oracle.validate_tree(g, root, depth, bad)[0] == 4
    bad = parent.copy(); bad[v] = -1
    assert oracle.validate_tree(g, root, depth, bad)[0] == 4


def test_algorithmic_bytes_formula():
    path, gold = load_golden("tiny")
    n, src, dst = oracle.read_edge_file(path)
    g = oracle.build_csr(n, src, dst)
    depth = oracle.bfs_serial(g, 2)
    # reached {0,1,2,3}: D_r = 2+3+2+3 = 10, V_r = 4  ->  4*10 + 8*4 + 8*6
    assert oracle.reach_stats(g, depth) == (4, 10)
    assert oracle.algorithmic_bytes(g, depth) == 40 + 32 + 48


@pytest.mark.parametrize("kind,scale", [(oracle.GEN_RMAT, 13), (oracle.GEN_KRON, 12), (oracle.GEN_GRID, 10)])
def test_multicore_port_equals_serial(kind, scale):
    """the all-cores port of bfsCPU gives the serial depths for any thread count."""
    n, s, d = oracle.generate_edges(kind, scale, 16, 2)
    g = oracle.build_csr(n, s, d)
    for root in (0, 2, n - 1):
        want = oracle.bfs_serial(g, root)
        for threads in (1, 3, 0):
            assert oracle.check_output(oracle.bfs_parallel(g, root, threads), want) == -1


@pytest.mark.parametrize("kind,scale", [(oracle.GEN_RMAT, 12), (oracle.GEN_KRON, 11), (oracle.GEN_GRID, 8)])
def test_parallel_csr_builder_holds_the_same_rows(kind, scale):
    n, src, dst = oracle.generate_edges(kind, scale, 8, 5)
    g = oracle.build_csr(n, src, dst)
    p = oracle.build_csr_parallel(n, src, dst, threads=4)
    assert np.array_equal(g.offsets, p.offsets) and np.array_equal(g.sizes, p.sizes)
    for v in list(range(0, n, max(1, n // 200))) + [int(np.argmax(g.sizes))]:
        a, b = g.offsets[v], g.offsets[v] + g.sizes[v]
        assert sorted(g.adj[a:b].tolist()) == sorted(p.adj[a:b].tolist())
    root = int(np.argmax(g.sizes))
    assert oracle.check_output(oracle.bfs_serial(p, root), oracle.bfs_serial(g, root)) == -1


@pytest.mark.parametrize("name", golden_names())
def test_edge_rule_check_accepts_reference_depths_on_golden(name):
    path, gold = load_golden(name)
    n, src, dst = oracle.read_edge_file(path)
    for s in gold["searches"]:
        assert oracle.check_depth_by_edges(n, src, dst, s["root"], np.array(s["depth"], np.int32)) == -1


def test_edge_rule_check_rejects_every_kind_of_wrong_depth():
    n, src, dst = oracle.generate_edges(oracle.GEN_RMAT, 10, 8, 3)
    g = oracle.build_csr(n, src, dst)
    root = int(np.argmax(g.sizes))
    good = oracle.bfs_serial(g, root)
    assert oracle.check_depth_by_edges(n, src, dst, root, good, threads=3) == -1
    reached = np.flatnonzero((good != INT_MAX) & (good > 0))
    rng = np.random.default_rng(0)
    for v in rng.choice(reached, 12, replace=False):
        for delta in (+1, -1):
            bad = good.copy()
            bad[v] += delta   # too deep: rule 3 or 2 trips; too shallow: rule 2 or 3 trips
            assert oracle.check_depth_by_edges(n, src, dst, root, bad) != -1, (v, delta)
        bad = good.copy()
        bad[v] = INT_MAX      # a reachable vertex reported unreached: rule 2
        assert oracle.check_depth_by_edges(n, src, dst, root, bad) != -1
    iso = np.flatnonzero(good == INT_MAX)
    if len(iso):
        bad = good.copy()
        bad[iso[0]] = 3       # an unreachable vertex given a depth: rule 2 (has edges) or rule 3 (has none)
        assert oracle.check_depth_by_edges(n, src, dst, root, bad) != -1
    bad = good.copy()
    bad[root] = 1
    assert oracle.check_depth_by_edges(n, src, dst, root, bad) != -1
    # all depths shifted by one keeps every edge rule but breaks the root rule
    bad = np.where(good == INT_MAX, INT_MAX, good + 1).astype(np.int32)
    assert oracle.check_depth_by_edges(n, src, dst, root, bad) != -1


def test_bounded_serial_search_is_a_prefix_of_the_full_one():
    n, src, dst = oracle.generate_edges(oracle.GEN_KRON, 12, 16, 1)
    g = oracle.build_csr(n, src, dst)
    root = int(np.argmax(g.sizes))
    full = oracle.bfs_serial(g, root)
    d, scanned, reached = oracle.bfs_serial_bounded(g, root, 1 << 60)
    assert oracle.check_output(d, full) == -1 and scanned == oracle.reach_stats(g, full)[1]
    assert reached == oracle.reach_stats(g, full)[0]
    d, scanned, reached = oracle.bfs_serial_bounded(g, root, 5000)
    assert 5000 <= scanned < 5000 + int(g.sizes.max()) + 1 and reached <= oracle.reach_stats(g, full)[0]
    seen = d != INT_MAX
    assert np.array_equal(d[seen], full[seen])  # whatever it labelled is labelled correctly


# ---- the reference's second reader, readGraph (bfs.cu:882-920): stdin form (directed) and random form
RG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "readgraph")


def _load_rg(name):
    import json
    with open(os.path.join(RG, name + ".json")) as f:
        return json.load(f)


def _same_csr(g, gold):
    assert g.n == gold["n"] and g.num_edges == gold["num_edges"]
    if gold["adj"] is not None:
        assert g.offsets.tolist() == gold["offsets"] and g.sizes.tolist() == gold["sizes"] and g.adj.tolist() == gold["adj"]
    chk = int((g.adj.astype(np.int64) * (np.arange(g.adj.size) % 1000003 + 1)).sum() % (2**61 - 1))
    assert chk == gold["csr_checksum"]


def test_directed_reader_restatement_matches_the_reference_vector():
    gold = _load_rg("stdin40")
    n, src, dst = oracle.read_edge_file(os.path.join(RG, "stdin40.txt"))
    g = oracle.build_csr_directed(n, src, dst)
    _same_csr(g, gold)  # rows in input order, arcs one way only (bfs.cu:901)
    assert g.num_edges == gold["m"]
    for s in gold["searches"]:
        assert oracle.check_output(oracle.bfs_serial(g, s["root"]), np.array(s["depth"], np.int32)) == -1
    assert (np.array(gold["searches"][-1]["depth"]) == INT_MAX).sum() == n - 1  # vertex 39 is a sink: it reaches nothing


@pytest.mark.parametrize("n,m", [(50, 120), (3000, 9000)])
def test_random_reader_restatement_matches_the_reference_vector(n, m):
    gold = _load_rg(f"random_{n}_{m}")
    src, dst = oracle.random_edges(n, m)  # srand(12345); rand() % n twice per edge (bfs.cu:892, 903-904)
    g = oracle.build_csr(n, src, dst)     # both directions pushed (bfs.cu:905-906), as readGraphFromFile does
    _same_csr(g, gold)
    for s in gold["searches"]:
        assert oracle.check_output(oracle.bfs_serial(g, s["root"]), np.array(s["depth"], np.int32)) == -1


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built (needs /root/reference)")
def test_second_reader_against_the_live_reference(tmp_path):
    rng = np.random.default_rng(5)
    n, m = 300, 1200
    s = rng.integers(0, n, m).astype(np.int32)
    d = rng.integers(0, n, m).astype(np.int32)
    p = str(tmp_path / "arcs.txt")
    oracle.write_edge_file(p, n, s, d)
    ref_g = oracle.ref_read_graph_stdin(p)
    mine = oracle.build_csr_directed(n, s, d)
    assert np.array_equal(ref_g.offsets, mine.offsets) and np.array_equal(ref_g.sizes, mine.sizes) and np.array_equal(ref_g.adj, mine.adj)
    for root in (0, 17, 299):
        depth, _ = oracle.ref_bfs_cpu(root, n)
        assert oracle.check_output(oracle.bfs_serial(mine, root), depth) == -1
    ref_g = oracle.ref_read_graph_random(777, 2500)
    rs, rd = oracle.random_edges(777, 2500)
    mine = oracle.build_csr(777, rs, rd)
    assert np.array_equal(ref_g.offsets, mine.offsets) and np.array_equal(ref_g.adj, mine.adj)
