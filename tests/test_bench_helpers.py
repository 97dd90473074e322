"""CPU tests of bench.py's host logic: both arms (ours, --impl reference) must time the SAME roots,
and the harmonic mean / root rule must behave as documented."""
import importlib.util
import os

import numpy as np

import oracle

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


def test_both_arms_pick_the_same_roots():
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 12, 16, 1)
    g = oracle.build_csr(n, s, d)
    # ours arm: degree in the deduplicated, self-loop-free CSR (b2bfs_copy_row_degrees)
    dedup_deg = np.array([len(set(g.adj[g.offsets[v]:g.offsets[v] + g.sizes[v]].tolist()) - {v}) for v in range(n)])
    # reference arm: tuples that are not self-loops, counted on the host
    nl = bench.nonloop_degree(n, s, d)
    assert np.array_equal(dedup_deg > 0, nl > 0)
    giant = oracle.bfs_serial(g, int(np.argmax(nl))) != oracle.INT_MAX
    a = bench.pick_roots(dedup_deg, 67, 2, giant)
    b = bench.pick_roots(nl, 67, 2, giant)
    assert a == b and len(set(a)) == 67
    assert all(giant[r] and nl[r] > 0 for r in a)
    assert bench.pick_roots(nl, 67, 3, giant) != a  # the seed matters


def test_root_rule_excludes_self_loop_only_and_tiny_components():
    # vertex 3 has only a self-loop; vertices 4-5 form a 2-vertex component; 0-1-2 is the big one
    s = np.array([0, 1, 3, 4], np.int32)
    d = np.array([1, 2, 3, 5], np.int32)
    nl = bench.nonloop_degree(6, s, d)
    assert nl.tolist() == [1, 2, 1, 0, 1, 1]
    g = oracle.build_csr(6, s, d)
    giant = oracle.bfs_serial(g, int(np.argmax(nl))) != oracle.INT_MAX
    roots = set(bench.pick_roots(nl, 3, 2, giant))
    assert roots == {0, 1, 2}


def test_harmonic_mean():
    assert abs(bench.hmean([1.0, 1.0, 1.0]) - 1.0) < 1e-12
    assert abs(bench.hmean([2.0, 6.0]) - 3.0) < 1e-12
    assert bench.hmean([]) == 0.0
    # one near-zero sample dominates: the reason roots are drawn from the largest component
    assert bench.hmean([300.0] * 63 + [3e-5]) < 0.002


def test_workload_defaults_follow_the_gpu_count(monkeypatch, capsys):
    import sys
    for gpus, want_kind, want_scale in ((1, "rmat", 22), (2, "kron", 23), (8, "kron", 25)):
        seen = {}
        monkeypatch.setattr(bench, "run_reference_arm", lambda args: seen.update(kind=args.kind, scale=args.scale))
        monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--gpus", str(gpus)])
        bench.main()
        assert (seen["kind"], seen["scale"]) == (want_kind, want_scale)
