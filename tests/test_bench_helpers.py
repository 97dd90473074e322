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
    for gpus, want_kind, want_scale in ((1, "rmat", 24), (2, "kron", 25), (4, "kron", 26), (8, "kron", 27)):
        seen = {}
        monkeypatch.setattr(bench, "run_reference_arm", lambda args: seen.update(kind=args.kind, scale=args.scale, legacy=args.legacy_scale))
        monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--gpus", str(gpus)])
        bench.main()
        assert (seen["kind"], seen["scale"]) == (want_kind, want_scale)
        assert seen["legacy"] == want_scale - 2  # round 1's sizes ride along under rmat22 / weak22


def test_both_arms_print_the_same_config():
    import argparse
    a = argparse.Namespace(kind="kron", scale=26, ef=16, seed=1, root_seed=2, mode="do", engine="fused", no_dedup=False,
                           multi_engine=0, gpus=4)
    c = bench.workload_config(a)
    assert "n=67108864, m=1073741824" in c["workload"] and c["multi_engine"] == "peer-memory"
    a.gpus, a.kind, a.scale = 1, "grid", 24
    assert "n=16777216, m=33546240" in bench.workload_config(a)["workload"]


def test_reference_arm_runs_end_to_end_on_a_small_graph(monkeypatch, capsys):
    import json
    import sys
    monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--kind", "rmat", "--scale", "12", "--steps", "3",
                                      "--warmup", "1", "--cpu-budget", "5"])
    bench.main()
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["steps"] == 3
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["e2e"]["value"] == line["value"]


def test_reference_arm_falls_back_to_the_bounded_port(monkeypatch, capsys):
    # what runs when 2m >= 2^31 (the reference's int offsets overflow): the 64-bit port, stopped early
    import json
    import sys
    monkeypatch.setattr(oracle, "ref_available", lambda: False)
    monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--kind", "kron", "--scale", "13", "--steps", "2",
                                      "--warmup", "1", "--cpu-budget", "2"])
    bench.main()
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert line["cpu_baseline"]["kind"] == "port" and line["value"] > 0
    assert "stopped after" in line["cpu_baseline"]["sample"]
