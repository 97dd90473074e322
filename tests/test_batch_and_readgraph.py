"""GPU tests (-m gpu) of SURVEY.md §8f row N4: the batched multi-source search (b2bfs_run_batch) and the
front end for the reference's second reader, readGraph (bfs.cu:882-920: directed graph from stdin,
random graph from srand(12345)).  Integer results: exact equality everywhere."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
INT_MAX = oracle.INT_MAX
RG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "readgraph")


def _gold(name):
    with open(os.path.join(RG, name + ".json")) as f:
        return json.load(f)


@pytest.mark.parametrize("kind,scale,ef,seed", [(0, 14, 16, 1), (1, 13, 16, 5), (2, 10, 0, 0), (0, 11, 2, 4)])
def test_batch_rows_equal_single_searches(b2, kind, scale, ef, seed):
    """Row i of the batch result = the depth array of a search from roots[i]: against the oracle, for
    64 roots that include duplicates, an isolated vertex and the hub, and for smaller batches."""
    n, s, d = oracle.generate_edges(kind, scale, ef, seed)
    g = oracle.build_csr(n, s, d)
    with b2.Bfs(0) as h:
        h.generate(kind, scale, ef, seed)
        rng = np.random.default_rng(seed)
        roots = list(rng.choice(np.flatnonzero(g.sizes > 0), 60, replace=False))
        roots += [int(np.argmax(g.sizes)), roots[0], roots[0]]
        roots.append(int(np.flatnonzero(g.sizes == 0)[0]) if (g.sizes == 0).any() else n - 1)
        assert len(roots) == 64
        want = {int(r): oracle.bfs_serial(g, int(r)) for r in set(roots)}
        for k in (64, 33, 7, 1):
            depth, st = h.run_batch(roots[:k])
            assert depth.shape == (k, n) and st["nroots"] == k
            for i, r in enumerate(roots[:k]):
                bad = oracle.check_output(depth[i], want[int(r)])
                assert bad == -1, f"batch of {k}, row {i} (root {r}): vertex {bad} got {depth[i][bad]} want {want[int(r)][bad]}"
                vr, dr = oracle.reach_stats(g, want[int(r)])
                assert (st["reached_vertices"][i], st["reached_entries"][i]) == (vr, dr)
            assert st["levels"] == max(int(want[int(r)][want[int(r)] != INT_MAX].max()) for r in roots[:k]) + 1
        # the single-source entry point still works on the same handle afterwards
        dsingle, _ = h.run(int(roots[5]))
        assert oracle.check_output(dsingle, want[int(roots[5])]) == -1


def test_batch_on_hub_rows_no_dedup_and_errors(b2):
    n = 20000  # a star whose centre row is cut into many items, plus duplicates and self-loops kept
    s = np.concatenate([np.zeros(n - 1, np.int32), [5, 5, 7]]).astype(np.int32)
    d = np.concatenate([np.arange(1, n, dtype=np.int32), [5, 6, 8]]).astype(np.int32)
    g = oracle.build_csr(n, s, d)
    for dedup in (0, 1):
        with b2.Bfs(0, dedup=dedup) as h:
            h.load_edges(n, s, d)
            roots = [0, 5, n - 1, 7, 8, 6]
            depth, _ = h.run_batch(roots)
            for i, r in enumerate(roots):
                assert oracle.check_output(depth[i], oracle.bfs_serial(g, r)) == -1
            with pytest.raises(b2.B2BfsError) as e:
                h.run_batch(list(range(65)))
            assert e.value.code == -1
            with pytest.raises(b2.B2BfsError) as e:
                h.run_batch([0, n])
            assert e.value.code == -1


def test_batch_large_graph_by_edge_rules(b2):
    """Scale-20: every one of 64 rows satisfies the three edge rules that pin BFS depths exactly."""
    with b2.Bfs(0) as h:
        h.generate(b2.KIND_KRON, 20, 16, 3)
        n, s, d = oracle.generate_edges(oracle.GEN_KRON, 20, 16, 3)
        deg = h.row_degrees()
        roots = np.random.default_rng(1).choice(np.flatnonzero(deg > 0), 64, replace=False)
        depth, st = h.run_batch(roots)
        for i, r in enumerate(roots):
            assert oracle.check_depth_by_edges(n, s, d, int(r), depth[i]) == -1, (i, r)
        single = h.run_device(int(roots[0]))
        assert st["reached_vertices"][0] == single["reached_vertices"]


def test_directed_graph_as_the_stdin_reader_builds_it(b2):
    gold = _gold("stdin40")
    n, s, d = oracle.read_edge_file(os.path.join(RG, "stdin40.txt"))
    for dedup in (0, 1):
        with b2.Bfs(0, directed=1, dedup=dedup) as h:
            h.load_edges(n, s, d)
            assert h.num_edges == gold["num_edges"]  # arcs, not 2m (bfs.cu:919)
            for engine in (b2.ENGINE_FUSED, b2.ENGINE_STEPWISE):
                h.set_option("engine", engine)
                for srch in gold["searches"]:
                    depth, st = h.run(srch["root"])
                    assert oracle.check_output(depth, np.array(srch["depth"], np.int32)) == -1
                    assert st["bu_levels"] == 0  # directed graphs are searched top-down only
    # a larger random digraph against the oracle's restatement
    rng = np.random.default_rng(11)
    n, m = 5000, 30000
    s = rng.integers(0, n, m).astype(np.int32)
    d = rng.integers(0, n, m).astype(np.int32)
    g = oracle.build_csr_directed(n, s, d)
    with b2.Bfs(0, directed=1) as h:
        h.load_edges(n, s, d)
        for root in (0, 1, 4999, int(s[0])):
            depth, _ = h.run(root)
            assert oracle.check_output(depth, oracle.bfs_serial(g, root)) == -1
        with pytest.raises(b2.B2BfsError):
            h.run_batch([0, 1])


def _cli(b2, args, stdin=None):
    return subprocess.run([b2.CLI_PATH] + args, input=stdin, capture_output=True, text=True, timeout=120)


def test_cli_front_end_for_the_second_reader(b2, tmp_path):
    # stdin form: `b2bfs <src> --stdin` and bare `b2bfs <src>` (argc <= 2, bfs.cu:888)
    gold = _gold("stdin40")
    text = open(os.path.join(RG, "stdin40.txt")).read()
    for srch in gold["searches"][:2]:
        for args in ([str(srch["root"]), "--stdin"], [str(srch["root"])]):
            dump = str(tmp_path / "d.txt")
            extra = ["--dump-depth", dump] if len(args) == 2 else []
            r = _cli(b2, args + extra, stdin=text)
            assert r.returncode == 0 and "Output OK!" in r.stdout, r.stdout + r.stderr
            assert f"Number of vertices {gold['n']}" in r.stdout and f"Number of edges {gold['num_edges']}" in r.stdout
            if extra:
                assert np.loadtxt(dump, dtype=np.int64).tolist() == srch["depth"]
    # random form: `--random N M` and readGraph's own argv layout `b2bfs <src> N M`
    for n, m in ((50, 120), (3000, 9000)):
        gold = _gold(f"random_{n}_{m}")
        for srch in gold["searches"]:
            for args in ([str(srch["root"]), "--random", str(n), str(m)], [str(srch["root"]), str(n), str(m)]):
                dump = str(tmp_path / "r.txt")
                r = _cli(b2, args + ["--dump-depth", dump])
                assert r.returncode == 0 and "Output OK!" in r.stdout, r.stdout + r.stderr
                assert f"Number of edges {gold['num_edges']}" in r.stdout
                assert np.loadtxt(dump, dtype=np.int64).tolist() == srch["depth"]
