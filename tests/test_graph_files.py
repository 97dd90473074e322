"""Wider ingest: MatrixMarket and SNAP readers and the binary edge cache.
The host-only parser (b2bfs_read_graph_file) runs without a GPU; the loaded graph must give the
depths of the same graph written in the reference's own format."""
import os

import numpy as np
import pytest

import oracle


def _write_mtx(path, n, src, dst, weights=False, crlf=False):
    nl = "\r\n" if crlf else "\n"
    with open(path, "w", newline="") as f:
        f.write("%%MatrixMarket matrix coordinate " + ("real" if weights else "pattern") + " general" + nl)
        f.write("% a comment" + nl + "%" + nl)
        f.write(f"{n} {n} {len(src)}" + nl)
        for i, (u, v) in enumerate(zip(src, dst)):
            f.write(f"{u + 1} {v + 1}" + (f" {0.5 + i:.3f}" if weights else "") + nl)


def _write_snap(path, src, dst):
    with open(path, "w") as f:
        f.write("# Directed graph (each unordered pair of nodes is saved once)\n# Nodes: ? Edges: ?\n# FromNodeId\tToNodeId\n")
        for i, (u, v) in enumerate(zip(src, dst)):
            if i == len(src) // 2:
                f.write("\n# a comment in the middle\n")
            f.write(f"{u}\t{v}\n")


@pytest.fixture(scope="module")
def graph():
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, 11, 8, 4)
    return n, s, d


@pytest.mark.parametrize("weights,crlf", [(False, False), (True, False), (True, True)])
def test_matrix_market_reader(b2, graph, tmp_path, weights, crlf):
    n, s, d = graph
    p = str(tmp_path / "g.mtx")
    _write_mtx(p, n, s, d, weights, crlf)
    for fmt in (b2.FORMAT_AUTO, b2.FORMAT_MTX):
        n2, s2, d2 = b2.read_graph_file(p, fmt)
        assert n2 == n and np.array_equal(s2, s) and np.array_equal(d2, d)


def test_snap_reader_infers_vertex_count(b2, graph, tmp_path):
    n, s, d = graph
    p = str(tmp_path / "g.snap")
    _write_snap(p, s, d)
    n2, s2, d2 = b2.read_graph_file(p)  # auto-detected from the leading '#'
    assert n2 == int(max(s.max(), d.max())) + 1
    assert np.array_equal(s2, s) and np.array_equal(d2, d)


def test_reference_format_and_errors(b2, graph, tmp_path):
    n, s, d = graph
    p = str(tmp_path / "g.txt")
    oracle.write_edge_file(p, n, s, d)
    n2, s2, d2 = b2.read_graph_file(p)
    assert n2 == n and np.array_equal(s2, s) and np.array_equal(d2, d)
    bad = tmp_path / "bad.mtx"
    bad.write_text("%%MatrixMarket matrix coordinate pattern general\n3 3 2\n1 2\n")  # fewer entries than declared
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(bad))
    assert e.value.code == -3
    bad.write_text("%%MatrixMarket matrix coordinate pattern general\n3 3 1\n1 4\n")  # id beyond the dimension
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(bad))
    assert e.value.code == -4
    bad.write_text("%%MatrixMarket matrix coordinate pattern general\n3 3 1\n0 1\n")  # MatrixMarket ids are 1-based
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(bad))
    assert e.value.code == -4
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(tmp_path / "missing.mtx"))
    assert e.value.code == -2


def test_crafted_headers_are_format_errors_not_allocations(b2, tmp_path):
    import struct
    # binary cache whose edge count would wrap the size check (m * 8 overflows 64 bits)
    c = tmp_path / "wrap.cache"
    c.write_bytes(b"B2BFSEDG" + struct.pack("<qq", 8, (1 << 61) + 1) + b"\0" * 64)
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(c), b2.FORMAT_CACHE)
    assert e.value.code == -3
    # binary cache with an id outside [0, n)
    c.write_bytes(b"B2BFSEDG" + struct.pack("<qq", 4, 2) + struct.pack("<4i", 0, 1, 2, 9))
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(c), b2.FORMAT_CACHE)
    assert e.value.code == -4 and "9" in str(e.value)
    c.write_bytes(b"B2BFSEDG" + struct.pack("<qq", 4, 2) + struct.pack("<4i", 0, 1, 2, -1))
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(c), b2.FORMAT_CACHE)
    assert e.value.code == -4
    # a well-formed cache still loads
    c.write_bytes(b"B2BFSEDG" + struct.pack("<qq", 4, 2) + struct.pack("<4i", 0, 1, 2, 3))
    n, s, d = b2.read_graph_file(str(c), b2.FORMAT_CACHE)
    assert (n, s.tolist(), d.tolist()) == (4, [0, 1], [2, 3])
    # reference-format header promising 2^40 pairs in a 20-byte file: FORMAT, not NOMEM
    t = tmp_path / "huge.txt"
    t.write_text("5 1099511627776\n0 1\n")
    with pytest.raises(b2.B2BfsError) as e:
        b2.read_graph_file(str(t), b2.FORMAT_REFERENCE)
    assert e.value.code == -3 and "too short" in str(e.value)


@pytest.mark.gpu
def test_formats_load_the_same_graph_on_the_gpu(b2, graph, tmp_path):
    n, s, d = graph
    g = oracle.build_csr(n, s, d)
    want = oracle.bfs_serial(g, 2)
    mtx, snap, cache = str(tmp_path / "g.mtx"), str(tmp_path / "g.snap"), str(tmp_path / "g.bin")
    _write_mtx(mtx, n, s, d, weights=True)
    _write_snap(snap, s, d)
    with b2.Bfs(0) as h:
        h.load_graph_file(mtx)
        depth, _ = h.run(2)
        assert oracle.check_output(depth, want) == -1
        h.save_edge_cache(cache)
        for path in (snap, cache):
            h.load_graph_file(path)
            nn = h.num_vertices
            depth, _ = h.run(2)
            assert oracle.check_output(depth, want[:nn]) == -1 and np.all(want[nn:] == oracle.INT_MAX)
        es, ed = h.edges()
        assert np.array_equal(es, s) and np.array_equal(ed, d)
    n3, s3, d3 = b2.read_graph_file(cache)
    assert n3 == n and np.array_equal(s3, s) and np.array_equal(d3, d)
