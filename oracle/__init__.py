"""oracle — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes/numpy doors into
  * oracle/liboracle.so            the plain-C restatement of the reference's hot path
                                   (oracle/bfs_oracle.c; cites /root/reference/bfs.cu line by line)
  * oracle/_ref/libref_oracle.so   the reference's own host code (readGraphFromFile, bfsCPU),
                                   compiled unmodified from /root/reference/bfs.cu by oracle/Makefile

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this package.  The product (distributed-cuda-bfs_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
INT_MAX = 2147483647

GEN_RMAT, GEN_KRON, GEN_GRID = 0, 1, 2

_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def build(verbose: bool = False) -> None:
    """Compile the checker (and oracle/_ref when /root/reference is present)."""
    r = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


_lib = None
_ref = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.oracle_read_edge_file.argtypes = [C.c_char_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
        L.oracle_read_edge_file.restype = C.c_int
        L.oracle_write_edge_file.argtypes = [C.c_char_p, C.c_int64, C.c_int64, _i32p, _i32p]
        L.oracle_write_edge_file.restype = C.c_int
        L.oracle_build_csr.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _i64p, _i32p, _i32p]
        L.oracle_build_csr.restype = C.c_int
        L.oracle_bfs_serial.argtypes = [C.c_int64, _i64p, _i32p, _i32p, C.c_int64, _i32p, C.c_void_p]
        L.oracle_bfs_serial.restype = C.c_int64
        L.oracle_bfs_kdev.argtypes = [C.c_int64, _i64p, _i32p, _i32p, C.c_int64, C.c_int, _i32p]
        L.oracle_bfs_kdev.restype = C.c_int64
        L.oracle_check_output.argtypes = [C.c_int64, _i32p, _i32p]
        L.oracle_check_output.restype = C.c_int64
        L.oracle_generate_edges.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int64, C.c_int64, _i32p, _i32p]
        L.oracle_generate_edges.restype = None
        L.oracle_gen_num_vertices.argtypes = [C.c_int, C.c_int]
        L.oracle_gen_num_vertices.restype = C.c_int64
        L.oracle_gen_num_edges.argtypes = [C.c_int, C.c_int, C.c_int]
        L.oracle_gen_num_edges.restype = C.c_int64
        L.oracle_validate_tree.argtypes = [C.c_int64, _i64p, _i32p, _i32p, C.c_int64, _i32p, _i32p, C.POINTER(C.c_int64)]
        L.oracle_validate_tree.restype = C.c_int
        L.oracle_reach_stats.argtypes = [C.c_int64, _i32p, _i32p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.oracle_reach_stats.restype = None
        L.oracle_bfs_parallel.argtypes = [C.c_int64, _i64p, _i32p, _i32p, C.c_int64, _i32p, C.c_int]
        L.oracle_bfs_parallel.restype = C.c_int64
        L.oracle_num_threads.restype = C.c_int
        L.oracle_build_csr_directed.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _i64p, _i32p, _i32p]
        L.oracle_build_csr_directed.restype = C.c_int
        L.oracle_random_edges.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p]
        L.oracle_random_edges.restype = None
        L.oracle_build_csr_parallel.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _i64p, _i32p, _i32p, C.c_int]
        L.oracle_build_csr_parallel.restype = C.c_int
        L.oracle_check_depth_by_edges.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, C.c_int64, _i32p, C.c_int]
        L.oracle_check_depth_by_edges.restype = C.c_int64
        L.oracle_bfs_serial_bounded.argtypes = [C.c_int64, _i64p, _i32p, _i32p, C.c_int64, _i32p, C.c_int64,
                                                C.POINTER(C.c_int64)]
        L.oracle_bfs_serial_bounded.restype = C.c_int64
        _lib = L
    return _lib


def ref_available() -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libref_oracle.so"))


def ref() -> C.CDLL:
    """The reference's own host code (None-safe: raises if it was never built)."""
    global _ref
    if _ref is None:
        path = os.path.join(_HERE, "_ref", "libref_oracle.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle/_ref/libref_oracle.so missing: run `make -C oracle` where /root/reference exists")
        R = C.CDLL(path)
        R.ref_load_file.argtypes = [C.c_char_p]
        R.ref_load_file.restype = C.c_int
        R.ref_set_graph.argtypes = [C.c_int, C.c_int, _i32p, _i32p, _i32p]
        R.ref_set_graph.restype = C.c_int
        R.ref_num_vertices.restype = C.c_int
        R.ref_num_edges.restype = C.c_int
        R.ref_get_csr.argtypes = [_i32p, _i32p, _i32p]
        R.ref_get_csr.restype = None
        R.ref_read_graph_stdin.argtypes = [C.c_char_p]
        R.ref_read_graph_stdin.restype = C.c_int
        R.ref_read_graph_random.argtypes = [C.c_int, C.c_int]
        R.ref_read_graph_random.restype = C.c_int
        R.ref_bfs_cpu.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        R.ref_bfs_cpu.restype = C.c_int64
        _ref = R
    return _ref


# ------------------------------------------------------------------ restatement, numpy-facing
class Csr:
    """The reference's Graph (bfs.cu:21-27): adjacencyList / edgesOffset / edgesSize."""

    def __init__(self, n, offsets, sizes, adj):
        self.n, self.offsets, self.sizes, self.adj = int(n), offsets, sizes, adj

    @property
    def num_edges(self):  # G.numEdges = 2m, bfs.cu:875
        return int(self.adj.shape[0])


def read_edge_file(path: str):
    n, m = C.c_int64(), C.c_int64()
    rc = lib().oracle_read_edge_file(path.encode(), C.byref(n), C.byref(m), None, None)
    if rc:
        raise ValueError(f"oracle_read_edge_file({path}) -> {rc}")
    src = np.empty(m.value, np.int32)
    dst = np.empty(m.value, np.int32)
    rc = lib().oracle_read_edge_file(path.encode(), C.byref(n), C.byref(m), src.ctypes.data, dst.ctypes.data)
    if rc:
        raise ValueError(f"oracle_read_edge_file({path}) -> {rc}")
    return n.value, src, dst


def write_edge_file(path: str, n: int, src, dst) -> None:
    src = np.ascontiguousarray(src, np.int32)
    dst = np.ascontiguousarray(dst, np.int32)
    rc = lib().oracle_write_edge_file(path.encode(), n, src.shape[0], src, dst)
    if rc:
        raise OSError(f"oracle_write_edge_file -> {rc}")


def build_csr(n: int, src, dst) -> Csr:
    src = np.ascontiguousarray(src, np.int32)
    dst = np.ascontiguousarray(dst, np.int32)
    m = src.shape[0]
    offsets = np.empty(max(n, 1), np.int64)[:n]
    sizes = np.empty(max(n, 1), np.int32)[:n]
    adj = np.empty(2 * m, np.int32)
    rc = lib().oracle_build_csr(n, m, src, dst, np.ascontiguousarray(offsets), np.ascontiguousarray(sizes), adj)
    if rc:
        raise ValueError(f"oracle_build_csr -> {rc}")
    return Csr(n, offsets, sizes, adj)


def build_csr_directed(n: int, src, dst) -> Csr:
    """The reference's stdin reader (readGraph, bfs.cu:882-920): every tuple is the single arc u -> v."""
    src = np.ascontiguousarray(src, np.int32)
    dst = np.ascontiguousarray(dst, np.int32)
    m = src.shape[0]
    offsets = np.empty(max(n, 1), np.int64)[:n]
    sizes = np.empty(max(n, 1), np.int32)[:n]
    adj = np.empty(max(m, 1), np.int32)[:m]
    rc = lib().oracle_build_csr_directed(n, m, src, dst, offsets, sizes, adj)
    if rc:
        raise ValueError(f"oracle_build_csr_directed -> {rc}")
    return Csr(n, offsets, sizes, adj)


def random_edges(n: int, m: int):
    """The tuples of the reference's random-graph reader: srand(12345), rand() % n twice per edge."""
    src = np.empty(m, np.int32)
    dst = np.empty(m, np.int32)
    lib().oracle_random_edges(n, m, src, dst)
    return src, dst


def build_csr_parallel(n: int, src, dst, threads: int = 0) -> Csr:
    """All-cores build: same rows as build_csr as multisets, order inside a row unspecified (depths
    do not depend on it).  For bench.py on graphs where the serial builder would take minutes."""
    src = np.ascontiguousarray(src, np.int32)
    dst = np.ascontiguousarray(dst, np.int32)
    m = src.shape[0]
    offsets = np.empty(max(n, 1), np.int64)[:n]
    sizes = np.empty(max(n, 1), np.int32)[:n]
    adj = np.empty(2 * m, np.int32)
    rc = lib().oracle_build_csr_parallel(n, m, src, dst, offsets, sizes, adj, threads)
    if rc:
        raise ValueError(f"oracle_build_csr_parallel -> {rc}")
    return Csr(n, offsets, sizes, adj)


def check_depth_by_edges(n: int, src, dst, root: int, depth, threads: int = 0) -> int:
    """-1 if `depth` is exactly the BFS depth array of `root` over the tuples (three edge rules, see
    bfs_oracle.c), else an offending vertex.  Needs no CSR: usable at any scale the tuples fit."""
    rc = int(lib().oracle_check_depth_by_edges(n, src.shape[0], np.ascontiguousarray(src, np.int32),
                                               np.ascontiguousarray(dst, np.int32), root,
                                               np.ascontiguousarray(depth, np.int32), threads))
    if rc < -1:
        raise ValueError(f"oracle_check_depth_by_edges -> {rc}")
    return rc


def bfs_serial_bounded(g: Csr, start: int, max_entries: int):
    """(depth, entries_scanned, vertices_reached) of a serial search stopped after max_entries."""
    depth = np.empty(g.n, np.int32)
    reached = C.c_int64()
    rc = lib().oracle_bfs_serial_bounded(g.n, g.offsets, g.sizes, g.adj, start, depth, max_entries, C.byref(reached))
    if rc < 0:
        raise ValueError(f"oracle_bfs_serial_bounded -> {rc}")
    return depth, int(rc), reached.value


def bfs_serial(g: Csr, start: int, want_parent: bool = False):
    depth = np.empty(g.n, np.int32)
    parent = np.empty(g.n, np.int64) if want_parent else None
    rc = lib().oracle_bfs_serial(g.n, g.offsets, g.sizes, g.adj, start, depth,
                                 parent.ctypes.data if want_parent else None)
    if rc < 0:
        raise ValueError(f"oracle_bfs_serial -> {rc}")
    return (depth, parent) if want_parent else depth


def bfs_parallel(g: Csr, start: int, threads: int = 0):
    """Multi-core level-synchronous port; same depths as bfs_serial.
    threads = 0: all online cores."""
    depth = np.empty(g.n, np.int32)
    rc = lib().oracle_bfs_parallel(g.n, g.offsets, g.sizes, g.adj, start, depth, threads)
    if rc < 0:
        raise ValueError(f"oracle_bfs_parallel -> {rc}")
    return depth


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def bfs_kdev(g: Csr, start: int, k: int):
    depth = np.empty(g.n, np.int32)
    rc = lib().oracle_bfs_kdev(g.n, g.offsets, g.sizes, g.adj, start, k, depth)
    if rc < 0:
        raise ValueError(f"oracle_bfs_kdev -> {rc}")
    return depth, int(rc)


def check_output(got, expected) -> int:
    got = np.ascontiguousarray(got, np.int32)
    expected = np.ascontiguousarray(expected, np.int32)
    assert got.shape == expected.shape
    return int(lib().oracle_check_output(got.shape[0], got, expected))


def generate_edges(kind: int, scale: int, edgefactor: int, seed: int, first: int = 0, count: int | None = None):
    m = int(lib().oracle_gen_num_edges(kind, scale, edgefactor))
    if count is None:
        count = m - first
    src = np.empty(count, np.int32)
    dst = np.empty(count, np.int32)
    lib().oracle_generate_edges(kind, scale, edgefactor, seed, first, count, src, dst)
    return int(lib().oracle_gen_num_vertices(kind, scale)), src, dst


def validate_tree(g: Csr, root: int, depth, parent):
    """Graph500 rules; returns (0, -1) when valid, else (rule, offending vertex)."""
    bad = C.c_int64(-1)
    rc = lib().oracle_validate_tree(g.n, g.offsets, g.sizes, g.adj, root,
                                    np.ascontiguousarray(depth, np.int32), np.ascontiguousarray(parent, np.int32),
                                    C.byref(bad))
    return int(rc), int(bad.value)


def reach_stats(g: Csr, depth):
    v, d = C.c_int64(), C.c_int64()
    lib().oracle_reach_stats(g.n, g.sizes, np.ascontiguousarray(depth, np.int32), C.byref(v), C.byref(d))
    return v.value, d.value


def algorithmic_bytes(g: Csr, depth) -> int:
    """B_alg(r) = 4*D_r + 8*V_r + 8*n  (SURVEY.md §8d / BASELINE.md §3)."""
    v, d = reach_stats(g, depth)
    return 4 * d + 8 * v + 8 * g.n


# ------------------------------------------------------------------ the reference's own host code
def ref_load_file(path: str) -> Csr:
    R = ref()
    n = R.ref_load_file(path.encode())
    e = R.ref_num_edges()
    offsets = np.empty(n, np.int32)
    sizes = np.empty(n, np.int32)
    adj = np.empty(e, np.int32)
    R.ref_get_csr(offsets, sizes, adj)
    return Csr(n, offsets.astype(np.int64), sizes, adj)


def _ref_current_graph() -> Csr:
    R = ref()
    n, e = R.ref_num_vertices(), R.ref_num_edges()
    offsets = np.empty(n, np.int32)
    sizes = np.empty(n, np.int32)
    adj = np.empty(max(e, 1), np.int32)[:e]
    R.ref_get_csr(offsets, sizes, adj)
    return Csr(n, offsets.astype(np.int64), sizes, adj)


def ref_read_graph_stdin(path: str) -> Csr:
    """The reference's readGraph in its stdin form, fed from `path` (bfs.cu:882-920)."""
    ref().ref_read_graph_stdin(path.encode())
    return _ref_current_graph()


def ref_read_graph_random(n: int, m: int) -> Csr:
    """The reference's readGraph in its random form: srand(12345), n = argv[2], m = argv[3]."""
    ref().ref_read_graph_random(n, m)
    return _ref_current_graph()


def ref_set_graph(g: Csr) -> None:
    assert g.num_edges < 2**31, "the reference's int edge count overflows here (SURVEY.md D3)"
    ref().ref_set_graph(g.n, g.num_edges, np.ascontiguousarray(g.offsets, dtype=np.int32),
                        np.ascontiguousarray(g.sizes, np.int32), np.ascontiguousarray(g.adj, np.int32))


def ref_bfs_cpu(start: int, n: int, want_parent: bool = False):
    """bfsCPU on whatever graph the reference currently holds; returns (depth, [parent,] ns)."""
    depth = np.empty(n, np.int32)
    parent = np.empty(n, np.int32) if want_parent else None
    ns = ref().ref_bfs_cpu(start, depth.ctypes.data, parent.ctypes.data if want_parent else None)
    return (depth, parent, ns) if want_parent else (depth, ns)
