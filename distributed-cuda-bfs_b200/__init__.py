"""Python mirror of the C-ABI in include/b2bfs.h (ctypes over lib/libb2bfs.so).

This is plumbing for tests/ and bench.py, not a second implementation: every method is one call
into the shared library, which in turn runs hand-written sm_100a kernels.  There is no fallback —
importing works anywhere (so symbol checks can run without a GPU), but creating a handle raises
unless a CUDA device is present, and a missing library is an ImportError.

The module name has hyphens (it is the directory the task names); import it with
    importlib.import_module("distributed-cuda-bfs_b200")
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2BFS_LIB", os.path.join(_HERE, "lib", "libb2bfs.so"))  # override: A/B builds only
CLI_PATH = os.path.join(_HERE, "bin", "b2bfs")
HEADER_PATH = os.path.abspath(os.path.join(_HERE, "..", "include", "b2bfs.h"))

INT_MAX = 2147483647
KIND_RMAT, KIND_KRON, KIND_GRID = 0, 1, 2
MODE_DIRECTION_OPT, MODE_TOP_DOWN = 0, 1
ENGINE_FUSED, ENGINE_STEPWISE, ENGINE_SCAN, ENGINE_PULL = 0, 1, 2, 3

ERR_NAMES = {0: "OK", -1: "ERR_ARG", -2: "ERR_IO", -3: "ERR_FORMAT", -4: "ERR_RANGE", -5: "ERR_CUDA",
             -6: "ERR_NOMEM", -7: "ERR_STATE", -8: "ERR_COMM", -9: "ERR_UNSUPPORTED"}


class B2BfsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


class Stats(C.Structure):
    _fields_ = [("levels", C.c_int32), ("td_levels", C.c_int32), ("bu_levels", C.c_int32), ("reserved0", C.c_int32),
                ("reached_vertices", C.c_int64), ("reached_entries", C.c_int64), ("kernel_launches", C.c_int64),
                ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64), ("search_ms", C.c_double), ("total_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved0"}


class BatchStats(C.Structure):
    _fields_ = [("nroots", C.c_int32), ("levels", C.c_int32), ("kernel_launches", C.c_int64), ("d2h_bytes", C.c_int64),
                ("search_ms", C.c_double), ("total_ms", C.c_double),
                ("reached_vertices", C.c_int64 * 64), ("reached_entries", C.c_int64 * 64)]


def build(verbose: bool = False) -> None:
    """Compile lib/libb2bfs.so and bin/b2bfs for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", _HERE, "-j8", "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("libb2bfs build failed")


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `make -C {_HERE}` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, i64, i32p = C.c_void_p, C.c_int64, C.c_void_p
        sig = {
            "b2bfs_version": ([], C.c_char_p),
            "b2bfs_create": ([C.POINTER(vp), C.c_int], C.c_int),
            "b2bfs_create_rank": ([C.POINTER(vp), C.c_int, C.c_int, C.c_int, vp], C.c_int),
            "b2bfs_comm_unique_id": ([vp], C.c_int),
            "b2bfs_set_option": ([vp, C.c_char_p, i64], C.c_int),
            "b2bfs_get_option": ([vp, C.c_char_p, C.POINTER(i64)], C.c_int),
            "b2bfs_load_edge_file": ([vp, C.c_char_p], C.c_int),
            "b2bfs_load_edges": ([vp, i64, i64, i32p, i32p], C.c_int),
            "b2bfs_load_graph_file": ([vp, C.c_char_p, C.c_int], C.c_int),
            "b2bfs_save_edge_cache": ([vp, C.c_char_p], C.c_int),
            "b2bfs_read_graph_file": ([C.c_char_p, C.c_int, C.POINTER(i64), C.POINTER(i64), C.POINTER(vp), C.POINTER(vp)], C.c_int),
            "b2bfs_free_edges": ([vp, vp], None),
            "b2bfs_generate": ([vp, C.c_int, C.c_int, C.c_int, C.c_uint64], C.c_int),
            "b2bfs_num_vertices": ([vp, C.POINTER(i64)], C.c_int),
            "b2bfs_num_edges": ([vp, C.POINTER(i64)], C.c_int),
            "b2bfs_num_input_edges": ([vp, C.POINTER(i64)], C.c_int),
            "b2bfs_csr_entries": ([vp, C.POINTER(i64)], C.c_int),
            "b2bfs_copy_degrees": ([vp, i32p], C.c_int),
            "b2bfs_copy_row_degrees": ([vp, i32p], C.c_int),
            "b2bfs_copy_edges": ([vp, i64, i64, i32p, i32p], C.c_int),
            "b2bfs_copy_csr": ([vp, vp, i32p], C.c_int),
            "b2bfs_run": ([vp, i64, i32p, i32p, C.POINTER(Stats)], C.c_int),
            "b2bfs_run_device": ([vp, i64, C.POINTER(Stats)], C.c_int),
            "b2bfs_run_slice": ([vp, i64, i32p, i32p, C.POINTER(Stats)], C.c_int),
            "b2bfs_local_range": ([vp, C.POINTER(i64), C.POINTER(i64)], C.c_int),
            "b2bfs_host_register": ([vp, i64], C.c_int),
            "b2bfs_host_unregister": ([vp], C.c_int),
            "b2bfs_run_batch": ([vp, C.c_int, vp, i32p, C.POINTER(BatchStats)], C.c_int),
            "b2bfs_device_batch_depth": ([vp, C.POINTER(vp)], C.c_int),
            "b2bfs_device_depth": ([vp, C.POINTER(vp)], C.c_int),
            "b2bfs_device_parent": ([vp, C.POINTER(vp)], C.c_int),
            "b2bfs_last_trace": ([vp, vp, vp, vp, C.c_int], C.c_int),
            "b2bfs_search_noalloc": ([C.c_int, i64, i64, vp, vp, i64, vp, vp, vp, C.POINTER(C.c_int32)], C.c_int),
            "b2bfs_noalloc_last_error": ([], C.c_char_p),
            "b2bfs_host_alloc": ([C.POINTER(vp), i64], C.c_int),
            "b2bfs_host_free": ([vp], C.c_int),
            "b2bfs_destroy": ([vp], C.c_int),
            "b2bfs_last_error": ([vp], C.c_char_p),
        }
        for name, (args, res) in sig.items():
            fn = getattr(L, name)
            fn.argtypes, fn.restype = args, res
        _lib = L
    return _lib


NOALLOC_MAX_VERTICES = 262144
NOALLOC_MAX_ENTRIES = 8388608


def search_noalloc(device: int, n: int, entries: int, rowptr_ptr: int, col_ptr: int, src: int, depth_ptr: int,
                   parent_ptr: int, stream_ptr: int | None = None) -> int:
    """b2bfs_search_noalloc over raw DEVICE pointers (ints).  Returns the number of levels."""
    levels = C.c_int32(0)
    rc = lib().b2bfs_search_noalloc(device, n, entries, rowptr_ptr, col_ptr, src, depth_ptr, parent_ptr, stream_ptr,
                                    C.byref(levels))
    if rc:
        raise B2BfsError(rc, lib().b2bfs_noalloc_last_error().decode())
    return levels.value


FORMAT_AUTO, FORMAT_REFERENCE, FORMAT_MTX, FORMAT_SNAP, FORMAT_CACHE = -1, 0, 1, 2, 3


def read_graph_file(path: str, fmt: int = FORMAT_AUTO):
    """Host-only parse of a graph file (no GPU): returns (n, src, dst) as numpy arrays."""
    n, m = C.c_int64(), C.c_int64()
    ps, pd = C.c_void_p(), C.c_void_p()
    rc = lib().b2bfs_read_graph_file(os.fsencode(path), fmt, C.byref(n), C.byref(m), C.byref(ps), C.byref(pd))
    if rc:
        raise B2BfsError(rc, lib().b2bfs_last_error(None).decode())
    try:
        src = np.ctypeslib.as_array(C.cast(ps, C.POINTER(C.c_int32)), shape=(max(m.value, 1),))[:m.value].copy()
        dst = np.ctypeslib.as_array(C.cast(pd, C.POINTER(C.c_int32)), shape=(max(m.value, 1),))[:m.value].copy()
    finally:
        lib().b2bfs_free_edges(ps, pd)
    return n.value, src, dst


def comm_unique_id() -> bytes:
    """128-byte NCCL unique id (call on rank 0, hand to every rank's Bfs(...))."""
    buf = C.create_string_buffer(128)
    rc = lib().b2bfs_comm_unique_id(C.cast(buf, C.c_void_p))
    if rc:
        raise B2BfsError(rc, lib().b2bfs_last_error(None).decode())
    return buf.raw


class PinnedArray:
    """n x dtype in page-locked host memory (b2bfs_host_alloc), exposed as a numpy view."""

    def __init__(self, n: int, dtype=np.int32):
        self._p = C.c_void_p()
        nbytes = int(n) * np.dtype(dtype).itemsize
        rc = lib().b2bfs_host_alloc(C.byref(self._p), nbytes)
        if rc:
            raise B2BfsError(rc, "b2bfs_host_alloc")
        buf = (C.c_char * max(nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(n))

    @property
    def ptr(self):
        return self._p.value

    def free(self):
        if self._p:
            self.array = None
            lib().b2bfs_host_free(self._p)
            self._p = C.c_void_p()


class SharedPinnedArray:
    """n x dtype in a POSIX shared-memory file that every rank of a node maps and page-locks: ranks
    write their slices of a result into it over their own PCIe links (Bfs.run_slice) and any rank can
    read the whole array afterwards.  Rank 0 creates (and finally unlinks) the file."""

    def __init__(self, name: str, n: int, dtype=np.int32, create: bool = False):
        self.path = os.path.join("/dev/shm", name)
        self.nbytes = max(int(n) * np.dtype(dtype).itemsize, 1)
        self.created = create
        if create:
            with open(self.path, "wb") as f:
                f.truncate(self.nbytes)
        self._mm = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(self.nbytes,))
        self._mm[::4096] = 0  # touch every page before locking it
        self.array = self._mm.view(dtype)[:int(n)]
        self._registered = lib().b2bfs_host_register(self._mm.ctypes.data, self.nbytes) == 0

    def free(self):
        if self._mm is not None:
            if self._registered:
                lib().b2bfs_host_unregister(self._mm.ctypes.data)
            self.array = None
            self._mm = None
            if self.created:
                try:
                    os.unlink(self.path)
                except OSError:
                    pass


class Bfs:
    """One handle = one GPU holding one graph.  Mirrors the reference's initCuda2 + runCudaQueueBfs
    pair (bfs.cu:308, 542): load a graph once, then `run(src)` any number of times."""

    def __init__(self, device: int = 0, rank: int = 0, world: int = 1, nccl_id: bytes | None = None, **options):
        """world > 1: one rank of a multi-GPU job (one process per GPU); `nccl_id` is the 128-byte
        id from comm_unique_id() on rank 0, delivered to every rank over any side channel."""
        self._h = C.c_void_p()
        self.rank, self.world = rank, world
        if world > 1:
            assert nccl_id is not None and len(nccl_id) == 128
            buf = C.create_string_buffer(bytes(nccl_id), 128)
            rc = lib().b2bfs_create_rank(C.byref(self._h), device, rank, world, C.cast(buf, C.c_void_p))
        else:
            rc = lib().b2bfs_create(C.byref(self._h), device)
        if rc:
            raise B2BfsError(rc, lib().b2bfs_last_error(None).decode())
        for k, v in options.items():
            self.set_option(k, v)

    # -- plumbing
    def _check(self, rc):
        if rc:
            raise B2BfsError(rc, lib().b2bfs_last_error(self._h).decode())

    def close(self):
        if self._h:
            lib().b2bfs_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key: str, value: int):
        self._check(lib().b2bfs_set_option(self._h, key.encode(), int(value)))

    def get_option(self, key: str) -> int:
        v = C.c_int64()
        self._check(lib().b2bfs_get_option(self._h, key.encode(), C.byref(v)))
        return v.value

    def _get_i64(self, fn):
        v = C.c_int64()
        self._check(fn(self._h, C.byref(v)))
        return v.value

    # -- graph
    def load_edge_file(self, path: str):
        self._check(lib().b2bfs_load_edge_file(self._h, os.fsencode(path)))

    def load_graph_file(self, path: str, fmt: int = FORMAT_AUTO):
        self._check(lib().b2bfs_load_graph_file(self._h, os.fsencode(path), fmt))

    def save_edge_cache(self, path: str):
        self._check(lib().b2bfs_save_edge_cache(self._h, os.fsencode(path)))

    def load_edges(self, n: int, src, dst):
        src = np.ascontiguousarray(src, np.int32)
        dst = np.ascontiguousarray(dst, np.int32)
        assert src.shape == dst.shape
        self._check(lib().b2bfs_load_edges(self._h, n, src.shape[0], src.ctypes.data, dst.ctypes.data))

    def generate(self, kind: int, scale: int, edgefactor: int = 16, seed: int = 1):
        self._check(lib().b2bfs_generate(self._h, kind, scale, edgefactor, seed))

    @property
    def num_vertices(self):
        return self._get_i64(lib().b2bfs_num_vertices)

    @property
    def num_edges(self):  # 2m, what the reference prints
        return self._get_i64(lib().b2bfs_num_edges)

    @property
    def num_input_edges(self):
        return self._get_i64(lib().b2bfs_num_input_edges)

    @property
    def csr_entries(self):
        return self._get_i64(lib().b2bfs_csr_entries)

    def degrees(self):
        out = np.empty(self.num_vertices, np.int32)
        self._check(lib().b2bfs_copy_degrees(self._h, out.ctypes.data))
        return out

    def row_degrees(self):
        """Degree in the resident (deduplicated) CSR for all n vertices; collective when world > 1."""
        out = np.empty(self.num_vertices, np.int32)
        self._check(lib().b2bfs_copy_row_degrees(self._h, out.ctypes.data))
        return out

    def edges(self, first: int | None = None, count: int | None = None):
        """Input tuples back to the host.  A rank of a multi-GPU job holds (and returns by default) only
        its own share, tuples [tuples_first, tuples_first + tuples_count)."""
        if first is None:
            first = self.get_option("tuples_first")
            if count is None:
                count = self.get_option("tuples_count")
        if count is None:
            count = self.num_input_edges - first
        s = np.empty(count, np.int32)
        d = np.empty(count, np.int32)
        self._check(lib().b2bfs_copy_edges(self._h, first, count, s.ctypes.data, d.ctypes.data))
        return s, d

    def csr(self, rowptr_only: bool = False):
        rows = self.get_option("local_rows") if self.world > 1 else self.num_vertices
        rowptr = np.empty(rows + 1, np.int64)
        if rowptr_only:
            self._check(lib().b2bfs_copy_csr(self._h, rowptr.ctypes.data, None))
            return rowptr
        col = np.empty(self.csr_entries, np.int32)
        self._check(lib().b2bfs_copy_csr(self._h, rowptr.ctypes.data, col.ctypes.data))
        return rowptr, col

    # -- search
    def run(self, src: int, want_parent: bool = False, depth_out=None, parent_out=None):
        """Depths (and parents) into host buffers.  Returns (depth, [parent,] stats dict)."""
        n = self.num_vertices
        depth = depth_out if depth_out is not None else np.empty(n, np.int32)
        parent = parent_out if parent_out is not None else (np.empty(n, np.int32) if want_parent else None)
        st = Stats()
        self._check(lib().b2bfs_run(self._h, int(src), depth.ctypes.data,
                                    parent.ctypes.data if parent is not None else None, C.byref(st)))
        return (depth, parent, st.as_dict()) if parent is not None else (depth, st.as_dict())

    def run_collective(self, src: int, depth_out=None):
        """Multi-rank: every rank calls this; only ranks passing a buffer receive the n depths."""
        st = Stats()
        self._check(lib().b2bfs_run(self._h, int(src), depth_out.ctypes.data if depth_out is not None else None,
                                    None, C.byref(st)))
        return st.as_dict()

    def local_range(self):
        """(first vertex, count) of the rows this rank owns (all of them on a single GPU)."""
        a, b = C.c_int64(), C.c_int64()
        self._check(lib().b2bfs_local_range(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def run_slice(self, src: int, depth_slice, parent_slice=None):
        """Multi-rank: every rank calls this; each receives only its own rows (no cross-GPU gather)."""
        st = Stats()
        self._check(lib().b2bfs_run_slice(self._h, int(src), depth_slice.ctypes.data,
                                          parent_slice.ctypes.data if parent_slice is not None else None, C.byref(st)))
        return st.as_dict()

    def run_batch(self, roots, depth_out=None, to_host: bool = True):
        """Up to 64 searches in one pass per level (b2bfs_run_batch).  Returns (depth[len(roots), n] or
        None, stats dict with per-root reached_vertices / reached_entries lists)."""
        roots = np.ascontiguousarray(roots, np.int64)
        k = int(roots.shape[0])
        depth = None
        if to_host:
            depth = depth_out if depth_out is not None else np.empty((k, self.num_vertices), np.int32)
        st = BatchStats()
        self._check(lib().b2bfs_run_batch(self._h, k, roots.ctypes.data, depth.ctypes.data if depth is not None else None,
                                          C.byref(st)))
        return depth, {"nroots": st.nroots, "levels": st.levels, "kernel_launches": st.kernel_launches,
                       "d2h_bytes": st.d2h_bytes, "search_ms": st.search_ms, "total_ms": st.total_ms,
                       "reached_vertices": list(st.reached_vertices[:k]), "reached_entries": list(st.reached_entries[:k])}

    def run_device(self, src: int):
        st = Stats()
        self._check(lib().b2bfs_run_device(self._h, int(src), C.byref(st)))
        return st.as_dict()

    def last_trace(self, cap: int = 96):
        """[(level, phase, microseconds since kernel start)] of the last fused search."""
        lv = np.empty(cap, np.int32)
        ph = np.empty(cap, np.int32)
        us = np.empty(cap, np.float64)
        n = lib().b2bfs_last_trace(self._h, lv.ctypes.data, ph.ctypes.data, us.ctypes.data, cap)
        if n < 0:
            raise B2BfsError(n, "b2bfs_last_trace")
        return [(int(lv[i]), int(ph[i]), float(us[i])) for i in range(n)]

    def device_depth_ptr(self) -> int:
        p = C.c_void_p()
        self._check(lib().b2bfs_device_depth(self._h, C.byref(p)))
        return p.value
