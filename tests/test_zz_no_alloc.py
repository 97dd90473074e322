"""GPU test of b2bfs_search_noalloc — the entry point for callers that cannot allocate device memory
at search time (they run under another framework's memory pool, or did all their allocation at
start-up).  The graph is handed over as device pointers; scratch is the array built into the library.

Collected last on purpose (the file name sorts after every other test file): if it fails under
`pytest -x`, everything else has already run.  It must pass alone in a fresh process and after the
other GPU tests — the built-in array is shared by every call and keeps whatever the last call left.

The check: depths and parents equal, element for element, what the ordinary handle-based entry point
(b2bfs_run) returns for the same graph and roots, on two calls in a row (the second starts from the
first one's leftovers), plus the Graph500 tree rules from the oracle.  Integer results: exact equality.
"""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


def _device_csr(b2, h):
    rowptr, col = h.csr()
    # torch only provides the device buffers here; the library itself allocates nothing in this call
    return (rowptr, col, torch.from_numpy(rowptr).cuda(), torch.from_numpy(col).cuda())


@pytest.mark.parametrize("kind,scale,ef", [(0, 14, 16), (2, 12, 0)])
def test_noalloc_matches_handle_based_search_twice_in_a_row(b2, kind, scale, ef):
    with b2.Bfs(0) as h:
        h.generate(kind, scale, ef, 3)
        n = h.num_vertices
        rowptr, col, d_rowptr, d_col = _device_csr(b2, h)
        d_depth = torch.empty(n, dtype=torch.int32, device="cuda")
        d_parent = torch.empty(n, dtype=torch.int32, device="cuda")
        deg = np.diff(rowptr)
        roots = [int(np.argmax(deg)), int(np.flatnonzero(deg > 0)[-1])]
        src, dst = h.edges()
        g = oracle.build_csr(n, src, dst)
        for root in roots:  # two calls back to back on the same shared scratch
            want_depth, want_parent, st = h.run(root, want_parent=True)
            d_depth.fill_(-7)
            d_parent.fill_(-7)
            torch.cuda.synchronize()
            levels = b2.search_noalloc(0, n, len(col), d_rowptr.data_ptr(), d_col.data_ptr(), root,
                                       d_depth.data_ptr(), d_parent.data_ptr())
            got_depth = d_depth.cpu().numpy()
            got_parent = d_parent.cpu().numpy()
            assert oracle.check_output(got_depth, want_depth) == -1
            assert oracle.check_output(got_depth, oracle.bfs_serial(g, root)) == -1  # and the oracle directly
            assert levels == st["levels"]
            # parents may legitimately differ between two valid trees; both must satisfy the rules
            assert oracle.validate_tree(g, root, got_depth, got_parent) == (0, -1)
            assert np.array_equal(got_parent >= 0, want_parent >= 0)


def test_noalloc_rejects_graphs_beyond_the_built_in_scratch(b2):
    d = torch.zeros(16, dtype=torch.int64, device="cuda")
    out = torch.zeros(16, dtype=torch.int32, device="cuda")
    with pytest.raises(b2.B2BfsError) as e:
        b2.search_noalloc(0, b2.NOALLOC_MAX_VERTICES + 1, 0, d.data_ptr(), d.data_ptr(), 0, out.data_ptr(), out.data_ptr())
    assert e.value.code == -4 and "exceeds the built-in scratch" in str(e.value)
    with pytest.raises(b2.B2BfsError) as e:
        b2.search_noalloc(0, 8, b2.NOALLOC_MAX_ENTRIES + 1, d.data_ptr(), d.data_ptr(), 0, out.data_ptr(), out.data_ptr())
    assert e.value.code == -4
    with pytest.raises(b2.B2BfsError):
        b2.search_noalloc(0, 8, 0, d.data_ptr(), d.data_ptr(), 9, out.data_ptr(), out.data_ptr())  # bad source


def test_noalloc_takes_no_device_memory_after_its_set_up_call(b2):
    """The promise to callers who cannot allocate: once the first (set-up) call on a device is behind
    us, a search leaves the amount of free device memory exactly where it was."""
    with b2.Bfs(0) as h:
        h.generate(0, 12, 8, 7)
        n = h.num_vertices
        rowptr, col, d_rowptr, d_col = _device_csr(b2, h)
        d_depth = torch.empty(n, dtype=torch.int32, device="cuda")
        d_parent = torch.empty(n, dtype=torch.int32, device="cuda")
        root = int(np.argmax(np.diff(rowptr)))
        call = lambda: b2.search_noalloc(0, n, len(col), d_rowptr.data_ptr(), d_col.data_ptr(), root,
                                         d_depth.data_ptr(), d_parent.data_ptr())
        call()  # set-up call (a no-op here when the tests above already ran)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(0)
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info(0)
        assert free_after == free_before
