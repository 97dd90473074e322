"""CPU test (gloo, world_size 2 and 3) of the multi-rank PROTOCOL that csrc/multi.cu drives on GPUs:
1-D partition with `part` rounded to 128 rows, top-down with a `seen` filter + (vertex, parent)
all-to-all-v to owners, bottom-up against an allgathered frontier bitmap, one allgathered stats
record per exchange from which every rank takes the same direction/termination decisions.

The per-rank "kernels" here are numpy restatements (test infrastructure); what is under test is the
exchange logic and the decision rule, i.e. that the protocol yields the serial depths of the
reference's bfsCPU (oracle) and valid parents on every rank — the property the reference's own
2-rank protocol is trusted for (bfs_mpi.cu:581-634) without ever being tested.
"""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

INT_MAX = oracle.INT_MAX
ALPHA, BETA = 14, 24


def partition(n, k):
    part = ((n + k - 1) // k + 127) // 128 * 128
    return part, [(min(n, r * part), min(n, (r + 1) * part)) for r in range(k)]


def rank_search(rank, world, n, src_e, dst_e, root, mode_do):
    """One rank of the protocol; returns (depth[n], parent[n]) after the final allgather."""
    part, ranges = partition(n, world)
    lo, hi = ranges[rank]
    nloc = hi - lo
    # this rank's rows of the symmetrised, deduplicated CSR (global ids in col)
    u = np.concatenate([src_e, dst_e]); v = np.concatenate([dst_e, src_e])
    keep = (u >= lo) & (u < hi) & (u != v)
    pairs = np.unique(np.stack([u[keep], v[keep]], 1), axis=0) if keep.any() else np.zeros((0, 2), np.int64)
    rowptr = np.zeros(nloc + 1, np.int64)
    np.add.at(rowptr, pairs[:, 0] - lo + 1, 1)
    rowptr = np.cumsum(rowptr)
    col = pairs[:, 1]
    depth = np.full(part, INT_MAX, np.int32); parent = np.full(part, -1, np.int32)
    visited = np.zeros(part, bool); seen = np.zeros(world * part, bool)
    frontier, src_deg = [], 0
    if root // part == rank:
        depth[root - lo] = 0; parent[root - lo] = root; visited[root - lo] = True
        frontier = [root - lo]
        src_deg = int(rowptr[root - lo + 1] - rowptr[root - lo])
    t = torch.tensor([len(col), src_deg], dtype=torch.int64)   # load-time allreduce + first stats record
    dist.all_reduce(t)
    unexplored = int(t[0] - t[1])                                # m_u: the same number on every rank
    level, mode, prev_nf, gfrontier = 0, 0, 1, None

    def allgather_records(rec):
        out = [None] * world
        dist.all_gather_object(out, rec)
        return out

    while True:
        if mode == 0:
            send = [[] for _ in range(world)]
            nxt, mf = [], 0
            for ul in frontier:
                for x in col[rowptr[ul]:rowptr[ul + 1]]:
                    x = int(x)
                    if lo <= x < hi:
                        if not visited[x - lo]:
                            visited[x - lo] = True; depth[x - lo] = level + 1; parent[x - lo] = ul + lo
                            nxt.append(x - lo); mf += int(rowptr[x - lo + 1] - rowptr[x - lo])
                    elif not seen[x]:
                        seen[x] = True
                        send[x // part].append((x, ul + lo))
            recs = allgather_records([len(s) for s in send])          # stats record #1: per-peer counts
            recv = None
            for d in range(world):                                       # all-to-all-v as k gathers (gloo)
                got = [None] * world if rank == d else None
                dist.gather_object(send[d], got, dst=d)
                if rank == d:
                    recv = got
            for src_rank in range(world):
                if src_rank == rank:
                    continue
                assert len(recv[src_rank]) == recs[src_rank][rank]     # counts announced == pairs delivered
                for x, p in recv[src_rank]:
                    if not visited[x - lo]:
                        visited[x - lo] = True; depth[x - lo] = level + 1; parent[x - lo] = p
                        nxt.append(x - lo); mf += int(rowptr[x - lo + 1] - rowptr[x - lo])
            recs = allgather_records((len(nxt), mf))                     # stats record #2
            nf = sum(r[0] for r in recs); mf_all = sum(r[1] for r in recs)
            frontier = nxt
        else:
            found = np.zeros(part, bool)
            for vl in np.flatnonzero(~visited[:nloc]):
                for x in col[rowptr[vl]:rowptr[vl + 1]]:
                    if gfrontier[int(x)]:
                        visited[vl] = True; depth[vl] = level + 1; parent[vl] = int(x); found[vl] = True
                        break
            slices = [None] * world
            dist.all_gather_object(slices, found)
            gfrontier = np.concatenate(slices)
            recs = allgather_records(int(found.sum()))
            nf, mf_all = sum(recs), 0
            frontier = list(np.flatnonzero(found))
        unexplored = max(0, unexplored - mf_all)
        level += 1
        if nf == 0:
            break
        if mode == 0:
            if mode_do and mf_all > unexplored // ALPHA and nf >= prev_nf:
                mode = 1
                mine = np.zeros(part, bool); mine[frontier] = True
                slices = [None] * world
                dist.all_gather_object(slices, mine)
                gfrontier = np.concatenate(slices)
        elif nf < n // BETA and nf < prev_nf:
            mode = 0
        prev_nf = nf
    alld, allp = [None] * world, [None] * world
    dist.all_gather_object(alld, depth); dist.all_gather_object(allp, parent)
    return np.concatenate(alld)[:n], np.concatenate(allp)[:n], level


def worker(rank, world, port, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        checked = 0
        for kind, scale, ef, seed in ((oracle.GEN_KRON, 9, 8, 1), (oracle.GEN_RMAT, 8, 16, 2), (oracle.GEN_GRID, 8, 0, 0)):
            n, s, d = oracle.generate_edges(kind, scale, ef, seed)
            g = oracle.build_csr(n, s, d)
            roots = [int(np.argmax(g.sizes)), n - 1, int(np.flatnonzero(g.sizes > 0)[3])]
            for root in roots:
                want = oracle.bfs_serial(g, root)
                for mode_do in (True, False):
                    depth, parent, levels = rank_search(rank, world, n, s.astype(np.int64), d.astype(np.int64), root, mode_do)
                    assert oracle.check_output(depth, want) == -1
                    assert oracle.validate_tree(g, root, depth, parent) == (0, -1)
                    assert levels == int(want[want != INT_MAX].max()) + 1
                    checked += 1
        results[rank] = checked
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,port", [(2, 29731), (3, 29733)])
def test_protocol_yields_serial_depths_on_every_rank(world, port):
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(worker, args=(world, port, results), nprocs=world, join=True)
    assert len(results) == world and all(v == 18 for v in results.values())


def test_partition_arithmetic():
    for n, k in ((1, 2), (97, 2), (4096, 8), (1 << 20, 3), (129, 8)):
        part, ranges = partition(n, k)
        assert part % 128 == 0 and part * k >= n
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
        assert all(0 <= hi - lo <= part for lo, hi in ranges)
        for v in (0, n - 1, n // 2):
            r = v // part
            assert ranges[r][0] <= v < ranges[r][1]


# ---------------------------------------------------------------------------------------------
# construction by owner exchange (api.cu build_partitioned / csr_build.cu owner_keys_kernel): every
# rank holds m/k of the tuples, buckets the two entries of each tuple by the rank that owns their
# row, the buckets are exchanged, and a rank keeps exactly the rows of the serial CSR that it owns
# ---------------------------------------------------------------------------------------------
def tuple_share(m, rank, world):
    first = m * rank // world
    return first, m * (rank + 1) // world - first


def exchange_worker(rank, world, port, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ok = 0
        for kind, scale, ef, seed in ((oracle.GEN_KRON, 10, 8, 1), (oracle.GEN_RMAT, 9, 16, 2), (oracle.GEN_GRID, 8, 0, 0),
                                      (oracle.GEN_RMAT, 3, 1, 5)):  # the last one: fewer tuples than some ranks' share needs
            n, s, d = oracle.generate_edges(kind, scale, ef, seed)
            g = oracle.build_csr(n, s, d)  # the reference's rows (multigraph), for every vertex
            m = len(s)
            part, ranges = partition(n, world)
            lo, hi = ranges[rank]
            first, cnt = tuple_share(m, rank, world)
            ls, ld = s[first:first + cnt].astype(np.int64), d[first:first + cnt].astype(np.int64)
            # degree of every vertex from the local tuples, summed over ranks (the hub-first class needs it)
            deg = torch.from_numpy(np.bincount(ls, minlength=n) + np.bincount(ld, minlength=n))
            dist.all_reduce(deg)
            assert np.array_equal(deg.numpy(), g.sizes.astype(np.int64))
            # both entries of every local tuple, bucketed by owner of the row
            rows = np.concatenate([ls, ld]); nbrs = np.concatenate([ld, ls])
            owner = rows // part
            send = [np.stack([rows[owner == r], nbrs[owner == r]], 1) for r in range(world)]
            counts = [None] * world
            dist.all_gather_object(counts, [len(x) for x in send])      # the k x k count matrix
            assert sum(counts[rank]) == 2 * cnt
            recv = [None] * world
            for r in range(world):                                      # all-to-all-v of the buckets
                out = [None] * world if r == rank else None
                dist.gather_object(send[r], out, dst=r)
                if r == rank:
                    recv = out
            mine = np.concatenate([x.reshape(-1, 2) for x in recv]) if recv else np.zeros((0, 2), np.int64)
            assert len(mine) == sum(c[rank] for c in counts)
            assert ((mine[:, 0] >= lo) & (mine[:, 0] < hi)).all() if len(mine) else True
            # what arrived is exactly this rank's rows of the serial CSR, as multisets
            for v in range(lo, hi):
                got = np.sort(mine[mine[:, 0] == v, 1])
                want = np.sort(g.adj[g.offsets[v]:g.offsets[v] + g.sizes[v]])
                assert np.array_equal(got, want), (rank, v)
            ok += 1
        results[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,port", [(2, 29741), (3, 29743)])
def test_owner_exchange_builds_each_ranks_rows(world, port):
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(exchange_worker, args=(world, port, results), nprocs=world, join=True)
    assert len(results) == world and all(v == 4 for v in results.values())


def test_tuple_shares_cover_every_tuple_once():
    for m, k in ((0, 2), (1, 8), (5, 8), (1000, 3), (1 << 33, 8)):
        shares = [tuple_share(m, r, k) for r in range(k)]
        assert shares[0][0] == 0 and sum(c for _, c in shares) == m
        assert all(a[0] + a[1] == b[0] for a, b in zip(shares, shares[1:]))
