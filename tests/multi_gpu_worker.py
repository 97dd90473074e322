"""Worker for the multi-GPU parity test: launched by torchrun, one process per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tests/multi_gpu_worker.py

Every rank builds its partition of the same graph, runs the same roots through the C-ABI (NCCL
exchange inside the library) and checks the gathered depth/parent arrays against the oracle.
torch.distributed (gloo) is only the side channel that carries the 128-byte NCCL id.
"""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from conftest import golden_names, load_golden  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    b2 = importlib.import_module("distributed-cuda-bfs_b200")
    ident = [b2.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ident, src=0)
    h = b2.Bfs(local, rank=rank, world=world, nccl_id=ident[0])
    checked = 0

    def check_slices(g, root, want):
        """b2bfs_run_slice: every rank receives only its own rows; together they are the whole array."""
        first, cnt = h.local_range()
        d = np.full(max(cnt, 1), -5, np.int32)[:cnt]
        q = np.full(max(cnt, 1), -5, np.int32)[:cnt]
        h.run_slice(int(root), d, q)
        pieces = [None] * world
        dist.all_gather_object(pieces, (first, d, q))
        depth = np.concatenate([x[1] for x in sorted(pieces, key=lambda x: x[0])])
        parent = np.concatenate([x[2] for x in sorted(pieces, key=lambda x: x[0])])
        assert depth.shape[0] == g.n and oracle.check_output(depth, want) == -1
        assert oracle.validate_tree(g, int(root), depth, parent) == (0, -1)

    def check(g, roots):
        nonlocal checked
        for root in roots:
            want = oracle.bfs_serial(g, int(root))
            for mode, engine in ((b2.MODE_DIRECTION_OPT, 0), (b2.MODE_TOP_DOWN, 0), (b2.MODE_DIRECTION_OPT, 1), (b2.MODE_TOP_DOWN, 1)):
                h.set_option("mode", mode)
                h.set_option("multi_engine", engine)  # 0: peer-memory fused kernel, 1: NCCL per level
                depth, parent, st = h.run(int(root), want_parent=True)
                bad = oracle.check_output(depth, want)
                assert bad == -1, f"rank {rank} root {root} mode {mode}: vertex {bad} got {depth[bad]} want {want[bad]}"
                assert oracle.validate_tree(g, int(root), depth, parent) == (0, -1)
                vr, dr = oracle.reach_stats(g, want)
                assert (st["reached_vertices"], st["reached_entries"]) == (vr, dr), (st, vr, dr)
                assert st["levels"] == int(want[want != oracle.INT_MAX].max()) + 1
                checked += 1
        h.set_option("mode", b2.MODE_DIRECTION_OPT)
        h.set_option("multi_engine", 0)
        check_slices(g, roots[0], oracle.bfs_serial(g, int(roots[0])))
        checked += 1

    # synthetic graphs, generated on every rank, each keeping its own rows
    for kind, scale, ef, seed in ((1, 16, 16, 1), (0, 14, 16, 3), (2, 12, 0, 0), (1, 9, 4, 5)):
        n, s, d = oracle.generate_edges(kind, scale, ef, seed)
        g = oracle.build_csr(n, s, d)
        h.generate(kind, scale, ef, seed)
        assert h.num_vertices == n
        assert np.array_equal(h.degrees(), g.sizes)           # allgathered edgesSize[]
        rdeg = h.row_degrees()
        assert rdeg.shape[0] == n and (rdeg > 0).sum() == (np.array([len(set(g.adj[g.offsets[v]:g.offsets[v] + g.sizes[v]]) - {v}) for v in range(min(n, 2000))]) > 0).sum() + (rdeg[2000:] > 0).sum()
        rng = np.random.default_rng(seed)
        roots = list(rng.choice(np.flatnonzero(g.sizes > 0), 4, replace=False)) + [0, n - 1]
        check(g, roots)
    # a graph large enough for every level kind (marking top-down, several bottom-up levels, tail levels)
    # to carry real traffic between the ranks; depths pinned by the edge rules (no host CSR needed)
    n, s, d = oracle.generate_edges(1, 20, 16, 2)
    h.generate(1, 20, 16, 2)
    rdeg = h.row_degrees()
    for root in np.random.default_rng(7).choice(np.flatnonzero(rdeg > 0), 6, replace=False):
        for engine in (0, 1):
            h.set_option("multi_engine", engine)
            depth, _ = h.run(int(root))
            bad = oracle.check_depth_by_edges(n, s, d, int(root), depth)
            assert bad == -1, f"rank {rank} kron-20 root {root} engine {engine}: vertex {bad} depth {depth[bad]}"
            checked += 1
    h.set_option("multi_engine", 0)
    del s, d
    # the reference's golden vectors through the text loader (every rank parses the file, as every
    # MPI rank of the reference does, bfs_mpi.cu:815)
    for name in golden_names():
        path, gold = load_golden(name)
        h.load_edge_file(path)
        for s in gold["searches"]:
            depth, _ = h.run(s["root"])
            assert oracle.check_output(depth, np.array(s["depth"], np.int32)) == -1, (name, s["root"])
            checked += 1
    peer = h.get_option("multi_engine") == 0
    h.close()
    dist.barrier()
    if rank == 0:
        print(f"MULTI_GPU_OK world={world} searches_checked={checked} peer_memory_engine={'yes' if peer else 'no (NCCL fallback)'}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
