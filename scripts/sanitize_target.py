"""Small searches through every single-GPU path of the library, for compute-sanitizer:
    compute-sanitizer --tool memcheck python scripts/sanitize_target.py
fused (direction-optimising, top-down, claim-only on/off), stepwise, the comparator engines, the batched
search, a directed graph, and the allocation-free entry point; every result is checked against the oracle."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import oracle
b2 = importlib.import_module("distributed-cuda-bfs_b200")

checked = 0
for kind, scale, ef in ((0, 10, 16), (1, 10, 8), (2, 8, 0)):
    n, s, d = oracle.generate_edges(kind, scale, ef, 1)
    g = oracle.build_csr(n, s, d)
    roots = [int(np.argmax(g.sizes)), int(np.flatnonzero(g.sizes > 0)[-1])]
    with b2.Bfs(0) as h:
        h.generate(kind, scale, ef, 1)
        for root in roots:
            want = oracle.bfs_serial(g, root)
            for engine in (b2.ENGINE_FUSED, b2.ENGINE_STEPWISE, b2.ENGINE_SCAN, b2.ENGINE_PULL):
                for mode in (b2.MODE_DIRECTION_OPT, b2.MODE_TOP_DOWN):
                    for gamma in (0, 2000):
                        h.set_option("engine", engine); h.set_option("mode", mode); h.set_option("gamma", gamma)
                        depth, parent, _ = h.run(root, want_parent=True)
                        assert oracle.check_output(depth, want) == -1
                        checked += 1
        h.set_option("engine", b2.ENGINE_FUSED); h.set_option("mode", b2.MODE_DIRECTION_OPT)
        depth, _ = h.run_batch(roots + roots[:1])
        for i, r in enumerate(roots + roots[:1]):
            assert oracle.check_output(depth[i], oracle.bfs_serial(g, r)) == -1
            checked += 1
        rowptr, col = h.csr()
        d_rowptr, d_col = torch.from_numpy(rowptr).cuda(), torch.from_numpy(col).cuda()
        d_depth = torch.empty(n, dtype=torch.int32, device="cuda")
        d_parent = torch.empty(n, dtype=torch.int32, device="cuda")
        for root in roots:
            b2.search_noalloc(0, n, len(col), d_rowptr.data_ptr(), d_col.data_ptr(), root, d_depth.data_ptr(), d_parent.data_ptr())
            assert oracle.check_output(d_depth.cpu().numpy(), oracle.bfs_serial(g, root)) == -1
            checked += 1
rng = np.random.default_rng(1)
n, m = 2000, 9000
s = rng.integers(0, n, m).astype(np.int32); d = rng.integers(0, n, m).astype(np.int32)
gd = oracle.build_csr_directed(n, s, d)
with b2.Bfs(0, directed=1) as h:
    h.load_edges(n, s, d)
    for root in (0, 5):
        depth, _ = h.run(root)
        assert oracle.check_output(depth, oracle.bfs_serial(gd, root)) == -1
        checked += 1
print(f"SANITIZE_TARGET_OK searches_checked={checked}")
