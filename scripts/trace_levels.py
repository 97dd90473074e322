"""Per-level trace of a few searches with the stepwise engine (B2BFS_TRACE=1 prints to stderr)."""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
b2 = importlib.import_module("distributed-cuda-bfs_b200")
kind, scale, nroots = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
h = b2.Bfs(0)
for a in sys.argv[4:]:
    k, v = a.split("=")
    h.set_option(k, int(v))
h.generate(kind, scale, 16, 1)
rowptr, _ = h.csr()
deg = np.diff(rowptr)
roots = np.random.default_rng(2).choice(np.flatnonzero(deg > 0), nroots, replace=False)
h.set_option("engine", b2.ENGINE_STEPWISE)
print("n", h.num_vertices, "entries", h.csr_entries, "nonisolated", int((deg > 0).sum()), "deg>=32", int((deg >= 32).sum()),
      "deg>=4096", int((deg >= 4096).sum()), "maxdeg", int(deg.max()), file=sys.stderr)
for r in roots:
    print(f"--- root {r} deg {deg[r]}", file=sys.stderr)
    st = h.run_device(int(r))
    print("   ", st, file=sys.stderr)
