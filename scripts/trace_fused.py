"""Per-phase timeline of a few fused searches (device %globaltimer marks of CTA 0)."""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
b2 = importlib.import_module("distributed-cuda-bfs_b200")
kind, scale, nroots = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
h = b2.Bfs(0)
for a in sys.argv[4:]:
    k, v = a.split("="); h.set_option(k, int(v))
h.generate(kind, scale, 16, 1)
rowptr, _ = h.csr(); deg = np.diff(rowptr)
roots = np.random.default_rng(2).choice(np.flatnonzero(deg > 0), nroots + 2, replace=False)
names = {-1: "start", 0: "reset", 1: "td", 2: "bu", 3: "q2b", 4: "b2q", 5: "grp", 6: "claim", 7: "cluster"}
for r in roots:
    st = h.run_device(int(r))
    tr = h.last_trace()
    prev = 0.0; parts = []
    for lv, ph, us in tr[:40]:
        parts.append(f"{names[ph]}@{lv}:{us - prev:.0f}"); prev = us
    print(f"root {r} deg {deg[r]} search_ms {st['search_ms']:.3f} levels {st['levels']} | " + " ".join(parts) + (f" ... ({len(tr)} marks)" if len(tr) > 40 else ""))
