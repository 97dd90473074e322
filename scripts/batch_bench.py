"""64 roots as ONE batched search (b2bfs_run_batch) against 64 single searches, same roots, one GPU.
    python scripts/batch_bench.py <kind> <scale> [reps]      prints one JSON line"""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
b2 = importlib.import_module("distributed-cuda-bfs_b200")
kind, scale = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
h = b2.Bfs(0)
h.generate(kind, scale, 16, 1)
n = h.num_vertices
deg = h.row_degrees()
depth, _ = h.run(int(np.argmax(deg)))
cand = np.flatnonzero((deg > 0) & (depth != b2.INT_MAX))
roots = [int(r) for r in np.random.default_rng(2).choice(cand, 64, replace=False)]
for r in roots[:3]:
    h.run_device(r)
single = [h.run_device(r) for r in roots]
single_ms = sum(s["search_ms"] for s in single)
edges = sum(s["reached_entries"] / 2 for s in single)
h.run_batch(roots, to_host=False)
best = None
for _ in range(reps):
    _, st = h.run_batch(roots, to_host=False)
    best = st if best is None or st["search_ms"] < best["search_ms"] else best
assert best["reached_entries"] == [s["reached_entries"] for s in single]
print(json.dumps({"kind": kind, "scale": scale, "roots": 64, "single_total_ms": round(single_ms, 3),
                  "single_gteps_aggregate": round(edges / (single_ms * 1e-3) / 1e9, 1),
                  "batch_ms": round(best["search_ms"], 3), "batch_levels": best["levels"], "batch_launches": best["kernel_launches"],
                  "batch_gteps_aggregate": round(edges / (best["search_ms"] * 1e-3) / 1e9, 1),
                  "speedup": round(single_ms / best["search_ms"], 2)}))
h.close()
