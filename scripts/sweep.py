"""Option sweep on one GPU: harmonic-mean GTEPS over the same roots for every value of one or more
options.    python scripts/sweep.py <kind> <scale> <nroots> gamma=0,8,24 alpha=14,28
Roots: seeded sample of vertices with an edge inside the largest component (bench.py's rule)."""
import importlib, itertools, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
b2 = importlib.import_module("distributed-cuda-bfs_b200")
kind, scale, nroots = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
axes = [(a.split("=")[0], [int(x) for x in a.split("=")[1].split(",")]) for a in sys.argv[4:]]
h = b2.Bfs(0)
h.generate(kind, scale, 16, 1)
n = h.num_vertices
deg = h.row_degrees()
depth, _ = h.run(int(np.argmax(deg)))
cand = np.flatnonzero((deg > 0) & (depth != b2.INT_MAX))
roots = [int(r) for r in np.random.default_rng(2).choice(cand, nroots + 3, replace=False)]
for combo in itertools.product(*[v for _, v in axes]):
    for (k, _), v in zip(axes, combo):
        h.set_option(k, v)
    for r in roots[:3]:
        h.run_device(r)
    st = [h.run_device(r) for r in roots[3:]]
    teps = np.array([s["reached_entries"] / 2 / (s["search_ms"] * 1e-3) / 1e9 for s in st])
    ms = np.array([s["search_ms"] for s in st])
    print(json.dumps({"kind": kind, "scale": scale, **{k: v for (k, _), v in zip(axes, combo)},
                      "gteps_hmean": round(float(len(teps) / np.sum(1 / teps)), 1), "ms_mean": round(float(ms.mean()), 4),
                      "ms_min": round(float(ms.min()), 4), "ms_max": round(float(ms.max()), 4),
                      "bu_levels_mean": float(np.mean([s["bu_levels"] for s in st])),
                      "levels_mean": float(np.mean([s["levels"] for s in st]))}), flush=True)
h.close()
