"""Generate tests/golden/readgraph/ from the REFERENCE'S OWN second reader, readGraph (bfs.cu:882-920),
and its bfsCPU (bfs.cu:923-945), through oracle/_ref/libref_oracle.so (build container only; the
vectors are committed because /root/reference does not exist on the GPU box).

  stdin form   "n m" + m pairs, every pair the single arc u -> v (bfs.cu:901): a DIRECTED graph
  random form  srand(12345); u = rand() % n; v = rand() % n per edge, both directions (bfs.cu:891-906)

    python tests/golden/readgraph/make_golden_readgraph.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..", "..")))
import oracle  # noqa: E402


def describe(g, roots):
    out = {"n": g.n, "num_edges": g.num_edges,
           "offsets": g.offsets.tolist() if g.n <= 64 else None,
           "sizes": g.sizes.tolist() if g.n <= 64 else None,
           "adj": g.adj.tolist() if g.n <= 64 else None,
           "csr_checksum": int((g.adj.astype(np.int64) * (np.arange(g.adj.size) % 1000003 + 1)).sum() % (2**61 - 1)),
           "searches": []}
    for r in roots:
        depth, _ = oracle.ref_bfs_cpu(int(r), g.n)
        out["searches"].append({"root": int(r), "depth": depth.tolist()})
    return out


def main():
    rng = np.random.default_rng(20261107)
    # stdin form: arcs with a duplicate, a self-loop, a sink (vertex 39 has no out-arc) and a source-only vertex
    n, m = 40, 90
    s = rng.integers(0, n - 1, m)
    d = rng.integers(0, n, m)
    s[0], d[0] = 3, 3
    s[1], d[1] = s[2], d[2]
    path = os.path.join(HERE, "stdin40.txt")
    with open(path, "w") as f:
        f.write(f"{n} {m}\n")
        for u, v in zip(s, d):
            f.write(f"{int(u)} {int(v)}\n")
    g = oracle.ref_read_graph_stdin(path)
    out = describe(g, [0, 3, int(s[5]), 39])
    out["m"] = m
    json.dump(out, open(os.path.join(HERE, "stdin40.json"), "w"), separators=(",", ":"))
    print("stdin40", g.n, g.num_edges)
    # random form
    for n, m, roots in ((50, 120, [0, 7, 49]), (3000, 9000, [0, 1234, 2999])):
        g = oracle.ref_read_graph_random(n, m)
        out = describe(g, roots)
        out["m"] = m
        json.dump(out, open(os.path.join(HERE, f"random_{n}_{m}.json"), "w"), separators=(",", ":"))
        print("random", n, m, g.num_edges)


if __name__ == "__main__":
    main()
