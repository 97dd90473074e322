"""Generate tests/golden/ from the REFERENCE'S OWN host code (run in the build container only).

Every fixture is an edge-list file in the reference's format plus the outputs that the reference's
unmodified readGraphFromFile (bfs.cu:829-880) and bfsCPU (bfs.cu:923-945) produce for it, obtained
through oracle/_ref/libref_oracle.so (compiled from /root/reference/bfs.cu by oracle/Makefile).
/root/reference does not exist on the GPU box, so the vectors are committed.

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
import oracle  # noqa: E402


def emit(name, n, src, dst, roots):
    path = os.path.join(HERE, name + ".txt")
    with open(path, "w") as f:
        f.write(f"{n} {len(src)}\n")
        for u, v in zip(src, dst):
            f.write(f"{int(u)} {int(v)}\n")
    g = oracle.ref_load_file(path)  # the reference's loader
    out = {"n": g.n, "m": len(src), "num_edges_printed": g.num_edges,
           "offsets": g.offsets.tolist() if g.n <= 64 else None,
           "sizes": g.sizes.tolist() if g.n <= 64 else None,
           "adj": g.adj.tolist() if g.n <= 64 else None,
           "csr_checksum": int((g.adj.astype(np.int64) * (np.arange(g.adj.size) % 1000003 + 1)).sum() % (2**61 - 1)),
           "searches": []}
    for r in roots:
        depth, parent, _ = oracle.ref_bfs_cpu(int(r), g.n, want_parent=True)  # the reference's bfsCPU
        out["searches"].append({"root": int(r), "depth": depth.tolist(), "parent_edge": parent.tolist()})
    with open(os.path.join(HERE, name + ".json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(name, "n", g.n, "E", g.num_edges)


def main():
    # the survey's vector: duplicate edge, self-loop, two isolated vertices
    emit("tiny", 6, [0, 1, 2, 3, 0], [1, 2, 3, 3, 1], [2, 0, 4])
    rng = np.random.default_rng(20261106)
    # sparse random multigraph with several components, odd n
    n = 97
    s = rng.integers(0, n, 110); d = rng.integers(0, n, 110)
    emit("random97", n, s, d, [0, 5, 96])
    # path (max diameter) and star (one hub)
    emit("path64", 64, list(range(63)), list(range(1, 64)), [0, 31, 63])
    emit("star65", 65, [0] * 64, list(range(1, 65)), [0, 7])
    # RMAT scale-8 from our generator: hubs, duplicates, self-loops, isolated vertices
    nn, s, d = oracle.generate_edges(oracle.GEN_RMAT, 8, 16, 1)
    emit("rmat8", nn, s, d, [2, 0, 255])
    # Kronecker (scrambled labels) scale-8 and a 16x16 grid
    nn, s, d = oracle.generate_edges(oracle.GEN_KRON, 8, 16, 1)
    deg = np.bincount(np.concatenate([s, d]), minlength=nn)
    emit("kron8", nn, s, d, [int(np.argmax(deg)), int(np.flatnonzero(deg)[3])])
    nn, s, d = oracle.generate_edges(oracle.GEN_GRID, 8, 0, 0)
    emit("grid16x16", nn, s, d, [0, 136, 255])
    # no edges at all
    emit("empty5", 5, [], [], [3])


if __name__ == "__main__":
    main()
