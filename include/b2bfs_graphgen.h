/*
 * b2bfs_graphgen.h — definition of the synthetic inputs (RMAT, Graph500-style Kronecker, 2-D grid).
 *
 * The reference ships no graph and no generator that `main` calls (its `readGraph`, bfs.cu:882-920,
 * is dead code using rand()); SURVEY.md §8(d) therefore defines the workloads.  This header IS that
 * definition: edge `i` of a graph is a pure function of (kind, scale, edgefactor, seed, i), written in
 * integer arithmetic only, so the host build (gcc, used by oracle/) and the device build (nvcc, used
 * by csrc/gen.cu) produce bit-identical edge lists.  tests/test_gpu_generators.py checks that.
 *
 * Edge ids are 0-based int32, as in the reference's file format (bfs.cu:851-861).
 */
#ifndef B2BFS_GRAPHGEN_H
#define B2BFS_GRAPHGEN_H

#include <stdint.h>

#ifdef __CUDACC__
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD static inline
#endif

#define B2BFS_GEN_RMAT 0 /* RMAT (0.57,0.19,0.19,0.05), no label permutation: low ids are hubs   */
#define B2BFS_GEN_KRON 1 /* same initiator + bijective vertex-label scramble (Graph500 style)    */
#define B2BFS_GEN_GRID 2 /* 2-D 4-neighbour grid, side = 2^(scale/2) x 2^(scale - scale/2)       */

/* splitmix64 finaliser */
B2_HD uint64_t b2_mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
  x ^= x >> 27; x *= 0x94D049BB133111EBULL;
  x ^= x >> 31;
  return x;
}

/* counter-based stream: word k of edge i under seed */
B2_HD uint64_t b2_rand64(uint64_t seed, uint64_t i, uint32_t k) {
  uint64_t x = b2_mix64(seed * 0x9E3779B97F4A7C15ULL + 0x632BE59BD9B4E019ULL);
  x = b2_mix64(x ^ (i * 0xD1342543DE82EF95ULL + 0x2545F4914F6CDD1DULL));
  return b2_mix64(x + (uint64_t)k * 0x9E3779B97F4A7C15ULL);
}

/* quadrant thresholds on a 16-bit uniform draw: a=0.57, a+b=0.76, a+b+c=0.95 */
#define B2_RMAT_A   37356u
#define B2_RMAT_AB  49807u
#define B2_RMAT_ABC 62259u

B2_HD void b2_rmat_edge(int scale, uint64_t seed, uint64_t i, uint32_t* u_out, uint32_t* v_out) {
  uint32_t u = 0, v = 0;
  uint64_t r = 0;
  for (int l = 0; l < scale; ++l) {
    if ((l & 3) == 0) r = b2_rand64(seed, i, (uint32_t)(l >> 2));
    uint32_t d = (uint32_t)(r & 0xFFFFu);
    r >>= 16;
    uint32_t ub = d >= B2_RMAT_AB;                                  /* lower half of the matrix   */
    uint32_t vb = (d >= B2_RMAT_A && d < B2_RMAT_AB) || d >= B2_RMAT_ABC; /* right half            */
    u = (u << 1) | ub;
    v = (v << 1) | vb;
  }
  *u_out = u;
  *v_out = v;
}

B2_HD uint32_t b2_bitrev(uint32_t x, int bits) {
  uint32_t r = 0;
  for (int b = 0; b < bits; ++b) { r = (r << 1) | (x & 1u); x >>= 1; }
  return r;
}

/* bijection on [0, 2^scale): odd multiplies, xor-shifts and a bit reversal are each invertible */
B2_HD uint32_t b2_scramble(uint32_t x, int scale, uint64_t seed) {
  if (scale <= 1) return x;
  const uint32_t mask = (scale >= 32) ? 0xFFFFFFFFu : ((1u << scale) - 1u);
  const uint64_t k = b2_mix64(seed ^ 0xA0761D6478BD642FULL);
  const uint32_t c1 = (uint32_t)(k) | 1u, c2 = (uint32_t)(k >> 32) | 1u, c3 = (uint32_t)(b2_mix64(k) >> 16) | 1u;
  const int s1 = (scale + 1) / 2, s2 = (scale + 2) / 3;
  x = (x * c1) & mask; x ^= x >> s1;
  x = (x * c2) & mask; x ^= x >> s2;
  x = b2_bitrev(x, scale);
  x = (x * c3) & mask; x ^= x >> s1;
  x = (x * c1) & mask;
  return x;
}

/* number of vertices / edge tuples of a generated graph */
B2_HD int64_t b2_gen_num_vertices(int kind, int scale) {
  (void)kind;
  return (int64_t)1 << scale;
}
B2_HD int64_t b2_gen_num_edges(int kind, int scale, int edgefactor) {
  if (kind == B2BFS_GEN_GRID) {
    const int64_t w = (int64_t)1 << (scale / 2), h = (int64_t)1 << (scale - scale / 2);
    return h * (w - 1) + w * (h - 1);
  }
  return ((int64_t)edgefactor) << scale;
}

/* edge tuple i (0 <= i < b2_gen_num_edges) */
B2_HD void b2_gen_edge(int kind, int scale, int edgefactor, uint64_t seed, uint64_t i,
                       uint32_t* u_out, uint32_t* v_out) {
  (void)edgefactor;
  if (kind == B2BFS_GEN_GRID) {
    /* id = y*w + x ; first h*(w-1) tuples are (id, id+1), the rest are (id, id+w) */
    const uint64_t w = (uint64_t)1 << (scale / 2), h = (uint64_t)1 << (scale - scale / 2);
    const uint64_t nh = h * (w - 1);
    if (i < nh) {
      const uint64_t y = i / (w - 1), x = i % (w - 1);
      *u_out = (uint32_t)(y * w + x);
      *v_out = (uint32_t)(y * w + x + 1);
    } else {
      const uint64_t j = i - nh; /* y in [0,h-1), x in [0,w) */
      *u_out = (uint32_t)j;
      *v_out = (uint32_t)(j + w);
    }
    return;
  }
  uint32_t u, v;
  b2_rmat_edge(scale, seed, i, &u, &v);
  if (kind == B2BFS_GEN_KRON) {
    u = b2_scramble(u, scale, seed);
    v = b2_scramble(v, scale, seed);
  }
  *u_out = u;
  *v_out = v;
}

#endif /* B2BFS_GRAPHGEN_H */
