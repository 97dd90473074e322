/*
 * oracle/bfs_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement, in plain C, of the reference's hot path:
 *   edge-list text -> symmetrised CSR -> serial BFS depth array.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's shared library; the product library (libb2bfs.so) never does.
 *
 * Parity status: PINNED.  Every function here is checked in tests/test_oracle.py against the
 * reference's own host code (readGraphFromFile + bfsCPU, /root/reference/bfs.cu) compiled unmodified
 * into oracle/_ref/libref_oracle.so, and against the golden vectors under tests/golden/ that were
 * generated from that library by tests/golden/make_golden.py.
 *
 * All citations are into /root/reference/.
 */
#include <limits.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/b2bfs_graphgen.h"

#define ORACLE_OK 0
#define ORACLE_ERR_IO -1
#define ORACLE_ERR_FORMAT -2
#define ORACLE_ERR_RANGE -3
#define ORACLE_ERR_NOMEM -4

/* ---------------------------------------------------------------------------------------------
 * Edge-list text reader.  Follows readGraphFromFile, bfs.cu:829-863: `f >> n >> m`, then m pairs
 * `f >> u >> v` of whitespace-separated decimal ints, 0-based.  The reference does not check that
 * the file holds m pairs or that ids are < n (defect D4 in SURVEY.md); the oracle reports those
 * cases as errors instead of reading uninitialised memory.
 * Two-call protocol: first with src==dst==NULL to learn n and m, then with buffers of m ints.
 * ------------------------------------------------------------------------------------------- */
int oracle_read_edge_file(const char* path, int64_t* n_out, int64_t* m_out, int32_t* src, int32_t* dst) {
  FILE* f = fopen(path, "r");
  if (!f) return ORACLE_ERR_IO;
  long long n, m;
  if (fscanf(f, "%lld %lld", &n, &m) != 2 || n < 0 || m < 0) { fclose(f); return ORACLE_ERR_FORMAT; }
  *n_out = n;
  *m_out = m;
  if (!src || !dst) { fclose(f); return ORACLE_OK; }
  for (long long i = 0; i < m; ++i) {
    long long u, v;
    if (fscanf(f, "%lld %lld", &u, &v) != 2) { fclose(f); return ORACLE_ERR_FORMAT; }
    if (u < 0 || v < 0 || u >= n || v >= n) { fclose(f); return ORACLE_ERR_RANGE; }
    src[i] = (int32_t)u;
    dst[i] = (int32_t)v;
  }
  fclose(f);
  return ORACLE_OK;
}

/* Writer for the same format (README.md:8-12): "n m" then one "u v" per line. */
int oracle_write_edge_file(const char* path, int64_t n, int64_t m, const int32_t* src, const int32_t* dst) {
  FILE* f = fopen(path, "w");
  if (!f) return ORACLE_ERR_IO;
  fprintf(f, "%lld %lld\n", (long long)n, (long long)m);
  for (int64_t i = 0; i < m; ++i) fprintf(f, "%d %d\n", src[i], dst[i]);
  return fclose(f) == 0 ? ORACLE_OK : ORACLE_ERR_IO;
}

/* ---------------------------------------------------------------------------------------------
 * CSR construction.  Follows bfs.cu:851-875: for every tuple in file order push v onto u's list
 * and u onto v's list (duplicates kept; a self-loop u==u is pushed twice), then flatten rows in
 * vertex order into adjacencyList with edgesOffset[v] / edgesSize[v].  A counting pass + cursor
 * scatter in file order reproduces the per-row push_back order exactly.
 * offsets has n entries (like edgesOffset), sizes has n entries, adj has 2*m entries.
 * 64-bit offsets so the restatement also covers E >= 2^31 (the reference overflows there, D3).
 * ------------------------------------------------------------------------------------------- */
int oracle_build_csr(int64_t n, int64_t m, const int32_t* src, const int32_t* dst, int64_t* offsets,
                     int32_t* sizes, int32_t* adj) {
  memset(sizes, 0, (size_t)n * sizeof(int32_t));
  for (int64_t i = 0; i < m; ++i) {
    if (src[i] < 0 || dst[i] < 0 || src[i] >= n || dst[i] >= n) return ORACLE_ERR_RANGE;
    sizes[src[i]]++;
    sizes[dst[i]]++;
  }
  int64_t run = 0;
  for (int64_t v = 0; v < n; ++v) { offsets[v] = run; run += sizes[v]; }
  int64_t* cur = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
  if (!cur) return ORACLE_ERR_NOMEM;
  memcpy(cur, offsets, (size_t)n * sizeof(int64_t));
  for (int64_t i = 0; i < m; ++i) {
    const int32_t u = src[i], v = dst[i];
    adj[cur[u]++] = v; /* bfs.cu:860 */
    adj[cur[v]++] = u; /* bfs.cu:861 */
  }
  free(cur);
  return ORACLE_OK;
}

/* ---------------------------------------------------------------------------------------------
 * Serial BFS.  Follows bfsCPU, bfs.cu:923-945: FIFO queue, distance[start]=0, parent[start]=start,
 * first discovery wins, parent[v] = index i of the discovering entry in adjacencyList (NOT a vertex
 * id).  depth/parent must be pre-filled by the caller with INT_MAX as main does (bfs.cu:793-794);
 * oracle_bfs_serial does that fill itself for convenience.  parent_edge is int64 so the restatement
 * holds beyond 2^31 entries; below that it equals the reference's int value.
 * Returns the number of vertices reached.
 * ------------------------------------------------------------------------------------------- */
int64_t oracle_bfs_serial(int64_t n, const int64_t* offsets, const int32_t* sizes, const int32_t* adj,
                          int64_t start, int32_t* depth, int64_t* parent_edge) {
  for (int64_t v = 0; v < n; ++v) depth[v] = INT_MAX;
  if (parent_edge) for (int64_t v = 0; v < n; ++v) parent_edge[v] = INT_MAX;
  if (start < 0 || start >= n) return ORACLE_ERR_RANGE;
  uint8_t* visited = (uint8_t*)calloc((size_t)n, 1);
  int32_t* q = (int32_t*)malloc((size_t)n * sizeof(int32_t));
  if (!visited || !q) { free(visited); free(q); return ORACLE_ERR_NOMEM; }
  int64_t head = 0, tail = 0;
  depth[start] = 0;
  if (parent_edge) parent_edge[start] = start; /* bfs.cu:926 */
  visited[start] = 1;
  q[tail++] = (int32_t)start;
  while (head < tail) {
    const int32_t u = q[head++];
    const int64_t b = offsets[u], e = b + sizes[u];
    for (int64_t i = b; i < e; ++i) { /* bfs.cu:935 */
      const int32_t v = adj[i];
      if (!visited[v]) {
        visited[v] = 1;
        depth[v] = depth[u] + 1;
        if (parent_edge) parent_edge[v] = i; /* bfs.cu:940 */
        q[tail++] = v;
      }
    }
  }
  free(visited);
  free(q);
  return tail;
}

/* ---------------------------------------------------------------------------------------------
 * K-device emulation of the reference's multi-GPU protocol — runCudaQueueBfs bfs.cu:569-619 (and the
 * 2-rank MPI loop bfs_mpi.cu:581-634) plus the min-merge gather bfs.cu:424-441 / bfs_mpi.cu:424-448.
 * Each emulated device has a PRIVATE full-length distance array and claims with "== INT_MAX then
 * min" (queueBfs, bfs.cu:146); discovered ids are bucketed by owner `v / part` (bfs.cu:148) with
 * part = n / k; the owner appends what every device sent it, in device order (bfs.cu:595-609);
 * nothing marks the owner's own distance array on receipt, so duplicates may re-enter queues; the
 * loop ends when all queues are empty (bfs.cu:569 / bfs_mpi.cu:621); the result is the element-wise
 * min over devices (bfs.cu:430-438).  Requires n % k == 0 (the reference writes out of bounds
 * otherwise, defect D2).  Within a device the frontier is processed sequentially in queue order,
 * one legal schedule of the kernel's threads; depths do not depend on the schedule.
 * Returns the number of levels executed, or a negative error.
 * ------------------------------------------------------------------------------------------- */
int64_t oracle_bfs_kdev(int64_t n, const int64_t* offsets, const int32_t* sizes, const int32_t* adj,
                        int64_t start, int k, int32_t* depth_out) {
  if (k < 1 || n % k != 0 || start < 0 || start >= n) return ORACLE_ERR_RANGE;
  const int64_t part = n / k;
  int32_t** dist = (int32_t**)calloc((size_t)k, sizeof(int32_t*));
  int32_t** curq = (int32_t**)calloc((size_t)k, sizeof(int32_t*));
  int32_t** nextq = (int32_t**)calloc((size_t)k, sizeof(int32_t*)); /* k buckets of n ints per device */
  int64_t* qsize = (int64_t*)calloc((size_t)k, sizeof(int64_t));
  int64_t* nsize = (int64_t*)calloc((size_t)k * k, sizeof(int64_t));
  size_t curcap = (size_t)n * (size_t)k + 1; /* an owner can receive up to n ids from each device */
  for (int d = 0; d < k; ++d) {
    dist[d] = (int32_t*)malloc((size_t)n * sizeof(int32_t));
    curq[d] = (int32_t*)malloc(curcap * sizeof(int32_t));
    nextq[d] = (int32_t*)malloc((size_t)n * (size_t)k * sizeof(int32_t) + 4);
    for (int64_t v = 0; v < n; ++v) dist[d][v] = INT_MAX; /* bfs.cu:404 */
    dist[d][start] = 0;                                   /* bfs.cu:406, copied to every device 411 */
  }
  const int owner = (int)(start / part); /* getDev, bfs.cu:29-32 */
  curq[owner][0] = (int32_t)start;
  qsize[owner] = 1;
  int64_t level = 0;
  for (;;) {
    int64_t total = 0;
    for (int d = 0; d < k; ++d) total += qsize[d];
    if (total == 0) break;
    for (int d = 0; d < k; ++d) { /* queueBfs on device d, bfs.cu:134-165 */
      for (int64_t t = 0; t < qsize[d]; ++t) {
        const int32_t u = curq[d][t];
        const int64_t b = offsets[u], e = b + sizes[u];
        for (int64_t i = b; i < e; ++i) {
          const int32_t v = adj[i];
          if (dist[d][v] == INT_MAX) {
            dist[d][v] = (int32_t)(level + 1);
            const int which = (int)(v / part);
            const int64_t pos = nsize[(size_t)d * k + which]++;
            nextq[d][pos + (int64_t)which * n] = v;
          }
        }
      }
    }
    level++;
    for (int i = 0; i < k; ++i) { /* exchange, bfs.cu:595-609 */
      int64_t off = 0;
      for (int j = 0; j < k; ++j) {
        const int64_t c = nsize[(size_t)j * k + i];
        memcpy(curq[i] + off, nextq[j] + (int64_t)i * n, (size_t)c * sizeof(int32_t));
        off += c;
      }
      qsize[i] = off; /* bfs.cu:617-618 */
    }
    memset(nsize, 0, (size_t)k * k * sizeof(int64_t)); /* bfs.cu:619 */
  }
  for (int64_t v = 0; v < n; ++v) { /* finalizeCudaBfs min-merge, bfs.cu:430-438 */
    int32_t best = dist[0][v];
    for (int d = 1; d < k; ++d) if (dist[d][v] < best) best = dist[d][v];
    depth_out[v] = best;
  }
  for (int d = 0; d < k; ++d) { free(dist[d]); free(curq[d]); free(nextq[d]); }
  free(dist); free(curq); free(nextq); free(qsize); free(nsize);
  return level;
}

/* ---------------------------------------------------------------------------------------------
 * checkOutput, bfs.cu:374-384: first index where two depth arrays differ, or -1 ("Output OK!").
 * ------------------------------------------------------------------------------------------- */
int64_t oracle_check_output(int64_t n, const int32_t* got, const int32_t* expected) {
  for (int64_t i = 0; i < n; ++i)
    if (got[i] != expected[i]) return i;
  return -1;
}

/* ---------------------------------------------------------------------------------------------
 * Synthetic inputs: tuples [first, first+count) of the generator defined in
 * include/b2bfs_graphgen.h.  No reference counterpart (the reference ships no live generator).
 * ------------------------------------------------------------------------------------------- */
void oracle_generate_edges(int kind, int scale, int edgefactor, uint64_t seed, int64_t first, int64_t count,
                           int32_t* src, int32_t* dst) {
  for (int64_t i = 0; i < count; ++i) {
    uint32_t u, v;
    b2_gen_edge(kind, scale, edgefactor, seed, (uint64_t)(first + i), &u, &v);
    src[i] = (int32_t)u;
    dst[i] = (int32_t)v;
  }
}
int64_t oracle_gen_num_vertices(int kind, int scale) { return b2_gen_num_vertices(kind, scale); }
int64_t oracle_gen_num_edges(int kind, int scale, int ef) { return b2_gen_num_edges(kind, scale, ef); }

/* ---------------------------------------------------------------------------------------------
 * Graph500 BFS-tree validation (the five rules of the Graph500 specification, "kernel 2
 * validation"), stated over the reference's symmetrised multigraph CSR.  parent[] holds VERTEX ids
 * (root -> root, -1 = not reached) as exported by the product's ABI; depth[] is the level array
 * with INT_MAX = not reached.  The reference itself never validates parents (bfs.cu:374-384
 * compares depth only), so this has no reference counterpart.
 * Returns 0 when valid, otherwise a positive rule number; *bad_vertex gets an offending vertex.
 *   1 root is its own parent with depth 0
 *   2 every reached non-root vertex has a reached parent exactly one level closer, and the pair
 *     (v, parent[v]) is an edge of the input graph
 *   3 every input edge joins vertices whose levels differ by at most one, or two unreached ones
 *   4 reached <=> has a parent (the tree spans exactly the root's component; with rule 3 no edge
 *     leaves the reached set)
 *   5 depth array equals the level implied by walking parents (no cycles)
 * ------------------------------------------------------------------------------------------- */
int oracle_validate_tree(int64_t n, const int64_t* offsets, const int32_t* sizes, const int32_t* adj,
                         int64_t root, const int32_t* depth, const int32_t* parent, int64_t* bad_vertex) {
  int64_t dummy;
  if (!bad_vertex) bad_vertex = &dummy;
  *bad_vertex = root;
  if (root < 0 || root >= n || parent[root] != root || depth[root] != 0) return 1;
  for (int64_t v = 0; v < n; ++v) {
    *bad_vertex = v;
    const int reached = depth[v] != INT_MAX;
    if (reached != (parent[v] >= 0)) return 4;
    if (!reached || v == root) continue;
    const int32_t p = parent[v];
    if (depth[v] < 1) return 5; /* only the root sits at level 0 */
    if (p >= n || depth[p] == INT_MAX || depth[p] != depth[v] - 1) return 2;
    int found = 0;
    for (int64_t i = offsets[v], e = offsets[v] + sizes[v]; i < e; ++i)
      if (adj[i] == p) { found = 1; break; }
    if (!found) return 2;
  }
  for (int64_t v = 0; v < n; ++v) {
    *bad_vertex = v;
    for (int64_t i = offsets[v], e = offsets[v] + sizes[v]; i < e; ++i) {
      const int32_t w = adj[i];
      const int rv = depth[v] != INT_MAX, rw = depth[w] != INT_MAX;
      if (rv != rw) return 3;
      if (rv) {
        const int32_t d = depth[v] - depth[w];
        if (d > 1 || d < -1) return 3;
      }
    }
  }
  /* rule 5 follows by induction from rule 2 + "depth >= 1 off the root"; the explicit walk is kept
     for small graphs as a belt-and-braces check (it is O(n * depth)). */
  for (int64_t v = 0; v < n && n <= (1 << 20); ++v) {
    if (depth[v] == INT_MAX) continue;
    *bad_vertex = v;
    int64_t x = v;
    int32_t steps = 0;
    while (x != root && steps <= depth[v]) { x = parent[x]; ++steps; }
    if (x != root || steps != depth[v]) return 5;
  }
  *bad_vertex = -1;
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Bookkeeping for the metric: reached vertices V_r and reached directed entries
 * D_r = sum of edgesSize over reached vertices = 2*m_r on the reference's CSR.
 * ------------------------------------------------------------------------------------------- */
void oracle_reach_stats(int64_t n, const int32_t* sizes, const int32_t* depth, int64_t* v_r, int64_t* d_r) {
  int64_t vr = 0, dr = 0;
  for (int64_t v = 0; v < n; ++v)
    if (depth[v] != INT_MAX) { vr++; dr += sizes[v]; }
  *v_r = vr;
  *d_r = dr;
}

/* ---------------------------------------------------------------------------------------------
 * Multi-core host BFS: a level-synchronous PORT of bfsCPU (bfs.cu:923-945) for all
 * host cores — same CSR, same depths; one frontier array per level, claims by compare-and-swap on
 * the depth array, per-thread chunks appended to the next frontier.  Not the reference's code
 * (which is serial); it exists so that the CPU baseline is not flattered by using one core.
 * POSIX threads + a barrier per level.  Returns the number of vertices reached.
 * ------------------------------------------------------------------------------------------- */
#include <pthread.h>
#include <unistd.h>

int oracle_num_threads(void) {
  long c = sysconf(_SC_NPROCESSORS_ONLN);
  return c < 1 ? 1 : (c > 64 ? 64 : (int)c);
}

typedef struct {
  int64_t n;
  const int64_t* offsets;
  const int32_t* sizes;
  const int32_t* adj;
  int32_t* depth;
  int32_t *cur, *nxt;
  int64_t ncur, nnext, cursor, reached;
  int32_t level;
  int threads, done;
  pthread_barrier_t bar;
} par_bfs_t;

typedef struct { par_bfs_t* s; int tid; } par_arg_t;

static void* par_bfs_worker(void* argp) {
  par_arg_t* a = (par_arg_t*)argp;
  par_bfs_t* s = a->s;
  enum { CHUNK = 1024, GRAB = 64 };
  int32_t local[CHUNK];
  for (int64_t v = (int64_t)a->tid; v < s->n; v += s->threads) s->depth[v] = INT_MAX;
  pthread_barrier_wait(&s->bar);
  for (;;) {
    pthread_barrier_wait(&s->bar); /* thread 0 has published the level */
    if (s->done) break;
    int cnt = 0;
    for (;;) {
      const int64_t t0 = __atomic_fetch_add(&s->cursor, (int64_t)GRAB, __ATOMIC_RELAXED);
      if (t0 >= s->ncur) break;
      const int64_t t1 = t0 + GRAB < s->ncur ? t0 + GRAB : s->ncur;
      for (int64_t t = t0; t < t1; ++t) {
        const int32_t u = s->cur[t];
        const int64_t b = s->offsets[u], e = b + s->sizes[u];
        for (int64_t i = b; i < e; ++i) {
          const int32_t v = s->adj[i];
          int32_t expect = INT_MAX;
          if (__atomic_load_n(&s->depth[v], __ATOMIC_RELAXED) == INT_MAX &&
              __atomic_compare_exchange_n(&s->depth[v], &expect, s->level + 1, 0, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {
            local[cnt++] = v;
            if (cnt == CHUNK) {
              const int64_t at = __atomic_fetch_add(&s->nnext, (int64_t)cnt, __ATOMIC_RELAXED);
              memcpy(s->nxt + at, local, (size_t)cnt * sizeof(int32_t));
              cnt = 0;
            }
          }
        }
      }
    }
    if (cnt) {
      const int64_t at = __atomic_fetch_add(&s->nnext, (int64_t)cnt, __ATOMIC_RELAXED);
      memcpy(s->nxt + at, local, (size_t)cnt * sizeof(int32_t));
    }
    pthread_barrier_wait(&s->bar); /* level complete */
    if (a->tid == 0) {
      int32_t* tmp = s->cur; s->cur = s->nxt; s->nxt = tmp;
      s->ncur = s->nnext;
      s->reached += s->nnext;
      s->nnext = 0;
      s->cursor = 0;
      s->level++;
      s->done = s->ncur == 0;
    }
  }
  return NULL;
}

int64_t oracle_bfs_parallel(int64_t n, const int64_t* offsets, const int32_t* sizes, const int32_t* adj,
                            int64_t start, int32_t* depth, int threads) {
  if (start < 0 || start >= n) return ORACLE_ERR_RANGE;
  if (threads < 1) threads = oracle_num_threads();
  par_bfs_t s;
  memset(&s, 0, sizeof(s));
  s.n = n; s.offsets = offsets; s.sizes = sizes; s.adj = adj; s.depth = depth; s.threads = threads;
  s.cur = (int32_t*)malloc((size_t)n * sizeof(int32_t));
  s.nxt = (int32_t*)malloc((size_t)n * sizeof(int32_t));
  if (!s.cur || !s.nxt) { free(s.cur); free(s.nxt); return ORACLE_ERR_NOMEM; }
  pthread_barrier_init(&s.bar, NULL, (unsigned)threads + 1);
  pthread_t* th = (pthread_t*)malloc((size_t)threads * sizeof(pthread_t));
  par_arg_t* args = (par_arg_t*)malloc((size_t)threads * sizeof(par_arg_t));
  for (int t = 0; t < threads; ++t) { args[t].s = &s; args[t].tid = t; pthread_create(&th[t], NULL, par_bfs_worker, &args[t]); }
  pthread_barrier_wait(&s.bar); /* depth array filled */
  depth[start] = 0;
  s.cur[0] = (int32_t)start;
  s.ncur = 1; s.reached = 1;
  /* the coordinator only takes part in the barriers; workers run the levels */
  for (;;) {
    pthread_barrier_wait(&s.bar);
    if (s.done) break;
    pthread_barrier_wait(&s.bar);
    /* thread 0 of the workers advances the level right after this barrier; wait for it at the next one */
  }
  for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
  pthread_barrier_destroy(&s.bar);
  const int64_t reached = s.reached;
  free(th); free(args); free(s.cur); free(s.nxt);
  return reached;
}

/* ---------------------------------------------------------------------------------------------
 * Helpers for bench.py's parity / baseline legs on graphs too large for the serial builders above.
 * None of them restates new reference arithmetic; each is checked in tests/test_oracle.py against
 * oracle_build_csr / oracle_bfs_serial.
 * ------------------------------------------------------------------------------------------- */

/* generic "run fn(tid) on T threads" */
typedef struct { void (*fn)(void*, int, int); void* ctx; int tid, threads; } par_job_t;
static void* par_job_entry(void* p) {
  par_job_t* j = (par_job_t*)p;
  j->fn(j->ctx, j->tid, j->threads);
  return NULL;
}
static void run_on_threads(void (*fn)(void*, int, int), void* ctx, int threads) {
  if (threads < 1) threads = oracle_num_threads();
  pthread_t* th = (pthread_t*)malloc((size_t)threads * sizeof(pthread_t));
  par_job_t* jobs = (par_job_t*)malloc((size_t)threads * sizeof(par_job_t));
  for (int t = 0; t < threads; ++t) {
    jobs[t].fn = fn; jobs[t].ctx = ctx; jobs[t].tid = t; jobs[t].threads = threads;
    pthread_create(&th[t], NULL, par_job_entry, &jobs[t]);
  }
  for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
  free(th);
  free(jobs);
}

/* All-cores CSR build: the same multiset per row as oracle_build_csr (bfs.cu:851-875: v pushed onto
 * u's list and u onto v's, duplicates and self-loops kept), but the ORDER inside a row is whatever
 * the threads' atomic cursors produce, not file order.  Depths do not depend on that order (parent
 * edge indices do), so this builder is used only where depths are compared: bench.py at scale >= 24. */
typedef struct {
  int64_t n, m;
  const int32_t *src, *dst;
  int64_t* offsets;
  int32_t* sizes;
  int32_t* adj;
  int64_t* cur;
  int bad;
} pcsr_t;
static void pcsr_count(void* c, int tid, int T) {
  pcsr_t* s = (pcsr_t*)c;
  const int64_t a = s->m * tid / T, b = s->m * (tid + 1) / T;
  for (int64_t i = a; i < b; ++i) {
    const int32_t u = s->src[i], v = s->dst[i];
    if (u < 0 || v < 0 || u >= s->n || v >= s->n) { s->bad = 1; return; }
    __atomic_fetch_add(&s->sizes[u], 1, __ATOMIC_RELAXED);
    __atomic_fetch_add(&s->sizes[v], 1, __ATOMIC_RELAXED);
  }
}
static void pcsr_fill(void* c, int tid, int T) {
  pcsr_t* s = (pcsr_t*)c;
  const int64_t a = s->m * tid / T, b = s->m * (tid + 1) / T;
  for (int64_t i = a; i < b; ++i) {
    const int32_t u = s->src[i], v = s->dst[i];
    s->adj[__atomic_fetch_add(&s->cur[u], 1, __ATOMIC_RELAXED)] = v;
    s->adj[__atomic_fetch_add(&s->cur[v], 1, __ATOMIC_RELAXED)] = u;
  }
}
int oracle_build_csr_parallel(int64_t n, int64_t m, const int32_t* src, const int32_t* dst, int64_t* offsets,
                              int32_t* sizes, int32_t* adj, int threads) {
  pcsr_t s = {n, m, src, dst, offsets, sizes, adj, NULL, 0};
  memset(sizes, 0, (size_t)n * sizeof(int32_t));
  run_on_threads(pcsr_count, &s, threads);
  if (s.bad) return ORACLE_ERR_RANGE;
  int64_t run = 0;
  for (int64_t v = 0; v < n; ++v) { offsets[v] = run; run += sizes[v]; }
  s.cur = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
  if (!s.cur) return ORACLE_ERR_NOMEM;
  memcpy(s.cur, offsets, (size_t)n * sizeof(int64_t));
  run_on_threads(pcsr_fill, &s, threads);
  free(s.cur);
  return ORACLE_OK;
}

/* Depth check straight from the edge tuples, no CSR needed.  A depth array equals bfsCPU's
 * (bfs.cu:923-945) for `root` if and only if
 *   (1) depth[root] == 0 and no other vertex has depth 0,
 *   (2) for every tuple (u,v): both endpoints reached or both not, and |depth[u]-depth[v]| <= 1,
 *   (3) every reached vertex other than the root has a tuple to a vertex one level up.
 * (2) gives depth[v] <= dist(root,v) by induction along a shortest path and keeps unreached vertices
 * out of the root's component; (3) gives a walk of depth[v] steps back to the root, so
 * depth[v] >= dist.  Returns -1 when all three hold, else the first offending vertex found
 * (a vertex id >= 0).  All host cores; two streaming passes over the tuples. */
typedef struct {
  int64_t n, m, root;
  const int32_t *src, *dst, *depth;
  uint8_t* up;      /* up[v] = 1: v has a neighbour one level closer to the root */
  int64_t bad;
} echk_t;
static void echk_edges(void* c, int tid, int T) {
  echk_t* s = (echk_t*)c;
  const int64_t a = s->m * tid / T, b = s->m * (tid + 1) / T;
  for (int64_t i = a; i < b; ++i) {
    const int32_t u = s->src[i], v = s->dst[i];
    const int32_t du = s->depth[u], dv = s->depth[v];
    if ((du == INT_MAX) != (dv == INT_MAX)) { __atomic_store_n(&s->bad, du == INT_MAX ? u : v, __ATOMIC_RELAXED); return; }
    if (du == INT_MAX) continue;
    if (du - dv > 1 || dv - du > 1) { __atomic_store_n(&s->bad, du > dv ? u : v, __ATOMIC_RELAXED); return; }
    if (du + 1 == dv) s->up[v] = 1;        /* racing stores of the same value */
    else if (dv + 1 == du) s->up[u] = 1;
  }
}
static void echk_vertices(void* c, int tid, int T) {
  echk_t* s = (echk_t*)c;
  const int64_t a = s->n * tid / T, b = s->n * (tid + 1) / T;
  for (int64_t v = a; v < b; ++v) {
    const int32_t d = s->depth[v];
    if (d == INT_MAX) continue;
    if (d < 0 || (d == 0) != (v == s->root) || (v != s->root && !s->up[v])) { __atomic_store_n(&s->bad, v, __ATOMIC_RELAXED); return; }
  }
}
int64_t oracle_check_depth_by_edges(int64_t n, int64_t m, const int32_t* src, const int32_t* dst, int64_t root,
                                    const int32_t* depth, int threads) {
  if (root < 0 || root >= n) return ORACLE_ERR_RANGE - 100;
  echk_t s = {n, m, root, src, dst, depth, (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1), -1};
  if (!s.up) return ORACLE_ERR_NOMEM - 100;
  if (depth[root] != 0) s.bad = root;
  if (s.bad < 0) run_on_threads(echk_edges, &s, threads);
  if (s.bad < 0) run_on_threads(echk_vertices, &s, threads);
  free(s.up);
  return s.bad;
}

/* bfsCPU (bfs.cu:923-945) stopped after it has scanned at least `max_entries` adjacency entries: a
 * BOUNDED SAMPLE of one serial search for graphs where a whole one takes minutes (bench.py
 * --impl reference at scale >= 26).  Same loop as oracle_bfs_serial; depth is complete only if the
 * return value (entries scanned) is below max_entries.  *reached_out = vertices dequeued + queued. */
int64_t oracle_bfs_serial_bounded(int64_t n, const int64_t* offsets, const int32_t* sizes, const int32_t* adj,
                                  int64_t start, int32_t* depth, int64_t max_entries, int64_t* reached_out) {
  for (int64_t v = 0; v < n; ++v) depth[v] = INT_MAX;
  if (start < 0 || start >= n) return ORACLE_ERR_RANGE;
  uint8_t* visited = (uint8_t*)calloc((size_t)n, 1);
  int32_t* q = (int32_t*)malloc((size_t)n * sizeof(int32_t));
  if (!visited || !q) { free(visited); free(q); return ORACLE_ERR_NOMEM; }
  int64_t head = 0, tail = 0, scanned = 0;
  depth[start] = 0;
  visited[start] = 1;
  q[tail++] = (int32_t)start;
  while (head < tail && scanned < max_entries) {
    const int32_t u = q[head++];
    const int64_t b = offsets[u], e = b + sizes[u];
    for (int64_t i = b; i < e; ++i) {
      const int32_t v = adj[i];
      if (!visited[v]) {
        visited[v] = 1;
        depth[v] = depth[u] + 1;
        q[tail++] = v;
      }
    }
    scanned += e - b;
  }
  free(visited);
  free(q);
  if (reached_out) *reached_out = tail;
  return scanned;
}

/* ---------------------------------------------------------------------------------------------
 * The reference's second reader, readGraph (bfs.cu:882-920; never called by its main).
 *   stdin form  (bfs.cu:888-890, 899-901): "n m" then m pairs "u v", each pushed onto u's list ONLY -
 *               a directed graph, not symmetrised.  oracle_build_csr_directed restates the flatten
 *               loop (bfs.cu:910-916) for such tuples; oracle_bfs_serial then runs on it unchanged.
 *   random form (bfs.cu:891-894, 902-906): srand(12345); u = rand() % n; v = rand() % n per edge, both
 *               directions pushed.  oracle_random_edges draws the same tuples from the C library's
 *               rand() - the stream is glibc's, the same library the reference links.
 * ------------------------------------------------------------------------------------------- */
int oracle_build_csr_directed(int64_t n, int64_t m, const int32_t* src, const int32_t* dst, int64_t* offsets,
                              int32_t* sizes, int32_t* adj) {
  memset(sizes, 0, (size_t)n * sizeof(int32_t));
  for (int64_t i = 0; i < m; ++i) {
    if (src[i] < 0 || dst[i] < 0 || src[i] >= n || dst[i] >= n) return ORACLE_ERR_RANGE;
    sizes[src[i]]++;
  }
  int64_t run = 0;
  for (int64_t v = 0; v < n; ++v) { offsets[v] = run; run += sizes[v]; }
  int64_t* cur = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
  if (!cur) return ORACLE_ERR_NOMEM;
  memcpy(cur, offsets, (size_t)n * sizeof(int64_t));
  for (int64_t i = 0; i < m; ++i) adj[cur[src[i]]++] = dst[i]; /* bfs.cu:901 */
  free(cur);
  return ORACLE_OK;
}

void oracle_random_edges(int64_t n, int64_t m, int32_t* src, int32_t* dst) {
  srand(12345); /* bfs.cu:892 */
  for (int64_t i = 0; i < m; ++i) {
    src[i] = (int32_t)(rand() % n); /* bfs.cu:903 */
    dst[i] = (int32_t)(rand() % n); /* bfs.cu:904 */
  }
}
