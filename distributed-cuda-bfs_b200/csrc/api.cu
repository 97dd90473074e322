// api.cu — the extern "C" boundary declared in include/b2bfs.h.
//
// The handle owns what the reference keeps in globals: the Graph (bfs.cu:28), the __managed__
// per-device pointer tables (bfs.cu:255-269) and the buffers initCuda2 allocates and never frees
// (bfs.cu:333-351, 362-372).
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/b2bfs_graphgen.h"
#include "batch.cuh"
#include "bfs.cuh"
#include "build.cuh"
#include "edge_file.hpp"
#include "multi.cuh"

using namespace b2;

struct b2bfs {
  int device = 0;
  int rank = 0, world = 1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  // options
  int dedup = 1, mode = B2BFS_MODE_DIRECTION_OPT, engine = B2BFS_ENGINE_FUSED;
  int alpha = 14, beta = 24, gamma = 2000, huge_min = 0, xfullpush = 0, keep_edges = 1, gather_parent = 1, hub_first = 1, directed = 0;
  // graph
  bool has_graph = false;
  int64_t n = 0, m = 0;
  int64_t m_first = 0, m_loc = 0;  // the tuples this rank holds: [m_first, m_first + m_loc) (all of them on one GPU)
  int32_t *d_src = nullptr, *d_dst = nullptr;
  BuiltCsr csr;
  // search state
  BatchState batch;  // allocated by the first b2bfs_run_batch on this graph
  unsigned int* cmp_scratch = nullptr;  // comparator engines: n + n/1024 + 16 words, allocated on first use
  SearchView sv{};
  Control* pinned_ctl = nullptr;
  int grid = 0;
  int grid_cluster = 0, cluster_size = 0;  // grid and CTAs per cluster of the cluster-mode kernel (0: not possible here)
  int use_cluster = 1;
  bool has_graph_rows_narrow() const { return has_graph && csr.max_degree < kNarrow; }
  int64_t build_launches = 0;
  // multi-rank (world > 1): 1-D partition of the rows, NCCL communicator, exchange buffers
  MultiContext mc;
  int64_t part = 0, vbase = 0, nloc = 0;
  uint32_t *gfrontier = nullptr, *seen = nullptr;
  int2* send_buf = nullptr;
  unsigned int* send_count = nullptr;
  int32_t* gather_tmp = nullptr;  // world * part ints: allgather staging for depth / parent / degrees
  // peer-memory engine: this rank's exchange region and the IPC mappings of the peers' regions
  int multi_engine = 0;            // 0: peer memory (falls back to 1 when IPC is unavailable), 1: NCCL per level
  bool p2p_ready = false;
  unsigned char* xregion = nullptr;
  unsigned char* xpeer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int grid_multi = 0;
};

static thread_local std::string g_create_error;

static int fail(b2bfs* h, const Status& s) {
  if (h) h->err = s.msg; else g_create_error = s.msg;
  return s.code;
}
static int fail(b2bfs* h, int code, const std::string& msg) { return fail(h, make_error(code, msg)); }

#define API_CUDA(h, expr)                                                                              \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess)                                                                             \
      return fail(h, _e == cudaErrorMemoryAllocation ? B2BFS_ERR_NOMEM : B2BFS_ERR_CUDA,               \
                  std::string(#expr) + ": " + cudaGetErrorString(_e));                                 \
  } while (0)

static void free_search_state(b2bfs* h) {
  cudaFree(h->sv.depth);
  cudaFree(h->sv.parent);
  cudaFree(h->sv.visited);
  cudaFree(h->sv.bitmap[0]);
  cudaFree(h->sv.bitmap[1]);
  cudaFree(h->sv.queue[0]);
  cudaFree(h->sv.queue[1]);
  cudaFree(h->sv.wideq[0]);
  cudaFree(h->sv.wideq[1]);
  cudaFree(h->sv.hugeq[0]);
  cudaFree(h->sv.hugeq[1]);
  cudaFree(h->sv.ctl);
  h->sv = SearchView{};
}

static void free_multi_state(b2bfs* h) {
  if (h->xregion) {
    // collective: unmap the peers' regions, make sure every rank has done so, then free our own
    for (int r = 0; r < h->world; ++r)
      if (r != h->rank && h->xpeer[r]) cudaIpcCloseMemHandle(h->xpeer[r]);
    if (h->mc.comm && h->mc.stats_send) {
      h->mc.nccl->AllReduce(h->mc.stats_send, h->mc.stats_all, 1, ncclUint64, ncclSum, h->mc.comm, h->stream);
      cudaStreamSynchronize(h->stream);
    }
    cudaFree(h->xregion);
    h->xregion = nullptr;
    for (int r = 0; r < 8; ++r) h->xpeer[r] = nullptr;
    h->p2p_ready = false;
  }
  cudaFree(h->gfrontier);
  cudaFree(h->seen);
  cudaFree(h->send_buf);
  cudaFree(h->send_count);
  cudaFree(h->gather_tmp);
  cudaFree(h->mc.recv_buf);
  h->gfrontier = h->seen = nullptr;
  h->send_buf = nullptr;
  h->send_count = nullptr;
  h->gather_tmp = nullptr;
  h->mc.recv_buf = nullptr;
}

static void free_graph(b2bfs* h) {
  free_multi_state(h);
  free_batch(&h->batch);
  cudaFree(h->cmp_scratch);
  h->cmp_scratch = nullptr;
  cudaFree(h->d_src);
  cudaFree(h->d_dst);
  h->d_src = h->d_dst = nullptr;
  free_csr(&h->csr);
  free_search_state(h);
  h->has_graph = false;
  h->n = h->m = 0;
}

// which tuples this rank holds: all of them on one GPU, an m / world share on several
static void set_tuple_share(b2bfs* h) {
  h->m_first = h->world > 1 ? h->m * h->rank / h->world : 0;
  h->m_loc = (h->world > 1 ? h->m * (h->rank + 1) / h->world : h->m) - h->m_first;
}

#define API_NCCL(h, expr)                                                                          \
  do {                                                                                             \
    ncclResult_t _r = (expr);                                                                      \
    if (_r != ncclSuccess)                                                                         \
      return fail(h, B2BFS_ERR_COMM, std::string(#expr) + ": " + (h)->mc.nccl->GetErrorString(_r)); \
  } while (0)

// Several ranks: this rank's share of the tuples -> the CSR of this rank's rows.  Each tuple is
// expanded where it lies, its two keys are bucketed by the rank that owns their row, the buckets are
// exchanged over NVLink (grouped ncclSend / ncclRecv), and a rank sorts only what it owns.  Replaces
// "every rank reads and keeps the whole graph" (bfs_mpi.cu:815, 832; bfs.cu:333-351).
static int build_partitioned(b2bfs* h) {
  const int k = h->world, me = h->rank;
  const KeySpec ks = key_spec(h->n, h->part, h->hub_first, 0);
  struct Temps {  // freed on every path
    uint32_t* degc = nullptr;
    unsigned long long* cursors = nullptr;
    int* bad = nullptr;
    uint64_t *send = nullptr, *ka = nullptr;
    ~Temps() { cudaFree(degc); cudaFree(cursors); cudaFree(bad); cudaFree(send); cudaFree(ka); }
  } t;
  API_CUDA(h, cudaMalloc(&t.cursors, 16 * sizeof(unsigned long long)));
  API_CUDA(h, cudaMalloc(&t.bad, sizeof(int)));
  API_CUDA(h, cudaMemsetAsync(t.cursors, 0, 16 * sizeof(unsigned long long), h->stream));
  API_CUDA(h, cudaMemsetAsync(t.bad, 0, sizeof(int), h->stream));
  Status s;
  if (ks.cbits) {  // degree class of every vertex: local counts, summed over the ranks
    API_CUDA(h, cudaMalloc(&t.degc, (size_t)h->n * sizeof(uint32_t)));
    API_CUDA(h, cudaMemsetAsync(t.degc, 0, (size_t)h->n * sizeof(uint32_t), h->stream));
    s = count_degrees_device(h->d_src, h->d_dst, h->m_loc, h->n, t.degc, h->stream);
    if (!s.ok()) return fail(h, s);
    API_NCCL(h, h->mc.nccl->AllReduce(t.degc, t.degc, (size_t)h->n, ncclUint32, ncclSum, h->mc.comm, h->stream));
    h->build_launches += 1;
  }
  // pass 1: how many keys for each owner; every rank learns the whole k x k matrix (+ a bad-id flag)
  s = owner_keys_device(false, h->d_src, h->d_dst, h->m_loc, h->n, h->part, k, ks, t.degc, t.cursors, nullptr, t.bad, h->stream);
  if (!s.ok()) return fail(h, s);
  unsigned long long mine[16] = {0};
  int bad_h = 0;
  API_CUDA(h, cudaMemcpyAsync(mine, t.cursors, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
  API_CUDA(h, cudaMemcpyAsync(&bad_h, t.bad, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  mine[8] = (unsigned long long)bad_h;
  API_CUDA(h, cudaMemcpyAsync(h->mc.stats_send, mine, 16 * sizeof(unsigned long long), cudaMemcpyHostToDevice, h->stream));
  API_NCCL(h, h->mc.nccl->AllGather(h->mc.stats_send, h->mc.stats_all, 16, ncclUint64, h->mc.comm, h->stream));
  API_CUDA(h, cudaMemcpyAsync(h->mc.stats_host, h->mc.stats_all, (size_t)k * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  const unsigned long long* all = h->mc.stats_host;  // all[s * 16 + d] = keys rank s holds for rank d
  for (int r = 0; r < k; ++r)
    if (all[r * 16 + 8]) return fail(h, B2BFS_ERR_RANGE, "edge list holds a vertex id outside [0, vertex_num)");
  // pass 2: keys into one bucket per owner
  unsigned long long start[16] = {0}, total_send = 0;
  for (int d = 0; d < k; ++d) { start[d] = total_send; total_send += mine[d]; }
  API_CUDA(h, cudaMalloc(&t.send, (size_t)(total_send > 0 ? total_send : 1) * sizeof(uint64_t)));
  API_CUDA(h, cudaMemcpyAsync(t.cursors, start, 8 * sizeof(unsigned long long), cudaMemcpyHostToDevice, h->stream));
  s = owner_keys_device(true, h->d_src, h->d_dst, h->m_loc, h->n, h->part, k, ks, t.degc, t.cursors, t.send, t.bad, h->stream);
  if (!s.ok()) return fail(h, s);
  h->build_launches += 2;
  // exchange: bucket d of rank s lands in rank d's key array at the offset of s
  unsigned long long N = 0, at[16] = {0};
  for (int r = 0; r < k; ++r) { at[r] = N; N += all[r * 16 + me]; }
  API_CUDA(h, cudaMalloc(&t.ka, (size_t)(N > 0 ? N : 1) * sizeof(uint64_t)));
  API_NCCL(h, h->mc.nccl->GroupStart());
  for (int r = 0; r < k; ++r) {
    if (r == me) continue;
    if (mine[r]) API_NCCL(h, h->mc.nccl->Send(t.send + start[r], (size_t)mine[r], ncclUint64, r, h->mc.comm, h->stream));
    if (all[r * 16 + me]) API_NCCL(h, h->mc.nccl->Recv(t.ka + at[r], (size_t)all[r * 16 + me], ncclUint64, r, h->mc.comm, h->stream));
  }
  API_NCCL(h, h->mc.nccl->GroupEnd());
  if (mine[me])
    API_CUDA(h, cudaMemcpyAsync(t.ka + at[me], t.send + start[me], (size_t)mine[me] * sizeof(uint64_t), cudaMemcpyDeviceToDevice, h->stream));
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  cudaFree(t.send); t.send = nullptr;
  cudaFree(t.degc); t.degc = nullptr;
  uint64_t* ka = t.ka;
  t.ka = nullptr;  // build_csr_from_keys owns it from here
  s = build_csr_from_keys(ka, (int64_t)N, h->n, h->vbase, h->vbase + h->nloc, ks, h->dedup, 0, h->part / 32, h->stream, &h->csr,
                          &h->build_launches);
  if (!s.ok()) return fail(h, s);
  return B2BFS_OK;
}

// edges are in d_src/d_dst: build the CSR (this rank's rows) and the per-search buffers
static int finish_load(b2bfs* h) {
  const bool multi = h->world > 1;
  if (multi) {
    // contiguous ranges of `part` rows; a multiple of 128 so that bitmap slices are whole 16-byte groups
    h->part = ((h->n + h->world - 1) / h->world + 127) / 128 * 128;
    h->vbase = std::min(h->n, (int64_t)h->rank * h->part);
    h->nloc = std::min(h->n, (int64_t)(h->rank + 1) * h->part) - h->vbase;
  } else {
    h->part = 0;
    h->vbase = 0;
    h->nloc = h->n;
  }
  Status s;
  if (!multi) {
    s = build_csr_device(h->d_src, h->d_dst, h->m, h->n, h->dedup, h->hub_first, h->directed, h->stream, &h->csr,
                         &h->build_launches);
    if (!s.ok()) {
      free_graph(h);
      return fail(h, s);
    }
  } else {
    const int rc = build_partitioned(h);
    if (rc) { free_graph(h); return rc; }
  }
  if (!h->keep_edges) {
    cudaFree(h->d_src);
    cudaFree(h->d_dst);
    h->d_src = h->d_dst = nullptr;
  }
  const int64_t rows = multi ? h->part : h->n, nw = h->csr.nwords;
  const size_t wide_cap = (size_t)h->csr.wide_items + 1;
  API_CUDA(h, cudaMalloc(&h->sv.depth, (size_t)(rows + 4) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.parent, (size_t)(rows + 4) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.visited, (size_t)nw * sizeof(uint32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.bitmap[0], (size_t)nw * sizeof(uint32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.bitmap[1], (size_t)nw * sizeof(uint32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.queue[0], (size_t)(rows > 0 ? rows : 1) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.queue[1], (size_t)(rows > 0 ? rows : 1) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.wideq[0], wide_cap * sizeof(int2)));
  API_CUDA(h, cudaMalloc(&h->sv.wideq[1], wide_cap * sizeof(int2)));
  if (!multi) {  // hub-row items: rows of >= 2 * kCtaChunk entries at the least, so E / kCtaChunk + E / (2 kCtaChunk) bounds them
    const size_t huge_cap = (size_t)h->csr.entries / kCtaChunk + (size_t)h->csr.entries / (2 * kCtaChunk) + 16;
    API_CUDA(h, cudaMalloc(&h->sv.hugeq[0], huge_cap * sizeof(int2)));
    API_CUDA(h, cudaMalloc(&h->sv.hugeq[1], huge_cap * sizeof(int2)));
  }
  API_CUDA(h, cudaMalloc(&h->sv.ctl, sizeof(Control)));
  API_CUDA(h, cudaMemsetAsync(h->sv.ctl, 0, sizeof(Control), h->stream));
  if (multi) {
    const size_t k = (size_t)h->world, part = (size_t)h->part;
    API_CUDA(h, cudaMalloc(&h->gfrontier, k * part / 8));
    API_CUDA(h, cudaMalloc(&h->seen, k * part / 8));
    API_CUDA(h, cudaMalloc(&h->send_buf, k * part * sizeof(int2)));
    API_CUDA(h, cudaMalloc(&h->send_count, 8 * 32 * sizeof(unsigned int)));  // one 128-byte line per possible peer
    API_CUDA(h, cudaMalloc(&h->gather_tmp, k * part * sizeof(int32_t)));
    API_CUDA(h, cudaMalloc(&h->mc.recv_buf, (k - 1) * part * sizeof(int2)));
    API_CUDA(h, cudaMemsetAsync(h->sv.depth, 0xff, (size_t)(rows + 4) * sizeof(int32_t), h->stream));
    // resident entries over all ranks (the switch heuristic's m_u starts from it)
    unsigned long long e = (unsigned long long)h->csr.entries;
    API_CUDA(h, cudaMemcpyAsync(h->mc.stats_send, &e, sizeof(e), cudaMemcpyHostToDevice, h->stream));
    API_NCCL(h, h->mc.nccl->AllReduce(h->mc.stats_send, h->mc.stats_all, 1, ncclUint64, ncclSum, h->mc.comm, h->stream));
    API_CUDA(h, cudaMemcpyAsync(h->mc.stats_host, h->mc.stats_all, sizeof(e), cudaMemcpyDeviceToHost, h->stream));
    API_CUDA(h, cudaStreamSynchronize(h->stream));
    h->mc.entries_global = (int64_t)h->mc.stats_host[0];
    h->mc.n_global = h->n;
    // peer-memory engine: one exchange region per rank, mapped into every peer with CUDA IPC
    // (the handles travel through the NCCL communicator, so no extra side channel is needed)
    unsigned long long ok = 1;
    cudaIpcMemHandle_t handles[8];
    const size_t xbytes = xchg_region_bytes(h->world, h->part);
    if (cudaMalloc(&h->xregion, xbytes) != cudaSuccess || cudaMemsetAsync(h->xregion, 0, xbytes, h->stream) != cudaSuccess ||
        cudaIpcGetMemHandle(&handles[h->rank], h->xregion) != cudaSuccess) {
      ok = 0;
      cudaGetLastError();
      memset(&handles[h->rank], 0, sizeof(cudaIpcMemHandle_t));
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    unsigned char* hbuf = reinterpret_cast<unsigned char*>(h->gather_tmp);
    API_CUDA(h, cudaMemcpyAsync(hbuf + 64 * h->rank, &handles[h->rank], 64, cudaMemcpyHostToDevice, h->stream));
    API_NCCL(h, h->mc.nccl->AllGather(hbuf + 64 * h->rank, hbuf, 64, ncclUint8, h->mc.comm, h->stream));
    API_CUDA(h, cudaMemcpyAsync(handles, hbuf, 64 * (size_t)k, cudaMemcpyDeviceToHost, h->stream));
    API_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int r = 0; r < h->world && ok; ++r) {
      if (r == h->rank) { h->xpeer[r] = h->xregion; continue; }
      void* ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, handles[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        ok = 0;
        cudaGetLastError();
      }
      h->xpeer[r] = static_cast<unsigned char*>(ptr);
    }
    // every rank must agree on the engine
    API_CUDA(h, cudaMemcpyAsync(h->mc.stats_send, &ok, sizeof(ok), cudaMemcpyHostToDevice, h->stream));
    API_NCCL(h, h->mc.nccl->AllReduce(h->mc.stats_send, h->mc.stats_all, 1, ncclUint64, ncclMin, h->mc.comm, h->stream));
    API_CUDA(h, cudaMemcpyAsync(h->mc.stats_host, h->mc.stats_all, sizeof(ok), cudaMemcpyDeviceToHost, h->stream));
    API_CUDA(h, cudaStreamSynchronize(h->stream));
    h->p2p_ready = h->mc.stats_host[0] == 1 && h->xregion != nullptr;
  }
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  h->has_graph = true;
  return B2BFS_OK;
}

// multi-rank: allgather one int32 per vertex (padded slices of `part`) and copy the first n to the host
static int gather_i32(b2bfs* h, const int32_t* local_dev, int32_t* host_out, int64_t* d2h) {
  API_NCCL(h, h->mc.nccl->AllGather(local_dev, h->gather_tmp, (size_t)h->part, ncclInt32, h->mc.comm, h->stream));
  if (host_out) {
    API_CUDA(h, cudaMemcpyAsync(host_out, h->gather_tmp, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    if (d2h) *d2h += h->n * (int64_t)sizeof(int32_t);
    API_CUDA(h, cudaStreamSynchronize(h->stream));  // gather_tmp is reused by the next gather
  }
  return B2BFS_OK;
}

static BfsParams make_params(const b2bfs* h) {
  BfsParams p;
  p.g.n = h->n;
  p.g.nwords = h->csr.nwords;
  p.g.entries = h->csr.entries;
  p.g.rowptr = h->csr.rowptr;
  p.g.col = h->csr.col;
  p.g.visited_init = h->csr.visited_init;
  p.g.first_nbr = h->csr.first_nbr;
  const bool multi = h->world > 1;
  p.g.n = multi ? h->nloc : h->n;
  p.g.vbase = h->vbase;
  p.g.part = multi ? h->part : h->csr.nwords * 32;
  p.g.rank = h->rank;
  p.g.world = h->world;
  p.s = h->sv;
  p.s.gfrontier = h->gfrontier;
  p.s.seen = h->seen;
  p.s.send_buf = h->send_buf;
  p.s.send_count = h->send_count;
  p.s.send_cap = h->part;
  const bool p2p = multi && h->p2p_ready && h->multi_engine == 0;
  for (int r = 0; r < 8; ++r) p.s.xbase[r] = p2p ? h->xpeer[r] : nullptr;
  p.g.n_global = h->n;
  p.g.entries_global = multi ? (unsigned long long)h->mc.entries_global : (unsigned long long)h->csr.entries;
  p.tn.mode = h->directed ? B2BFS_MODE_TOP_DOWN : h->mode;  // bottom-up needs in-edges = out-edges
  p.tn.alpha = h->alpha;
  p.tn.beta = h->beta;
  p.tn.gamma = h->gamma;
  p.tn.huge_min = (!multi && h->sv.hugeq[0]) ? h->huge_min : 0;  // no hub lists, no hub bin
  // graphs of narrow rows only (grids, road-like meshes): the kernel instance with the cluster mode
  p.tn.cluster = (h->world == 1 && h->use_cluster && h->grid_cluster > 0 && h->has_graph_rows_narrow()) ? h->cluster_size : 0;
  p.tn.xfullpush = h->xfullpush;
  return p;
}

// ---------------------------------------------------------------------------------------------
// graph files of any supported format -> host tuple arrays (malloc or page-locked)
// ---------------------------------------------------------------------------------------------
static int alloc_tuples(bool pinned, int64_t m, int32_t** src, int32_t** dst) {
  const size_t bytes = (size_t)(m > 0 ? m : 1) * sizeof(int32_t);
  *src = *dst = nullptr;
  if (pinned) {
    if (cudaMallocHost(src, bytes) != cudaSuccess || cudaMallocHost(dst, bytes) != cudaSuccess) {
      cudaFreeHost(*src);
      *src = nullptr;
      return B2BFS_ERR_NOMEM;
    }
  } else {
    *src = static_cast<int32_t*>(malloc(bytes));
    *dst = static_cast<int32_t*>(malloc(bytes));
    if (!*src || !*dst) { free(*src); free(*dst); *src = *dst = nullptr; return B2BFS_ERR_NOMEM; }
  }
  return B2BFS_OK;
}
static void free_tuples(bool pinned, int32_t* src, int32_t* dst) {
  if (pinned) { cudaFreeHost(src); cudaFreeHost(dst); } else { free(src); free(dst); }
}

static int read_graph_file(const char* path, int format, bool pinned, int64_t* n_out, int64_t* m_out, int32_t** src,
                           int32_t** dst, std::string* err) {
  MappedFile f;
  int rc = map_file(path, &f, err);
  if (rc) return rc;
  if (format == kFormatAuto) format = detect_format(f);
  unsigned hc = std::thread::hardware_concurrency();
  const int threads = hc ? (int)hc : 4;
  int64_t n = 0, m = 0;
  *src = *dst = nullptr;
  if (format == kFormatReference) {
    size_t body = 0;
    rc = parse_header(f, &n, &m, &body, err);
    if (!rc && n < 1) { rc = B2BFS_ERR_FORMAT; *err = "vertex_num must be at least 1"; }
    if (!rc) rc = alloc_tuples(pinned, m, src, dst);
    if (!rc) rc = parse_body(f, body, n, m, *src, *dst, threads, err);
  } else if (format == kFormatMatrixMarket || format == kFormatSnap) {
    size_t body = 0;
    int64_t declared = 0;
    if (format == kFormatMatrixMarket) rc = parse_mtx_header(f, &n, &declared, &body, err);
    LinePlan plan;
    if (!rc) rc = plan_lines(f, body, format == kFormatSnap ? '#' : '%', format == kFormatMatrixMarket ? 1 : 0, threads, &plan, err);
    if (!rc && format == kFormatMatrixMarket && plan.m != declared) {
      rc = B2BFS_ERR_FORMAT;
      *err = "MatrixMarket file declares " + std::to_string(declared) + " entries but holds " + std::to_string(plan.m);
    }
    m = plan.m;
    if (!rc) rc = alloc_tuples(pinned, m, src, dst);
    int64_t max_id = -1;
    if (!rc) rc = fill_lines(f, plan, format == kFormatMatrixMarket ? n : 0, *src, *dst, &max_id, err);
    if (!rc && format == kFormatSnap) n = max_id + 1;
    if (!rc && n < 1) { rc = B2BFS_ERR_FORMAT; *err = "the file holds no edges"; }
  } else if (format == kFormatBinaryCache) {
    if (f.size < 24 || memcmp(f.data, "B2BFSEDG", 8) != 0) { rc = B2BFS_ERR_FORMAT; *err = "not a b2bfs edge cache"; }
    if (!rc) {
      memcpy(&n, f.data + 8, 8);
      memcpy(&m, f.data + 16, 8);
      // m is bounded by what the file can hold BEFORE any multiplication (a crafted m >= 2^61 would wrap)
      if (n < 1 || n > 0x7fffffffLL || m < 0 || (uint64_t)m > (uint64_t)(f.size - 24) / 8) { rc = B2BFS_ERR_FORMAT; *err = "edge cache is truncated"; }
    }
    if (!rc) rc = alloc_tuples(pinned, m, src, dst);
    if (!rc) {
      memcpy(*src, f.data + 24, (size_t)m * 4);
      memcpy(*dst, f.data + 24 + (size_t)m * 4, (size_t)m * 4);
      for (int64_t i = 0; i < m; ++i) {  // the text parsers range-check while they parse; do the same here
        if ((uint32_t)(*src)[i] >= (uint32_t)n || (uint32_t)(*dst)[i] >= (uint32_t)n) {
          rc = B2BFS_ERR_RANGE;
          *err = "edge cache holds vertex id " + std::to_string((uint32_t)(*src)[i] >= (uint32_t)n ? (*src)[i] : (*dst)[i]) +
                 " >= vertex_num " + std::to_string(n);
          break;
        }
      }
    }
  } else {
    rc = B2BFS_ERR_ARG;
    *err = "unknown graph file format";
  }
  unmap_file(&f);
  if (rc == B2BFS_ERR_NOMEM && err->empty()) *err = "cannot allocate host memory for the edge list";
  if (rc) { free_tuples(pinned, *src, *dst); *src = *dst = nullptr; return rc; }
  *n_out = n;
  *m_out = m;
  return B2BFS_OK;
}

extern "C" {

const char* b2bfs_version(void) { return "b2bfs 0.1 (sm_100a)"; }

int b2bfs_create(b2bfs_t** out, int device) {
  if (!out) return fail(nullptr, B2BFS_ERR_ARG, "b2bfs_create: null handle pointer");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(nullptr, B2BFS_ERR_CUDA,
                std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                    " (this library has no CPU path)");
  if (device < 0 || device >= count) return fail(nullptr, B2BFS_ERR_ARG, "b2bfs_create: device index out of range");
  b2bfs* h = new b2bfs();
  h->device = device;
  // any failure from here on releases what was created so far (b2bfs_destroy copes with a half-built handle)
  auto create = [&]() -> int {
    API_CUDA(nullptr, cudaSetDevice(device));
    int coop = 0;
    API_CUDA(nullptr, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
    if (!coop) return fail(nullptr, B2BFS_ERR_UNSUPPORTED, "device lacks cooperative launch");
    API_CUDA(nullptr, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    API_CUDA(nullptr, cudaEventCreate(&h->ev0));
    API_CUDA(nullptr, cudaEventCreate(&h->ev1));
    API_CUDA(nullptr, cudaMallocHost(&h->pinned_ctl, sizeof(Control)));
    Status s = search_grid_size(device, &h->grid);
    if (s.ok()) s = search_cluster_grid_size(h->grid, &h->grid_cluster, &h->cluster_size);
    return s.ok() ? B2BFS_OK : fail(nullptr, s);
  };
  const int rc = create();
  if (rc) {
    const std::string msg = g_create_error;  // b2bfs_destroy must not disturb the message
    b2bfs_destroy(h);
    g_create_error = msg;
    return rc;
  }
  *out = h;
  return B2BFS_OK;
}

int b2bfs_comm_unique_id(void* out128) {
  if (!out128) return fail(nullptr, B2BFS_ERR_ARG, "b2bfs_comm_unique_id: null buffer");
  const char* err = nullptr;
  const NcclApi* nccl = nccl_api(&err);
  if (!nccl) return fail(nullptr, B2BFS_ERR_COMM, err);
  static_assert(sizeof(ncclUniqueId) == 128, "the ABI promises a 128-byte id");
  ncclUniqueId id;
  const ncclResult_t r = nccl->GetUniqueId(&id);
  if (r != ncclSuccess) return fail(nullptr, B2BFS_ERR_COMM, std::string("ncclGetUniqueId: ") + nccl->GetErrorString(r));
  memcpy(out128, &id, sizeof(id));
  return B2BFS_OK;
}

int b2bfs_create_rank(b2bfs_t** out, int device, int rank, int world, const void* nccl_id) {
  if (world == 1 && rank == 0) return b2bfs_create(out, device);
  if (!out || !nccl_id || world < 2 || world > 8 || rank < 0 || rank >= world)
    return fail(nullptr, B2BFS_ERR_ARG, "b2bfs_create_rank: need 2 <= world <= 8, 0 <= rank < world and a unique id");
  const char* err = nullptr;
  const NcclApi* nccl = nccl_api(&err);
  if (!nccl) return fail(nullptr, B2BFS_ERR_COMM, err);
  int rc = b2bfs_create(out, device);
  if (rc) return rc;
  b2bfs* h = *out;
  h->rank = rank;
  h->world = world;
  h->mc.nccl = nccl;
  h->mc.rank = rank;
  h->mc.world = world;
  ncclUniqueId id;
  memcpy(&id, nccl_id, sizeof(id));
  const ncclResult_t r = nccl->CommInitRank(&h->mc.comm, world, id, rank);
  if (r != ncclSuccess) {
    const std::string msg = std::string("ncclCommInitRank: ") + nccl->GetErrorString(r);
    b2bfs_destroy(h);
    *out = nullptr;
    return fail(nullptr, B2BFS_ERR_COMM, msg);
  }
  Status gs = multi_grid_size(device, &h->grid_multi);
  if (!gs.ok()) {
    b2bfs_destroy(h);
    *out = nullptr;
    return fail(nullptr, gs);
  }
  cudaError_t e = cudaMalloc(&h->mc.stats_send, 16 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&h->mc.stats_all, (size_t)world * 16 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMallocHost(&h->mc.stats_host, (size_t)world * 16 * sizeof(unsigned long long));
  if (e != cudaSuccess) {
    b2bfs_destroy(h);
    *out = nullptr;
    return fail(nullptr, B2BFS_ERR_NOMEM, "cannot allocate the per-level exchange records");
  }
  return B2BFS_OK;
}

int b2bfs_set_option(b2bfs_t* h, const char* key, int64_t value) {
  if (!h || !key) return fail(h, B2BFS_ERR_ARG, "b2bfs_set_option: null argument");
  const std::string k(key);
  const bool graph_time = (k == "dedup" || k == "keep_edges" || k == "hub_first" || k == "directed");
  if (graph_time && h->has_graph) return fail(h, B2BFS_ERR_STATE, "option '" + k + "' must be set before loading a graph");
  if (k == "dedup" && (value == 0 || value == 1)) h->dedup = (int)value;
  else if (k == "keep_edges" && (value == 0 || value == 1)) h->keep_edges = (int)value;
  else if (k == "hub_first" && (value == 0 || value == 1)) h->hub_first = (int)value;
  else if (k == "directed" && (value == 0 || value == 1)) {
    if (value && h->world > 1) return fail(h, B2BFS_ERR_UNSUPPORTED, "directed graphs: single-GPU handles only");
    h->directed = (int)value;
  }
  else if (k == "mode" && (value == B2BFS_MODE_DIRECTION_OPT || value == B2BFS_MODE_TOP_DOWN)) h->mode = (int)value;
  else if (k == "engine" && value >= B2BFS_ENGINE_FUSED && value <= B2BFS_ENGINE_PULL) h->engine = (int)value;
  else if (k == "gather_parent" && (value == 0 || value == 1)) h->gather_parent = (int)value;
  else if (k == "multi_engine" && (value == 0 || value == 1)) h->multi_engine = (int)value;
  else if (k == "alpha" && value >= 1 && value <= (1 << 20)) h->alpha = (int)value;
  else if (k == "beta" && value >= 1 && value <= (1 << 20)) h->beta = (int)value;
  else if (k == "gamma" && value >= 0 && value <= (1 << 20)) h->gamma = (int)value;
  else if (k == "xfullpush" && (value == 0 || value == 1)) h->xfullpush = (int)value;
  else if (k == "cluster" && (value == 0 || value == 1)) h->use_cluster = (int)value;
  else if (k == "huge_min" && (value == 0 || (value >= 2 * kCtaChunk && value <= (1 << 30)))) h->huge_min = (int)value;
  else return fail(h, B2BFS_ERR_ARG, "unknown option or value out of range: " + k);
  return B2BFS_OK;
}

int b2bfs_get_option(const b2bfs_t* h, const char* key, int64_t* value) {
  if (!h || !key || !value) return B2BFS_ERR_ARG;
  const std::string k(key);
  if (k == "dedup") *value = h->dedup;
  else if (k == "keep_edges") *value = h->keep_edges;
  else if (k == "hub_first") *value = h->hub_first;
  else if (k == "directed") *value = h->directed;
  else if (k == "mode") *value = h->mode;
  else if (k == "engine") *value = h->engine;
  else if (k == "alpha") *value = h->alpha;
  else if (k == "beta") *value = h->beta;
  else if (k == "gamma") *value = h->gamma;
  else if (k == "xfullpush") *value = h->xfullpush;
  else if (k == "huge_min") *value = h->huge_min;
  else if (k == "cluster") *value = (h->world == 1 && h->use_cluster && h->grid_cluster > 0 && h->has_graph_rows_narrow()) ? h->cluster_size : 0;
  else if (k == "grid_cluster") *value = h->grid_cluster;
  else if (k == "grid") *value = h->grid;
  else if (k == "gather_parent") *value = h->gather_parent;
  else if (k == "multi_engine") *value = (h->world > 1 && h->p2p_ready && h->multi_engine == 0) ? 0 : 1;
  else if (k == "rank") *value = h->rank;
  else if (k == "world") *value = h->world;
  else if (k == "local_rows") *value = h->nloc;
  else if (k == "part") *value = h->part;
  else if (k == "tuples_first") *value = h->m_first;
  else if (k == "tuples_count") *value = h->m_loc;
  else if (k == "wide_items") *value = h->csr.wide_items;
  else if (k == "max_degree") *value = h->csr.max_degree;
  else if (k == "build_launches") *value = h->build_launches;
  else return B2BFS_ERR_ARG;
  return B2BFS_OK;
}

int b2bfs_load_edges(b2bfs_t* h, int64_t n, int64_t m, const int32_t* src, const int32_t* dst) {
  if (!h) return fail(h, B2BFS_ERR_ARG, "null handle");
  if (n < 1 || n > 0x7fffffffLL || m < 0 || (m > 0 && (!src || !dst)))
    return fail(h, B2BFS_ERR_ARG, "b2bfs_load_edges: need 1 <= n < 2^31, m >= 0 and non-null arrays");
  API_CUDA(h, cudaSetDevice(h->device));
  free_graph(h);
  h->n = n;
  h->m = m;
  set_tuple_share(h);  // several ranks: every rank is handed all m tuples (as every MPI rank of the reference parses
                       // the whole file, bfs_mpi.cu:815) but takes only its share of them to the device
  const size_t bytes = (size_t)(h->m_loc > 0 ? h->m_loc : 1) * sizeof(int32_t);
  API_CUDA(h, cudaMalloc(&h->d_src, bytes));
  API_CUDA(h, cudaMalloc(&h->d_dst, bytes));
  if (h->m_loc > 0) {
    API_CUDA(h, cudaMemcpyAsync(h->d_src, src + h->m_first, (size_t)h->m_loc * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    API_CUDA(h, cudaMemcpyAsync(h->d_dst, dst + h->m_first, (size_t)h->m_loc * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
  }
  return finish_load(h);
}

int b2bfs_load_graph_file(b2bfs_t* h, const char* path, int format) {
  if (!h || !path) return fail(h, B2BFS_ERR_ARG, "b2bfs_load_graph_file: null argument");
  int64_t n = 0, m = 0;
  int32_t *src = nullptr, *dst = nullptr;
  std::string err;
  int rc = read_graph_file(path, format, true, &n, &m, &src, &dst, &err);
  if (rc) return fail(h, rc, err);
  rc = b2bfs_load_edges(h, n, m, src, dst);
  free_tuples(true, src, dst);
  return rc;
}

int b2bfs_load_edge_file(b2bfs_t* h, const char* path) { return b2bfs_load_graph_file(h, path, kFormatReference); }

int b2bfs_read_graph_file(const char* path, int format, int64_t* n, int64_t* m, int32_t** src, int32_t** dst) {
  if (!path || !n || !m || !src || !dst) return fail(nullptr, B2BFS_ERR_ARG, "b2bfs_read_graph_file: null argument");
  std::string err;
  const int rc = read_graph_file(path, format, false, n, m, src, dst, &err);
  return rc ? fail(nullptr, rc, err) : B2BFS_OK;
}

void b2bfs_free_edges(int32_t* src, int32_t* dst) { free_tuples(false, src, dst); }

int b2bfs_save_edge_cache(b2bfs_t* h, const char* path) {
  if (!h || !path) return fail(h, B2BFS_ERR_ARG, "b2bfs_save_edge_cache: null argument");
  if (!h->has_graph || !h->d_src) return fail(h, B2BFS_ERR_STATE, "input tuples are not resident (keep_edges=0?)");
  if (h->world > 1) return fail(h, B2BFS_ERR_UNSUPPORTED, "a rank of a multi-GPU job holds only its share of the tuples: save the cache from a single-GPU handle");
  API_CUDA(h, cudaSetDevice(h->device));
  FILE* f = fopen(path, "wb");
  if (!f) return fail(h, B2BFS_ERR_IO, std::string("cannot write ") + path);
  std::vector<int32_t> buf((size_t)(h->m > 0 ? h->m : 1));
  bool ok = fwrite("B2BFSEDG", 1, 8, f) == 8 && fwrite(&h->n, 8, 1, f) == 1 && fwrite(&h->m, 8, 1, f) == 1;
  for (int pass = 0; pass < 2 && ok; ++pass) {
    if (cudaMemcpy(buf.data(), pass ? h->d_dst : h->d_src, (size_t)h->m * 4, cudaMemcpyDeviceToHost) != cudaSuccess) ok = false;
    else ok = h->m == 0 || fwrite(buf.data(), 4, (size_t)h->m, f) == (size_t)h->m;
  }
  ok = (fclose(f) == 0) && ok;
  return ok ? B2BFS_OK : fail(h, B2BFS_ERR_IO, std::string("short write to ") + path);
}

int b2bfs_generate(b2bfs_t* h, int kind, int scale, int edgefactor, uint64_t seed) {
  if (!h) return fail(h, B2BFS_ERR_ARG, "null handle");
  if (kind < B2BFS_KIND_RMAT || kind > B2BFS_KIND_GRID || scale < 2 || scale > 30 ||
      (kind != B2BFS_KIND_GRID && (edgefactor < 1 || edgefactor > 1024)))
    return fail(h, B2BFS_ERR_ARG, "b2bfs_generate: need kind in {RMAT,KRON,GRID}, 2 <= scale <= 30, 1 <= edgefactor <= 1024");
  API_CUDA(h, cudaSetDevice(h->device));
  free_graph(h);
  h->n = b2_gen_num_vertices(kind, scale);
  h->m = b2_gen_num_edges(kind, scale, edgefactor);
  set_tuple_share(h);  // several ranks: each generates its own m / world tuples of the stream
  API_CUDA(h, cudaMalloc(&h->d_src, (size_t)(h->m_loc > 0 ? h->m_loc : 1) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->d_dst, (size_t)(h->m_loc > 0 ? h->m_loc : 1) * sizeof(int32_t)));
  Status s = generate_edges_device(kind, scale, edgefactor, seed, h->m_first, h->m_loc, h->d_src, h->d_dst, h->stream);
  if (!s.ok()) { free_graph(h); return fail(h, s); }
  h->build_launches += 1;
  return finish_load(h);
}

int b2bfs_num_vertices(const b2bfs_t* h, int64_t* n) {
  if (!h || !n) return B2BFS_ERR_ARG;
  *n = h->n;
  return B2BFS_OK;
}
int b2bfs_num_edges(const b2bfs_t* h, int64_t* two_m) {
  if (!h || !two_m) return B2BFS_ERR_ARG;
  *two_m = h->directed ? h->m : 2 * h->m;  // G.numEdges: adjacency entries (bfs.cu:875; readGraph bfs.cu:919)
  return B2BFS_OK;
}
int b2bfs_num_input_edges(const b2bfs_t* h, int64_t* m) {
  if (!h || !m) return B2BFS_ERR_ARG;
  *m = h->m;
  return B2BFS_OK;
}
int b2bfs_csr_entries(const b2bfs_t* h, int64_t* entries) {
  if (!h || !entries) return B2BFS_ERR_ARG;
  *entries = h->csr.entries;
  return B2BFS_OK;
}

int b2bfs_copy_degrees(b2bfs_t* h, int32_t* deg_out) {
  if (!h || !deg_out) return fail(h, B2BFS_ERR_ARG, "null argument");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "no graph resident");
  API_CUDA(h, cudaSetDevice(h->device));
  if (h->world > 1)  // collective: the reference's edgesSize[] for all n vertices, on every rank that asks
    return gather_i32(h, reinterpret_cast<const int32_t*>(h->csr.orig_deg), deg_out, nullptr);
  API_CUDA(h, cudaMemcpy(deg_out, h->csr.orig_deg, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return B2BFS_OK;
}

int b2bfs_copy_row_degrees(b2bfs_t* h, int32_t* deg_out) {
  if (!h) return fail(h, B2BFS_ERR_ARG, "null argument");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "no graph resident");
  API_CUDA(h, cudaSetDevice(h->device));
  if (h->world > 1) {
    int32_t* mine = h->gather_tmp + (int64_t)h->rank * h->part;  // in-place allgather: own slice first
    API_CUDA(h, cudaMemsetAsync(mine, 0, (size_t)h->part * sizeof(int32_t), h->stream));
    Status s = row_degrees_device(h->csr.rowptr, h->nloc, mine, h->stream);
    if (!s.ok()) return fail(h, s);
    return gather_i32(h, mine, deg_out, nullptr);
  }
  if (!deg_out) return fail(h, B2BFS_ERR_ARG, "null argument");
  int32_t* tmp = nullptr;
  API_CUDA(h, cudaMalloc(&tmp, (size_t)h->n * sizeof(int32_t)));
  Status s = row_degrees_device(h->csr.rowptr, h->n, tmp, h->stream);
  if (s.ok() && cudaMemcpyAsync(deg_out, tmp, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess)
    s = make_error(B2BFS_ERR_CUDA, "copy of row degrees failed");
  cudaStreamSynchronize(h->stream);
  cudaFree(tmp);
  return s.ok() ? B2BFS_OK : fail(h, s);
}

int b2bfs_copy_edges(b2bfs_t* h, int64_t first, int64_t count, int32_t* src_out, int32_t* dst_out) {
  if (!h || !src_out || !dst_out) return fail(h, B2BFS_ERR_ARG, "null argument");
  if (!h->has_graph || !h->d_src) return fail(h, B2BFS_ERR_STATE, "input tuples are not resident (keep_edges=0?)");
  if (first < 0 || count < 0 || first + count > h->m) return fail(h, B2BFS_ERR_ARG, "tuple range out of bounds");
  if (first < h->m_first || first + count > h->m_first + h->m_loc)
    return fail(h, B2BFS_ERR_ARG, "tuple range not held by this rank: it keeps tuples [" + std::to_string(h->m_first) + ", " +
                                      std::to_string(h->m_first + h->m_loc) + ") (options \"tuples_first\" / \"tuples_count\")");
  API_CUDA(h, cudaSetDevice(h->device));
  API_CUDA(h, cudaMemcpy(src_out, h->d_src + (first - h->m_first), (size_t)count * sizeof(int32_t), cudaMemcpyDeviceToHost));
  API_CUDA(h, cudaMemcpy(dst_out, h->d_dst + (first - h->m_first), (size_t)count * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return B2BFS_OK;
}

int b2bfs_copy_csr(b2bfs_t* h, int64_t* rowptr_out, int32_t* col_out) {
  if (!h || !rowptr_out) return fail(h, B2BFS_ERR_ARG, "null argument");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "no graph resident");
  API_CUDA(h, cudaSetDevice(h->device));
  // multi-rank: this rank's rows only (local_rows + 1 offsets; col holds global ids)
  API_CUDA(h, cudaMemcpy(rowptr_out, h->csr.rowptr, (size_t)(h->nloc + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
  if (col_out && h->csr.entries > 0)
    API_CUDA(h, cudaMemcpy(col_out, h->csr.col, (size_t)h->csr.entries * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return B2BFS_OK;
}

enum GatherMode { kGatherNone = 0, kGatherAll = 1, kGatherSlice = 2 };

static int run_impl(b2bfs_t* h, int64_t src, int32_t* depth_out, int32_t* parent_out, b2bfs_stats* stats,
                    GatherMode gather) {
  if (!h) return fail(h, B2BFS_ERR_ARG, "null handle");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "b2bfs_run: no graph resident");
  if (src < 0 || src >= h->n) return fail(h, B2BFS_ERR_ARG, "b2bfs_run: source vertex outside [0, vertex_num)");
  const auto t0 = std::chrono::steady_clock::now();
  API_CUDA(h, cudaSetDevice(h->device));
  const BfsParams p = make_params(h);
  const bool multi = h->world > 1;
  int64_t launches = 0;
  MultiResult mr;
  API_CUDA(h, cudaEventRecord(h->ev0, h->stream));
  const bool p2p = multi && p.s.xbase[0] != nullptr;
  const bool comparator = !multi && (h->engine == B2BFS_ENGINE_SCAN || h->engine == B2BFS_ENGINE_PULL);
  const size_t cmp_totals_at = (size_t)h->n + 4;
  if (comparator && !h->cmp_scratch)
    API_CUDA(h, cudaMalloc(&h->cmp_scratch, (cmp_totals_at + (size_t)h->n / 1024 + 16) * sizeof(unsigned int)));
  Status s = comparator ? (h->engine == B2BFS_ENGINE_SCAN
                               ? run_scan_engine(p, src, h->grid, h->stream, h->pinned_ctl, h->cmp_scratch,
                                                 h->cmp_scratch + cmp_totals_at, &launches)
                               : run_pull_engine(p, src, h->grid, h->stream, h->pinned_ctl,
                                                 reinterpret_cast<int*>(h->cmp_scratch), &launches))
           : p2p ? run_fused_multi(p, src, h->grid_multi, h->stream, &mr.launches)
           : multi ? run_multi(p, h->mc, src, h->grid, h->stream, &mr)
           : (h->engine == B2BFS_ENGINE_FUSED) ? run_fused(p, src, p.tn.cluster ? h->grid_cluster : h->grid, h->stream, &launches)
                                               : run_stepwise(p, src, h->grid, h->stream, h->pinned_ctl, &launches);
  if (!s.ok()) return fail(h, s);
  API_CUDA(h, cudaEventRecord(h->ev1, h->stream));
  int64_t d2h = 0;
  if (multi && gather == kGatherSlice) {
    // this rank's rows only, over this GPU's own PCIe link: no cross-GPU gather at all
    launches = mr.launches;
    if (depth_out && h->nloc > 0) {
      API_CUDA(h, cudaMemcpyAsync(depth_out, h->sv.depth, (size_t)h->nloc * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
      d2h += h->nloc * (int64_t)sizeof(int32_t);
    }
    if (parent_out && h->nloc > 0) {
      API_CUDA(h, cudaMemcpyAsync(parent_out, h->sv.parent, (size_t)h->nloc * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
      d2h += h->nloc * (int64_t)sizeof(int32_t);
    }
  } else if (multi) {
    launches = mr.launches;
    if (gather == kGatherAll) {  // collective: every rank takes part, whoever passed a buffer gets all n entries
      int rc = gather_i32(h, h->sv.depth, depth_out, &d2h);
      if (rc) return rc;
      if (h->gather_parent) {
        rc = gather_i32(h, h->sv.parent, parent_out, &d2h);
        if (rc) return rc;
      }
    }
  } else {
    if (depth_out) {
      API_CUDA(h, cudaMemcpyAsync(depth_out, h->sv.depth, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
      d2h += h->n * (int64_t)sizeof(int32_t);
    }
    if (parent_out) {
      API_CUDA(h, cudaMemcpyAsync(parent_out, h->sv.parent, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
      d2h += h->n * (int64_t)sizeof(int32_t);
    }
  }
  if (stats) {
    s = run_reach_stats(p, h->csr.orig_deg, h->stream, &launches);
    if (!s.ok()) return fail(h, s);
    if (multi)
      API_NCCL(h, h->mc.nccl->AllReduce(&h->sv.ctl->reached_vertices, &h->sv.ctl->reached_vertices, 2, ncclUint64, ncclSum,
                                        h->mc.comm, h->stream));
    API_CUDA(h, cudaMemcpyAsync(h->pinned_ctl, h->sv.ctl, sizeof(Control), cudaMemcpyDeviceToHost, h->stream));
    d2h += (int64_t)sizeof(Control);
  }
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  API_CUDA(h, cudaGetLastError());
  if (stats) {
    float ms = 0.f;
    API_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    const Control* c = h->pinned_ctl;
    if (p2p && c->error) return fail(h, B2BFS_ERR_COMM, "a cross-GPU wait timed out inside the search kernel (peer rank missing?)");
    const bool host_loop = multi && !p2p;
    stats->levels = host_loop ? mr.levels : c->result.levels;
    stats->td_levels = host_loop ? mr.td_levels : c->result.td_levels;
    stats->bu_levels = host_loop ? mr.bu_levels : c->result.bu_levels;
    stats->reserved0 = 0;
    stats->reached_vertices = (int64_t)c->reached_vertices;
    stats->reached_entries = (int64_t)c->reached_entries;
    stats->kernel_launches = launches;
    stats->h2d_bytes = (!multi && h->engine == B2BFS_ENGINE_STEPWISE) ? (int64_t)sizeof(SearchResult) : 0;
    stats->d2h_bytes = d2h;
    stats->search_ms = ms;
    stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  return B2BFS_OK;
}

int b2bfs_run(b2bfs_t* h, int64_t src, int32_t* depth_out, int32_t* parent_out, b2bfs_stats* stats) {
  if (h && h->world == 1 && !depth_out)
    return fail(h, B2BFS_ERR_ARG, "b2bfs_run: depth_out is null (use b2bfs_run_device to keep results in HBM)");
  return run_impl(h, src, depth_out, parent_out, stats, kGatherAll);
}

int b2bfs_run_device(b2bfs_t* h, int64_t src, b2bfs_stats* stats) {
  return run_impl(h, src, nullptr, nullptr, stats, kGatherNone);
}

int b2bfs_run_slice(b2bfs_t* h, int64_t src, int32_t* depth_slice_out, int32_t* parent_slice_out, b2bfs_stats* stats) {
  if (h && !depth_slice_out) return fail(h, B2BFS_ERR_ARG, "b2bfs_run_slice: depth_slice_out is null");
  return run_impl(h, src, depth_slice_out, parent_slice_out, stats, h && h->world > 1 ? kGatherSlice : kGatherAll);
}

int b2bfs_local_range(const b2bfs_t* h, int64_t* first_vertex, int64_t* count) {
  if (!h || !first_vertex || !count || !h->has_graph) return B2BFS_ERR_ARG;
  *first_vertex = h->vbase;
  *count = h->nloc;
  return B2BFS_OK;
}

int b2bfs_run_batch(b2bfs_t* h, int nroots, const int64_t* roots, int32_t* depth_out, b2bfs_batch_stats* stats) {
  if (!h || !roots) return fail(h, B2BFS_ERR_ARG, "b2bfs_run_batch: null argument");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "b2bfs_run_batch: no graph resident");
  if (nroots < 1 || nroots > B2BFS_BATCH_MAX) return fail(h, B2BFS_ERR_ARG, "b2bfs_run_batch: need 1 <= nroots <= 64");
  if (h->world > 1) return fail(h, B2BFS_ERR_UNSUPPORTED, "b2bfs_run_batch: single-GPU handles only");
  if (h->directed) return fail(h, B2BFS_ERR_UNSUPPORTED, "b2bfs_run_batch pulls over in-edges: not available for a directed graph");
  for (int i = 0; i < nroots; ++i)
    if (roots[i] < 0 || roots[i] >= h->n) return fail(h, B2BFS_ERR_ARG, "b2bfs_run_batch: source vertex outside [0, vertex_num)");
  const auto t0 = std::chrono::steady_clock::now();
  API_CUDA(h, cudaSetDevice(h->device));
  int64_t launches = 0;
  Status s = prepare_batch(&h->batch, h->n, h->csr.wide_items, h->csr.rowptr, nroots, h->stream, &launches);
  if (!s.ok()) { free_batch(&h->batch); return fail(h, s); }
  int levels = 0;
  s = run_batch(&h->batch, h->n, h->csr.rowptr, h->csr.col, nroots, roots, h->grid, h->stream, h->ev0, h->ev1, &levels, &launches);
  if (!s.ok()) return fail(h, s);
  int64_t d2h = 0;
  if (depth_out) {
    API_CUDA(h, cudaMemcpyAsync(depth_out, h->batch.depth, (size_t)nroots * (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    d2h += (int64_t)nroots * h->n * (int64_t)sizeof(int32_t);
  }
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  if (stats) {
    unsigned long long vd[128];
    s = batch_reach_stats(&h->batch, h->n, h->csr.orig_deg, h->stream, vd, &launches);
    if (!s.ok()) return fail(h, s);
    float ms = 0.f;
    API_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    memset(stats, 0, sizeof(*stats));
    stats->nroots = nroots;
    stats->levels = levels;
    stats->kernel_launches = launches;
    stats->d2h_bytes = d2h;
    stats->search_ms = ms;
    for (int i = 0; i < nroots; ++i) {
      stats->reached_vertices[i] = (int64_t)vd[i];
      stats->reached_entries[i] = (int64_t)vd[64 + i];
    }
    stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  return B2BFS_OK;
}

int b2bfs_device_batch_depth(const b2bfs_t* h, const int32_t** depth_dev) {
  if (!h || !depth_dev || !h->batch.depth) return B2BFS_ERR_ARG;
  *depth_dev = h->batch.depth;
  return B2BFS_OK;
}

int b2bfs_device_depth(const b2bfs_t* h, const int32_t** depth_dev) {
  if (!h || !depth_dev || !h->has_graph) return B2BFS_ERR_ARG;
  *depth_dev = h->sv.depth;
  return B2BFS_OK;
}
int b2bfs_device_parent(const b2bfs_t* h, const int32_t** parent_dev) {
  if (!h || !parent_dev || !h->has_graph) return B2BFS_ERR_ARG;
  *parent_dev = h->sv.parent;
  return B2BFS_OK;
}

int b2bfs_last_trace(b2bfs_t* h, int32_t* level_out, int32_t* phase_out, double* us_out, int cap) {
  if (!h || !h->has_graph || cap < 0 || (cap > 0 && (!level_out || !phase_out || !us_out))) return B2BFS_ERR_ARG;
  if (cudaSetDevice(h->device) != cudaSuccess) return B2BFS_ERR_CUDA;
  if (cudaMemcpy(h->pinned_ctl, h->sv.ctl, sizeof(Control), cudaMemcpyDeviceToHost) != cudaSuccess) return B2BFS_ERR_CUDA;
  const Control* c = h->pinned_ctl;
  int n = c->trace_len < cap ? c->trace_len : cap;
  if (n > kTraceCap) n = kTraceCap;
  for (int i = 0; i < n; ++i) {
    level_out[i] = c->trace[i].level;
    phase_out[i] = c->trace[i].phase;
    us_out[i] = (double)(c->trace[i].t_ns - c->trace[0].t_ns) * 1e-3;
  }
  return n;
}

int b2bfs_host_alloc(void** p, int64_t bytes) {
  if (!p || bytes < 0) return B2BFS_ERR_ARG;
  return cudaMallocHost(p, (size_t)(bytes > 0 ? bytes : 1)) == cudaSuccess ? B2BFS_OK : B2BFS_ERR_NOMEM;
}
int b2bfs_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? B2BFS_OK : B2BFS_ERR_CUDA; }
int b2bfs_host_register(void* p, int64_t bytes) {
  if (!p || bytes <= 0) return B2BFS_ERR_ARG;
  const cudaError_t e = cudaHostRegister(p, (size_t)bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) cudaGetLastError();
  return e == cudaSuccess ? B2BFS_OK : (e == cudaErrorMemoryAllocation ? B2BFS_ERR_NOMEM : B2BFS_ERR_CUDA);
}
int b2bfs_host_unregister(void* p) { return cudaHostUnregister(p) == cudaSuccess ? B2BFS_OK : B2BFS_ERR_CUDA; }

int b2bfs_destroy(b2bfs_t* h) {
  if (!h) return B2BFS_OK;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  free_graph(h);
  if (h->mc.comm) h->mc.nccl->CommDestroy(h->mc.comm);
  cudaFree(h->mc.stats_send);
  cudaFree(h->mc.stats_all);
  if (h->mc.stats_host) cudaFreeHost(h->mc.stats_host);
  if (h->pinned_ctl) cudaFreeHost(h->pinned_ctl);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return B2BFS_OK;
}

const char* b2bfs_last_error(const b2bfs_t* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

}  // extern "C"
