// api.cu — the extern "C" boundary declared in include/b2bfs.h.
//
// The handle owns what the reference keeps in globals: the Graph (bfs.cu:28), the __managed__
// per-device pointer tables (bfs.cu:255-269) and the buffers initCuda2 allocates and never frees
// (bfs.cu:333-351, 362-372).
#include <chrono>
#include <cstring>
#include <thread>

#include "../../include/b2bfs_graphgen.h"
#include "bfs.cuh"
#include "build.cuh"
#include "edge_file.hpp"

using namespace b2;

struct b2bfs {
  int device = 0;
  int rank = 0, world = 1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  // options
  int dedup = 1, mode = B2BFS_MODE_DIRECTION_OPT, engine = B2BFS_ENGINE_FUSED;
  int alpha = 14, beta = 24, keep_edges = 1;
  // graph
  bool has_graph = false;
  int64_t n = 0, m = 0;
  int32_t *d_src = nullptr, *d_dst = nullptr;
  BuiltCsr csr;
  // search state
  SearchView sv{};
  Control* pinned_ctl = nullptr;
  int grid = 0;
  int64_t build_launches = 0;
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
  cudaFree(h->sv.ctl);
  h->sv = SearchView{};
}

static void free_graph(b2bfs* h) {
  cudaFree(h->d_src);
  cudaFree(h->d_dst);
  h->d_src = h->d_dst = nullptr;
  free_csr(&h->csr);
  free_search_state(h);
  h->has_graph = false;
  h->n = h->m = 0;
}

// edges are in d_src/d_dst: build the CSR and the per-search buffers
static int finish_load(b2bfs* h) {
  Status s = build_csr_device(h->d_src, h->d_dst, h->m, h->n, 0, h->n, false, h->dedup, h->stream,
                              &h->csr, &h->build_launches);
  if (!s.ok()) {
    free_graph(h);
    return fail(h, s);
  }
  if (!h->keep_edges) {
    cudaFree(h->d_src);
    cudaFree(h->d_dst);
    h->d_src = h->d_dst = nullptr;
  }
  const int64_t n = h->n, nw = h->csr.nwords;
  const size_t wide_cap = (size_t)h->csr.wide_items + 1;
  API_CUDA(h, cudaMalloc(&h->sv.depth, (size_t)(n + 4) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.parent, (size_t)(n + 4) * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.visited, (size_t)nw * sizeof(uint32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.bitmap[0], (size_t)nw * sizeof(uint32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.bitmap[1], (size_t)nw * sizeof(uint32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.queue[0], (size_t)n * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.queue[1], (size_t)n * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->sv.wideq[0], wide_cap * sizeof(int2)));
  API_CUDA(h, cudaMalloc(&h->sv.wideq[1], wide_cap * sizeof(int2)));
  API_CUDA(h, cudaMalloc(&h->sv.ctl, sizeof(Control)));
  API_CUDA(h, cudaMemsetAsync(h->sv.ctl, 0, sizeof(Control), h->stream));
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  h->has_graph = true;
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
  p.s = h->sv;
  p.tn.mode = h->mode;
  p.tn.alpha = h->alpha;
  p.tn.beta = h->beta;
  return p;
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
  API_CUDA(nullptr, cudaSetDevice(device));
  int coop = 0;
  API_CUDA(nullptr, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
  if (!coop) { delete h; return fail(nullptr, B2BFS_ERR_UNSUPPORTED, "device lacks cooperative launch"); }
  API_CUDA(nullptr, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  API_CUDA(nullptr, cudaEventCreate(&h->ev0));
  API_CUDA(nullptr, cudaEventCreate(&h->ev1));
  API_CUDA(nullptr, cudaMallocHost(&h->pinned_ctl, sizeof(Control)));
  Status s = search_grid_size(device, &h->grid);
  if (!s.ok()) { delete h; return fail(nullptr, s); }
  *out = h;
  return B2BFS_OK;
}

int b2bfs_comm_unique_id(void* out128) {
  (void)out128;
  return fail(nullptr, B2BFS_ERR_UNSUPPORTED, "multi-rank support is not built into this library yet");
}

int b2bfs_create_rank(b2bfs_t** out, int device, int rank, int world, const void* nccl_id) {
  (void)nccl_id;
  if (world == 1 && rank == 0) return b2bfs_create(out, device);
  return fail(nullptr, B2BFS_ERR_UNSUPPORTED, "multi-rank support is not built into this library yet");
}

int b2bfs_set_option(b2bfs_t* h, const char* key, int64_t value) {
  if (!h || !key) return fail(h, B2BFS_ERR_ARG, "b2bfs_set_option: null argument");
  const std::string k(key);
  const bool graph_time = (k == "dedup" || k == "keep_edges");
  if (graph_time && h->has_graph) return fail(h, B2BFS_ERR_STATE, "option '" + k + "' must be set before loading a graph");
  if (k == "dedup" && (value == 0 || value == 1)) h->dedup = (int)value;
  else if (k == "keep_edges" && (value == 0 || value == 1)) h->keep_edges = (int)value;
  else if (k == "mode" && (value == B2BFS_MODE_DIRECTION_OPT || value == B2BFS_MODE_TOP_DOWN)) h->mode = (int)value;
  else if (k == "engine" && (value == B2BFS_ENGINE_FUSED || value == B2BFS_ENGINE_STEPWISE)) h->engine = (int)value;
  else if (k == "alpha" && value >= 1 && value <= (1 << 20)) h->alpha = (int)value;
  else if (k == "beta" && value >= 1 && value <= (1 << 20)) h->beta = (int)value;
  else return fail(h, B2BFS_ERR_ARG, "unknown option or value out of range: " + k);
  return B2BFS_OK;
}

int b2bfs_get_option(const b2bfs_t* h, const char* key, int64_t* value) {
  if (!h || !key || !value) return B2BFS_ERR_ARG;
  const std::string k(key);
  if (k == "dedup") *value = h->dedup;
  else if (k == "keep_edges") *value = h->keep_edges;
  else if (k == "mode") *value = h->mode;
  else if (k == "engine") *value = h->engine;
  else if (k == "alpha") *value = h->alpha;
  else if (k == "beta") *value = h->beta;
  else if (k == "grid") *value = h->grid;
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
  const size_t bytes = (size_t)(m > 0 ? m : 1) * sizeof(int32_t);
  API_CUDA(h, cudaMalloc(&h->d_src, bytes));
  API_CUDA(h, cudaMalloc(&h->d_dst, bytes));
  if (m > 0) {
    API_CUDA(h, cudaMemcpyAsync(h->d_src, src, (size_t)m * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    API_CUDA(h, cudaMemcpyAsync(h->d_dst, dst, (size_t)m * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
  }
  return finish_load(h);
}

int b2bfs_load_edge_file(b2bfs_t* h, const char* path) {
  if (!h || !path) return fail(h, B2BFS_ERR_ARG, "b2bfs_load_edge_file: null argument");
  MappedFile f;
  std::string err;
  int rc = map_file(path, &f, &err);
  if (rc) return fail(h, rc, err);
  int64_t n = 0, m = 0;
  size_t body = 0;
  rc = parse_header(f, &n, &m, &body, &err);
  if (rc) { unmap_file(&f); return fail(h, rc, err); }
  if (n < 1) { unmap_file(&f); return fail(h, B2BFS_ERR_FORMAT, "vertex_num must be at least 1"); }
  int32_t *src = nullptr, *dst = nullptr;
  const size_t bytes = (size_t)(m > 0 ? m : 1) * sizeof(int32_t);
  if (cudaMallocHost(&src, bytes) != cudaSuccess || cudaMallocHost(&dst, bytes) != cudaSuccess) {
    cudaFreeHost(src);
    unmap_file(&f);
    return fail(h, B2BFS_ERR_NOMEM, "cannot allocate page-locked staging for the edge list");
  }
  unsigned hc = std::thread::hardware_concurrency();
  rc = parse_body(f, body, n, m, src, dst, hc ? (int)hc : 4, &err);
  unmap_file(&f);
  if (rc == B2BFS_OK) rc = b2bfs_load_edges(h, n, m, src, dst);
  else fail(h, rc, err);
  cudaFreeHost(src);
  cudaFreeHost(dst);
  return rc;
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
  API_CUDA(h, cudaMalloc(&h->d_src, (size_t)h->m * sizeof(int32_t)));
  API_CUDA(h, cudaMalloc(&h->d_dst, (size_t)h->m * sizeof(int32_t)));
  Status s = generate_edges_device(kind, scale, edgefactor, seed, 0, h->m, h->d_src, h->d_dst, h->stream);
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
  *two_m = 2 * h->m;
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
  API_CUDA(h, cudaMemcpy(deg_out, h->csr.orig_deg, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return B2BFS_OK;
}

int b2bfs_copy_edges(b2bfs_t* h, int64_t first, int64_t count, int32_t* src_out, int32_t* dst_out) {
  if (!h || !src_out || !dst_out) return fail(h, B2BFS_ERR_ARG, "null argument");
  if (!h->has_graph || !h->d_src) return fail(h, B2BFS_ERR_STATE, "input tuples are not resident (keep_edges=0?)");
  if (first < 0 || count < 0 || first + count > h->m) return fail(h, B2BFS_ERR_ARG, "tuple range out of bounds");
  API_CUDA(h, cudaSetDevice(h->device));
  API_CUDA(h, cudaMemcpy(src_out, h->d_src + first, (size_t)count * sizeof(int32_t), cudaMemcpyDeviceToHost));
  API_CUDA(h, cudaMemcpy(dst_out, h->d_dst + first, (size_t)count * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return B2BFS_OK;
}

int b2bfs_copy_csr(b2bfs_t* h, int64_t* rowptr_out, int32_t* col_out) {
  if (!h || !rowptr_out) return fail(h, B2BFS_ERR_ARG, "null argument");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "no graph resident");
  API_CUDA(h, cudaSetDevice(h->device));
  API_CUDA(h, cudaMemcpy(rowptr_out, h->csr.rowptr, (size_t)(h->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
  if (col_out && h->csr.entries > 0)
    API_CUDA(h, cudaMemcpy(col_out, h->csr.col, (size_t)h->csr.entries * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return B2BFS_OK;
}

static int run_impl(b2bfs_t* h, int64_t src, int32_t* depth_out, int32_t* parent_out, b2bfs_stats* stats) {
  if (!h) return fail(h, B2BFS_ERR_ARG, "null handle");
  if (!h->has_graph) return fail(h, B2BFS_ERR_STATE, "b2bfs_run: no graph resident");
  if (src < 0 || src >= h->n) return fail(h, B2BFS_ERR_ARG, "b2bfs_run: source vertex outside [0, vertex_num)");
  const auto t0 = std::chrono::steady_clock::now();
  API_CUDA(h, cudaSetDevice(h->device));
  const BfsParams p = make_params(h);
  int64_t launches = 0;
  API_CUDA(h, cudaEventRecord(h->ev0, h->stream));
  Status s = (h->engine == B2BFS_ENGINE_FUSED) ? run_fused(p, src, h->grid, h->stream, &launches)
                                                : run_stepwise(p, src, h->grid, h->stream, h->pinned_ctl, &launches);
  if (!s.ok()) return fail(h, s);
  API_CUDA(h, cudaEventRecord(h->ev1, h->stream));
  int64_t d2h = 0;
  if (depth_out) {
    API_CUDA(h, cudaMemcpyAsync(depth_out, h->sv.depth, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    d2h += h->n * (int64_t)sizeof(int32_t);
  }
  if (parent_out) {
    API_CUDA(h, cudaMemcpyAsync(parent_out, h->sv.parent, (size_t)h->n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    d2h += h->n * (int64_t)sizeof(int32_t);
  }
  if (stats) {
    s = run_reach_stats(p, h->csr.orig_deg, h->stream, &launches);
    if (!s.ok()) return fail(h, s);
    API_CUDA(h, cudaMemcpyAsync(h->pinned_ctl, h->sv.ctl, sizeof(Control), cudaMemcpyDeviceToHost, h->stream));
    d2h += (int64_t)sizeof(Control);
  }
  API_CUDA(h, cudaStreamSynchronize(h->stream));
  API_CUDA(h, cudaGetLastError());
  if (stats) {
    float ms = 0.f;
    API_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    const Control* c = h->pinned_ctl;
    stats->levels = c->result.levels;
    stats->td_levels = c->result.td_levels;
    stats->bu_levels = c->result.bu_levels;
    stats->reserved0 = 0;
    stats->reached_vertices = (int64_t)c->reached_vertices;
    stats->reached_entries = (int64_t)c->reached_entries;
    stats->kernel_launches = launches;
    stats->h2d_bytes = (h->engine == B2BFS_ENGINE_STEPWISE) ? (int64_t)sizeof(SearchResult) : 0;
    stats->d2h_bytes = d2h;
    stats->search_ms = ms;
    stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  return B2BFS_OK;
}

int b2bfs_run(b2bfs_t* h, int64_t src, int32_t* depth_out, int32_t* parent_out, b2bfs_stats* stats) {
  if (h && !depth_out) return fail(h, B2BFS_ERR_ARG, "b2bfs_run: depth_out is null (use b2bfs_run_device to keep results in HBM)");
  return run_impl(h, src, depth_out, parent_out, stats);
}

int b2bfs_run_device(b2bfs_t* h, int64_t src, b2bfs_stats* stats) { return run_impl(h, src, nullptr, nullptr, stats); }

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
  if (!h || !h->has_graph || cap < 0) return B2BFS_ERR_ARG;
  if (cudaSetDevice(h->device) != cudaSuccess) return B2BFS_ERR_CUDA;
  if (cudaMemcpy(h->pinned_ctl, h->sv.ctl, sizeof(Control), cudaMemcpyDeviceToHost) != cudaSuccess) return B2BFS_ERR_CUDA;
  const Control* c = h->pinned_ctl;
  int n = c->trace_len < cap ? c->trace_len : cap;
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

int b2bfs_destroy(b2bfs_t* h) {
  if (!h) return B2BFS_OK;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  free_graph(h);
  if (h->pinned_ctl) cudaFreeHost(h->pinned_ctl);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return B2BFS_OK;
}

const char* b2bfs_last_error(const b2bfs_t* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

}  // extern "C"
