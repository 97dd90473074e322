// csr_build.cu — edge tuples in HBM -> CSR in HBM, entirely on the GPU.
//
// Stands in for the reference's host-side loader tail and flatten loop (bfs.cu:851-875: push v onto
// u's list and u onto v's, then concatenate rows) and for the H2D half of initCuda2
// (bfs.cu:333-351).  Instead of n heap vectors on one host core, the 2m directed entries become
// 64-bit keys (row | degree class of the neighbour | neighbour), are radix-sorted (radix_sort.cu), optionally de-duplicated
// (duplicates and self-loops never change a BFS depth), and compacted into col[] while the same pass
// derives rowptr[] from the row boundaries of the sorted keys — no per-vertex atomics.
//
// The reference's per-vertex degree (edgesSize, bfs.cu:868 — duplicates and self-loops included) is
// kept as orig_deg[] because the metric counts input tuples (Graph500 rule), not CSR entries.
#include "../../include/b2bfs_graphgen.h"
#include "build.cuh"
#include "sort.cuh"

namespace b2 {

// ---------------------------------------------------------------------------------------------
// synthetic edge streams: tuple i is a pure function of (kind, scale, ef, seed, i)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) generate_edges_kernel(int kind, int scale, int ef, uint64_t seed,
                                                                int64_t first, int64_t count,
                                                                int32_t* __restrict__ src, int32_t* __restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < count; i += (int64_t)gridDim.x * kBlock) {
    uint32_t u, v;
    b2_gen_edge(kind, scale, ef, seed, (uint64_t)(first + i), &u, &v);
    src[i] = (int32_t)u;
    dst[i] = (int32_t)v;
  }
}

Status generate_edges_device(int kind, int scale, int ef, uint64_t seed, int64_t first, int64_t count, int32_t* src,
                             int32_t* dst, cudaStream_t stream) {
  if (count <= 0) return Status();
  const int64_t blocks = (count + kBlock - 1) / kBlock;
  const int grid = (int)(blocks < (int64_t)kSMs * 16 ? blocks : (int64_t)kSMs * 16);
  generate_edges_kernel<<<grid, kBlock, 0, stream>>>(kind, scale, ef, seed, first, count, src, dst);
  B2_CUDA(cudaGetLastError());
  return Status();
}

// ---------------------------------------------------------------------------------------------
// tuples -> keys.  Every tuple (u,v) yields (u -> v) and (v -> u), as bfs.cu:860-861 does; a
// self-loop yields two identical keys, which keeps orig_deg equal to the reference's edgesSize.
// (One GPU.  Several ranks expand their share of the tuples with owner_keys_kernel below; the
// cursor path here - keep only rows in [row_begin, row_end) - is the round-1 scheme, kept for callers
// that hold all tuples.)  bad_flag is raised for ids outside [0,n).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) expand_edges_kernel(const int32_t* __restrict__ src,
                                                              const int32_t* __restrict__ dst, int64_t m, int64_t n,
                                                              int vbits, int rshift, const uint32_t* __restrict__ degc,
                                                              int64_t row_begin, int64_t row_end,
                                                              uint64_t* __restrict__ keys,
                                                              unsigned long long* cursor, int* bad_flag, int directed) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  const int64_t rounds = (m + stride - 1) / stride;
  for (int64_t r = 0; r < rounds; ++r) {
    const int64_t i = r * stride + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    bool ok = i < m;
    int64_t u = 0, v = 0;
    if (ok) {
      u = src[i];
      v = dst[i];
      if (u < 0 || v < 0 || u >= n || v >= n) { *bad_flag = 1; ok = false; }
    }
    // hub-first rows: neighbours sort by descending floor(log2(degree)) before ascending id
    uint64_t cu = 0, cv = 0;
    if (ok && degc) {
      cu = (uint64_t)(15 - min(15, 31 - __clz((int)max(degc[u], 1u)))) << vbits;
      cv = (uint64_t)(15 - min(15, 31 - __clz((int)max(degc[v], 1u)))) << vbits;
    }
    if (directed) {  // the reference's stdin reader: one arc per tuple (bfs.cu:901)
      if (ok) keys[i] = ((uint64_t)u << rshift) | cv | (uint64_t)v;
      else if (i < m) keys[i] = 0;
    } else if (cursor == nullptr) {
      if (ok) {
        keys[2 * i] = ((uint64_t)u << rshift) | cv | (uint64_t)v;
        keys[2 * i + 1] = ((uint64_t)v << rshift) | cu | (uint64_t)u;
      }
    } else {
      const bool a = ok && u >= row_begin && u < row_end;
      const bool b = ok && v >= row_begin && v < row_end;
      const unsigned ma = __ballot_sync(0xffffffffu, a), mb = __ballot_sync(0xffffffffu, b);
      const int na = __popc(ma), nb = __popc(mb);
      unsigned long long base = 0;
      if (lane == 0 && na + nb) base = atomicAdd(cursor, (unsigned long long)(na + nb));
      base = __shfl_sync(0xffffffffu, base, 0);
      const unsigned lt = (1u << lane) - 1u;
      if (a) keys[base + __popc(ma & lt)] = ((uint64_t)(u - row_begin) << rshift) | cv | (uint64_t)v;
      if (b) keys[base + na + __popc(mb & lt)] = ((uint64_t)(v - row_begin) << rshift) | cu | (uint64_t)u;
    }
  }
}

// multigraph degree of every vertex straight from the tuples (only its log2 class is used)
__global__ void __launch_bounds__(kBlock) degree_count_kernel(const int32_t* __restrict__ src,
                                                              const int32_t* __restrict__ dst, int64_t m, int64_t n,
                                                              uint32_t* __restrict__ degc) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < m; i += (int64_t)gridDim.x * kBlock) {
    const int64_t u = src[i], v = dst[i];
    if (u >= 0 && u < n && v >= 0 && v < n && u != v) {
      atomicAdd(&degc[u], 1u);
      atomicAdd(&degc[v], 1u);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// several ranks: every rank holds only ITS share of the tuples (m / world of them).  A tuple (u,v)
// yields the key of entry u -> v for owner(u) and of v -> u for owner(v); pass 1 counts the keys per
// owner, pass 2 writes them into one bucket per owner (one warp-aggregated cursor atomic per owner
// and warp), and the buckets travel to their owners over NVLink (multi.cu, exchange_owner_keys).
// Replaces "every rank reads the whole graph and keeps what it owns" (bfs_mpi.cu:815, and round 1 here).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t class_bits(const uint32_t* degc, int64_t v, int vbits) {
  return degc ? (uint64_t)(15 - min(15, 31 - __clz((int)max(degc[v], 1u)))) << vbits : 0ull;
}

template <bool kWrite>
__global__ void __launch_bounds__(kBlock) owner_keys_kernel(const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                                                            int64_t m, int64_t n, int64_t part, int world, int vbits, int rshift,
                                                            const uint32_t* __restrict__ degc,
                                                            unsigned long long* __restrict__ cursors /* world, zeroed (count) or bucket starts (write) */,
                                                            uint64_t* __restrict__ keys, int* bad_flag) {
  const unsigned lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  const int64_t rounds = (2 * m + stride - 1) / stride;
  for (int64_t r = 0; r < rounds; ++r) {  // one (tuple, direction) per thread and round; warp-uniform trip count
    const int64_t t = r * stride + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    int owner = -1;
    uint64_t key = 0;
    if (t < 2 * m) {
      const int64_t i = t >> 1;
      int64_t u = src[i], v = dst[i];
      if (u < 0 || v < 0 || u >= n || v >= n) {
        *bad_flag = 1;
      } else {
        if (t & 1) { const int64_t x = u; u = v; v = x; }
        owner = (int)(u / part);
        if (kWrite) key = ((uint64_t)(u - (int64_t)owner * part) << rshift) | class_bits(degc, v, vbits) | (uint64_t)v;
      }
    }
    for (int d = 0; d < world; ++d) {
      const unsigned mk = __ballot_sync(0xffffffffu, owner == d);
      if (!mk) continue;
      unsigned long long base = 0;
      const int leader = __ffs(mk) - 1;
      if ((int)lane == leader) base = atomicAdd(&cursors[d], (unsigned long long)__popc(mk));
      if (kWrite) {
        base = __shfl_sync(0xffffffffu, base, leader);
        if (owner == d) keys[base + __popc(mk & ((1u << lane) - 1u))] = key;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// sorted keys -> col[] + rowptr[] (+ the undeduplicated row boundaries)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool keep_entry(const uint64_t key, const uint64_t prev, bool first, int dedup, int vbits,
                                           int rshift, int64_t row_begin) {
  if (!dedup) return true;
  const uint64_t vmask = (1ull << vbits) - 1ull;
  const int64_t row = (int64_t)(key >> rshift) + row_begin;
  if ((int64_t)(key & vmask) == row) return false;  // self-loop
  return first || key != prev;
}

__global__ void __launch_bounds__(kBlock) count_keep_kernel(const uint64_t* __restrict__ keys, int64_t N,
                                                            int64_t chunk, int dedup, int vbits, int rshift,
                                                            int64_t row_begin, uint64_t* __restrict__ block_counts) {
  __shared__ unsigned long long total_s;
  if (threadIdx.x == 0) total_s = 0;
  __syncthreads();
  const int64_t begin = (int64_t)blockIdx.x * chunk, end = min(begin + chunk, N);
  unsigned long long c = 0;
  for (int64_t i = begin + threadIdx.x; i < end; i += kBlock) {
    const uint64_t key = keys[i];
    const uint64_t prev = i > 0 ? keys[i - 1] : 0;
    c += keep_entry(key, prev, i == 0, dedup, vbits, rshift, row_begin);
  }
  for (int o = 16; o; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&total_s, c);
  __syncthreads();
  if (threadIdx.x == 0) block_counts[blockIdx.x] = total_s;
}

__global__ void __launch_bounds__(kBlock) compact_kernel(const uint64_t* __restrict__ keys, int64_t N, int64_t chunk,
                                                         int dedup, int vbits, int rshift, int64_t row_begin,
                                                         int64_t nrows,
                                                         const uint64_t* __restrict__ block_offsets,
                                                         int32_t* __restrict__ col, int64_t* __restrict__ rowptr,
                                                         int64_t* __restrict__ orig_rowptr) {
  __shared__ unsigned int wtot[kWarpsPerBlock];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const uint64_t vmask = (1ull << vbits) - 1ull;
  const int64_t begin = (int64_t)blockIdx.x * chunk, end = min(begin + chunk, N);
  uint64_t carry = block_offsets[blockIdx.x];
  for (int64_t tile = begin; tile < end; tile += kBlock) {
    const int64_t i = tile + threadIdx.x;
    const bool valid = i < end;
    const uint64_t key = valid ? keys[i] : 0;
    const uint64_t prev = (valid && i > 0) ? keys[i - 1] : 0;
    const bool keep = valid && keep_entry(key, prev, i == 0, dedup, vbits, rshift, row_begin);
    const unsigned b = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) wtot[warp] = __popc(b);
    __syncthreads();
    unsigned before = 0, total = 0;
    for (int w = 0; w < kWarpsPerBlock; ++w) {
      const unsigned c = wtot[w];
      if (w < warp) before += c;
      total += c;
    }
    const uint64_t P = carry + before + __popc(b & lt);
    if (keep) col[P] = (int32_t)(key & vmask);
    if (valid) {
      const int64_t row = (int64_t)(key >> rshift);
      const int64_t prow = i > 0 ? (int64_t)(prev >> rshift) : -1;
      for (int64_t r = prow + 1; r <= row; ++r) {  // rows (prow, row] start here; gaps are empty rows
        rowptr[r] = (int64_t)P;
        orig_rowptr[r] = i;
      }
      if (i == N - 1) {
        for (int64_t r = row + 1; r <= nrows; ++r) {
          rowptr[r] = (int64_t)(P + (keep ? 1 : 0));
          orig_rowptr[r] = N;
        }
      }
    }
    carry += total;
    __syncthreads();
  }
}

// orig_deg, the "never scan bottom-up" bitmap (degree-0 rows and the padding past n), wide-item count
__global__ void __launch_bounds__(kBlock) finish_rows_kernel(int64_t nrows, int64_t nwords,
                                                             const int64_t* __restrict__ rowptr,
                                                             const int64_t* __restrict__ orig_rowptr,
                                                             uint32_t* __restrict__ orig_deg,
                                                             const int32_t* __restrict__ col,
                                                             uint32_t* __restrict__ visited_init,
                                                             int32_t* __restrict__ first_nbr,
                                                             unsigned long long* wide_items,
                                                             unsigned long long* max_degree, int directed) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * kWarpsPerBlock;
  unsigned long long mx = 0;
  for (int64_t w = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); w < nwords; w += warps) {
    const int64_t v = w * 32 + lane;
    bool skip = true;
    unsigned long long items = 0;
    if (v >= nrows) first_nbr[v] = -1;
    if (v < nrows) {
      const int64_t d = rowptr[v + 1] - rowptr[v];
      first_nbr[v] = d > 0 ? col[rowptr[v]] : -1;
      orig_deg[v] = (uint32_t)(orig_rowptr[v + 1] - orig_rowptr[v]);
      skip = d == 0 && !directed;  // a vertex without out-arcs can still be reached over an in-arc
      if (d >= kNarrow) items = (unsigned long long)((d + kChunk - 1) / kChunk);
      mx = d > (int64_t)mx ? (unsigned long long)d : mx;
    }
    const unsigned sm = __ballot_sync(0xffffffffu, skip);
    for (int o = 16; o; o >>= 1) items += __shfl_down_sync(0xffffffffu, items, o);
    if (lane == 0) {
      visited_init[w] = sm;
      if (items) atomicAdd(wide_items, items);
    }
  }
  for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_down_sync(0xffffffffu, mx, o));
  if (lane == 0 && mx) atomicMax(max_degree, mx);
}

// temporaries of one build: freed on every exit path
struct DeviceTemps {
  void* ptrs[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int count = 0;
  template <typename T>
  cudaError_t alloc(T** p, size_t bytes) {
    const cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaSuccess && count < 8) ptrs[count++] = *p;
    return e;
  }
  ~DeviceTemps() {
    for (int i = 0; i < count; ++i) cudaFree(ptrs[i]);
  }
};

static int bits_for(int64_t n) {
  int b = 1;
  while (((int64_t)1 << b) < n) ++b;
  return b;
}

KeySpec key_spec(int64_t n, int64_t max_rows_per_rank, int hub_first, int directed) {
  KeySpec k;
  k.vbits = bits_for(n);
  const int rbits = bits_for(max_rows_per_rank > 0 ? max_rows_per_rank : 1);
  k.cbits = (hub_first && !directed && k.vbits + rbits + 4 <= 64) ? 4 : 0;
  k.rshift = k.vbits + k.cbits;
  return k;
}

static int tuple_grid(int64_t m) {
  const int64_t eblocks = (m + kBlock - 1) / kBlock;
  return (int)(eblocks < (int64_t)kSMs * 16 ? (eblocks > 0 ? eblocks : 1) : (int64_t)kSMs * 16);
}

Status count_degrees_device(const int32_t* src, const int32_t* dst, int64_t m, int64_t n, uint32_t* degc, cudaStream_t stream) {
  degree_count_kernel<<<tuple_grid(m), kBlock, 0, stream>>>(src, dst, m, n, degc);
  B2_CUDA(cudaGetLastError());
  return Status();
}

Status owner_keys_device(bool write, const int32_t* src, const int32_t* dst, int64_t m, int64_t n, int64_t part, int world,
                         const KeySpec& ks, const uint32_t* degc, unsigned long long* cursors, uint64_t* keys, int* bad_flag,
                         cudaStream_t stream) {
  const int grid = tuple_grid(2 * m);
  if (write)
    owner_keys_kernel<true><<<grid, kBlock, 0, stream>>>(src, dst, m, n, part, world, ks.vbits, ks.rshift, degc, cursors, keys, bad_flag);
  else
    owner_keys_kernel<false><<<grid, kBlock, 0, stream>>>(src, dst, m, n, part, world, ks.vbits, ks.rshift, degc, cursors, keys, bad_flag);
  B2_CUDA(cudaGetLastError());
  return Status();
}

// one GPU: all 2m (or m, directed) keys of the tuples, then build_csr_from_keys
Status build_csr_device(const int32_t* src, const int32_t* dst, int64_t m, int64_t n, int dedup, int hub_first, int directed,
                        cudaStream_t stream, BuiltCsr* out, int64_t* launches) {
  *out = BuiltCsr();
  const KeySpec ks = key_spec(n, n, hub_first, directed);
  DeviceTemps temps;
  int* bad = nullptr;
  B2_CUDA(temps.alloc(&bad, sizeof(int)));
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), stream));
  const int64_t N = directed ? m : 2 * m;
  uint64_t* ka = nullptr;
  B2_CUDA(cudaMalloc(&ka, (size_t)(N > 0 ? N : 1) * sizeof(uint64_t)));  // handed over to build_csr_from_keys
  uint32_t* degc = nullptr;
  if (ks.cbits) {
    if (cudaMalloc(&degc, (size_t)n * sizeof(uint32_t)) != cudaSuccess) { cudaFree(ka); return make_error(B2BFS_ERR_NOMEM, "degree classes"); }
    temps.ptrs[temps.count++] = degc;
    cudaMemsetAsync(degc, 0, (size_t)n * sizeof(uint32_t), stream);
    degree_count_kernel<<<tuple_grid(m), kBlock, 0, stream>>>(src, dst, m, n, degc);
    if (launches) *launches += 1;
  }
  expand_edges_kernel<<<tuple_grid(m), kBlock, 0, stream>>>(src, dst, m, n, ks.vbits, ks.rshift, degc, 0, n, ka, nullptr, bad, directed);
  if (launches) *launches += 1;
  int bad_h = 0;
  cudaError_t e = cudaMemcpyAsync(&bad_h, bad, sizeof(int), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess || bad_h) {
    cudaFree(ka);
    if (e != cudaSuccess) return make_error(B2BFS_ERR_CUDA, std::string("key expansion: ") + cudaGetErrorString(e));
    return make_error(B2BFS_ERR_RANGE, "edge list holds a vertex id outside [0, vertex_num)");
  }
  return build_csr_from_keys(ka, N, n, 0, n, ks, dedup, directed, 0, stream, out, launches);
}

// keys of this rank's rows (row id relative to row_begin in the top bits) -> CSR.  Takes ownership of
// `ka` (freed on every path).
Status build_csr_from_keys(uint64_t* ka, int64_t N, int64_t n, int64_t row_begin, int64_t row_end, const KeySpec& ks, int dedup,
                           int directed, int64_t nwords_override, cudaStream_t stream, BuiltCsr* out, int64_t* launches) {
  *out = BuiltCsr();
  const int64_t nrows = row_end - row_begin;
  const int vbits = ks.vbits, rshift = ks.rshift;
  const int rbits = bits_for(nrows > 0 ? nrows : 1);
  const int64_t nwords = nwords_override > 0 ? nwords_override : ((nrows + 31) / 32 + 3) / 4 * 4;
  out->nrows = nrows;
  out->nwords = nwords;
  (void)n;

  DeviceTemps temps;
  temps.ptrs[temps.count++] = ka;
  unsigned long long* counters = nullptr;  // [1] wide items, [2] max degree
  B2_CUDA(temps.alloc(&counters, 4 * sizeof(unsigned long long)));
  B2_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned long long), stream));

  uint64_t *kb = nullptr, *sorted = nullptr;
  void* scratch = nullptr;
  B2_CUDA(temps.alloc(&kb, (size_t)(N > 0 ? N : 1) * sizeof(uint64_t)));
  const int G = sort_grid_size(N);
  B2_CUDA(temps.alloc(&scratch, sort_scratch_bytes(N) + (size_t)(G + 1) * sizeof(uint64_t)));

  B2_TRY(radix_sort_u64(ka, kb, N, rshift + rbits, scratch, stream, &sorted, launches));

  // chunked two-pass compaction over the sorted keys
  int64_t chunk = (N + G - 1) / G;
  chunk = (chunk + kBlock - 1) / kBlock * kBlock;
  if (chunk == 0) chunk = kBlock;
  uint64_t* block_counts = reinterpret_cast<uint64_t*>(scratch);
  uint64_t* total_d = block_counts + G;
  count_keep_kernel<<<G, kBlock, 0, stream>>>(sorted, N, chunk, dedup, vbits, rshift, row_begin, block_counts);
  scan_u64_single_block<<<1, 1024, 0, stream>>>(block_counts, G, total_d);
  uint64_t entries = 0;
  B2_CUDA(cudaMemcpyAsync(&entries, total_d, sizeof(entries), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  if (launches) *launches += 2;

  out->entries = (int64_t)entries;
  out->orig_entries = N;
  int64_t* orig_rowptr = nullptr;
  B2_CUDA(cudaMalloc(&out->rowptr, (size_t)(nrows + 1) * sizeof(int64_t)));
  B2_CUDA(temps.alloc(&orig_rowptr, (size_t)(nrows + 1) * sizeof(int64_t)));
  // col is read in 16-byte vectors and 16-byte-aligned bulk copies: pad both ends
  B2_CUDA(cudaMalloc(&out->col, ((size_t)entries + 64) * sizeof(int32_t)));
  B2_CUDA(cudaMalloc(&out->orig_deg, (size_t)nwords * 32 * sizeof(uint32_t)));
  B2_CUDA(cudaMemsetAsync(out->orig_deg, 0, (size_t)nwords * 32 * sizeof(uint32_t), stream));
  B2_CUDA(cudaMalloc(&out->visited_init, (size_t)nwords * sizeof(uint32_t)));
  B2_CUDA(cudaMalloc(&out->first_nbr, (size_t)nwords * 32 * sizeof(int32_t)));
  B2_CUDA(cudaMemsetAsync(out->col + entries, 0xff, 64 * sizeof(int32_t), stream));
  if (N == 0) {
    B2_CUDA(cudaMemsetAsync(out->rowptr, 0, (size_t)(nrows + 1) * sizeof(int64_t), stream));
    B2_CUDA(cudaMemsetAsync(orig_rowptr, 0, (size_t)(nrows + 1) * sizeof(int64_t), stream));
  } else {
    compact_kernel<<<G, kBlock, 0, stream>>>(sorted, N, chunk, dedup, vbits, rshift, row_begin, nrows, block_counts,
                                             out->col, out->rowptr, orig_rowptr);
    if (launches) *launches += 1;
  }
  const int64_t fblocks = (nwords + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const int fgrid = (int)(fblocks < (int64_t)kSMs * 8 ? fblocks : (int64_t)kSMs * 8);
  finish_rows_kernel<<<fgrid, kBlock, 0, stream>>>(nrows, nwords, out->rowptr, orig_rowptr, out->orig_deg, out->col,
                                                   out->visited_init, out->first_nbr, counters + 1, counters + 2, directed);
  if (launches) *launches += 1;
  unsigned long long hc[2] = {0, 0};
  B2_CUDA(cudaMemcpyAsync(hc, counters + 1, sizeof(hc), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  B2_CUDA(cudaGetLastError());
  out->wide_items = (int64_t)hc[0];
  out->max_degree = (int64_t)hc[1];

  return Status();
}

__global__ void __launch_bounds__(kBlock) row_degrees_kernel(const int64_t* __restrict__ rowptr, int64_t nrows,
                                                             int32_t* __restrict__ deg) {
  for (int64_t v = (int64_t)blockIdx.x * kBlock + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * kBlock)
    deg[v] = (int32_t)(rowptr[v + 1] - rowptr[v]);
}

Status row_degrees_device(const int64_t* rowptr, int64_t nrows, int32_t* deg, cudaStream_t stream) {
  if (nrows <= 0) return Status();
  const int64_t blocks = (nrows + kBlock - 1) / kBlock;
  row_degrees_kernel<<<(int)(blocks < (int64_t)kSMs * 8 ? blocks : (int64_t)kSMs * 8), kBlock, 0, stream>>>(rowptr, nrows, deg);
  B2_CUDA(cudaGetLastError());
  return Status();
}

void free_csr(BuiltCsr* c) {
  cudaFree(c->rowptr);
  cudaFree(c->col);
  cudaFree(c->orig_deg);
  cudaFree(c->visited_init);
  cudaFree(c->first_nbr);
  *c = BuiltCsr();
}

}  // namespace b2
