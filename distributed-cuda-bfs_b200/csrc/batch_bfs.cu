// batch_bfs.cu — b2bfs_run_batch: up to 64 searches over the resident graph in ONE pass per level
// (SURVEY.md §8f, N4).  The reference has no batched search; what this serves is the way its metric
// is taken - 64 roots over one graph (BASELINE.json) - where 64 separate searches read the same
// adjacency 64 times.
//
// Bit-parallel multi-source BFS: per vertex three 64-bit words, bit i belonging to search i —
//   seen[v]   search i has reached v
//   fcur[v]   v entered search i's frontier in the previous level
//   fnext[v]  v enters it in this level
// A level is a PULL over all vertices that still miss some search: OR the frontier words of the
// neighbours, keep the bits not seen yet.  One read of col[] serves all 64 searches, there is no
// atomic per edge and no queue; a vertex stops scanning as soon as every search it still misses has
// been found among its neighbours (rows are hub-first, so that is usually the first few entries once
// the hubs are in the frontiers).  Work is split like the single-source top-down phase: rows narrower
// than kNarrow one per lane, wider rows cut into kChunk-entry items that any warp may take, the
// items OR-ing their finds into fnext[v] (one atomic per item at most).
//
// algorithmic bytes per level: 4 B per entry scanned + 8 B (fcur word, L2 resident up to scale 23) per
// entry + 24 B per vertex (seen, fnext) ; once per search and vertex 4 B of depth (+ the 4 B fill).
#include <vector>

#include "batch.cuh"

namespace b2 {

namespace {

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ uint64_t warp_or64(uint64_t x) {
  const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)x);
  const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(x >> 32));
  return ((uint64_t)hi << 32) | lo;
}

// depth rows of the searches whose bit is set in `fresh`; warp-uniform loop over the union of the
// lanes' bits, so the lanes that share a bit write one coalesced piece of that search's row
__device__ __forceinline__ void write_depths(int32_t* depth, int64_t n, int64_t v, uint64_t fresh, int32_t value) {
  uint64_t all = warp_or64(fresh);
  while (all) {
    const int b = __ffsll((long long)all) - 1;
    all &= all - 1;
    if ((fresh >> b) & 1ull) depth[(int64_t)b * n + v] = value;
  }
}

__global__ void __launch_bounds__(kBlock) batch_fill_kernel(int64_t n, int nroots, int32_t* __restrict__ depth,
                                                            uint64_t* __restrict__ seen, uint64_t* __restrict__ fcur,
                                                            uint64_t* __restrict__ fnext) {
  const int64_t tid = (int64_t)blockIdx.x * kBlock + threadIdx.x, nth = (int64_t)gridDim.x * kBlock;
  const int64_t total4 = ((int64_t)nroots * n) >> 2;
  int4* d4 = reinterpret_cast<int4*>(depth);
  const int4 far = make_int4(B2BFS_UNREACHED_DEPTH, B2BFS_UNREACHED_DEPTH, B2BFS_UNREACHED_DEPTH, B2BFS_UNREACHED_DEPTH);
  for (int64_t i = tid; i < total4; i += nth) d4[i] = far;
  for (int64_t i = (total4 << 2) + tid; i < (int64_t)nroots * n; i += nth) depth[i] = B2BFS_UNREACHED_DEPTH;
  for (int64_t i = tid; i < n; i += nth) { seen[i] = 0; fcur[i] = 0; fnext[i] = 0; }
}

__global__ void batch_seed_kernel(int64_t n, int nroots, const int64_t* __restrict__ roots, int32_t* __restrict__ depth,
                                  uint64_t* __restrict__ seen, uint64_t* __restrict__ fcur) {
  const int i = threadIdx.x;
  if (i < nroots) {
    const int64_t v = roots[i];
    atomicOr(reinterpret_cast<unsigned long long*>(seen + v), 1ull << i);
    atomicOr(reinterpret_cast<unsigned long long*>(fcur + v), 1ull << i);
    depth[(int64_t)i * n + v] = 0;
  }
}

// vertices with wide rows, and the (vertex, chunk >= 1) items of those rows; chunk 0 of a wide row
// is implied by the vertex list, so that all chunk-0 items can be handed out first
__global__ void __launch_bounds__(kBlock) batch_list_wide_kernel(int64_t n, const int64_t* __restrict__ rowptr,
                                                                 int32_t* __restrict__ wide_list, int2* __restrict__ rest,
                                                                 unsigned long long* counters) {
  const unsigned lane = lane_id();
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  const int64_t rounds = (n + stride - 1) / stride;
  for (int64_t r = 0; r < rounds; ++r) {
    const int64_t v = r * stride + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const int64_t d = v < n ? rowptr[v + 1] - rowptr[v] : 0;
    const bool wide = d >= kNarrow;
    const unsigned m = __ballot_sync(0xffffffffu, wide);
    if (!m) continue;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(&counters[0], (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (wide) {
      wide_list[base + __popc(m & ((1u << lane) - 1u))] = (int32_t)v;
      const unsigned more = (unsigned)((d + kChunk - 1) / kChunk) - 1u;
      if (more) {
        const unsigned long long at = atomicAdd(&counters[1], (unsigned long long)more);
        for (unsigned c = 0; c < more; ++c) rest[at + c] = make_int2((int)v, (int)c + 1);
      }
    }
  }
}

struct BatchView {
  int64_t n;
  int nroots;
  const int64_t* rowptr;
  const int32_t* col;
  uint64_t* seen;
  const uint64_t* fcur;
  uint64_t* fnext;
  int32_t* depth;
  const int32_t* wide_list;
  int64_t n_wide;
  const int2* rest;
  int64_t n_rest;
  unsigned long long* fresh_count;  // bits newly set by this level (narrow rows; wide rows add theirs when finalised)
};

// one wide item: <= kChunk entries of row v, 4 per lane
__device__ __forceinline__ void pull_wide_item(const BatchView& b, int32_t v, int chunk, uint64_t full) {
  const unsigned lane = lane_id();
  // what the row still misses, narrowed by what other items of the same row have found so far
  const uint64_t need = full & ~b.seen[v] & ~__ldcg(b.fnext + v);
  if (!need) return;
  const int64_t rs = __ldg(b.rowptr + v) + (int64_t)chunk * kChunk;
  const int64_t re = min(__ldg(b.rowptr + v + 1), rs + kChunk);
  uint64_t acc = 0;
  for (int64_t b0 = rs & ~(int64_t)3; b0 < re; b0 += 128) {  // warp-uniform; col is padded and 16-byte aligned
    const int64_t base = b0 + 4 * lane;
    if (base < re) {
      const int4 x = __ldg(reinterpret_cast<const int4*>(b.col + base));
      const uint64_t f0 = (base >= rs) ? __ldg(b.fcur + x.x) : 0;
      const uint64_t f1 = (base + 1 >= rs && base + 1 < re) ? __ldg(b.fcur + x.y) : 0;
      const uint64_t f2 = (base + 2 >= rs && base + 2 < re) ? __ldg(b.fcur + x.z) : 0;
      const uint64_t f3 = (base + 3 >= rs && base + 3 < re) ? __ldg(b.fcur + x.w) : 0;
      acc |= f0 | f1 | f2 | f3;
    }
  }
  const uint64_t found = warp_or64(acc & need);
  if (found && lane == 0) atomicOr(reinterpret_cast<unsigned long long*>(b.fnext + v), found);
}

__global__ void __launch_bounds__(kBlock) batch_pull_kernel(BatchView b, int level) {
  const unsigned lane = lane_id();
  const uint64_t full = b.nroots == 64 ? ~0ull : ((1ull << b.nroots) - 1ull);
  const int64_t tiles = (b.n + 31) >> 5;
  const int64_t work = b.n_wide + tiles + b.n_rest;
  const int64_t gw = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  unsigned long long fresh_bits = 0;
  for (int64_t item = gw; item < work; item += nwarps) {
    if (item < b.n_wide) {
      pull_wide_item(b, __ldg(b.wide_list + item), 0, full);
    } else if (item >= b.n_wide + tiles) {
      const int2 it = __ldg(b.rest + (item - b.n_wide - tiles));
      pull_wide_item(b, it.x, it.y, full);
    } else {
      // 32 consecutive vertices, one per lane; only the narrow rows are handled here
      const int64_t v = (item - b.n_wide) * 32 + lane;
      int64_t rs = 0;
      int d = kNarrow;  // lanes past n and wide rows: nothing to do here
      if (v < b.n) {
        rs = __ldg(b.rowptr + v);
        const int64_t dd = __ldg(b.rowptr + v + 1) - rs;
        d = dd < kNarrow ? (int)dd : kNarrow;
      }
      uint64_t fresh = 0;
      if (d < kNarrow) {
        const uint64_t need = full & ~b.seen[v];
        if (need && d > 0) {
          uint64_t acc = 0;
          for (int j = 0; j < d; j += 4) {  // four independent col -> fcur chains per round
            const int32_t u0 = __ldg(b.col + rs + j);
            const int32_t u1 = j + 1 < d ? __ldg(b.col + rs + j + 1) : u0;
            const int32_t u2 = j + 2 < d ? __ldg(b.col + rs + j + 2) : u0;
            const int32_t u3 = j + 3 < d ? __ldg(b.col + rs + j + 3) : u0;
            acc |= __ldg(b.fcur + u0) | __ldg(b.fcur + u1) | __ldg(b.fcur + u2) | __ldg(b.fcur + u3);
            if ((acc & need) == need) break;  // every search this vertex still missed has reached it
          }
          fresh = acc & need;
          if (fresh) b.seen[v] |= fresh;
        }
        b.fnext[v] = fresh;  // plain store: a narrow row has exactly one writer
      }
      write_depths(b.depth, b.n, v, fresh, level + 1);
      fresh_bits += (unsigned long long)__popcll(fresh);
    }
  }
  for (int o = 16; o; o >>= 1) fresh_bits += __shfl_down_sync(0xffffffffu, fresh_bits, o);
  if (lane == 0 && fresh_bits) atomicAdd(b.fresh_count, fresh_bits);
}

// after the pull: wide rows take over what their items found; the word that will collect the next
// level's finds (today's fcur) is cleared for them
__global__ void __launch_bounds__(kBlock) batch_finish_wide_kernel(BatchView b, uint64_t* fcur_to_clear, int level) {
  const unsigned lane = lane_id();
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  const int64_t rounds = (b.n_wide + stride - 1) / stride;
  unsigned long long fresh_bits = 0;
  for (int64_t r = 0; r < rounds; ++r) {  // warp-uniform trip count (write_depths votes)
    const int64_t i = r * stride + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    uint64_t fresh = 0;
    int64_t v = 0;
    if (i < b.n_wide) {
      v = b.wide_list[i];
      fresh = b.fnext[v] & ~b.seen[v];
      if (fresh) b.seen[v] |= fresh;
      b.fnext[v] = fresh;
      fcur_to_clear[v] = 0;
    }
    write_depths(b.depth, b.n, v, fresh, level + 1);
    fresh_bits += (unsigned long long)__popcll(fresh);
  }
  for (int o = 16; o; o >>= 1) fresh_bits += __shfl_down_sync(0xffffffffu, fresh_bits, o);
  if (lane == 0 && fresh_bits) atomicAdd(b.fresh_count, fresh_bits);
}

// V_r and D_r of every search, for the metric (outside the timed window)
__global__ void __launch_bounds__(kBlock) batch_stats_kernel(int64_t n, const uint64_t* __restrict__ seen,
                                                             const uint32_t* __restrict__ orig_deg,
                                                             unsigned long long* __restrict__ out /* [64] V_r, [64] D_r */) {
  __shared__ unsigned long long vr[64], dr[64];
  if (threadIdx.x < 64) { vr[threadIdx.x] = 0; dr[threadIdx.x] = 0; }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  const int64_t rounds = (n + stride - 1) / stride;
  for (int64_t r = 0; r < rounds; ++r) {
    const int64_t v = r * stride + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const uint64_t s = v < n ? seen[v] : 0;
    const unsigned long long dg = v < n ? orig_deg[v] : 0;
    uint64_t all = warp_or64(s);
    while (all) {
      const int bit = __ffsll((long long)all) - 1;
      all &= all - 1;
      const bool has = (s >> bit) & 1ull;
      const unsigned m = __ballot_sync(0xffffffffu, has);
      unsigned long long x = has ? dg : 0;
      for (int o = 16; o; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
      if (lane_id() == 0) {
        atomicAdd(&vr[bit], (unsigned long long)__popc(m));
        atomicAdd(&dr[bit], x);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 64) {
    if (vr[threadIdx.x]) atomicAdd(&out[threadIdx.x], vr[threadIdx.x]);
    if (dr[threadIdx.x]) atomicAdd(&out[64 + threadIdx.x], dr[threadIdx.x]);
  }
}

}  // namespace

void free_batch(BatchState* bs) {
  cudaFree(bs->seen);
  cudaFree(bs->fa);
  cudaFree(bs->fb);
  cudaFree(bs->depth);
  cudaFree(bs->wide_list);
  cudaFree(bs->rest);
  cudaFree(bs->counters);
  cudaFree(bs->roots_dev);
  if (bs->host) cudaFreeHost(bs->host);
  *bs = BatchState();
}

// per graph, once: the state words, the wide-row lists, room for `nroots` depth rows
Status prepare_batch(BatchState* bs, int64_t n, int64_t wide_items_cap, const int64_t* rowptr, int nroots,
                     cudaStream_t stream, int64_t* launches) {
  if (!bs->seen) {
    const size_t words = (size_t)(n > 0 ? n : 1) * sizeof(uint64_t);
    B2_CUDA(cudaMalloc(&bs->seen, words));
    B2_CUDA(cudaMalloc(&bs->fa, words));
    B2_CUDA(cudaMalloc(&bs->fb, words));
    B2_CUDA(cudaMalloc(&bs->counters, 256 * sizeof(unsigned long long)));
    B2_CUDA(cudaMalloc(&bs->roots_dev, 64 * sizeof(int64_t)));
    B2_CUDA(cudaMallocHost(&bs->host, 256 * sizeof(unsigned long long)));
    B2_CUDA(cudaMalloc(&bs->wide_list, (size_t)(wide_items_cap + 1) * sizeof(int32_t)));
    B2_CUDA(cudaMalloc(&bs->rest, (size_t)(wide_items_cap + 1) * sizeof(int2)));
    B2_CUDA(cudaMemsetAsync(bs->counters, 0, 256 * sizeof(unsigned long long), stream));
    const int64_t blocks = (n + kBlock - 1) / kBlock;
    batch_list_wide_kernel<<<(int)(blocks < (int64_t)kSMs * 8 ? (blocks > 0 ? blocks : 1) : (int64_t)kSMs * 8), kBlock, 0, stream>>>(
        n, rowptr, bs->wide_list, bs->rest, bs->counters);
    B2_CUDA(cudaMemcpyAsync(bs->host, bs->counters, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    B2_CUDA(cudaGetLastError());
    bs->n_wide = (int64_t)bs->host[0];
    bs->n_rest = (int64_t)bs->host[1];
    if (launches) *launches += 1;
  }
  if (bs->depth_rows < nroots) {
    cudaFree(bs->depth);
    bs->depth = nullptr;
    bs->depth_rows = 0;
    B2_CUDA(cudaMalloc(&bs->depth, ((size_t)nroots * (size_t)n + 4) * sizeof(int32_t)));
    bs->depth_rows = nroots;
  }
  return Status();
}

Status run_batch(BatchState* bs, int64_t n, const int64_t* rowptr, const int32_t* col, int nroots,
                 const int64_t* roots_host, int grid, cudaStream_t stream, cudaEvent_t ev0, cudaEvent_t ev1,
                 int* levels_out, int64_t* launches) {
  B2_CUDA(cudaMemcpyAsync(bs->roots_dev, roots_host, (size_t)nroots * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
  B2_CUDA(cudaEventRecord(ev0, stream));
  batch_fill_kernel<<<grid, kBlock, 0, stream>>>(n, nroots, bs->depth, bs->seen, bs->fa, bs->fb);
  batch_seed_kernel<<<1, 64, 0, stream>>>(n, nroots, bs->roots_dev, bs->depth, bs->seen, bs->fa);
  int64_t nl = 2;
  BatchView b;
  b.n = n;
  b.nroots = nroots;
  b.rowptr = rowptr;
  b.col = col;
  b.seen = bs->seen;
  b.depth = bs->depth;
  b.wide_list = bs->wide_list;
  b.n_wide = bs->n_wide;
  b.rest = bs->rest;
  b.n_rest = bs->n_rest;
  uint64_t *cur = bs->fa, *nxt = bs->fb;
  const int64_t wblocks = (bs->n_wide + kBlock - 1) / kBlock;
  const int wgrid = (int)(wblocks < (int64_t)grid ? (wblocks > 0 ? wblocks : 1) : (int64_t)grid);
  int level = 0;
  for (;;) {
    unsigned long long* fresh = bs->counters + 8 + (level & 7);
    B2_CUDA(cudaMemsetAsync(fresh, 0, sizeof(unsigned long long), stream));
    b.fcur = cur;
    b.fnext = nxt;
    b.fresh_count = fresh;
    batch_pull_kernel<<<grid, kBlock, 0, stream>>>(b, level);
    ++nl;
    if (bs->n_wide > 0) {
      batch_finish_wide_kernel<<<wgrid, kBlock, 0, stream>>>(b, cur, level);
      ++nl;
    }
    B2_CUDA(cudaMemcpyAsync(bs->host, fresh, sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    ++level;
    if (bs->host[0] == 0) break;  // no search gained a vertex: all frontiers are empty
    uint64_t* t = cur; cur = nxt; nxt = t;
  }
  B2_CUDA(cudaEventRecord(ev1, stream));
  B2_CUDA(cudaGetLastError());
  *levels_out = level;  // as for one search: levels executed = largest eccentricity + 1
  if (launches) *launches += nl;
  return Status();
}

Status batch_reach_stats(BatchState* bs, int64_t n, const uint32_t* orig_deg, cudaStream_t stream,
                         unsigned long long* vr_dr_host /* 128 */, int64_t* launches) {
  unsigned long long* out = bs->counters + 16;
  B2_CUDA(cudaMemsetAsync(out, 0, 128 * sizeof(unsigned long long), stream));
  const int64_t blocks = (n + kBlock - 1) / kBlock;
  batch_stats_kernel<<<(int)(blocks < (int64_t)kSMs * 8 ? (blocks > 0 ? blocks : 1) : (int64_t)kSMs * 8), kBlock, 0, stream>>>(
      n, bs->seen, orig_deg, out);
  B2_CUDA(cudaMemcpyAsync(bs->host + 16, out, 128 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  B2_CUDA(cudaGetLastError());
  for (int i = 0; i < 128; ++i) vr_dr_host[i] = bs->host[16 + i];
  if (launches) *launches += 1;
  return Status();
}

}  // namespace b2
