// comparators.cu — the reference's two ALTERNATIVE strategies as selectable engines.
// Both are dead code in the reference (call sites commented out of main, bfs.cu:805-819); they are
// built here, done correctly, so that the production engines can be measured against them on the
// same graphs through the same C-ABI (option "engine" = B2BFS_ENGINE_SCAN / B2BFS_ENGINE_PULL).
//
//   engine scan   runCudaScanBfs, bfs.cu:748-781 (kernels nextLayer / countDegrees / scanDegrees /
//                 assignVerticesNextQueue live in a .ptx the repository does not ship): the next
//                 queue is generated WITHOUT a tail atomic — every frontier vertex counts the
//                 neighbours it discovered, an exclusive prefix sum over those counts gives it a
//                 private output range, and it writes its discoveries there.  The reference scans
//                 inside 1024-thread blocks on the device and adds the block sums in a HOST loop
//                 (bfs.cu:733-736, through pinned memory); here the block sums are scanned on the
//                 device too and the host reads one number per level (the new queue length).
//   engine pull   multiBfs + runCudaSimpleBfsMulti, bfs.cu:101-130, 475-539: topology-driven - every
//                 level looks at ALL vertices, those whose depth equals the level relax their
//                 neighbours, a `changed` flag ends the loop.  No queue at all; O(n) work per level
//                 on top of the edges.  The reference's relax is a racy read-modify-write through
//                 getLevel(), which indexes the CALLER's device instead of the neighbour's owner
//                 (bfs.cu:41-46, 119-122); here a neighbour is claimed with one atomicCAS on its depth
//                 word, so exactly one parent is recorded, and (single GPU) the owner question does
//                 not arise.
// Depths are the same as every other engine's (tests/test_comparators.py); parents form a valid
// Graph500 tree.  Single-GPU handles only.
#include <algorithm>

#include "bfs.cuh"

namespace b2 {

namespace {

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }

// ------------------------------------------------------------------------------------------ scan
// nextLayer (bfs.cu:706-713 launches it): one warp per frontier vertex, 32 neighbours per step; a
// neighbour is claimed through the visited bitmap, the winner records depth and parent
__global__ void __launch_bounds__(kBlock) scan_next_layer_kernel(BfsParams p, const int32_t* __restrict__ queue,
                                                                 unsigned int qsize, int level) {
  const unsigned lane = lane_id();
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  for (int64_t i = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); i < (int64_t)qsize; i += nwarps) {
    const int32_t u = queue[i];
    const int64_t rs = __ldg(p.g.rowptr + u), re = __ldg(p.g.rowptr + u + 1);
    for (int64_t e = rs + lane; e < re; e += 32) {
      const int32_t v = __ldg(p.g.col + e);
      const uint32_t bit = 1u << (v & 31);
      uint32_t* w = p.s.visited + ((uint32_t)v >> 5);
      if (!(*w & bit) && !(atomicOr(w, bit) & bit)) {
        p.s.depth[v] = level + 1;
        p.s.parent[v] = u;
      }
    }
  }
}

// was entry e of u's row the one through which u discovered col[e] in this level?  (a neighbour
// listed twice in a row - multigraph, dedup = 0 - counts once: equal ids are adjacent in a row)
__device__ __forceinline__ bool discovered_here(const BfsParams& p, int32_t u, int64_t rs, int64_t e, int level) {
  const int32_t v = __ldg(p.g.col + e);
  if (e > rs && __ldg(p.g.col + e - 1) == v) return false;
  return p.s.parent[v] == u && p.s.depth[v] == level + 1 && v != u;
}

// countDegrees (bfs.cu:715-722): how many neighbours each frontier vertex discovered
__global__ void __launch_bounds__(kBlock) scan_count_kernel(BfsParams p, const int32_t* __restrict__ queue,
                                                            unsigned int qsize, int level, unsigned int* __restrict__ counts) {
  const unsigned lane = lane_id();
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  for (int64_t i = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); i < (int64_t)qsize; i += nwarps) {
    const int32_t u = queue[i];
    const int64_t rs = __ldg(p.g.rowptr + u), re = __ldg(p.g.rowptr + u + 1);
    unsigned c = 0;
    for (int64_t e = rs + lane; e < re; e += 32) c += discovered_here(p, u, rs, e, level);
    for (int o = 16; o; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if (lane == 0) counts[i] = c;
  }
}

// scanDegrees (bfs.cu:724-737): exclusive prefix sums inside blocks of 1024 counts + the block totals
constexpr int kScanBlock = 1024;
__global__ void __launch_bounds__(kScanBlock) scan_blocks_kernel(unsigned int* __restrict__ counts, unsigned int qsize,
                                                                 unsigned int* __restrict__ block_totals) {
  __shared__ unsigned int wsum[32];
  const unsigned i = blockIdx.x * kScanBlock + threadIdx.x;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  const unsigned x = i < qsize ? counts[i] : 0u;
  unsigned incl = x;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= (unsigned)o) incl += y;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    unsigned s = wsum[lane], t = s;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned y = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= (unsigned)o) t += y;
    }
    wsum[lane] = t - s;  // exclusive over warps
    if (lane == 31) block_totals[blockIdx.x] = t;
  }
  __syncthreads();
  if (i < qsize) counts[i] = wsum[warp] + incl - x;
}

// ... the block totals, scanned by one block on the device (the reference does this loop on the host)
__global__ void __launch_bounds__(kScanBlock) scan_totals_kernel(unsigned int* __restrict__ block_totals, unsigned int nblocks,
                                                                 unsigned int* __restrict__ total_out) {
  __shared__ unsigned int wsum[32];
  __shared__ unsigned int carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  for (unsigned base = 0; base < nblocks; base += kScanBlock) {
    const unsigned i = base + threadIdx.x;
    const unsigned x = i < nblocks ? block_totals[i] : 0u;
    unsigned incl = x;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= (unsigned)o) incl += y;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned s = wsum[lane], t = s;
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= (unsigned)o) t += y;
      }
      wsum[lane] = t - s;
    }
    __syncthreads();
    const unsigned carry = carry_s;
    if (i < nblocks) block_totals[i] = carry + wsum[warp] + incl - x;
    __syncthreads();
    if (threadIdx.x == kScanBlock - 1) carry_s = carry + wsum[warp] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry_s;
}

// assignVerticesNextQueue (bfs.cu:739-746): every frontier vertex writes what it discovered into its
// own range of the next queue - no atomic on a shared tail
__global__ void __launch_bounds__(kBlock) scan_assign_kernel(BfsParams p, const int32_t* __restrict__ queue, unsigned int qsize,
                                                             int level, const unsigned int* __restrict__ offsets,
                                                             const unsigned int* __restrict__ block_offsets,
                                                             int32_t* __restrict__ next_queue) {
  const unsigned lane = lane_id();
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  for (int64_t i = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); i < (int64_t)qsize; i += nwarps) {
    const int32_t u = queue[i];
    const int64_t rs = __ldg(p.g.rowptr + u), re = __ldg(p.g.rowptr + u + 1);
    unsigned out = offsets[i] + block_offsets[i / kScanBlock];
    for (int64_t e0 = rs; e0 < re; e0 += 32) {  // warp-uniform
      const int64_t e = e0 + lane;
      const bool mine = e < re && discovered_here(p, u, rs, e, level);
      const unsigned m = __ballot_sync(0xffffffffu, mine);
      if (mine) next_queue[out + __popc(m & ((1u << lane) - 1u))] = __ldg(p.g.col + e);
      out += __popc(m);
    }
  }
}

// ------------------------------------------------------------------------------------------ pull
// multiBfs (bfs.cu:101-130): all vertices are inspected every level; 32 consecutive vertices per warp,
// narrow rows one per lane, wide rows by the whole warp
__global__ void __launch_bounds__(kBlock) pull_level_kernel(BfsParams p, int level, int* __restrict__ changed) {
  const unsigned lane = lane_id();
  const int64_t tiles = (p.g.n + 31) >> 5;
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  bool any = false;
  for (int64_t t = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); t < tiles; t += nwarps) {
    const int64_t u = t * 32 + lane;
    const bool active = u < p.g.n && p.s.depth[u] == level;  // bfs.cu:112 `if (*getLevel(u) == level)`
    int64_t rs = 0, re = 0;
    if (active) { rs = __ldg(p.g.rowptr + u); re = __ldg(p.g.rowptr + u + 1); }
    const bool wide = active && re - rs >= kNarrow;
    if (active && !wide) {
      for (int64_t e = rs; e < re; ++e) {
        const int32_t v = __ldg(p.g.col + e);
        if (p.s.depth[v] == B2BFS_UNREACHED_DEPTH &&
            atomicCAS(p.s.depth + v, B2BFS_UNREACHED_DEPTH, level + 1) == B2BFS_UNREACHED_DEPTH) {
          p.s.parent[v] = (int32_t)u;
          any = true;
        }
      }
    }
    unsigned wm = __ballot_sync(0xffffffffu, wide);
    while (wm) {
      const int l = __ffs(wm) - 1;
      wm &= wm - 1;
      const int64_t wrs = __shfl_sync(0xffffffffu, rs, l), wre = __shfl_sync(0xffffffffu, re, l);
      const int32_t wu = (int32_t)(t * 32 + l);
      for (int64_t e = wrs + lane; e < wre; e += 32) {
        const int32_t v = __ldg(p.g.col + e);
        if (p.s.depth[v] == B2BFS_UNREACHED_DEPTH &&
            atomicCAS(p.s.depth + v, B2BFS_UNREACHED_DEPTH, level + 1) == B2BFS_UNREACHED_DEPTH) {
          p.s.parent[v] = wu;
          any = true;
        }
      }
    }
  }
  if (__any_sync(0xffffffffu, any) && lane == 0) *changed = 1;  // the reference's managed `changed` flag, bfs.cu:123
}

}  // namespace

// scratch: queue[0/1] as the two queues, wideq[0] reinterpreted as the per-vertex counts, wideq[1] as
// the block totals + the level's total (the caller guarantees wideq capacity >= n ints, see api.cu)
Status run_scan_engine(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                       unsigned int* counts, unsigned int* block_totals, int64_t* launches) {
  launch_init(p, src, grid, stream);
  int64_t nl = 1;
  unsigned int qsize = 1;
  int level = 0;
  unsigned int* total_dev = block_totals + ((size_t)(p.g.n + kScanBlock - 1) / kScanBlock + 1);
  unsigned int* total_host = reinterpret_cast<unsigned int*>(pinned);
  // launch_init seeds queue[0] only for narrow sources; this engine has one kind of queue
  int32_t s32 = (int32_t)src;
  B2_CUDA(cudaMemcpyAsync(p.s.queue[0], &s32, sizeof(s32), cudaMemcpyHostToDevice, stream));
  while (qsize) {  // bfs.cu:760
    const int32_t* q = p.s.queue[level & 1];
    int32_t* nq = p.s.queue[(level + 1) & 1];
    const unsigned nblocks = (qsize + kScanBlock - 1) / kScanBlock;
    const int wgrid = (int)std::min<int64_t>(grid, ((int64_t)qsize + kWarpsPerBlock - 1) / kWarpsPerBlock);
    scan_next_layer_kernel<<<wgrid, kBlock, 0, stream>>>(p, q, qsize, level);
    scan_count_kernel<<<wgrid, kBlock, 0, stream>>>(p, q, qsize, level, counts);
    scan_blocks_kernel<<<nblocks, kScanBlock, 0, stream>>>(counts, qsize, block_totals);
    scan_totals_kernel<<<1, kScanBlock, 0, stream>>>(block_totals, nblocks, total_dev);
    scan_assign_kernel<<<wgrid, kBlock, 0, stream>>>(p, q, qsize, level, counts, block_totals, nq);
    nl += 5;
    B2_CUDA(cudaMemcpyAsync(total_host, total_dev, sizeof(unsigned int), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    qsize = *total_host;  // bfs.cu:768
    ++level;
  }
  B2_CUDA(cudaGetLastError());
  SearchResult r{level, level, 0, 0};
  B2_CUDA(cudaMemcpyAsync(&p.s.ctl->result, &r, sizeof(r), cudaMemcpyHostToDevice, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  if (launches) *launches += nl;
  return Status();
}

Status run_pull_engine(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                       int* changed_dev, int64_t* launches) {
  launch_init(p, src, grid, stream);
  int64_t nl = 1;
  int level = 0;
  int* changed_host = reinterpret_cast<int*>(pinned);
  for (;;) {  // bfs.cu:510-530: `while (changed)`
    B2_CUDA(cudaMemsetAsync(changed_dev, 0, sizeof(int), stream));
    pull_level_kernel<<<grid, kBlock, 0, stream>>>(p, level, changed_dev);
    ++nl;
    B2_CUDA(cudaMemcpyAsync(changed_host, changed_dev, sizeof(int), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    ++level;
    if (!*changed_host) break;
  }
  B2_CUDA(cudaGetLastError());
  SearchResult r{level, level, 0, 0};
  B2_CUDA(cudaMemcpyAsync(&p.s.ctl->result, &r, sizeof(r), cudaMemcpyHostToDevice, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  if (launches) *launches += nl;
  return Status();
}

}  // namespace b2
