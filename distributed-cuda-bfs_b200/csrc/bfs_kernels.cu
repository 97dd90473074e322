// bfs_kernels.cu — the search itself: per-search reset, top-down expansion, bottom-up step,
// frontier conversions, and the level loop, as ONE persistent cooperative kernel (fused engine) or
// as one launch per phase (stepwise engine; also what the multi-rank driver interleaves with NCCL).
//
// What it replaces in the reference (all bfs.cu):
//   search_init      initializeCudaBfs2, 402-422 (host std::fill + two n-int H2D copies per device)
//   top_down phase   queueBfs, 134-165 (thread per frontier vertex, serial adjacency walk, one
//                    atomicAdd per discovered vertex on a managed counter)
//   bottom_up phase  nothing (the reference is top-down only); north_star item (4)
//   level loop       runCudaQueueBfs, 569-619 (launch + cudaDeviceSynchronize + managed counters
//                    + cudaMemset per level)
//
// Frontier: two work lists (narrow rows + wide-row items) while top-down, two bitmaps while bottom-up, and
// a visited bitmap that is the claim target (atomicOr) — 2 MiB at scale-24, L2-resident.
// Every phase is a __device__ function over (blockIdx, gridDim) so both engines run the same code.
#include <cooperative_groups.h>

#include <cstdlib>

#include "bfs.cuh"

namespace cg = cooperative_groups;

namespace b2 {

constexpr int kBuProbe = 4;          // neighbours a bottom-up lane checks alone before its warp helps
constexpr int kStageFlush = 160;     // a warp flushes its staging buffer once it holds this many ids
constexpr int kStageCap = kStageFlush + 128;  // + at most 4 discoveries per lane between two checks

// Per-warp staging of discovered vertices in shared memory: lanes append with shared-memory atomics
// (no warp-wide vote per edge), and the warp pays ONE global atomic on the queue tail per
// ~kStageFlush vertices.  A single hot tail counter sustains only ~1 atomic/ns chip-wide
// (measured: 400 k warp-aggregated appends cost 0.4 ms), so the tail must not be touched per edge.
struct WarpStage {
  int32_t narrow[kStageCap];  // rows with degree <  kNarrow
  int32_t wide[kStageCap];    // rows with degree >= kNarrow ...
  int32_t witems[kStageCap];  // ... and how many kChunk-entry work items each becomes
  int ncnt, wcnt;
};

// ---------------------------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ bool test_bit(const uint32_t* bm, int32_t v) {
  return (bm[(uint32_t)v >> 5] >> (v & 31)) & 1u;
}

// block-level sum of a 64-bit per-thread value -> one atomic per CTA
__device__ __forceinline__ void block_accumulate(unsigned long long x, unsigned long long* target) {
  __shared__ unsigned long long part[kWarpsPerBlock];
  for (int o = 16; o; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
  if (lane_id() == 0) part[threadIdx.x >> 5] = x;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int w = 0; w < kWarpsPerBlock; ++w) t += part[w];
    if (t) atomicAdd(target, t);
  }
  __syncthreads();
}

// Route a newly discovered vertex of degree d to the next level's work lists (lane-local).
__device__ __forceinline__ void route_vertex(WarpStage& sg, int32_t v, int64_t d) {
  if (d < kNarrow) {
    sg.narrow[atomicAdd(&sg.ncnt, 1)] = v;
  } else {
    const int pos = atomicAdd(&sg.wcnt, 1);
    sg.wide[pos] = v;
    sg.witems[pos] = (int32_t)((d + kChunk - 1) / kChunk);
  }
}

// Warp-uniform: drain the staging buffers into the global queues when they are full enough.
__device__ __forceinline__ void stage_flush(WarpStage& sg, LevelSlot* ns, int32_t* nq, int2* nwq, bool force) {
  __syncwarp();
  const unsigned lane = lane_id();
  const int nc = *reinterpret_cast<volatile int*>(&sg.ncnt);
  const int wc = *reinterpret_cast<volatile int*>(&sg.wcnt);
  if (nc >= kStageFlush || (force && nc > 0)) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(&ns->q_count, (unsigned)nc);
    base = __shfl_sync(0xffffffffu, base, 0);
    for (int i = lane; i < nc; i += 32) nq[base + i] = sg.narrow[i];
    __syncwarp();
    if (lane == 0) sg.ncnt = 0;
  }
  if (wc >= kStageFlush || (force && wc > 0)) {
    for (int i0 = 0; i0 < wc; i0 += 32) {  // one tail atomic per 32 staged rows
      const int i = i0 + (int)lane;
      const unsigned items = i < wc ? (unsigned)sg.witems[i] : 0u;
      unsigned incl = items;
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += y;
      }
      unsigned base = 0;
      if (lane == 31) base = atomicAdd(&ns->wide_count, incl);
      base = __shfl_sync(0xffffffffu, base, 31) + incl - items;
      if (items) {
        const int32_t v = sg.wide[i];
        for (unsigned c = 0; c < items; ++c) nwq[base + c] = make_int2(v, (int)c);
      }
    }
    __syncwarp();
    if (lane == 0) sg.wcnt = 0;
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// per-search reset (fused: depth/parent fill, visited := visited_init, bitmaps := 0, seed)
// algorithmic bytes: 8n written (depth + parent) + 3n/8 bitmap bytes
// ---------------------------------------------------------------------------------------------
__device__ void phase_init(const BfsParams& p, int64_t src) {
  const int64_t tid = (int64_t)blockIdx.x * kBlock + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * kBlock;
  const int64_t n = p.g.n, n4 = n >> 2;
  int4* d4 = reinterpret_cast<int4*>(p.s.depth);
  int4* p4 = reinterpret_cast<int4*>(p.s.parent);
  const int4 dmax = make_int4(B2BFS_UNREACHED_DEPTH, B2BFS_UNREACHED_DEPTH, B2BFS_UNREACHED_DEPTH,
                              B2BFS_UNREACHED_DEPTH);
  const int4 pnone = make_int4(-1, -1, -1, -1);
  for (int64_t i = tid; i < n4; i += nthreads) {
    d4[i] = dmax;
    p4[i] = pnone;
    if ((src >> 2) == i) {  // same thread, program order: the seed lands after the fill
      p.s.depth[src] = 0;
      p.s.parent[src] = (int32_t)src;
    }
  }
  for (int64_t i = (n4 << 2) + tid; i < n; i += nthreads) {
    p.s.depth[i] = (i == src) ? 0 : B2BFS_UNREACHED_DEPTH;
    p.s.parent[i] = (i == src) ? (int32_t)src : -1;
  }
  const int64_t w4 = p.g.nwords >> 2;
  const uint4* vi4 = reinterpret_cast<const uint4*>(p.g.visited_init);
  uint4* v4 = reinterpret_cast<uint4*>(p.s.visited);
  uint4* b0 = reinterpret_cast<uint4*>(p.s.bitmap[0]);
  uint4* b1 = reinterpret_cast<uint4*>(p.s.bitmap[1]);
  const uint4 zero = make_uint4(0, 0, 0, 0);
  for (int64_t i = tid; i < w4; i += nthreads) {
    uint4 x = vi4[i];
    if (((src >> 5) >> 2) == i) {
      const uint32_t bit = 1u << (src & 31);
      switch ((src >> 5) & 3) {
        case 0: x.x |= bit; break;
        case 1: x.y |= bit; break;
        case 2: x.z |= bit; break;
        default: x.w |= bit; break;
      }
    }
    v4[i] = x;
    b0[i] = zero;
    b1[i] = zero;
  }
  if (tid == 0) {
    Control* c = p.s.ctl;
    for (int k = 0; k < 4; ++k) {
      c->slot[k].q_count = 0;
      c->slot[k].wide_count = 0;
      c->slot[k].edges = 0;
      c->slot[k].bu_count = 0;
    }
    const int64_t deg = p.g.rowptr[src + 1] - p.g.rowptr[src];
    if (deg < kNarrow) {
      p.s.queue[0][0] = (int32_t)src;
      c->slot[0].q_count = 1;
    } else {
      const unsigned items = (unsigned)((deg + kChunk - 1) / kChunk);
      for (unsigned k = 0; k < items; ++k) p.s.wideq[0][k] = make_int2((int)src, (int)k);
      c->slot[0].wide_count = items;
    }
    c->result.levels = c->result.td_levels = c->result.bu_levels = 0;
    c->src_degree = deg;
  }
}

__host__ __device__ inline LoopState initial_state(const BfsParams& p, int64_t src_degree) {
  LoopState st;
  st.level = 0;
  st.mode = 0;
  st.fb = 0;
  st.done = 0;
  const bool wide = src_degree >= kNarrow;
  st.cur_q = wide ? 0u : 1u;
  st.cur_wide = wide ? (unsigned)((src_degree + kChunk - 1) / kChunk) : 0u;
  st.prev_nf = 1;
  st.unexplored = (unsigned long long)(p.g.entries - src_degree);
  st.td_levels = st.bu_levels = 0;
  return st;
}

// ---------------------------------------------------------------------------------------------
// top-down.  Work items, strided over ALL warps of the grid so that a frontier of a few heavy rows
// (the neighbours of any root in a power-law graph) still spreads over the whole chip:
//   narrow tile  32 frontier rows of degree < kNarrow: warp scan over the degrees, then 32 entries
//                per step regardless of row boundaries
//   wide item    <= kChunk consecutive entries of one row: 16-byte vector loads, 128 entries per step
// Each lane then works alone: probe the visited bitmap, atomicOr to claim, write depth/parent, read
// the new vertex's degree and stage it for the next level.
// algorithmic bytes per level: 4 B per frontier vertex / 8 B per item + 16 B row extent + 4 B per
// neighbour entry; per discovered vertex 8 B (depth + parent) + 16 B its row extent + 4-8 B queue
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void td_visit(const BfsParams& p, int level, WarpStage& sg, int32_t v, int32_t u,
                                         unsigned long long& edges_acc) {
  const uint32_t w = (uint32_t)v >> 5, bit = 1u << (v & 31);
  // cheap read first; a stale "clear" only costs one extra atomic (bits only go 0 -> 1 in a search)
  if (p.s.visited[w] & bit) return;
  if (atomicOr(&p.s.visited[w], bit) & bit) return;
  p.s.depth[v] = level + 1;
  p.s.parent[v] = u;
  const int64_t d = __ldg(p.g.rowptr + v + 1) - __ldg(p.g.rowptr + v);
  edges_acc += (unsigned long long)d;
  route_vertex(sg, v, d);
}

__device__ void phase_top_down(const BfsParams& p, const LoopState& st, WarpStage* stage) {
  const int level = st.level;
  const int32_t* q = p.s.queue[level & 1];
  const int2* wq = p.s.wideq[level & 1];
  int32_t* nq = p.s.queue[(level + 1) & 1];
  int2* nwq = p.s.wideq[(level + 1) & 1];
  LevelSlot* ns = &p.s.ctl->slot[(level + 1) & 3];
  const unsigned lane = lane_id();
  WarpStage& sg = stage[threadIdx.x >> 5];
  const int64_t gw = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  unsigned long long edges_acc = 0;

  if (lane == 0) { sg.ncnt = 0; sg.wcnt = 0; }
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // recycle the slot two levels ahead
    LevelSlot* z = &p.s.ctl->slot[(level + 2) & 3];
    z->q_count = 0; z->wide_count = 0; z->edges = 0; z->bu_count = 0;
  }
  __syncwarp();

  const int64_t tiles = ((int64_t)st.cur_q + 31) >> 5;
  const int64_t work = tiles + (int64_t)st.cur_wide;
  for (int64_t item = gw; item < work; item += nwarps) {
    if (item < tiles) {
      const int64_t idx = item * 32 + lane;
      const bool has = idx < (int64_t)st.cur_q;
      const int32_t u = has ? q[idx] : 0;
      int64_t rs = 0;
      uint32_t d = 0;
      if (has) {
        rs = __ldg(p.g.rowptr + u);
        d = (uint32_t)(__ldg(p.g.rowptr + u + 1) - rs);
      }
      uint32_t incl = d;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += y;
      }
      const uint32_t excl = incl - d;
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t k = base + lane;
        int lo = 0, hi = 31;  // owner = largest lane whose exclusive prefix is <= k
#pragma unroll
        for (int s = 0; s < 5; ++s) {
          const int mid = (lo + hi + 1) >> 1;
          const uint32_t val = __shfl_sync(0xffffffffu, excl, mid);
          if (val <= k) lo = mid; else hi = mid - 1;
        }
        const uint32_t oe = __shfl_sync(0xffffffffu, excl, lo);
        const int64_t ors = __shfl_sync(0xffffffffu, rs, lo);
        const int32_t uu = __shfl_sync(0xffffffffu, u, lo);
        if (k < total) td_visit(p, level, sg, __ldg(p.g.col + ors + (k - oe)), uu, edges_acc);
        stage_flush(sg, ns, nq, nwq, false);
      }
    } else {
      const int2 it = wq[item - tiles];
      const int32_t u = it.x;
      const int64_t rs = __ldg(p.g.rowptr + u) + (int64_t)it.y * kChunk;
      const int64_t re = min(__ldg(p.g.rowptr + u + 1), rs + kChunk);
      // col is 16-byte aligned at index 0 and padded past its end: align the window down to 4 entries
      for (int64_t b0 = rs & ~(int64_t)3; b0 < re; b0 += 128) {  // warp-uniform trip count
        const int64_t base = b0 + 4 * lane;
        if (base < re) {
          const int4 x = __ldg(reinterpret_cast<const int4*>(p.g.col + base));
          if (base >= rs) td_visit(p, level, sg, x.x, u, edges_acc);
          if (base + 1 >= rs && base + 1 < re) td_visit(p, level, sg, x.y, u, edges_acc);
          if (base + 2 >= rs && base + 2 < re) td_visit(p, level, sg, x.z, u, edges_acc);
          if (base + 3 >= rs && base + 3 < re) td_visit(p, level, sg, x.w, u, edges_acc);
        }
        stage_flush(sg, ns, nq, nwq, false);
      }
    }
  }
  stage_flush(sg, ns, nq, nwq, true);
  block_accumulate(edges_acc, &ns->edges);
}

// ---------------------------------------------------------------------------------------------
// bottom-up: one warp per visited-bitmap word (32 consecutive vertices), one lane per vertex.
// A lane probes its first kBuProbe neighbours alone; rows still unresolved after that are finished
// by the whole warp with coalesced reads and an early exit on the first frontier neighbour.
// The next-frontier word and the visited word are written once per warp with plain stores.
// algorithmic bytes per level: n/8 (visited) + n/8 (next) + 16 B per unvisited vertex (row extent)
// + 4 B per neighbour entry inspected + 8 B per discovered vertex; frontier probes hit L2.
// ---------------------------------------------------------------------------------------------
__device__ void phase_bottom_up(const BfsParams& p, const LoopState& st) {
  const int level = st.level;
  const uint32_t* fr = p.s.bitmap[st.fb];
  uint32_t* nx = p.s.bitmap[st.fb ^ 1];
  LevelSlot* ns = &p.s.ctl->slot[(level + 1) & 3];
  const unsigned lane = lane_id();
  const int64_t gw = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  const int64_t n = p.g.n;
  unsigned long long edges_acc = 0, count_acc = 0;

  if (blockIdx.x == 0 && threadIdx.x == 0) {
    LevelSlot* z = &p.s.ctl->slot[(level + 2) & 3];
    z->q_count = 0; z->wide_count = 0; z->edges = 0; z->bu_count = 0;
  }

  for (int64_t w = gw; w < p.g.nwords; w += nwarps) {
    const uint32_t vis = p.s.visited[w];
    if (vis == 0xffffffffu) {
      if (lane == 0) nx[w] = 0;
      continue;
    }
    const int64_t v = w * 32 + lane;
    const bool act = !((vis >> lane) & 1u);
    const int64_t vc = v < n ? v : n - 1;  // padding lanes are never active; keep their loads in range
    const int64_t rs = __ldg(p.g.rowptr + vc), re = __ldg(p.g.rowptr + vc + 1);
    bool found = false;
    int32_t par = -1;
    if (act) {
      const int64_t lim = min(re, rs + kBuProbe);
      for (int64_t i = rs; i < lim; ++i) {
        const int32_t u = __ldg(p.g.col + i);
        if (test_bit(fr, u)) { found = true; par = u; break; }
      }
    }
    unsigned pend = __ballot_sync(0xffffffffu, act && !found && rs + kBuProbe < re);
    while (pend) {
      const int l = __ffs(pend) - 1;
      pend &= pend - 1;
      const int64_t ss = __shfl_sync(0xffffffffu, rs, l) + kBuProbe, ee = __shfl_sync(0xffffffffu, re, l);
      int32_t hit_par = -1;
      for (int64_t base = ss; base < ee; base += 32) {
        const int64_t i = base + lane;
        const bool a = i < ee;
        const int32_t u = a ? __ldg(p.g.col + i) : 0;
        const unsigned hm = __ballot_sync(0xffffffffu, a && test_bit(fr, u));
        if (hm) { hit_par = __shfl_sync(0xffffffffu, u, __ffs(hm) - 1); break; }
      }
      if ((int)lane == l && hit_par >= 0) { found = true; par = hit_par; }
    }
    const unsigned fm = __ballot_sync(0xffffffffu, found);
    if (found) {
      p.s.depth[v] = level + 1;
      p.s.parent[v] = par;
      edges_acc += (unsigned long long)(re - rs);
    }
    if (lane == 0) {
      nx[w] = fm;
      if (fm) p.s.visited[w] = vis | fm;
      count_acc += __popc(fm);
    }
  }
  block_accumulate(edges_acc, &ns->edges);
  block_accumulate(count_acc, &ns->bu_count);
}

// ---------------------------------------------------------------------------------------------
// representation changes at a direction switch
// ---------------------------------------------------------------------------------------------
// work lists of the new frontier (level st.level) -> bitmap[0]; both bitmaps are all-zero on entry
__device__ void phase_queue_to_bitmap(const BfsParams& p, const LoopState& st) {
  const int32_t* q = p.s.queue[st.level & 1];
  const int2* wq = p.s.wideq[st.level & 1];
  uint32_t* bm = p.s.bitmap[0];
  const int64_t tid = (int64_t)blockIdx.x * kBlock + threadIdx.x, nthreads = (int64_t)gridDim.x * kBlock;
  for (int64_t i = tid; i < (int64_t)st.cur_q; i += nthreads) {
    const int32_t v = q[i];
    atomicOr(&bm[(uint32_t)v >> 5], 1u << (v & 31));
  }
  for (int64_t i = tid; i < (int64_t)st.cur_wide; i += nthreads) {
    const int2 it = wq[i];
    if (it.y == 0) atomicOr(&bm[(uint32_t)it.x >> 5], 1u << (it.x & 31));
  }
}

// next bitmap (written by the bottom-up level that just ended) -> work lists of level st.level,
// routed by degree through the same per-warp staging; clears both bitmaps so the "all-zero while
// top-down" invariant holds again
__device__ void phase_bitmap_to_queue(const BfsParams& p, const LoopState& st, WarpStage* stage) {
  uint32_t* nx = p.s.bitmap[st.fb ^ 1];
  uint32_t* fr = p.s.bitmap[st.fb];
  int32_t* q = p.s.queue[st.level & 1];
  int2* wq = p.s.wideq[st.level & 1];
  LevelSlot* cs = &p.s.ctl->slot[st.level & 3];
  const unsigned lane = lane_id();
  WarpStage& sg = stage[threadIdx.x >> 5];
  if (lane == 0) { sg.ncnt = 0; sg.wcnt = 0; }
  __syncwarp();
  const int64_t gw = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
  for (int64_t w = gw; w < p.g.nwords; w += nwarps) {
    uint32_t bits = 0;
    if (lane == 0) {
      bits = nx[w];
      nx[w] = 0;
      fr[w] = 0;
    }
    bits = __shfl_sync(0xffffffffu, bits, 0);
    if (!bits) continue;
    if ((bits >> lane) & 1u) {
      const int64_t v = w * 32 + lane;
      route_vertex(sg, (int32_t)v, __ldg(p.g.rowptr + v + 1) - __ldg(p.g.rowptr + v));
    }
    stage_flush(sg, cs, q, wq, false);
  }
  stage_flush(sg, cs, q, wq, true);
}

// ---------------------------------------------------------------------------------------------
// stepwise engine: one kernel per phase
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) search_init_kernel(BfsParams p, int64_t src) { phase_init(p, src); }
__global__ void __launch_bounds__(kBlock) top_down_kernel(BfsParams p, LoopState st) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  phase_top_down(p, st, stage);
}
__global__ void __launch_bounds__(kBlock) bottom_up_kernel(BfsParams p, LoopState st) { phase_bottom_up(p, st); }
__global__ void __launch_bounds__(kBlock) queue_to_bitmap_kernel(BfsParams p, LoopState st) {
  phase_queue_to_bitmap(p, st);
}
__global__ void __launch_bounds__(kBlock) bitmap_to_queue_kernel(BfsParams p, LoopState st) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  phase_bitmap_to_queue(p, st, stage);
}

// ---------------------------------------------------------------------------------------------
// fused engine: the whole search in one persistent cooperative kernel.  One grid-wide barrier per
// level (two at a direction switch); every thread derives the same LoopState from the level
// counters, so no thread has to publish decisions.  Replaces the reference's per-level
// launch / cudaDeviceSynchronize / managed-counter read / cudaMemset round trips (bfs.cu:569-619).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int ld_volatile(const unsigned int* p) {
  return *reinterpret_cast<const volatile unsigned int*>(p);
}
__device__ __forceinline__ unsigned long long ld_volatile(const unsigned long long* p) {
  return *reinterpret_cast<const volatile unsigned long long*>(p);
}

__global__ void __launch_bounds__(kBlock) bfs_fused_kernel(BfsParams p, int64_t src) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  cg::grid_group grid = cg::this_grid();
  phase_init(p, src);
  LoopState st = initial_state(p, __ldg(p.g.rowptr + src + 1) - __ldg(p.g.rowptr + src));
  grid.sync();
  while (!st.done) {
    if (st.mode == 0) phase_top_down(p, st, stage); else phase_bottom_up(p, st);
    grid.sync();
    const LevelSlot* ns = &p.s.ctl->slot[(st.level + 1) & 3];
    const Conversion cv = advance(st, p.tn, p.g.n, ld_volatile(&ns->q_count), ld_volatile(&ns->wide_count),
                                  ld_volatile(&ns->bu_count), ld_volatile(&ns->edges));
    if (cv == kQueueToBitmap) {
      phase_queue_to_bitmap(p, st);
      grid.sync();
    } else if (cv == kBitmapToQueue) {
      phase_bitmap_to_queue(p, st, stage);
      grid.sync();
      const LevelSlot* cs = &p.s.ctl->slot[st.level & 3];
      st.cur_q = ld_volatile(&cs->q_count);
      st.cur_wide = ld_volatile(&cs->wide_count);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    p.s.ctl->result.levels = st.level;
    p.s.ctl->result.td_levels = st.td_levels;
    p.s.ctl->result.bu_levels = st.bu_levels;
  }
}

// ---------------------------------------------------------------------------------------------
// after the search (outside the timed window): V_r and D_r for the metric
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) reach_stats_kernel(int64_t n, const int32_t* __restrict__ depth,
                                                             const uint32_t* __restrict__ orig_deg, Control* ctl) {
  unsigned long long vr = 0, dr = 0;
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    if (depth[i] != B2BFS_UNREACHED_DEPTH) { vr++; dr += orig_deg[i]; }
  }
  block_accumulate(vr, &ctl->reached_vertices);
  block_accumulate(dr, &ctl->reached_entries);
}

// ---------------------------------------------------------------------------------------------
// host drivers
// ---------------------------------------------------------------------------------------------
Status search_grid_size(int device, int* grid_out) {
  int per_sm = 0, sms = 0;
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bfs_fused_kernel, kBlock, 0));
  B2_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  if (per_sm < 1) return make_error(B2BFS_ERR_CUDA, "bfs_fused_kernel does not fit on an SM");
  *grid_out = per_sm * sms;
  return Status();
}

Status run_fused(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, int64_t* launches) {
  BfsParams pp = p;
  int64_t s = src;
  void* args[] = {&pp, &s};
  B2_CUDA(cudaLaunchCooperativeKernel((const void*)bfs_fused_kernel, dim3(grid), dim3(kBlock), args, 0, stream));
  if (launches) *launches += 1;
  return Status();
}

// Host level loop.  `pinned` is a page-locked Control the counters are copied into each level.
Status run_stepwise(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                    int64_t* launches) {
  search_init_kernel<<<grid, kBlock, 0, stream>>>(p, src);
  B2_CUDA(cudaMemcpyAsync(pinned, p.s.ctl, sizeof(Control), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  LoopState st = initial_state(p, pinned->src_degree);
  int64_t nl = 1;
  static const bool trace = getenv("B2BFS_TRACE") != nullptr;
  while (!st.done) {
    if (trace)
      fprintf(stderr, "[b2bfs] level %d %s frontier q=%u wide=%u unexplored=%llu\n", st.level,
              st.mode ? "bottom-up" : "top-down", st.cur_q, st.cur_wide, st.unexplored);
    if (st.mode == 0) top_down_kernel<<<grid, kBlock, 0, stream>>>(p, st);
    else bottom_up_kernel<<<grid, kBlock, 0, stream>>>(p, st);
    ++nl;
    B2_CUDA(cudaMemcpyAsync(pinned, p.s.ctl, sizeof(Control), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    const LevelSlot& ns = pinned->slot[(st.level + 1) & 3];
    if (trace)
      fprintf(stderr, "[b2bfs]   found q=%u wide=%u bu=%llu edges=%llu\n", ns.q_count, ns.wide_count, ns.bu_count,
              ns.edges);
    const Conversion cv = advance(st, p.tn, p.g.n, ns.q_count, ns.wide_count, ns.bu_count, ns.edges);
    if (cv == kQueueToBitmap) {
      queue_to_bitmap_kernel<<<grid, kBlock, 0, stream>>>(p, st);
      ++nl;
    } else if (cv == kBitmapToQueue) {
      bitmap_to_queue_kernel<<<grid, kBlock, 0, stream>>>(p, st);
      ++nl;
      B2_CUDA(cudaMemcpyAsync(pinned, p.s.ctl, sizeof(Control), cudaMemcpyDeviceToHost, stream));
      B2_CUDA(cudaStreamSynchronize(stream));
      st.cur_q = pinned->slot[st.level & 3].q_count;
      st.cur_wide = pinned->slot[st.level & 3].wide_count;
    }
  }
  B2_CUDA(cudaGetLastError());
  pinned->result.levels = st.level;
  pinned->result.td_levels = st.td_levels;
  pinned->result.bu_levels = st.bu_levels;
  B2_CUDA(cudaMemcpyAsync(&p.s.ctl->result, &pinned->result, sizeof(SearchResult), cudaMemcpyHostToDevice, stream));
  if (launches) *launches += nl;
  return Status();
}

Status run_reach_stats(const BfsParams& p, const uint32_t* orig_deg, cudaStream_t stream, int64_t* launches) {
  B2_CUDA(cudaMemsetAsync(&p.s.ctl->reached_vertices, 0, 2 * sizeof(unsigned long long), stream));
  const int64_t blocks = (p.g.n + kBlock - 1) / kBlock;
  const int grid = (int)(blocks < (int64_t)kSMs * 8 ? (blocks > 0 ? blocks : 1) : (int64_t)kSMs * 8);
  reach_stats_kernel<<<grid, kBlock, 0, stream>>>(p.g.n, p.s.depth, orig_deg, p.s.ctl);
  B2_CUDA(cudaGetLastError());
  if (launches) *launches += 1;
  return Status();
}

}  // namespace b2
