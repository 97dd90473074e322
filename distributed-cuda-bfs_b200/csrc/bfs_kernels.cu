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

// Multi-rank top-down only: (vertex, parent) pairs whose vertex another rank owns, staged per warp and
// drained into per-owner send buckets (one tail atomic per owner per 32 staged pairs).
struct RemoteStage {
  int2 pair[kStageCap];
  int cnt;
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

__device__ __forceinline__ unsigned int ld_volatile(const unsigned int* p) {
  return *reinterpret_cast<const volatile unsigned int*>(p);
}
__device__ __forceinline__ unsigned long long ld_volatile(const unsigned long long* p) {
  return *reinterpret_cast<const volatile unsigned long long*>(p);
}

// Route a newly discovered vertex of degree d to the work lists of level `items_level` (lane-local).
// Hub rows (the CTA-level bin; Tuning::huge_min == 0 switches it off - the host sets 0 whenever the
// hub lists do not exist) are rare: everything they need is fetched inside their branch, so the
// common path carries no extra registers.
__device__ __forceinline__ void route_vertex(WarpStage& sg, int32_t v, int64_t d, const BfsParams& p, int items_level) {
  if (p.tn.huge_min > 0 && d >= p.tn.huge_min) {  // a handful of vertices per search: one tail atomic each is fine
    const unsigned items = (unsigned)((d + kCtaChunk - 1) / kCtaChunk);
    const unsigned base = atomicAdd(&p.s.ctl->slot[items_level & 3].huge_count, items);
    int2* hq = p.s.hugeq[items_level & 1];
    for (unsigned c = 0; c < items; ++c) hq[base + c] = make_int2(v, (int)c);
  } else if (d < kNarrow) {
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
    // one tail atomic for everything staged: total first, then per-round prefixes from a running base
    unsigned mine = 0;
    for (int i = lane; i < wc; i += 32) mine += (unsigned)sg.witems[i];
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    unsigned run = 0;
    if (lane == 0) run = atomicAdd(&ns->wide_count, mine);
    run = __shfl_sync(0xffffffffu, run, 0);
    for (int i0 = 0; i0 < wc; i0 += 32) {
      const int i = i0 + (int)lane;
      const unsigned items = i < wc ? (unsigned)sg.witems[i] : 0u;
      const int32_t v = i < wc ? sg.wide[i] : 0;
      unsigned incl = items;
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += y;
      }
      const unsigned base = run + incl - items;
      run += __shfl_sync(0xffffffffu, incl, 31);
      if (items && items <= 8) {
        for (unsigned c = 0; c < items; ++c) nwq[base + c] = make_int2(v, (int)c);
      }
      unsigned big = __ballot_sync(0xffffffffu, items > 8);  // long rows: the whole warp writes the items
      while (big) {
        const int l = __ffs(big) - 1;
        big &= big - 1;
        const unsigned bi = __shfl_sync(0xffffffffu, items, l), bb = __shfl_sync(0xffffffffu, base, l);
        const int32_t bv = __shfl_sync(0xffffffffu, v, l);
        for (unsigned c = lane; c < bi; c += 32) nwq[bb + c] = make_int2(bv, (int)c);
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
__device__ __forceinline__ void phase_init(const BfsParams& p, int64_t src) {
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
      p.s.parent[src] = (int32_t)(src + p.g.vbase);
    }
  }
  for (int64_t i = (n4 << 2) + tid; i < n; i += nthreads) {
    p.s.depth[i] = (i == src) ? 0 : B2BFS_UNREACHED_DEPTH;
    p.s.parent[i] = (i == src) ? (int32_t)(src + p.g.vbase) : -1;
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
      c->slot[k].huge_count = 0;
      c->slot[k].edges = 0;
      c->slot[k].bu_count = 0;
      c->slot[k].bu_cursor = 0;
    }
    const int64_t deg = src >= 0 ? p.g.rowptr[src + 1] - p.g.rowptr[src] : 0;  // src < 0: another rank owns it
    if (src < 0) {
    } else if (p.tn.huge_min > 0 && deg >= p.tn.huge_min) {
      const unsigned items = (unsigned)((deg + kCtaChunk - 1) / kCtaChunk);
      for (unsigned k = 0; k < items; ++k) p.s.hugeq[0][k] = make_int2((int)src, (int)k);
      c->slot[0].huge_count = items;
    } else if (deg < kNarrow) {
      p.s.queue[0][0] = (int32_t)src;
      c->slot[0].q_count = 1;
    } else {
      const unsigned items = (unsigned)((deg + kChunk - 1) / kChunk);
      for (unsigned k = 0; k < items; ++k) p.s.wideq[0][k] = make_int2((int)src, (int)k);
      c->slot[0].wide_count = items;
    }
    c->result.levels = c->result.td_levels = c->result.bu_levels = 0;
    c->src_degree = deg;
    c->group_barrier = 0;
    c->group_exit_seq = 0;
  }
}

__host__ __device__ inline LoopState initial_state(const BfsParams& p, int64_t src_degree) {
  LoopState st;
  st.level = 0;
  st.mode = 0;
  st.fb = 0;
  st.done = 0;
  const bool hub = p.tn.huge_min > 0 && src_degree >= p.tn.huge_min;
  const bool wide = src_degree >= kNarrow && !hub;
  st.cur_q = (wide || hub) ? 0u : 1u;
  st.cur_wide = wide ? (unsigned)((src_degree + kChunk - 1) / kChunk) : 0u;
  st.prev_nf = 1;
  st.unexplored = (unsigned long long)(p.g.entries - src_degree);
  st.frontier_edges = (unsigned long long)src_degree;
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
// Warp-uniform: drain staged remote pairs into the send bucket of each owner.
__device__ __forceinline__ void remote_flush(const BfsParams& p, RemoteStage& rs, bool force) {
  __syncwarp();
  const unsigned lane = lane_id();
  const int cnt = *reinterpret_cast<volatile int*>(&rs.cnt);
  if (cnt < kStageFlush && !(force && cnt > 0)) return;
  for (int i0 = 0; i0 < cnt; i0 += 32) {
    const int i = i0 + (int)lane;
    const bool has = i < cnt;
    const int2 e = has ? rs.pair[i] : make_int2(0, 0);
    const int owner = has ? (int)(e.x / p.g.part) : -1;
    for (int d = 0; d < p.g.world; ++d) {
      const unsigned m = __ballot_sync(0xffffffffu, owner == d);
      if (!m) continue;
      unsigned base = 0;
      const int leader = __ffs(m) - 1;
      if ((int)lane == leader) base = atomicAdd(&p.s.send_count[d * 32], (unsigned)__popc(m));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (owner == d) {
        const int64_t pos = base + __popc(m & ((1u << lane) - 1u));
        if (p.s.xbase[0])  // straight into rank d's inbox bucket for this rank, over NVLink
          reinterpret_cast<int2*>(p.s.xbase[d] + xchg_off_inbox(p.g.world, p.g.part))[(int64_t)p.g.rank * p.s.send_cap + pos] = e;
        else
          p.s.send_buf[(int64_t)d * p.s.send_cap + pos] = e;
      }
    }
  }
  if (p.s.xbase[0]) __threadfence_system();  // the pairs are in the owners' memory before this warp moves on
  __syncwarp();
  if (lane == 0) rs.cnt = 0;
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// CTA-level bin: the adjacency of a hub row reaches shared memory through the asynchronous copy
// engine (cp.async.bulk, 1-D TMA; UBLKCP in SASS) behind an mbarrier, two buffers deep, so the
// next chunk of kCtaChunk entries is in flight while the eight warps of the CTA probe and claim
// the current one - no warp ever waits on a col[] load of its own.
// ---------------------------------------------------------------------------------------------
struct HubStage {
  alignas(128) int32_t buf[2][kCtaChunk + 4];  // + 4: a chunk starts at a 16-byte boundary below its first entry
  alignas(8) unsigned long long mbar[2];
  int64_t rs[2], re[2];                        // entries [rs, re) of col are the chunk; buf holds [rs & ~3, ...)
  int32_t u[2];
  unsigned int seq;                            // items this CTA has consumed so far (buffer = seq & 1, parity = seq >> 1 & 1)
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void hub_stage_init(HubStage* hs) {  // one thread, once per kernel
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&hs->mbar[0])));
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&hs->mbar[1])));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  hs->seq = 0;
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t mbar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
               : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
  return ok != 0;
}
// one thread: start the copy of item `it` of the frontier's hub list into buffer b
__device__ __forceinline__ void hub_issue(const BfsParams& p, HubStage* hs, int b, const int2 it) {
  const int64_t rs = __ldg(p.g.rowptr + it.x) + (int64_t)it.y * kCtaChunk;
  const int64_t re = min(__ldg(p.g.rowptr + it.x + 1), rs + kCtaChunk);
  const int64_t a0 = rs & ~(int64_t)3;
  const uint32_t bytes = (uint32_t)(((re - a0 + 3) & ~(int64_t)3) * (int64_t)sizeof(int32_t));  // multiple of 16; col is padded
  hs->u[b] = it.x;
  hs->rs[b] = rs;
  hs->re[b] = re;
  const uint32_t mbar = smem_addr(&hs->mbar[b]);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_addr(&hs->buf[b][0])), "l"(p.g.col + a0), "r"(bytes), "r"(mbar) : "memory");
}

// Four (u -> v) entries per lane at a time, phase by phase, so that the four visited probes, the
// four claims and the four row-extent reads are each in flight together instead of forming one
// long dependent chain per entry.  v and u are GLOBAL ids; with kMulti a vertex another rank owns is
// filtered through the `seen` bitmap and staged for its owner instead of being claimed here.
template <bool kMulti>
__device__ __forceinline__ void td_visit4(const BfsParams& p, int level, WarpStage& sg, RemoteStage* rs,
                                          const int32_t (&v)[4], const int32_t (&u)[4], const bool (&valid)[4],
                                          unsigned long long& edges_acc, bool mark, bool claim_only) {
  uint32_t word[4];
  bool c[4], local[4];
  int32_t vl[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    vl[k] = v[k] - (int32_t)p.g.vbase;
    local[k] = !kMulti || (vl[k] >= 0 && vl[k] < (int32_t)p.g.n);
  }
  // cheap read first; a stale "clear" only costs one extra atomic (bits only go 0 -> 1 in a search)
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t* bm = local[k] ? p.s.visited : p.s.seen;
    const uint32_t idx = (uint32_t)(local[k] ? vl[k] : v[k]) >> 5;
    word[k] = valid[k] ? bm[idx] : 0xffffffffu;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) c[k] = !((word[k] >> (v[k] & 31)) & 1u);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (c[k]) {
      uint32_t* bm = local[k] ? p.s.visited : p.s.seen;
      const uint32_t idx = (uint32_t)(local[k] ? vl[k] : v[k]) >> 5;
      word[k] = atomicOr(&bm[idx], 1u << (v[k] & 31));
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) c[k] = c[k] && !((word[k] >> (v[k] & 31)) & 1u);
  if (claim_only) {  // no work lists for the next level: claim, mark, count (edges_acc counts CLAIMS here)
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (c[k]) {
        if (local[k]) {
          p.s.depth[vl[k]] = level + 1;
          p.s.parent[vl[k]] = u[k];
          edges_acc++;
          atomicOr(&p.s.bitmap[0][(uint32_t)vl[k] >> 5], 1u << (vl[k] & 31));
        } else if (kMulti) {
          rs->pair[atomicAdd(&rs->cnt, 1)] = make_int2(v[k], u[k]);
        }
      }
    }
    return;
  }
  int64_t d[4];
#pragma unroll
  for (int k = 0; k < 4; ++k)
    d[k] = (c[k] && local[k]) ? __ldg(p.g.rowptr + vl[k] + 1) - __ldg(p.g.rowptr + vl[k]) : 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (c[k]) {
      if (local[k]) {
        p.s.depth[vl[k]] = level + 1;
        p.s.parent[vl[k]] = u[k];
        edges_acc += (unsigned long long)d[k];
        route_vertex(sg, vl[k], d[k], p, level + 1);
        // the next level will most likely run bottom-up: leave the frontier bit behind now and
        // save the queue->bitmap pass and its grid-wide barrier
        if (mark) atomicOr(&p.s.bitmap[0][(uint32_t)vl[k] >> 5], 1u << (vl[k] & 31));
      } else if (kMulti) {
        rs->pair[atomicAdd(&rs->cnt, 1)] = make_int2(v[k], u[k]);
      }
    }
  }
}

template <bool kMulti>
__device__ __forceinline__ void phase_top_down(const BfsParams& p, const LoopState& st, WarpStage* stage, int bid, int nb,
                                               RemoteStage* rstage = nullptr, bool mark = false, bool claim_only = false,
                                               HubStage* hs = nullptr) {
  const int level = st.level;
  const int32_t* q = p.s.queue[level & 1];
  const int2* wq = p.s.wideq[level & 1];
  int32_t* nq = p.s.queue[(level + 1) & 1];
  int2* nwq = p.s.wideq[(level + 1) & 1];
  LevelSlot* ns = &p.s.ctl->slot[(level + 1) & 3];
  const unsigned lane = lane_id();
  WarpStage& sg = stage[threadIdx.x >> 5];
  const int64_t gw = (int64_t)bid * kWarpsPerBlock + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)nb * kWarpsPerBlock;
  unsigned long long edges_acc = 0;

  RemoteStage* rst = kMulti ? &rstage[threadIdx.x >> 5] : nullptr;
  const int32_t vbase = (int32_t)p.g.vbase;
  if (lane == 0) { sg.ncnt = 0; sg.wcnt = 0; }
  if (kMulti && lane == 0) rst->cnt = 0;
  if (bid == 0 && threadIdx.x == 0) {  // recycle the slot two levels ahead
    LevelSlot* z = &p.s.ctl->slot[(level + 2) & 3];
    z->q_count = 0; z->wide_count = 0; z->huge_count = 0; z->edges = 0; z->bu_count = 0; z->bu_cursor = 0;
  }
  __syncwarp();

  // ---- hub rows first (they are the long poles): CTA items, block-uniform control flow
  if (!kMulti && hs != nullptr) {
    const int64_t nh = (int64_t)ld_volatile(&p.s.ctl->slot[level & 3].huge_count);  // final since the last barrier
    const int2* hq = p.s.hugeq[level & 1];
    unsigned seq = hs->seq;  // uniform: written back below, behind a block barrier
    if (threadIdx.x == 0 && bid < nh) hub_issue(p, hs, seq & 1, __ldcg(hq + bid));
    for (int64_t ci = bid; ci < nh; ci += nb) {
      const int b = seq & 1;
      if (threadIdx.x == 0 && ci + nb < nh) hub_issue(p, hs, b ^ 1, __ldcg(hq + ci + nb));  // next chunk in flight
      while (!mbar_try_wait(smem_addr(&hs->mbar[b]), (seq >> 1) & 1)) {}
      const int64_t rs = hs->rs[b], re = hs->re[b], a0 = rs & ~(int64_t)3;
      const int32_t u = hs->u[b];
      // the buffer holds re - a0 <= kCtaChunk + 3 positions: 128 per warp and round, rounds strided over the warps
      for (int64_t off = (int64_t)(threadIdx.x >> 5) * 128; off < re - a0; off += kWarpsPerBlock * 128) {  // warp-uniform
        const int64_t idx = off + 4 * lane;  // position in the buffer; entry a0 + idx of col
        if (a0 + idx < re) {
          const int4 x = *reinterpret_cast<const int4*>(&hs->buf[b][idx]);
          const int32_t v[4] = {x.x, x.y, x.z, x.w};
          const int32_t uu[4] = {u + vbase, u + vbase, u + vbase, u + vbase};
          const int64_t e = a0 + idx;
          const bool valid[4] = {e >= rs, e + 1 >= rs && e + 1 < re, e + 2 >= rs && e + 2 < re, e + 3 >= rs && e + 3 < re};
          td_visit4<kMulti>(p, level, sg, rst, v, uu, valid, edges_acc, mark, claim_only);
        }
        stage_flush(sg, ns, nq, nwq, false);
      }
      __syncthreads();  // every warp is done with buffer b before it is refilled
      ++seq;
    }
    if (threadIdx.x == 0) hs->seq = seq;
    __syncthreads();
  }

  const int64_t tiles = ((int64_t)st.cur_q + 31) >> 5;
  const int64_t work = tiles + (int64_t)st.cur_wide;
  for (int64_t item = gw; item < work; item += nwarps) {
    if (item < tiles) {
      const int64_t idx = item * 32 + lane;
      const bool has = idx < (int64_t)st.cur_q;
      const int32_t u = has ? __ldcg(q + idx) : 0;  // written last level by other SMs: read at L2
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
      for (uint32_t base = 0; base < total; base += 128) {  // 4 entries per lane per step
        int32_t v[4], uu[4];
        bool valid[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t k = base + 32 * j + lane;
          int lo = 0, hi = 31;  // owner = largest lane whose exclusive prefix is <= k
#pragma unroll
          for (int s = 0; s < 5; ++s) {
            const int mid = (lo + hi + 1) >> 1;
            const uint32_t val = __shfl_sync(0xffffffffu, excl, mid);
            if (val <= k) lo = mid; else hi = mid - 1;
          }
          const uint32_t oe = __shfl_sync(0xffffffffu, excl, lo);
          const int64_t ors = __shfl_sync(0xffffffffu, rs, lo);
          uu[j] = __shfl_sync(0xffffffffu, u, lo) + vbase;
          valid[j] = k < total;
          v[j] = valid[j] ? __ldg(p.g.col + ors + (k - oe)) : 0;
        }
        td_visit4<kMulti>(p, level, sg, rst, v, uu, valid, edges_acc, mark, claim_only);
        stage_flush(sg, ns, nq, nwq, false);
        if (kMulti) remote_flush(p, *rst, false);
      }
    } else {
      const int2 it = __ldcg(wq + (item - tiles));
      const int32_t u = it.x;
      const int64_t rs = __ldg(p.g.rowptr + u) + (int64_t)it.y * kChunk;
      const int64_t re = min(__ldg(p.g.rowptr + u + 1), rs + kChunk);
      // col is 16-byte aligned at index 0 and padded past its end: align the window down to 4 entries
      for (int64_t b0 = rs & ~(int64_t)3; b0 < re; b0 += 128) {  // warp-uniform trip count
        const int64_t base = b0 + 4 * lane;
        if (base < re) {
          const int4 x = __ldg(reinterpret_cast<const int4*>(p.g.col + base));
          const int32_t v[4] = {x.x, x.y, x.z, x.w};
          const int32_t uu[4] = {u + vbase, u + vbase, u + vbase, u + vbase};
          const bool valid[4] = {base >= rs, base + 1 >= rs && base + 1 < re, base + 2 >= rs && base + 2 < re,
                                 base + 3 >= rs && base + 3 < re};
          td_visit4<kMulti>(p, level, sg, rst, v, uu, valid, edges_acc, mark, claim_only);
        }
        stage_flush(sg, ns, nq, nwq, false);
        if (kMulti) remote_flush(p, *rst, false);
      }
    }
  }
  stage_flush(sg, ns, nq, nwq, true);
  if (kMulti) remote_flush(p, *rst, true);
  block_accumulate(edges_acc, claim_only ? &ns->bu_count : &ns->edges);
}

// Multi-rank: claim the (vertex, parent) pairs other ranks discovered in this rank's rows and queue
// the winners for the next level (same slot counters and staging as the local discoveries).  The
// pairs sit in `nb` buckets of `stride` pairs each (bucket s holds counts[s] of them); all buckets
// are walked in ONE pass, so the phase pays its staging flush and its block reduction once.
__device__ __forceinline__ void phase_apply_remote(const BfsParams& p, const LoopState& st, WarpStage* stage,
                                                   const int2* recv, int64_t stride, const unsigned int* counts, int nb,
                                                   bool mark = false, bool claim_only = false) {
  const int level = st.level;
  int32_t* nq = p.s.queue[(level + 1) & 1];
  int2* nwq = p.s.wideq[(level + 1) & 1];
  LevelSlot* ns = &p.s.ctl->slot[(level + 1) & 3];
  WarpStage& sg = stage[threadIdx.x >> 5];
  if (lane_id() == 0) { sg.ncnt = 0; sg.wcnt = 0; }
  __syncwarp();
  unsigned long long edges_acc = 0;
  int64_t count = 0;
  for (int b = 0; b < nb; ++b) count += counts[b];
  const int64_t nthreads = (int64_t)gridDim.x * kBlock;
  const int64_t rounds = (count + nthreads - 1) / nthreads;
  for (int64_t r = 0; r < rounds; ++r) {  // warp-uniform trip count
    int64_t i = r * nthreads + (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (i < count) {
      int b = 0;
      while (i >= (int64_t)counts[b]) { i -= counts[b]; ++b; }
      const int2 e = __ldcg(recv + (int64_t)b * stride + i);  // written by another GPU (or by NCCL): read at L2
      const int32_t vl = e.x - (int32_t)p.g.vbase;
      const uint32_t w = (uint32_t)vl >> 5, bit = 1u << (vl & 31);
      if ((uint32_t)vl < (uint32_t)p.g.n && !(p.s.visited[w] & bit) && !(atomicOr(&p.s.visited[w], bit) & bit)) {
        p.s.depth[vl] = level + 1;
        p.s.parent[vl] = e.y;
        if (claim_only) {
          edges_acc++;  // counts claims
        } else {
          const int64_t d = __ldg(p.g.rowptr + vl + 1) - __ldg(p.g.rowptr + vl);
          edges_acc += (unsigned long long)d;
          route_vertex(sg, vl, d, p, level + 1);
        }
        if (mark) atomicOr(&p.s.bitmap[0][w], bit);
      }
    }
    stage_flush(sg, ns, nq, nwq, false);
  }
  stage_flush(sg, ns, nq, nwq, true);
  block_accumulate(edges_acc, claim_only ? &ns->bu_count : &ns->edges);
}

// ---------------------------------------------------------------------------------------------
// bottom-up, in three tiers of decreasing population and increasing cost per vertex.
//   tier 1  one lane per VERTEX of a visited-bitmap word that still has unvisited bits: read
//           first_nbr[v] (coalesced, 4 B per vertex, no rowptr/col access) and probe the frontier
//           bitmap.  In a power-law graph the lowest-id neighbour is usually a hub, so most of the
//           first bottom-up level ends here.  Software-pipelined over the words of a 32-word batch
//           (load for word j+2, probe for word j+1, evaluate word j).  Batches are claimed from an
//           atomic cursor, so warps that draw sparse regions simply take more batches.
//   tier 2  misses are compacted into a per-warp list in shared memory and taken 32 at a time, one
//           lane per vertex: row extent -> neighbours 2..5 -> four probes, as a four-stage pipeline
//           across rounds, so the rowptr -> col -> bitmap chain of one vertex overlaps its
//           neighbours' in the list.
//   tier 3  rows still unresolved after five neighbours: 8 lanes per vertex, 32-byte reads, early exit.
// The warp that owns a batch clears its next-frontier words first; all finds are then OR-ed into the
// next-frontier and visited words (one atomic per word in tier 1, per vertex in tiers 2 and 3 —
// distinct addresses, no hot line).
// algorithmic bytes per level: n/8 (visited) + n/8 (next) + 4 B per unvisited vertex (first_nbr)
// + 16 B row extent and 4 B per further neighbour inspected for tier-2/3 vertices + 8 B per
// discovered vertex; frontier probes hit L1/L2.
// ---------------------------------------------------------------------------------------------
constexpr int kBuT2 = 4;                       // neighbours 2..5 are probed in tier 2
constexpr int kT2Cap = 2 * kStageCap;          // WarpStage::narrow + ::wide, contiguous
constexpr int kT3Cap = kStageCap;              // WarpStage::witems

__device__ __forceinline__ void bu_found(const BfsParams& p, int level, uint32_t* nx, int32_t v, int32_t par) {
  p.s.depth[v] = level + 1;
  p.s.parent[v] = par;
  atomicOr(&nx[(uint32_t)v >> 5], 1u << (v & 31));
  atomicOr(&p.s.visited[(uint32_t)v >> 5], 1u << (v & 31));
}

// tier 3: finish the listed vertices from their 6th neighbour on, 8 lanes per vertex
__device__ __forceinline__ void bu_tier3(const BfsParams& p, int level, const uint32_t* fr, uint32_t* nx,
                                         int32_t* list, int* cnt_p, unsigned long long& count_acc) {
  __syncwarp();
  const unsigned lane = lane_id();
  const int cnt = *reinterpret_cast<volatile int*>(cnt_p);
  const int g = lane >> 3, gl = lane & 7;
  for (int base = 0; base < cnt; base += 4) {
    const int idx = base + g;
    const bool has = idx < cnt;
    const int32_t v = has ? list[idx] : 0;
    int64_t pos = 0, re = 0;
    if (has) {
      pos = __ldg(p.g.rowptr + v) + 1 + kBuT2;
      re = __ldg(p.g.rowptr + v + 1);
    }
    bool active = has && pos < re;
    int32_t par = -1;
    while (__ballot_sync(0xffffffffu, active)) {
      const int64_t i = pos + gl;
      const bool a = active && i < re;
      const int32_t u = a ? __ldg(p.g.col + i) : 0;
      const unsigned bm = __ballot_sync(0xffffffffu, a && test_bit(fr, u));
      const unsigned gm = (bm >> (g * 8)) & 0xffu;
      const int32_t pu = __shfl_sync(0xffffffffu, u, gm ? g * 8 + __ffs(gm) - 1 : (int)lane);
      if (active) {
        if (gm) {
          par = pu;
          active = false;
        } else {
          pos += 8;
          active = pos < re;
        }
      }
    }
    if (has && gl == 0 && par >= 0) {
      bu_found(p, level, nx, v, par);
      count_acc++;
    }
  }
  __syncwarp();
  if (lane == 0) *cnt_p = 0;
  __syncwarp();
}

// tier 2: listed vertices, one per lane, 4-stage pipeline across rounds of 32
__device__ __forceinline__ void bu_tier2(const BfsParams& p, int level, const uint32_t* fr, uint32_t* nx,
                                         WarpStage& sg, unsigned long long& count_acc) {
  __syncwarp();
  const unsigned lane = lane_id();
  int32_t* list = sg.narrow;  // kT2Cap entries (narrow + wide)
  const int cnt = *reinterpret_cast<volatile int*>(&sg.ncnt);
  const int rounds = (cnt + 31) >> 5;
  int32_t vA = -1, vB = -1, vC = -1;
  int64_t rsA = 0, reA = 0;
  int nB = 0, nC = 0;  // neighbours 2..5 actually present
  bool moreB = false, moreC = false;
  int32_t uB[kBuT2], uC[kBuT2];
  uint32_t pw[kBuT2];
#pragma unroll
  for (int k = 0; k < kBuT2; ++k) { uB[k] = 0; uC[k] = 0; pw[k] = 0; }
  for (int r = 0; r < rounds + 3; ++r) {
    // D: evaluate
    if (vC >= 0) {
      int32_t par = -1;
#pragma unroll
      for (int k = kBuT2 - 1; k >= 0; --k)
        if (k < nC && ((pw[k] >> (uC[k] & 31)) & 1u)) par = uC[k];
      if (par >= 0) {
        bu_found(p, level, nx, vC, par);
        count_acc++;
      } else if (moreC) {
        sg.witems[atomicAdd(&sg.wcnt, 1)] = vC;
      }
    }
    // C: probes for the neighbours that arrived
    vC = vB; nC = nB; moreC = moreB;
#pragma unroll
    for (int k = 0; k < kBuT2; ++k) uC[k] = uB[k];
    if (vC >= 0) {
#pragma unroll
      for (int k = 0; k < kBuT2; ++k) pw[k] = fr[(uint32_t)uC[k] >> 5];
    }
    // B: neighbours 2..5 of the row whose extent arrived
    vB = vA;
    if (vB >= 0) {
      const int64_t d = reA - rsA;
      nB = (int)min((int64_t)kBuT2, d - 1);
      moreB = d > 1 + kBuT2;
      if (nB <= 0) {
        vB = -1;  // degree 1: its only neighbour was probed in tier 1
      } else {
#pragma unroll
        for (int k = 0; k < kBuT2; ++k) uB[k] = __ldg(p.g.col + rsA + 1 + (k < nB ? k : 0));
      }
    }
    // A: next list entry
    const int i = r * 32 + (int)lane;
    if (r < rounds && i < cnt) {
      vA = list[i];
      rsA = __ldg(p.g.rowptr + vA);
      reA = __ldg(p.g.rowptr + vA + 1);
    } else {
      vA = -1;
    }
    __syncwarp();
    if (*reinterpret_cast<volatile int*>(&sg.wcnt) > kT3Cap - 32) bu_tier3(p, level, fr, nx, sg.witems, &sg.wcnt, count_acc);
  }
  __syncwarp();
  if (lane == 0) sg.ncnt = 0;
  __syncwarp();
}

__device__ __forceinline__ void phase_bottom_up(const BfsParams& p, const LoopState& st, WarpStage* stage,
                                                const uint32_t* frontier_override = nullptr) {
  const int level = st.level;
  // multi-rank: probe the allgathered frontier of all ranks (global ids), write this rank's slice
  const uint32_t* fr = frontier_override ? frontier_override : p.s.gfrontier ? p.s.gfrontier : p.s.bitmap[st.fb];
  uint32_t* nx = p.s.bitmap[st.fb ^ 1];
  LevelSlot* ns = &p.s.ctl->slot[(level + 1) & 3];
  const unsigned lane = lane_id();
  WarpStage& sg = stage[threadIdx.x >> 5];
  unsigned long long count_acc = 0;

  if (blockIdx.x == 0 && threadIdx.x == 0) {
    LevelSlot* z = &p.s.ctl->slot[(level + 2) & 3];
    z->q_count = 0; z->wide_count = 0; z->huge_count = 0; z->edges = 0; z->bu_count = 0; z->bu_cursor = 0;
  }
  if (lane == 0) { sg.ncnt = 0; sg.wcnt = 0; }
  __syncwarp();
  int t2n = 0;  // warp-uniform mirror of sg.ncnt while tier 1 appends

  for (;;) {
    unsigned long long w0u = 0;
    if (lane == 0) w0u = atomicAdd(&ns->bu_cursor, 32ull);
    const int64_t w0 = (int64_t)__shfl_sync(0xffffffffu, w0u, 0);
    if (w0 >= p.g.nwords) break;
    const int64_t w = w0 + lane;
    const bool inr = w < p.g.nwords;
    const uint32_t vis = inr ? p.s.visited[w] : 0xffffffffu;
    const uint32_t todo = ~vis;
    uint32_t found_bits = 0;
    // clear the stale word now; every find of this level is OR-ed in afterwards (tier 2 may run
    // in the middle of a batch and must not be overwritten by a later whole-word store)
    if (inr) nx[w] = 0;
    unsigned words = __ballot_sync(0xffffffffu, todo != 0);
    // tier 1, pipelined over the words of this batch that have unvisited vertices
    int jA = -1, jB = -1;
    bool myA = false, myB = false;
    int32_t fA = 0, fB = 0;
    uint32_t pwB = 0;
    while (words || jA >= 0 || jB >= 0) {  // warp-uniform
      // evaluate word jB
      if (jB >= 0) {
        const bool hit = myB && ((pwB >> (fB & 31)) & 1u);
        const int64_t v = (w0 + jB) * 32 + lane;
        if (hit) {
          p.s.depth[v] = level + 1;
          p.s.parent[v] = fB;
        }
        const unsigned fm = __ballot_sync(0xffffffffu, hit);
        if ((int)lane == jB) found_bits |= fm;
        const unsigned mm = __ballot_sync(0xffffffffu, myB && !hit);
        if (myB && !hit) sg.narrow[t2n + __popc(mm & ((1u << lane) - 1u))] = (int32_t)v;
        t2n += __popc(mm);
      }
      // probe for word jA
      jB = jA; myB = myA; fB = fA;
      if (jB >= 0 && myB) pwB = fr[(uint32_t)fB >> 5];
      // first neighbours of the next word with unvisited vertices
      if (words) {
        jA = __ffs(words) - 1;
        words &= words - 1;
        const uint32_t tw = __shfl_sync(0xffffffffu, todo, jA);
        myA = (tw >> lane) & 1u;
        fA = myA ? __ldg(p.g.first_nbr + (w0 + jA) * 32 + lane) : 0;
      } else {
        jA = -1;
      }
      if (t2n > kT2Cap - 32) {
        __syncwarp();
        if (lane == 0) sg.ncnt = t2n;
        bu_tier2(p, level, fr, nx, sg, count_acc);
        t2n = 0;
      }
    }
    if (inr && found_bits) {
      atomicOr(&nx[w], found_bits);
      atomicOr(&p.s.visited[w], found_bits);
      count_acc += __popc(found_bits);
    }
    __syncwarp();
  }
  if (lane == 0) sg.ncnt = t2n;
  bu_tier2(p, level, fr, nx, sg, count_acc);
  bu_tier3(p, level, fr, nx, sg.witems, &sg.wcnt, count_acc);
  block_accumulate(count_acc, &ns->bu_count);
}

// ---------------------------------------------------------------------------------------------
// representation changes at a direction switch
// ---------------------------------------------------------------------------------------------
// work lists of the new frontier (level st.level) -> bitmap[0]; both bitmaps are all-zero on entry
__device__ __forceinline__ void phase_queue_to_bitmap(const BfsParams& p, const LoopState& st) {
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
  if (p.s.hugeq[0]) {
    const int2* hq = p.s.hugeq[st.level & 1];
    const int64_t nh = (int64_t)ld_volatile(&p.s.ctl->slot[st.level & 3].huge_count);
    for (int64_t i = tid; i < nh; i += nthreads) {
      const int2 it = hq[i];
      if (it.y == 0) atomicOr(&bm[(uint32_t)it.x >> 5], 1u << (it.x & 31));
    }
  }
}

// take back the frontier bits a "marking" top-down level left in bitmap[0] when the search stays
// top-down after all (the bitmaps must be all-zero while top-down runs)
__device__ __forceinline__ void phase_unmark(const BfsParams& p, const LoopState& st) {
  const int32_t* q = p.s.queue[st.level & 1];
  const int2* wq = p.s.wideq[st.level & 1];
  uint32_t* bm = p.s.bitmap[0];
  const int64_t tid = (int64_t)blockIdx.x * kBlock + threadIdx.x, nthreads = (int64_t)gridDim.x * kBlock;
  for (int64_t i = tid; i < (int64_t)st.cur_q; i += nthreads) bm[(uint32_t)__ldcg(q + i) >> 5] = 0;
  for (int64_t i = tid; i < (int64_t)st.cur_wide; i += nthreads) {
    const int2 it = __ldcg(wq + i);
    if (it.y == 0) bm[(uint32_t)it.x >> 5] = 0;
  }
  if (p.s.hugeq[0]) {
    const int2* hq = p.s.hugeq[st.level & 1];
    const int64_t nh = (int64_t)ld_volatile(&p.s.ctl->slot[st.level & 3].huge_count);
    for (int64_t i = tid; i < nh; i += nthreads) {
      const int2 it = __ldcg(hq + i);
      if (it.y == 0) bm[(uint32_t)it.x >> 5] = 0;
    }
  }
}

// next bitmap (written by the bottom-up level that just ended) -> work lists of level st.level,
// routed by degree through the same per-warp staging; clears both bitmaps so the "all-zero while
// top-down" invariant holds again
__device__ __forceinline__ void phase_bitmap_to_queue(const BfsParams& p, const LoopState& st, WarpStage* stage) {
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
  // one bitmap word per LANE (coalesced 128-byte reads and clears); set bits are peeled four per
  // lane per round so a warp never stages more than 128 ids between two flush checks
  for (int64_t w0 = gw * 32; w0 < p.g.nwords; w0 += nwarps * 32) {
    const int64_t w = w0 + lane;
    uint32_t bits = 0;
    if (w < p.g.nwords) {
      bits = nx[w];
      nx[w] = 0;
      fr[w] = 0;
    }
    while (__ballot_sync(0xffffffffu, bits != 0)) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (bits) {
          const int b = __ffs(bits) - 1;
          bits &= bits - 1;
          const int64_t v = w * 32 + b;
          route_vertex(sg, (int32_t)v, __ldg(p.g.rowptr + v + 1) - __ldg(p.g.rowptr + v), p, st.level);
        }
      }
      stage_flush(sg, cs, q, wq, false);
    }
  }
  stage_flush(sg, cs, q, wq, true);
}

// ---------------------------------------------------------------------------------------------
// stepwise engine: one kernel per phase
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) search_init_kernel(BfsParams p, int64_t src) { phase_init(p, src); }
__global__ void __launch_bounds__(kBlock) top_down_kernel(BfsParams p, LoopState st, bool claim_only) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  phase_top_down<false>(p, st, stage, blockIdx.x, gridDim.x, nullptr, claim_only, claim_only);
}
// the instance with the hub bin (its 16 KiB of staging buffers stay out of the plain kernels: shared
// memory taken from every CTA is L1 taken from the bitmap probes)
__global__ void __launch_bounds__(kBlock) top_down_hub_kernel(BfsParams p, LoopState st, bool claim_only) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  __shared__ HubStage hub;
  if (threadIdx.x == 0) hub_stage_init(&hub);
  __syncthreads();
  phase_top_down<false>(p, st, stage, blockIdx.x, gridDim.x, nullptr, claim_only, claim_only, &hub);
}
__global__ void __launch_bounds__(kBlock) bottom_up_kernel(BfsParams p, LoopState st) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  phase_bottom_up(p, st, stage);
}
__global__ void __launch_bounds__(kBlock) queue_to_bitmap_kernel(BfsParams p, LoopState st) {
  phase_queue_to_bitmap(p, st);
}
__global__ void __launch_bounds__(kBlock) bitmap_to_queue_kernel(BfsParams p, LoopState st) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  phase_bitmap_to_queue(p, st, stage);
}

// ---------------------------------------------------------------------------------------------
// multi-rank engine: the same phases, one launch each, with NCCL exchanges in between (multi.cu)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) top_down_multi_kernel(BfsParams p, LoopState st) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  __shared__ RemoteStage rstage[kWarpsPerBlock];
  phase_top_down<true>(p, st, stage, blockIdx.x, gridDim.x, rstage);
}
__global__ void __launch_bounds__(kBlock) apply_remote_kernel(BfsParams p, LoopState st, const int2* recv, int64_t count) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  const unsigned int counts[1] = {(unsigned int)count};  // NCCL delivered the pairs back to back: one bucket
  phase_apply_remote(p, st, stage, recv, 0, counts, 1);
}
// one 16-word record per rank per exchange: [0] narrow queue tail, [1] wide tail, [2] degree sum,
// [3] bottom-up finds, [4] source degree, [8 + d] pairs staged for rank d
__global__ void pack_stats_kernel(BfsParams p, int slot, unsigned long long* out) {
  const LevelSlot* s = &p.s.ctl->slot[slot & 3];
  if (threadIdx.x == 0) {
    out[0] = s->q_count;
    out[1] = s->wide_count;
    out[2] = s->edges;
    out[3] = s->bu_count;
    out[4] = (unsigned long long)p.s.ctl->src_degree;
    out[5] = out[6] = out[7] = 0;
  }
  if (threadIdx.x < 8) out[8 + threadIdx.x] = (int)threadIdx.x < p.g.world ? p.s.send_count[threadIdx.x * 32] : 0;
}

void launch_init(const BfsParams& p, int64_t src_local, int grid, cudaStream_t s) {
  search_init_kernel<<<grid, kBlock, 0, s>>>(p, src_local);
}
void launch_top_down_multi(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s) {
  top_down_multi_kernel<<<grid, kBlock, 0, s>>>(p, st);
}
void launch_apply_remote(const BfsParams& p, const LoopState& st, const int2* recv, int64_t count, int grid,
                         cudaStream_t s) {
  apply_remote_kernel<<<grid, kBlock, 0, s>>>(p, st, recv, count);
}
void launch_bottom_up(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s) {
  bottom_up_kernel<<<grid, kBlock, 0, s>>>(p, st);
}
void launch_queue_to_bitmap(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s) {
  queue_to_bitmap_kernel<<<grid, kBlock, 0, s>>>(p, st);
}
void launch_bitmap_to_queue(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s) {
  bitmap_to_queue_kernel<<<grid, kBlock, 0, s>>>(p, st);
}
void launch_pack_stats(const BfsParams& p, int slot, unsigned long long* out, cudaStream_t s) {
  pack_stats_kernel<<<1, 32, 0, s>>>(p, slot, out);
}

// ---------------------------------------------------------------------------------------------
// fused engine: the whole search in one persistent cooperative kernel.  One grid-wide barrier per
// level (two at a direction switch); every thread derives the same LoopState from the level
// counters, so no thread has to publish decisions.  Replaces the reference's per-level
// launch / cudaDeviceSynchronize / managed-counter read / cudaMemset round trips (bfs.cu:569-619).
// ---------------------------------------------------------------------------------------------

// Small frontiers (the first and last levels of a power-law search, EVERY level of a grid or road
// graph): a full-grid barrier per level costs more than the level.  The first kGroupBlocks CTAs
// then run consecutive top-down levels on their own, separated by a 16-party barrier on one global
// counter, while the other CTAs wait at the next grid-wide barrier.
#ifndef B2_FUSED_MIN_BLOCKS
#define B2_FUSED_MIN_BLOCKS 4  // resident CTAs per SM the fused kernel is register-capped for
#endif
constexpr int kGroupBlocks = 16;
constexpr unsigned kGroupMaxWork = kGroupBlocks * kWarpsPerBlock * 2;  // tiles + items per level

__device__ __forceinline__ unsigned frontier_work(const LoopState& st) { return ((st.cur_q + 31u) >> 5) + st.cur_wide; }

__device__ __forceinline__ void group_barrier(unsigned long long* counter, unsigned long long& epoch, int parties) {
  epoch += (unsigned long long)parties;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();  // publish this CTA's queue writes
    atomicAdd(counter, 1ull);
    while (ld_volatile(counter) < epoch) {}
    __threadfence();  // and drop stale L1 lines before anyone reads what the others wrote
  }
  __syncthreads();
}

// one record per call, from whichever thread calls (slots are handed out by an atomic counter that
// phase_init's caller zeroes before the first mark of a search)
__device__ __forceinline__ void trace_put(Control* c, int level, int phase) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const int i = atomicAdd(&c->trace_len, 1);
  if (i < kTraceCap) {
    c->trace[i].level = level;
    c->trace[i].phase = phase;
    c->trace[i].t_ns = t;
  }
}
__device__ __forceinline__ void trace_mark(Control* c, int& n, int level, int phase) {
  (void)n;
  if (blockIdx.x == 0 && threadIdx.x == 0) trace_put(c, level, phase);
}

// ---------------------------------------------------------------------------------------------
// Cluster mode - high-diameter graphs (grids, road networks: thousands of levels of a few thousand
// vertices each).  What such a level costs is its chain of dependent memory round trips, and on the
// global-memory path most links of that chain are bookkeeping: the queue read, the tail atomic, the
// barrier's arrival atomic and spin, the counter reads.  Here the first cluster of the grid keeps the
// frontier in its CTAs' SHARED memory (a discovery is appended to the queue of the CTA its id hashes
// to, through distributed shared memory), the level barrier is the hardware cluster barrier, and the
// counters every CTA needs for the level decision are read from the other CTAs' shared memory.  What
// is left of the chain: row extent -> neighbour ids -> the claim's atomicOr.
// Only frontiers of narrow rows qualify (no vertex of kNarrow or more entries).  On leaving, the
// frontier is handed back through the ordinary work lists, routed by degree.
// Replaces, per level, the reference's launch + cudaDeviceSynchronize + peer copies + managed-counter
// round trips (bfs.cu:569-619).
// ---------------------------------------------------------------------------------------------
constexpr int kClusterQCap = 2048;                     // per CTA and queue; three queues (ring indexed by level % 3)
constexpr unsigned kClusterMaxQPerBlock = kClusterQCap / 2;   // leave above this many frontier vertices per CTA of the cluster
constexpr unsigned kClusterEnterQ = 4096;
struct ClusterQ {
  int32_t q[3][kClusterQCap];
  unsigned int cnt[3], edges[3];
  unsigned int base;  // scratch for the dump
};
static_assert(sizeof(ClusterQ) <= sizeof(WarpStage) * kWarpsPerBlock, "the cluster queues overlay the staging buffers");

// The cluster variant of the kernel is chosen per GRAPH (api.cu): only graphs without a single row of
// kNarrow or more entries (grids, road-like meshes) get it, so a frontier here never holds a wide row.
__device__ __forceinline__ bool cluster_eligible(const LoopState& st) {
  return st.mode == 0 && st.cur_wide == 0 && st.cur_q <= kClusterEnterQ;
}

// run by the CTAs of cluster 0; on return st describes the level that follows; unless st.done the
// frontier of level st.level is back in the global work lists with its slot counters set
__device__ __forceinline__ Conversion cluster_levels(const BfsParams& p, LoopState& st, WarpStage* stage) {
  cg::cluster_group cluster = cg::this_cluster();
  ClusterQ* cq = reinterpret_cast<ClusterQ*>(stage);
  const unsigned c = cluster.block_rank(), nc = cluster.num_blocks();
  const unsigned t = threadIdx.x, lane = t & 31;
  Control* ctl = p.s.ctl;
  unsigned cur = 0;  // ring position of the current frontier's queue
  if (t == 0) {
    cq->cnt[0] = st.cur_q > c ? (st.cur_q - c + nc - 1) / nc : 0u;
    cq->cnt[1] = cq->cnt[2] = 0;
    cq->edges[0] = cq->edges[1] = cq->edges[2] = 0;
    if (c == 0) { ctl->dump_count = 0; ctl->cluster_overflow = 0; }
  }
  {  // the frontier, dealt round-robin to the CTAs
    const int32_t* gq = p.s.queue[st.level & 1];
    for (unsigned j = t; c + nc * j < st.cur_q; j += kBlock) cq->q[0][j] = __ldcg(gq + c + nc * j);
  }
  __threadfence();
  cluster.sync();
  Conversion cv = kNoConversion;
#ifdef B2_CLUSTER_PROFILE
  unsigned long long tp[4] = {0, 0, 0, 0}, t0p = 0;
#define B2_TP(i) do { unsigned long long _n; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_n)); tp[i] += _n - t0p; t0p = _n; } while (0)
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0p));
#else
#define B2_TP(i)
#endif
  for (;;) {
    const unsigned nxt = cur == 2 ? 0 : cur + 1, spare = nxt == 2 ? 0 : nxt + 1;
    const unsigned cnt = cq->cnt[cur];
    const int level = st.level;
    // recycle the ring: the count that collects the level after next, the degree sum of the next level
    // (both were last read before the previous barrier, by CTAs that have all passed it since)
    if (t == 0) { cq->cnt[spare] = 0; cq->edges[nxt] = 0; }
    unsigned my_edges = 0;
    for (unsigned j = t; j < cnt; j += kBlock) {
      const int32_t u = cq->q[cur][j];
      const int64_t rs = __ldg(p.g.rowptr + u), re = __ldg(p.g.rowptr + u + 1);
      my_edges += (unsigned)(re - rs);
      for (int64_t e0 = rs; e0 < re; e0 += 4) {  // four neighbours at a time: loads, claims and appends each in flight together
        int32_t v[4];
        uint32_t old[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = e0 + k < re ? __ldg(p.g.col + e0 + k) : -1;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          old[k] = v[k] >= 0 ? atomicOr(&p.s.visited[(uint32_t)v[k] >> 5], 1u << (v[k] & 31)) : 0xffffffffu;
        ClusterQ* o[4];
        unsigned pos[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          pos[k] = 0xffffffffu;
          if (v[k] >= 0 && !((old[k] >> (v[k] & 31)) & 1u)) {
            p.s.depth[v[k]] = level + 1;
            p.s.parent[v[k]] = u;
            o[k] = cluster.map_shared_rank(cq, ((uint32_t)v[k] ^ ((uint32_t)v[k] >> 9)) & (nc - 1));  // nc is 8 or 16
            pos[k] = atomicAdd(&o[k]->cnt[nxt], 1u);
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (pos[k] == 0xffffffffu) continue;
          if (pos[k] < (unsigned)kClusterQCap) {
            o[k]->q[nxt][pos[k]] = v[k];
          } else {  // that queue is full: the vertex goes to the global list, the mode ends after this level
            atomicSub(&o[k]->cnt[nxt], 1u);
            p.s.queue[level & 1][atomicAdd(&ctl->dump_count, 1u)] = v[k];
            atomicExch(&ctl->cluster_overflow, 1u);
          }
        }
      }
    }
    for (int o2 = 16; o2; o2 >>= 1) my_edges += __shfl_xor_sync(0xffffffffu, my_edges, o2);
    if (lane == 0 && my_edges) atomicAdd(&cq->edges[cur], my_edges);
    B2_TP(0);
    cluster.sync();  // every append has landed, every count and degree sum is final
    B2_TP(1);
    unsigned nq_t = 0, mf_cur = 0;
    const unsigned cur_t = st.cur_q;  // what the previous level (or the entry) counted
    {
      unsigned a[kClusterMaxBlocks], b[kClusterMaxBlocks];
#pragma unroll
      for (unsigned r = 0; r < (unsigned)kClusterMaxBlocks; ++r) {  // all remote loads in flight together
        const ClusterQ* oq = cluster.map_shared_rank(cq, r < nc ? r : 0);
        a[r] = *reinterpret_cast<const volatile unsigned int*>(&oq->cnt[nxt]);
        b[r] = *reinterpret_cast<const volatile unsigned int*>(&oq->edges[cur]);
      }
#pragma unroll
      for (unsigned r = 0; r < (unsigned)kClusterMaxBlocks; ++r) {
        if (r < nc) { nq_t += a[r]; mf_cur += b[r]; }
      }
    }
    // the degree sum of the NEW frontier is not known yet (its rows are read next level): the mean
    // degree of the level just expanded stands in for it - it only feeds the switch heuristic
    const unsigned long long mf_est = (unsigned long long)nq_t * (cur_t ? (mf_cur + cur_t - 1) / cur_t : 1u);
    cv = advance(st, p.tn, p.g.n, nq_t, 0, 0, mf_est);
    cur = nxt;
    B2_TP(2);
    const bool overflow = *reinterpret_cast<volatile unsigned int*>(&ctl->cluster_overflow) != 0;
    if (st.done || cv != kNoConversion || overflow || nq_t > kClusterMaxQPerBlock * nc) break;
  }
#ifdef B2_CLUSTER_PROFILE
  if (c == 0 && t == 0) printf("[cluster profile, CTA 0 thread 0] levels %d: expand %.1f us, barrier wait %.1f us, totals+advance %.1f us\n",
                               st.td_levels, tp[0] * 1e-3, tp[1] * 1e-3, tp[2] * 1e-3);
#endif
  if (st.done) return cv;
  // ---- leave: my part of the frontier -> the global list behind the spills, then all of it -> the
  // ordinary work lists of level st.level
  int32_t* tmp = p.s.queue[(st.level + 1) & 1];
  const unsigned cnt = cq->cnt[cur];
  if (t == 0) {
    cq->base = atomicAdd(&ctl->dump_count, cnt);
    if (c == 0) {
      for (int k = 0; k < 2; ++k) {
        LevelSlot* z = &ctl->slot[(st.level + k) & 3];
        z->q_count = 0; z->wide_count = 0; z->huge_count = 0; z->edges = 0; z->bu_count = 0; z->bu_cursor = 0;
      }
    }
  }
  __syncthreads();
  for (unsigned j = t; j < cnt; j += kBlock) tmp[cq->base + j] = cq->q[cur][j];
  __threadfence();
  cluster.sync();  // the list is complete; the queues in shared memory are no longer needed
  const unsigned total = *reinterpret_cast<volatile unsigned int*>(&ctl->dump_count);
  {
    LevelSlot* cs = &ctl->slot[st.level & 3];
      WarpStage& sg = stage[t >> 5];
    if (lane == 0) { sg.ncnt = 0; sg.wcnt = 0; }
    __syncwarp();
    const unsigned stride = nc * kBlock;
    for (unsigned i0 = 0; i0 < total; i0 += stride) {  // warp-uniform trip count
      const unsigned i = i0 + c * kBlock + t;
      if (i < total) {
        const int32_t v = __ldcg(tmp + i);
        route_vertex(sg, v, __ldg(p.g.rowptr + v + 1) - __ldg(p.g.rowptr + v), p, st.level);
      }
      stage_flush(sg, cs, p.s.queue[st.level & 1], p.s.wideq[st.level & 1], false);
    }
    stage_flush(sg, cs, p.s.queue[st.level & 1], p.s.wideq[st.level & 1], true);
  }
  return cv;
}

// work items of the current frontier that only a full grid should take on: hub-row items
__device__ __forceinline__ unsigned hub_items(const BfsParams& p, int level) {
  return p.s.hugeq[0] ? ld_volatile(&p.s.ctl->slot[level & 3].huge_count) : 0u;
}

template <bool kCluster, bool kHub>
__device__ __forceinline__ void fused_search(BfsParams& p, int64_t src) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  HubStage* hs = nullptr;
  if constexpr (kHub) {
    __shared__ HubStage hub;
    if (threadIdx.x == 0) hub_stage_init(&hub);
    hs = &hub;
  }
  __syncthreads();
  cg::grid_group grid = cg::this_grid();
  if (blockIdx.x == 0 && threadIdx.x == 0) p.s.ctl->trace_len = 0;
  phase_init(p, src);
  LoopState st = initial_state(p, __ldg(p.g.rowptr + src + 1) - __ldg(p.g.rowptr + src));
  const int group = min((int)gridDim.x, kGroupBlocks);
  unsigned long long epoch = 0;
  unsigned int group_seq = 0;
  int ntrace = 0;
  trace_mark(p.s.ctl, ntrace, 0, -1);
  grid.sync();
  trace_mark(p.s.ctl, ntrace, 0, 0);
  while (!st.done) {
    Conversion cv = kNoConversion;
    bool marked = false;
    const unsigned hubs = hub_items(p, st.level);
    if (kCluster && cluster_eligible(st)) {
      ++group_seq;
      if ((int)blockIdx.x < p.tn.cluster) {
        cv = cluster_levels(p, st, stage);
        if (blockIdx.x == 0 && threadIdx.x == 0) {
          p.s.ctl->resume_state = st;
          p.s.ctl->resume_conversion = (int)cv;
          __threadfence();
          *reinterpret_cast<volatile unsigned int*>(&p.s.ctl->group_exit_seq) = group_seq;
        }
      } else {
        if (threadIdx.x == 0)
          while (ld_volatile(&p.s.ctl->group_exit_seq) < group_seq) __nanosleep(2000);
        __syncthreads();
      }
      grid.sync();
      if ((int)blockIdx.x >= p.tn.cluster) {  // catch up with what the cluster decided
        const volatile LoopState* rs = &p.s.ctl->resume_state;
        st.level = rs->level; st.mode = rs->mode; st.fb = rs->fb; st.done = rs->done;
        st.prev_nf = rs->prev_nf; st.unexplored = rs->unexplored; st.td_levels = rs->td_levels;
        st.bu_levels = rs->bu_levels; st.frontier_edges = rs->frontier_edges;
        cv = (Conversion)(*reinterpret_cast<const volatile int*>(&p.s.ctl->resume_conversion));
      }
      if (!st.done) {  // the frontier came back through the work lists, routed by degree
        const LevelSlot* cs = &p.s.ctl->slot[st.level & 3];
        st.cur_q = ld_volatile(&cs->q_count);
        st.cur_wide = ld_volatile(&cs->wide_count);
      }
      trace_mark(p.s.ctl, ntrace, st.level, 7);
    } else if (st.mode == 0 && frontier_work(st) <= kGroupMaxWork && hubs == 0) {
      ++group_seq;
      if ((int)blockIdx.x < group) {
        do {
          phase_top_down<false>(p, st, stage, blockIdx.x, group);
          group_barrier(&p.s.ctl->group_barrier, epoch, group);
          const LevelSlot* ns = &p.s.ctl->slot[(st.level + 1) & 3];
          cv = advance(st, p.tn, p.g.n, ld_volatile(&ns->q_count), ld_volatile(&ns->wide_count),
                       ld_volatile(&ns->bu_count), ld_volatile(&ns->edges), false, ld_volatile(&ns->huge_count));
        } while (!st.done && cv == kNoConversion && frontier_work(st) <= kGroupMaxWork && hub_items(p, st.level) == 0);
        if (blockIdx.x == 0 && threadIdx.x == 0) {
          p.s.ctl->resume_state = st;
          p.s.ctl->resume_conversion = (int)cv;
          __threadfence();
          *reinterpret_cast<volatile unsigned int*>(&p.s.ctl->group_exit_seq) = group_seq;
        }
      } else {
        // the other CTAs sleep on a flag instead of hammering the grid barrier's L2 line for
        // what can be thousands of levels
        if (threadIdx.x == 0)
          while (ld_volatile(&p.s.ctl->group_exit_seq) < group_seq) __nanosleep(2000);
        __syncthreads();
      }
      grid.sync();
      if ((int)blockIdx.x >= group) {  // catch up with what the group decided
        const volatile LoopState* rs = &p.s.ctl->resume_state;
        st.level = rs->level; st.mode = rs->mode; st.fb = rs->fb; st.done = rs->done;
        st.cur_q = rs->cur_q; st.cur_wide = rs->cur_wide; st.prev_nf = rs->prev_nf;
        st.unexplored = rs->unexplored; st.td_levels = rs->td_levels; st.bu_levels = rs->bu_levels;
        st.frontier_edges = rs->frontier_edges;  // every field: all CTAs must take the same decisions from here on
        cv = (Conversion)(*reinterpret_cast<const volatile int*>(&p.s.ctl->resume_conversion));
      }
      trace_mark(p.s.ctl, ntrace, st.level, 5);
    } else {
      const int ph = st.mode == 0 ? 1 : 2;
      marked = st.mode == 0 && p.tn.mode == B2BFS_MODE_DIRECTION_OPT && st.bu_levels == 0 && st.level >= 1;
      const bool claim = claim_only_level(st, p.tn);  // implies marked
      if (st.mode == 0) phase_top_down<false>(p, st, stage, blockIdx.x, gridDim.x, nullptr, marked, claim, hs);
      else phase_bottom_up(p, st, stage);
      grid.sync();
      trace_mark(p.s.ctl, ntrace, st.level, claim ? 6 : ph);
      const LevelSlot* ns = &p.s.ctl->slot[(st.level + 1) & 3];
      cv = advance(st, p.tn, p.g.n, ld_volatile(&ns->q_count), ld_volatile(&ns->wide_count),
                   ld_volatile(&ns->bu_count), ld_volatile(&ns->edges), claim, ld_volatile(&ns->huge_count));
    }
    if (cv == kQueueToBitmap) {
      if (!marked) {  // a marking top-down level has already left the new frontier in bitmap[0]
        phase_queue_to_bitmap(p, st);
        grid.sync();
        trace_mark(p.s.ctl, ntrace, st.level, 3);
      }
    } else if (marked && !st.done) {
      phase_unmark(p, st);
      grid.sync();
      trace_mark(p.s.ctl, ntrace, st.level, 3);
    } else if (cv == kBitmapToQueue) {
      phase_bitmap_to_queue(p, st, stage);
      grid.sync();
      trace_mark(p.s.ctl, ntrace, st.level, 4);
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

// two instances of the search: the plain one, and one that also holds the cluster mode (chosen per
// graph by the host: only graphs whose every row is narrower than kNarrow run it, launched in
// clusters of Tuning::cluster CTAs).  Separate instances keep the extra code's register pressure
// away from the power-law path.
__global__ void __launch_bounds__(kBlock, B2_FUSED_MIN_BLOCKS) bfs_fused_kernel(BfsParams p, int64_t src) {
  fused_search<false, false>(p, src);
}
// ... and the instance with the CTA-level hub bin (option huge_min > 0).  Its two 8 KiB staging
// buffers cost every CTA shared memory, i.e. every SM 64 KiB of L1: measured against the plain
// instance the bin loses more through that than bulk-copy staging wins (profiles/r2), so it is its
// own kernel and off by default.
__global__ void __launch_bounds__(kBlock, B2_FUSED_MIN_BLOCKS) bfs_fused_hub_kernel(BfsParams p, int64_t src) {
  fused_search<false, true>(p, src);
}
// (two CTAs per SM are plenty for graphs of narrow rows: the registers go to the level loop instead)
__global__ void __launch_bounds__(kBlock, 2) bfs_fused_cluster_kernel(BfsParams p, int64_t src) {
  fused_search<true, false>(p, src);
}

// ---------------------------------------------------------------------------------------------
// peer-memory engine: the whole MULTI-RANK search in one persistent cooperative kernel per GPU.
// Each rank's kernel writes what its peers need straight into their memory over NVLink (CUDA IPC
// mappings of one exchange region per rank) — frontier-bitmap slices after a bottom-up level,
// (vertex, parent) pairs during a top-down level — and the per-level collective shrinks to a record
// exchange: every rank stores a 128-byte record (queue tails, degree sum, per-peer pair counts) into
// every peer's record table, the last word being the epoch number, and waits until the records of
// all peers for this epoch have landed in its own table.  Records carry the counters, so every rank
// takes the same decisions from the same numbers without a host in the loop.
// Requires one rank per physical GPU (the kernels of different ranks wait for one another).
// Replaces the per-level MPI_Sendrecv + MPI_Allreduce round trips of bfs_mpi.cu:612-621.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ volatile unsigned long long* xrecord(const BfsParams& p, int on_rank, unsigned long long epoch,
                                                                int from_rank) {
  return reinterpret_cast<volatile unsigned long long*>(p.s.xbase[on_rank] + xchg_off_records(p.g.world, p.g.part)) +
         ((epoch & 3) * 8 + from_rank) * 32;
}

// All ranks call this the same number of times.  On return sums[0..4] = totals over ranks of
// (narrow tail, wide tail, degree sum, bottom-up finds, source degree) of `slot`, sums[5..6] = this
// rank's own tails and incoming[s] = pairs rank s staged for this rank.
//
// One exchange = one arrival atomic per CTA, one burst of NVLink stores, one poll:
//   * every CTA bumps an arrival counter (nth = number of this exchange in the kernel, 1-based; the
//     host zeroes the counter before the launch);
//   * the CTA that arrives LAST knows that everything this GPU wrote for its peers during the phase
//     is on its way (the issuing warps fenced their own peer stores: remote_flush and
//     push_frontier_slice end with __threadfence_system) and stores this rank's record into every
//     rank's table.  Every 64-bit word of a record carries the epoch in its upper 16 bits, so a word
//     IS its own arrival flag: 13 independent stores per peer, no fence and no separate flag store
//     behind them (a store + system fence + flag chain costs two more NVLink crossings);
//   * every CTA then polls ITS OWN table, one thread per word, until the words of all ranks carry
//     this epoch - nobody waits for a second, local release step.
// `epoch` runs on across searches (Control::xepoch), so a slot of the 4-deep record ring is never
// rewritten while a slower rank may still be reading it: to post epoch e+4 a rank must have
// completed e+1..e+3, and completing e+1 needs every CTA of every rank to have arrived there, i.e.
// to have finished reading e.  A stale word of epoch e-4 cannot pass for e (tags differ by 4).
constexpr unsigned long long kRecValueMask = (1ull << 48) - 1ull;  // counters stay far below 2^48

__device__ __forceinline__ void cross_exchange(const BfsParams& p, unsigned long long epoch, unsigned long long nth,
                                               int slot, unsigned long long* sums, unsigned int* incoming, int* failed,
                                               int trace_level = -1) {
  __shared__ unsigned long long rec_s[8][16];
  __shared__ int last_s;
  const int k = p.g.world, me = p.g.rank;
  Control* ctl = p.s.ctl;
  const unsigned long long tag = (epoch & 0xffffull) << 48;
  const int r = threadIdx.x >> 4, w = threadIdx.x & 15;  // (rank, word) this thread posts / polls
  const bool used = r < k && (w < 5 || w >= 8);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();  // this CTA's counter atomics and local writes are ordered before its arrival
    const unsigned long long before = atomicAdd(&ctl->xarrive, 1ull);
    last_s = (before + 1 == nth * gridDim.x);
    if (last_s) __threadfence();
  }
  __syncthreads();
  if (last_s) {
    if (threadIdx.x == 0 && trace_level >= 0) trace_put(ctl, trace_level, 11);  // every CTA of this GPU has arrived
    if (used) {
      const LevelSlot* s = &ctl->slot[slot & 3];
      unsigned long long v;
      switch (w) {
        case 0: v = ld_volatile(&s->q_count); break;
        case 1: v = ld_volatile(&s->wide_count); break;
        case 2: v = ld_volatile(&s->edges); break;
        case 3: v = ld_volatile(&s->bu_count); break;
        case 4: v = (unsigned long long)ctl->src_degree; break;
        default: v = (w - 8) < k ? ld_volatile(&p.s.send_count[(w - 8) * 32]) : 0u; break;
      }
      xrecord(p, r, epoch, me)[w] = (v & kRecValueMask) | tag;
    }
  }
  if (used) {
    volatile unsigned long long* src = xrecord(p, me, epoch, r) + w;
    const unsigned long long t0 = global_ns();
    unsigned long long x;
    while (((x = *src) & ~kRecValueMask) != tag) {
      __nanosleep(40);
      if (global_ns() - t0 > 4000000000ull) {  // 4 s: a peer died or never launched
        ctl->error = 1;
        break;
      }
    }
    rec_s[r][w] = x & kRecValueMask;
    // acquire: what the peers wrote before their records is visible to this SM's later loads (a
    // gpu-scope fence also drops the L1 lines that predate it: the frontier buffers are re-read)
    if (w == 0) __threadfence();
  }
  __syncthreads();
  for (int i = 0; i < 5; ++i) sums[i] = 0;
  for (int q = 0; q < k; ++q) {
    for (int i = 0; i < 5; ++i) sums[i] += rec_s[q][i];
    incoming[q] = q == me ? 0u : (unsigned int)rec_s[q][8 + me];
  }
  sums[5] = rec_s[me][0];
  sums[6] = rec_s[me][1];
  *failed = *reinterpret_cast<volatile int*>(&ctl->error);
  __syncthreads();
}

// copy this rank's next-frontier slice (bitmap[0]) into buffer `which` of every rank's frontier bitmap.
// Only the 16-byte groups that DIFFER from what this rank stored there the last time are sent: every
// rank's copy of my slice equals my own copy (all copies receive the same stores), so my own copy
// tells what the peers hold.  Sparse levels (the first and the last bottom-up levels, a marking
// top-down level) then move a few kilobytes instead of part/8 bytes to each of the k ranks.
__device__ __forceinline__ void push_frontier_slice(const BfsParams& p, int which) {
  const int64_t words4 = p.g.part / 128;  // uint4 groups in a slice
  const uint4* mine = reinterpret_cast<const uint4*>(p.s.bitmap[0]);
  const int64_t tid = (int64_t)blockIdx.x * kBlock + threadIdx.x, nthreads = (int64_t)gridDim.x * kBlock;
  const size_t off = (size_t)which * xchg_frontier_bytes(p.g.world, p.g.part) + (size_t)p.g.rank * (size_t)p.g.part / 8;
  const uint4* held = reinterpret_cast<const uint4*>(p.s.xbase[p.g.rank] + off);
  const bool full = p.tn.xfullpush != 0;
  for (int64_t j = tid; j < words4; j += nthreads) {
    const uint4 x = mine[j];
    if (!full) {
      const uint4 o = __ldcg(held + j);
      if (x.x == o.x && x.y == o.y && x.z == o.z && x.w == o.w) continue;
    }
    for (int d = 0; d < p.g.world; ++d) reinterpret_cast<uint4*>(p.s.xbase[d] + off)[j] = x;
  }
  __threadfence_system();  // the slice is in every peer's memory before this thread reaches the exchange
}

__global__ void __launch_bounds__(kBlock, B2_FUSED_MIN_BLOCKS) bfs_fused_multi_kernel(BfsParams p, int64_t src) {
  __shared__ WarpStage stage[kWarpsPerBlock];
  __shared__ RemoteStage rstage[kWarpsPerBlock];
  cg::grid_group grid = cg::this_grid();
  const int k = p.g.world, me = p.g.rank;
  const int64_t src_local = (src / p.g.part == me) ? src - p.g.vbase : -1;
  unsigned long long epoch = ld_volatile(&p.s.ctl->xepoch);  // exchanges completed by earlier searches (same on all ranks)
  if (blockIdx.x == 0 && threadIdx.x == 0) p.s.ctl->trace_len = 0;
  {  // reset: local state as on one GPU, plus the replicated `seen` filter
    uint4* s4 = reinterpret_cast<uint4*>(p.s.seen);
    const int64_t n4 = (int64_t)k * p.g.part / 128;
    for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kBlock)
      s4[i] = make_uint4(0, 0, 0, 0);
    if (blockIdx.x == 0 && (int)threadIdx.x < p.g.world) p.s.send_count[threadIdx.x * 32] = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) p.s.ctl->error = 0;
  }
  phase_init(p, src_local);
  unsigned long long sums[7];
  unsigned int incoming[8];
  int failed = 0;
  int ntrace = 0;
  unsigned long long nth = 0;
  trace_mark(p.s.ctl, ntrace, 0, -1);
  grid.sync();  // no exchange needed yet: nobody reads what a peer has not written in THIS search
  trace_mark(p.s.ctl, ntrace, 0, 0);

  LoopState st;
  st.level = 0; st.mode = 0; st.fb = 1; st.done = 0;  // fb = 1: bottom-up writes bitmap[0], this rank's slice
  st.cur_q = ld_volatile(&p.s.ctl->slot[0].q_count);
  st.cur_wide = ld_volatile(&p.s.ctl->slot[0].wide_count);
  st.prev_nf = 1;
  st.unexplored = p.g.entries_global;  // the source's own degree is noise at this size
  st.frontier_edges = 0;
  st.td_levels = st.bu_levels = 0;
  int gcur = 0;  // frontier bitmap buffer bottom-up reads
  const int2* inbox = reinterpret_cast<const int2*>(p.s.xbase[me] + xchg_off_inbox(k, p.g.part));

  while (!st.done) {
    unsigned long long nf = 0, mf = 0;
    bool claim = false;
    bool pushed = false;  // this level's discoveries already sit in everyone's frontier buffer gcur ^ 1
    if (st.mode == 0) {
      // a top-down level that may well be the last one before bottom-up leaves the frontier bit of
      // every discovery behind (as on one GPU) and ships the slice together with exchange #2, so
      // the switch costs neither a queue->bitmap pass nor an exchange of its own
      const bool marked = p.tn.mode == B2BFS_MODE_DIRECTION_OPT && st.bu_levels == 0 && st.level >= 1;
      claim = claim_only_level(st, p.tn);  // implies marked; same decision on every rank (global counters)
      phase_top_down<true>(p, st, stage, blockIdx.x, gridDim.x, rstage, marked, claim);
      trace_mark(p.s.ctl, ntrace, st.level, 6);
      cross_exchange(p, ++epoch, ++nth, st.level + 1, sums, incoming, &failed, st.level);  // #1: pair counts
      trace_mark(p.s.ctl, ntrace, st.level, 7);
      phase_apply_remote(p, st, stage, inbox, p.s.send_cap, incoming, k, marked, claim);
      if (blockIdx.x == 0 && (int)threadIdx.x < p.g.world) p.s.send_count[threadIdx.x * 32] = 0;  // posted in #1
      if (marked) {
        grid.sync();  // bits set by any CTA (local finds and arrivals) are all in bitmap[0]
        push_frontier_slice(p, gcur ^ 1);
        pushed = true;
      }
      cross_exchange(p, ++epoch, ++nth, st.level + 1, sums, incoming, &failed, st.level);  // #2: tails after the arrivals
      trace_mark(p.s.ctl, ntrace, st.level, 8);
      nf = claim ? sums[3] : sums[0] + sums[1];
      mf = claim ? 0 : sums[2];
      st.cur_q = (unsigned)sums[5];
      st.cur_wide = (unsigned)sums[6];
      st.td_levels++;
    } else {
      phase_bottom_up(p, st, stage,
                      reinterpret_cast<const uint32_t*>(p.s.xbase[me] + (size_t)gcur * xchg_frontier_bytes(k, p.g.part)));
      grid.sync();  // the slice is complete (late finds are OR-ed in by any CTA)
      trace_mark(p.s.ctl, ntrace, st.level, 9);
      push_frontier_slice(p, gcur ^ 1);
      cross_exchange(p, ++epoch, ++nth, st.level + 1, sums, incoming, &failed, st.level);
      trace_mark(p.s.ctl, ntrace, st.level, 10);
      gcur ^= 1;
      nf = sums[3];
      st.bu_levels++;
    }
    st.unexplored = mf > st.unexplored ? 0ull : st.unexplored - mf;
    st.frontier_edges = mf;
    st.level++;
    bool switched = false;
    if (nf == 0 || failed) {
      st.done = 1;
    } else if (st.mode == 0) {
      if (claim || (p.tn.mode == B2BFS_MODE_DIRECTION_OPT && mf > st.unexplored / (unsigned)p.tn.alpha && nf >= st.prev_nf)) {
        st.mode = 1;
        switched = true;
        if (pushed) {
          gcur ^= 1;  // the new frontier of every rank arrived with exchange #2
        } else {
          phase_queue_to_bitmap(p, st);  // own new frontier -> bitmap[0] (all-zero while top-down runs)
          grid.sync();
          push_frontier_slice(p, gcur ^ 1);
          cross_exchange(p, ++epoch, ++nth, st.level, sums, incoming, &failed);
          trace_mark(p.s.ctl, ntrace, st.level, 3);
          gcur ^= 1;
          if (failed) st.done = 1;
        }
      }
    } else if (nf < (unsigned long long)p.g.n_global / (unsigned)p.tn.beta && nf < st.prev_nf) {
      st.mode = 0;
      phase_bitmap_to_queue(p, st, stage);
      grid.sync();
      trace_mark(p.s.ctl, ntrace, st.level, 4);
      const LevelSlot* cs = &p.s.ctl->slot[st.level & 3];
      st.cur_q = ld_volatile(&cs->q_count);
      st.cur_wide = ld_volatile(&cs->wide_count);
    }
    if (pushed && !switched && !st.done) {  // stayed top-down after all: take the frontier bits back
      phase_unmark(p, st);
      grid.sync();
    }
    st.prev_nf = nf;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    p.s.ctl->result.levels = st.level;
    p.s.ctl->result.td_levels = st.td_levels;
    p.s.ctl->result.bu_levels = st.bu_levels;
    p.s.ctl->xepoch = epoch;
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
  if (p.tn.cluster) {  // cooperative AND in clusters (grid is a multiple of the cluster size)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kBlock);
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = (unsigned)p.tn.cluster;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    B2_CUDA(cudaLaunchKernelExC(&cfg, (const void*)bfs_fused_cluster_kernel, args));
  } else {
    const void* kernel = p.tn.huge_min > 0 ? (const void*)bfs_fused_hub_kernel : (const void*)bfs_fused_kernel;
    B2_CUDA(cudaLaunchCooperativeKernel(kernel, dim3(grid), dim3(kBlock), args, 0, stream));
  }
  if (launches) *launches += 1;
  return Status();
}

// Cluster size and grid for the cluster-mode kernel: 16 CTAs per cluster where the device admits it
// (a non-portable size, opt-in), else 8; the grid is the largest co-resident multiple.  *size_out = 0
// if the device (or this driver) cannot run the kernel in clusters on at least a quarter of the plain grid.
Status search_cluster_grid_size(int plain_grid, int* grid_out, int* size_out) {
  *grid_out = 0;
  *size_out = 0;
  for (int size = kClusterMaxBlocks; size >= 8 && *size_out == 0; size /= 2) {
    if (size > 8 &&
        cudaFuncSetAttribute((const void*)bfs_fused_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
      cudaGetLastError();
      continue;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(plain_grid / size * size);
    cfg.blockDim = dim3(kBlock);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)size;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, (const void*)bfs_fused_cluster_kernel, &cfg) != cudaSuccess) {
      cudaGetLastError();
      continue;
    }
    int g = clusters * size;
    if (g > plain_grid / size * size) g = plain_grid / size * size;
    // graphs that run this kernel have little parallel work per level: a quarter of the plain grid is plenty
    if (g * 4 >= plain_grid) {
      *grid_out = g;
      *size_out = size;
    }
  }
  return Status();
}

Status run_fused_multi(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, int64_t* launches) {
  BfsParams pp = p;
  int64_t s = src;
  void* args[] = {&pp, &s};
  // the arrival counter of the in-kernel exchanges starts from zero in every search
  B2_CUDA(cudaMemsetAsync(&p.s.ctl->xarrive, 0, sizeof(unsigned long long), stream));
  B2_CUDA(cudaLaunchCooperativeKernel((const void*)bfs_fused_multi_kernel, dim3(grid), dim3(kBlock), args, 0, stream));
  if (launches) *launches += 1;
  return Status();
}

Status multi_grid_size(int device, int* grid_out) {
  int per_sm = 0, sms = 0;
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bfs_fused_multi_kernel, kBlock, 0));
  B2_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  if (per_sm < 1) return make_error(B2BFS_ERR_CUDA, "bfs_fused_multi_kernel does not fit on an SM");
  *grid_out = per_sm * sms;
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
    const bool claim = claim_only_level(st, p.tn);
    if (st.mode == 0 && p.tn.huge_min > 0) top_down_hub_kernel<<<grid, kBlock, 0, stream>>>(p, st, claim);
    else if (st.mode == 0) top_down_kernel<<<grid, kBlock, 0, stream>>>(p, st, claim);
    else bottom_up_kernel<<<grid, kBlock, 0, stream>>>(p, st);
    ++nl;
    B2_CUDA(cudaMemcpyAsync(pinned, p.s.ctl, sizeof(Control), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    const LevelSlot& ns = pinned->slot[(st.level + 1) & 3];
    if (trace)
      fprintf(stderr, "[b2bfs]   found q=%u wide=%u bu=%llu edges=%llu\n", ns.q_count, ns.wide_count, ns.bu_count,
              ns.edges);
    const Conversion cv = advance(st, p.tn, p.g.n, ns.q_count, ns.wide_count, ns.bu_count, ns.edges, claim, ns.huge_count);
    if (cv == kQueueToBitmap) {
      if (!claim) {  // a claim-only level has left the new frontier in bitmap[0] itself
        queue_to_bitmap_kernel<<<grid, kBlock, 0, stream>>>(p, st);
        ++nl;
      }
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
