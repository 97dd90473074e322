// common.cuh — shared declarations of libb2bfs.so (host + device).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <string>

#include "../../include/b2bfs.h"

namespace b2 {

constexpr int kBlock = 256;              // threads per CTA for every kernel in this library
constexpr int kWarpsPerBlock = kBlock / 32;
constexpr int kSMs = 148;                // B200: 2 dies x 74 SMs; grids are sized in multiples of it
constexpr int kNarrow = 32;              // rows narrower than this are gathered 32 at a time by one warp
constexpr int kChunk = 128;              // wider rows are cut into work items of this many entries
constexpr int kClusterMaxBlocks = 16;     // cluster size of the cluster-mode kernel: 16 where the device admits it, else 8 (portable)
constexpr int kCtaChunk = 2048;          // rows of at least Tuning::huge_min entries: items of this many entries, one CTA
                                         // each, staged into shared memory by a bulk asynchronous copy (TMA, 1-D)

// ---------------------------------------------------------------------------------------------
// error plumbing: every CUDA call is checked (the reference discards all return codes, bfs.cu:336)
// ---------------------------------------------------------------------------------------------
struct Status {
  int code = B2BFS_OK;
  std::string msg;
  bool ok() const { return code == B2BFS_OK; }
};

inline Status make_error(int code, const std::string& msg) {
  Status s;
  s.code = code;
  s.msg = msg;
  return s;
}

#define B2_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      return b2::make_error(_e == cudaErrorMemoryAllocation ? B2BFS_ERR_NOMEM : B2BFS_ERR_CUDA,    \
                            std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                                ":" + std::to_string(__LINE__) + ")");                             \
    }                                                                                              \
  } while (0)

#define B2_TRY(expr)           \
  do {                         \
    b2::Status _s = (expr);    \
    if (!_s.ok()) return _s;   \
  } while (0)

// ---------------------------------------------------------------------------------------------
// resident graph (this rank's rows) and per-search state, as the kernels see them
// ---------------------------------------------------------------------------------------------
struct GraphView {
  int64_t n;                 // rows resident on this GPU (= all vertices on a single GPU)
  int64_t nwords;            // bitmap words covering n (multiple of 4)
  int64_t entries;           // resident CSR entries
  int64_t vbase;             // global id of local row 0 (0 on a single GPU); col[] holds GLOBAL ids
  int64_t part;              // rows per rank of the 1-D partition (multiple of 128); owner(v) = v / part
  int rank, world;
  int64_t n_global;          // all vertices
  unsigned long long entries_global;  // resident entries over all ranks
  const int64_t* rowptr;     // n+1 offsets into col
  const int32_t* col;        // neighbour ids, sorted ascending inside a row
  const uint32_t* visited_init;  // bit set = never scan this vertex bottom-up (degree 0, or id >= n)
  const int32_t* first_nbr;      // col[rowptr[v]] (or -1): lets bottom-up probe one neighbour without touching rowptr/col
};

// Per-level counters.  Every field that many warps hit with atomics sits on its own 128-byte line
// (B300_MICROARCH.md "Atomics": distinct lines are ~60x faster than one).
struct alignas(128) LevelSlot {
  alignas(128) unsigned int q_count;          // tail of the regular next queue
  alignas(128) unsigned int wide_count;       // tail of the wide-item next queue
  alignas(128) unsigned int huge_count;       // tail of the CTA-item next queue (hub rows)
  alignas(128) unsigned long long edges;      // sum of degrees of the vertices found this level
  alignas(128) unsigned long long bu_count;   // vertices found by a bottom-up level
  alignas(128) unsigned long long bu_cursor;  // next unclaimed batch of bitmap words (dynamic balance)
};

struct SearchResult {
  int32_t levels, td_levels, bu_levels, pad;
};

// Level-loop state; a pure function of the counters, so the host (stepwise engine) and every
// thread of the persistent kernel (fused engine) compute the same decisions.
struct LoopState {
  int level;
  int mode;                  // 0 top-down, 1 bottom-up
  int fb;                    // index of the frontier bitmap while bottom-up
  int done;
  unsigned int cur_q, cur_wide;
  unsigned long long prev_nf;
  unsigned long long unexplored;  // directed entries whose source is still unvisited
  unsigned long long frontier_edges;  // degree sum of the current frontier (what this level will scan top-down)
  int td_levels, bu_levels;
};

static_assert(sizeof(LoopState) == 56, "new LoopState field: also copy it where the waiting CTAs catch up (bfs_fused_kernel)");

constexpr int kTraceCap = 96;
struct TraceRec {
  int level, phase;            // phase: 0 reset, 1 top-down, 2 bottom-up, 3 queue->bitmap, 4 bitmap->queue, 5 group levels
  unsigned long long t_ns;     // %globaltimer when CTA 0 left the phase's barrier
};

struct Control {
  LevelSlot slot[4];         // ring indexed by level & 3
  SearchResult result;
  int64_t src_degree;        // written by the reset phase (the stepwise host loop starts from it)
  unsigned long long reached_vertices, reached_entries;  // filled by reach_stats_kernel
  alignas(128) unsigned long long group_barrier;         // monotonic arrival counter of the small-frontier group
  alignas(128) LoopState resume_state;                   // what the group decided, for the CTAs that waited
  int resume_conversion;
  int error;                                             // peer-memory engine: a cross-GPU wait timed out
  alignas(128) unsigned long long xarrive;               // peer-memory engine: CTAs arrived at exchanges so far (this launch)
  alignas(128) unsigned long long xepoch;                // ... exchanges completed by all searches on this graph so far
  unsigned int group_exit_seq;                           // bumped when the group leaves its level loop
  alignas(128) unsigned int dump_count;                  // cluster mode: vertices spilled / dumped to the global list so far
  unsigned int cluster_overflow;                         // cluster mode: some CTA's shared-memory queue ran full
  int trace_len;                                         // fused engine: per-phase timestamps of CTA 0
  TraceRec trace[kTraceCap];
};

struct SearchView {
  int32_t* depth;            // n
  int32_t* parent;           // n (vertex ids)
  uint32_t* visited;         // nwords
  uint32_t* bitmap[2];       // frontier / next frontier, nwords each; all-zero while top-down runs
  int32_t* queue[2];         // narrow rows (degree < kNarrow): vertex ids, n each
  int2* wideq[2];            // wide rows: (vertex, chunk) work items of <= kChunk entries
  int2* hugeq[2];            // hub rows: (vertex, chunk) work items of <= kCtaChunk entries (nullptr: bin not in use)
  Control* ctl;
  // multi-rank only (nullptr / 0 on a single GPU)
  uint32_t* gfrontier;       // world * part bits: the frontier of ALL ranks (allgathered), probed bottom-up
  uint32_t* seen;            // world * part bits: remote vertices this rank has already sent to their owner
  int2* send_buf;            // world buckets of send_cap (vertex, parent) pairs, indexed by owner
  unsigned int* send_count;  // world counters, 32 uints (128 bytes) apart
  int64_t send_cap;
  // peer-memory engine: base of every rank's exchange region (own one included), mapped with CUDA
  // IPC; all nullptr when exchanges go through NCCL.  Layout of a region (same on every rank):
  //   [2][world * part/8 B]  frontier bitmaps, ping-pong
  //   [world][part] int2     inbox: bucket s holds the (vertex, parent) pairs rank s found for me
  //   [4][8][32] u64         per-exchange records of every rank: words 0-4 counters, 8-15 pair counts
  //                          per peer, word 31 = epoch (the arrival flag, written last)
  unsigned char* xbase[8];
};

__host__ __device__ inline size_t xchg_frontier_bytes(int world, int64_t part) { return (size_t)world * (size_t)part / 8; }
__host__ __device__ inline size_t xchg_off_inbox(int world, int64_t part) { return 2 * xchg_frontier_bytes(world, part); }
__host__ __device__ inline size_t xchg_off_records(int world, int64_t part) {
  return xchg_off_inbox(world, part) + (size_t)world * (size_t)part * sizeof(int2);
}
__host__ __device__ inline size_t xchg_region_bytes(int world, int64_t part) {
  return xchg_off_records(world, part) + 4 * 8 * 32 * sizeof(unsigned long long);
}

struct Tuning {
  int mode;                  // B2BFS_MODE_*
  int alpha;                 // top-down -> bottom-up when m_f > m_u / alpha
  int beta;                  // bottom-up -> top-down when n_f < n / beta
  int gamma;                 // claim-only top-down level + forced switch when m_f * gamma > m_u / alpha (0: never)
  int cluster;               // CTAs per cluster the search kernel was launched with, 0 = no clusters (small frontiers then
                             // run inside cluster 0: queues in distributed shared memory, hardware cluster barrier)
  int huge_min;              // rows of at least this many entries go to the CTA-level bin (0: bin off)
  int xfullpush;              // peer-memory engine: 1 = push whole frontier slices, 0 = only the 16-byte groups that changed
};

enum Conversion { kNoConversion = 0, kQueueToBitmap = 1, kBitmapToQueue = 2 };

// A top-down level (before the first bottom-up level) whose frontier is already so heavy that the NEXT
// level is all but certain to run bottom-up does not build the next level's work lists at all: it
// only claims (visited bit, depth, parent) and leaves the frontier bit of every discovery in
// bitmap[0]; the switch to bottom-up is then forced.  Skipped per discovery: the row-extent read, the
// staging and the queue writes.  A wrong guess costs one bottom-up level the heuristic would have
// run top-down, never a wrong depth.
__host__ __device__ inline bool claim_only_level(const LoopState& st, const Tuning& tn) {
  return st.mode == 0 && tn.mode == B2BFS_MODE_DIRECTION_OPT && tn.gamma > 0 && st.level >= 1 && st.bu_levels == 0 &&
         st.frontier_edges * (unsigned long long)tn.gamma > st.unexplored / (unsigned)tn.alpha;
}

// `claimed`: the level that just ended was a claim-only one (nbu = its claims, nq/nwide/mf unused).
__host__ __device__ inline Conversion advance(LoopState& st, const Tuning& tn, int64_t n, unsigned int nq,
                                              unsigned int nwide, unsigned long long nbu, unsigned long long mf,
                                              bool claimed = false, unsigned int nhuge = 0) {
  const unsigned long long nf = (st.mode == 0 && !claimed) ? (unsigned long long)nq + nwide + nhuge : nbu;
  if (st.mode == 0) st.td_levels++; else st.bu_levels++;
  if (claimed) mf = 0;  // degrees of the discoveries were not read
  st.unexplored = (mf > st.unexplored) ? 0ull : st.unexplored - mf;
  st.frontier_edges = mf;
  st.level++;
  Conversion cv = kNoConversion;
  if (nf == 0) {
    st.done = 1;
  } else if (st.mode == 0) {
    st.cur_q = nq;
    st.cur_wide = nwide;
    if (claimed || (tn.mode == B2BFS_MODE_DIRECTION_OPT && mf > st.unexplored / (unsigned)tn.alpha && nf >= st.prev_nf)) {
      st.mode = 1;
      st.fb = 0;
      cv = kQueueToBitmap;  // a claim-only (or marking) level has already left the bits in bitmap[0]
    }
  } else {
    if (nf < (unsigned long long)n / (unsigned)tn.beta && nf < st.prev_nf) {
      st.mode = 0;
      cv = kBitmapToQueue;  // cur_q / cur_wide are read back after the conversion
    } else {
      st.fb ^= 1;
    }
  }
  st.prev_nf = nf;
  return cv;
}

}  // namespace b2
