// multi.cu — the level loop across ranks (one process per GPU), NCCL over NVLink.
//
// Replaces the 2-rank CUDA-aware-MPI loop of bfs_mpi.cu:581-634 (kernel, cudaDeviceSynchronize,
// MPI_Sendrecv of device buffers with MPI_Get_count, MPI_Allreduce of the next frontier size) and the
// peer-copy loop of bfs.cu:595-619.  Differences that matter:
//   * the CSR is PARTITIONED by contiguous vertex ranges (each rank holds only its rows), not
//     replicated per device as at bfs.cu:333-351 / bfs_mpi.cu:832;
//   * a vertex discovered for another rank is sent to its owner ONCE per search (`seen` filter) with
//     its parent, and the owner claims it on arrival — the reference ships bare ids, never marks them
//     on receipt and repairs the duplicates with a min-merge of full distance arrays (bfs.cu:430-438);
//   * dense levels run bottom-up on each rank's own rows against the allgathered frontier bitmap
//     (n/8 bytes over NVLink per level) instead of pushing ids at all;
//   * any number of ranks, not `(i+1)%2` (bfs_mpi.cu:576, 615).
// Per level the host reads ONE small allgathered record per rank (queue tails, degree sums, per-peer
// send counts), so every rank takes the same decisions from the same numbers.
#include <cstdlib>

#include "multi.cuh"

namespace b2 {

namespace {

struct Totals {
  unsigned long long nq = 0, nwide = 0, edges = 0, bu = 0, src_degree = 0;
};

// pack this rank's counters, allgather the records, bring them to the host
Status exchange_stats(const BfsParams& p, const MultiContext& mc, int slot, cudaStream_t stream, MultiResult* out) {
  launch_pack_stats(p, slot, mc.stats_send, stream);
  B2_NCCL(mc, mc.nccl->AllGather(mc.stats_send, mc.stats_all, 16, ncclUint64, mc.comm, stream));
  B2_CUDA(cudaMemcpyAsync(mc.stats_host, mc.stats_all, (size_t)mc.world * 16 * sizeof(unsigned long long),
                          cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  out->launches += 1;
  out->nccl_calls += 1;
  return Status();
}

Totals sum_stats(const MultiContext& mc) {
  Totals t;
  for (int r = 0; r < mc.world; ++r) {
    const unsigned long long* s = mc.stats_host + (size_t)r * 16;
    t.nq += s[0]; t.nwide += s[1]; t.edges += s[2]; t.bu += s[3]; t.src_degree += s[4];
  }
  return t;
}

}  // namespace

Status run_multi(const BfsParams& p, const MultiContext& mc, int64_t src, int grid, cudaStream_t stream,
                 MultiResult* out) {
  static const bool trace = getenv("B2BFS_TRACE") != nullptr;
  const int me = mc.rank, k = mc.world;
  const int64_t part = p.g.part;
  const int64_t slice_words = part / 32;
  const int owner = (int)(src / part);
  const int64_t src_local = owner == me ? src - p.g.vbase : -1;

  B2_CUDA(cudaMemsetAsync(p.s.seen, 0, (size_t)k * slice_words * sizeof(uint32_t), stream));
  launch_init(p, src_local, grid, stream);
  out->launches += 1;
  B2_TRY(exchange_stats(p, mc, 0, stream, out));
  const Totals t0 = sum_stats(mc);

  LoopState st;
  st.level = 0; st.mode = 0; st.fb = 1; st.done = 0;  // fb = 1: bottom-up always writes bitmap[0], this rank's slice
  st.cur_q = (unsigned)mc.stats_host[(size_t)me * 16 + 0];
  st.cur_wide = (unsigned)mc.stats_host[(size_t)me * 16 + 1];
  st.prev_nf = 1;
  st.unexplored = (unsigned long long)mc.entries_global - t0.src_degree;
  st.td_levels = st.bu_levels = 0;

  while (!st.done) {
    unsigned long long nf = 0, mf = 0;
    if (st.mode == 0) {
      // ---- top-down: expand own frontier, ship foreign discoveries to their owners, claim arrivals
      B2_CUDA(cudaMemsetAsync(p.s.send_count, 0, (size_t)k * 32 * sizeof(unsigned int), stream));
      launch_top_down_multi(p, st, grid, stream);
      out->launches += 1;
      B2_TRY(exchange_stats(p, mc, st.level + 1, stream, out));
      int64_t recv_total = 0;
      B2_NCCL(mc, mc.nccl->GroupStart());
      for (int d = 0; d < k; ++d) {
        if (d == me) continue;
        const size_t to_d = (size_t)mc.stats_host[(size_t)me * 16 + 8 + d];
        const size_t from_d = (size_t)mc.stats_host[(size_t)d * 16 + 8 + me];
        if (to_d) B2_NCCL(mc, mc.nccl->Send(p.s.send_buf + (int64_t)d * p.s.send_cap, to_d * 2, ncclInt32, d, mc.comm, stream));
        if (from_d) B2_NCCL(mc, mc.nccl->Recv(mc.recv_buf + recv_total, from_d * 2, ncclInt32, d, mc.comm, stream));
        recv_total += (int64_t)from_d;
        out->exchanged_bytes += (int64_t)to_d * 8;
      }
      B2_NCCL(mc, mc.nccl->GroupEnd());
      out->nccl_calls += 1;
      if (recv_total) {
        launch_apply_remote(p, st, mc.recv_buf, recv_total, grid, stream);
        out->launches += 1;
      }
      B2_TRY(exchange_stats(p, mc, st.level + 1, stream, out));
      const Totals t = sum_stats(mc);
      nf = t.nq + t.nwide;
      mf = t.edges;
      st.cur_q = (unsigned)mc.stats_host[(size_t)me * 16 + 0];
      st.cur_wide = (unsigned)mc.stats_host[(size_t)me * 16 + 1];
      st.td_levels++;
    } else {
      // ---- bottom-up: own unvisited rows against the global frontier; share the new slice
      launch_bottom_up(p, st, grid, stream);
      out->launches += 1;
      B2_NCCL(mc, mc.nccl->AllGather(p.s.bitmap[0], p.s.gfrontier, (size_t)slice_words, ncclUint32, mc.comm, stream));
      out->nccl_calls += 1;
      out->exchanged_bytes += slice_words * 4;
      B2_TRY(exchange_stats(p, mc, st.level + 1, stream, out));
      nf = sum_stats(mc).bu;
      st.bu_levels++;
    }
    if (trace && me == 0)
      fprintf(stderr, "[b2bfs x%d] level %d %s found %llu (degree sum %llu) unexplored %llu\n", k, st.level,
              st.mode ? "bottom-up" : "top-down", nf, mf, st.unexplored);
    st.unexplored = mf > st.unexplored ? 0ull : st.unexplored - mf;
    st.level++;
    if (nf == 0) {
      st.done = 1;
    } else if (st.mode == 0) {
      if (p.tn.mode == B2BFS_MODE_DIRECTION_OPT && mf > st.unexplored / (unsigned)p.tn.alpha && nf >= st.prev_nf) {
        // top-down -> bottom-up: own new frontier as a bitmap slice, then everyone's
        st.mode = 1;
        B2_CUDA(cudaMemsetAsync(p.s.bitmap[0], 0, (size_t)slice_words * sizeof(uint32_t), stream));
        launch_queue_to_bitmap(p, st, grid, stream);
        B2_NCCL(mc, mc.nccl->AllGather(p.s.bitmap[0], p.s.gfrontier, (size_t)slice_words, ncclUint32, mc.comm, stream));
        out->launches += 1;
        out->nccl_calls += 1;
        out->exchanged_bytes += slice_words * 4;
      }
    } else if (nf < (unsigned long long)mc.n_global / (unsigned)p.tn.beta && nf < st.prev_nf) {
      // bottom-up -> top-down: purely local (each rank queues the vertices of its own slice)
      st.mode = 0;
      launch_bitmap_to_queue(p, st, grid, stream);
      out->launches += 1;
      B2_TRY(exchange_stats(p, mc, st.level, stream, out));
      st.cur_q = (unsigned)mc.stats_host[(size_t)me * 16 + 0];
      st.cur_wide = (unsigned)mc.stats_host[(size_t)me * 16 + 1];
    }
    st.prev_nf = nf;
  }
  B2_CUDA(cudaGetLastError());
  out->levels = st.level;
  out->td_levels = st.td_levels;
  out->bu_levels = st.bu_levels;
  return Status();
}

}  // namespace b2
