#!/usr/bin/env python
"""bench.py — GTEPS (Graph500 harmonic mean over the timed roots) of the BFS hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scale S] ...

A "step" is one search from one root over the resident synthetic graph.  Default workloads:
  --gpus 1   RMAT scale-24 edgefactor-16, the graph BASELINE.json's north_star quotes the >= 50 % HBM
             roofline target on (2 GiB of column indices; BASELINE configs[1], RMAT-22, is measured in the
             same run and reported under the key "rmat22")
  --gpus N   Kronecker scale 24 + log2 N (weak scaling, 2^24 vertices per GPU; N = 4 is BASELINE
             configs[2], Kronecker-26 on 4 GPUs); round 1's weak series (scale 22 + log2 N) is kept
             under the key "weak22"
Column indices alone are far larger than the 126 MB L2, so consecutive searches cannot live in
cache.  One JSON line on stdout.

  value      harmonic-mean GTEPS, device-timed (CUDA events on the library's stream, inside
             b2bfs_run_device), graph resident in HBM, results left in HBM
  e2e        the same metric through b2bfs_run with HOST buffers: root id in, n x int32 depth array
             copied device->host into page-locked memory inside the timed region (host wall clock)
  roofline   the dominant kernel (the fused search kernel = the whole step): algorithmic bytes
             B_alg(r) = 4*D_r + 8*V_r + 8*n / its CUDA-event time, against the
             measured HBM copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline  the reference's own bfsCPU (oracle/_ref, else the oracle port) on a bounded sample
             of the same roots, on this box's host cores; it is also the parity check of the run
  parity     what was compared with the CPU side inside this very run.  One GPU: depth arrays of the
             sampled roots against bfsCPU, element for element.  Several GPUs: rank 0 regenerates the
             tuples on the host and checks the gathered depth arrays of a few timed roots with the
             three edge rules that pin BFS depths exactly (oracle_check_depth_by_edges), and - when
             the graph fits the reference's int offsets and the time budget - against bfsCPU too.
             Any mismatch ends the run with a non-zero exit code.
  strict_graph500_rule  the same measurement over roots drawn by the bare Graph500 rule (degree >= 1,
             no component filter), next to the headline whose roots lie in the largest component

`--impl reference` times that CPU implementation alone (the reference has no other runnable
implementation of the path on this box: its GPU kernels need its own main, its MPI variant needs
MPI).  The oracle is only ever the checker / baseline here, never the measured product.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

KINDS = {"rmat": 0, "kron": 1, "grid": 2}
METRIC = "GTEPS (Graph500 harmonic mean over timed roots)"


def hmean(x):
    x = np.asarray(x, dtype=np.float64)
    return float(len(x) / np.sum(1.0 / x)) if len(x) and np.all(x > 0) else 0.0


def pick_roots(deg, count, seed=2, member=None):
    """Seeded sample without replacement among vertices with at least one edge that is not a
    self-loop (Graph500 rule) AND inside the largest connected component.
    The second condition is ours: RMAT/Kronecker graphs carry a sprinkle of 2-3-vertex components,
    a root there measures the O(n) state reset and launch latency (tens of microseconds for one
    edge), and a single such root drags a 64-root harmonic mean down by five orders of magnitude
    for every implementation alike.  `member` is the depth != INT_MAX mask of one search from the
    highest-degree vertex."""
    ok = deg > 0
    if member is not None:
        ok &= member
    cand = np.flatnonzero(ok)
    rng = np.random.default_rng(seed)
    if len(cand) <= count:
        return [int(c) for c in rng.permutation(cand)] * (count // max(len(cand), 1) + 1)
    return [int(r) for r in rng.choice(cand, count, replace=False)]


def nonloop_degree(n, src, dst):
    """Per-vertex count of incident tuples that are not self-loops (host side of the root rule)."""
    keep = src != dst
    return np.bincount(src[keep], minlength=n) + np.bincount(dst[keep], minlength=n)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device=0):
        self.device, self.rows, self._stop, self._t = device, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                return
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return None
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------------
# CPU side: the reference's bfsCPU (oracle/_ref) or the oracle port — baseline + parity checker
# ------------------------------------------------------------------------------------------------
def mem_available_bytes():
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    return int(ln.split()[1]) * 1024
    except Exception:
        pass
    return 1 << 62


def cpu_graph_bytes(n, m):
    """Host bytes of tuples + CSR + one search's scratch, as CpuReference holds them at its peak."""
    return 8 * m + 8 * m + 12 * n + 4 * (2 * m if 2 * m < 2**31 else 0) + 16 * n


class CpuReference:
    def __init__(self, n, src, dst, scale):
        import oracle
        self.oracle = oracle
        t0 = time.perf_counter()
        # rows as readGraphFromFile builds them (bfs.cu:851-875).  From scale 23 up the all-cores
        # builder is used: same rows as multisets, order inside a row unspecified - depths (the only
        # thing compared or timed here) do not depend on it, and the serial builder needs minutes.
        self.builder = "file-order rows" if scale < 23 else "all-cores builder (row order unspecified)"
        self.g = oracle.build_csr(n, src, dst) if scale < 23 else oracle.build_csr_parallel(n, src, dst)
        self.kind = "reference" if oracle.ref_available() and self.g.num_edges < 2**31 else "port"
        if self.kind == "reference":
            oracle.ref_set_graph(self.g)
        self.build_s = time.perf_counter() - t0

    def search(self, root):
        """(depth, seconds) of one serial search, timed around bfsCPU alone as runCpu does."""
        if self.kind == "reference":
            depth, ns = self.oracle.ref_bfs_cpu(root, self.g.n)
            return depth, ns * 1e-9
        t0 = time.perf_counter()
        depth = self.oracle.bfs_serial(self.g, root)
        return depth, time.perf_counter() - t0

    def search_bounded(self, root, max_entries):
        """(GTEPS, seconds, complete) of a serial search stopped after max_entries adjacency entries
        (port only: the reference's bfsCPU cannot be stopped)."""
        t0 = time.perf_counter()
        depth, scanned, _ = self.oracle.bfs_serial_bounded(self.g, root, max_entries)
        sec = time.perf_counter() - t0
        return (scanned / 2) / sec / 1e9, sec, scanned < max_entries

    def gteps(self, root, depth, seconds):
        _, dr = self.oracle.reach_stats(self.g, depth)
        return (dr / 2) / seconds / 1e9


def generate_edges_threaded(oracle, kind, scale, ef, seed):
    """The host generator over disjoint tuple ranges on all cores (input synthesis, untimed)."""
    from concurrent.futures import ThreadPoolExecutor
    m = int(oracle.lib().oracle_gen_num_edges(kind, scale, ef))
    n = int(oracle.lib().oracle_gen_num_vertices(kind, scale))
    src = np.empty(m, np.int32)
    dst = np.empty(m, np.int32)
    threads = max(1, min(os.cpu_count() or 1, 64))
    cuts = [m * i // (threads * 8) for i in range(threads * 8 + 1)]

    def work(i):
        a, b = cuts[i], cuts[i + 1]
        if b > a:
            oracle.lib().oracle_generate_edges(kind, scale, ef, seed, a, b - a, src[a:b], dst[a:b])

    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(threads * 8)))
    return n, src, dst


def giant_component_mask(oracle, g, deg):
    """depth != INT_MAX of one all-cores search from the highest-degree vertex (untimed)."""
    return oracle.bfs_parallel(g, int(np.argmax(deg))) != oracle.INT_MAX


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path, host cores only."""
    import oracle
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind = KINDS[args.kind]
    scale, note = args.scale, ""
    n = int(oracle.lib().oracle_gen_num_vertices(kind, scale))
    m = int(oracle.lib().oracle_gen_num_edges(kind, scale, args.ef))
    avail = mem_available_bytes()
    while cpu_graph_bytes(n, m) > 0.8 * avail and scale > 16:  # the host cannot hold the graph: say so, shrink
        scale -= 1
        n = int(oracle.lib().oracle_gen_num_vertices(kind, scale))
        m = int(oracle.lib().oracle_gen_num_edges(kind, scale, args.ef))
        note = (f"; host RAM ({avail / 2**30:.0f} GiB available) cannot hold scale-{args.scale}: "
                f"sampled on scale-{scale} of the same generator")
    n, src, dst = generate_edges_threaded(oracle, kind, scale, args.ef, args.seed)
    cpu = CpuReference(n, src, dst, scale)
    nl = nonloop_degree(n, src, dst)
    giant = giant_component_mask(oracle, cpu.g, nl)  # untimed: membership of the largest component
    roots = pick_roots(nl, args.steps + args.warmup, args.root_seed, giant)
    del src, dst
    # one serial search takes ~1 s at scale 22, ~5 s at scale 24, minutes at scale 27: keep the arm to a few minutes
    budget_s = args.cpu_budget if args.cpu_budget > 0 else 120.0
    teps, times, t_start = [], [], time.perf_counter()
    if cpu.kind == "reference":
        cpu.search(roots[0])  # warm-up: page the CSR in
        for r in roots[args.warmup:]:
            depth, sec = cpu.search(r)
            teps.append(cpu.gteps(r, depth, sec))
            times.append(sec)
            if time.perf_counter() - t_start > budget_s:
                break
        sample = (f"{len(teps)} of {args.steps} roots, one whole serial bfsCPU search each over the full {args.kind} "
                  f"scale-{scale} graph (bounded to ~{budget_s:.0f} s; {cpu.builder})")
    else:
        # E >= 2^31: the reference's int offsets overflow (bfs.cu:22-26), the 64-bit port of bfsCPU runs
        # instead, each step stopped after a fixed number of adjacency entries (a bounded sample)
        per_step = max(int(budget_s * 1.6e8 / max(args.steps, 1)), 1 << 24)
        cpu.search_bounded(roots[0], 1 << 24)
        complete = 0
        for r in roots[args.warmup:]:
            g, sec, done = cpu.search_bounded(r, per_step)
            teps.append(g)
            times.append(sec)
            complete += int(done)
            if time.perf_counter() - t_start > budget_s:
                break
        sample = (f"{len(teps)} of {args.steps} roots, 64-bit port of bfsCPU over the {args.kind} scale-{scale} graph, each "
                  f"search stopped after {per_step} adjacency entries ({complete} ran to completion); TEPS = entries "
                  f"scanned / 2 / time ({cpu.builder})")
    value = hmean(teps)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "GTEPS", "n_gpus": args.gpus,
            "steps": len(teps), "warmup": 1, "ms_per_step": 1e3 * float(np.mean(times)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": workload_config(args),
            "cpu_baseline": {"value": value, "unit": "GTEPS", "cores": 1, "kind": cpu.kind, "sample": sample + note},
            "e2e": {"value": value, "unit": "GTEPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "host_csr_build_s": cpu.build_s}
    print(json.dumps(line), flush=True)


def workload_config(args):
    """Identical for both arms: a function of the command line only."""
    kind = KINDS[args.kind]
    if kind == 2:
        side = 1 << (args.scale // 2)
        n, m = side * side, 2 * side * (side - 1)
    else:
        n, m = 1 << args.scale, args.ef << args.scale
    return {"workload": f"{args.kind} scale-{args.scale} edgefactor-{args.ef} seed-{args.seed}: n={n}, m={m} input tuples, "
                        f"roots = seeded sample (seed {args.root_seed}) of vertices with a non-self-loop edge inside the largest component",
            "mode": args.mode, "engine": args.engine, "dedup": int(not args.no_dedup),
            "multi_engine": ("nccl-per-level" if args.multi_engine else "peer-memory") if args.gpus > 1 else "n/a",
            "l2": "inputs larger than L2 (col array alone exceeds 126 MB at scale>=21); no flush between searches",
            "parallelism": f"1d-vertex-partition x{args.gpus}" if args.gpus > 1 else "single-gpu"}


def run_ours(args):
    import torch
    b2 = importlib.import_module("distributed-cuda-bfs_b200")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU baseline)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    # stdout carries ONE line, the JSON: whatever libraries write to file descriptor 1 meanwhile (NCCL
    # prints a version banner there under NCCL_DEBUG=VERSION) goes to stderr instead
    sys.stdout.flush()
    stdout_fd = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
        ident = [b2.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)  # side channel for the library's own communicator
        h = b2.Bfs(dev, rank=rank, world=world, nccl_id=ident[0], dedup=int(not args.no_dedup), gather_parent=0,
                   multi_engine=args.multi_engine)
    else:
        h = b2.Bfs(dev, dedup=int(not args.no_dedup))
    kind = KINDS[args.kind]
    h.set_option("mode", b2.MODE_TOP_DOWN if args.mode == "td" else b2.MODE_DIRECTION_OPT)
    h.set_option("engine", {"fused": b2.ENGINE_FUSED, "stepwise": b2.ENGINE_STEPWISE, "scan": b2.ENGINE_SCAN,
                            "pull": b2.ENGINE_PULL}[args.engine])
    peak, peak_kind = measured_peaks()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(values):
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().numpy()

    def load(scale):
        t0 = time.perf_counter()
        h.generate(kind, scale, args.ef, args.seed)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    def choose_roots(pinned, count):
        """(filtered, strict) root lists, the same on every rank: row degrees and the giant-component
        mask are gathered collectively."""
        n = h.num_vertices
        if args.no_dedup and world == 1:
            es, ed = h.edges()
            deg = nonloop_degree(n, es, ed)
            del es, ed
        elif args.no_dedup:
            deg = h.degrees()  # the multigraph's degrees (self-loops included), gathered over the ranks
        else:
            deg = h.row_degrees()
        h.run_collective(int(np.argmax(deg)), depth_out=pinned.array)  # untimed: membership of the largest component
        member = pinned.array != b2.INT_MAX
        return pick_roots(deg, count, args.root_seed, member), pick_roots(deg, count, args.root_seed, None), member

    def device_timed(roots):
        """Searches with results left in HBM; per-step device time is the max over ranks."""
        barrier()
        w0 = time.perf_counter()
        stats = [h.run_device(r) for r in roots]
        barrier()
        wall = time.perf_counter() - w0
        ms = max_over_ranks([s["search_ms"] for s in stats])
        m_r = np.array([s["reached_entries"] / 2 for s in stats])
        return stats, ms, m_r, wall

    def roofline_of(stats, ms, n):
        b_alg = np.array([4 * s["reached_entries"] + 8 * s["reached_vertices"] + 8 * n for s in stats], dtype=np.float64)
        achieved = float(b_alg.sum() / (ms.sum() * 1e-3) / 1e9)
        return b_alg, achieved

    # ---------------------------------------------------------------- the headline workload
    build_s = load(args.scale)
    n, m = h.num_vertices, h.num_input_edges
    pinned = b2.PinnedArray(n, np.int32)
    roots, strict_roots, member = choose_roots(pinned, args.steps + args.warmup)
    warm, timed = roots[:args.warmup], roots[args.warmup:args.warmup + args.steps]
    strict_timed = strict_roots[args.warmup:args.warmup + args.steps]
    out_buf = pinned.array if rank == 0 else None
    shared = my_slice = None
    if world > 1:
        # end-to-end on several GPUs: every rank copies ITS rows over its own PCIe link into one
        # page-locked shared-memory array (b2bfs_run_slice); rank 0 reads the whole result there
        name = f"b2bfs_bench_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
        if rank == 0:
            shared = b2.SharedPinnedArray(name, n, np.int32, create=True)
        barrier()
        if rank != 0:
            shared = b2.SharedPinnedArray(name, n, np.int32)
        first, cnt = h.local_range()
        my_slice = shared.array[first:first + cnt]

    def e2e_call(r):
        return h.run_slice(r, my_slice) if world > 1 else h.run_collective(r, depth_out=out_buf)

    for r in warm:
        h.run_device(r)
        e2e_call(r)

    with ClockSampler(dev) as clk:
        # ---- device-resident region: value + roofline
        stats, ms, m_r, wall_dev = device_timed(timed)
        # ---- end-to-end region: host buffers, gather + D2H inside the timed call
        e2e = []
        barrier()
        for r in timed:
            t1 = time.perf_counter()
            st = e2e_call(r)
            e2e.append((time.perf_counter() - t1, st))
        barrier()
        # ---- the same measurement over roots drawn by the bare Graph500 rule
        s_stats, s_ms, s_mr, _ = device_timed(strict_timed)
    e2e_s = max_over_ranks([sec for sec, _ in e2e])
    wall_dev = float(max_over_ranks([wall_dev])[0])
    launches = int(max_over_ranks([float(sum(s["kernel_launches"] for s in stats))])[0])
    teps = m_r / (ms * 1e-3) / 1e9
    value = hmean(teps)
    b_alg, achieved = roofline_of(stats, ms, n)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
            if t.get("scale") == args.scale and t.get("kind") == args.kind and world == 1:
                traffic = t.get("dram_bytes_per_launch")
    except Exception:
        pass
    e2e_teps = m_r / e2e_s / 1e9
    d2h_mine = float(np.mean([st["d2h_bytes"] for _, st in e2e]))
    if dist is not None:  # bytes over all ranks' links
        t = torch.tensor([d2h_mine], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        d2h_mine = float(t.item())
    d2h_total = int(d2h_mine)
    multi_engine = h.get_option("multi_engine") if world > 1 else None
    kernel = {"fused": "bfs_fused_kernel", "stepwise": "stepwise phases", "scan": "scan engine (next_layer/count/scan/assign)",
              "pull": "pull engine (pull_level_kernel)"}[args.engine] if world == 1 else \
        ("bfs_fused_multi_kernel (peer-memory engine, exchanges inside the kernel)" if multi_engine == 0 else
         "multi-rank phases (top_down_multi / bottom_up / apply_remote) + NCCL per level")
    s_teps = s_mr / (s_ms * 1e-3) / 1e9
    outside = int(sum(1 for r in strict_timed if not member[r]))

    line = {"metric": METRIC, "value": value, "unit": "GTEPS", "n_gpus": world, "steps": len(timed), "warmup": len(warm),
            "ms_per_step": float(ms.mean()), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic", "config": workload_config(args),
            "gteps_min": float(teps.min()), "gteps_median": float(np.median(teps)), "gteps_max": float(teps.max()),
            "levels_mean": float(np.mean([s["levels"] for s in stats])),
            "bu_levels_mean": float(np.mean([s["bu_levels"] for s in stats])),
            "build_s": build_s, "csr_entries_rank0": h.csr_entries, "wall_ms_per_step": 1e3 * wall_dev / len(timed),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak * world, "unit": "GB/s",
                         "frac": achieved / (peak * world), "traffic": traffic, "peak_kind": f"of {peak_kind} x{world} GPUs",
                         "kernel": kernel, "algorithmic_bytes_per_launch": float(b_alg.mean())},
            "e2e": {"value": hmean(e2e_teps), "unit": "GTEPS", "h2d_bytes_per_step": 8,
                    "d2h_bytes_per_step": d2h_total,
                    "ms_per_step": 1e3 * float(e2e_s.mean()),
                    "call": "b2bfs_run (depth array into page-locked host memory)" if world == 1 else
                            "b2bfs_run_slice on every rank (each copies its own rows into one shared page-locked array; "
                            "per-step time = slowest rank)"},
            "strict_graph500_rule": {"value": hmean(s_teps[s_mr > 0]) if np.any(s_mr > 0) else 0.0, "unit": "GTEPS",
                                     "roots": len(strict_timed), "roots_outside_largest_component": outside,
                                     "ms_per_step": float(s_ms.mean()),
                                     "note": "roots = seeded sample of vertices with degree >= 1, no component filter; a root in a "
                                             "2-3-vertex component contributes (1-2 edges) / (reset + launch time)"},
            "gpu_launches": launches, "clocks": clk.summary()}
    if world > 1:
        line["multi_engine"] = int(multi_engine)

    # ---------------------------------------------------------------- parity + CPU baseline
    if not args.no_cpu_baseline and world == 1:
        src, dst = h.edges()
        need = cpu_graph_bytes(n, m)
        if need > 0.8 * mem_available_bytes():
            line["cpu_baseline"] = {"value": None, "unit": "GTEPS", "cores": 1, "kind": "port",
                                    "sample": f"skipped: the host has {mem_available_bytes() / 2**30:.0f} GiB available, the CSR needs {need / 2**30:.0f}"}
            cpu = None
        else:
            cpu = CpuReference(n, src, dst, args.scale)
        del src, dst
        if cpu is not None:
            cteps, checked, t_start = [], 0, time.perf_counter()
            for r in timed:
                depth_cpu, sec = cpu.search(r)
                cteps.append(cpu.gteps(r, depth_cpu, sec))
                depth_gpu, _ = h.run(r, depth_out=pinned.array)
                bad = cpu.oracle.check_output(depth_gpu, depth_cpu)
                if bad != -1:
                    raise SystemExit(f"bench.py: PARITY FAILURE root {r} vertex {bad}: gpu {depth_gpu[bad]} cpu {depth_cpu[bad]}")
                checked += 1
                if time.perf_counter() - t_start > args.cpu_budget:
                    break
            line["cpu_baseline"] = {"value": hmean(cteps), "unit": "GTEPS", "cores": 1, "kind": cpu.kind,
                                    "sample": f"{checked} of the {len(timed)} timed roots, one serial bfsCPU search each over the "
                                              f"full graph (~{args.cpu_budget:.0f} s budget; {cpu.builder}); depths bit-exact vs the GPU on all {checked}"}
            line["parity"] = {"roots_checked": checked, "against": f"bfsCPU ({cpu.kind}), all n depths per root", "result": "bit-exact"}
            # B2: the same searches on ALL host cores with the level-synchronous port of bfsCPU,
            # so that the CPU side is not flattered by the reference being serial
            mteps, t_start = [], time.perf_counter()
            for r in timed[:checked]:
                t1 = time.perf_counter()
                depth_mc = cpu.oracle.bfs_parallel(cpu.g, r)
                sec = time.perf_counter() - t1
                _, dr = cpu.oracle.reach_stats(cpu.g, depth_mc)
                mteps.append((dr / 2) / sec / 1e9)
                if time.perf_counter() - t_start > args.cpu_budget / 2:
                    break
            line["cpu_baseline_multicore"] = {"value": hmean(mteps), "unit": "GTEPS", "cores": cpu.oracle.num_threads(),
                                              "kind": "port", "sample": f"{len(mteps)} of the same roots, level-synchronous "
                                              "all-cores port of bfsCPU (oracle_bfs_parallel)"}
            del cpu
    elif not args.no_cpu_baseline:
        # several GPUs: every rank takes part in the searches, rank 0 alone checks the gathered depths
        k = min(args.parity_roots, len(timed))
        kept = []
        for r in timed[:k]:  # through the same call the end-to-end figure times
            h.run_slice(r, my_slice)
            barrier()  # every rank's rows are in the shared array
            if rank == 0:
                kept.append(shared.array.copy())
            barrier()
        if rank == 0:
            line["parity"] = multi_gpu_parity(args, kind, n, timed[:k], kept, line)
            line["parity"]["multi_engine"] = int(multi_engine)
        barrier()

    # ---------------------------------------------------------------- last round's sizes, like for like
    if args.legacy_scale and not args.no_legacy:
        key = "rmat22" if world == 1 else "weak22"
        pinned.free()
        load(args.legacy_scale)
        n2 = h.num_vertices
        pinned = b2.PinnedArray(n2, np.int32)
        r2, _, _ = choose_roots(pinned, args.steps + args.warmup)
        for r in r2[:args.warmup]:
            h.run_device(r)
        st2, ms2, mr2, _ = device_timed(r2[args.warmup:args.warmup + args.steps])
        _, ach2 = roofline_of(st2, ms2, n2)
        line[key] = {"workload": f"{args.kind} scale-{args.legacy_scale} edgefactor-{args.ef}", "value": hmean(mr2 / (ms2 * 1e-3) / 1e9),
                     "unit": "GTEPS", "ms_per_step": float(ms2.mean()), "frac": ach2 / (peak * world), "steps": len(st2)}
    pinned.free()
    if shared is not None:
        my_slice = None
        shared.free()
    h.close()
    sys.stdout.flush()
    os.dup2(stdout_fd, 1)
    os.close(stdout_fd)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def multi_gpu_parity(args, kind, n, roots, depths, line):
    """Rank 0, host cores only: are the depth arrays the N-GPU job gathered the BFS depths?"""
    import oracle
    t_start = time.perf_counter()
    m = int(oracle.lib().oracle_gen_num_edges(kind, args.scale, args.ef))
    if 8 * m + 6 * n > 0.8 * mem_available_bytes():
        return {"roots_checked": 0, "result": f"skipped: the host cannot hold the {8 * m / 2**30:.0f} GiB tuple list"}
    _, src, dst = generate_edges_threaded(oracle, kind, args.scale, args.ef, args.seed)  # host generator, not the GPU's tuples
    checked = 0
    for r, d in zip(roots, depths):
        bad = oracle.check_depth_by_edges(n, src, dst, r, d)
        if bad != -1:
            raise SystemExit(f"bench.py: PARITY FAILURE (edge rules) root {r} vertex {bad}: gpu depth {d[bad]}")
        checked += 1
    out = {"roots_checked": checked, "result": "bit-exact",
           "against": "edge rules over the host-generated tuples (depth[root]=0 alone; |depth[u]-depth[v]|<=1 and same "
                      "reachability on every tuple; every reached vertex has a neighbour one level up) - together they "
                      "admit exactly one depth array, bfsCPU's",
           "edge_rule_s": time.perf_counter() - t_start}
    # bfsCPU itself on as many of those roots as the budget allows (the reference's own code while
    # 2m < 2^31, its 64-bit port beyond)
    left = args.cpu_budget - (time.perf_counter() - t_start)
    est_build = 4.0 * (m / 2**26) * 8.0 / max(os.cpu_count() or 8, 1)  # ~4 s per 2^26 tuples on 8 cores, measured
    if cpu_graph_bytes(n, m) < 0.8 * mem_available_bytes() and est_build < left:
        cpu = CpuReference(n, src, dst, args.scale)
        del src, dst
        cteps, done = [], 0
        for r, d in zip(roots, depths):
            if done and time.perf_counter() - t_start > args.cpu_budget:
                break
            depth_cpu, sec = cpu.search(r)
            cteps.append(cpu.gteps(r, depth_cpu, sec))
            bad = oracle.check_output(d, depth_cpu)
            if bad != -1:
                raise SystemExit(f"bench.py: PARITY FAILURE root {r} vertex {bad}: gpu {d[bad]} cpu {depth_cpu[bad]}")
            done += 1
        out["bfs_cpu_roots"] = done
        out["against"] += f"; plus bfsCPU ({cpu.kind}) element for element on {done} of them"
        line["cpu_baseline"] = {"value": hmean(cteps), "unit": "GTEPS", "cores": 1, "kind": cpu.kind,
                                "sample": f"{done} of the timed roots, one serial bfsCPU search each over the full graph ({cpu.builder})"}
    else:
        out["bfs_cpu_roots"] = 0
        out["bfs_cpu_skipped"] = "host CSR build does not fit the time/RAM budget of this run"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default=None, choices=list(KINDS),
                    help="default: rmat on 1 GPU (north_star's target graph); kron (RMAT with scrambled labels, so a contiguous "
                         "1-D partition is balanced) on several")
    ap.add_argument("--scale", type=int, default=None, help="default: 24 + log2(gpus) - weak scaling, 2^24 vertices per GPU")
    ap.add_argument("--ef", type=int, default=16)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--root-seed", type=int, default=2)
    ap.add_argument("--mode", default="do", choices=["do", "td"])
    ap.add_argument("--engine", default="fused", choices=["fused", "stepwise", "scan", "pull"],
                    help="scan / pull: the reference's alternative strategies as comparators (one GPU)")
    ap.add_argument("--multi-engine", type=int, default=0, choices=[0, 1], help="several GPUs: 0 peer-memory kernel, 1 NCCL per level")
    ap.add_argument("--no-dedup", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU baseline AND the parity check")
    ap.add_argument("--no-legacy", action="store_true", help="skip the rmat22 / weak22 re-measurement of round 1's sizes")
    ap.add_argument("--parity-roots", type=int, default=3, help="several GPUs: timed roots whose gathered depths rank 0 checks")
    ap.add_argument("--cpu-budget", type=float, default=None,
                    help="seconds of CPU baseline work (default: 20 on one GPU, 60 on several, 120 in the reference arm)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    if args.impl == "ours" and world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py: --gpus N > 1 must be launched with torch.distributed.run (one process per GPU)")
    args.gpus = world if args.impl == "ours" else args.gpus
    lg = max(0, int(np.log2(max(args.gpus, 1))))
    defaults = args.kind is None and args.scale is None
    if args.kind is None:
        args.kind = "rmat" if args.gpus == 1 else "kron"
    if args.scale is None:
        args.scale = 24 + lg
    args.legacy_scale = 22 + lg if defaults else 0
    if args.cpu_budget is None:
        args.cpu_budget = 0.0 if args.impl == "reference" else (20.0 if args.gpus == 1 else 60.0)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
