#!/usr/bin/env python
"""bench.py — GTEPS (Graph500 harmonic mean over the timed roots) of the BFS hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scale S] ...

A "step" is one search from one root over the resident synthetic graph.  Default workload =
BASELINE.json configs[1]: RMAT scale-22 edgefactor-16 on one B200 (536 MB of column indices — larger
than the 126 MB L2, so consecutive searches cannot live in cache).  One JSON line on stdout.

  value      harmonic-mean GTEPS, device-timed (CUDA events on the library's stream, inside
             b2bfs_run_device), graph resident in HBM, results left in HBM
  e2e        the same metric through b2bfs_run with HOST buffers: root id in, n x int32 depth array
             copied device->host into page-locked memory inside the timed region (host wall clock)
  roofline   the dominant kernel (the fused search kernel = the whole step): algorithmic bytes
             B_alg(r) = 4*D_r + 8*V_r + 8*n / its CUDA-event time, against the
             measured HBM copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline  the reference's own bfsCPU (oracle/_ref, else the oracle port) on a bounded sample
             of the same roots, on this box's host cores; it is also the parity check of the run

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
class CpuReference:
    def __init__(self, n, src, dst):
        import oracle
        self.oracle = oracle
        t0 = time.perf_counter()
        self.g = oracle.build_csr(n, src, dst)  # restates readGraphFromFile's CSR, bfs.cu:851-875
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
    threads = max(1, min(os.cpu_count() or 1, 32))
    cuts = [m * i // threads for i in range(threads + 1)]

    def work(i):
        a, b = cuts[i], cuts[i + 1]
        if b > a:
            oracle.lib().oracle_generate_edges(kind, scale, ef, seed, a, b - a, src[a:b], dst[a:b])

    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(threads)))
    return n, src, dst


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path, host cores only."""
    import oracle
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind = KINDS[args.kind]
    n, src, dst = generate_edges_threaded(oracle, kind, args.scale, args.ef, args.seed)
    cpu = CpuReference(n, src, dst)
    nl = nonloop_degree(n, src, dst)
    giant, _ = cpu.search(int(np.argmax(nl)))  # untimed: membership of the largest component
    roots = pick_roots(nl, args.steps + args.warmup, args.root_seed, giant != oracle.INT_MAX)
    # one serial search takes ~1 s at scale 22 and ~20 s at scale 25: keep the whole arm to a few minutes
    budget_s = args.cpu_budget if args.cpu_budget > 0 else (150.0 if args.scale <= 23 else 60.0)
    # warm-up: the membership search above already ran the whole loop once
    teps, times, t_start = [], [], time.perf_counter()
    for r in roots[args.warmup:]:
        depth, sec = cpu.search(r)
        teps.append(cpu.gteps(r, depth, sec))
        times.append(sec)
        if time.perf_counter() - t_start > budget_s:
            break
    value = hmean(teps)
    sample = (f"{len(teps)} of {args.steps} roots, one serial search each over the full {args.kind} scale-{args.scale} "
              f"graph (bounded to ~{budget_s:.0f} s)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "GTEPS", "n_gpus": args.gpus,
            "steps": len(teps), "warmup": 1, "ms_per_step": 1e3 * float(np.mean(times)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": workload_config(args, n, len(src)),
            "cpu_baseline": {"value": value, "unit": "GTEPS", "cores": 1, "kind": cpu.kind, "sample": sample},
            "e2e": {"value": value, "unit": "GTEPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, n, m):
    return {"workload": f"{args.kind} scale-{args.scale} edgefactor-{args.ef} seed-{args.seed}: n={n}, m={m} input tuples, "
                        f"roots = seeded sample (seed {args.root_seed}) of vertices with a non-self-loop edge inside the largest component",
            "mode": args.mode, "engine": args.engine, "dedup": int(not args.no_dedup),
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
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
        ident = [b2.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)  # side channel for the library's own communicator
        h = b2.Bfs(dev, rank=rank, world=world, nccl_id=ident[0], dedup=int(not args.no_dedup), gather_parent=0)
    else:
        h = b2.Bfs(dev, dedup=int(not args.no_dedup))
    kind = KINDS[args.kind]
    h.set_option("mode", b2.MODE_TOP_DOWN if args.mode == "td" else b2.MODE_DIRECTION_OPT)
    h.set_option("engine", b2.ENGINE_STEPWISE if args.engine == "stepwise" else b2.ENGINE_FUSED)
    t0 = time.perf_counter()
    h.generate(kind, args.scale, args.ef, args.seed)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    n, m = h.num_vertices, h.num_input_edges

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(values):
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().numpy()

    # roots: the same on every rank (row degrees and the giant-component mask are gathered collectively)
    if args.no_dedup:
        es, ed = h.edges()
        deg = nonloop_degree(n, es, ed)
        del es, ed
    else:
        deg = h.row_degrees()
    pinned = b2.PinnedArray(n, np.int32)
    h.run_collective(int(np.argmax(deg)), depth_out=pinned.array)  # untimed: membership of the largest component
    roots = pick_roots(deg, args.steps + args.warmup, args.root_seed, pinned.array != b2.INT_MAX)
    warm, timed = roots[:args.warmup], roots[args.warmup:args.warmup + args.steps]
    out_buf = pinned.array if rank == 0 else None

    for r in warm:
        h.run_device(r)
        h.run_collective(r, depth_out=out_buf)

    # ---- device-resident region: value + roofline
    stats = []
    with ClockSampler(dev) as clk:
        barrier()
        w0 = time.perf_counter()
        for r in timed:
            stats.append(h.run_device(r))
        barrier()
        wall_dev = time.perf_counter() - w0
        # ---- end-to-end region: host buffers, gather + D2H inside the timed call
        e2e = []
        barrier()
        for r in timed:
            t1 = time.perf_counter()
            st = h.run_collective(r, depth_out=out_buf)
            e2e.append((time.perf_counter() - t1, st))
        barrier()
    ms = max_over_ranks([s["search_ms"] for s in stats])          # device time, max over ranks per step
    e2e_s = max_over_ranks([sec for sec, _ in e2e])
    wall_dev = float(max_over_ranks([wall_dev])[0])
    launches = int(max_over_ranks([float(sum(s["kernel_launches"] for s in stats))])[0])
    m_r = np.array([s["reached_entries"] / 2 for s in stats])
    teps = m_r / (ms * 1e-3) / 1e9
    value = hmean(teps)
    b_alg = np.array([4 * s["reached_entries"] + 8 * s["reached_vertices"] + 8 * n for s in stats], dtype=np.float64)
    peak, peak_kind = measured_peaks()
    achieved = float(b_alg.sum() / (ms.sum() * 1e-3) / 1e9)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
            if t.get("scale") == args.scale and t.get("kind") == args.kind and world == 1:
                traffic = t.get("dram_bytes_per_launch")
    except Exception:
        pass
    e2e_teps = m_r / e2e_s / 1e9
    kernel = ("bfs_fused_kernel" if args.engine == "fused" else "stepwise phases") if world == 1 else \
        "multi-rank phases (top_down_multi / bottom_up / apply_remote) + NCCL"

    line = {"metric": METRIC, "value": value, "unit": "GTEPS", "n_gpus": world, "steps": len(timed), "warmup": len(warm),
            "ms_per_step": float(ms.mean()), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic", "config": workload_config(args, n, m),
            "gteps_min": float(teps.min()), "gteps_median": float(np.median(teps)), "gteps_max": float(teps.max()),
            "levels_mean": float(np.mean([s["levels"] for s in stats])),
            "bu_levels_mean": float(np.mean([s["bu_levels"] for s in stats])),
            "build_s": build_s, "csr_entries_rank0": h.csr_entries, "wall_ms_per_step": 1e3 * wall_dev / len(timed),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak * world, "unit": "GB/s",
                         "frac": achieved / (peak * world), "traffic": traffic, "peak_kind": f"of {peak_kind} x{world} GPUs",
                         "kernel": kernel, "algorithmic_bytes_per_launch": float(b_alg.mean())},
            "e2e": {"value": hmean(e2e_teps), "unit": "GTEPS", "h2d_bytes_per_step": 8,
                    "d2h_bytes_per_step": int(np.mean([st["d2h_bytes"] for _, st in e2e])),
                    "ms_per_step": 1e3 * float(e2e_s.mean())},
            "gpu_launches": launches, "clocks": clk.summary()}

    # ---- CPU baseline + parity check on a bounded sample of the same roots (N = 1 only)
    if not args.no_cpu_baseline and world == 1:
        src, dst = h.edges()
        cpu = CpuReference(n, src, dst)
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
                                          f"full graph (~{args.cpu_budget:.0f} s budget); depths bit-exact vs the GPU on all {checked}"}
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
    pinned.free()
    h.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default=None, choices=list(KINDS),
                    help="default: rmat on 1 GPU (BASELINE configs[1]); kron (RMAT with scrambled labels, so a contiguous "
                         "1-D partition is balanced) on several")
    ap.add_argument("--scale", type=int, default=None, help="default: 22 + log2(gpus) — weak scaling, fixed tuples per GPU")
    ap.add_argument("--ef", type=int, default=16)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--root-seed", type=int, default=2)
    ap.add_argument("--mode", default="do", choices=["do", "td"])
    ap.add_argument("--engine", default="fused", choices=["fused", "stepwise"])
    ap.add_argument("--no-dedup", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="seconds of CPU baseline work (ours arm); 0 = default 150 s in the reference arm")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    if args.impl == "ours" and world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py: --gpus N > 1 must be launched with torch.distributed.run (one process per GPU)")
    args.gpus = world if args.impl == "ours" else args.gpus
    if args.kind is None:
        args.kind = "rmat" if args.gpus == 1 else "kron"
    if args.scale is None:
        args.scale = 22 + max(0, int(np.log2(max(args.gpus, 1))))
    if args.impl == "reference":
        if args.cpu_budget == 20.0:
            args.cpu_budget = 0
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
