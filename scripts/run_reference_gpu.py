"""Time the reference's GPU build (oracle/_ref/ref_gpu_bfs, D1-patched, see oracle/Makefile) beside
ours on the same edge-list files: BASELINE.md rows B3.  Also records what the AS-SHIPPED binary does."""
import importlib, os, re, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import oracle
b2 = importlib.import_module("distributed-cuda-bfs_b200")
REF = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_bfs_2dev" if os.environ.get("REF_2DEV") else "ref_gpu_bfs")
ASIS = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_bfs_asshipped")
scales = [int(x) for x in sys.argv[1:]] or [16, 18]
tmp = tempfile.mkdtemp()
for scale in scales:
    n, s, d = oracle.generate_edges(oracle.GEN_RMAT, scale, 16, 1)
    path = os.path.join(tmp, f"rmat{scale}.txt")
    oracle.write_edge_file(path, n, s, d)
    m_r = None
    t0 = time.perf_counter()
    r = subprocess.run([REF, "2", path], capture_output=True, text=True, timeout=900)
    wall = time.perf_counter() - t0
    t = [int(x) for x in re.findall(r"Elapsed time in milliseconds : (\d+) ms", r.stdout)]
    ok = "Output OK!" in r.stdout
    with b2.Bfs(0) as h:
        h.load_edge_file(path)
        for _ in range(3):
            h.run_device(2)
        st = [h.run_device(2) for _ in range(10)]
        ours = float(np.median([x["search_ms"] for x in st]))
        m_r = st[0]["reached_entries"] / 2
    ref_gpu_ms = t[1] if len(t) > 1 else None
    print(f"RMAT-{scale} src=2: reference GPU build level loop {ref_gpu_ms} ms ({m_r / (max(ref_gpu_ms, 1) * 1e6):.3f} GTEPS, "
          f"'Output OK!': {ok}, rc {r.returncode}, whole process {wall:.1f} s incl. its loader and bfsCPU {t[0] if t else None} ms); "
          f"ours {ours:.3f} ms ({m_r / (ours * 1e6):.1f} GTEPS)", flush=True)
    if scale == scales[0] and os.path.exists(ASIS):
        try:
            a = subprocess.run([ASIS, "2", path], capture_output=True, text=True, timeout=300)
            tail = [l for l in a.stdout.strip().splitlines() if l][-2:]
            print(f"  as-shipped binary (DeviceNum 1, unpatched): rc {a.returncode}, last lines: {tail}", flush=True)
        except subprocess.TimeoutExpired:
            print("  as-shipped binary: timed out", flush=True)
