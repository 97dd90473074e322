import sys, time, importlib, faulthandler
faulthandler.enable()
sys.path.insert(0, '.')
t0=time.time()
def log(*a): print(f"[{time.time()-t0:7.2f}]", *a, flush=True)
import numpy as np, oracle
b2 = importlib.import_module("distributed-cuda-bfs_b200")
log("imports done")
h = b2.Bfs(0); log("handle, grid", h.get_option("grid"))
h.generate(0, 12, 16, 1); log("generated, entries", h.csr_entries, "hubs", h.get_option("num_hubs"), "maxdeg", h.get_option("max_degree"))
n,s,d = oracle.generate_edges(0,12,16,1); g = oracle.build_csr(n,s,d); want = oracle.bfs_serial(g,2)
for mode,eng in ((1,1),(0,1),(1,0),(0,0)):
    h.set_option("mode",mode); h.set_option("engine",eng)
    log("run mode",mode,"engine",eng)
    depth,st = h.run(2)
    log("  done", st, "mismatch at", oracle.check_output(depth,want))
