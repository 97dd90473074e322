"""Option sweep on several GPUs (run under torchrun): harmonic-mean GTEPS over the same roots for every
value of the listed options.   torchrun ... scripts/sweep_multi.py <kind> <scale> <nroots> gamma=0,400 alpha=14,28"""
import importlib, itertools, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
b2 = importlib.import_module("distributed-cuda-bfs_b200")
ident = [b2.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ident, src=0)
h = b2.Bfs(local, rank=rank, world=world, nccl_id=ident[0], gather_parent=0)
kind, scale, nroots = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
axes = [(a.split("=")[0], [int(x) for x in a.split("=")[1].split(",")]) for a in sys.argv[4:]]
h.generate(kind, scale, 16, 1)
n = h.num_vertices
deg = h.row_degrees()
depth = np.empty(n, np.int32)
h.run_collective(int(np.argmax(deg)), depth_out=depth)
cand = np.flatnonzero((deg > 0) & (depth != b2.INT_MAX))
roots = [int(r) for r in np.random.default_rng(2).choice(cand, nroots + 3, replace=False)]
for combo in itertools.product(*[v for _, v in axes]):
    for (k, _), v in zip(axes, combo):
        h.set_option(k, v)
    for r in roots[:3]:
        h.run_device(r)
    dist.barrier(); torch.cuda.synchronize()
    st = [h.run_device(r) for r in roots[3:]]
    t = torch.tensor([s["search_ms"] for s in st], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.cpu().numpy()
    teps = np.array([s["reached_entries"] / 2 for s in st]) / (ms * 1e-3) / 1e9
    if rank == 0:
        print(json.dumps({"gpus": world, "kind": kind, "scale": scale, **{k: v for (k, _), v in zip(axes, combo)},
                          "gteps_hmean": round(float(len(teps) / np.sum(1 / teps)), 1), "ms_mean": round(float(ms.mean()), 4),
                          "ms_min": round(float(ms.min()), 4), "ms_max": round(float(ms.max()), 4),
                          "bu_levels_mean": float(np.mean([s["bu_levels"] for s in st])),
                          "levels_mean": float(np.mean([s["levels"] for s in st]))}), flush=True)
h.close()
dist.barrier()
dist.destroy_process_group()
