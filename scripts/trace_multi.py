"""Per-phase timeline of the peer-memory multi-rank kernel (run under torchrun); prints rank 0 and the last rank."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
b2 = importlib.import_module("distributed-cuda-bfs_b200")
ident = [b2.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ident, src=0)
h = b2.Bfs(local, rank=rank, world=world, nccl_id=ident[0], gather_parent=0)
kind, scale, nroots = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
for a in sys.argv[4:]:
    k, v = a.split("="); h.set_option(k, int(v))
h.generate(kind, scale, 16, 1)
deg = h.row_degrees()
roots = np.random.default_rng(2).choice(np.flatnonzero(deg > 0), nroots + 2, replace=False)
names = {-1: "start", 0: "reset", 1: "td", 2: "bu", 3: "q2b+push+x", 4: "b2q", 5: "grp", 6: "td", 7: "x1", 8: "apply+x2", 9: "bu", 10: "push+x", 11: "[all-arrived]", 12: "[peers-in]"}
for r in roots:
    st = h.run_device(int(r))
    tr = h.last_trace()
    prev = 0.0; parts = []
    for lv, ph, us in sorted(tr, key=lambda t: t[2])[:90]:
        parts.append(f"{names[ph]}@{lv}:{us - prev:.0f}"); prev = us
    if rank in (0, world - 1):
        print(f"[rank {rank}] root {r} search_ms {st['search_ms']:.3f} levels {st['levels']} | " + " ".join(parts), flush=True)
    dist.barrier()
h.close()
dist.destroy_process_group()
