import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import oracle
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
b2 = importlib.import_module("distributed-cuda-bfs_b200")
ident = [b2.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ident, src=0)
n, s, d = oracle.generate_edges(0, 14, 16, 3)
g = oracle.build_csr(n, s, d)
want = oracle.bfs_serial(g, 3146)
for hub_first in (1, 0):
    h = b2.Bfs(local, rank=rank, world=world, nccl_id=ident[0], hub_first=hub_first)
    h.generate(0, 14, 16, 3)
    part = h.get_option("part")
    deg = h.degrees()
    sl = slice(rank * part, (rank + 1) * part)
    exp = int(g.sizes[sl][want[sl] != oracle.INT_MAX].sum())
    for engine in (0, 1):
        h.set_option("multi_engine", engine)
        depth, st = h.run(3146)
        ok = oracle.check_output(depth, want)
        print(f"rank {rank} hub_first {hub_first} engine {engine}: depth_ok {ok==-1} local reached_entries {st['reached_entries']} expected {exp} reached_v {st['reached_vertices']} exp_v {(want[sl]!=oracle.INT_MAX).sum()} deg_ok {np.array_equal(deg, g.sizes)}", flush=True)
    dist.barrier()
    h.close()
    dist.barrier()
ident2 = None
dist.destroy_process_group()
