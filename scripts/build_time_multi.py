"""Graph construction time on several GPUs, first and second load on the same communicator (torchrun)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
b2 = importlib.import_module("distributed-cuda-bfs_b200")
ident = [b2.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ident, src=0)
h = b2.Bfs(local, rank=rank, world=world, nccl_id=ident[0], gather_parent=0)
scale = int(sys.argv[1])
for i in range(3):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    h.generate(1, scale, 16, 1 + i)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if rank == 0:
        print(f"world {world} kron scale {scale} load #{i + 1}: {dt:.3f} s", flush=True)
h.close()
dist.destroy_process_group()
