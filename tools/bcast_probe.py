"""Developer probe (torchrun): time dist.broadcast of the replicated cloud (12 B/point, 30 M points = 361 MB) under the current NCCL
environment; also all_gather of the same bytes."""
import os, sys, torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", "0")); torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
n = 30_071_989
buf = torch.empty((n, 3), device=dev)
if rank == 0:
    buf.uniform_()
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
ms = timed(lambda: dist.broadcast(buf, 0))
rpr = (n + world - 1) // world
sl = torch.empty((rpr, 3), device=dev); full = torch.empty((world * rpr, 3), device=dev)
ms2 = timed(lambda: dist.all_gather_into_tensor(full, sl))
if rank == 0:
    tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("NCCL_"))
    print(f"[{tag or 'default'}] broadcast {n * 12 / 1e6:.0f} MB: {ms:.3f} ms = {n * 12 / ms / 1e6:.0f} GB/s | all_gather: {ms2:.3f} ms", flush=True)
dist.destroy_process_group()
