"""2+ rank check of the fused query + NVLink scatter (run under torchrun; used by tests/test_gpu_scatter.py):
every rank answers its Morton range with ccg_pca_knn_features_scatter, writing rows in original order straight into
rank 0's arrays through IPC peer memory; rank 0 compares them bit for bit with its own single-GPU full call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import cloudclean_b200 as cc
from cloudclean_b200 import sharding, synth

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
if rank == 0:
    pts = synth.config_scan(1, device=dev, shrink=4)["points"].contiguous()
    nt = torch.tensor([pts.shape[0]], device=dev)
else:
    nt = torch.zeros(1, dtype=torch.int64, device=dev)
dist.broadcast(nt, 0)
n = int(nt.item())
if rank != 0:
    pts = torch.empty((n, 8), device=dev)
ctx = cc.Context(local)
peer = sharding.PeerRows(ctx, n, (3, 4, 7), collector=0)
token = torch.zeros(1, device=dev)
for it in range(2):
    dist.broadcast(pts, 0)
    torch.cuda.synchronize()
    ctx.set_cloud(pts)
    b, e, per = sharding.shard_range(n, rank, world)
    ctx.pca_knn_features_scatter(20, cc.COV_NM1, 0, b, e, *peer.bufs)
    torch.cuda.synchronize()
    dist.all_reduce(token)
    torch.cuda.synchronize()
ok = True
if rank == 0:
    got = [b_.tensor(dev).clone() for b_ in peer.bufs]
    eig = torch.empty((n, 3), device=dev); nrm = torch.empty((n, 4), device=dev); feat = torch.empty((n, 7), device=dev)
    ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat)
    for a, g in zip((eig, nrm, feat), got):
        same = torch.equal(a.view(torch.int32), g.view(torch.int32))
        ok = ok and same
    print(f"scatter check: n={n} world={world} bit-identical={ok}", flush=True)
dist.barrier()
peer.close()
ctx.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
