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
npad = (n + 31) // 32 * 32
peer = sharding.PeerRows(ctx, npad, (3, 4, 7), collector=0)
token = torch.zeros(1, device=dev)
ok = True
names = ["contiguous range, original-order rows", "interleaved chunks, original-order rows", "interleaved chunks, Morton-order rows + un-permute"]
for it in range(3):
    if rank == 0:
        for b_ in peer.bufs:
            b_.tensor(dev).fill_(-3.0)
    dist.broadcast(pts, 0)
    torch.cuda.synchronize()
    ctx.set_cloud(pts)
    b, e, per = sharding.shard_range(n, rank, world)
    if it == 0:
        ctx.pca_knn_features_scatter(20, cc.COV_NM1, 0, b, e, *peer.bufs)
    elif it == 1:
        ctx.pca_knn_features_shard(20, cc.COV_NM1, 0, rank, world, *peer.bufs, chunk_points=1024)
    else:
        ctx.pca_knn_features_shard(20, cc.COV_NM1, 0, rank, world, *peer.bufs, chunk_points=1024, sorted_rows=True)
    torch.cuda.synchronize()
    dist.all_reduce(token)
    torch.cuda.synchronize()
    if rank == 0:
        got = []
        for b_ in peer.bufs:
            if it == 2:                        # rows at their Morton position: un-permute on the collector
                o = torch.empty((n, b_.shape[1]), device=dev)
                ctx.unpermute(b_.tensor(dev), o)
            else:
                o = b_.tensor(dev)[:n].clone()
            got.append(o)
        eig = torch.empty((n, 3), device=dev); nrm = torch.empty((n, 4), device=dev); feat = torch.empty((n, 7), device=dev)
        ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat)
        same = all(torch.equal(a_.view(torch.int32), g.view(torch.int32)) for a_, g in zip((eig, nrm, feat), got))
        print(f"scatter check [{names[it]}]: n={n} world={world} same={same}", flush=True)
        ok = ok and same
if rank == 0:
    print(f"scatter check: n={n} world={world} bit-identical={ok}", flush=True)
dist.barrier()
peer.close()
ctx.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
