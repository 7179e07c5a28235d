"""Multi-GPU host logic: replicated cloud, query shards by Morton range.

One process per GPU.  Rank 0 owns the scan; `broadcast_cloud` replicates it (NCCL over NVLink on the GPU box,
gloo in the CPU tests), every rank builds the identical Morton order (the device radix sort is deterministic),
answers the contiguous range `shard_range(n, rank, world)` of the SORTED order, and `gather_rows` all-gathers
the per-point rows; `unpermute` (ccg_unpermute on the GPU) scatters them back to original order.  There is no
other exchange on the path: queries are independent given read-only access to the whole cloud.

This module holds no compute: the per-range work is a callable so that the CPU tests (gloo, world_size 2) can
drive the same logic with the oracle standing in for the GPU.
"""
import torch
import torch.distributed as dist

LEAF = 32   # ccg ranges must begin on a leaf boundary (include/ccgpu.h, ccg_*_range)


def shard_range(n, rank, world):
    """-> (begin, end, per): sorted positions [begin, end) of `rank`; per = rows every rank contributes to the
    all-gather (a multiple of 32; the last ranks may own fewer or no points)."""
    nleaf = (n + LEAF - 1) // LEAF
    per = (nleaf + world - 1) // world * LEAF
    if rank * per >= n:                      # more ranks than leaves: an empty range on a leaf boundary
        b = n // LEAF * LEAF
        return b, b, per
    return rank * per, min(n, (rank + 1) * per), per


def broadcast_cloud(points, src=0):
    """replicate the [n, S] float32 cloud from `src` (in place on the other ranks)"""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(points, src)
    return points


def gather_rows(local_rows, per, world):
    """all-gather equal-sized [per, C] row blocks -> [world*per, C] (sorted order, tail rows beyond n unused)"""
    if world == 1:
        return local_rows
    out = torch.empty((world * per,) + tuple(local_rows.shape[1:]), dtype=local_rows.dtype, device=local_rows.device)
    dist.all_gather_into_tensor(out, local_rows.contiguous())
    return out


def sharded_rows(n, compute_range, width, dtype=torch.float32, device="cpu"):
    """Run `compute_range(begin, end, out_rows)` on this rank's shard and return the gathered [world*per, width]
    rows in sorted order.  `out_rows` is a [per, width] buffer whose first end-begin rows must be filled."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    begin, end, per = shard_range(n, rank, world)
    local = torch.zeros((per, width), dtype=dtype, device=device)
    if end > begin:
        compute_range(begin, end, local)
    return gather_rows(local, per, world)
