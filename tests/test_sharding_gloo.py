"""The N>1 host logic on CPU: world_size-2 gloo processes, the oracle standing in for the GPU per-range compute.
Checks that replicate -> per-rank Morton range -> all-gather -> un-permute reproduces the unsharded rows exactly."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cloudclean_b200 import sharding


def test_shard_range_partitions_exactly():
    for n in (0, 1, 31, 32, 33, 1000, 4097, 30_000_001):
        for world in (1, 2, 3, 4, 8):
            covered = 0
            prev_end = 0
            for r in range(world):
                b, e, per = sharding.shard_range(n, r, world)
                assert b % 32 == 0 and per % 32 == 0 and 0 <= b <= e <= n and e - b <= per
                assert b == prev_end or b == e
                prev_end = e
                covered += e - b
            assert covered == n


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, seed, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    rng = np.random.default_rng(seed)
    if rank == 0:
        pts = torch.from_numpy(np.ascontiguousarray(rng.normal(size=(n, 8)).astype(np.float32)))
    else:
        pts = torch.zeros((n, 8), dtype=torch.float32)
    sharding.broadcast_cloud(pts, 0)
    cloud = pts.numpy()
    # every rank derives the same order from the same cloud (stand-in for the deterministic device sort)
    order = np.lexsort((cloud[:, 2], cloud[:, 1], cloud[:, 0])).astype(np.int64)
    T = oracle.Tree(cloud)

    def compute_range(begin, end, out_rows):
        rows = np.zeros((end - begin, 4), np.float32)
        for j, oi in enumerate(order[begin:end]):
            i, d = T.knn(5, int(oi), int(oi) + 1, threads=1)
            rows[j, 0] = oi
            rows[j, 1:] = d[0, 1:4]
        out_rows[: end - begin] = torch.from_numpy(rows)

    gathered = sharding.sharded_rows(n, compute_range, 4)
    assert gathered.shape[0] >= n
    full = np.zeros((n, 4), np.float32)
    full[order] = gathered.numpy()[:n]                       # un-permute (ccg_unpermute on the GPU)
    i, d = T.knn(5)
    expect = np.concatenate([np.arange(n, dtype=np.float32)[:, None], d[:, 1:4]], 1)
    ok = np.array_equal(full, expect)
    if rank == 0:
        with open(result_path, "w") as f:
            f.write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_unpermute(tmp_path):
    res = tmp_path / "res.txt"
    mp.spawn(_worker, args=(2, _free_port(), 1000, 4, str(res)), nprocs=2, join=True)
    assert res.read_text() == "ok"
