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


def _route_worker(rank, world, port, n, seed, result_path):
    """The end-to-end N-rank flow of bench.py on CPU: one shared host copy of the scan and of the outputs (SharedPinned),
    every rank takes ITS 1/N of the rows, the ranks all-gather the cloud, every rank answers its interleaved chunks of the
    common order and delivers each row to the rank that OWNS its original index (an all-to-all stands in for the NVLink peer
    stores of ccg_pca_knn_features_route), and every rank writes its own contiguous slice of the shared output."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    tag = f"ccg_test_{port}"
    rpr = (n + world - 1) // world
    if rank == 0:
        src = sharding.SharedPinned(f"{tag}_pts", (n, 8), create=True)
        out = sharding.SharedPinned(f"{tag}_out", (n, 4), create=True)
        src.t.copy_(torch.from_numpy(np.random.default_rng(seed).normal(size=(n, 8)).astype(np.float32)))
        out.t.fill_(-1.0)
    dist.barrier()
    if rank != 0:
        src = sharding.SharedPinned(f"{tag}_pts", (n, 8))
        out = sharding.SharedPinned(f"{tag}_out", (n, 4))
    a, b = rank * rpr, min(n, (rank + 1) * rpr)
    mine = torch.zeros((rpr, 4))
    mine[: b - a] = src.t[a:b, :4]                                   # this rank's "upload"
    full = torch.empty((world * rpr, 4))
    dist.all_gather_into_tensor(full, mine)
    cloud = np.ascontiguousarray(full[:n].numpy())
    cloud8 = np.zeros((n, 8), np.float32); cloud8[:, :4] = cloud
    order = np.lexsort((cloud[:, 2], cloud[:, 1], cloud[:, 0])).astype(np.int64)
    T = oracle.Tree(cloud8)
    chunk = 64
    sends = [[] for _ in range(world)]
    for c0 in range(rank * chunk, n, world * chunk):                 # interleaved chunks of the common order
        for oi in order[c0:c0 + chunk]:
            _, d = T.knn(4, int(oi), int(oi) + 1, threads=1)
            sends[int(oi) // rpr].append([float(oi), d[0, 1], d[0, 2], d[0, 3]])
    send_t = [torch.tensor(s_, dtype=torch.float32).reshape(-1, 4) for s_ in sends]
    counts = torch.tensor([t.shape[0] for t in send_t], dtype=torch.int64)
    all_counts = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(all_counts, counts)
    recv_t = [torch.empty((int(all_counts[r][rank]), 4)) for r in range(world)]
    dist.all_to_all(recv_t, send_t) if dist.get_backend() != "gloo" else None
    if dist.get_backend() == "gloo":                                 # gloo has no all_to_all: pairwise exchange
        for r in range(world):
            if r == rank:
                recv_t[r] = send_t[r]
        reqs = [dist.isend(send_t[r], r) for r in range(world) if r != rank] + [dist.irecv(recv_t[r], r) for r in range(world) if r != rank]
        for q in reqs:
            q.wait()
    rows = torch.cat(recv_t)
    idx = rows[:, 0].long()
    assert ((idx >= a) & (idx < b)).all() and idx.numel() == b - a   # exactly the rows this rank owns
    out.t[idx] = rows                                                # "download" of the own slice to its final host rows
    dist.barrier()
    if rank == 0:
        _, d = T.knn(4)
        expect = np.concatenate([np.arange(n, dtype=np.float32)[:, None], d[:, 1:4]], 1)
        ok = np.array_equal(out.t.numpy(), expect)
        with open(result_path, "w") as f:
            f.write("ok" if ok else "mismatch")
    dist.barrier()
    src.close(); out.close()
    dist.destroy_process_group()


def test_two_rank_shared_host_and_owner_routing(tmp_path):
    res = tmp_path / "res2.txt"
    mp.spawn(_route_worker, args=(2, _free_port(), 777, 9, str(res)), nprocs=2, join=True)
    assert res.read_text() == "ok"
