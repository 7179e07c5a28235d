"""Fused query + scatter (ccg_pca_knn_features_scatter): rows of a Morton range written in ORIGINAL order from
inside the query kernel, into own memory (1 GPU) or the collector's memory over NVLink peer mappings (>= 2 GPUs)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import cloudclean_b200 as cc
from cloudclean_b200 import sharding, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_scatter_ranges_fill_the_full_result():
    import torch
    pts = np.ascontiguousarray(synth.make_scan_np(150, 110, scene="room", seed=9)["points"])
    n = pts.shape[0]
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        eig, nrm = ctx.pca_knn(20, cc.COV_NM1, device_out=True)
        feat = ctx.features(eig)
        rows = sharding.PeerRows(ctx, n, (3, 4, 7))
        for b_ in rows.bufs:
            b_.tensor("cuda:0").fill_(-5.0)
        torch.cuda.synchronize()      # torch's stream and the ctx's own (non-blocking) stream are not ordered
        world = 3
        for r in range(world):
            b, e, _ = sharding.shard_range(n, r, world)
            ctx.pca_knn_features_scatter(20, cc.COV_NM1, 0, b, e, *rows.bufs)
        got = [b_.tensor("cuda:0").clone() for b_ in rows.bufs]
        rows.close()
    for a, g in zip((eig, nrm, feat), got):
        assert torch.equal(torch.as_tensor(a).view(torch.int32), g.view(torch.int32))


def test_scatter_rejects_host_outputs():
    pts = np.ascontiguousarray(synth.make_scan_np(32, 32, scene="room", seed=2)["points"])
    n = pts.shape[0]
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        with pytest.raises(cc.CcgError):
            ctx.pca_knn_features_scatter(8, cc.COV_NM1, 0, 0, n, np.empty((n, 3), np.float32), None, None)


def test_scatter_over_nvlink_peer_memory_two_ranks():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "tools", "scatter_check.py")], capture_output=True, text=True, timeout=600)
    print(r.stdout[-2000:], r.stderr[-2000:])
    assert r.returncode == 0
    assert "bit-identical=True" in r.stdout


@pytest.mark.parametrize("world,chunk", [(2, 32), (3, 256), (8, 8192), (5, 64)])
def test_interleaved_shards_fill_the_full_result(world, chunk):
    import torch
    pts = np.ascontiguousarray(synth.make_scan_np(150, 110, scene="room", seed=9)["points"])
    n = pts.shape[0]
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        eig, nrm = ctx.pca_knn(20, cc.COV_NM1, device_out=True)
        feat = ctx.features(eig)
        rows = sharding.PeerRows(ctx, n, (3, 4, 7))
        for b_ in rows.bufs:
            b_.tensor("cuda:0").fill_(-5.0)
        torch.cuda.synchronize()      # torch's stream and the ctx's own (non-blocking) stream are not ordered
        for r in range(world):
            ctx.pca_knn_features_shard(20, cc.COV_NM1, 0, r, world, *rows.bufs, chunk_points=chunk)
        got = [b_.tensor("cuda:0").clone() for b_ in rows.bufs]
        rows.close()
    for a, g in zip((eig, nrm, feat), got):
        assert torch.equal(torch.as_tensor(a).view(torch.int32), g.view(torch.int32))


@pytest.mark.parametrize("world,chunk", [(2, 32), (3, 256), (8, 8192)])
def test_owner_routed_rows_fill_every_ranks_slice(world, chunk):
    """ccg_pca_knn_features_route, `world` ranks emulated on one GPU: every emulated rank answers its interleaved chunks and
    stores each row into the slice of the rank owning its ORIGINAL index; the concatenated slices are the full result."""
    import torch
    pts = np.ascontiguousarray(synth.make_scan_np(150, 110, scene="room", seed=9)["points"])
    n = pts.shape[0]
    rpr = (n + world - 1) // world
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        eig, nrm = ctx.pca_knn(20, cc.COV_NM1, device_out=True)
        feat = ctx.features(eig)
        slices = [[torch.full((rpr, w), -5.0, device="cuda:0") for _ in range(world)] for w in (3, 4, 7)]
        torch.cuda.synchronize()
        for r in range(world):
            ctx.pca_knn_features_route(20, cc.COV_NM1, 0, r, world, rpr, slices[0], slices[1], slices[2], chunk_points=chunk)
        got = [torch.cat(s_)[:n] for s_ in slices]
        with pytest.raises(cc.CcgError):                               # slices that do not cover the cloud are refused
            ctx.pca_knn_features_route(20, cc.COV_NM1, 0, 0, world, rpr - 1 if world > 1 else 0, slices[0], slices[1], slices[2])
    for a, g in zip((eig, nrm, feat), got):
        assert torch.equal(torch.as_tensor(a).view(torch.int32), g.view(torch.int32))


def test_async_mode_enqueues_and_sync_collects():
    """ccg_set_async: build and query are only enqueued (device buffers), ccg_sync makes results and deferred errors valid."""
    import torch
    pts = torch.from_numpy(np.ascontiguousarray(synth.make_scan_np(200, 150, scene="room", seed=14)["points"])).cuda()
    n = pts.shape[0]
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        eig, nrm = ctx.pca_knn(20, cc.COV_NM1, device_out=True)
        feat = ctx.features(eig)
        e2 = torch.empty((n, 3), device="cuda"); n2 = torch.empty((n, 4), device="cuda"); f2 = torch.empty((n, 7), device="cuda")
        torch.cuda.synchronize()
        ctx.set_async(True)
        for _ in range(3):                                   # three steps back to back without a host synchronisation
            ctx.set_cloud(pts)
            ctx.pca_knn_features_into(20, cc.COV_NM1, 0, e2, n2, f2)
        ctx.sync()
        st = ctx.stats()
        assert st["ms_build"] > 0 and st["ms_query"] > 0
        for a, g in zip((eig, nrm, feat), (e2, n2, f2)):
            assert torch.equal(torch.as_tensor(a).view(torch.int32), g.view(torch.int32))
        bad = pts.clone(); bad[17, 1] = float("nan")
        torch.cuda.synchronize()
        ctx.set_cloud(bad)                                   # accepted now ...
        with pytest.raises(cc.CcgError) as e:
            ctx.sync()                                       # ... refused here
        assert e.value.code == -4
        ctx.set_async(False)
        with pytest.raises(cc.CcgError):
            ctx.knn(4)                                       # no cloud after the refused one
        ctx.set_cloud(pts)
        idx, _ = ctx.knn(4)
        assert idx.shape == (n, 4)
