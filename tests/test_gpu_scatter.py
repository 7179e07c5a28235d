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
