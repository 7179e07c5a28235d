"""GPU parity: every C-ABI entry point of libccgpu.so against the CPU oracle and the committed golden vectors.

Bars (BASELINE.json north_star): neighbour index sets / distances / counts / outlier masks / stage-(ii) grid
normals BIT-EXACT; eigenvalues |d lambda_i| <= 1e-4 * lambda_0 against the fp64 truth; normals |cos| >= 1 - 1e-5
wherever the normal is defined ((l1-l2)/l0 above GAP_MIN); curvature within 1e-5 absolute."""
import ctypes
import os

import numpy as np
import pytest

import cloudclean_b200 as cc
import oracle
from cloudclean_b200 import synth
from conftest import clouds_small

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EIG_TOL = 1e-4          # relative to lambda_0
COS_TOL = 1e-5
GAP_MIN = 1e-3          # below this (l1-l2)/l0 the smallest eigenvector is ill-conditioned in fp32: normal not compared
SMALL = clouds_small()


def gold(name):
    return dict(np.load(os.path.join(GOLD, name + ".npz")))


def check_pca(eig, nrm, ref, check_normals=True):
    truth = ref["eig64"]
    scale = np.maximum(truth[:, :1], 1e-30)
    marker = (truth[:, 0] == 0) & (truth[:, 2] == 1)
    err = np.abs(eig - truth) / scale
    assert err[~marker].max(initial=0) <= EIG_TOL
    assert np.array_equal(eig[marker], truth[marker].astype(np.float32))
    if check_normals and nrm is not None:
        ok = (ref["gap64"] > GAP_MIN) & ~marker
        cos = np.abs((nrm[:, :3].astype(np.float64) * ref["nrm64"][:, :3]).sum(1))
        assert (1 - cos[ok]).max(initial=0) <= COS_TOL
        assert np.abs(nrm[ok, 3] - ref["nrm64"][ok, 3]).max(initial=0) <= 1e-5
        assert np.isnan(nrm[marker]).all()
        unit = np.linalg.norm(nrm[~marker, :3].astype(np.float64), axis=1)
        assert np.abs(unit - 1).max(initial=0) <= 1e-5


# ---- a1 kNN --------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["plane4x5", "room40x36"])
def test_knn_golden(gpu_ctx, name):
    g = gold(name)
    gpu_ctx.set_cloud(g["points"])
    for k in (1, 8, 20, 50):
        idx, d2 = gpu_ctx.knn(k)
        assert np.array_equal(idx, g[f"knn{k}_idx"])
        assert np.array_equal(d2, g[f"knn{k}_d2"])


@pytest.mark.parametrize("stride_floats", [3, 4, 8, 12])
def test_knn_strides_and_device_pointers(gpu_ctx, stride_floats):
    import torch
    xyz = SMALL["lattice_ties"][:, :3]
    pts = np.zeros((xyz.shape[0], stride_floats), np.float32)
    pts[:, :3] = xyz
    ref_i, ref_d = oracle.Tree(pts).knn(20)
    gpu_ctx.set_cloud(pts)
    idx, d2 = gpu_ctx.knn(20)
    assert np.array_equal(idx, ref_i) and np.array_equal(d2, ref_d)
    dev = torch.from_numpy(pts).cuda()
    gpu_ctx.set_cloud(dev)
    idx, d2 = gpu_ctx.knn(20)
    assert idx.is_cuda and np.array_equal(idx.cpu().numpy(), ref_i) and np.array_equal(d2.cpu().numpy(), ref_d)


def test_order_is_a_permutation(gpu_ctx):
    pts = SMALL["uniform"]
    gpu_ctx.set_cloud(pts)
    o = gpu_ctx.order()
    assert np.array_equal(np.sort(o), np.arange(pts.shape[0], dtype=np.int32))


# ---- a2 / a3 radius ------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["uniform", "lattice_ties", "coplanar", "dups_only", "n2", "clusters_outliers"])
@pytest.mark.parametrize("r,cap", [(0.05, 8), (0.25, 20), (0.125, 64)])
@pytest.mark.parametrize("strict", [True, False])
def test_radius_lists(gpu_ctx, name, r, cap, strict):
    pts = SMALL[name]
    T = oracle.Tree(pts)
    ref = T.radius(r, cap=cap, strict_lt=strict)
    gpu_ctx.set_cloud(pts)
    got = gpu_ctx.radius(r, max_nn=cap, strict_lt=strict)
    assert np.array_equal(got["count"], ref["count"])
    assert np.array_equal(got["idx"], ref["idx"])
    assert np.array_equal(got["d2"], ref["d2"])


@pytest.mark.parametrize("name", ["uniform", "lattice_ties", "clusters_outliers", "n1"])
@pytest.mark.parametrize("strict", [True, False])
def test_radius_counts(gpu_ctx, name, strict):
    pts = SMALL[name]
    T = oracle.Tree(pts)
    gpu_ctx.set_cloud(pts)
    for r in (0.0, 0.125, 0.5):            # 0.125 and 0.5 are lattice distances: the strict/inclusive edge is hit
        total = T.radius(r, strict_lt=strict, lists=False)["total"]
        got = gpu_ctx.radius(r, max_nn=0, strict_lt=strict, lists=False)["count"]
        assert np.array_equal(got, total)
        got = gpu_ctx.radius(r, max_nn=7, strict_lt=strict, lists=False)["count"]
        assert np.array_equal(got, np.minimum(total, 7))


@pytest.mark.parametrize("name", ["plane4x5", "room40x36"])
def test_radius_golden(gpu_ctx, name):
    g = gold(name)
    gpu_ctx.set_cloud(g["points"])
    for r, cap in ((0.3, 20), (0.12, 8)):
        tag = f"r{int(r * 100):03d}_c{cap}"
        got = gpu_ctx.radius(r, max_nn=cap, strict_lt=True)
        assert np.array_equal(got["idx"], g[f"rad_{tag}_idx"])
        assert np.array_equal(got["d2"], g[f"rad_{tag}_d2"])
        assert np.array_equal(got["count"], g[f"rad_{tag}_count"])
        assert np.array_equal(gpu_ctx.radius(r, lists=False)["count"], g[f"rad_{tag}_total"])
        assert np.array_equal(gpu_ctx.radius(r, strict_lt=False, lists=False)["count"], g[f"rad_{tag}_total_incl"])


# ---- a12 kNN + PCA -------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", list(SMALL.keys()))
@pytest.mark.parametrize("k,cov", [(20, cc.COV_NM1), (10, cc.COV_N), (50, cc.COV_SCATTER)])
def test_pca_knn(gpu_ctx, name, k, cov):
    pts = SMALL[name]
    gpu_ctx.set_cloud(pts)
    eig, nrm = gpu_ctx.pca_knn(k, cov)
    ri, _ = oracle.Tree(pts).knn(k)
    ref = oracle.pca_lists(pts, ri, cov_norm=cov)
    check_pca(eig, nrm, ref)
    assert (eig[:, 0] >= eig[:, 1]).all() and ((eig[:, 1] >= eig[:, 2]) | (eig[:, 2] == 1)).all()


def test_pca_knn_golden_and_viewpoint(gpu_ctx):
    g = gold("room40x36")
    pts = g["points"]
    gpu_ctx.set_cloud(pts)
    for cn in (0, 1, 2):
        eig, nrm = gpu_ctx.pca_knn(20, cn)
        truth = g[f"pca20_cov{cn}_eig"]
        assert (np.abs(eig - truth) / truth[:, :1]).max() <= EIG_TOL
        gap = (truth[:, 1] - truth[:, 2]) / truth[:, 0]
        cos = np.abs((nrm[:, :3] * g[f"pca20_cov{cn}_nrm"][:, :3]).sum(1))
        assert (1 - cos[gap > GAP_MIN]).max() <= COS_TOL
    vp = np.array([0.0, 0.0, 0.0], np.float32)            # the scanner: every flipped normal faces it
    eig, nrm = gpu_ctx.pca_knn(20, cc.COV_NM1, viewpoint=vp)
    assert ((nrm[:, :3] * (vp[None, :] - pts[:, :3])).sum(1) >= 0).all()


def test_pca_knn_fused_features_equal_standalone(gpu_ctx):
    pts = SMALL["uniform"]
    n = pts.shape[0]
    gpu_ctx.set_cloud(pts)
    eig, nrm = gpu_ctx.pca_knn(20, cc.COV_NM1)
    for fs in (0, 1):
        e2, n2, f2 = np.empty((n, 3), np.float32), np.empty((n, 4), np.float32), np.empty((n, 7), np.float32)
        gpu_ctx.pca_knn_features_into(20, cc.COV_NM1, fs, e2, n2, f2)
        assert np.array_equal(e2, eig) and np.array_equal(n2, nrm)
        assert np.array_equal(f2, gpu_ctx.features(eig, fs), equal_nan=True)


# ---- a4 / a5 / a6 radius + PCA (getPCA / getHist) -------------------------------------------------------
@pytest.mark.parametrize("name", ["uniform", "coplanar", "clusters_outliers", "n2", "lattice_ties"])
@pytest.mark.parametrize("r,max_nn", [(0.2, 20), (0.1, 64), (0.15, 0), (0.2, 1000), (0.3, 100), (0.5, 700)])
def test_pca_radius_getpca(gpu_ctx, name, r, max_nn):
    pts = SMALL[name]
    n = pts.shape[0]
    T = oracle.Tree(pts)
    total = T.radius(r, strict_lt=True, lists=False)["total"]
    gpu_ctx.set_cloud(pts)
    cap = n if (max_nn == 0 or max_nn >= n) else max_nn          # FLANN caps only when 0 < max_nn < n
    lists = T.radius(r, cap=cap, strict_lt=True)
    ref = oracle.pca_lists(pts, lists["idx"], count=lists["count"], cov_norm=cc.COV_NM1)
    eig, nrm, count = gpu_ctx.pca_radius(r, max_nn, cc.COV_NM1, want_normals=True)
    assert np.array_equal(count, lists["count"])
    check_pca(eig, nrm, ref)
    # getHist (cv.cpp:293): the same eigenvalues L2-normalised
    hist, _, _ = gpu_ctx.pca_radius(r, max_nn, cc.COV_NM1, normalise_hist=True)
    t = ref["eig64"] / np.maximum(np.linalg.norm(ref["eig64"], axis=1, keepdims=True), 1e-300)
    assert np.abs(hist - t).max() <= 2e-4


def test_pca_radius_golden(gpu_ctx):
    g = gold("room40x36")
    gpu_ctx.set_cloud(g["points"])
    for r, cap in ((0.3, 20), (0.12, 8)):
        tag = f"r{int(r * 100):03d}_c{cap}"
        eig, _, count = gpu_ctx.pca_radius(r, cap, cc.COV_NM1)
        assert np.array_equal(count, g[f"rad_{tag}_count"])
        truth = g[f"pcarad_{tag}_eig"]
        marker = count < 3
        assert np.array_equal(eig[marker], truth[marker].astype(np.float32))
        assert (np.abs(eig - truth)[~marker] / truth[~marker, :1]).max() <= EIG_TOL


# ---- a9 features ---------------------------------------------------------------------------------------
def test_features_and_scalars(gpu_ctx):
    pts = SMALL["uniform"]
    gpu_ctx.set_cloud(pts)
    eig, _, _ = gpu_ctx.pca_radius(0.12, 20)              # includes (0,0,1) markers
    assert ((eig[:, 0] == 0) & (eig[:, 2] == 1)).any()
    for fs in (0, 1):
        f = gpu_ctx.features(eig, fs)
        ref = oracle.features(eig, fs)
        assert np.array_equal(f[:, :4], ref[:, :4], equal_nan=True)          # +,-,/ only: bit-exact
        assert np.allclose(f, ref, rtol=4e-6, atol=1e-7, equal_nan=True)     # cbrtf / logf: libm vs CUDA, few ulp
    assert np.array_equal(gpu_ctx.eigen_scalar(eig, 0), oracle.eigen_ratio(eig))
    assert np.array_equal(gpu_ctx.eigen_scalar(eig, 1), oracle.plane_similarity(eig))
    g = gold("room40x36")
    e32 = g["pca20_cov1_eig"].astype(np.float32)
    for fs in (0, 1):
        assert np.allclose(gpu_ctx.features(e32, fs), g[f"feat_set{fs}"], rtol=4e-6, atol=1e-7, equal_nan=True)


# ---- a11 SOR -------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", list(SMALL.keys()))
@pytest.mark.parametrize("k,alpha", [(8, 1.0), (30, 1.0), (63, 0.0)])
def test_sor_mask(gpu_ctx, name, k, alpha):
    pts = SMALL[name]
    gpu_ctx.set_cloud(pts)
    mask, md = gpu_ctx.sor_mask(k, alpha)
    rmask, rmd, _, _ = oracle.Tree(pts).sor(k, alpha)
    assert np.array_equal(md, rmd)
    assert np.array_equal(mask, rmask)


def test_sor_golden(gpu_ctx):
    for name in ("plane4x5", "room40x36"):
        g = gold(name)
        gpu_ctx.set_cloud(g["points"])
        for k, a in ((8, 1.0), (30, 0.5)):
            mask, md = gpu_ctx.sor_mask(k, a)
            assert np.array_equal(md, g[f"sor{k}_mean_d"]) and np.array_equal(mask, g[f"sor{k}_mask"])


# ---- a10 radius / min-neighbour filter -------------------------------------------------------------------
@pytest.mark.parametrize("name", ["uniform", "lattice_ties", "clusters_outliers", "n33"])
@pytest.mark.parametrize("r,mn", [(0.1, 10), (0.5, 10), (1.0, 3)])
@pytest.mark.parametrize("prefilter", [False, True])
def test_radius_count_mask(gpu_ctx, name, r, mn, prefilter):
    pts = SMALL[name]
    gpu_ctx.set_cloud(pts)
    got = gpu_ctx.radius_count_mask(r, mn, prefilter)
    ref = oracle.Tree(pts).radius_outliers(r, mn, prefilter)
    assert np.array_equal(got, ref)


# ---- a7 grid normals -------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["plane4x5", "room40x36"])
def test_grid_normals_golden(gpu_ctx, name):
    g = gold(name)
    w, h = int(g["scan_w"]), int(g["scan_h"])
    gpu_ctx.set_cloud(g["points"])
    s2, miss2 = gpu_ctx.grid_normals(g["cloud_to_grid"], w, h, fallback=False)
    assert np.array_equal(s2, g["gridn_stage2"], equal_nan=True) and miss2 == len(g["gridn_missing"])
    s3, miss3 = gpu_ctx.grid_normals(g["cloud_to_grid"], w, h, fallback=True)
    assert np.array_equal(s3, g["gridn_final"], equal_nan=True) and miss3 == int(g["gridn_still"])


@pytest.mark.parametrize("w,h,dropout,seed", [(64, 48, 0.05, 1), (96, 64, 0.45, 2), (33, 7, 0.6, 3)])
def test_grid_normals_scan(gpu_ctx, w, h, dropout, seed):
    s = synth.make_scan_np(w, h, scene="room", seed=seed, dropout=dropout)
    pts, c2g = s["points"], s["cloud_to_grid"]
    gpu_ctx.set_cloud(pts)
    for fb in (False, True):
        got, miss = gpu_ctx.grid_normals(c2g, w, h, fallback=fb)
        ref, rmiss = oracle.grid_normals(pts, c2g, w, h, fallback=fb)
        assert np.array_equal(got, ref, equal_nan=True)
        assert miss == rmiss
    with pytest.raises(cc.CcgError):
        gpu_ctx.grid_normals(c2g + w * h, w, h)            # grid index outside the scan


# ---- sharding entry points ---------------------------------------------------------------------------------
def test_ranges_and_unpermute_reproduce_the_full_call(gpu_ctx):
    from cloudclean_b200 import sharding
    pts = SMALL["uniform"]
    n = pts.shape[0]
    gpu_ctx.set_cloud(pts)
    eig, nrm = gpu_ctx.pca_knn(20, cc.COV_NM1)
    feat = gpu_ctx.features(eig, 0)
    _, md = gpu_ctx.sor_mask(30, 1.0)
    world = 3
    per = sharding.shard_range(n, 0, world)[2]
    g_e, g_n = np.zeros((world * per, 3), np.float32), np.zeros((world * per, 4), np.float32)
    g_f, g_m = np.zeros((world * per, 7), np.float32), np.zeros((world * per, 1), np.float32)
    for r in range(world):
        b, e, _ = sharding.shard_range(n, r, world)
        se, sn, sf = np.zeros((per, 3), np.float32), np.zeros((per, 4), np.float32), np.zeros((per, 7), np.float32)
        sm = np.zeros((per, 1), np.float32)
        gpu_ctx.pca_knn_features_range(20, cc.COV_NM1, 0, b, e, se, sn, sf)
        gpu_ctx.sor_mean_range(30, b, e, sm)
        g_e[r * per:(r + 1) * per], g_n[r * per:(r + 1) * per], g_f[r * per:(r + 1) * per] = se, sn, sf
        g_m[r * per:(r + 1) * per] = sm
    for gathered, full in ((g_e, eig), (g_n, nrm), (g_f, feat), (g_m, md[:, None])):
        out = np.empty_like(full)
        gpu_ctx.unpermute(gathered, out)
        assert np.array_equal(out, full, equal_nan=True)
    order = gpu_ctx.order()
    assert np.array_equal(g_e[:n], eig[order])


# ---- error behaviour ---------------------------------------------------------------------------------------
def test_errors_and_empty_cloud():
    ctx = cc.Context(0)
    with pytest.raises(cc.CcgError) as e:
        ctx.n = 4
        ctx.on_device = False
        ctx.knn(3)
    assert e.value.code == -3                                 # CCG_ERR_NO_CLOUD
    bad = SMALL["n33"].copy()
    bad[5, 1] = np.nan
    with pytest.raises(cc.CcgError) as e:
        ctx.set_cloud(bad)
    assert e.value.code == -4                                 # CCG_ERR_NONFINITE (PCL's kd-tree rejects these too)
    ctx.set_cloud(SMALL["n33"])
    for k in (0, 65):
        with pytest.raises(cc.CcgError) as e:
            ctx.knn(k)
        assert e.value.code == -5
    with pytest.raises(cc.CcgError) as e:
        ctx.radius(-1.0, 5)
    assert e.value.code == -1
    with pytest.raises(cc.CcgError) as e:
        ctx.pca_knn(20, 7)
    assert e.value.code == -1
    flag = ctypes.c_int(1)
    ctx.L.ccg_set_cancel_flag(ctx.h, ctypes.byref(flag))
    with pytest.raises(cc.CcgError) as e:
        ctx.knn(5)
    assert e.value.code == -6                                 # honours PointCloud::deleting_
    flag.value = 0
    idx, _ = ctx.knn(5)
    assert idx.shape == (33, 5)
    ctx.L.ccg_set_cancel_flag(ctx.h, None)
    ctx.set_cloud(np.zeros((0, 8), np.float32))               # empty cloud: every call is a no-op
    idx, d2 = ctx.knn(4)
    assert idx.shape == (0, 4)
    mask, md = ctx.sor_mask(5, 1.0)
    assert mask.shape == (0,)
    ctx.close()


# ---- host-output pipelining (chunked D2H behind the query) ------------------------------------------------
@pytest.mark.parametrize("shuffle", [False, True])
@pytest.mark.parametrize("chunks", [2, 5, 16])
def test_pipelined_host_outputs_are_identical(monkeypatch, shuffle, chunks):
    """ccg_pca_knn_features / ccg_pca_radius with HOST outputs ship rows in chunks while later tiles are computed
    (tiles ordered by their smallest original index).  Every row must equal the unpipelined call bit for bit —
    for a scan in file order (row front follows the tile front) and for a shuffled cloud (all rows in the last copy)."""
    scan = synth.make_scan_np(160, 120, scene="room", seed=21)
    pts = np.ascontiguousarray(scan["points"])
    if shuffle:
        pts = np.ascontiguousarray(pts[np.random.default_rng(5).permutation(pts.shape[0])])
    n = pts.shape[0]

    def run():
        with cc.Context(0) as ctx:
            ctx.set_cloud(pts)
            eig = np.full((n, 3), -7.0, np.float32); nrm = np.full((n, 4), -7.0, np.float32); feat = np.full((n, 7), -7.0, np.float32)
            ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat)
            e2, _, cnt = ctx.pca_radius(0.08, 20, device_out=False)
            mask, md = ctx.sor_mask(12, 1.0, device_out=False)        # host sum of the mean distances rides the same chunks
            return eig, nrm, feat, e2, cnt, mask, md

    monkeypatch.setenv("CCG_PIPELINE_CHUNKS", "1")
    ref = run()
    monkeypatch.setenv("CCG_PIPELINE_CHUNKS", str(chunks))
    monkeypatch.setenv("CCG_PIPELINE_MIN_POINTS", "0")
    got = run()
    for a, b in zip(ref, got):
        assert np.array_equal(a, b, equal_nan=True)
    # and the reference run itself is right
    T = oracle.Tree(pts)
    ri, _ = T.knn(20)
    check_pca(ref[0], ref[1], oracle.pca_lists(pts, ri, cov_norm=1))
    rmask, rmd, _, _ = T.sor(12, 1.0)
    assert np.array_equal(got[5], rmask) and np.array_equal(got[6], rmd)


def test_cancel_inside_a_pipelined_call_leaves_nothing_in_flight(monkeypatch):
    """The cancel flag (PointCloud::deleting_, normalestimation.cpp:305-306) raised from another thread while a
    host-output call is shipping chunks: the call ends with CCG_OK or CCG_ERR_CANCELLED, no copy is still running when it
    returns (the output buffers are dropped right away), and the context answers the next call correctly."""
    import threading
    monkeypatch.setenv("CCG_PIPELINE_CHUNKS", "8")
    monkeypatch.setenv("CCG_PIPELINE_MIN_POINTS", "0")
    scan = synth.make_scan_np(640, 480, scene="room", seed=22)
    pts = np.ascontiguousarray(scan["points"])
    n = pts.shape[0]
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        flag = ctypes.c_int(0)
        ctx.L.ccg_set_cancel_flag(ctx.h, ctypes.byref(flag))
        seen = set()
        for delay in (0.0, 0.0005, 0.002, 0.005):
            eig = np.empty((n, 3), np.float32); nrm = np.empty((n, 4), np.float32); feat = np.empty((n, 7), np.float32)
            flag.value = 0
            t = threading.Timer(delay, lambda: setattr(flag, "value", 1))
            t.start()
            try:
                ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat)
                seen.add(0)
            except cc.CcgError as e:
                assert e.code == -6
                seen.add(-6)
            t.join()
            del eig, nrm, feat                                    # a caller frees its buffers as soon as the call is back
        flag.value = 0
        ctx.L.ccg_set_cancel_flag(ctx.h, None)
        idx, _ = ctx.knn(8)
        ri, _ = oracle.Tree(pts).knn(8)
        assert np.array_equal(idx, ri)
        assert seen                                               # at least one of the two outcomes happened, none crashed


# ---- f1: voxel downsample (octreeDownsample, utils.h:16-62) -------------------------------------------------
@pytest.mark.parametrize("res", [0.02, 0.1, 0.5])
@pytest.mark.parametrize("as_written", [True, False])
@pytest.mark.parametrize("device", [False, True])
def test_voxel_downsample(gpu_ctx, res, as_written, device):
    """Voxel membership (big_to_small, voxel count, numbering) bit-exact; the per-voxel means within 1e-5 relative of
    the oracle's sequential fp32 sums (the GPU sums lane-strided partials, so the fp32 rounding order differs)."""
    scan = synth.make_scan_np(120, 90, scene="room", seed=4)
    pts = np.ascontiguousarray(scan["points"])
    want, wmap = oracle.voxel_downsample(pts, res, as_written=as_written)
    if device:
        import torch
        dpts = torch.as_tensor(pts, device="cuda")
        gpu_ctx.set_cloud(dpts)
        got, gmap = gpu_ctx.voxel_downsample(dpts, res, as_written=as_written)
        got, gmap = got.cpu().numpy(), gmap.cpu().numpy()
    else:
        gpu_ctx.set_cloud(pts)
        got, gmap = gpu_ctx.voxel_downsample(pts, res, as_written=as_written, device_out=False)
    assert got.shape == want.shape
    assert np.array_equal(gmap, wmap)
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    counts = np.bincount(wmap, minlength=want.shape[0])
    w3 = (counts + (1 if as_written else 0)) / counts * pts[0, 3]
    assert np.allclose(got[:, 3], w3, rtol=1e-5)


def test_voxel_downsample_xyz_only_and_errors(gpu_ctx):
    rng = np.random.default_rng(8)
    pts = rng.random((5000, 3)).astype(np.float32)
    gpu_ctx.set_cloud(pts)
    got, gmap = gpu_ctx.voxel_downsample(pts, 0.2, as_written=False, device_out=False)
    want, wmap = oracle.voxel_downsample(pts, 0.2, as_written=False)
    assert np.array_equal(gmap, wmap) and np.allclose(got, want, rtol=1e-5, atol=1e-6)
    with pytest.raises(cc.CcgError):
        gpu_ctx.voxel_downsample(pts, 0.0, device_out=False)
    with pytest.raises(cc.CcgError):
        gpu_ctx.voxel_downsample(pts, 0.1, n_floats=4, device_out=False)      # 4 floats do not fit a 12-byte stride


# ---- f3: kNN graph with MinCut binary potentials (graphcut/mincut.cpp:458-475, 574-585) ---------------------
@pytest.mark.parametrize("name", ["uniform", "lattice_ties", "clusters_outliers", "n2"])
@pytest.mark.parametrize("k,inv_sigma", [(14, 4.0), (8, 250.0)])
def test_knn_graph(gpu_ctx, name, k, inv_sigma):
    """Neighbour table bit-exact (same contract as ccg_knn); edge weights within 4 ulp of the numpy restatement
    (CUDA's double exp is within 1 ulp, numpy/glibc's is correctly rounded)."""
    pts = SMALL[name]
    gpu_ctx.set_cloud(pts)
    idx, w = gpu_ctx.knn_graph(k, inv_sigma, device_out=False)
    ri, _ = oracle.Tree(pts).knn(k)
    assert np.array_equal(idx, ri)
    want = oracle.binary_potential(pts, ri, inv_sigma)
    assert np.allclose(w, want, rtol=1e-15 * 4, atol=0)
    assert np.all(w[idx[:, 0] >= 0, 0] == 1.0)          # self edge


# ---- f2: radius-neighbourhood reductions (cv.h:123-206, cv.cpp:166-230) --------------------------------------
def _scan_with_normals():
    scan = synth.make_scan_np(100, 80, scene="room", seed=6)
    pts = np.ascontiguousarray(scan["points"])
    rng = np.random.default_rng(12)
    nrm = np.zeros((pts.shape[0], 8), np.float32)                 # pcl::Normal layout, stride 32
    v = rng.normal(size=(pts.shape[0], 3)).astype(np.float32)
    v[:, 2] += 3.0                                                # mostly "up", with spread
    nrm[:, :3] = v / np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    nrm[::97, :3] = np.nan                                        # the NaN islands the reference skips (cv.h:161-162)
    return pts, nrm


@pytest.mark.parametrize("r,max_nn", [(0.1, 20), (0.3, 64), (0.02, 8)])
def test_normal_stdev(gpu_ctx, r, max_nn):
    """Same neighbour lists as the oracle (bit-exact search); the reduction is fp32 as written — acosf differs by an ulp
    between CUDA and glibc and sqrt(a - b) amplifies that near 0, so the VARIANCE is compared (1e-6 absolute)."""
    pts, nrm = _scan_with_normals()
    gpu_ctx.set_cloud(pts)
    got = gpu_ctx.normal_stdev(r, max_nn, nrm, device_out=False)
    want = oracle.normal_stdev(pts, nrm, r, max_nn)
    assert got.shape == want.shape
    assert np.all(got[::97] == 0)                                 # NaN normal -> skipped -> 0
    assert np.abs(got.astype(np.float64) ** 2 - want.astype(np.float64) ** 2).max() <= 1e-6
    if r >= 0.3:
        assert (want > 0).sum() > want.size // 4


@pytest.mark.parametrize("r,max_nn", [(0.1, 20), (0.25, 64)])
@pytest.mark.parametrize("use_depth", [False, True])
def test_stdev_dist(gpu_ctx, r, max_nn, use_depth):
    pts, _ = _scan_with_normals()
    gpu_ctx.set_cloud(pts)
    centre = np.array([0.1, -0.2, 0.3], np.float32)
    got = gpu_ctx.stdev_dist(r, max_nn, use_depth, centre, device_out=False)
    want = oracle.stdev_dist(pts, r, max_nn, use_depth, centre)
    assert np.array_equal(got, want)       # fp32 +,*,/,sqrt in the same order on both sides: identical bits


@pytest.mark.parametrize("r,max_nn", [(0.1, 0), (0.3, 100), (0.3, 0), (0.5, 1000)])
def test_reductions_without_a_list(gpu_ctx, r, max_nn):
    """max_nn == 0 (every point inside r) and caps beyond the list kernels (the reference's max_nn_ is whatever the GUI
    holds: featureeval.cpp:816,999,1032): the neighbour SET is the oracle's (counts and histograms bit-exact), the two fp32
    sums run in traversal order instead of ascending distance and are held to 1e-5 of the value range."""
    pts, nrm = _scan_with_normals()
    gpu_ctx.set_cloud(pts)
    got = gpu_ctx.normal_stdev(r, max_nn, nrm, device_out=False)
    want = oracle.normal_stdev(pts, nrm, r, max_nn)
    assert np.all(got[::97] == 0)
    assert np.abs(got.astype(np.float64) ** 2 - want.astype(np.float64) ** 2).max() <= 1e-5
    centre = np.array([0.1, -0.2, 0.3], np.float32)
    for use_depth in (False, True):
        g = gpu_ctx.stdev_dist(r, max_nn, use_depth, centre, device_out=False)
        w = oracle.stdev_dist(pts, r, max_nn, use_depth, centre)
        # (sum_sq - sum^2/m)/(m-1) in fp32 as written: the two fp32 sums of m terms up to D^2 = (|q - ref| + r)^2 are added in a
        # different order on the two sides; after the division by m-1 the rounding noise is a few 2^-24 D^2 sqrt(m).
        # Stated bound: 64 * 2^-24 * D^2 (1e-5 of r^2 for the query-relative flavour, more when the reference is the far-away sensor)
        ref = centre[None, :].astype(np.float64) if use_depth else pts[:, :3].astype(np.float64)
        D2 = (np.linalg.norm(pts[:, :3].astype(np.float64) - ref, axis=1) + r) ** 2
        tol = 64.0 * 2.0 ** -24 * D2 + 1e-12
        assert (np.abs(g.astype(np.float64) - w.astype(np.float64)) <= tol).all()
    for bins in (20, 70):
        gd = gpu_ctx.neighbour_hist(r, max_nn, bins, device_out=False)
        assert np.array_equal(gd, oracle.neighbour_hist(pts, r, max_nn, bins))
        inten = pts[:, 4]
        gi = gpu_ctx.neighbour_hist(r, max_nn, bins, attr=inten, device_out=False)
        assert np.array_equal(gi, oracle.neighbour_hist(pts, r, max_nn, bins, attr=inten))


@pytest.mark.parametrize("name", ["uniform", "lattice_ties", "clusters_outliers", "n2"])
@pytest.mark.parametrize("r,cap,strict", [(0.2, 0, True), (0.3, 100, True), (0.15, 20, False), (0.5, 700, True)])
def test_radius_csr_lists(gpu_ctx, name, r, cap, strict):
    """ccg_radius_csr: the variable-length rows radiusSearch fills into its vectors, for every cap (0 = all inside r), bit-exact
    in indices AND distances against the oracle's kd-tree search, rows ascending (d2, index)."""
    pts = SMALL[name]
    n = pts.shape[0]
    T = oracle.Tree(pts)
    gpu_ctx.set_cloud(pts)
    off, idx, d2 = gpu_ctx.radius_csr(r, cap, strict_lt=strict, device_out=False)
    total = T.radius(r, strict_lt=strict, lists=False)["total"]
    eff = total if (cap == 0 or cap >= n) else np.minimum(total, cap)          # FLANN caps only when 0 < max_nn < n
    assert off[0] == 0 and np.array_equal(np.diff(off), eff)
    width = int(max(1, eff.max()))
    ref = T.radius(r, cap=width, strict_lt=strict)
    rows = np.repeat(np.arange(n), eff)
    cols = np.arange(off[-1]) - np.repeat(off[:-1], eff)
    assert np.array_equal(idx, ref["idx"][rows, cols])
    assert np.array_equal(d2, ref["d2"][rows, cols])


def test_capped_radius_ties_follow_the_index_order(gpu_ctx):
    """A cap that cuts through a class of equidistant candidates (integer lattice): the admitted ones are those with the
    smallest indices, exactly as the (d2, index) order of the oracle's kd-tree search says."""
    pts = SMALL["lattice_ties"]
    T = oracle.Tree(pts)
    gpu_ctx.set_cloud(pts)
    for r, cap in ((0.35, 70), (0.5, 90), (0.5, 200)):
        lists = T.radius(r, cap=cap, strict_lt=True)
        ref = oracle.pca_lists(pts, lists["idx"], count=lists["count"], cov_norm=cc.COV_NM1)
        eig, nrm, count = gpu_ctx.pca_radius(r, cap, cc.COV_NM1, want_normals=True)
        assert np.array_equal(count, lists["count"])
        check_pca(eig, nrm, ref)


def _radius_pca_truth(pts, r):
    """fp64 brute force over the fp32-contract membership mask d2 < float32(r*r): counts, eigenvalues (desc) of the 1/(m-1) covariance"""
    xyz = pts[:, :3].astype(np.float32)
    n = xyz.shape[0]
    r2 = np.float32(r * r)
    P = xyz.astype(np.float64)
    cnt = np.empty(n, np.int32)
    eig = np.zeros((n, 3))
    for a in range(0, n, 500):
        q = xyz[a:a + 500]
        dx = q[:, None, 0] - xyz[None, :, 0]
        dy = q[:, None, 1] - xyz[None, :, 1]
        dz = q[:, None, 2] - xyz[None, :, 2]
        mask = ((dx * dx + dy * dy) + dz * dz) < r2                     # every product and sum rounded to fp32, as the contract says
        m = mask.sum(1)
        cnt[a:a + 500] = m
        w = mask.astype(np.float64)
        for i in range(q.shape[0]):
            if m[i] >= 3:
                nb = P[mask[i]]
                eig[a + i] = np.linalg.eigvalsh(np.cov(nb.T, ddof=1))[::-1]
            else:
                eig[a + i] = (0.0, 0.0, 1.0)
    return cnt, eig


@pytest.mark.parametrize("kind", ["surface", "volume", "far_from_origin"])
def test_uncapped_radius_sums_take_whole_boxes(kind):
    """Uncapped radius PCA / counts over neighbourhoods of thousands of points: boxes of the hierarchy that lie inside the radius
    contribute their aggregated count and moments (StreamNodes, query.cu) — counts stay bit-exact, eigenvalues within the bar, and
    the result agrees with the per-point walk (CCG_NODE_MOMENTS=0) to the same bar."""
    rng = np.random.default_rng(5)
    n = 12000
    xyz = rng.random((n, 3)) - 0.5
    if kind == "surface":
        xyz[:, 2] = 0.02 * np.sin(6 * xyz[:, 0]) + 0.002 * rng.random(n)
    if kind == "far_from_origin":
        xyz += np.array([250.0, -80.0, 40.0])                            # cancellation test: sums must not be taken about the origin
    pts = np.zeros((n, 4), np.float32)
    pts[:, :3] = xyz
    r = 0.3
    cnt, eig64 = _radius_pca_truth(pts, r)
    T = oracle.Tree(pts)
    assert np.array_equal(cnt, T.radius(r, strict_lt=True, lists=False)["total"])
    mn = int(np.median(cnt))                                             # a min-neighbour count that binds for half the points
    ref_mask = T.radius_outliers(r, mn, False)
    res = {}
    for knob in ("1", "0"):
        os.environ["CCG_NODE_MOMENTS"] = knob
        try:
            ctx = cc.Context(0)
        finally:
            del os.environ["CCG_NODE_MOMENTS"]
        ctx.set_cloud(pts)
        eig, nrm, count = ctx.pca_radius(r, 0, cc.COV_NM1, want_normals=True)
        assert np.array_equal(count, cnt)
        scale = np.maximum(eig64[:, :1], 1e-30)
        assert (np.abs(eig - eig64) / scale).max() <= EIG_TOL
        assert np.array_equal(ctx.radius(r, 0, strict_lt=True, lists=False)["count"], cnt)
        assert np.array_equal(ctx.radius(r, mn, strict_lt=True, lists=False)["count"], np.minimum(cnt, mn))
        assert np.array_equal(ctx.radius_count_mask(r, mn, prefilter=False), ref_mask)
        res[knob] = (eig, nrm)
        ctx.close()
    assert (np.abs(res["1"][0] - res["0"][0]) / np.maximum(eig64[:, :1], 1e-30)).max() <= EIG_TOL


@pytest.mark.parametrize("kind", ["surface", "volume", "lattice", "pile"])
def test_capped_radius_far_below_the_neighbourhood_size(gpu_ctx, kind):
    """Caps 10-300x below the neighbourhood size (the reference's pca_max_nn = 1000 inside r = 0.2 m holds 24 k points on a dense
    scan): the counting select drops its bound while it walks (radsel.cu, HistV) — rows still bit-exact in indices and distances."""
    rng = np.random.default_rng(77)
    n = 24000
    xyz = rng.random((n, 3), dtype=np.float32) - 0.5
    if kind == "surface":
        xyz[:, 2] *= 0.01
    elif kind == "lattice":
        xyz = np.round(xyz * np.float32(16)) / np.float32(16)
    elif kind == "pile":
        xyz[::3] = np.float32([0.01, 0.02, 0.0])
    pts = np.zeros((n, 4), np.float32)
    pts[:, :3] = xyz
    T = oracle.Tree(pts)
    gpu_ctx.set_cloud(pts)
    for r, cap in ((0.6, 65), (0.45, 300), (0.25, 1000)):
        off, idx, d2 = gpu_ctx.radius_csr(r, cap, strict_lt=True, device_out=False)
        total = T.radius(r, strict_lt=True, lists=False)["total"]
        eff = np.minimum(total, cap)
        assert np.array_equal(np.diff(off), eff)
        ref = T.radius(r, cap=int(max(1, eff.max())), strict_lt=True)
        rows = np.repeat(np.arange(n), eff)
        cols = np.arange(off[-1]) - np.repeat(off[:-1], eff)
        assert np.array_equal(d2, ref["d2"][rows, cols])
        assert np.array_equal(idx, ref["idx"][rows, cols])


@pytest.mark.parametrize("r,max_nn,bins", [(0.1, 20, 20), (0.3, 64, 8), (0.05, 10, 33)])
def test_neighbour_histograms(gpu_ctx, r, max_nn, bins):
    """calcDistHist / calcIntensityHist (featureeval/utils.h:18-195): bit-exact (same increments, same bins)."""
    pts, _ = _scan_with_normals()
    gpu_ctx.set_cloud(pts)
    got_d = gpu_ctx.neighbour_hist(r, max_nn, bins, device_out=False)
    assert np.array_equal(got_d, oracle.neighbour_hist(pts, r, max_nn, bins))
    inten = pts[:, 4]                                    # strided view into the PointXYZI array: stride 32 bytes
    assert inten.strides[0] == 32
    got_i = gpu_ctx.neighbour_hist(r, max_nn, bins, attr=inten, device_out=False)
    assert np.array_equal(got_i, oracle.neighbour_hist(pts, r, max_nn, bins, attr=inten))
    filled = got_d.sum(1) > 0
    assert np.allclose(got_d[filled].sum(1), 1.0, atol=1e-5) and np.allclose(got_i[filled].sum(1), 1.0, atol=1e-5)


def test_difference_of_normals(gpu_ctx):
    """don() (featureeval.cpp:694-739): uncapped radius means of the neighbours' normals; 1e-5 absolute (fp32 sums in
    traversal order vs the oracle's fp64 sums); the neighbour COUNTS behind the means are checked bit-exact by the oracle."""
    scan = synth.make_scan_np(70, 50, scene="room", seed=16)
    pts = np.ascontiguousarray(scan["points"])
    rng = np.random.default_rng(3)
    nrm = np.zeros((pts.shape[0], 8), np.float32)
    v = rng.normal(size=(pts.shape[0], 3)).astype(np.float32); v[:, 2] += 2.0
    nrm[:, :3] = v / np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    gpu_ctx.set_cloud(pts)
    got = gpu_ctx.don(nrm, 0.2, 0.5, device_out=False)
    want = oracle.don(pts, nrm, 0.2, 0.5)
    assert np.abs(got[:, :3] - want[:, :3]).max() <= 1e-5
    assert np.all(got[:, 3] == 0)
    assert np.abs(want[:, :3]).max() > 0.01
    cnt = gpu_ctx.radius(0.2, 0, strict_lt=True, lists=False, device_out=False)     # the density count of markov.cpp:359-370


@pytest.mark.parametrize("kind", ["estimated", "noisy"])
def test_principal_curvatures(gpu_ctx, kind):
    """pcl::PrincipalCurvaturesEstimation (radius search; markov.cpp:389-402, featureeval.cpp:947-956).  pc1/pc2 against
    the two-pass fp32 restatement: 2e-4 relative to the row's pc1 plus 1e-7 absolute (one-pass shifted moments on the GPU);
    the principal direction up to sign where the two curvatures are separated.  "estimated" = the kNN-PCA normals of the
    same scan (smooth: the case where unshifted one-pass moments would cancel), "noisy" = random normals."""
    scan = synth.make_scan_np(70, 50, scene="room", seed=21)
    pts = np.ascontiguousarray(scan["points"])
    n = pts.shape[0]
    gpu_ctx.set_cloud(pts)
    nrm = np.zeros((n, 8), np.float32)
    if kind == "estimated":
        _, n4 = gpu_ctx.pca_knn(12, viewpoint=(0.0, 0.0, 0.0), device_out=False)[:2]
        nrm[:, :3] = n4[:, :3]
    else:
        rng = np.random.default_rng(5)
        v = rng.normal(size=(n, 3)).astype(np.float32); v[:, 2] += 1.5
        nrm[:, :3] = v / np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    assert np.isfinite(nrm).all()
    r = 0.25
    got = gpu_ctx.principal_curvatures(nrm, r, device_out=False)
    want = oracle.principal_curvatures(pts, nrm, r)
    assert got.shape == (n, 5) and np.isfinite(got).all() and np.isfinite(want).all()
    tol = 2e-4 * np.abs(want[:, 3]) + 1e-7
    assert (np.abs(got[:, 3] - want[:, 3]) <= tol).all()
    assert (np.abs(got[:, 4] - want[:, 4]) <= tol).all()
    assert (got[:, 3] >= got[:, 4]).all() and (got[:, 4] >= -1e-7).all()
    sep = (want[:, 3] - want[:, 4]) > 0.05 * want[:, 3]
    sep &= want[:, 3] > 1e-6
    assert sep.sum() > n // 4
    dots = np.abs(np.einsum("ij,ij->i", got[sep, :3].astype(np.float64), want[sep, :3].astype(np.float64)))
    assert dots.min() > 1 - 1e-3
    assert np.abs(np.linalg.norm(got[:, :3].astype(np.float64), axis=1) - 1).max() < 1e-5
    # the principal direction lies in the tangent plane
    tang = np.abs(np.einsum("ij,ij->i", got[sep, :3].astype(np.float64), nrm[sep, :3].astype(np.float64)))
    assert tang.max() < 1e-3


def test_fpfh(gpu_ctx):
    """pcl::FPFHEstimation (radius search; featureeval.cpp:925-932) against the oracle's restatement.  SPFH rows are integer
    counts x 100/(N-1): a row either matches to fp32 rounding or a pair sat within rounding of a bin edge (sinf/atan2f/
    acosf differ by an ulp between CUDA and numpy) — such rows must be rare and differ by whole increments only.  FPFH rows
    (weighted sums of ~N SPFH rows) then agree to a small L1 distance."""
    scan = synth.make_scan_np(60, 40, scene="room", seed=23)
    pts = np.ascontiguousarray(scan["points"])
    n = pts.shape[0]
    gpu_ctx.set_cloud(pts)
    _, n4 = gpu_ctx.pca_knn(12, viewpoint=(0.0, 0.0, 0.0), device_out=False)[:2]
    nrm = np.zeros((n, 8), np.float32)
    nrm[:, :3] = n4[:, :3]
    r = 0.3
    got, got_s = gpu_ctx.fpfh(nrm, r, want_spfh=True, device_out=False)
    want, want_s = oracle.fpfh(pts, nrm, r)
    assert got.shape == (n, 33) and np.isfinite(got).all()
    cnt = gpu_ctx.radius(r, 0, strict_lt=True, lists=False, device_out=False)["count"]
    lone = cnt < 2                                            # nothing but the point itself inside r: all-zero rows, as PCL
    assert lone.mean() < 0.5 and np.all(got[lone] == 0) and np.all(got_s[lone] == 0)
    # every group of a non-empty row sums to 100 (SPFH: up to skipped degenerate pairs)
    for g in range(3):
        assert np.abs(got[~lone, 11 * g:11 * g + 11].sum(1) - 100).max() < 1e-2
        assert (got_s[:, 11 * g:11 * g + 11].sum(1) <= 100 + 1e-3).all()
    incr = 100.0 / np.maximum(cnt - 1, 1)
    diff_s = np.abs(got_s.astype(np.float64) - want_s)
    rows_off = (diff_s > 1e-3 * incr[:, None]).any(1)
    assert rows_off.mean() < 0.05, rows_off.mean()
    # where a row differs it does so by whole pair increments moved to an adjacent bin: L1 per group is a multiple of 2*incr
    assert (diff_s.sum(1) <= 6.5 * 3 * incr).all()
    l1 = np.abs(got.astype(np.float64) - want).sum(1)
    # a flipped pair in a neighbour's SPFH moves 100/(N_j - 1) of that row; it shows in the FPFH in proportion to that
    # neighbour's weight, so only sparse neighbourhoods can be far off: bound the dense ones, count the rest
    dense = cnt >= 30
    assert np.median(l1) < 1e-2 and np.quantile(l1, 0.99) < 0.5, (np.median(l1), np.quantile(l1, 0.99))
    assert dense.sum() > 200 and l1[dense].max() < 5.0, (dense.sum(), l1[dense].max())
    assert (l1 > 1.0).mean() < 0.01, (l1 > 1.0).mean()
    # the descriptor is informative on this scene (not all mass in one bin)
    assert (got[~lone] > 1.0).sum(1).mean() > 3.3


# ---- the A/B switches of DESIGN.md §3.1 change speed, never results ----------------------------------------------------
@pytest.mark.parametrize("knob", [{"CCG_SEL_STAGE": "ldg"}, {"CCG_CURVE": "morton"}, {"CCG_KEY_BITS": "13"}, {"CCG_KEY_BITS": "21"},
                                  {"CCG_SEL_NET": "0"}, {"CCG_SEL_DRAIN_AT": "4"}, {"CCG_SEL_MINB": "12"}, {"CCG_POINT_OVERLAP": "0"},
                                  {"CCG_SEL_CHUNKS": "1"}])
def test_knobs_do_not_change_results(monkeypatch, knob):
    """Every measurement switch (staging by LDG instead of TMA, Morton instead of Hilbert order, key resolution, own-leaf
    network off, mid-list drains, 80-register build, point kernel on the main stream, one launch) must give the oracle's
    neighbours bit for bit and the same eigenvalues / SOR mask as the default build."""
    scan = synth.make_scan_np(220, 160, scene="room", seed=31)
    pts = np.ascontiguousarray(scan["points"])
    lat = SMALL["lattice_ties"]
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        ref_eig, ref_nrm = ctx.pca_knn(20, cc.COV_NM1)
        ref_mask, ref_md = ctx.sor_mask(30, 1.0)
    for k_, v_ in knob.items():
        monkeypatch.setenv(k_, v_)
    with cc.Context(0) as ctx:
        ctx.set_cloud(pts)
        idx, d2 = ctx.knn(20)
        ri, rd = oracle.Tree(pts).knn(20)
        assert np.array_equal(idx, ri) and np.array_equal(d2, rd)
        eig, nrm = ctx.pca_knn(20, cc.COV_NM1)
        if "CCG_CURVE" in knob or "CCG_KEY_BITS" in knob:     # another leaf order = another fp32 summation order of the moments
            assert np.abs(eig - ref_eig).max(1).max() <= 1e-4 * ref_eig[:, 0].max()
            check_pca(eig, nrm, oracle.pca_lists(pts, ri, cov_norm=cc.COV_NM1))
        else:
            assert np.array_equal(eig, ref_eig) and np.array_equal(nrm, ref_nrm, equal_nan=True)
        mask, md = ctx.sor_mask(30, 1.0)
        assert np.array_equal(mask, ref_mask) and np.array_equal(md, ref_md)
        ctx.set_cloud(lat)                                   # ties and duplicates: the point kernel decides
        idx, d2 = ctx.knn(14)
        ri, rd = oracle.Tree(lat).knn(14)
        assert np.array_equal(idx, ri) and np.array_equal(d2, rd)


def test_sor_finish_reproduces_sor_mask(gpu_ctx):
    """ccg_sor_mean_range over the whole curve order + ccg_unpermute + ccg_sor_finish (the sharded path of SURVEY §8e, here on one
    rank in three ranges) is ccg_sor_mask bit for bit: mean distances and mask."""
    import torch
    scan = synth.make_scan_np(200, 150, scene="room", seed=8)
    pts = np.ascontiguousarray(scan["points"])
    n = pts.shape[0]
    gpu_ctx.set_cloud(pts)
    mask, md = gpu_ctx.sor_mask(16, 1.5, device_out=False)
    npad = (n + 31) // 32 * 32
    md_sorted = torch.zeros((npad,), device="cuda")
    torch.cuda.synchronize()
    cuts = [0, (n // 3) // 32 * 32, (2 * n // 3) // 32 * 32, n]
    for a, b in zip(cuts[:-1], cuts[1:]):
        gpu_ctx.sor_mean_range(16, a, b, md_sorted[a:])
    md_orig = torch.empty((n,), device="cuda")
    gpu_ctx.unpermute(md_sorted[:n].view(n, 1), md_orig.view(n, 1))
    out_dev = torch.empty((n,), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    gpu_ctx.sor_finish(md_orig, 1.5, out_dev)
    assert np.array_equal(md_orig.cpu().numpy(), md)
    assert np.array_equal(out_dev.cpu().numpy(), mask)
    out_host = np.empty(n, np.uint8)
    gpu_ctx.sor_finish(md, 1.5, out_host)                    # host in, host out
    assert np.array_equal(out_host, mask)
