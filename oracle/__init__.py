"""ctypes front end of the CPU oracle (oracle/ccoracle.c).

TEST INFRASTRUCTURE ONLY — "parity unpinned" (see the header of ccoracle.c).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
package; the product (cloudclean_b200/, adapters/) never does.

Clouds are numpy float32 arrays of shape (n, S) with S >= 3 (x, y, z first) — e.g. S=8 mirrors
pcl::PointXYZI (32 B), S=12 pcl::PointXYZINormal (48 B), S=4 a packed float4.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libccoracle.so")


def build(force=False):
    src = os.path.join(_HERE, "ccoracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libccoracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        vp, i32, f64 = C.c_void_p, C.c_int, C.c_double
        L.cco_tree_build.restype = vp
        L.cco_tree_build.argtypes = [vp, i32, i32]
        L.cco_tree_free.argtypes = [vp]
        L.cco_knn.argtypes = [vp, i32, i32, i32, vp, vp, i32]
        L.cco_knn_brute.argtypes = [vp, i32, i32, i32, i32, i32, vp, vp]
        L.cco_radius.argtypes = [vp, f64, i32, i32, i32, i32, vp, vp, vp, vp, i32]
        L.cco_pca_lists.argtypes = [vp, i32, vp, vp, i32, i32, i32, i32, vp, vp, vp, vp]
        L.cco_hist_normalise.argtypes = [vp, i32]
        L.cco_features.argtypes = [vp, i32, i32, vp]
        L.cco_eigen_ratio.argtypes = [vp, i32, vp]
        L.cco_plane_similarity.argtypes = [vp, i32, vp]
        L.cco_sor.argtypes = [vp, i32, f64, vp, vp, vp, vp, i32]
        L.cco_radius_outliers.argtypes = [vp, f64, i32, i32, vp, i32]
        L.cco_grid_normals.restype = i32
        L.cco_grid_normals.argtypes = [vp, i32, i32, vp, i32, i32, i32, i32, vp]
        L.cco_voxel_downsample.restype = i32
        L.cco_voxel_downsample.argtypes = [vp, i32, i32, i32, C.c_float, i32, vp, vp]
        L.cco_eig3_f64_export.argtypes = [vp, vp, vp]
        L.cco_eig3_f32_export.argtypes = [vp, vp]
        L.cco_max_threads.restype = i32
        L.cco_knn_pca_features.argtypes = [vp, i32, i32, i32, i32, i32, vp, vp, vp, i32]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _cloud(points):
    pts = np.ascontiguousarray(points, dtype=np.float32)
    assert pts.ndim == 2 and pts.shape[1] >= 3
    return pts


def max_threads():
    return lib().cco_max_threads()


class Tree:
    """Exact single kd-tree over a cloud — stands in for pcl::KdTreeFLANN (cv.h:83-84)."""

    def __init__(self, points):
        self.pts = _cloud(points)
        self.n = self.pts.shape[0]
        self.stride = self.pts.strides[0]
        self.h = lib().cco_tree_build(_p(self.pts), self.n, self.stride)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.cco_tree_free(self.h)
            self.h = None

    def knn(self, k, q_begin=0, q_end=None, threads=0):
        """a1 — (idx[int32 rows,k], d2[float32 rows,k]) ascending (d2, idx), self included."""
        q_end = self.n if q_end is None else q_end
        rows = q_end - q_begin
        idx = np.empty((rows, k), np.int32)
        d2 = np.empty((rows, k), np.float32)
        lib().cco_knn(self.h, k, q_begin, q_end, _p(idx), _p(d2), threads)
        return idx, d2

    def radius(self, r, cap=0, strict_lt=True, q_begin=0, q_end=None, lists=True, threads=0):
        """a2/a3 — lists=True (needs cap > 0): dict(count, idx, d2) with the `cap` nearest in radius;
        lists=False: dict(count, total) where total = every point inside the radius (exact)."""
        q_end = self.n if q_end is None else q_end
        rows = q_end - q_begin
        count = np.empty(rows, np.int32)
        total = np.empty(rows, np.int32)
        idx = d2 = None
        if lists:
            assert cap > 0, "lists need a cap"
            idx = np.empty((rows, cap), np.int32)
            d2 = np.empty((rows, cap), np.float32)
        lib().cco_radius(self.h, float(r), cap, 1 if strict_lt else 0, q_begin, q_end,
                         _p(idx), _p(d2), _p(count), _p(total), threads)
        if lists:
            return {"count": count, "idx": idx, "d2": d2}
        return {"count": count, "total": total}

    def sor(self, mean_k, stddev_mul, threads=0):
        """a11 — (outlier[uint8 n], mean_d[float32 n], mu, sigma)."""
        out = np.empty(self.n, np.uint8)
        md = np.empty(self.n, np.float32)
        mu, sg = C.c_double(), C.c_double()
        lib().cco_sor(self.h, mean_k, float(stddev_mul), _p(out), _p(md), C.byref(mu), C.byref(sg), threads)
        return out, md, mu.value, sg.value

    def radius_outliers(self, r, min_neighbours, prefilter=False, threads=0):
        """a10 — outlier[uint8 n] (outlierfilter.cpp:120-179)."""
        out = np.empty(self.n, np.uint8)
        lib().cco_radius_outliers(self.h, float(r), min_neighbours, 1 if prefilter else 0, _p(out), threads)
        return out


def knn_pca_features(tree, k, cov_norm=1, formula_set=0, threads=0, normals=True):
    """The bench metric's CPU baseline: per point kNN(k) -> fp32 PCA -> 7 features, fused (the reference's getPCA loop shape,
    cv.h:96-113 + markov.cpp:449-475), OpenMP over queries (threads=1: serial like the reference).  -> eig3, normal4, feat7"""
    n = tree.n
    eig = np.empty((n, 3), np.float32)
    nrm = np.empty((n, 4), np.float32) if normals else None
    feat = np.empty((n, 7), np.float32)
    lib().cco_knn_pca_features(tree.h, k, cov_norm, formula_set, 0, n, _p(eig), _p(nrm) if normals else None, _p(feat), threads)
    return eig, nrm, feat


def knn_brute(points, k, q_begin=0, q_end=None):
    pts = _cloud(points)
    n = pts.shape[0]
    q_end = n if q_end is None else q_end
    idx = np.empty((q_end - q_begin, k), np.int32)
    d2 = np.empty((q_end - q_begin, k), np.float32)
    lib().cco_knn_brute(_p(pts), n, pts.strides[0], k, q_begin, q_end, _p(idx), _p(d2))
    return idx, d2


def knn_numpy(points, k, queries=None):
    """Independent numpy restatement (small n only): same fp32 op order, lexsort on (d2, idx)."""
    pts = _cloud(points)[:, :3]
    q = pts if queries is None else pts[queries]
    dx = q[:, None, 0] - pts[None, :, 0]
    dy = q[:, None, 1] - pts[None, :, 1]
    dz = q[:, None, 2] - pts[None, :, 2]
    d2 = ((dx * dx) + (dy * dy)) + (dz * dz)
    assert d2.dtype == np.float32
    order = np.argsort(d2, axis=1, kind="stable")[:, :k]      # stable => ties by index
    return order.astype(np.int32), np.take_along_axis(d2, order, axis=1)


def pca_lists(points, nbr, count=None, cov_norm=1, min_pts=3):
    """a4/a6/a12 — dict(eig32, eig64, nrm64, gap64) for given neighbour lists."""
    pts = _cloud(points)
    nbr = np.ascontiguousarray(nbr, np.int32)
    rows, cap = nbr.shape
    cnt = None if count is None else np.ascontiguousarray(count, np.int32)
    eig32 = np.empty((rows, 3), np.float32)
    eig64 = np.empty((rows, 3), np.float64)
    nrm64 = np.empty((rows, 4), np.float64)
    gap64 = np.empty(rows, np.float64)
    lib().cco_pca_lists(_p(pts), pts.strides[0], _p(nbr), _p(cnt), rows, cap, cov_norm, min_pts,
                        _p(eig32), _p(eig64), _p(nrm64), _p(gap64))
    return {"eig32": eig32, "eig64": eig64, "nrm64": nrm64, "gap64": gap64}


def hist_normalise(eig):
    e = np.ascontiguousarray(eig, np.float32).copy()
    lib().cco_hist_normalise(_p(e), e.shape[0])
    return e


def features(eig3, formula_set=0):
    e = np.ascontiguousarray(eig3, np.float32)
    out = np.empty((e.shape[0], 7), np.float32)
    with np.errstate(all="ignore"):
        lib().cco_features(_p(e), e.shape[0], formula_set, _p(out))
    return out


def eigen_ratio(eig3):
    e = np.ascontiguousarray(eig3, np.float32)
    out = np.empty(e.shape[0], np.float32)
    lib().cco_eigen_ratio(_p(e), e.shape[0], _p(out))
    return out


def plane_similarity(eig3):
    e = np.ascontiguousarray(eig3, np.float32)
    out = np.empty(e.shape[0], np.float32)
    lib().cco_plane_similarity(_p(e), e.shape[0], _p(out))
    return out


def grid_normals(points, cloud_to_grid, scan_w, scan_h, fallback=True, k_fallback=50):
    """a7 — (normal4[float32 n,4], still_missing)."""
    pts = _cloud(points)
    c2g = np.ascontiguousarray(cloud_to_grid, np.int32)
    out = np.empty((pts.shape[0], 4), np.float32)
    still = lib().cco_grid_normals(_p(pts), pts.shape[0], pts.strides[0], _p(c2g), scan_w, scan_h,
                                   1 if fallback else 0, k_fallback, _p(out))
    return out, still


def voxel_downsample(points, resolution, n_floats=None, as_written=True):
    """f1 (utils.h:16-62) — (means[float32 nvox, nf], big_to_small[int32 n]); sequential fp32 sums in index order."""
    pts = np.ascontiguousarray(points, dtype=np.float32)
    nf = pts.shape[1] if n_floats is None else n_floats
    out = np.empty((pts.shape[0], nf), np.float32)
    b2s = np.empty(pts.shape[0], np.int32)
    nv = lib().cco_voxel_downsample(_p(pts), pts.shape[0], pts.strides[0], nf, float(resolution), 1 if as_written else 0,
                                    _p(out), _p(b2s))
    return out[:nv].copy(), b2s


def binary_potential(points, idx, inverse_sigma):
    """f3 — MinCut::calculateBinaryPotential (graphcut/mincut.cpp:574-585) for every edge (i, idx[i, j]):
    float differences and float products, summed in double, * inverse_sigma, exp(-x) in double; idx < 0 -> 0."""
    p = _cloud(points)[:, :3]
    idx = np.asarray(idx)
    j = np.where(idx < 0, 0, idx)
    d = (p[:, None, :] - p[j]).astype(np.float32)
    sq = (d * d).astype(np.float32)
    dist = sq[..., 0].astype(np.float64) + sq[..., 1].astype(np.float64) + sq[..., 2].astype(np.float64)
    w = np.exp(-(dist * float(inverse_sigma)))
    return np.where(idx < 0, 0.0, w)


def _eigen_sum3(a0, a1, a2):
    """Sum of a fixed-size 3-vector expression as Eigen <= 3.2 evaluates it for Vector3f / Map<Vector3f> (.dot(), .norm(),
    .sum()): not vectorised (12 bytes), completely unrolled by redux_novec_unroller, which halves the range —
    func(a0, func(a1, a2)) [EXT Eigen/src/Core/Redux.h].  float32 in, float32 out."""
    F = np.float32
    return (a0.astype(F) + (a1.astype(F) + a2.astype(F)).astype(F)).astype(F)


def _capped_lists(pts, r, max_nn, strict_lt):
    """radiusSearch(i, r, idx, d2, max_nn) for every point: the max_nn nearest inside r in ascending (d2, index) order;
    max_nn == 0 or >= n means every point inside r (FLANN caps only when 0 < max_nn < n)."""
    T = Tree(pts)
    n = pts.shape[0]
    if max_nn <= 0 or max_nn >= n:
        max_nn = int(max(1, T.radius(r, strict_lt=strict_lt, lists=False)["total"].max()))
    return T.radius(r, cap=max_nn, strict_lt=strict_lt)


def normal_stdev(points, normals, r, max_nn):
    """f2 — normal_stdev (cv.h:123-206) as written: strict radius, the max_nn nearest in ascending (d2, index) order,
    cosine = dot / |neighbour| * |current| (sic), clamp, angle = float(acosf / pi [double]), float sums in neighbour
    order, std = float(sqrt(double(sumsq/m) - double(sum/m)^2)); 0 for NaN normals, < 3 neighbours, NaN results."""
    F = np.float32
    pts = _cloud(points)
    nr = np.ascontiguousarray(normals, dtype=F)[:, :3]
    res = _capped_lists(pts, r, max_nn, True)
    idx, cnt = res["idx"], res["count"]
    n = pts.shape[0]
    cur = nr
    with np.errstate(all="ignore"):
        cn = np.sqrt(_eigen_sum3(cur[:, 0] * cur[:, 0], cur[:, 1] * cur[:, 1], cur[:, 2] * cur[:, 2])).astype(F)
        s = np.zeros(n, F); sq = np.zeros(n, F)
        for j in range(idx.shape[1]):
            live = j < cnt
            nb = nr[np.where(live, idx[:, j], 0)]
            dot = _eigen_sum3(nb[:, 0] * cur[:, 0], nb[:, 1] * cur[:, 1], nb[:, 2] * cur[:, 2])
            nn = np.sqrt(_eigen_sum3(nb[:, 0] * nb[:, 0], nb[:, 1] * nb[:, 1], nb[:, 2] * nb[:, 2])).astype(F)
            cosine = ((dot / nn).astype(F) * cn).astype(F)
            cosine = np.where(cosine < 0, F(0), np.where(cosine > 1, F(1), cosine)).astype(F)       # clamp(): NaN passes through
            angle = (np.arccos(cosine).astype(F).astype(np.float64) / np.pi).astype(F)
            s = np.where(live, (s + angle).astype(F), s)
            sq = np.where(live, (sq + (angle * angle).astype(F)).astype(F), sq)
        m = cnt.astype(F)
        mean = (s / m).astype(F).astype(np.float64)
        sd = np.sqrt((sq / m).astype(F).astype(np.float64) - mean * mean).astype(F)
    out = np.where((cur[:, 0] == cur[:, 0]) & (cnt >= 3) & (sd == sd), sd, F(0)).astype(F)
    return out


def stdev_dist(points, r, max_nn, use_depth=False, centre=None):
    """f2 — stdev_dist (cv.cpp:166-230) as written: inclusive radius, the max_nn nearest (see include/ccgpu.h on the
    octree's own cap), float sums of |p - q| (or |p - centre|), (sum_sq - sum^2/m)/(m-1) for m > 2 else 0."""
    F = np.float32
    pts = _cloud(points)
    xyz = pts[:, :3]
    res = _capped_lists(pts, r, max_nn, False)
    idx, cnt = res["idx"], res["count"]
    n = pts.shape[0]
    ref = np.broadcast_to(np.asarray(centre, F), (n, 3)) if use_depth else xyz
    s = np.zeros(n, F); sq = np.zeros(n, F)
    for j in range(idx.shape[1]):
        live = j < cnt
        p = xyz[np.where(live, idx[:, j], 0)]
        d = (p - ref).astype(F)
        dist = np.sqrt(_eigen_sum3(d[:, 0] * d[:, 0], d[:, 1] * d[:, 1], d[:, 2] * d[:, 2])).astype(F)
        s = np.where(live, (s + dist).astype(F), s)
        sq = np.where(live, (sq + (dist * dist).astype(F)).astype(F), sq)
    m = cnt.astype(F)
    with np.errstate(all="ignore"):
        var = ((sq - ((s * s).astype(F) / m).astype(F)).astype(F) / (m - F(1)).astype(F)).astype(F)
    return np.where(cnt > 2, var, F(0)).astype(F)


def neighbour_hist(points, r, max_nn, bins, attr=None):
    """f2 — calcDistHist (attr None) / calcIntensityHist (featureeval/utils.h:18-195) as written: strict radius, the
    max_nn nearest, < 3 neighbours -> zeros, every neighbour adds float(1.0/m) to its bin (clamped to [0, bins-1])."""
    F = np.float32
    pts = _cloud(points)
    res = _capped_lists(pts, r, max_nn, True)
    idx, cnt, d2 = res["idx"], res["count"], res["d2"]
    n = pts.shape[0]
    hist = np.zeros((n, bins), F)
    inc = (1.0 / np.maximum(cnt, 1).astype(np.float64)).astype(F)
    rows = np.arange(n)
    for j in range(idx.shape[1]):
        live = (j < cnt) & (cnt >= 3)
        if attr is None:
            b = ((np.sqrt(np.where(live, d2[:, j], 0).astype(F)).astype(np.float64) / float(r)) * (bins - 1)).astype(np.int64)
        else:
            a = np.asarray(attr, F)[np.where(live, idx[:, j], 0)]
            b = (a * F(bins - 1)).astype(F).astype(np.int64)
        b = np.clip(b, 0, bins - 1)
        hist[rows[live], b[live]] = (hist[rows[live], b[live]] + inc[live]).astype(F)
    return hist


def don(points, normals, r1=0.2, r2=0.5):
    """f2 — don() (featureeval.cpp:694-739): (mean1 - mean2)/2 of the neighbours' normals inside the two strict, uncapped
    radii (float radii, as the reference's signature), the point's own normal counted twice.  fp64 sums here: the
    reference adds in its tree's (unspecified) order in fp32, the GPU in traversal order — compared at 1e-5."""
    from scipy.spatial import cKDTree
    pts = _cloud(points)[:, :3]
    nr = np.ascontiguousarray(normals, np.float32)[:, :3].astype(np.float64)
    n = pts.shape[0]
    T = Tree(_cloud(points))
    out = np.zeros((n, 4), np.float32)
    means = []
    for r in (np.float32(r1), np.float32(r2)):
        cnt = T.radius(float(r), strict_lt=True, lists=False)["total"]
        kd = cKDTree(pts.astype(np.float64))
        nb = kd.query_ball_point(pts.astype(np.float64), float(r) * 1.0001)
        rb2 = np.float32(float(r) * float(r))
        m = np.zeros((n, 3))
        for i in range(n):
            j = np.asarray(nb[i])
            d = (pts[i] - pts[j]).astype(np.float32)
            d2 = ((d[:, 0] * d[:, 0]).astype(np.float32) + (d[:, 1] * d[:, 1]).astype(np.float32)).astype(np.float32) + (d[:, 2] * d[:, 2]).astype(np.float32)
            j = j[d2.astype(np.float32) < rb2]
            assert len(j) == cnt[i]
            m[i] = (nr[j].sum(0) + nr[i]) / (len(j) + 1)
        means.append(m)
    out[:, :3] = ((means[0] - means[1]) / 2.0).astype(np.float32)
    return out


def principal_curvatures(points, normals, r):
    """f4 — pcl::PrincipalCurvaturesEstimation::computePointPrincipalCurvatures [EXT] with setRadiusSearch(r)
    (markov.cpp:389-402, featureeval.cpp:947-956), restated from PCL's published algorithm (PCL is absent: unpinned):
    M = I - n n^T; p_j = M n_j over the strict, uncapped radius neighbourhood (self included); centroid; scatter
    matrix of the demeaned p_j (two passes, fp32 as PCL; fp64 eigensolve here, PCL uses its closed-form eigen33);
    row = (eigenvector of the largest eigenvalue, l_max / N, l_mid / N).  Compared at 1e-5-ish, eigenvector up to sign."""
    from scipy.spatial import cKDTree
    pts = _cloud(points)[:, :3]
    nr = np.ascontiguousarray(np.asarray(normals, np.float32)[:, :3])
    n = pts.shape[0]
    T = Tree(_cloud(points))
    cnt = T.radius(float(r), strict_lt=True, lists=False)["total"]
    kd = cKDTree(pts.astype(np.float64))
    nb = kd.query_ball_point(pts.astype(np.float64), float(r) * 1.0001)
    rb2 = np.float32(float(r) * float(r))
    out = np.full((n, 5), np.nan, np.float32)
    f32 = np.float32
    for i in range(n):
        j = np.sort(np.asarray(nb[i]))
        d = (pts[i] - pts[j]).astype(f32)
        d2 = ((d[:, 0] * d[:, 0]).astype(f32) + (d[:, 1] * d[:, 1]).astype(f32)).astype(f32) + (d[:, 2] * d[:, 2]).astype(f32)
        j = j[d2.astype(f32) < rb2]
        assert len(j) == cnt[i]
        if len(j) == 0:
            continue
        ni = nr[i]
        M = (np.eye(3, dtype=f32) - np.outer(ni, ni).astype(f32)).astype(f32)
        proj = (nr[j] @ M.T).astype(f32)
        cen = (proj.sum(0, dtype=f32) / f32(len(j))).astype(f32)
        dm = (proj - cen).astype(f32)
        cov = (dm.T @ dm).astype(f32)
        w, v = np.linalg.eigh(cov.astype(np.float64))
        out[i, :3] = v[:, 2]
        out[i, 3] = w[2] / len(j)
        out[i, 4] = w[1] / len(j)
    return out


def _radius_lists(points, r):
    """strict, uncapped radius neighbourhoods under the contract's fp32 d2 (index order) -> (list of index arrays, list of d2)"""
    from scipy.spatial import cKDTree
    pts = _cloud(points)[:, :3]
    n = pts.shape[0]
    T = Tree(_cloud(points))
    cnt = T.radius(float(r), strict_lt=True, lists=False)["total"]
    p64 = pts.astype(np.float64)
    nb = cKDTree(p64).query_ball_point(p64, float(r) * 1.0001)
    rb2 = np.float32(float(r) * float(r))
    f32 = np.float32
    idxs, d2s = [], []
    for i in range(n):
        j = np.sort(np.asarray(nb[i]))
        d = (pts[i] - pts[j]).astype(f32)
        d2 = ((d[:, 0] * d[:, 0]).astype(f32) + (d[:, 1] * d[:, 1]).astype(f32)).astype(f32) + (d[:, 2] * d[:, 2]).astype(f32)
        d2 = d2.astype(f32)
        keep = d2 < rb2
        assert int(keep.sum()) == cnt[i]
        idxs.append(j[keep]); d2s.append(d2[keep])
    return idxs, d2s


def fpfh(points, normals, r):
    """f4 — pcl::FPFHEstimation with setRadiusSearch(r) [EXT] (featureeval.cpp:925-932), restated from PCL's published
    algorithm (PCL is absent: unpinned): computePairFeatures, computeSPFHSignature (increments of 100/(N-1), bins
    floor(11*(f1+pi)/(2pi)), floor(11*(f2+1)/2), floor(11*(f3+1)/2) clamped), weightPointSPFHSignature (weights 1/d2,
    d2 == 0 skipped, groups normalised to 100).  fp32 pair arithmetic, vectorised over a point's neighbours.
    -> (fpfh float32[n, 33], spfh float32[n, 33])"""
    f32 = np.float32
    pts = _cloud(points)[:, :3]
    nr = np.ascontiguousarray(np.asarray(normals, f32)[:, :3])
    n = pts.shape[0]
    idxs, d2s = _radius_lists(points, r)
    spfh = np.zeros((n, 33), f32)
    d_pi = f32(1.0) / (f32(2.0) * f32(np.pi))
    for i in range(n):
        j = idxs[i]
        N = len(j)
        j = j[j != i]
        if len(j) == 0:
            continue
        dp = (pts[j] - pts[i]).astype(f32)
        f4 = np.sqrt((dp * dp).sum(1, dtype=f32)).astype(f32)
        ok = f4 != 0
        j, dp, f4 = j[ok], dp[ok], f4[ok]
        n1 = np.broadcast_to(nr[i], dp.shape).copy()
        n2 = nr[j].copy()
        a1 = ((n1 * dp).sum(1, dtype=f32) / f4).astype(f32)
        a2 = ((n2 * dp).sum(1, dtype=f32) / f4).astype(f32)
        sw = np.arccos(np.abs(a1)).astype(f32) > np.arccos(np.abs(a2)).astype(f32)
        n1s = np.where(sw[:, None], n2, n1); n2s = np.where(sw[:, None], n1, n2)
        dps = np.where(sw[:, None], -dp, dp)
        f3 = np.where(sw, -a2, a1).astype(f32)
        v = np.cross(dps, n1s).astype(f32)
        vn = np.sqrt((v * v).sum(1, dtype=f32)).astype(f32)
        ok = vn != 0
        v = (v[ok] / vn[ok, None]).astype(f32); n1s, n2s, f3 = n1s[ok], n2s[ok], f3[ok]
        w = np.cross(n1s, v).astype(f32)
        f2 = (v * n2s).sum(1, dtype=f32)
        f1 = np.arctan2((w * n2s).sum(1, dtype=f32), (n1s * n2s).sum(1, dtype=f32)).astype(f32)
        b1 = np.clip(np.floor(11.0 * ((f1.astype(np.float64) + np.pi) * float(d_pi))).astype(int), 0, 10)
        b2 = np.clip(np.floor(11.0 * ((f2.astype(np.float64) + 1.0) * 0.5)).astype(int), 0, 10)
        b3 = np.clip(np.floor(11.0 * ((f3.astype(np.float64) + 1.0) * 0.5)).astype(int), 0, 10)
        incr = f32(100.0) / f32(N - 1)
        spfh[i, :11] = np.bincount(b1, minlength=11).astype(f32) * incr
        spfh[i, 11:22] = np.bincount(b2, minlength=11).astype(f32) * incr
        spfh[i, 22:] = np.bincount(b3, minlength=11).astype(f32) * incr
    out = np.zeros((n, 33), f32)
    for i in range(n):
        j, d2 = idxs[i], d2s[i]
        ok = d2 != 0
        j, d2 = j[ok], d2[ok]
        if len(j) == 0:
            continue
        wgt = (f32(1.0) / d2).astype(f32)
        acc = (spfh[j] * wgt[:, None]).sum(0, dtype=np.float64)
        for g in range(3):
            sm = acc[11 * g:11 * g + 11].sum()
            if sm != 0:
                out[i, 11 * g:11 * g + 11] = (acc[11 * g:11 * g + 11] * (100.0 / sm)).astype(f32)
    return out, spfh


def load_ptx(filename, decimation_factor=1):
    """f4 — PointCloud::load_ptx (pointcloud.cpp:127-330) through the same stdio calls (oracle/ccoracle.c cco_load_ptx):
    -> dict(points float32[n, 8], cloud_to_grid int32[n], file_width, file_height, has_rgb, sampled, origin[3], matrix[9])"""
    import ctypes as C
    L = lib()
    L.cco_load_ptx.restype = C.c_longlong
    L.cco_load_ptx.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
    hdr = np.zeros(4, np.int64)
    pose = np.zeros(12, np.float32)
    cap = 1
    for _ in range(2):
        pts = np.zeros((cap, 8), np.float32)
        c2g = np.zeros(cap, np.int32)
        n = L.cco_load_ptx(str(filename).encode(), int(decimation_factor), _p(pts), _p(c2g), cap, _p(hdr), _p(pose))
        if n < 0:
            raise IOError("cco_load_ptx failed (%d)" % n)
        if n <= cap:
            break
        cap = int(n)
    return {"points": pts[:n].copy(), "cloud_to_grid": c2g[:n].copy(), "file_width": int(hdr[0]), "file_height": int(hdr[1]),
            "has_rgb": bool(hdr[2]), "sampled": int(hdr[3]), "origin": pose[:3].copy(), "matrix": pose[3:].copy()}


def write_ptx(filename, rows8, width, height):
    """bench/test infrastructure: a PTX file of width*height lines from [n, 8] float32 rows (NaN x or i >= n -> "0 0 0")"""
    import ctypes as C
    L = lib()
    L.cco_write_ptx.restype = C.c_longlong
    L.cco_write_ptx.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int]
    rows8 = np.ascontiguousarray(rows8, np.float32)
    assert rows8.ndim == 2 and rows8.shape[1] == 8
    return L.cco_write_ptx(str(filename).encode(), _p(rows8), rows8.shape[0], int(width), int(height))


def strtof_batch(tokens):
    """glibc strtof on a list of decimal strings -> float32[n] (what fscanf("%f") computes)"""
    import ctypes as C
    L = lib()
    L.cco_strtof_batch.restype = None
    L.cco_strtof_batch.argtypes = [C.c_char_p, C.c_int, C.c_void_p]
    blob = b"\0".join(t.encode() for t in tokens) + b"\0"
    out = np.zeros(len(tokens), np.float32)
    L.cco_strtof_batch(blob, len(tokens), _p(out))
    return out


def eig3_f64(a6):
    a = np.ascontiguousarray(a6, np.float64)
    w = np.empty(3, np.float64)
    v = np.empty(9, np.float64)
    lib().cco_eig3_f64_export(_p(a), _p(w), _p(v))
    return w, v.reshape(3, 3)


def eig3_f32(a6):
    a = np.ascontiguousarray(a6, np.float32)
    w = np.empty(3, np.float32)
    lib().cco_eig3_f32_export(_p(a), _p(w))
    return w
