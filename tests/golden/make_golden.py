"""Generates the committed golden vectors under tests/golden/ — run once, by hand:  python tests/golden/make_golden.py

The reference holds NO golden vectors or known-answer tests for this path (its only test, test/main.cpp:32-50,
prints and asserts nothing; its fixture plane4x5.ptx is not in the tree) and PCL/FLANN/Eigen cannot be built or
imported here, so these vectors are produced by restatements that are INDEPENDENT of oracle/ccoracle.c:
  * kNN / radius : dense numpy brute force, fp32 d2 = ((dx*dx)+(dy*dy))+(dz*dz), stable argsort => (d2, index) order;
                   cross-checked against scipy.spatial.cKDTree (fp64) at generation time
  * PCA          : fp64 centroid/covariance + numpy.linalg.eigh
  * SOR          : python-float sequential sums in index order
  * grid normals : a line-by-line numpy-float32 restatement of normalestimation.cpp:268-470
tests/test_oracle_cpu.py holds the C oracle to them; tests/test_gpu_parity.py holds the CUDA path to them.
"""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

F = np.float32


def d2_matrix(xyz):
    x, y, z = (xyz[:, i].astype(F) for i in range(3))
    dx = x[:, None] - x[None, :]
    dy = y[:, None] - y[None, :]
    dz = z[:, None] - z[None, :]
    d2 = ((dx * dx) + (dy * dy)) + (dz * dz)
    assert d2.dtype == F
    return d2


def knn(d2, k):
    order = np.argsort(d2, axis=1, kind="stable")[:, :k].astype(np.int32)
    return order, np.take_along_axis(d2, order, axis=1)


def radius_lists(d2, r, cap, strict):
    n = d2.shape[0]
    r2d = float(r) * float(r)
    inside = (d2 < F(r2d)) if strict else (d2.astype(np.float64) <= r2d)
    total = inside.sum(1).astype(np.int32)
    idx = np.full((n, cap), -1, np.int32)
    dd = np.full((n, cap), np.inf, F)
    order = np.argsort(np.where(inside, d2, np.inf), axis=1, kind="stable")
    for i in range(n):
        m = min(cap, int(total[i]))
        idx[i, :m] = order[i, :m]
        dd[i, :m] = d2[i, order[i, :m]]
    return idx, dd, np.minimum(total, cap).astype(np.int32), total


def pca_truth(xyz, idx, count, cov_norm):
    n = idx.shape[0]
    eig = np.zeros((n, 3))
    nrm = np.full((n, 4), np.nan)
    for i in range(n):
        m = int(count[i])
        if m < 3:
            eig[i] = (0, 0, 1)                                  # cv.h:105-109
            continue
        p = xyz[idx[i, :m], :3].astype(np.float64)
        c = p - p.mean(0)
        S = c.T @ c
        S *= {0: 1.0, 1: 1.0 / (m - 1), 2: 1.0 / m}[cov_norm]
        w, v = np.linalg.eigh(S)
        eig[i] = w[::-1]
        nrm[i, :3] = v[:, 0]
        tr = w.sum()
        nrm[i, 3] = abs(w[0] / tr) if tr > 0 else 0.0
    return eig, nrm


def sor(d2k, mean_k, alpha):
    n = d2k.shape[0]
    md = np.zeros(n, F)
    for i in range(n):
        s = 0.0
        for j in range(1, mean_k + 1):
            if j < d2k.shape[1] and np.isfinite(d2k[i, j]):
                s += float(np.sqrt(F(d2k[i, j])))
        md[i] = F(s / mean_k)
    sm = sq = 0.0
    for i in range(n):
        sm += float(md[i])
        sq += float(md[i]) * float(md[i])
    mu = sm / n
    var = (sq - sm * sm / n) / (n - 1.0) if n > 1 else 0.0
    sigma = math.sqrt(max(var, 0.0))
    thr = mu + alpha * sigma
    return md, (md.astype(np.float64) > thr).astype(np.uint8), mu, sigma


RING = [(1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1), (1, 1)]     # normalestimation.cpp:293-302


def grid_normals(xyz, c2g, scan_w, scan_h, knn50=None):
    """normalestimation.cpp:268-470 in numpy float32 scalars; Eigen 3.2 Vector3f::squaredNorm = a0 + (a1 + a2)."""
    n = xyz.shape[0]
    height, width = scan_w, scan_h                                                # sic, :288-289
    length = width * height
    g2c = np.full(length, -1, np.int64)
    for i in range(n):
        g2c[c2g[i]] = i
    out = np.zeros((n, 4), F)
    missing = []

    def valid(rel, index):
        cx, cy = index % width, index // width
        x, y = cx + rel[0], cy + rel[1]
        if x > width - 1 or x < 0 or y > height - 1 or y < 0:
            return False
        return g2c[y * width + x] != -1

    def normalized(v):
        with np.errstate(all="ignore"):
            nn = np.sqrt(F(v[0] * v[0]) + F(F(v[1] * v[1]) + F(v[2] * v[2])))
            return np.array([v[0] / nn, v[1] / nn, v[2] / nn], F)

    for i in range(length):
        if g2c[i] == -1:
            continue
        agg = np.zeros(3, F)
        face_count, nxt, first = 0, 0, -1
        while True:
            p1 = p2 = -1
            while nxt < 8:
                if valid(RING[nxt], i):
                    p1 = nxt
                    nxt += 1
                    break
                nxt += 1
            while nxt < 9:
                if nxt == 8 and face_count == 1:
                    break
                if nxt == 8 and face_count > 1:
                    p2 = first
                    break
                if valid(RING[nxt % 8], i):
                    p2 = nxt % 8
                    break
                nxt += 1
            if p1 == -1 or p2 == -1:
                break
            if first == -1:
                first = p1
            cx, cy = i % width, i // width
            i1 = (cy + RING[p1][1]) * width + (cx + RING[p1][0])
            i2 = (cy + RING[p2][1]) * width + (cx + RING[p2][0])
            p0, q1, q2 = xyz[g2c[i], :3].astype(F), xyz[g2c[i1], :3].astype(F), xyz[g2c[i2], :3].astype(F)
            v1, v2 = q1 - p0, q2 - p0
            cr = np.array([F(v1[1] * v2[2]) - F(v1[2] * v2[1]), F(v1[2] * v2[0]) - F(v1[0] * v2[2]),
                           F(v1[0] * v2[1]) - F(v1[1] * v2[0])], F)
            t = normalized(cr)
            if np.isnan(t).any():
                continue
            agg = agg + t
            face_count += 1
        ci = g2c[i]
        if face_count == 0:
            missing.append(int(ci))
            out[ci] = np.nan
            continue
        out[ci, :3] = normalized(agg)
        out[ci, 3] = 1.0
    stage2 = out.copy()
    still = len(missing)
    if knn50 is not None:
        for idx in missing:                                                        # :436-459, in place, in order
            s = np.zeros(3, F)
            cnt = 0
            for nb in knn50[idx]:
                if nb < 0 or np.isnan(out[nb, 0]):
                    continue
                cnt += 1
                s = s + out[nb, :3]
            if cnt == 0:
                continue
            out[idx, :3] = s / F(cnt)
            still -= 1
    return stage2, out, np.array(missing, np.int32), still


def features(eig, formula_set):
    with np.errstate(all="ignore"):
        l0, l1, l2 = (eig[:, i].astype(F) for i in range(3))
        o = np.zeros((eig.shape[0], 7), F)
        o[:, 0] = (l0 - l2) / l0
        o[:, 1] = (l1 - l2) / l0
        o[:, 2] = l2 / l0
        o[:, 3] = (l0 - l1) / l0
        if formula_set == 0:                                                       # markov.cpp:467-475 as written
            o[:, 4] = np.cbrt(l0 * l1 * l0)
            o[:, 5] = -(l0 * np.log(l0) + l1 * np.log(l1) + l2 * np.log(l0))
            o[:, 6] = l0 / (l0 + l1 + l2)
        else:
            o[:, 4] = np.cbrt(l0 * l1 * l2)
            o[:, 5] = -(l0 * np.log(l0) + l1 * np.log(l1) + l2 * np.log(l2))
            o[:, 6] = l2 / (l0 + l1 + l2)
    return o


def scan_with_holes(seed):
    """A 40 x 36 organised room scan with dropouts, an isolated island (no valid ring neighbour) and duplicates."""
    from cloudclean_b200 import synth
    s = synth.make_scan_np(40, 36, scene="room", seed=seed, dropout=0.15)
    pts, c2g = s["points"], s["cloud_to_grid"]
    # carve an island: drop the 8 ring cells around one interior valid cell (reference's height/width swap applies)
    width = s["scan_h"]
    g = set(int(v) for v in c2g)
    centre = None
    for cand in c2g[len(c2g) // 3:]:
        cx, cy = int(cand) % width, int(cand) // width
        if 2 <= cx < width - 2 and 2 <= cy < s["scan_w"] - 2:
            centre = int(cand)
            break
    ring = {(centre // width + dy) * width + (centre % width + dx) for dx, dy in RING}
    keep = np.array([int(v) not in ring for v in c2g])
    pts, c2g = pts[keep].copy(), c2g[keep].copy()
    # PTX is written with 6 decimals (pointcloud.cpp:113): quantise => exact distance ties and duplicates appear
    pts[:, :3] = np.round(pts[:, :3].astype(np.float64), 2).astype(F)
    return pts, c2g.astype(np.int32), s["scan_w"], s["scan_h"]


def plane4x5():
    """Echo of the reference's missing fixture plane4x5.ptx (test/main.cpp:36): a 4 x 5 organised plane, one hole."""
    w, h = 4, 5
    pts, c2g = [], []
    for col in range(w):
        for row in range(h):
            gi = col * h + row
            if gi == 7:
                continue
            pts.append([1.0 + 0.1 * col, -0.2 + 0.1 * row, 0.5 + 0.01 * col, 1, 0.5, 0, 0, 0])
            c2g.append(gi)
    return np.array(pts, F), np.array(c2g, np.int32), w, h


def build(name, pts, c2g, w, h):
    from scipy.spatial import cKDTree
    n = pts.shape[0]
    d2 = d2_matrix(pts)
    out = dict(points=pts, cloud_to_grid=c2g, scan_w=np.int32(w), scan_h=np.int32(h))
    for k in (1, 8, 20, 50):
        kk = min(k, n)
        i, d = knn(d2, kk)
        if kk < k:
            i = np.concatenate([i, np.full((n, k - kk), -1, np.int32)], 1)
            d = np.concatenate([d, np.full((n, k - kk), np.inf, F)], 1)
        out[f"knn{k}_idx"], out[f"knn{k}_d2"] = i, d
    # independent exact search (fp64): the neighbour SETS must agree wherever the k-th distance is not tied
    kk = min(20, n)
    dd, ii = cKDTree(pts[:, :3].astype(np.float64)).query(pts[:, :3].astype(np.float64), k=kk)
    ii = ii.reshape(n, kk)
    same = 0
    for r in range(n):
        if set(ii[r]) == set(out["knn20_idx"][r, :kk]):
            same += 1
    out["ckdtree_rows_equal"] = np.int32(same)
    for r, cap in ((0.3, 20), (0.12, 8)):
        tag = f"r{int(r * 100):03d}_c{cap}"
        i, d, c, t = radius_lists(d2, r, cap, True)
        out[f"rad_{tag}_idx"], out[f"rad_{tag}_d2"], out[f"rad_{tag}_count"], out[f"rad_{tag}_total"] = i, d, c, t
        _, _, _, t2 = radius_lists(d2, r, cap, False)
        out[f"rad_{tag}_total_incl"] = t2
        e, nr = pca_truth(pts, i, c, 1)
        out[f"pcarad_{tag}_eig"], out[f"pcarad_{tag}_nrm"] = e, nr
    cnt = np.minimum(np.full(n, 20, np.int32), n)
    for cn in (0, 1, 2):
        e, nr = pca_truth(pts, out["knn20_idx"], cnt, cn)
        out[f"pca20_cov{cn}_eig"], out[f"pca20_cov{cn}_nrm"] = e, nr
    md, mask, mu, sg = sor(out["knn50_d2"], 8, 1.0)
    out["sor8_mean_d"], out["sor8_mask"], out["sor8_mu_sigma"] = md, mask, np.array([mu, sg])
    md, mask, mu, sg = sor(out["knn50_d2"], 30, 0.5)
    out["sor30_mean_d"], out["sor30_mask"], out["sor30_mu_sigma"] = md, mask, np.array([mu, sg])
    s2, s3, missing, still = grid_normals(pts, c2g, w, h, out["knn50_idx"])
    out["gridn_stage2"], out["gridn_final"], out["gridn_missing"], out["gridn_still"] = s2, s3, missing, np.int32(still)
    e32 = out["pca20_cov1_eig"].astype(F)
    out["feat_set0"], out["feat_set1"] = features(e32, 0), features(e32, 1)
    path = os.path.join(HERE, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(name, "n =", n, "missing normals:", len(missing), "still:", still, "cKDTree-equal rows:", same, "/", n,
          os.path.getsize(path), "bytes")


if __name__ == "__main__":
    build("plane4x5", *plane4x5())
    build("room40x36", *scan_with_holes(5))
