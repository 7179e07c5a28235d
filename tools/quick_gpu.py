"""Developer probe (not part of the product): time build + kNN(20)+PCA + SOR(30) on the bench scans and, for the
small ones, check parity against the oracle.  CCG_COLLECT_STATS=1 adds leaves scanned per tile."""
import sys, time, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cloudclean_b200 as cc
from cloudclean_b200 import synth
import oracle
cfgs = [(1, 1), (3, 2), (3, 1), (5, 1)]
if len(sys.argv) > 1:
    cfgs = [tuple(int(x) for x in a.split(":")) for a in sys.argv[1:]]
ctx = cc.Context(0)
for cfg, shrink in cfgs:
    s = synth.config_scan(cfg, device="cuda", shrink=shrink)
    pts = s["points"]
    n = pts.shape[0]
    ctx.set_cloud(pts); torch.cuda.synchronize()
    st = ctx.stats()
    eig = torch.empty((n, 3), device="cuda"); nrm = torch.empty((n, 4), device="cuda")
    for it in range(3):
        ctx.pca_knn_into(20, 1, eig, nrm)
        q = ctx.stats()
    print(f"cfg{cfg}/{shrink} n={n} build_ms={st['ms_build']:.2f} knn20pca_ms={q['ms_query']:.2f} -> {n/ (q['ms_query']*1e-3)/1e6:.1f} Mpts/s  leaves/tile={q['leaves_scanned']/max(1,st['n_leaves']):.2f}", flush=True)
    if n <= 3_000_000:
        h = pts.cpu().numpy()
        T = oracle.Tree(h); t0 = time.time(); ri, rd = T.knn(20); t1 = time.time()
        idx, d2 = ctx.knn(20)
        print("  oracle knn s:", round(t1 - t0, 2), "idx exact:", bool((idx.cpu().numpy() == ri).all()), "d2 exact:", bool((d2.cpu().numpy() == rd).all()), flush=True)
        res = oracle.pca_lists(h, ri, cov_norm=1)
        e = eig.cpu().numpy(); nn = nrm.cpu().numpy()
        err = np.abs(e - res["eig64"]).max(1) / res["eig64"][:, 0]
        cos = np.abs((nn[:, :3] * res["nrm64"][:, :3]).sum(1))
        ok = res["gap64"] > 1e-3
        print("  eig err max/ 99.9%:", err.max(), np.quantile(err, 0.999), " 1-|cos| max (gap>1e-3):", (1 - cos[ok]).max(), flush=True)
    out, md = ctx.sor_mask(30, 1.0)
    out, md = ctx.sor_mask(30, 1.0)      # second call: pinned bounce buffer already grown
    q = ctx.stats(); print(f"  sor k=30 query_ms={q['ms_query']:.2f} outliers={int(out.sum())}", flush=True)
    for r in (0.05, 0.1, 0.2):
        e3, _, cnt = ctx.pca_radius(r, 20)
        q = ctx.stats(); print(f"  pca_radius r={r} max_nn=20 query_ms={q['ms_query']:.2f} mean count={float(cnt.float().mean()):.1f}", flush=True)
    for res in (0.02, 0.1):
        import time as _t
        torch.cuda.synchronize(); t0 = _t.perf_counter()
        small, b2s = ctx.voxel_downsample(pts, res)
        torch.cuda.synchronize(); t1 = _t.perf_counter()
        q = ctx.stats(); print(f"  voxel_downsample res={res} query_ms={q['ms_query']:.2f} wall_ms={(t1-t0)*1e3:.2f} voxels={small.shape[0]}", flush=True)
    for r in (0.05, 0.1):
        cnt = ctx.radius(r, 0, strict_lt=True, lists=False)
        mean_cnt = float(cnt["count"].float().mean())
        pc = ctx.principal_curvatures(nrm, r)
        q = ctx.stats(); print(f"  principal_curvatures r={r} query_ms={q['ms_query']:.2f} mean neighbours={mean_cnt:.1f} finite={bool(torch.isfinite(pc).all())}", flush=True)
        dn = ctx.don(nrm, r, 2 * r)
        q = ctx.stats(); print(f"  don r1={r} r2={2*r} query_ms={q['ms_query']:.2f}", flush=True)
        if r <= 0.05:
            fp = ctx.fpfh(nrm, r)
            q = ctx.stats(); print(f"  fpfh r={r} query_ms={q['ms_query']:.2f}", flush=True)
            del fp
        del pc, dn, cnt
    if n <= 8_000_000:
        for k in (14, 20, 50):
            idx, d2 = ctx.knn(k)
            q = ctx.stats(); print(f"  knn lists k={k} query_ms={q['ms_query']:.2f} -> {n/(q['ms_query']*1e-3)/1e6:.1f} Mpts/s", flush=True)
            del idx, d2
    del s, pts, eig, nrm
    torch.cuda.empty_cache()
