"""Developer probe: time ccg_set_cloud + ccg_pca_knn_features(k=20) on a bench scan under the k_tile_sel A/B knobs
(one Context per variant; the knobs are read by ccg_create).  usage: sel_variants.py [cfg:shrink] [stats]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cloudclean_b200 as cc
from cloudclean_b200 import synth

cfg, shrink = (int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "3:1").split(":"))
variants = [
    {},
    {"CCG_SEL_NET": "0"},
    {"CCG_SEL_MINB": "12"},
    {"CCG_SEL_MINB": "12", "CCG_SEL_NET": "0"},
    {"CCG_SEL_STAGE": "ldg"},
    {"CCG_SEL_CHUNKS": "1"},
]
knobs = ["CCG_SEL_NET", "CCG_SEL_MINB", "CCG_CURVE", "CCG_SEL_STAGE", "CCG_SEL_DRAIN_AT", "CCG_POINT_OVERLAP", "CCG_SEL_CHUNKS", "CCG_COLLECT_STATS", "CCG_VERBOSE_STATS"]
s = synth.config_scan(cfg, device="cuda", shrink=shrink)
pts = s["points"]; n = pts.shape[0]
eig = torch.empty((n, 3), device="cuda"); nrm = torch.empty((n, 4), device="cuda"); feat = torch.empty((n, 7), device="cuda")
ref = None
for stats in ([False, True] if "stats" in sys.argv else [False]):
    for v in variants:
        for k in knobs:
            os.environ.pop(k, None)
        os.environ.update(v)
        if stats:
            os.environ["CCG_COLLECT_STATS"] = "1"; os.environ["CCG_VERBOSE_STATS"] = "1"
        ctx = cc.Context(0)
        bms, qms = [], []
        for it in range(2 if stats else 5):
            ctx.set_cloud(pts); bms.append(ctx.stats()["ms_build"])
            ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); ev0.record()
            ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat)
            ev1.record(); torch.cuda.synchronize()
            qms.append(ev0.elapsed_time(ev1))
        st = ctx.stats()
        chk = int(eig.view(torch.int32).sum(dtype=torch.int64).item())
        if ref is None:
            ref = chk
        print(f"{'stats ' if stats else ''}{v or 'default'}: n={n} build_ms={min(bms):.2f} query_ms min={min(qms):.2f} mean={sum(qms[1:])/len(qms[1:]):.2f} "
              f"-> {n/min(qms)/1e3:.1f} Mpts/s deferred={st['deferred_queries']} leaves/tile={st['leaves_scanned']/max(1,st['n_leaves']):.2f} "
              f"eig checksum {'same' if chk == ref else 'DIFFERENT'}", flush=True)
        ctx.close()
