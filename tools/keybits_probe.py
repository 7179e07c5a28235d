"""Developer probe: build + kNN(20)+PCA+features time against the curve-key resolution (CCG_KEY_BITS) on configs 3 / 4, and the
adversarial duplicate pile of config 5.  usage: keybits_probe.py [cfgs, e.g. 3,4] [dups]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cloudclean_b200 as cc
from cloudclean_b200 import synth

cfgs = [int(c) for c in (sys.argv[1] if len(sys.argv) > 1 else "3,4").split(",") if c]
def run(pts, tag):
    n = pts.shape[0]
    eig = torch.empty((n, 3), device="cuda"); nrm = torch.empty((n, 4), device="cuda"); feat = torch.empty((n, 7), device="cuda")
    for bits in os.environ.get("CCG_KEY_BITS_LIST", "13,16").split(","):
        os.environ["CCG_KEY_BITS"] = bits
        ctx = cc.Context(0)
        b, q = [], []
        for _ in range(3):
            ctx.set_cloud(pts); b.append(ctx.stats()["ms_build"])
            ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat); st = ctx.stats(); q.append(st["ms_query"])
        print(f"{tag} n={n} key_bits={bits}: build_ms={min(b):.2f} query_ms={min(q):.2f} -> {n / (min(b) + min(q)) / 1e3:.1f} Mpts/s "
              f"deferred={st['deferred_queries']} checksum={int(eig.view(torch.int32).sum(dtype=torch.int64).item())}", flush=True)
        ctx.close()
    del eig, nrm, feat
for cfg in cfgs:
    s = synth.config_scan(cfg, device="cuda")
    pts = s["points"]; del s
    run(pts, f"cfg{cfg}")
    del pts; torch.cuda.empty_cache()
if "dups" in sys.argv:
    a = dict(synth.CONFIGS[5]); a["scan_w"] //= 4; a["scan_h"] //= 4
    for dup in (100_000, 1_000_000):
        s = synth.make_scan(device="cuda", duplicate=dup, **a)
        run(s["points"], f"cfg5/16 + {dup} duplicates")
        del s; torch.cuda.empty_cache()
