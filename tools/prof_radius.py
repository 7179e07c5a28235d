"""Developer probe for ncu: the capped radius search (counting select, k_tile_kth + k_tile_rule) on the config-3 scan, or the
uncapped radius PCA (k_tile_stream<StreamNodes>) on the config-5 scan.  argv: kth|stream [shrink]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cloudclean_b200 as cc
from cloudclean_b200 import synth
what = sys.argv[1] if len(sys.argv) > 1 else "kth"
shrink = int(sys.argv[2]) if len(sys.argv) > 2 else 1
s = synth.config_scan(3 if what == "kth" else 5, device="cuda", shrink=shrink)
pts = s["points"]; n = pts.shape[0]
ctx = cc.Context(0)
ctx.set_cloud(pts)
for _ in range(2):
    if what == "kth":
        eig, _, cnt = ctx.pca_radius(0.2, 1000, cc.COV_NM1)
    else:
        eig, _, cnt = ctx.pca_radius(0.1, 0, cc.COV_NM1)
torch.cuda.synchronize()
print(n, float(cnt.float().mean()), ctx.stats())
