"""Developer probe for ncu: one kNN(20)+PCA pass over a shrunk config-3 scan (captured with -k regex:k_tile_topk)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cloudclean_b200 as cc
from cloudclean_b200 import synth
shrink = int(sys.argv[1]) if len(sys.argv) > 1 else 2
s = synth.config_scan(3, device="cuda", shrink=shrink)
pts = s["points"]; n = pts.shape[0]
ctx = cc.Context(0)
ctx.set_cloud(pts)
eig = torch.empty((n, 3), device="cuda"); nrm = torch.empty((n, 4), device="cuda"); feat = torch.empty((n, 7), device="cuda")
for _ in range(2):
    ctx.pca_knn_features_into(20, 1, 0, eig, nrm, feat)      # the bench call
print(n, ctx.stats())
