#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python tools/lib_variants.py build/variants/ref1.so build/variants/featno.so build/variants/pf1.so 2>&1 | tee gpurun_out/lib_variants.log
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 600 python - <<'PY'
import sys, os, torch
sys.path.insert(0, os.getcwd())
import cloudclean_b200 as cc
from cloudclean_b200 import synth
s = synth.config_scan(3, device="cuda"); pts = s["points"]
ctx = cc.Context(0); ctx.set_cloud(pts)
for r in (0.05, 0.1):
    for it in range(2):
        e, _, c = ctx.pca_radius(r, 1000, device_out=True)
    print("pca_radius r=%.2f max_nn=1000: %.1f ms  mean used %.1f" % (r, ctx.stats()["ms_query"], float(c.float().mean())))
PY
