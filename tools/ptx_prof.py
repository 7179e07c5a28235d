"""PTX parse of a resident byte buffer, for the ncu launch list: python tools/ptx_prof.py [width height]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cloudclean_b200 as cc
import oracle

w = int(sys.argv[1]) if len(sys.argv) > 2 else 4096
h = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
n = w * h
rng = np.random.default_rng(1)
rows = np.zeros((n, 8), np.float32)
rows[:, :3] = rng.normal(size=(n, 3)).astype(np.float32) * np.array([15.0, 12.0, 3.0], np.float32)
rows[:, 4] = rng.random(n, dtype=np.float32)
path = "/tmp/ptx_prof.ptx"
oracle.write_ptx(path, rows, w, h)
raw = torch.from_numpy(np.fromfile(path, np.uint8)).cuda()
os.remove(path)
ctx = cc.Context(0)
for it in range(3):
    got = ctx.ptx_parse(raw, 1, device_out=True)
    print(f"parse {ctx.stats()['ms_query']:.2f} ms, {raw.numel()/1e6:.0f} MB, {n} lines", flush=True)
