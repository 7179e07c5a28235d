"""Full-size PTX ingest: writes a synthetic scan as a PTX file, times the GPU loader (ccg_ptx_load: mmap + H2D + parse)
and the oracle's stdio restatement of PointCloud::load_ptx on the same file, checks bit equality.
usage: python tools/ptx_bench.py [width height] (default 4096 x 2048 = 8.4 M lines)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cloudclean_b200 as cc
import oracle

w = int(sys.argv[1]) if len(sys.argv) > 2 else 4096
h = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
n = w * h
rng = np.random.default_rng(1)
rows = np.zeros((n, 8), np.float32)
rows[:, :3] = rng.normal(size=(n, 3)).astype(np.float32) * np.array([15.0, 12.0, 3.0], np.float32)
rows[:, 4] = rng.random(n, dtype=np.float32)
rows[rng.random(n) < 0.1, 0] = np.nan                # 10 % no-return lines
path = "/tmp/ptx_bench.ptx"
t0 = time.perf_counter(); oracle.write_ptx(path, rows, w, h); t1 = time.perf_counter()
size = os.path.getsize(path)
print(f"wrote {path}: {n} lines, {size/1e6:.1f} MB in {t1-t0:.1f} s", flush=True)
ctx = cc.Context(0)
for it in range(3):
    t0 = time.perf_counter(); got = ctx.ptx_load(path, 1); t1 = time.perf_counter()
    st = ctx.stats()
    print(f"gpu ptx_load: wall {1e3*(t1-t0):.1f} ms (incl. mmap + fetch to host), h2d {st['ms_upload']:.2f} ms, device parse {st['ms_query']:.2f} ms "
          f"-> {n/(t1-t0)/1e6:.1f} M lines/s wall, {size/st['ms_query']/1e6:.1f} GB/s parse", flush=True)
import torch
raw = torch.from_numpy(np.fromfile(path, np.uint8)).cuda()
for it in range(2):
    got_dev = ctx.ptx_parse(raw, 1, device_out=True)
    st = ctx.stats()
    print(f"gpu ptx_parse (bytes resident): device parse {st['ms_query']:.2f} ms -> {n/st['ms_query']/1e3:.1f} M lines/s, {size/st['ms_query']/1e6:.1f} GB/s", flush=True)
t0 = time.perf_counter(); want = oracle.load_ptx(path, 1); t1 = time.perf_counter()
print(f"oracle load_ptx (fscanf, 1 core): {t1-t0:.2f} s -> {n/(t1-t0)/1e6:.2f} M lines/s", flush=True)
print("points:", got["points"].shape[0], "equal:", bool(np.array_equal(got["points"].view(np.int32), want["points"].view(np.int32))
      and np.array_equal(got["cloud_to_grid"], want["cloud_to_grid"])), flush=True)
os.remove(path)
