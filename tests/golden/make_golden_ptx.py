"""Generates tests/golden/scan12x9.ptx and scan12x9_ptx.npz — run once, by hand:  python tests/golden/make_golden_ptx.py

Golden vectors for the PTX ingest row (PointCloud::load_ptx, pointcloud.cpp:127-330).  The reference ships no PTX
fixture (test/main.cpp names plane4x5.ptx, which is not in the tree), so the expected rows are produced here by a route
independent of both oracle/ccoracle.c (stdio) and the CUDA parser:
  * lines are split in Python; the decimation rule (:219-230), the (0,0,0) drop (:263-265) and the row layout (:284-287)
    are applied line by line;
  * every decimal token becomes a float32 by EXACT rational arithmetic (fractions.Fraction): the nearest float32,
    ties to even — no libc strtof, no double rounding.
tests/test_oracle_cpu.py holds the oracle to these vectors, tests/test_gpu_ptx.py holds the CUDA path to them."""
import os
import sys
from fractions import Fraction

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from cloudclean_b200 import synth


def to_f32_exact(tok):
    """nearest float32 to the decimal string, ties to even, by exact rational comparison"""
    v = Fraction(tok)                        # Fraction parses '1.5e-3', '-0', '+2.25' exactly
    if v == 0:
        return np.float32(-0.0) if tok.strip().startswith("-") else np.float32(0.0)
    c = np.float32(float(v))                 # a candidate within one ulp
    best = None
    for cand in (np.nextafter(c, np.float32(-np.inf)), c, np.nextafter(c, np.float32(np.inf))):
        err = abs(Fraction(float(cand)) - v)
        even = (int(np.float32(cand).view(np.uint32)) & 1) == 0
        key = (err, 0 if even else 1)
        if best is None or key < best[0]:
            best = (key, cand)
    return np.float32(best[1])


def read(path, dec):
    lines = open(path).read().split("\n")
    w, h = int(lines[0]), int(lines[1])
    origin = [to_f32_exact(t) for t in lines[2].split()]
    matrix = [to_f32_exact(t) for ln in lines[3:6] for t in ln.split()]
    rows, c2g, sampled = [], [], 0
    for ln, text in enumerate(lines[10:10 + w * h]):
        row, col = ln // w, ln % w
        if (row + 1) % dec or (col + 1) % dec:
            continue
        x, y, z, i = (to_f32_exact(t) for t in text.split()[:4])
        sampled += 1
        if x == 0 and y == 0 and z == 0:
            continue
        rows.append([x, y, z, 1.0, i, 0.0, float(len(rows)), float(sampled - 1)])
        c2g.append(sampled - 1)
    return (np.array(rows, np.float32).reshape(-1, 8), np.array(c2g, np.int32), sampled,
            np.array(origin, np.float32), np.array(matrix, np.float32), w, h)


if __name__ == "__main__":
    path = os.path.join(HERE, "scan12x9.ptx")
    synth.write_ptx(path, 12, 9, seed=42, style="mixed", zero_fraction=0.15)
    out = {}
    for dec in (1, 2):
        pts, c2g, sampled, origin, matrix, w, h = read(path, dec)
        out[f"points_dec{dec}"] = pts
        out[f"c2g_dec{dec}"] = c2g
        out[f"sampled_dec{dec}"] = np.int64(sampled)
    out["origin"] = origin
    out["matrix"] = matrix
    out["size"] = np.array([w, h], np.int32)
    np.savez(os.path.join(HERE, "scan12x9_ptx.npz"), **out)
    print("wrote", path, {k: getattr(v, "shape", v) for k, v in out.items()})
