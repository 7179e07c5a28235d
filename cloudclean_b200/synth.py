"""Synthetic terrestrial-laser-scan generator.

An organised spherical scan from the origin: `scan_w` columns (azimuth) x `scan_h` rows (elevation),
column-major file order exactly like the PTX dialect the reference reads
(src/model/pointcloud.cpp:149-298: header rows/cols, `x y z i` rows, `0 0 0` rows dropped at :275-277,
grid index = file order :294-297).  Scenes are analytic (ray-cast planes, rectangles, spheres,
volumetric "vegetation" blobs, isolated outliers) with Gaussian range noise and dropouts.

Uses torch only as an array library so the same code runs on the CPU (tests, goldens) and on the
GPU (30M / 100M bench clouds without minutes of numpy).  Output layout mirrors pcl::PointXYZI:
float32 [n, 8] = x, y, z, 1.0 | intensity, 0, 0, 0   (32 B stride).
"""
import math

import numpy as np
import torch

SCENES = ("room", "outdoor", "planar")


def _rects_room(lo, hi):
    """6 inward faces of an axis-aligned box as (axis, offset, lo2, hi2) rectangles."""
    rects = []
    for ax in range(3):
        o = [a for a in range(3) if a != ax]
        for c in (lo[ax], hi[ax]):
            rects.append((ax, c, (lo[o[0]], lo[o[1]]), (hi[o[0]], hi[o[1]])))
    return rects


def _scene(name, scale, rng):
    """-> dict(rects=[(axis, offset, (lo_a, lo_b), (hi_a, hi_b))], spheres=[(c, r)], blobs=(centres, sigma),
    outlier_frac, noise, max_range)."""
    s = scale
    if name == "room":
        rects = _rects_room((-5 * s, -4 * s, -1.5 * s), (5 * s, 4 * s, 1.5 * s))
        rects.append((0, 3.0 * s, (-1.0 * s, -1.5 * s), (2.5 * s, 0.5 * s)))     # interior plane x=3
        rects.append((1, -2.0 * s, (-4.0 * s, -1.5 * s), (-1.0 * s, 1.0 * s)))   # interior plane y=-2
        return dict(rects=rects, spheres=[((2.0 * s, 1.0 * s, -0.5 * s), 0.5 * s)], blobs=None,
                    outlier_frac=0.0, noise=0.002, max_range=100.0 * s)
    if name == "outdoor":
        rects = [(2, -1.8 * s, (-80 * s, -80 * s), (80 * s, 80 * s))]            # ground
        fac = [(0, 12.0, (-10, -1.8), (14, 9)), (0, -18.0, (-20, -1.8), (6, 12)), (1, 15.0, (-18, -1.8), (12, 7)),
               (1, -9.0, (-6, -1.8), (10, 6)), (0, 30.0, (-30, -1.8), (30, 15)), (1, 40.0, (-40, -1.8), (40, 20))]
        for ax, c, l2, h2 in fac:
            rects.append((ax, c * s, (l2[0] * s, l2[1] * s), (h2[0] * s, h2[1] * s)))
        cen = rng.uniform(-1, 1, size=(200, 3)) * np.array([28.0, 28.0, 0.0]) + np.array([0, 0, -0.8])
        cen[:, 2] += rng.uniform(0, 2.5, size=200)
        # keep blobs out of the scanner's immediate surroundings
        too_close = np.linalg.norm(cen[:, :2], axis=1) < 3.0
        cen[too_close, :2] *= 3.0 / np.maximum(np.linalg.norm(cen[too_close, :2], axis=1, keepdims=True), 0.3)
        return dict(rects=rects, spheres=[], blobs=(cen * s, 0.3 * s), outlier_frac=0.002, noise=0.003,
                    max_range=70.0 * s)
    if name == "planar":
        rects = _rects_room((-20 * s, -20 * s, -0.3 * s), (20 * s, 20 * s, 2.2 * s))
        return dict(rects=rects, spheres=[], blobs=None, outlier_frac=0.0, noise=0.001, max_range=100.0 * s)
    raise ValueError(f"unknown scene {name!r}; choose from {SCENES}")


def make_scan(scan_w, scan_h, scene="room", seed=1, scale=1.0, dropout=0.05, device="cpu", chunk_cols=None,
              el_min_deg=-60.0, el_max_deg=90.0, duplicate=0, max_cols=None):
    """Ray-cast an organised scan.

    Returns dict(points float32[n,8] (PointXYZI layout), cloud_to_grid int32[n], scan_w, scan_h, scene).
    `max_cols` stops after that many scan columns (a file-order prefix of the same scan — the bounded CPU sample).
    `duplicate` > 0 appends that many exact copies of one point (adversarial cell occupancy, config 5)."""
    dev = torch.device(device)
    rng = np.random.default_rng(seed)
    sc = _scene(scene, scale, rng)
    g = torch.Generator(device=dev)
    g.manual_seed(int(seed) * 7919 + 17)
    if chunk_cols is None:
        chunk_cols = max(1, (262144 if dev.type == "cpu" else 2097152) // scan_h)
    el = torch.linspace(math.radians(el_min_deg), math.radians(el_max_deg), scan_h + 2, device=dev,
                        dtype=torch.float64)[1:-1]
    pts_chunks, grid_chunks = [], []
    blobs = None
    if sc["blobs"] is not None:
        blobs = (torch.as_tensor(sc["blobs"][0], device=dev, dtype=torch.float32), float(sc["blobs"][1]))
    last_col = scan_w if max_cols is None else min(scan_w, int(max_cols))
    for c0 in range(0, last_col, chunk_cols):
        c1 = min(last_col, c0 + chunk_cols)
        az = (torch.arange(c0, c1, device=dev, dtype=torch.float64) + 0.5) * (2.0 * math.pi / scan_w)
        ce, se = torch.cos(el), torch.sin(el)
        d = torch.stack([(torch.cos(az)[:, None] * ce[None, :]), (torch.sin(az)[:, None] * ce[None, :]),
                         se[None, :].expand(c1 - c0, scan_h)], dim=-1).reshape(-1, 3).to(torch.float32)
        m = d.shape[0]
        t_hit = torch.full((m,), float("inf"), device=dev)
        inten = torch.zeros((m,), device=dev)
        # rectangles (axis-aligned)
        for ridx, (ax, c, l2, h2) in enumerate(sc["rects"]):
            o = [a for a in range(3) if a != ax]
            t = c / d[:, ax]
            pa, pb = t * d[:, o[0]], t * d[:, o[1]]
            ok = (t > 1e-3) & (pa >= l2[0]) & (pa <= h2[0]) & (pb >= l2[1]) & (pb <= h2[1]) & (t < t_hit)
            t_hit = torch.where(ok, t, t_hit)
            inten = torch.where(ok, torch.abs(d[:, ax]) * (0.4 + 0.07 * (ridx % 7)), inten)
        for (cen, rad) in sc["spheres"]:
            cv = torch.tensor(cen, device=dev, dtype=torch.float32)
            b = d @ cv
            disc = b * b - (cv @ cv - rad * rad)
            t = b - torch.sqrt(torch.clamp(disc, min=0))
            ok = (disc > 0) & (t > 1e-3) & (t < t_hit)
            t_hit = torch.where(ok, t, t_hit)
            inten = torch.where(ok, torch.full_like(inten, 0.8), inten)
        if blobs is not None:
            cen, sig = blobs
            rad = 2.0 * sig
            for r0 in range(0, m, 1 << 18):
                r1 = min(m, r0 + (1 << 18))
                dd = d[r0:r1]
                b = dd @ cen.T                                        # [rays, blobs]
                disc = b * b - ((cen * cen).sum(1)[None, :] - rad * rad)
                sq = torch.sqrt(torch.clamp(disc, min=0))
                u = torch.rand((r1 - r0, cen.shape[0]), device=dev, generator=g)
                take = (disc > 0) & (u < 0.6)                         # porous: some rays pass through
                depth = torch.rand((r1 - r0, cen.shape[0]), device=dev, generator=g)
                t = torch.where(take, (b - sq) + depth * 2.0 * sq, torch.full_like(b, float("inf")))
                t = torch.where(t > 1e-3, t, torch.full_like(t, float("inf")))
                tb, _ = t.min(dim=1)
                ok = tb < t_hit[r0:r1]
                t_hit[r0:r1] = torch.where(ok, tb, t_hit[r0:r1])
                inten[r0:r1] = torch.where(ok, torch.full_like(tb, 0.25), inten[r0:r1])
        # isolated outliers: a random range in free space in front of the first hit
        if sc["outlier_frac"] > 0:
            u = torch.rand((m,), device=dev, generator=g)
            frac = torch.rand((m,), device=dev, generator=g)
            is_out = u < sc["outlier_frac"]
            t_free = torch.where(torch.isfinite(t_hit), t_hit, torch.full_like(t_hit, sc["max_range"]))
            t_hit = torch.where(is_out, 1.0 + frac * torch.clamp(t_free - 1.0, min=0.0), t_hit)
        noise = torch.randn((m,), device=dev, generator=g) * sc["noise"]
        t_n = t_hit + noise
        drop = torch.rand((m,), device=dev, generator=g) < dropout
        valid = torch.isfinite(t_hit) & (t_hit < sc["max_range"]) & (~drop) & (t_n > 1e-3)
        p = d * t_n[:, None]
        gi = torch.arange(c0 * scan_h, c1 * scan_h, device=dev, dtype=torch.int64)
        sel = torch.nonzero(valid).squeeze(1)
        out = torch.zeros((sel.numel(), 8), device=dev, dtype=torch.float32)
        out[:, 0:3] = p[sel]
        out[:, 3] = 1.0
        out[:, 4] = torch.clamp(inten[sel], 0.0, 1.0)
        pts_chunks.append(out)
        grid_chunks.append(gi[sel].to(torch.int32))
    points = torch.cat(pts_chunks, dim=0)
    c2g = torch.cat(grid_chunks, dim=0)
    if duplicate > 0 and points.shape[0] > 0:
        # exact copies of one point; they get no grid cell of their own, so they reuse the source's
        src = points[points.shape[0] // 2: points.shape[0] // 2 + 1]
        points = torch.cat([points, src.expand(duplicate, 8)], dim=0).contiguous()
        c2g = torch.cat([c2g, c2g[c2g.shape[0] // 2: c2g.shape[0] // 2 + 1].expand(duplicate)], dim=0).contiguous()
    return dict(points=points, cloud_to_grid=c2g, scan_w=scan_w, scan_h=scan_h, scene=scene)


def make_scan_np(*args, **kw):
    """make_scan on the CPU, returned as numpy arrays."""
    kw["device"] = "cpu"
    s = make_scan(*args, **kw)
    s["points"] = s["points"].numpy()
    s["cloud_to_grid"] = s["cloud_to_grid"].numpy()
    return s


# The five BASELINE.json configs restated as generator arguments.  `small` variants keep
# the scene but shrink the grid so the CPU oracle finishes in seconds.
CONFIGS = {
    1: dict(scan_w=1024, scan_h=1024, scene="room", seed=1, dropout=0.05),
    # outdoor rays that leave the scene (sky, beyond max_range) return nothing — about 57 % of the rays are valid
    # returns — so the grids are sized for the POINT counts BASELINE.json names (10 M / 30 M / 100 M valid points).
    2: dict(scan_w=5632, scan_h=3200, scene="outdoor", seed=2, dropout=0.05),
    3: dict(scan_w=10240, scan_h=5120, scene="outdoor", seed=3, dropout=0.02),
    4: dict(scan_w=18688, scan_h=9344, scene="outdoor", seed=4, dropout=0.02),
    5: dict(scan_w=8192, scan_h=4096, scene="planar", seed=5, dropout=0.02),
}


def config_scan(cfg, device="cpu", shrink=1):
    a = dict(CONFIGS[cfg])
    a["scan_w"] = max(8, a["scan_w"] // shrink)
    a["scan_h"] = max(8, a["scan_h"] // shrink)
    return make_scan(device=device, **a)


def write_ptx(path, scan, origin=(0.0, 0.0, 0.0)):
    """Write the scan in the reference's PTX dialect (pointcloud.cpp:68-115): cols, rows, origin,
    3x3 rotation, 4x4 transform, then one `x y z i` row per grid cell, `0 0 0 0` where empty."""
    pts = scan["points"]
    pts = pts.cpu().numpy() if hasattr(pts, "cpu") else pts
    c2g = scan["cloud_to_grid"]
    c2g = c2g.cpu().numpy() if hasattr(c2g, "cpu") else c2g
    w, h = scan["scan_w"], scan["scan_h"]
    grid = np.zeros((w * h, 4), np.float32)
    grid[c2g, 0:3] = pts[: c2g.shape[0], 0:3]
    grid[c2g, 3] = pts[: c2g.shape[0], 4]
    with open(path, "w") as f:
        f.write(f"{w}\n{h}\n{origin[0]:f} {origin[1]:f} {origin[2]:f}\n")
        f.write("1 0 0\n0 1 0\n0 0 1\n")
        f.write("1 0 0 0\n0 1 0 0\n0 0 1 0\n")
        f.write(f"{origin[0]:f} {origin[1]:f} {origin[2]:f} 1\n")
        for r in grid:
            if r[0] == 0 and r[1] == 0 and r[2] == 0:
                f.write("0 0 0 0\n")
            else:
                f.write(f"{r[0]:f} {r[1]:f} {r[2]:f} {r[3]:f}\n")


def read_ptx(path):
    """Read the PTX dialect back the way load_ptx does (pointcloud.cpp:149-298): `0 0 0` rows dropped,
    grid index = file order."""
    with open(path) as f:
        w = int(f.readline())
        h = int(f.readline())
        for _ in range(8):
            f.readline()
        rows = np.loadtxt(f, dtype=np.float32, ndmin=2)
    keep = ~((rows[:, 0] == 0) & (rows[:, 1] == 0) & (rows[:, 2] == 0))
    n = int(keep.sum())
    pts = np.zeros((n, 8), np.float32)
    pts[:, 0:3] = rows[keep, 0:3]
    pts[:, 3] = 1.0
    pts[:, 4] = rows[keep, 3]
    return dict(points=pts, cloud_to_grid=np.nonzero(keep)[0].astype(np.int32), scan_w=w, scan_h=h, scene="ptx")


def write_ptx(path, width, height, seed=0, rgb=False, zero_fraction=0.1, crlf=False, style="fixed", missing_lines=0,
              trailing_blank_lines=0, final_newline=True):
    """A synthetic Leica PTX file (the format PointCloud::load_ptx reads, pointcloud.cpp:127-330): 10 header lines
    (columns, rows, scanner position, 3x3 axes, 4x4 transform) and width*height body lines "x y z intensity [r g b]".
    style: "fixed" = %.6f like scanner exports, "mixed" = a mix of %.6f / %g / %e / integers / signs / padding.
    zero_fraction of the lines are "0 0 0 0.5" (no return).  Returns the number of body lines written."""
    rng = np.random.default_rng(seed)
    eol = "\r\n" if crlf else "\n"
    n = width * height - missing_lines
    xyz = rng.normal(size=(n, 3)) * np.array([12.0, 9.0, 2.5])
    inten = rng.random(n)
    zero = rng.random(n) < zero_fraction
    lines = []
    head = [f"{width}", f"{height}", "1.250000 -2.500000 0.125000", "1.000000 0.000000 0.000000", "0.000000 1.000000 0.000000",
            "0.000000 0.000000 1.000000", "1.000000 0.000000 0.000000 0.000000", "0.000000 1.000000 0.000000 0.000000",
            "0.000000 0.000000 1.000000 0.000000", "1.250000 -2.500000 0.125000 1.000000"]
    fmts = ["%.6f", "%g", "%e", "%.3f", "%+.5f", "%.9f", "%.12g"]
    for i in range(n):
        if zero[i]:
            toks = ["0", "0", "0", "0.500000"] if style == "fixed" else ["0.000000", "-0", "0e0", "0.5"]
        elif style == "fixed":
            toks = ["%.6f" % v for v in (*xyz[i], inten[i])]
        else:
            toks = [fmts[int(rng.integers(len(fmts)))] % v for v in (*xyz[i], inten[i])]
            if i % 17 == 3:
                toks[0] = str(int(xyz[i, 0] * 1000))            # integers
            if i % 29 == 5:
                toks[1] = "%.4E" % (xyz[i, 1] * 1e-7)            # upper-case exponent, small magnitude
        if rgb:
            toks += [str(int(c)) for c in rng.integers(0, 256, 3)]
        sep = " " if style == "fixed" or i % 13 else "  "
        lines.append(sep.join(toks) + (" " if style == "mixed" and not rgb and i % 11 == 1 else ""))
    body = eol.join(lines)
    text = eol.join(head) + eol + body + (eol if final_newline else "") + eol * trailing_blank_lines
    with open(path, "w", newline="") as f:
        f.write(text)
    return n
