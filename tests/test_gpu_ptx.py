"""f4 — PTX ingest: the GPU parser (ccg_ptx_parse / ccg_ptx_load) against the oracle's stdio restatement of
PointCloud::load_ptx (pointcloud.cpp:127-330) on the same files, bit for bit; the decimal -> float conversion against
glibc's strtof on random and adversarial tokens."""
import os
from decimal import Decimal, getcontext
from fractions import Fraction

import numpy as np
import pytest

import cloudclean_b200 as cc
import oracle
from cloudclean_b200 import synth

pytestmark = pytest.mark.gpu


def _same(got, want):
    assert got["file_width"] == want["file_width"] and got["file_height"] == want["file_height"]
    assert got["has_rgb"] == want["has_rgb"]
    assert got["sampled"] == want["sampled"]
    assert np.array_equal(got["origin"].view(np.int32), want["origin"].view(np.int32))
    assert np.array_equal(got["matrix"].view(np.int32), want["matrix"].view(np.int32))
    assert got["points"].shape == want["points"].shape
    assert np.array_equal(got["points"].view(np.int32), want["points"].view(np.int32))
    assert np.array_equal(got["cloud_to_grid"], want["cloud_to_grid"])


@pytest.mark.parametrize("rgb", [False, True])
@pytest.mark.parametrize("style,crlf", [("fixed", False), ("mixed", False), ("fixed", True), ("mixed", True)])
@pytest.mark.parametrize("dec", [1, 2, 4])
def test_ptx_matches_the_stdio_restatement(gpu_ctx, tmp_path, rgb, style, crlf, dec):
    path = tmp_path / "scan.ptx"
    synth.write_ptx(path, 37, 50, seed=dec + 7 * rgb, rgb=rgb, crlf=crlf, style=style)
    got = gpu_ctx.ptx_load(path, dec)
    want = oracle.load_ptx(path, dec)
    _same(got, want)
    assert got["scan_width"] == 37 // dec and got["scan_height"] == 50 // dec
    assert got["lines"] == 37 * 50
    assert 0 < got["points"].shape[0] <= got["sampled"]
    # rows are pcl::PointXYZI as load_ptx leaves them
    p = got["points"]
    assert np.all(p[:, 3] == 1.0) and np.all(p[:, 5] == 0.0)
    assert np.array_equal(p[:, 6], np.arange(p.shape[0], dtype=np.float32))
    assert np.array_equal(p[:, 7], got["cloud_to_grid"].astype(np.float32))
    assert not np.any((p[:, 0] == 0) & (p[:, 1] == 0) & (p[:, 2] == 0))


def test_ptx_short_file_blank_tail_and_missing_final_newline(gpu_ctx, tmp_path):
    for kw in (dict(trailing_blank_lines=3), dict(final_newline=False)):
        path = tmp_path / "scan.ptx"
        written = synth.write_ptx(path, 40, 25, seed=3, **kw)
        got = gpu_ctx.ptx_load(path, 1)
        _same(got, oracle.load_ptx(path, 1))
        assert got["lines"] == written
    # a file that ends before the last sampled line: the reference returns false (pointcloud.cpp:232-235), so do both sides
    for kw, dec in ((dict(missing_lines=123), 1), (dict(missing_lines=5, final_newline=False), 1), (dict(missing_lines=81), 2),
                    (dict(missing_lines=1), 1), (dict(missing_lines=41, final_newline=False), 2)):
        path = tmp_path / "short.ptx"
        synth.write_ptx(path, 40, 25, seed=3, **kw)
        with pytest.raises(IOError):
            oracle.load_ptx(path, dec)
        with pytest.raises(cc.CcgError) as ei:
            gpu_ctx.ptx_load(path, dec)
        assert ei.value.code == cc.ERR_INVALID and "File read fail" in str(ei.value)
    # ... but lines missing after the last sampled one do not matter: 40 x 25 with decimation 2 samples up to line 23*40+39
    # and ONE missing last sampled line slips through as a failed fscanf when end-of-file was not reported before it
    # (:240-246): the cloud is a point short.  rgb, decimation 1: the last fgets stops at the newline, no EOF yet.
    for kw, dec in ((dict(missing_lines=40), 2), (dict(missing_lines=41), 2), (dict(missing_lines=1, rgb=True), 1)):
        path = tmp_path / "short_ok.ptx"
        synth.write_ptx(path, 40, 25, seed=3, **kw)
        _same(gpu_ctx.ptx_load(path, dec), oracle.load_ptx(path, dec))
    # more lines than the header announces: the surplus is ignored (the loop runs file_width*file_height times)
    path = tmp_path / "long.ptx"
    synth.write_ptx(path, 40, 25, seed=4)
    text = path.read_text().replace("40\n25\n", "40\n20\n", 1)
    path.write_text(text)
    got = gpu_ctx.ptx_load(path, 1)
    _same(got, oracle.load_ptx(path, 1))
    assert got["sampled"] == 800 and got["lines"] == 1000


def test_ptx_from_memory_and_device_bytes_and_set_cloud(gpu_ctx, tmp_path):
    import torch
    path = tmp_path / "scan.ptx"
    synth.write_ptx(path, 64, 64, seed=9, zero_fraction=0.2)
    raw = path.read_bytes()
    want = oracle.load_ptx(path, 1)
    _same(gpu_ctx.ptx_parse(raw, 1), want)
    dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
    got = gpu_ctx.ptx_parse(dev, 1, device_out=True)
    assert got["points"].is_cuda
    assert np.array_equal(got["points"].cpu().numpy().view(np.int32), want["points"].view(np.int32))
    # parsed rows -> cloud without a host trip: same kNN-PCA as the oracle-parsed points through set_cloud
    n = got["points"].shape[0]
    gpu_ctx.ptx_set_cloud(n)
    e1, n1 = gpu_ctx.pca_knn(12, device_out=False)
    gpu_ctx.set_cloud(want["points"])
    e2, n2 = gpu_ctx.pca_knn(12, device_out=False)
    assert np.array_equal(e1.view(np.int32), e2.view(np.int32)) and np.array_equal(n1.view(np.int32), n2.view(np.int32))


def _ptx_of_tokens(path, tokens):
    """a PTX whose body carries the tokens four per line (padded with 1s); x of line 0 is kept non-zero"""
    assert len(tokens) % 4 == 0
    n = len(tokens) // 4
    head = [str(n), "1", "0 0 0", "1 0 0", "0 1 0", "0 0 1", "1 0 0 0", "0 1 0 0", "0 0 1 0", "0 0 0 1"]
    lines = [" ".join(tokens[4 * i:4 * i + 4]) for i in range(n)]
    with open(path, "w") as f:
        f.write("\n".join(head + lines) + "\n")
    return n


def _random_tokens(rng, count):
    toks = []
    for _ in range(count):
        nd = int(rng.integers(1, 20))
        digits = "".join(str(int(d)) for d in rng.integers(0, 10, nd))
        digits = str(int(rng.integers(1, 10))) + digits[1:]                  # no leading zero: value is never 0
        kind = int(rng.integers(0, 4))
        if kind == 0:
            pos = int(rng.integers(0, nd + 1))
            t = digits[:pos] + "." + digits[pos:]
            if t.startswith("."):
                t = ("0" if rng.random() < 0.5 else "") + t
        elif kind == 1:
            t = digits[:10]
        else:
            e = int(rng.integers(-30, 20)) - (nd - 1)
            e = max(min(e, 37 - nd), -36)
            t = digits[0] + ("." + digits[1:] if nd > 1 else "") + ("e" if kind == 2 else "E") + ("%+d" % e if rng.random() < 0.7 else str(e))
        if rng.random() < 0.4:
            t = "-" + t
        elif rng.random() < 0.1:
            t = "+" + t
        toks.append(t)
    return toks


def _midpoint_tokens(rng, count):
    """decimal strings at, just below and just above the midpoint of two adjacent floats (19 significant digits)"""
    getcontext().prec = 60
    toks = []
    while len(toks) < count:
        f = np.float32(rng.normal() * 10.0 ** float(rng.integers(-6, 9)))
        if f == 0 or not np.isfinite(f):
            continue
        f = abs(f)
        g = np.nextafter(f, np.float32(np.inf))
        mid = (Fraction(float(f)) + Fraction(float(g))) / 2
        d = Decimal(mid.numerator) / Decimal(mid.denominator)
        s = format(d, ".18E")                              # 19 significant digits, rounded
        lo = format(d * (1 - Decimal(10) ** -18), ".18E")
        hi = format(d * (1 + Decimal(10) ** -18), ".18E")
        toks += [s, lo, hi]
    # exact ties with few digits: floats in [2^23, 2^24) are spaced 1 apart, in [2^24, 2^25) 2 apart
    toks += ["8388608.5", "8388609.5", "16777217", "16777219", "16777217.000001", "16777216.999999", "33554434", "33554438",
             "0.1", "0.3", "1e-37", "2.9e38", "123456789012345678", "9999999999999999999", "1234567890123456789e-19"]
    return toks


@pytest.mark.parametrize("force_exact", [False, True])
def test_decimal_conversion_equals_strtof(gpu_ctx, tmp_path, monkeypatch, force_exact):
    """every conversion correctly rounded: random tokens (1..19 digits, exponents) and midpoints of adjacent floats; with
    CCG_PTX_FORCE_EXACT=1 every number is decided by the wide-integer comparison instead of the double estimate"""
    if force_exact:
        monkeypatch.setenv("CCG_PTX_FORCE_EXACT", "1")
    rng = np.random.default_rng(11 + force_exact)
    toks = _random_tokens(rng, 40000) + _midpoint_tokens(rng, 6000)
    toks = toks[: len(toks) - len(toks) % 4]
    path = tmp_path / "tokens.ptx"
    n = _ptx_of_tokens(path, toks)
    got = gpu_ctx.ptx_load(path, 1)
    want = oracle.strtof_batch(toks).reshape(n, 4)
    assert got["points"].shape[0] == n                     # no token is zero, so no row is dropped
    g = np.concatenate([got["points"][:, :3], got["points"][:, 4:5]], axis=1)
    bad = np.nonzero(g.view(np.int32) != want.view(np.int32))
    assert bad[0].size == 0, [(toks[4 * i + j], g[i, j], want[i, j]) for i, j in zip(*bad)][:5]
    _same(got, oracle.load_ptx(path, 1))


def test_ptx_refuses_what_it_cannot_parse_exactly(gpu_ctx, tmp_path):
    path = tmp_path / "scan.ptx"
    synth.write_ptx(path, 20, 10, seed=1, zero_fraction=0.0)
    good = path.read_text().split("\n")
    for bad_line, code in (("1.0 2.0 abc 0.5", cc.ERR_INVALID), ("1.0 2.0 3.0", cc.ERR_INVALID), ("nan 1 2 0.5", cc.ERR_INVALID),
                           ("1.0 2.0 3.0 0.5 7", cc.ERR_INVALID), ("1e-45 1 2 0.5", cc.ERR_UNSUPPORTED),
                           ("1.00000000000000000001 1 2 0.5", cc.ERR_UNSUPPORTED)):
        lines = list(good)
        lines[10 + 57] = bad_line
        path.write_text("\n".join(lines))
        with pytest.raises(cc.CcgError) as ei:
            gpu_ctx.ptx_load(path, 1)
        assert ei.value.code == code, (bad_line, str(ei.value))
        assert "line 57" in str(ei.value) and "file line 68" in str(ei.value)
        # the same bad line is harmless when the decimation rule skips it (the reference fgets()s over it)
        got = gpu_ctx.ptx_load(path, 2)
        _same(got, oracle.load_ptx(path, 2))
    with pytest.raises(cc.CcgError):
        gpu_ctx.ptx_load(path, 3)                              # pointcloud.cpp:130: 1 or even
    with pytest.raises(cc.CcgError):
        gpu_ctx.ptx_load(tmp_path / "missing.ptx", 1)
    # the ctx stays usable
    path.write_text("\n".join(good))
    _same(gpu_ctx.ptx_load(path, 1), oracle.load_ptx(path, 1))


def test_ptx_lines_longer_than_the_staged_tail(gpu_ctx, tmp_path):
    """lines padded with kilobytes of trailing blanks (legal for fscanf: whitespace) run past the shared-memory image
    of their 16 KB segment and are parsed from global memory: same rows as the reference restatement"""
    path = tmp_path / "scan.ptx"
    synth.write_ptx(path, 40, 60, seed=5)
    lines = path.read_text().split("\n")
    for i in range(10, len(lines) - 1, 37):
        lines[i] = lines[i] + " " * (700 + 13 * (i % 50)) + ("\t" if i % 2 else "")
    path.write_text("\n".join(lines))
    _same(gpu_ctx.ptx_load(path, 1), oracle.load_ptx(path, 1))
    # with decimation the reference skips lines with fgets(buff, 1024) (pointcloud.cpp:228), which splits a line longer
    # than 1023 characters in two and loses step; the GPU loader counts real lines: compare against the unpadded file
    want = None
    clean = tmp_path / "clean.ptx"
    synth.write_ptx(clean, 40, 60, seed=5)
    want = oracle.load_ptx(clean, 2)
    _same(gpu_ctx.ptx_load(path, 2), want)


def test_ptx_flood_of_blank_lines(gpu_ctx, tmp_path):
    """thousands of consecutive blank lines (whitespace to fscanf) put more line starts into one 15 KB segment than its
    shared list holds: the kernel's per-thread path must give the same rows"""
    path = tmp_path / "scan.ptx"
    synth.write_ptx(path, 40, 60, seed=6)
    lines = path.read_text().split("\n")
    lines[700:700] = [""] * 5000
    lines[1500:1500] = [" ", "\t", ""] * 900
    path.write_text("\n".join(lines))
    got = gpu_ctx.ptx_load(path, 1)
    _same(got, oracle.load_ptx(path, 1))
    assert got["lines"] == 2400


def test_ptx_device_bytes_at_odd_addresses(gpu_ctx, tmp_path):
    """the segment grid is laid on 16-byte aligned ADDRESSES: a device buffer that starts at any byte offset (a view into
    a larger tensor) must parse to the same rows"""
    import torch
    path = tmp_path / "scan.ptx"
    synth.write_ptx(path, 50, 41, seed=8, style="mixed")
    raw = np.frombuffer(path.read_bytes(), np.uint8)
    want = oracle.load_ptx(path, 1)
    for off in (1, 2, 3, 5, 8, 11, 15, 16, 31):
        buf = torch.full((off + raw.size + 64,), ord("\n"), dtype=torch.uint8, device="cuda")      # newlines around: must not leak in
        buf[off:off + raw.size] = torch.from_numpy(raw.copy()).cuda()
        got = gpu_ctx.ptx_parse(buf[off:off + raw.size], 1)
        _same(got, want)


def test_ptx_golden(gpu_ctx):
    """the CUDA parser against the committed golden vectors (exact-rational float32 rounding, no libc)"""
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g = np.load(os.path.join(here, "scan12x9_ptx.npz"))
    for dec in (1, 2):
        r = gpu_ctx.ptx_load(os.path.join(here, "scan12x9.ptx"), dec)
        assert (r["file_width"], r["file_height"]) == tuple(g["size"])
        assert r["sampled"] == int(g[f"sampled_dec{dec}"])
        assert np.array_equal(r["points"].view(np.int32), g[f"points_dec{dec}"].view(np.int32))
        assert np.array_equal(r["cloud_to_grid"], g[f"c2g_dec{dec}"])
        assert np.array_equal(r["origin"], g["origin"]) and np.array_equal(r["matrix"], g["matrix"])
