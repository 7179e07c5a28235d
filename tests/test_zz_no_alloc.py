"""ccg_grid_normals_static (include/ccgpu.h): the scan-grid normals with their scratch in a static __device__ array.

No counterpart in the reference (added for callers that cannot call cudaMalloc).  The
file sorts last on purpose: a failure here must not hide the parity tests.

The check: through the C ABI, on two calls in a row (the second on a DIFFERENT scan, so that stale scratch of the first call
would show), the output equals ccg_grid_normals(do_fallback = 0) bit for bit — the tolerance the grid-normal parity tests
use — and the device's free memory does not move across the call."""
import numpy as np
import pytest

import cloudclean_b200 as cc
from cloudclean_b200 import synth

pytestmark = pytest.mark.gpu


def _scan(w, h, seed, dropout):
    s = synth.make_scan_np(w, h, scene="room", seed=seed, dropout=dropout)
    return np.ascontiguousarray(s["points"], np.float32), np.ascontiguousarray(s["cloud_to_grid"], np.int32)


def test_static_scratch_entry_matches_grid_normals_twice_in_a_row(gpu_ctx):
    import torch
    dev = torch.device("cuda:0")
    # (w, h, seed, dropout): a large scan first, then a smaller, sparser one that reuses the low cells of the scratch
    cases = [(96, 64, 11, 0.30), (33, 7, 12, 0.60), (96, 64, 11, 0.30)]
    nmax = 96 * 64
    # every device buffer exists before the first call; the calls themselves must not allocate
    bufs = []
    for w, h, seed, dropout in cases:
        pts, c2g = _scan(w, h, seed, dropout)
        bufs.append((torch.from_numpy(pts).to(dev), torch.from_numpy(c2g).to(dev)))
    out = torch.empty((nmax, 4), dtype=torch.float32, device=dev)
    status = torch.empty(2, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()

    for call, ((w, h, seed, dropout), (pts_d, c2g_d)) in enumerate(zip(cases, bufs)):
        n = pts_d.shape[0]
        out.fill_(7.0)                                   # poison: every row must be written by the call
        status.fill_(-9)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        gpu_ctx.grid_normals_static(pts_d, c2g_d, w, h, out, status, stream=stream.cuda_stream)
        stream.synchronize()
        free_after = torch.cuda.mem_get_info(dev)[0]
        got = out[:n].cpu().numpy()
        st = status.cpu().numpy()
        # the existing entry point on the same scan (host buffers, the context's own scratch)
        gpu_ctx.set_cloud(pts_d.cpu().numpy())
        ref, miss = gpu_ctx.grid_normals(c2g_d.cpu().numpy(), w, h, fallback=False)
        assert np.array_equal(got, ref, equal_nan=True), f"call {call}: rows differ"
        assert int(st[0]) == miss and int(st[1]) == 0, f"call {call}: status {st} vs missing {miss}"
        if call > 0:                                     # the first call loads the module (and with it the static array)
            assert free_after >= free_before, f"call {call}: free device memory fell by {free_before - free_after} bytes"


def test_static_scratch_entry_limits_and_bad_indices(gpu_ctx):
    import torch
    dev = torch.device("cuda:0")
    pts, c2g = _scan(33, 7, 3, 0.2)
    pts_d, c2g_d = torch.from_numpy(pts).to(dev), torch.from_numpy(c2g).to(dev)
    out = torch.zeros((pts.shape[0], 4), dtype=torch.float32, device=dev)
    status = torch.zeros(2, dtype=torch.int32, device=dev)
    # beyond CCG_STATIC_MAX_CELLS = 4194304 cells: a clear error, nothing launched
    with pytest.raises(cc.CcgError, match="static scratch"):
        gpu_ctx.grid_normals_static(pts_d, c2g_d, 4096, 1025, out, status)
    # an index outside the grid is reported through the status word (the call cannot synchronise to raise)
    bad = c2g_d.clone()
    bad[5] = 33 * 7
    gpu_ctx.grid_normals_static(pts_d, bad, 33, 7, out, status)
    torch.cuda.synchronize()
    assert int(status.cpu()[1]) == 1
    # NULL status is allowed
    gpu_ctx.grid_normals_static(pts_d, c2g_d, 33, 7, out, None)
    torch.cuda.synchronize()
    gpu_ctx.set_cloud(pts)
    ref, _ = gpu_ctx.grid_normals(c2g, 33, 7, fallback=False)
    assert np.array_equal(out.cpu().numpy(), ref, equal_nan=True)
