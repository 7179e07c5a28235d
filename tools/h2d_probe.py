"""Developer probe: pinned-host -> device upload of the xyz(w) half of a 32-byte PointXYZI array — contiguous copy of
everything (960 MB at 30 M points) against a 2-D copy of the 16 used bytes per point (cudaMemcpy2DAsync, width 16, pitch 32)."""
import ctypes, sys, torch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30_000_000
rt = ctypes.CDLL("libcudart.so.12")
h = torch.empty((n, 8), dtype=torch.float32, pin_memory=True); h.uniform_()
d_full = torch.empty((n, 8), device="cuda"); d_half = torch.empty((n, 4), device="cuda")
st = torch.cuda.current_stream().cuda_stream
def timed(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
rt.cudaMemcpy2DAsync.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
ms = timed(lambda: rt.cudaMemcpyAsync(d_full.data_ptr(), h.data_ptr(), n * 32, 1, st))
print(f"contiguous 32 B/pt: {ms:.2f} ms  {n*32/ms/1e6:.1f} GB/s")
ms = timed(lambda: rt.cudaMemcpy2DAsync(d_half.data_ptr(), 16, h.data_ptr(), 32, 16, n, 1, st))
print(f"2-D 16 of 32 B/pt: {ms:.2f} ms  {n*16/ms/1e6:.1f} GB/s (useful bytes)")
ok = torch.equal(d_half.cpu(), h[:, :4])
print("2-D copy correct:", ok)
# chunked: 8 chunks
def chunked():
    c = (n + 7) // 8
    for i in range(8):
        a, b = i * c, min(n, (i + 1) * c)
        rt.cudaMemcpy2DAsync(d_half.data_ptr() + a * 16, 16, h.data_ptr() + a * 32, 32, 16, b - a, 1, st)
print(f"2-D in 8 chunks: {timed(chunked):.2f} ms")
