"""cloudclean_b200 — B200-native (sm_100a) kNN / radius / PCA / outlier hot path of CloudClean.

The product is `libccgpu.so` (hand-written CUDA behind the C ABI of include/ccgpu.h); the C++ adapters in
adapters/ keep the reference's call surface (getPCA, getHist, NormalEstimator, OutlierFilter::filter).
This package is only the ctypes binding used by tests/, bench.py and Python callers: it contains no
compute and no fallback — if the library or a GPU is missing, it raises.

Arrays: numpy arrays are host buffers; torch CUDA tensors are passed as device pointers (no copies).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CCG_LIBRARY") or os.path.join(_HERE, "libccgpu.so")     # CCG_LIBRARY: an experiment build of the same ABI

COV_SCATTER, COV_NM1, COV_N = 0, 1, 2
MAX_K = 64

ERR_INVALID, ERR_CUDA, ERR_NO_CLOUD, ERR_NONFINITE, ERR_UNSUPPORTED, ERR_CANCELLED = -1, -2, -3, -4, -5, -6
ERRORS = {0: "CCG_OK", -1: "CCG_ERR_INVALID", -2: "CCG_ERR_CUDA", -3: "CCG_ERR_NO_CLOUD", -4: "CCG_ERR_NONFINITE",
          -5: "CCG_ERR_UNSUPPORTED", -6: "CCG_ERR_CANCELLED"}

# every symbol include/ccgpu.h declares (tests/test_abi.py checks the header against this list and the .so)
SYMBOLS = [
    "ccg_create", "ccg_destroy", "ccg_last_error", "ccg_version", "ccg_get_stats", "ccg_set_cancel_flag", "ccg_set_stream",
    "ccg_set_cloud", "ccg_cloud_size", "ccg_get_order", "ccg_knn", "ccg_radius", "ccg_pca_knn", "ccg_pca_radius",
    "ccg_features", "ccg_eigen_scalar", "ccg_sor_mask", "ccg_radius_count_mask", "ccg_grid_normals",
    "ccg_pca_knn_range", "ccg_sor_mean_range", "ccg_unpermute", "ccg_pca_knn_features", "ccg_pca_knn_features_range",
    "ccg_pca_knn_features_scatter", "ccg_dev_alloc", "ccg_dev_free", "ccg_ipc_export", "ccg_ipc_open", "ccg_ipc_close",
    "ccg_voxel_downsample", "ccg_voxel_downsample_fetch", "ccg_pca_knn_features_shard", "ccg_knn_graph", "ccg_normal_stdev", "ccg_stdev_dist", "ccg_neighbour_hist", "ccg_don", "ccg_principal_curvatures", "ccg_fpfh", "ccg_ptx_parse", "ccg_ptx_load", "ccg_ptx_fetch", "ccg_ptx_set_cloud",
    "ccg_set_async", "ccg_sync", "ccg_pca_knn_features_route", "ccg_radius_csr", "ccg_sor_finish",
    "ccg_grid_normals_static",
]


class CcgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERRORS.get(code, code)}: {msg}")
        self.code = code


class PtxInfo(C.Structure):
    """ccg_ptx_info (include/ccgpu.h)"""
    _fields_ = [("file_width", C.c_int32), ("file_height", C.c_int32), ("scan_width", C.c_int32), ("scan_height", C.c_int32),
                ("has_rgb", C.c_int32), ("sensor_origin", C.c_float * 3), ("orientation", C.c_float * 9),
                ("lines", C.c_int64), ("sampled", C.c_int64), ("points", C.c_int64), ("body_offset", C.c_size_t)]


class Stats(C.Structure):
    _fields_ = [("ms_upload", C.c_float), ("ms_build", C.c_float), ("ms_query", C.c_float), ("ms_download", C.c_float),
                ("launches", C.c_int32), ("launches_build", C.c_int32), ("n_points", C.c_int64), ("n_leaves", C.c_int64),
                ("leaves_scanned", C.c_int64), ("deferred_queries", C.c_int64)]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


_lib = None


def load_library():
    """dlopen libccgpu.so and declare the ABI.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i32, f64, sz = C.c_void_p, C.c_int, C.c_double, C.c_size_t
    L.ccg_create.argtypes = [C.POINTER(vp), i32]
    L.ccg_destroy.argtypes = [vp]
    L.ccg_destroy.restype = None
    L.ccg_last_error.argtypes = [vp]
    L.ccg_last_error.restype = C.c_char_p
    L.ccg_version.restype = i32
    L.ccg_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.ccg_set_cancel_flag.argtypes = [vp, vp]
    L.ccg_set_stream.argtypes = [vp, vp]
    L.ccg_set_async.argtypes = [vp, i32]
    L.ccg_sync.argtypes = [vp]
    L.ccg_set_cloud.argtypes = [vp, vp, sz, sz]
    L.ccg_cloud_size.argtypes = [vp]
    L.ccg_cloud_size.restype = C.c_int64
    L.ccg_get_order.argtypes = [vp, vp]
    L.ccg_knn.argtypes = [vp, i32, vp, vp]
    L.ccg_radius.argtypes = [vp, f64, i32, i32, vp, vp, vp]
    L.ccg_pca_knn.argtypes = [vp, i32, i32, vp, vp, vp]
    L.ccg_pca_radius.argtypes = [vp, f64, i32, i32, i32, vp, vp, vp, vp]
    L.ccg_features.argtypes = [vp, vp, sz, i32, vp]
    L.ccg_eigen_scalar.argtypes = [vp, vp, sz, i32, vp]
    L.ccg_sor_mask.argtypes = [vp, i32, f64, vp, vp]
    L.ccg_radius_count_mask.argtypes = [vp, f64, i32, i32, vp]
    L.ccg_grid_normals.argtypes = [vp, vp, i32, i32, i32, vp, vp]
    L.ccg_grid_normals_static.argtypes = [vp, vp, vp, sz, sz, vp, i32, i32, vp, vp]
    L.ccg_pca_knn_range.argtypes = [vp, i32, i32, vp, sz, sz, vp, vp]
    L.ccg_sor_mean_range.argtypes = [vp, i32, sz, sz, vp]
    L.ccg_sor_finish.argtypes = [vp, vp, sz, f64, vp]
    L.ccg_unpermute.argtypes = [vp, vp, sz, vp]
    L.ccg_pca_knn_features.argtypes = [vp, i32, i32, vp, i32, vp, vp, vp]
    L.ccg_pca_knn_features_range.argtypes = [vp, i32, i32, vp, i32, sz, sz, vp, vp, vp]
    L.ccg_pca_knn_features_scatter.argtypes = [vp, i32, i32, vp, i32, sz, sz, vp, vp, vp]
    L.ccg_pca_knn_features_shard.argtypes = [vp, i32, i32, vp, i32, i32, i32, i32, i32, vp, vp, vp]
    L.ccg_pca_knn_features_route.argtypes = [vp, i32, i32, vp, i32, i32, i32, i32, sz, vp, vp, vp]
    L.ccg_radius_csr.argtypes = [vp, f64, i32, i32, vp, vp, vp, C.c_int64, C.POINTER(C.c_int64)]
    L.ccg_knn_graph.argtypes = [vp, i32, f64, vp, vp]
    L.ccg_normal_stdev.argtypes = [vp, f64, i32, vp, sz, vp]
    L.ccg_stdev_dist.argtypes = [vp, f64, i32, i32, vp, vp]
    L.ccg_neighbour_hist.argtypes = [vp, f64, i32, i32, i32, vp, sz, vp]
    L.ccg_don.argtypes = [vp, vp, sz, f64, f64, vp]
    L.ccg_principal_curvatures.argtypes = [vp, f64, vp, sz, vp]
    L.ccg_fpfh.argtypes = [vp, f64, vp, sz, vp, vp]
    L.ccg_ptx_parse.argtypes = [vp, vp, sz, i32, C.POINTER(PtxInfo)]
    L.ccg_ptx_load.argtypes = [vp, C.c_char_p, i32, C.POINTER(PtxInfo)]
    L.ccg_ptx_fetch.argtypes = [vp, vp, vp]
    L.ccg_ptx_set_cloud.argtypes = [vp]
    L.ccg_dev_alloc.argtypes = [vp, sz, C.POINTER(vp)]
    L.ccg_dev_free.argtypes = [vp, vp]
    L.ccg_ipc_export.argtypes = [vp, vp, vp]
    L.ccg_ipc_open.argtypes = [vp, vp, C.POINTER(vp)]
    L.ccg_ipc_close.argtypes = [vp, vp]
    L.ccg_voxel_downsample.argtypes = [vp, vp, sz, i32, C.c_float, i32, vp, C.POINTER(C.c_int64)]
    L.ccg_voxel_downsample_fetch.argtypes = [vp, vp]
    _lib = L
    return L


def _is_torch(a):
    return a is not None and type(a).__module__.startswith("torch")


def _addr(a):
    """tensor / array / DevBuffer / raw integer device address -> c_void_p"""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if _is_torch(a) or isinstance(a, np.ndarray):          # (torch.Tensor has an unrelated .addr method)
        return _ptr(a)
    if hasattr(a, "addr"):
        return C.c_void_p(a.addr)
    return _ptr(a)


def _ptr(a):
    if a is None:
        return None
    if _is_torch(a):
        assert a.is_contiguous()
        return C.c_void_p(a.data_ptr())
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def _rowptr(a):
    """first byte of a [n, S] float32 array whose rows may be strided (a column-sliced view of a wider struct array): the C ABI
    takes the row stride separately, only the elements of a row must be adjacent"""
    if a is None:
        return None
    if _is_torch(a):
        assert a.dim() == 1 or a.stride(-1) == 1
        return C.c_void_p(a.data_ptr())
    assert a.ndim == 1 or a.strides[-1] == a.itemsize
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One GPU, one cloud.  Thin wrapper over ccg_ctx; every method is one C-ABI call."""

    def __init__(self, device=0):
        self.L = load_library()
        h = C.c_void_p()
        rc = self.L.ccg_create(C.byref(h), int(device))
        if rc != 0:
            raise CcgError(rc, "ccg_create failed (no usable CUDA device?) — there is no CPU fallback")
        self.h = h
        self.device = int(device)
        self.n = 0
        self._keep = None

    def close(self):
        if getattr(self, "h", None):
            self.L.ccg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise CcgError(rc, self.L.ccg_last_error(self.h).decode())

    def stats(self):
        s = Stats()
        self._check(self.L.ccg_get_stats(self.h, C.byref(s)))
        return s.as_dict()

    def _out(self, like_device, shape, dtype):
        """allocate an output next to the inputs: torch CUDA tensor if the cloud came from one, else numpy"""
        if like_device:
            import torch
            if getattr(self, "_own_stream", True):
                # the ctx stream is not ordered after torch's: a block the caching allocator hands out may still be read or
                # written by work queued on torch's current stream
                torch.cuda.current_stream(self.device).synchronize()
            tdt = {np.float32: torch.float32, np.int32: torch.int32, np.uint8: torch.uint8, np.float64: torch.float64}[dtype]
            return torch.empty(shape, dtype=tdt, device=f"cuda:{self.device}")
        return np.empty(shape, dtype)

    def set_stream(self, cuda_stream):
        """run later calls on `cuda_stream` (an int cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream); None = own"""
        self._check(self.L.ccg_set_stream(self.h, C.c_void_p(cuda_stream) if cuda_stream else None))
        self._own_stream = not cuda_stream

    def set_async(self, on=True):
        """entry points enqueue on the ctx stream and return; results / deferred errors / stats are valid after sync()"""
        self._check(self.L.ccg_set_async(self.h, 1 if on else 0))

    def sync(self):
        self._check(self.L.ccg_sync(self.h))

    def _settle(self, *tensors):
        """The ctx's own stream is non-blocking, i.e. NOT ordered after torch's streams: when a torch CUDA tensor is an
        INPUT and no shared stream was set, wait for torch's current stream so pending writes to it have landed."""
        if getattr(self, "_own_stream", True) and any(_is_torch(t) and t.is_cuda for t in tensors):
            import torch
            torch.cuda.current_stream(self.device).synchronize()

    # ---- cloud -------------------------------------------------------------------------------------
    def set_cloud(self, points):
        """points: float32 [n, S>=3] (x, y, z first; S=8 mirrors pcl::PointXYZI).  numpy (host) or torch CUDA."""
        if _is_torch(points):
            assert points.dtype.is_floating_point and points.element_size() == 4 and points.dim() == 2
            assert points.is_contiguous() and points.shape[1] >= 3
            n, stride = points.shape[0], points.shape[1] * 4
            self.on_device = points.is_cuda
            if not points.is_cuda:
                points = points.numpy()
        else:
            points = np.ascontiguousarray(points, np.float32)
            assert points.ndim == 2 and points.shape[1] >= 3
            n, stride = points.shape[0], points.strides[0]
            self.on_device = False
        self._settle(points)
        self._check(self.L.ccg_set_cloud(self.h, _ptr(points), n, stride))
        self.n = n
        return self

    def order(self):
        o = np.empty(self.n, np.int32)
        self._check(self.L.ccg_get_order(self.h, _ptr(o)))
        return o

    # ---- a1 ----------------------------------------------------------------------------------------
    def knn(self, k, want_d2=True, device_out=None):
        dev = self.on_device if device_out is None else device_out
        idx = self._out(dev, (self.n, k), np.int32)
        d2 = self._out(dev, (self.n, k), np.float32) if want_d2 else None
        self._check(self.L.ccg_knn(self.h, k, _ptr(idx), _ptr(d2)))
        return idx, d2

    # ---- a2 / a3 -----------------------------------------------------------------------------------
    def radius(self, r, max_nn=0, strict_lt=True, lists=True, device_out=None):
        dev = self.on_device if device_out is None else device_out
        count = self._out(dev, (self.n,), np.int32)
        idx = d2 = None
        if lists:
            idx = self._out(dev, (self.n, max_nn), np.int32)
            d2 = self._out(dev, (self.n, max_nn), np.float32)
        self._check(self.L.ccg_radius(self.h, float(r), max_nn, 1 if strict_lt else 0, _ptr(idx), _ptr(d2), _ptr(count)))
        return {"count": count, "idx": idx, "d2": d2}

    # ---- a12 ---------------------------------------------------------------------------------------
    def pca_knn(self, k, cov_norm=COV_NM1, viewpoint=None, want_normals=True, device_out=None):
        dev = self.on_device if device_out is None else device_out
        eig = self._out(dev, (self.n, 3), np.float32)
        nrm = self._out(dev, (self.n, 4), np.float32) if want_normals else None
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn(self.h, k, cov_norm, _ptr(vp), _ptr(eig), _ptr(nrm)))
        return eig, nrm

    def pca_knn_into(self, k, cov_norm, eig, nrm, viewpoint=None):
        """same, into caller-provided buffers (host numpy or torch CUDA) — what bench.py times"""
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn(self.h, k, cov_norm, _ptr(vp), _ptr(eig), _ptr(nrm)))

    def pca_knn_features_into(self, k, cov_norm, formula_set, eig, nrm, feat, viewpoint=None):
        """kNN + PCA + fused eigen features into caller buffers (host numpy or torch CUDA); any output may be None"""
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn_features(self.h, k, cov_norm, _ptr(vp), formula_set, _ptr(eig), _ptr(nrm), _ptr(feat)))

    def pca_knn_features_range(self, k, cov_norm, formula_set, begin, end, eig_sorted, nrm_sorted, feat_sorted, viewpoint=None):
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn_features_range(self.h, k, cov_norm, _ptr(vp), formula_set, begin, end, _ptr(eig_sorted),
                                                      _ptr(nrm_sorted), _ptr(feat_sorted)))

    def pca_knn_features_scatter(self, k, cov_norm, formula_set, begin, end, eig, nrm, feat, viewpoint=None):
        """range query whose rows land in ORIGINAL order in eig/nrm/feat — tensors, DevBuffers or raw device addresses,
        possibly peer memory of the collecting GPU (fused query + NVLink scatter, include/ccgpu.h)"""
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn_features_scatter(self.h, k, cov_norm, _ptr(vp), formula_set, begin, end, _addr(eig),
                                                        _addr(nrm), _addr(feat)))

    def pca_knn_features_shard(self, k, cov_norm, formula_set, rank, world, eig, nrm, feat, chunk_points=8192, sorted_rows=False,
                               viewpoint=None):
        """interleaved shard of the query (chunks rank, rank+world, ... of the Morton order), rows in ORIGINAL order into
        own or peer device memory — the load-balanced form of pca_knn_features_scatter.  sorted_rows=True: rows at
        their Morton position instead (coalesced NVLink stores; the collector un-permutes)"""
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn_features_shard(self.h, k, cov_norm, _ptr(vp), formula_set, rank, world, chunk_points,
                                                      1 if sorted_rows else 0, _addr(eig), _addr(nrm), _addr(feat)))

    def pca_knn_features_route(self, k, cov_norm, formula_set, rank, world, rows_per_rank, eig_of_rank, nrm_of_rank, feat_of_rank,
                               chunk_points=8192, viewpoint=None):
        """answer this rank's interleaved chunks and store every row into the arrays of the rank that OWNS its original
        index (lists of device addresses / DevBuffers / tensors, one per rank; own memory or ccg_ipc_open'ed peer memory)"""
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)

        def table(lst):
            if lst is None:
                return None
            arr = (C.c_void_p * world)()
            for i, a in enumerate(lst):
                arr[i] = _addr(a).value
            return arr
        te, tn, tf = table(eig_of_rank), table(nrm_of_rank), table(feat_of_rank)
        self._check(self.L.ccg_pca_knn_features_route(self.h, k, cov_norm, _ptr(vp), formula_set, rank, world, chunk_points,
                                                      rows_per_rank, te, tn, tf))

    def normal_stdev(self, r, max_nn, normals, device_out=None):
        """normal_stdev (cv.h:123-206): normals = [n, S>=3] float32 (host numpy or torch CUDA) -> float32[n]"""
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n,), np.float32)
        stride = normals.stride(0) * 4 if _is_torch(normals) else normals.strides[0]
        self._settle(normals)
        self._check(self.L.ccg_normal_stdev(self.h, float(r), max_nn, _rowptr(normals), stride, _ptr(out)))
        return out

    def stdev_dist(self, r, max_nn, use_depth=False, centre=None, device_out=None):
        """stdev_dist (cv.cpp:166-230) -> float32[n]"""
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n,), np.float32)
        c = None if centre is None else np.ascontiguousarray(centre, np.float32)
        self._check(self.L.ccg_stdev_dist(self.h, float(r), max_nn, 1 if use_depth else 0, _ptr(c), _ptr(out)))
        return out

    def neighbour_hist(self, r, max_nn, bins, attr=None, device_out=None):
        """calcDistHist (attr=None) / calcIntensityHist (attr = float32[n] or a strided column view, host numpy or torch
        CUDA) of featureeval/utils.h -> float32[n, bins]"""
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n, bins), np.float32)
        stride = 0
        if attr is not None:
            stride = attr.stride(0) * 4 if _is_torch(attr) else attr.strides[0]
            self._settle(attr)
        p = None if attr is None else (C.c_void_p(attr.data_ptr()) if _is_torch(attr) else attr.ctypes.data_as(C.c_void_p))
        self._check(self.L.ccg_neighbour_hist(self.h, float(r), max_nn, bins, 0 if attr is None else 1, p, stride, _ptr(out)))
        return out

    def don(self, normals, r1=0.2, r2=0.5, device_out=None):
        """difference of normals (featureeval.cpp:694-739): normals [n, S>=3] float32 -> float32[n, 4]"""
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n, 4), np.float32)
        stride = normals.stride(0) * 4 if _is_torch(normals) else normals.strides[0]
        self._settle(normals)
        self._check(self.L.ccg_don(self.h, _rowptr(normals), stride, float(r1), float(r2), _ptr(out)))
        return out

    def principal_curvatures(self, normals, r, device_out=None):
        """pcl::PrincipalCurvaturesEstimation, radius search (markov.cpp:389-402, featureeval.cpp:947-956):
        normals [n, S>=3] float32 -> float32[n, 5] = (pcx, pcy, pcz, pc1, pc2)"""
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n, 5), np.float32)
        stride = normals.stride(0) * 4 if _is_torch(normals) else normals.strides[0]
        self._settle(normals)
        self._check(self.L.ccg_principal_curvatures(self.h, float(r), _rowptr(normals), stride, _ptr(out)))
        return out

    def fpfh(self, normals, r, want_spfh=False, device_out=None):
        """pcl::FPFHEstimation, radius search (featureeval.cpp:925-932): normals [n, S>=3] float32 -> float32[n, 33]
        (and the SPFH rows [n, 33] when want_spfh)"""
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n, 33), np.float32)
        sp = self._out(dev, (self.n, 33), np.float32) if want_spfh else None
        stride = normals.stride(0) * 4 if _is_torch(normals) else normals.strides[0]
        self._settle(normals)
        self._check(self.L.ccg_fpfh(self.h, float(r), _rowptr(normals), stride, _ptr(out), _ptr(sp)))
        return (out, sp) if want_spfh else out

    def knn_graph(self, k, inverse_sigma, device_out=None):
        """kNN table + MinCut binary potentials (graphcut/mincut.cpp:458-475,574-585): (idx int32[n,k], weight float64[n,k])"""
        dev = self.on_device if device_out is None else device_out
        idx = self._out(dev, (self.n, k), np.int32)
        w = self._out(dev, (self.n, k), np.float64)
        self._check(self.L.ccg_knn_graph(self.h, k, float(inverse_sigma), _ptr(idx), _ptr(w)))
        return idx, w

    def voxel_downsample(self, points, resolution, n_floats=None, as_written=True, device_out=None):
        """octreeDownsample (utils.h:16-62): points = the [n, S] float32 array given to set_cloud ->
        (means [nvox, n_floats], big_to_small int32[n])"""
        nf = points.shape[1] if n_floats is None else n_floats
        stride = points.stride(0) * 4 if _is_torch(points) else points.strides[0]
        dev = self.on_device if device_out is None else device_out
        b2s = self._out(dev, (self.n,), np.int32)
        nv = C.c_int64(0)
        self._settle(points)
        self._check(self.L.ccg_voxel_downsample(self.h, _rowptr(points), stride, nf, float(resolution), 1 if as_written else 0,
                                                _ptr(b2s), C.byref(nv)))
        out = self._out(dev, (nv.value, nf), np.float32)
        if nv.value:
            self._check(self.L.ccg_voxel_downsample_fetch(self.h, _ptr(out)))
        return out, b2s

    # ---- f4: PTX ingest (PointCloud::load_ptx, pointcloud.cpp:127-330) ---------------------------------
    def _ptx_result(self, info, device_out):
        n = int(info.points)
        pts = self._out(device_out, (n, 8), np.float32)
        c2g = self._out(device_out, (n,), np.int32)
        if n:
            self._check(self.L.ccg_ptx_fetch(self.h, _ptr(pts), _ptr(c2g)))
        return {"points": pts, "cloud_to_grid": c2g, "file_width": info.file_width, "file_height": info.file_height,
                "scan_width": info.scan_width, "scan_height": info.scan_height, "has_rgb": bool(info.has_rgb),
                "origin": np.array(info.sensor_origin[:], np.float32), "matrix": np.array(info.orientation[:], np.float32),
                "lines": int(info.lines), "sampled": int(info.sampled), "body_offset": int(info.body_offset)}

    def ptx_parse(self, data, decimation_factor=1, device_out=False):
        """data: the file's bytes (bytes / numpy uint8 / torch uint8 CUDA tensor) -> dict (points [n, 8], cloud_to_grid, header)"""
        info = PtxInfo()
        if isinstance(data, (bytes, bytearray, memoryview)):
            data = np.frombuffer(data, np.uint8)
        nbytes = data.numel() if _is_torch(data) else data.size
        self._settle(data)
        self._check(self.L.ccg_ptx_parse(self.h, _ptr(data), nbytes, decimation_factor, C.byref(info)))
        return self._ptx_result(info, device_out)

    def ptx_load(self, filename, decimation_factor=1, device_out=False):
        """load_ptx(filename, decimation_factor): mmap + GPU parse"""
        info = PtxInfo()
        self._check(self.L.ccg_ptx_load(self.h, str(filename).encode(), decimation_factor, C.byref(info)))
        return self._ptx_result(info, device_out)

    def ptx_set_cloud(self, n):
        """make the rows of the last ptx_parse / ptx_load the ctx's cloud (device-resident, no host trip)"""
        self._check(self.L.ccg_ptx_set_cloud(self.h))
        self.n = n
        self.on_device = True
        return self

    def dev_alloc(self, nbytes):
        p = C.c_void_p()
        self._check(self.L.ccg_dev_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def dev_free(self, addr):
        self._check(self.L.ccg_dev_free(self.h, C.c_void_p(addr)))

    def ipc_export(self, addr):
        h = (C.c_ubyte * 64)()
        self._check(self.L.ccg_ipc_export(self.h, C.c_void_p(addr), h))
        return bytes(h)

    def ipc_open(self, handle):
        p = C.c_void_p()
        buf = (C.c_ubyte * 64).from_buffer_copy(handle)
        self._check(self.L.ccg_ipc_open(self.h, buf, C.byref(p)))
        return p.value

    def ipc_close(self, addr):
        self._check(self.L.ccg_ipc_close(self.h, C.c_void_p(addr)))

    def radius_csr(self, r, max_nn=0, strict_lt=True, want_d2=True, device_out=None):
        """variable-length radius lists (any cap, 0 = uncapped): (offsets int64[n+1], idx int32[total], d2 float32[total]),
        every row ascending (d2, index) — what radiusSearch fills into its vectors"""
        dev = self.on_device if device_out is None else device_out
        _np = np
        if dev:
            import torch
            off = torch.empty((self.n + 1,), dtype=torch.int64, device=f"cuda:{self.device}")
        else:
            off = np.empty(self.n + 1, np.int64)
        tot = C.c_int64(0)
        self._check(self.L.ccg_radius_csr(self.h, float(r), max_nn, 1 if strict_lt else 0, _ptr(off), None, None, 0, C.byref(tot)))
        idx = self._out(dev, (tot.value,), _np.int32)
        d2 = self._out(dev, (tot.value,), _np.float32) if want_d2 else None
        if tot.value:
            self._check(self.L.ccg_radius_csr(self.h, float(r), max_nn, 1 if strict_lt else 0, _ptr(off), _ptr(idx), _ptr(d2), tot.value,
                                              C.byref(tot)))
        return off, idx, d2

    # ---- a4 / a5 -----------------------------------------------------------------------------------
    def pca_radius(self, r, max_nn=0, cov_norm=COV_NM1, normalise_hist=False, viewpoint=None, want_normals=False,
                   device_out=None):
        """getPCA (cv.h:76-116) when normalise_hist=False, getHist (cv.cpp:249-308) when True."""
        dev = self.on_device if device_out is None else device_out
        eig = self._out(dev, (self.n, 3), np.float32)
        nrm = self._out(dev, (self.n, 4), np.float32) if want_normals else None
        count = self._out(dev, (self.n,), np.int32)
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_radius(self.h, float(r), max_nn, cov_norm, 1 if normalise_hist else 0, _ptr(vp),
                                          _ptr(eig), _ptr(nrm), _ptr(count)))
        return eig, nrm, count

    # ---- a9 ----------------------------------------------------------------------------------------
    def features(self, eig3, formula_set=0):
        n = eig3.shape[0]
        out = self._out(_is_torch(eig3) and eig3.is_cuda, (n, 7), np.float32)
        self._settle(eig3)
        self._check(self.L.ccg_features(self.h, _ptr(eig3), n, formula_set, _ptr(out)))
        return out

    def eigen_scalar(self, eig3, which):
        n = eig3.shape[0]
        out = self._out(_is_torch(eig3) and eig3.is_cuda, (n,), np.float32)
        self._settle(eig3)
        self._check(self.L.ccg_eigen_scalar(self.h, _ptr(eig3), n, which, _ptr(out)))
        return out

    # ---- a11 / a10 ---------------------------------------------------------------------------------
    def sor_mask(self, mean_k, stddev_mul, want_mean_d=True, device_out=None):
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n,), np.uint8)
        md = self._out(dev, (self.n,), np.float32) if want_mean_d else None
        self._check(self.L.ccg_sor_mask(self.h, mean_k, float(stddev_mul), _ptr(out), _ptr(md)))
        return out, md

    def radius_count_mask(self, r, min_neighbours, prefilter=False, device_out=None):
        dev = self.on_device if device_out is None else device_out
        out = self._out(dev, (self.n,), np.uint8)
        self._check(self.L.ccg_radius_count_mask(self.h, float(r), min_neighbours, 1 if prefilter else 0, _ptr(out)))
        return out

    # ---- a7 ----------------------------------------------------------------------------------------
    def grid_normals(self, cloud_to_grid, scan_w, scan_h, fallback=True, device_out=None):
        dev = self.on_device if device_out is None else device_out
        if not _is_torch(cloud_to_grid):
            cloud_to_grid = np.ascontiguousarray(cloud_to_grid, np.int32)
        out = self._out(dev, (self.n, 4), np.float32)
        self._settle(cloud_to_grid)
        missing = C.c_int32(0)
        self._check(self.L.ccg_grid_normals(self.h, _ptr(cloud_to_grid), scan_w, scan_h, 1 if fallback else 0, _ptr(out),
                                            C.byref(missing)))
        return out, missing.value

    def grid_normals_static(self, xyz_dev, cloud_to_grid_dev, scan_w, scan_h, normal4_dev, status2_dev=None, stream=None):
        """ccg_grid_normals_static: stages (i) + (ii) on DEVICE tensors, nothing allocated, nothing synchronised — the
        caller waits on `stream` (an integer cudaStream_t; None = the context's stream).  The wrapper allocates nothing either."""
        n = int(xyz_dev.shape[0])
        stride = int(xyz_dev.stride(0)) * xyz_dev.element_size()
        self._check(self.L.ccg_grid_normals_static(self.h, stream, xyz_dev.data_ptr(), stride, n, cloud_to_grid_dev.data_ptr(),
                                                   scan_w, scan_h, normal4_dev.data_ptr(),
                                                   None if status2_dev is None else status2_dev.data_ptr()))

    # ---- sharding ----------------------------------------------------------------------------------
    def pca_knn_range(self, k, cov_norm, begin, end, eig_sorted, nrm_sorted, viewpoint=None):
        vp = None if viewpoint is None else np.ascontiguousarray(viewpoint, np.float32)
        self._check(self.L.ccg_pca_knn_range(self.h, k, cov_norm, _ptr(vp), begin, end, _ptr(eig_sorted), _ptr(nrm_sorted)))

    def sor_mean_range(self, mean_k, begin, end, mean_d_sorted):
        self._check(self.L.ccg_sor_mean_range(self.h, mean_k, begin, end, _ptr(mean_d_sorted)))

    def sor_finish(self, mean_d, stddev_mul, outlier):
        """gathered mean distances (original order, host or device) -> outlier mask into `outlier` (uint8[n], host or device)"""
        self._settle(mean_d)
        self._check(self.L.ccg_sor_finish(self.h, _ptr(mean_d), mean_d.shape[0], float(stddev_mul), _ptr(outlier)))

    def unpermute(self, rows_sorted, rows_original):
        row_bytes = rows_sorted.shape[1] * 4 if rows_sorted.ndim == 2 else 4
        if _is_torch(rows_sorted):
            row_bytes = (rows_sorted.shape[1] if rows_sorted.dim() == 2 else 1) * rows_sorted.element_size()
        else:
            row_bytes = (rows_sorted.shape[1] if rows_sorted.ndim == 2 else 1) * rows_sorted.itemsize
        self._settle(rows_sorted)
        self._check(self.L.ccg_unpermute(self.h, _ptr(rows_sorted), row_bytes, _ptr(rows_original)))
