// noalloc.cu — ccg_grid_normals_static: stages (i) and (ii) of the scan-grid normals (gridnormals.cu) for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  The arithmetic is the fan of
// normalestimation.cpp:304-424 exactly as k_fan_normals runs it, so the output equals ccg_grid_normals(do_fallback = 0)
// bit for bit.
//
// The kernel's scratch — grid_to_cloud, int32 per grid cell, and two counters — lives in a file-scope __device__ array of
// CCG_STATIC_MAX_CELLS cells (16 MiB; every GPU that loads the library holds it until exit).  Host code never uses the
// array's NAME as a pointer (in host code it is a host-side stand-in): the device address comes from cudaGetSymbolAddress,
// once per device with that device current, and is passed to the kernels.  The entry point calls no cudaMalloc / cudaFree /
// allocating library routine and does not synchronise.  Every call shares the array: it is cleared at the start of each call,
// and two calls that overlap on the GPU (different streams, different threads) race.
#include "common.cuh"
#include <mutex>

namespace ccg {

constexpr size_t kStaticMaxCells = CCG_STATIC_MAX_CELLS;
constexpr int kStaticMaxDevices = 64;

// [0, kStaticMaxCells): grid_to_cloud; then 64 ints of counters ([0] normals missing, [1] cloud_to_grid index outside the grid)
static __device__ __align__(256) int32_t g_static_scratch[kStaticMaxCells + 64];

static std::mutex g_static_mu;
static int32_t *g_static_ptr[kStaticMaxDevices];       // device address of g_static_scratch, per device

// set-up, once per device, with that device current
static int static_scratch(ccg_ctx *ctx, int32_t **out) {
    const int dev = ctx->device;
    if (dev < 0 || dev >= kStaticMaxDevices) return set_error(ctx, CCG_ERR_UNSUPPORTED, "device %d outside [0,%d)", dev, kStaticMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void *p = nullptr;
        CCG_CUDA(ctx, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return set_error(ctx, CCG_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_static_ptr[dev] = static_cast<int32_t *>(p);
    }
    *out = g_static_ptr[dev];
    return CCG_OK;
}

// clears what must start clean: grid_to_cloud = -1 for the cells of this call, the counters = 0, and the status words
__global__ void __launch_bounds__(256) k_static_clear(int32_t *g2c, int len, int32_t *counters, int32_t *status2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < len) g2c[i] = -1;
    if (i < 4) counters[i] = 0;
    if (i < 2 && status2) status2[i] = 0;
}

// grid_to_cloud[cloud_to_grid[i]] = i, the largest i wins a collision (normalestimation.cpp:279-283 runs in index order)
__global__ void __launch_bounds__(256) k_static_scatter(const int32_t *__restrict__ c2g, int n, int len, int32_t *g2c, int32_t *counters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int g = c2g[i];
    if (g < 0 || g >= len) { counters[1] = 1; return; }
    atomicMax(&g2c[g], i);
}

struct P3 { float x, y, z; };
__device__ __forceinline__ P3 load_p3(const char *xyz, size_t stride, int i) {
    const float *p = reinterpret_cast<const float *>(xyz + (size_t)i * stride);
    return P3{p[0], p[1], p[2]};
}

// one thread per grid cell: the fan of triangles over the valid ring cells (normalestimation.cpp:304-424)
__global__ void __launch_bounds__(256) k_static_fan(const char *__restrict__ xyz, size_t stride, const int32_t *__restrict__ g2c, int width,
                                                    int height, float4 *__restrict__ normal4, int32_t *counters, int32_t *status2) {
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    bool missing = false;
    if (cell < width * height) {
        const int ci = g2c[cell];
        if (ci != -1) {
            const int cx = cell % width, cy = cell / width;
            // ring cell r (anticlockwise from (1,0), :293-302) as an offset; its validity as a bit
            const int rx[8] = {1, 1, 0, -1, -1, -1, 0, 1};
            const int ry[8] = {0, -1, -1, -1, 0, 1, 1, 1};
            int ring[8];
            unsigned valid = 0;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int x = cx + rx[r], y = cy + ry[r];
                ring[r] = -1;
                if (x >= 0 && x < width && y >= 0 && y < height) ring[r] = g2c[y * width + x];
                if (ring[r] != -1) valid |= 1u << r;
            }
            const P3 c = load_p3(xyz, stride, ci);
            float sx = 0.f, sy = 0.f, sz = 0.f;
            int faces = 0, next = 0, first = -1;
            for (;;) {                                             // :320-390: consecutive valid ring cells span a face
                int a = -1, b = -1;
                for (; next < 8; ++next)
                    if ((valid >> next) & 1u) { a = next++; break; }
                for (; next < 9; ++next) {
                    if (next == 8) {                               // back at ring cell 0: the fan closes only when it has two faces
                        if (faces > 1) b = first;
                        else if (faces == 0 && (valid & 1u)) b = 0;
                        break;
                    }
                    if ((valid >> next) & 1u) { b = next; break; }
                }
                if (a < 0 || b < 0) break;
                if (first < 0) first = a;
                int ia = 0, ib = 0;
#pragma unroll
                for (int r = 0; r < 8; ++r) { if (r == a) ia = ring[r]; if (r == b) ib = ring[r]; }
                const P3 pa = load_p3(xyz, stride, ia), pb = load_p3(xyz, stride, ib);
                const float ux = __fsub_rn(pa.x, c.x), uy = __fsub_rn(pa.y, c.y), uz = __fsub_rn(pa.z, c.z);
                const float wx = __fsub_rn(pb.x, c.x), wy = __fsub_rn(pb.y, c.y), wz = __fsub_rn(pb.z, c.z);
                const float kx = __fsub_rn(__fmul_rn(uy, wz), __fmul_rn(uz, wy));
                const float ky = __fsub_rn(__fmul_rn(uz, wx), __fmul_rn(ux, wz));
                const float kz = __fsub_rn(__fmul_rn(ux, wy), __fmul_rn(uy, wx));
                // Eigen 3.2 squaredNorm of a Vector3f: a0 + (a1 + a2)
                const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(kx, kx), __fadd_rn(__fmul_rn(ky, ky), __fmul_rn(kz, kz))));
                const float tx = __fdiv_rn(kx, len), ty = __fdiv_rn(ky, len), tz = __fdiv_rn(kz, len);
                if (tx != tx || ty != ty || tz != tz) continue;    // :376-382, a degenerate face is skipped
                sx = __fadd_rn(sx, tx); sy = __fadd_rn(sy, ty); sz = __fadd_rn(sz, tz);
                ++faces;
            }
            if (faces == 0) {                                      // :406-414
                normal4[ci] = make_float4(NAN, NAN, NAN, NAN);
                missing = true;
            } else {                                               // :416-422
                const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(sx, sx), __fadd_rn(__fmul_rn(sy, sy), __fmul_rn(sz, sz))));
                normal4[ci] = make_float4(__fdiv_rn(sx, len), __fdiv_rn(sy, len), __fdiv_rn(sz, len), 1.0f);
            }
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, missing);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(&counters[0], __popc(m));
    // the last block to finish publishes the counters to the caller's status words
    if (status2) {
        __shared__ bool s_last;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            s_last = atomicAdd(&counters[2], 1) == (int)gridDim.x - 1;
        }
        __syncthreads();
        if (s_last && threadIdx.x < 2) {
            __threadfence();
            status2[threadIdx.x] = atomicAdd(&counters[threadIdx.x], 0);
            if (threadIdx.x == 0) counters[2] = 0;
        }
    }
}

}  // namespace ccg

using namespace ccg;

extern "C" int ccg_grid_normals_static(ccg_ctx *ctx, void *cuda_stream, const void *xyz_dev, size_t stride_bytes, size_t n,
                                       const int32_t *cloud_to_grid_dev, int scan_w, int scan_h, float *normal4_dev,
                                       int32_t *status2_dev) {
    if (!ctx) return CCG_ERR_INVALID;
    ctx->err[0] = 0;
    if (!xyz_dev || !cloud_to_grid_dev || !normal4_dev) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (stride_bytes < 12 || (stride_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be a multiple of 4, at least 12");
    if (scan_w < 1 || scan_h < 1) return set_error(ctx, CCG_ERR_INVALID, "scan grid must be at least 1 x 1");
    const size_t len = (size_t)scan_w * (size_t)scan_h;
    if (len > kStaticMaxCells)
        return set_error(ctx, CCG_ERR_UNSUPPORTED, "scan grid %d x %d = %zu cells exceeds the static scratch of %zu cells (CCG_STATIC_MAX_CELLS)",
                         scan_w, scan_h, len, kStaticMaxCells);
    if (n > 0x7fffffffu) return set_error(ctx, CCG_ERR_UNSUPPORTED, "n = %zu exceeds 2^31 - 1", n);
    CCG_CUDA(ctx, cudaSetDevice(ctx->device));
    int32_t *scratch = nullptr;
    CCG_TRY(static_scratch(ctx, &scratch));
    cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->stream;
    int32_t *g2c = scratch, *counters = scratch + kStaticMaxCells;
    const int height = scan_w, width = scan_h;                     // sic, normalestimation.cpp:288-289
    const unsigned cells = (unsigned)((len + 255) / 256);
    k_static_clear<<<cells, 256, 0, st>>>(g2c, (int)len, counters, status2_dev);
    if (n > 0) {
        k_static_scatter<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(cloud_to_grid_dev, (int)n, (int)len, g2c, counters);
        CCG_CUDA(ctx, cudaMemsetAsync(normal4_dev, 0, n * 16, st));    // points outside every cell keep 0,0,0,0 like the reference's resize
        k_static_fan<<<cells, 256, 0, st>>>(static_cast<const char *>(xyz_dev), stride_bytes, g2c, width, height,
                                            reinterpret_cast<float4 *>(normal4_dev), counters, status2_dev);
    }
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}
