// gridnormals.cu — scan-grid triangle-fan normals (a7).
//
// Replaces NormalEstimator::estimateNormals (src/plugins/normalestimation/normalestimation.cpp:268-470):
//   (i)   grid_to_cloud scatter                              (:279-283)
//   (ii)  per grid cell: fan of triangles over the valid 8-neighbours, sum of the normalised cross
//         products, normalise; no face -> NaN x4 + remember   (:304-424)
//   (iii) remembered points, in grid order: kNN(k=50) mean of the neighbours' non-NaN normals, written in
//         place so later points see earlier fixes              (:428-459)
// One thread per grid cell for (ii) — a 3x3 stencil through grid_to_cloud (HBM/L2 gather bound, 196
// algorithmic bytes per point); (iii) uses the warp-per-query kNN kernel and one warp that walks the
// (few) remembered points in order.  All arithmetic is fp32 round-to-nearest without fma, so stage (ii)
// is bit-identical to the oracle's restatement.
#include "common.cuh"
#include <cub/device/device_select.cuh>

namespace ccg {

__global__ void __launch_bounds__(256) k_fill_i32(int32_t *p, size_t n, int32_t v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// grid_to_cloud[cloud_to_grid[i]] = i; the reference's sequential loop lets the largest i win a collision
__global__ void __launch_bounds__(256) k_scatter_g2c(const int32_t *__restrict__ c2g, int n, int len, int32_t *g2c, int *bad) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int g = c2g[i];
    if (g < 0 || g >= len) { *bad = 1; return; }
    atomicMax(&g2c[g], i);
}

__device__ __forceinline__ bool fan_valid(int dx, int dy, int width, int height, int cx, int cy, const int32_t *g2c) {
    int x = cx + dx, y = cy + dy;                              // normalestimation.cpp:246-266
    if (x > width - 1 || x < 0) return false;
    if (y > height - 1 || y < 0) return false;
    return g2c[y * width + x] != -1;
}

__global__ void __launch_bounds__(256) k_fan_normals(const float4 *__restrict__ pts, const int32_t *__restrict__ g2c,
                                                     int width, int height, float4 *__restrict__ normal4,
                                                     uint8_t *__restrict__ missing_flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int length = width * height;
    if (i >= length) return;
    const int ci = g2c[i];
    missing_flag[i] = 0;
    if (ci == -1) return;
    const int RX[8] = {1, 1, 0, -1, -1, -1, 0, 1};             // :293-302, anticlockwise from (1,0)
    const int RY[8] = {0, -1, -1, -1, 0, 1, 1, 1};
    const int cx = i % width, cy = i / width;
    // validity of the 8 ring cells as a bit mask
    unsigned vmask = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) vmask |= fan_valid(RX[r], RY[r], width, height, cx, cy, g2c) ? (1u << r) : 0u;
    const float4 p0 = pts[ci];
    float ax = 0.f, ay = 0.f, az = 0.f;
    int face_count = 0, next = 0, first = -1;
    while (true) {                                             // :320-390, same control flow
        int p1 = -1, p2 = -1;
        while (next < 8) {
            if ((vmask >> next) & 1u) { p1 = next++; break; }
            next++;
        }
        while (next < 9) {
            if (next == 8 && face_count == 1) break;
            if (next == 8 && face_count > 1) { p2 = first; break; }
            int id = next % 8;
            if ((vmask >> id) & 1u) { p2 = id; break; }
            next++;
        }
        if (p1 == -1 || p2 == -1) break;
        if (first == -1) first = p1;
        const float4 q1 = pts[g2c[(cy + RY[p1]) * width + (cx + RX[p1])]];
        const float4 q2 = pts[g2c[(cy + RY[p2]) * width + (cx + RX[p2])]];
        const float v1x = __fsub_rn(q1.x, p0.x), v1y = __fsub_rn(q1.y, p0.y), v1z = __fsub_rn(q1.z, p0.z);
        const float v2x = __fsub_rn(q2.x, p0.x), v2y = __fsub_rn(q2.y, p0.y), v2z = __fsub_rn(q2.z, p0.z);
        const float cxp = __fsub_rn(__fmul_rn(v1y, v2z), __fmul_rn(v1z, v2y));
        const float cyp = __fsub_rn(__fmul_rn(v1z, v2x), __fmul_rn(v1x, v2z));
        const float czp = __fsub_rn(__fmul_rn(v1x, v2y), __fmul_rn(v1y, v2x));
        // Eigen 3.2 squaredNorm of a Vector3f: fixed-size redux halves the range, a0 + (a1 + a2) [EXT]
        const float nn = __fsqrt_rn(__fadd_rn(__fmul_rn(cxp, cxp), __fadd_rn(__fmul_rn(cyp, cyp), __fmul_rn(czp, czp))));
        const float tx = __fdiv_rn(cxp, nn), ty = __fdiv_rn(cyp, nn), tz = __fdiv_rn(czp, nn);
        if (tx != tx || ty != ty || tz != tz) continue;        // :376-382
        ax = __fadd_rn(ax, tx); ay = __fadd_rn(ay, ty); az = __fadd_rn(az, tz);
        face_count++;
    }
    if (face_count == 0) {                                     // :406-414
        normal4[ci] = make_float4(NAN, NAN, NAN, NAN);
        missing_flag[i] = 1;
        return;
    }
    const float nn = __fsqrt_rn(__fadd_rn(__fmul_rn(ax, ax), __fadd_rn(__fmul_rn(ay, ay), __fmul_rn(az, az))));
    normal4[ci] = make_float4(__fdiv_rn(ax, nn), __fdiv_rn(ay, nn), __fdiv_rn(az, nn), 1.0f);   // :416-422
}

// stage (iii): one warp walks the remembered points in order (sequentially dependent, :436-459)
__global__ void __launch_bounds__(32) k_fallback_seq(const int32_t *__restrict__ missing, int nm, const int32_t *__restrict__ nbr,
                                                     int k, float4 *normal4, int *still_missing) {
    const int lane = threadIdx.x;
    int still = nm;
    for (int m = 0; m < nm; ++m) {
        const int idx = missing[m];
        float sx = 0.f, sy = 0.f, sz = 0.f;
        int count = 0;
        for (int j0 = 0; j0 < k; j0 += 32) {
            const int j = j0 + lane;
            float4 nb = make_float4(NAN, NAN, NAN, NAN);
            if (j < k) {
                const int id = nbr[(size_t)m * k + j];
                if (id >= 0) nb = __ldcg(&normal4[id]);
            }
            const int lim = min(32, k - j0);
            for (int t = 0; t < lim; ++t) {                    // neighbour order, like the reference's range-for
                const float x = __shfl_sync(kFull, nb.x, t), y = __shfl_sync(kFull, nb.y, t), z = __shfl_sync(kFull, nb.z, t);
                if (x != x) continue;
                ++count;
                sx = __fadd_rn(sx, x); sy = __fadd_rn(sy, y); sz = __fadd_rn(sz, z);
            }
        }
        if (count > 0) {
            if (lane == 0) {
                const float c = (float)count;
                float4 old = __ldcg(&normal4[idx]);
                __stcg(&normal4[idx], make_float4(__fdiv_rn(sx, c), __fdiv_rn(sy, c), __fdiv_rn(sz, c), old.w));
            }
            --still;
        }
        __threadfence_block();
        __syncwarp();
    }
    if (lane == 0) *still_missing = still;
}

// c2g, normal4: device pointers.  Returns the number of normals still missing through *n_missing_host.
int launch_grid_normals(ccg_ctx *ctx, const int32_t *c2g, int scan_w, int scan_h, int do_fallback, float *normal4,
                        int32_t *n_missing_host) {
    const size_t n = ctx->n;
    cudaStream_t st = ctx->stream;
    const int height = scan_w, width = scan_h;                 // sic, normalestimation.cpp:288-289
    const size_t len = (size_t)width * (size_t)height;
    if (n_missing_host) *n_missing_host = 0;
    if (n == 0) return CCG_OK;
    if (len == 0 || len > 0x7fffffff) return set_error(ctx, CCG_ERR_INVALID, "scan grid %d x %d is empty or too large", scan_w, scan_h);
    // tmp[0]: g2c int32[len]; tmp[1]: missing flags u8[len]; tmp[2]: missing list int32[len] + count; tmp[3]: select scratch / nbr rows
    CCG_CUDA(ctx, ctx->tmp[0].reserve(len * 4));
    CCG_CUDA(ctx, ctx->tmp[1].reserve(len));
    CCG_CUDA(ctx, ctx->tmp[2].reserve(len * 4 + 64));
    int32_t *g2c = ctx->tmp[0].as<int32_t>();
    uint8_t *flags = ctx->tmp[1].as<uint8_t>();
    int32_t *missing = ctx->tmp[2].as<int32_t>();
    int *d_counts = (int *)((char *)ctx->scratch.p + 2048);   // [0] bad flag, [1] n selected, [2] still missing
    CCG_CUDA(ctx, cudaMemsetAsync(d_counts, 0, 16, st));
    k_fill_i32<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(g2c, len, -1);
    k_scatter_g2c<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(c2g, (int)n, (int)len, g2c, d_counts);
    CCG_CUDA(ctx, cudaMemsetAsync(normal4, 0, n * 16, st));
    k_fan_normals<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(ctx->orig.as<float4>(), g2c, width, height,
                                                                 reinterpret_cast<float4 *>(normal4), flags);
    ctx->stats.launches += 3;
    // ordered compaction of the cloud indices whose fan had no face (grid order, like missing_normals.push_back)
    size_t sel_bytes = 0;
    CCG_CUDA(ctx, cub::DeviceSelect::Flagged(nullptr, sel_bytes, g2c, flags, missing, d_counts + 1, (int)len, st));
    CCG_CUDA(ctx, ctx->sort_tmp.reserve(sel_bytes));
    CCG_CUDA(ctx, cub::DeviceSelect::Flagged(ctx->sort_tmp.p, sel_bytes, g2c, flags, missing, d_counts + 1, (int)len, st));
    ctx->stats.launches += 2;
    int h_counts[4] = {0, 0, 0, 0};
    CCG_CUDA(ctx, cudaMemcpyAsync(h_counts, d_counts, 16, cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaStreamSynchronize(st));
    if (h_counts[0]) return set_error(ctx, CCG_ERR_INVALID, "cloud_to_grid holds an index outside the %d x %d scan grid", scan_w, scan_h);
    int nm = h_counts[1];
    int still = nm;
    if (do_fallback && nm > 0) {
        const int k = 50;                                      // normalestimation.cpp:430
        CCG_CUDA(ctx, ctx->tmp[3].reserve((size_t)nm * k * 4));
        PointArgs a{};
        a.queries = missing; a.nq = nm; a.mode = MODE_KNN; a.k = k; a.rb = INFINITY; a.list_rows = 1;
        a.idx = ctx->tmp[3].as<int32_t>();
        CCG_TRY(launch_point_query(ctx, a));
        k_fallback_seq<<<1, 32, 0, st>>>(missing, nm, ctx->tmp[3].as<int32_t>(), k, reinterpret_cast<float4 *>(normal4), d_counts + 2);
        ctx->stats.launches++;
        CCG_CUDA(ctx, cudaMemcpyAsync(&still, d_counts + 2, 4, cudaMemcpyDeviceToHost, st));
        CCG_CUDA(ctx, cudaStreamSynchronize(st));
    }
    if (n_missing_host) *n_missing_host = still;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

}  // namespace ccg
