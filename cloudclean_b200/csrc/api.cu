// api.cu — the extern "C" surface of libccgpu.so (include/ccgpu.h): argument checking, host/device pointer
// detection, output staging, and the few O(n) glue kernels (SOR mask, voxel prefilter).
#include "tile.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <stdarg.h>
#include <stdlib.h>
#include <new>

namespace ccg {

int set_error(ccg_ctx *ctx, int code, const char *fmt, ...) {
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

static bool is_dev(const ccg_ctx *ctx, const void *p) {
    if (!p) return false;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// An output the caller wants at `user` (host or device, may be NULL): kernels write to dev().
struct OutBuf {
    ccg_ctx *ctx; void *user; size_t bytes; void *d; bool staged;
    OutBuf() : ctx(nullptr), user(nullptr), bytes(0), d(nullptr), staged(false) {}
    int open(ccg_ctx *c, int slot, void *user_ptr, size_t nbytes) {
        ctx = c; user = user_ptr; bytes = nbytes; d = nullptr; staged = false;
        if (!user_ptr || nbytes == 0) return CCG_OK;
        if (is_dev(c, user_ptr)) { d = user_ptr; return CCG_OK; }
        CCG_CUDA(c, c->out[slot].reserve(nbytes));
        d = c->out[slot].p; staged = true;
        return CCG_OK;
    }
    template <class T> T *dev() const { return reinterpret_cast<T *>(d); }
    int close() {
        if (staged) CCG_CUDA(ctx, cudaMemcpyAsync(user, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        return CCG_OK;
    }
};

static int begin_call(ccg_ctx *ctx, bool need_cloud) {
    if (!ctx) return CCG_ERR_INVALID;
    ctx->err[0] = 0;
    CCG_CUDA(ctx, cudaSetDevice(ctx->device));
    if (need_cloud && !ctx->has_cloud)
        return set_error(ctx, CCG_ERR_NO_CLOUD, "no cloud: call ccg_set_cloud first");
    if (ctx->cancel && *ctx->cancel) return set_error(ctx, CCG_ERR_CANCELLED, "cancelled by caller");
    ctx->stats.launches = 0;
    ctx->stats.leaves_scanned = 0;
    return CCG_OK;
}
static int query_begin(ccg_ctx *ctx) { CCG_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream)); return CCG_OK; }
static int query_end(ccg_ctx *ctx) { CCG_TRY(join_side(ctx)); CCG_CUDA(ctx, cudaEventRecord(ctx->ev[4], ctx->stream)); return CCG_OK; }
static int finish_call(ccg_ctx *ctx) {
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[5], ctx->stream));
    if (ctx->async_mode) { ctx->pending_stats = 1; return CCG_OK; }      // ccg_sync collects
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    CCG_CUDA(ctx, cudaGetLastError());
    cudaEventElapsedTime(&ctx->stats.ms_query, ctx->ev[3], ctx->ev[4]);
    cudaEventElapsedTime(&ctx->stats.ms_download, ctx->ev[4], ctx->ev[5]);
    return CCG_OK;
}

// strict: d2 < (float)(r*r)  <=>  d2 <= pred((float)(r*r));  inclusive: (double)d2 <= r*r  <=>  d2 <= rounddown(r*r)
static float radius_bound(double r, int strict_lt) {
    double r2 = r * r;
    float f = (float)r2;
    if (strict_lt) return f > 0.f ? nextafterf(f, -INFINITY) : -1.f;
    if ((double)f > r2) f = nextafterf(f, -INFINITY);
    return f;
}

// ---- glue kernels --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_sor_mask(const float *__restrict__ mean_d, size_t n, double thr, uint8_t *__restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)mean_d[i] > thr ? 1 : 0;
}
__global__ void __launch_bounds__(256) k_count_mask(const int32_t *__restrict__ count, const uint8_t *__restrict__ cand, size_t n,
                                                    int min_neighbours, uint8_t *__restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ((cand == nullptr || cand[i]) && count[i] < min_neighbours) ? 1 : 0;
}
// after sorting by voxel key: candidate[i] = (points in i's voxel) < min_neighbours
__global__ void __launch_bounds__(256) k_voxel_candidates(const unsigned long long *__restrict__ keys, const unsigned *__restrict__ vals,
                                                          size_t n, int min_neighbours, uint8_t *__restrict__ cand) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long k = keys[i];
    // a voxel is sparse iff the run of equal keys around i is shorter than min_neighbours: look at most that far
    int run = 1;
    for (long long j = (long long)i - 1; j >= 0 && run < min_neighbours && keys[j] == k; --j) ++run;
    for (size_t j = i + 1; j < n && run < min_neighbours && keys[j] == k; ++j) ++run;
    cand[vals[i]] = run < min_neighbours ? 1 : 0;
}

// ---- output pipelining ---------------------------------------------------------------------------------
// With HOST output pointers the D2H of the per-point rows (56 B/point for eig3+normal4+feat7) costs as much as
// the query itself.  Rows are in ORIGINAL order while the kernel works in Morton tiles, so a prefix of rows is
// complete only when every tile holding one of them has run.  Ordering the tiles by their smallest original
// index makes that a prefix property: after the first T tiles of that order every row below the (T+1)-th
// tile's smallest index is final.  The tile list is cut into chunks; the copy stream ships rows
// [bound[c], bound[c+1]) while chunk c+1 is being computed.  For a scan in file order (column-major grid) the
// row front follows the tile front closely; for a shuffled cloud the bounds stay near 0 and the last copy
// moves everything — still correct, just not overlapped.
__global__ void __launch_bounds__(256) k_tile_minidx(const float4 *__restrict__ sorted, int nleaf, unsigned *__restrict__ key,
                                                     int32_t *__restrict__ tile) {
    const int leaf = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (leaf >= nleaf) return;
    const int w = __float_as_int(sorted[(size_t)leaf * kLeaf + lane].w);       // pads carry INT_MAX
    const int m = __reduce_min_sync(kFull, w);
    if (lane == 0) { key[leaf] = (unsigned)m; tile[leaf] = leaf; }
}

struct PipeOut { OutBuf *o; size_t row_bytes; };

static int pipelined_query(ccg_ctx *ctx, QueryArgs a, PipeOut *outs, int nouts, int chunks, size_t *bound_out = nullptr,
                           int *nch_out = nullptr) {
    const int nleaf = ctx->nleaf;
    const size_t n = ctx->n;
    cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
    if (chunks > 16) chunks = 16;
    for (int c = 0; c <= 16; ++c) {
        if (!ctx->ev_chunk[c]) CCG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_chunk[c], cudaEventDisableTiming));
        if (!ctx->ev_copy[c]) CCG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_copy[c], cudaEventDisableTiming));
    }
    // tiles ordered by their smallest original index
    const size_t stride = ((size_t)nleaf + 63) & ~(size_t)63;
    CCG_CUDA(ctx, ctx->tmp[3].reserve(stride * 16));
    unsigned *k0 = ctx->tmp[3].as<unsigned>(), *k1 = k0 + stride;
    int32_t *v0 = (int32_t *)(k1 + stride), *v1 = v0 + stride;
    k_tile_minidx<<<(nleaf + 7) / 8, 256, 0, st>>>(ctx->sorted.as<float4>(), nleaf, k0, v0);
    int bits = 1;
    while (bits < 32 && (n >> bits)) ++bits;
    cub::DoubleBuffer<unsigned> dk(k0, k1);
    cub::DoubleBuffer<int32_t> dv(v0, v1);
    size_t tmp_bytes = 0;
    CCG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, nleaf, 0, bits, st));
    CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
    CCG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tmp_bytes, dk, dv, nleaf, 0, bits, st));
    ctx->stats.launches += 1 + (bits + 7) / 8 + 1;
    // chunk c = slots [slot[c], slot[c+1]); rows below bound[c+1] are final once it has run
    // `chunks` equal chunks over the first 88 % of the tiles, then three short ones (6, 4, 2 %): what cannot overlap is the
    // copy of the LAST chunk's rows, so the last chunk is kept small without paying launch tails on all of them
    if (chunks > 11) chunks = 11;
    int slot[18]; size_t bound[18];
    int nch = 0;
    slot[0] = 0;
    {
        const bool taper = nleaf >= 64;
        // two short chunks first (2 %, 6 %: the copy stream starts early), `chunks` equal ones up to 88 %, three short ones last
        const int head[2] = {taper ? (int)((long long)nleaf * 2 / 100) & ~1 : 0, taper ? (int)((long long)nleaf * 8 / 100) & ~1 : 0};
        for (int i = 0; i < 2; ++i)
            if (head[i] > slot[nch]) { slot[nch + 1] = head[i]; ++nch; }
        const int main_tiles = taper ? (int)((long long)nleaf * 88 / 100) : nleaf;
        const int main_left = main_tiles - slot[nch];
        int per = ((main_left + chunks - 1) / chunks + 1) & ~1;
        if (per < 2) per = 2;
        while (slot[nch] < main_tiles) { slot[nch + 1] = slot[nch] + per < main_tiles ? slot[nch] + per : main_tiles; ++nch; }
        if (taper) {
            const int cut[3] = {(int)((long long)nleaf * 94 / 100) & ~1, (int)((long long)nleaf * 98 / 100) & ~1, nleaf};
            slot[nch] &= ~1;                                  // chunk starts stay even: two tiles per block
            for (int i = 0; i < 3; ++i)
                if (cut[i] > slot[nch]) { slot[nch + 1] = cut[i]; ++nch; }
        }
    }
    CCG_CUDA(ctx, ctx->hpin[1].reserve(64 * sizeof(unsigned)));
    unsigned *hb = ctx->hpin[1].as<unsigned>();
    for (int c = 1; c < nch; ++c)
        CCG_CUDA(ctx, cudaMemcpyAsync(hb + c, dk.Current() + slot[c], sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaStreamSynchronize(st));
    bound[0] = 0;
    for (int c = 1; c < nch; ++c) bound[c] = hb[c] < n ? hb[c] : n;
    bound[nch] = n;
    a.tile_order = dv.Current();
    a.async_side = 1;
    int launches = ctx->stats.launches;
    // An early exit must not leave copies of earlier chunks in flight: the caller may free its (pinned) output buffers as
    // soon as the call returns, and the next call may overwrite ctx->out[] while the copy stream still reads it.
    auto bail = [&](int rc) {
        cudaStreamSynchronize(cs);
        cudaStreamSynchronize(st);
        if (ctx->side_stream) cudaStreamSynchronize(ctx->side_stream);
        ctx->side_pending = 0;
        return rc;
    };
    // CCG_PIPELINE_TRACE=1 (measurement only): when each chunk's kernels and copies ended, relative to the first launch
    const bool trace = getenv("CCG_PIPELINE_TRACE") != nullptr;
    cudaEvent_t tr0 = nullptr, trk[17] = {}, trc[17] = {};
    if (trace) {
        cudaEventCreate(&tr0);
        for (int c = 0; c < nch; ++c) { cudaEventCreate(&trk[c]); cudaEventCreate(&trc[c]); }
        cudaEventRecord(tr0, st);
    }
    for (int c = 0; c < nch; ++c) {
        if (ctx->cancel && *ctx->cancel) return bail(set_error(ctx, CCG_ERR_CANCELLED, "cancelled by caller"));
        a.begin = (size_t)slot[c] * kLeaf;
        a.end = (size_t)slot[c + 1] * kLeaf;
        ctx->stats.launches = 0;
        { const int rc = launch_tile_query(ctx, a); if (rc != CCG_OK) return bail(rc); }
        launches += ctx->stats.launches;
        CCG_CUDA(ctx, cudaEventRecord(ctx->ev_chunk[c], st));
        CCG_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_chunk[c], 0));
        if (ctx->side_pending)                                       // this chunk's deferred queries (point kernel, side stream)
            for (int i = 0; i < 2; ++i)
                if (ctx->ev_side[i]) CCG_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_side[i], 0));
        const size_t r0 = bound[c], r1 = bound[c + 1];
        if (r1 > r0)
            for (int i = 0; i < nouts; ++i)
                CCG_CUDA(ctx, cudaMemcpyAsync((char *)outs[i].o->user + r0 * outs[i].row_bytes, (char *)outs[i].o->d + r0 * outs[i].row_bytes,
                                              (r1 - r0) * outs[i].row_bytes, cudaMemcpyDeviceToHost, cs));
        CCG_CUDA(ctx, cudaEventRecord(ctx->ev_copy[c], cs));       // rows below bound[c + 1] are on the host after this
        if (trace) { cudaEventRecord(trk[c], st); cudaEventRecord(trc[c], cs); }
    }
    if (trace) {
        cudaStreamSynchronize(st); cudaStreamSynchronize(cs);
        size_t row_bytes = 0;
        for (int i = 0; i < nouts; ++i) row_bytes += outs[i].row_bytes;
        for (int c = 0; c < nch; ++c) {
            float tk = 0.f, tc = 0.f;
            cudaEventElapsedTime(&tk, tr0, trk[c]); cudaEventElapsedTime(&tc, tr0, trc[c]);
            fprintf(stderr, "[ccg pipeline] chunk %2d tiles %8d..%8d kernels end %7.3f ms, rows %10zu..%10zu (%7.1f MB) on the host at %7.3f ms\n", c,
                    slot[c], slot[c + 1], tk, bound[c], bound[c + 1], (double)(bound[c + 1] - bound[c]) * row_bytes / 1e6, tc);
            cudaEventDestroy(trk[c]); cudaEventDestroy(trc[c]);
        }
        cudaEventDestroy(tr0);
    }
    if (bound_out) for (int c = 0; c <= nch; ++c) bound_out[c] = bound[c];
    if (nch_out) *nch_out = nch;
    ctx->stats.launches = launches;
    for (int i = 0; i < nouts; ++i) outs[i].o->staged = false;          // already shipped: close() has nothing left to do
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev_chunk[16], cs));
    CCG_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_chunk[16], 0));
    return CCG_OK;
}

// f3: MinCut::calculateBinaryPotential (graphcut/mincut.cpp:574-585) for every kNN edge: the three squared float
// differences are accumulated in double, scaled by inverse_sigma and passed through exp(-x) in double.
__global__ void __launch_bounds__(256) k_binary_potential(const float4 *__restrict__ orig, const int32_t *__restrict__ idx, size_t n, int k,
                                                          double inverse_sigma, double *__restrict__ weight) {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * (size_t)k) return;
    const int j = idx[e];
    if (j < 0) { weight[e] = 0.0; return; }
    const float4 a = orig[e / (size_t)k], b = orig[j];
    const float dx = __fsub_rn(a.x, b.x), dy = __fsub_rn(a.y, b.y), dz = __fsub_rn(a.z, b.z);
    double d = 0.0;
    d += (double)__fmul_rn(dx, dx);
    d += (double)__fmul_rn(dy, dy);
    d += (double)__fmul_rn(dz, dz);
    d *= inverse_sigma;
    weight[e] = exp(-d);
}

// f2: epilogues over capped radius neighbour lists (rows of `k` original indices, ascending (d2, index), `count` valid).
// normal_stdev (cv.h:123-206), arithmetic as written there: cosine = dot / |neighbour| * |current| (sic, precedence),
// clamp to [0,1], angle = float(acosf(cosine) / M_PI) [double division], float sums in neighbour order,
// std = float(sqrt(double(sumsq / m) - pow(double(sum / m), 2))); NaN normals, < 3 neighbours and NaN results give 0.
__global__ void __launch_bounds__(256) k_normal_stdev(const int32_t *__restrict__ idx, const int32_t *__restrict__ count, size_t n, int k,
                                                      const char *__restrict__ normals, size_t stride, float *__restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *cur = reinterpret_cast<const float *>(normals + i * stride);
    const float cx = cur[0], cy = cur[1], cz = cur[2];
    const int m = count[i];
    float res = 0.f;
    if (cx == cx && m >= 3) {
        const float cn = __fsqrt_rn(eigen_dot3(cx, cy, cz, cx, cy, cz));
        float sum = 0.f, sumsq = 0.f;
        for (int j = 0; j < m; ++j) {
            const float *nb = reinterpret_cast<const float *>(normals + (size_t)idx[i * k + j] * stride);
            const float x = nb[0], y = nb[1], z = nb[2];
            const float dot = eigen_dot3(x, y, z, cx, cy, cz);
            const float nn = __fsqrt_rn(eigen_dot3(x, y, z, x, y, z));
            float cosine = __fmul_rn(__fdiv_rn(dot, nn), cn);
            cosine = cosine < 0.f ? 0.f : (cosine > 1.f ? 1.f : cosine);
            const float angle = (float)((double)acosf(cosine) / 3.14159265358979323846);
            sum = __fadd_rn(sum, angle);
            sumsq = __fadd_rn(sumsq, __fmul_rn(angle, angle));
        }
        const float fm = (float)m;
        const double mean = (double)__fdiv_rn(sum, fm);
        const float sd = (float)sqrt((double)__fdiv_rn(sumsq, fm) - mean * mean);
        if (sd == sd) res = sd;
    }
    out[i] = res;
}
// stdev_dist (cv.cpp:166-230): float sums of the neighbour distances (to the query, or to `centre` when use_depth),
// result = (sum_sq - sum*sum/m) / (m-1) for m > 2 else 0 — a VARIANCE despite the name, kept as written.
__global__ void __launch_bounds__(256) k_stdev_dist(const float4 *__restrict__ orig, const int32_t *__restrict__ idx,
                                                    const int32_t *__restrict__ count, size_t n, int k, int use_depth, float3 centre,
                                                    float *__restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int m = count[i];
    const float4 q = orig[i];
    const float rx = use_depth ? centre.x : q.x, ry = use_depth ? centre.y : q.y, rz = use_depth ? centre.z : q.z;
    float sum = 0.f, sumsq = 0.f;
    for (int j = 0; j < m; ++j) {
        const float4 p = orig[idx[i * k + j]];
        const float dx = __fsub_rn(p.x, rx), dy = __fsub_rn(p.y, ry), dz = __fsub_rn(p.z, rz);
        const float dist = __fsqrt_rn(eigen_dot3(dx, dy, dz, dx, dy, dz));
        sum = __fadd_rn(sum, dist);
        sumsq = __fadd_rn(sumsq, __fmul_rn(dist, dist));
    }
    const float fm = (float)m;
    out[i] = m > 2 ? __fdiv_rn(__fsub_rn(sumsq, __fdiv_rn(__fmul_rn(sum, sum), fm)), __fsub_rn(fm, 1.f)) : 0.f;
}

// f2: calcDistHist / calcIntensityHist (featureeval/utils.h:18-110, 113-195): per point a `bins`-bin histogram over its
// capped strict-radius neighbourhood, every neighbour adding norm_increment = float(1.0 / m).  Distance flavour:
// bin = int((double(sqrtf(d2)) / radius) * (bins-1)); intensity flavour: bin = int(intensity * float(bins-1)).
// Fewer than 3 neighbours leave the row at 0.  A bin's value is the same increment added count times, so the result
// does not depend on the neighbour order.
__global__ void __launch_bounds__(256) k_neighbour_hist(const float4 *__restrict__ orig, const int32_t *__restrict__ idx,
                                                        const int32_t *__restrict__ count, size_t n, int k, int bins, int which, double radius,
                                                        const char *__restrict__ attr, size_t attr_stride, float *__restrict__ hist) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float *row = hist + i * (size_t)bins;
    for (int b = 0; b < bins; ++b) row[b] = 0.f;
    const int m = count[i];
    if (m < 3) return;
    const float inc = (float)(1.0 / (double)m);
    const float4 q = orig[i];
    for (int j = 0; j < m; ++j) {
        const int t = idx[i * k + j];
        int bin;
        if (which == 0) {
            const float4 p = orig[t];
            const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y), dz = __fsub_rn(q.z, p.z);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            bin = (int)(((double)__fsqrt_rn(d2) / radius) * (double)(bins - 1));
        } else {
            const float a = *reinterpret_cast<const float *>(attr + (size_t)t * attr_stride);
            bin = (int)__fmul_rn(a, (float)(bins - 1));
        }
        bin = bin < 0 ? 0 : (bin > bins - 1 ? bins - 1 : bin);       // the reference indexes out of bounds for intensities outside [0,1]
        row[bin] = __fadd_rn(row[bin], inc);
    }
}

// f2: don() of featureeval.cpp:694-739 — per point and radius the mean of the neighbours' normals, the point's own
// normal counted twice (radiusSearch returns it and :729 pushes it again), then (mean(r1) - mean(r2)) / 2.
__global__ void __launch_bounds__(256) k_don_combine(const float *__restrict__ s1, const int32_t *__restrict__ c1, const float *__restrict__ s2,
                                                     const int32_t *__restrict__ c2, const char *__restrict__ normals, size_t stride, size_t n,
                                                     float *__restrict__ out4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *self = reinterpret_cast<const float *>(normals + i * stride);
    const float m1 = (float)(c1[i] + 1), m2 = (float)(c2[i] + 1);
    float o[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float a1 = __fdiv_rn(__fadd_rn(s1[3 * i + d], self[d]), m1);
        const float a2 = __fdiv_rn(__fadd_rn(s2[3 * i + d], self[d]), m2);
        o[d] = (float)((double)__fsub_rn(a1, a2) / 2.0);
    }
    out4[4 * i] = o[0]; out4[4 * i + 1] = o[1]; out4[4 * i + 2] = o[2]; out4[4 * i + 3] = 0.f;
}

static QueryArgs base_args(ccg_ctx *ctx) {
    QueryArgs a{};
    a.rb = INFINITY; a.begin = 0; a.end = ctx->n; a.sorted_rows = 0;
    a.cov_norm = CCG_COV_NM1; a.min_pts = 3;
    return a;
}

}  // namespace ccg

using namespace ccg;

extern "C" {

int ccg_version(void) { return CCG_VERSION; }

int ccg_create(ccg_ctx **out, int device_id) {
    if (!out) return CCG_ERR_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return CCG_ERR_CUDA; }
    if (device_id < 0 || device_id >= ndev) return CCG_ERR_INVALID;
    ccg_ctx *ctx = new (std::nothrow) ccg_ctx();
    if (!ctx) return CCG_ERR_INVALID;
    ctx->device = device_id;
    if (cudaSetDevice(device_id) != cudaSuccess) { delete ctx; return CCG_ERR_CUDA; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) { delete ctx; return CCG_ERR_CUDA; }
    ctx->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return CCG_ERR_CUDA; }
    ctx->stream = ctx->own_stream;
    for (int i = 0; i < 8; ++i) if (cudaEventCreate(&ctx->ev[i]) != cudaSuccess) { delete ctx; return CCG_ERR_CUDA; }
    const char *cs = getenv("CCG_COLLECT_STATS");
    ctx->collect_stats = (cs && cs[0] == '1') ? 1 : 0;
    const char *sp = getenv("CCG_SPLIT_SPARSE");
    ctx->split_sparse = (sp && sp[0] == '0') ? 0 : 1;
    const char *kbits = getenv("CCG_KEY_BITS");
    if (kbits) ctx->key_bits = atoi(kbits);
    const char *cv = getenv("CCG_CURVE");
    ctx->curve_hilbert = (cv && cv[0] == 'm') ? 0 : 1;
    const char *da = getenv("CCG_SEL_DRAIN_AT");
    if (da) ctx->sel_drain_at = atoi(da);
    const char *sn = getenv("CCG_SEL_NET");
    if (sn) ctx->sel_net = atoi(sn);
    const char *mb = getenv("CCG_SEL_MINB");
    if (mb) ctx->sel_minb = atoi(mb);
    const char *sg = getenv("CCG_SEL_STAGE");
    ctx->sel_stage = (sg && sg[0] == 'l') ? 0 : 1;
    const char *sc = getenv("CCG_SEL_CHUNKS");
    if (sc) ctx->sel_chunks = atoi(sc);
    const char *po = getenv("CCG_POINT_OVERLAP");
    ctx->point_overlap = (po && po[0] == '0') ? 0 : 1;
    const char *nm = getenv("CCG_NODE_MOMENTS");
    if (nm) ctx->node_moments = atoi(nm) != 0;
    const char *kt = getenv("CCG_KNN_TOPK");
    ctx->knn_use_topk = (kt && kt[0] == '1') ? 1 : 0;
    const char *pc = getenv("CCG_PIPELINE_CHUNKS");
    if (pc) ctx->pipeline_chunks = atoi(pc);
    const char *pm = getenv("CCG_PIPELINE_MIN_POINTS");
    if (pm) ctx->pipeline_min_points = atoi(pm);
    *out = ctx;
    return CCG_OK;
}

void ccg_destroy(ccg_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    ctx->raw.release(); ctx->orig.release(); ctx->sorted.release(); ctx->rel.release();
    for (int i = 0; i < 2; ++i) { ctx->keys[i].release(); ctx->vals[i].release(); ctx->hpin[i].release(); }
    ctx->vox_out.release();
    ctx->ptx_bytes.release(); ctx->ptx_starts.release(); ctx->ptx_points.release(); ctx->ptx_c2g.release();
    ctx->sort_tmp.release(); ctx->boxes_lo.release(); ctx->boxes_hi.release(); ctx->scratch.release(); ctx->defer.release(); ctx->defer_b.release(); ctx->rule_T.release(); ctx->rule_I.release(); ctx->node_mom.release();
    for (int i = 0; i < 2; ++i) if (ctx->ev_side[i]) cudaEventDestroy(ctx->ev_side[i]);
    if (ctx->ev_tile) cudaEventDestroy(ctx->ev_tile);
    if (ctx->side_stream) cudaStreamDestroy(ctx->side_stream);
    for (int i = 0; i < 4; ++i) { ctx->out[i].release(); ctx->tmp[i].release(); }
    for (int i = 0; i < 8; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    for (int i = 0; i < 17; ++i) if (ctx->ev_chunk[i]) cudaEventDestroy(ctx->ev_chunk[i]);
    for (int i = 0; i < 17; ++i) if (ctx->ev_copy[i]) cudaEventDestroy(ctx->ev_copy[i]);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
}

const char *ccg_last_error(const ccg_ctx *ctx) { return ctx ? ctx->err : "invalid context"; }

int ccg_get_stats(const ccg_ctx *ctx, ccg_stats *out) {
    if (!ctx || !out) return CCG_ERR_INVALID;
    *out = ctx->stats;
    return CCG_OK;
}

int ccg_set_cancel_flag(ccg_ctx *ctx, const volatile int *flag) {
    if (!ctx) return CCG_ERR_INVALID;
    ctx->cancel = flag;
    return CCG_OK;
}

int ccg_set_stream(ccg_ctx *ctx, void *cuda_stream) {
    if (!ctx) return CCG_ERR_INVALID;
    if (cudaSetDevice(ctx->device) == cudaSuccess && ctx->stream) cudaStreamSynchronize(ctx->stream);
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return CCG_OK;
}

int ccg_set_async(ccg_ctx *ctx, int on) {
    if (!ctx) return CCG_ERR_INVALID;
    if (ctx->async_mode && !on) { const int rc = ccg_sync(ctx); ctx->async_mode = 0; return rc; }
    ctx->async_mode = on ? 1 : 0;
    return CCG_OK;
}

int ccg_sync(ccg_ctx *ctx) {
    if (!ctx) return CCG_ERR_INVALID;
    CCG_CUDA(ctx, cudaSetDevice(ctx->device));
    CCG_TRY(join_side(ctx));
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->copy_stream) CCG_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    CCG_CUDA(ctx, cudaGetLastError());
    if (ctx->pending_stats) {
        ctx->pending_stats = 0;
        cudaEventElapsedTime(&ctx->stats.ms_query, ctx->ev[3], ctx->ev[4]);
        cudaEventElapsedTime(&ctx->stats.ms_download, ctx->ev[4], ctx->ev[5]);
        cudaGetLastError();
    }
    return finish_cloud_check(ctx);
}

int ccg_set_cloud(ccg_ctx *ctx, const void *xyz, size_t n, size_t stride_bytes) {
    CCG_TRY(begin_call(ctx, false));
    if (n > 0 && !xyz) return set_error(ctx, CCG_ERR_INVALID, "xyz is NULL");
    return build_cloud(ctx, xyz, n, stride_bytes);
}

int64_t ccg_cloud_size(const ccg_ctx *ctx) { return ctx ? (int64_t)ctx->n : -1; }

int ccg_get_order(ccg_ctx *ctx, int32_t *order) {
    CCG_TRY(begin_call(ctx, true));
    if (!order) return set_error(ctx, CCG_ERR_INVALID, "order is NULL");
    if (ctx->n == 0) return CCG_OK;
    CCG_CUDA(ctx, cudaMemcpyAsync(order, ctx->vals[ctx->order_sel].p, ctx->n * 4, cudaMemcpyDefault, ctx->stream));
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return CCG_OK;
}

int ccg_knn(ccg_ctx *ctx, int k, int32_t *idx, float *d2) {
    CCG_TRY(begin_call(ctx, true));
    if (!idx) return set_error(ctx, CCG_ERR_INVALID, "idx is NULL");
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    OutBuf oi, od;
    CCG_TRY(oi.open(ctx, 0, idx, ctx->n * (size_t)k * 4));
    CCG_TRY(od.open(ctx, 1, d2, ctx->n * (size_t)k * 4));
    QueryArgs a = base_args(ctx);
    a.mode = MODE_KNN; a.k = k; a.idx = oi.dev<int32_t>(); a.d2 = od.dev<float>();
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_tile_query(ctx, a));
    CCG_TRY(query_end(ctx));
    CCG_TRY(oi.close()); CCG_TRY(od.close());
    return finish_call(ctx);
}

/* ---- f3: kNN graph for the graph-cut / flood consumers ------------------------------------------------- */
int ccg_knn_graph(ccg_ctx *ctx, int k, double inverse_sigma, int32_t *idx, double *weight) {
    CCG_TRY(begin_call(ctx, true));
    if (!idx || !weight) return set_error(ctx, CCG_ERR_INVALID, "idx and weight are required");
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    OutBuf oi, ow;
    CCG_TRY(oi.open(ctx, 0, idx, n * (size_t)k * 4));
    CCG_TRY(ow.open(ctx, 1, weight, n * (size_t)k * 8));
    QueryArgs a = base_args(ctx);
    a.mode = MODE_KNN; a.k = k; a.idx = oi.dev<int32_t>(); a.d2 = nullptr;
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_tile_query(ctx, a));
    const size_t edges = n * (size_t)k;
    k_binary_potential<<<(unsigned)((edges + 255) / 256), 256, 0, ctx->stream>>>(ctx->orig.as<float4>(), oi.dev<int32_t>(), n, k, inverse_sigma,
                                                                                 ow.dev<double>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oi.close()); CCG_TRY(ow.close());
    return finish_call(ctx);
}

/* ---- f2: radius-neighbourhood reductions ------------------------------------------------------------------ */
static int radius_lists_internal(ccg_ctx *ctx, double r, int max_nn, int strict_lt, int32_t **idx, int32_t **count) {
    if (!(r >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be >= 0");
    if (max_nn < 1 || max_nn > CCG_MAX_K)
        return set_error(ctx, CCG_ERR_UNSUPPORTED, "this reduction needs a neighbour cap 1 <= max_nn <= %d (got %d)", CCG_MAX_K, max_nn);
    const size_t n = ctx->n;
    CCG_CUDA(ctx, ctx->tmp[2].reserve(n * (size_t)max_nn * 4 + 64));
    CCG_CUDA(ctx, ctx->tmp[0].reserve(n * 4 + 64));
    *idx = ctx->tmp[2].as<int32_t>();
    *count = ctx->tmp[0].as<int32_t>();
    QueryArgs a = base_args(ctx);
    a.mode = MODE_KNN; a.k = max_nn; a.rb = radius_bound(r, strict_lt); a.idx = *idx; a.d2 = nullptr; a.count = *count;
    return launch_tile_query(ctx, a);
}

int ccg_normal_stdev(ccg_ctx *ctx, double r, int max_nn, const void *normals, size_t stride_bytes, float *out) {
    CCG_TRY(begin_call(ctx, true));
    if (!normals || !out) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (stride_bytes < 12 || (stride_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be >= 12 and a multiple of 4");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    const void *nd = normals;
    if (!is_dev(ctx, normals)) {
        const size_t bytes = (n - 1) * stride_bytes + 12;
        CCG_CUDA(ctx, ctx->tmp[1].reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, normals, bytes, cudaMemcpyHostToDevice, ctx->stream));
        nd = ctx->tmp[1].p;
    }
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, out, n * 4));
    CCG_TRY(query_begin(ctx));
    if (max_nn == 0 || max_nn > CCG_MAX_K) {                  // no list: streaming epilogue in traversal order (radsel.cu)
        if (max_nn < 0) return set_error(ctx, CCG_ERR_INVALID, "max_nn must be >= 0");
        const float rb = radius_bound(r, 1);
        const float *Tp; const int32_t *Ip;
        CCG_TRY(launch_radius_rule(ctx, rb, max_nn, &Tp, &Ip));
        CCG_TRY(launch_rule_normal_stdev(ctx, rb, Tp, Ip, nd, stride_bytes, oo.dev<float>()));
        CCG_TRY(query_end(ctx));
        CCG_TRY(oo.close());
        return finish_call(ctx);
    }
    int32_t *idx, *cnt;
    CCG_TRY(radius_lists_internal(ctx, r, max_nn, 1, &idx, &cnt));
    k_normal_stdev<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(idx, cnt, n, max_nn, (const char *)nd, stride_bytes, oo.dev<float>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_stdev_dist(ccg_ctx *ctx, double r, int max_nn, int use_depth, const float *centre3, float *out) {
    CCG_TRY(begin_call(ctx, true));
    if (!out) return set_error(ctx, CCG_ERR_INVALID, "out is NULL");
    if (use_depth && !centre3) return set_error(ctx, CCG_ERR_INVALID, "use_depth needs the sensor origin");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, out, n * 4));
    CCG_TRY(query_begin(ctx));
    if (max_nn == 0 || max_nn > CCG_MAX_K) {
        if (max_nn < 0) return set_error(ctx, CCG_ERR_INVALID, "max_nn must be >= 0");
        const float rb = radius_bound(r, 0);
        const float zero[3] = {0.f, 0.f, 0.f};
        const float *Tp; const int32_t *Ip;
        CCG_TRY(launch_radius_rule(ctx, rb, max_nn, &Tp, &Ip));
        CCG_TRY(launch_rule_stdev_dist(ctx, rb, Tp, Ip, use_depth ? 1 : 0, centre3 ? centre3 : zero, oo.dev<float>()));
        CCG_TRY(query_end(ctx));
        CCG_TRY(oo.close());
        return finish_call(ctx);
    }
    int32_t *idx, *cnt;
    CCG_TRY(radius_lists_internal(ctx, r, max_nn, 0, &idx, &cnt));
    const float3 c = centre3 ? make_float3(centre3[0], centre3[1], centre3[2]) : make_float3(0.f, 0.f, 0.f);
    k_stdev_dist<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->orig.as<float4>(), idx, cnt, n, max_nn, use_depth ? 1 : 0, c,
                                                                       oo.dev<float>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_neighbour_hist(ccg_ctx *ctx, double r, int max_nn, int bins, int which, const void *attr, size_t attr_stride_bytes, float *hist) {
    CCG_TRY(begin_call(ctx, true));
    if (!hist) return set_error(ctx, CCG_ERR_INVALID, "hist is NULL");
    if (bins < 1 || bins > 4096) return set_error(ctx, CCG_ERR_INVALID, "bins must be in [1,4096]");
    if (which != 0 && which != 1) return set_error(ctx, CCG_ERR_INVALID, "which must be 0 (distance) or 1 (attribute)");
    if (which == 1 && (!attr || attr_stride_bytes < 4 || (attr_stride_bytes & 3))) return set_error(ctx, CCG_ERR_INVALID, "attribute array / stride invalid");
    if (which == 0 && !(r > 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be > 0");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    const void *ad = attr;
    if (which == 1 && !is_dev(ctx, attr)) {
        const size_t bytes = (n - 1) * attr_stride_bytes + 4;
        CCG_CUDA(ctx, ctx->tmp[1].reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, attr, bytes, cudaMemcpyHostToDevice, ctx->stream));
        ad = ctx->tmp[1].p;
    }
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, hist, n * (size_t)bins * 4));
    CCG_TRY(query_begin(ctx));
    if (max_nn == 0 || max_nn > CCG_MAX_K) {                  // counts per bin are order-independent: bit-exact without a list
        if (max_nn < 0) return set_error(ctx, CCG_ERR_INVALID, "max_nn must be >= 0");
        if (!(r >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be >= 0");
        const float rb = radius_bound(r, 1);
        const float *Tp; const int32_t *Ip;
        CCG_TRY(launch_radius_rule(ctx, rb, max_nn, &Tp, &Ip));
        CCG_TRY(launch_rule_hist(ctx, rb, Tp, Ip, bins, which, r, ad, attr_stride_bytes, oo.dev<float>()));
        CCG_TRY(query_end(ctx));
        CCG_TRY(oo.close());
        return finish_call(ctx);
    }
    int32_t *idx, *cnt;
    CCG_TRY(radius_lists_internal(ctx, r, max_nn, 1, &idx, &cnt));
    k_neighbour_hist<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->orig.as<float4>(), idx, cnt, n, max_nn, bins, which, r,
                                                                           (const char *)ad, attr_stride_bytes, oo.dev<float>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_don(ccg_ctx *ctx, const void *normals, size_t stride_bytes, double r1, double r2, float *don4) {
    CCG_TRY(begin_call(ctx, true));
    if (!normals || !don4) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (stride_bytes < 12 || (stride_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be >= 12 and a multiple of 4");
    if (!(r1 >= 0.0) || !(r2 >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radii must be >= 0");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    const void *nd = normals;
    if (!is_dev(ctx, normals)) {
        const size_t bytes = (n - 1) * stride_bytes + 12;
        CCG_CUDA(ctx, ctx->tmp[1].reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, normals, bytes, cudaMemcpyHostToDevice, ctx->stream));
        nd = ctx->tmp[1].p;
    }
    CCG_CUDA(ctx, ctx->tmp[2].reserve(n * 24 + 64));
    CCG_CUDA(ctx, ctx->tmp[0].reserve(n * 8 + 64));
    float *s1 = ctx->tmp[2].as<float>(), *s2 = s1 + 3 * n;
    int32_t *c1 = ctx->tmp[0].as<int32_t>(), *c2 = c1 + n;
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, don4, n * 16));
    CCG_TRY(query_begin(ctx));
    // the reference passes float radii to pcl::search::KdTree (don(cloud, normals, float radius1, float radius2))
    // both radii in one traversal (the walk is pruned by the larger one)
    CCG_TRY(launch_tile_attrsum2(ctx, radius_bound((double)(float)r1, 1), radius_bound((double)(float)r2, 1), nd, stride_bytes, s1, c1, s2, c2));
    k_don_combine<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(s1, c1, s2, c2, (const char *)nd, stride_bytes, n, oo.dev<float>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_principal_curvatures(ccg_ctx *ctx, double r, const void *normals, size_t stride_bytes, float *pc5) {
    CCG_TRY(begin_call(ctx, true));
    if (!normals || !pc5) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (stride_bytes < 12 || (stride_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be >= 12 and a multiple of 4");
    if (!(r >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be >= 0");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    const void *nd = normals;
    if (!is_dev(ctx, normals)) {
        const size_t bytes = (n - 1) * stride_bytes + 12;
        CCG_CUDA(ctx, ctx->tmp[1].reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, normals, bytes, cudaMemcpyHostToDevice, ctx->stream));
        nd = ctx->tmp[1].p;
    }
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, pc5, n * 20));
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_tile_principal_curv(ctx, radius_bound(r, 1), nd, stride_bytes, oo.dev<float>()));
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_fpfh(ccg_ctx *ctx, double r, const void *normals, size_t stride_bytes, float *fpfh33, float *spfh33) {
    CCG_TRY(begin_call(ctx, true));
    if (!normals || !fpfh33) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (stride_bytes < 12 || (stride_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be >= 12 and a multiple of 4");
    if (!(r >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be >= 0");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    const void *nd = normals;
    if (!is_dev(ctx, normals)) {
        const size_t bytes = (n - 1) * stride_bytes + 12;
        CCG_CUDA(ctx, ctx->tmp[1].reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, normals, bytes, cudaMemcpyHostToDevice, ctx->stream));
        nd = ctx->tmp[1].p;
    }
    OutBuf of, os;
    CCG_TRY(of.open(ctx, 0, fpfh33, n * 132));
    float *spfh_dev = nullptr;
    if (spfh33) { CCG_TRY(os.open(ctx, 1, spfh33, n * 132)); spfh_dev = os.dev<float>(); }
    else { CCG_CUDA(ctx, ctx->tmp[2].reserve(n * 132 + 64)); spfh_dev = ctx->tmp[2].as<float>(); }
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_tile_fpfh(ctx, radius_bound(r, 1), nd, stride_bytes, spfh_dev, of.dev<float>()));
    CCG_TRY(query_end(ctx));
    CCG_TRY(of.close());
    if (spfh33) CCG_TRY(os.close());
    return finish_call(ctx);
}

int ccg_radius(ccg_ctx *ctx, double r, int max_nn, int strict_lt, int32_t *idx, float *d2, int32_t *count) {
    CCG_TRY(begin_call(ctx, true));
    if (!(r >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be >= 0");
    const bool lists = idx || d2;
    if (lists && (max_nn < 1 || max_nn > CCG_MAX_K))
        return set_error(ctx, CCG_ERR_UNSUPPORTED, "neighbour lists need 1 <= max_nn <= %d (got %d)", CCG_MAX_K, max_nn);
    if (max_nn < 0) return set_error(ctx, CCG_ERR_INVALID, "max_nn must be >= 0");
    OutBuf oi, od, oc;
    CCG_TRY(oc.open(ctx, 2, count, ctx->n * 4));
    const float rb = radius_bound(r, strict_lt);
    CCG_TRY(query_begin(ctx));
    if (lists) {
        CCG_TRY(oi.open(ctx, 0, idx, ctx->n * (size_t)max_nn * 4));
        CCG_TRY(od.open(ctx, 1, d2, ctx->n * (size_t)max_nn * 4));
        QueryArgs a = base_args(ctx);
        a.mode = MODE_KNN; a.k = max_nn; a.rb = rb; a.idx = oi.dev<int32_t>(); a.d2 = od.dev<float>(); a.count = oc.dev<int32_t>();
        CCG_TRY(launch_tile_query(ctx, a));
    } else {
        StreamArgs s{};
        s.mode = 0; s.rb = rb; s.count_cap = max_nn; s.begin = 0; s.end = ctx->n; s.count = oc.dev<int32_t>();
        CCG_TRY(launch_tile_stream(ctx, s));
    }
    CCG_TRY(query_end(ctx));
    CCG_TRY(oi.close()); CCG_TRY(od.close()); CCG_TRY(oc.close());
    return finish_call(ctx);
}

/* ---- a2: variable-length neighbour lists (what radiusSearch fills into std::vector<int> / std::vector<float>) ---- */
int ccg_radius_csr(ccg_ctx *ctx, double r, int max_nn, int strict_lt, int64_t *offsets, int32_t *idx, float *d2, int64_t capacity,
                   int64_t *total) {
    CCG_TRY(begin_call(ctx, true));
    if (!(r >= 0.0) || max_nn < 0) return set_error(ctx, CCG_ERR_INVALID, "need r >= 0 and max_nn >= 0");
    if (!offsets || !total) return set_error(ctx, CCG_ERR_INVALID, "offsets and total are required");
    const size_t n = ctx->n;
    *total = 0;
    if (n == 0) { offsets[0] = 0; return CCG_OK; }
    static_assert(sizeof(long long) == sizeof(int64_t), "int64");
    const float rb = radius_bound(r, strict_lt);
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, offsets, (n + 1) * 8));
    CCG_TRY(query_begin(ctx));
    const float *Tp; const int32_t *Ip;
    CCG_TRY(launch_radius_rule(ctx, rb, max_nn, &Tp, &Ip));
    long long tot = 0;
    CCG_TRY(launch_radius_csr_count(ctx, rb, Tp, Ip, oo.dev<long long>(), &tot));
    *total = tot;
    if (idx || d2) {
        if (tot > capacity) return set_error(ctx, CCG_ERR_INVALID, "radius lists hold %lld entries, capacity is %lld", tot, (long long)capacity);
        OutBuf oi, od;
        CCG_TRY(oi.open(ctx, 1, idx, (size_t)tot * 4));
        CCG_TRY(od.open(ctx, 2, d2, (size_t)tot * 4));
        CCG_TRY(launch_radius_csr_fill(ctx, rb, Tp, Ip, oo.dev<long long>(), tot, oi.dev<int32_t>(), od.dev<float>()));
        CCG_TRY(query_end(ctx));
        CCG_TRY(oi.close()); CCG_TRY(od.close());
    } else {
        CCG_TRY(query_end(ctx));
    }
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

static int pca_common(ccg_ctx *ctx, int k, float rb, int cov_norm, int normalise_hist, const float *vp, size_t begin, size_t end,
                      int sorted_rows, float *eig3, float *normal4, int32_t *count, bool uncapped, float *feat7 = nullptr,
                      int formula_set = 0) {
    if (cov_norm < 0 || cov_norm > 2) return set_error(ctx, CCG_ERR_INVALID, "cov_norm must be 0, 1 or 2");
    size_t rows = sorted_rows ? (end > begin ? end - begin : 0) : ctx->n;
    if (formula_set < 0 || formula_set > 1) return set_error(ctx, CCG_ERR_INVALID, "formula_set must be 0 or 1");
    OutBuf oe, on, oc, of;
    CCG_TRY(oe.open(ctx, 0, eig3, rows * 12));
    CCG_TRY(on.open(ctx, 1, normal4, rows * 16));
    CCG_TRY(oc.open(ctx, 2, count, rows * 4));
    CCG_TRY(of.open(ctx, 3, feat7, rows * 28));
    CCG_TRY(query_begin(ctx));
    if (!uncapped) {
        QueryArgs a = base_args(ctx);
        a.mode = MODE_PCA; a.k = k; a.rb = rb; a.begin = begin; a.end = end; a.sorted_rows = sorted_rows;
        a.cov_norm = cov_norm; a.min_pts = 3; a.normalise_hist = normalise_hist; a.has_vp = vp ? 1 : 0;
        if (vp) for (int i = 0; i < 3; ++i) a.vp[i] = vp[i];
        a.eig3 = oe.dev<float>(); a.normal4 = on.dev<float>(); a.count = oc.dev<int32_t>();
        a.feat7 = of.dev<float>(); a.formula_set = formula_set;
        a.shard_world = ctx->shard_world; a.shard_rank = ctx->shard_rank; a.shard_chunk = ctx->shard_chunk;
        a.route_world = ctx->route_world; a.route_rows = ctx->route_rows;
        a.r_eig = ctx->route_ptr[0]; a.r_nrm = ctx->route_ptr[1]; a.r_feat = ctx->route_ptr[2];
        PipeOut po[4];
        int npo = 0;
        if (oe.staged) po[npo++] = PipeOut{&oe, 12};
        if (on.staged) po[npo++] = PipeOut{&on, 16};
        if (oc.staged) po[npo++] = PipeOut{&oc, 4};
        if (of.staged) po[npo++] = PipeOut{&of, 28};
        if (npo > 0 && !sorted_rows && a.shard_world <= 1 && ctx->pipeline_chunks > 1 && !ctx->collect_stats && !ctx->async_mode && ctx->n >= (size_t)ctx->pipeline_min_points)
            CCG_TRY(pipelined_query(ctx, a, po, npo, ctx->pipeline_chunks));
        else
            CCG_TRY(launch_tile_query(ctx, a));
    } else {
        StreamArgs s{};
        s.mode = 1; s.rb = rb; s.count_cap = 0; s.begin = begin; s.end = end; s.sorted_rows = sorted_rows;
        s.cov_norm = cov_norm; s.min_pts = 3; s.normalise_hist = normalise_hist; s.has_vp = vp ? 1 : 0;
        if (vp) for (int i = 0; i < 3; ++i) s.vp[i] = vp[i];
        s.eig3 = oe.dev<float>(); s.normal4 = on.dev<float>(); s.count = oc.dev<int32_t>();
        s.feat7 = of.dev<float>(); s.formula_set = formula_set;
        CCG_TRY(launch_tile_stream(ctx, s));
    }
    CCG_TRY(query_end(ctx));
    CCG_TRY(oe.close()); CCG_TRY(on.close()); CCG_TRY(oc.close()); CCG_TRY(of.close());
    return finish_call(ctx);
}

int ccg_pca_knn_features(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, float *eig3, float *normal4,
                         float *feat7) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    return pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, 0, ctx->n, 0, eig3, normal4, nullptr, false, feat7, formula_set);
}

int ccg_pca_knn_features_range(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, size_t begin, size_t end,
                               float *eig3_sorted, float *normal4_sorted, float *feat7_sorted) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    if (begin % kLeaf) return set_error(ctx, CCG_ERR_INVALID, "begin must be a multiple of 32");
    if (end > ctx->n) end = ctx->n;
    return pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, begin, end, 1, eig3_sorted, normal4_sorted, nullptr, false,
                      feat7_sorted, formula_set);
}

int ccg_pca_knn_features_scatter(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, size_t begin, size_t end,
                                 float *eig3, float *normal4, float *feat7) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    if (begin % kLeaf) return set_error(ctx, CCG_ERR_INVALID, "begin must be a multiple of 32");
    if (end > ctx->n) end = ctx->n;
    if ((eig3 && !is_dev(ctx, eig3)) || (normal4 && !is_dev(ctx, normal4)) || (feat7 && !is_dev(ctx, feat7)))
        return set_error(ctx, CCG_ERR_INVALID, "scatter outputs must be device (own or peer) memory");
    return pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, begin, end, 0, eig3, normal4, nullptr, false, feat7, formula_set);
}

int ccg_pca_knn_features_shard(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, int rank, int world,
                               int chunk_points, int sorted_rows, float *eig3, float *normal4, float *feat7) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    if (world < 1 || rank < 0 || rank >= world) return set_error(ctx, CCG_ERR_INVALID, "need 0 <= rank < world");
    if (chunk_points < 32 || chunk_points % 32) return set_error(ctx, CCG_ERR_INVALID, "chunk_points must be a positive multiple of 32");
    if ((eig3 && !is_dev(ctx, eig3)) || (normal4 && !is_dev(ctx, normal4)) || (feat7 && !is_dev(ctx, feat7)))
        return set_error(ctx, CCG_ERR_INVALID, "shard outputs must be device (own or peer) memory");
    ctx->shard_world = world > 1 ? world : 0; ctx->shard_rank = rank; ctx->shard_chunk = chunk_points / 32;
    const int rc = pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, 0, ctx->n, sorted_rows ? 1 : 0, eig3, normal4, nullptr, false, feat7,
                              formula_set);
    ctx->shard_world = 0; ctx->shard_rank = 0; ctx->shard_chunk = 0;
    return rc;
}

int ccg_pca_knn_features_route(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, int rank, int world,
                               int chunk_points, size_t rows_per_rank, float *const *eig3_of_rank, float *const *normal4_of_rank,
                               float *const *feat7_of_rank) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    if (world < 1 || world > 16 || rank < 0 || rank >= world) return set_error(ctx, CCG_ERR_INVALID, "need 0 <= rank < world <= 16");
    if (chunk_points < 32 || chunk_points % 32) return set_error(ctx, CCG_ERR_INVALID, "chunk_points must be a positive multiple of 32");
    if (rows_per_rank == 0 || rows_per_rank * (size_t)world < ctx->n) return set_error(ctx, CCG_ERR_INVALID, "rows_per_rank * world must cover the cloud");
    for (int r = 0; r < world; ++r) {
        ctx->route_ptr[0][r] = eig3_of_rank ? eig3_of_rank[r] : nullptr;
        ctx->route_ptr[1][r] = normal4_of_rank ? normal4_of_rank[r] : nullptr;
        ctx->route_ptr[2][r] = feat7_of_rank ? feat7_of_rank[r] : nullptr;
        for (int a = 0; a < 3; ++a)
            if (ctx->route_ptr[a][r] && !is_dev(ctx, ctx->route_ptr[a][r])) return set_error(ctx, CCG_ERR_INVALID, "route outputs must be device (own or peer) memory");
    }
    ctx->shard_world = world > 1 ? world : 0; ctx->shard_rank = rank; ctx->shard_chunk = chunk_points / 32;
    ctx->route_world = world; ctx->route_rows = rows_per_rank;
    if (world == 1) ctx->route_world = 0;
    // non-NULL markers so that pca_common opens the three outputs as "device" buffers; with world == 1 they ARE the outputs
    float *e = ctx->route_ptr[0][rank], *nn = ctx->route_ptr[1][rank], *f = ctx->route_ptr[2][rank];
    const int rc = pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, 0, ctx->n, 0, e, nn, nullptr, false, f, formula_set);
    ctx->shard_world = 0; ctx->shard_rank = 0; ctx->shard_chunk = 0; ctx->route_world = 0; ctx->route_rows = 0;
    return rc;
}

int ccg_dev_alloc(ccg_ctx *ctx, size_t bytes, void **out) {
    CCG_TRY(begin_call(ctx, false));
    if (!out) return set_error(ctx, CCG_ERR_INVALID, "out is NULL");
    *out = nullptr;
    CCG_CUDA(ctx, cudaMalloc(out, bytes ? bytes : 1));
    return CCG_OK;
}
int ccg_dev_free(ccg_ctx *ctx, void *ptr) {
    CCG_TRY(begin_call(ctx, false));
    if (ptr) CCG_CUDA(ctx, cudaFree(ptr));
    return CCG_OK;
}
int ccg_ipc_export(ccg_ctx *ctx, const void *ptr, void *handle64) {
    CCG_TRY(begin_call(ctx, false));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    if (!ptr || !handle64) return set_error(ctx, CCG_ERR_INVALID, "NULL argument");
    cudaIpcMemHandle_t h;
    CCG_CUDA(ctx, cudaIpcGetMemHandle(&h, const_cast<void *>(ptr)));
    memcpy(handle64, &h, 64);
    return CCG_OK;
}
int ccg_ipc_open(ccg_ctx *ctx, const void *handle64, void **out) {
    CCG_TRY(begin_call(ctx, false));
    if (!handle64 || !out) return set_error(ctx, CCG_ERR_INVALID, "NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    *out = nullptr;
    CCG_CUDA(ctx, cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return CCG_OK;
}
int ccg_ipc_close(ccg_ctx *ctx, void *ptr) {
    CCG_TRY(begin_call(ctx, false));
    if (ptr) CCG_CUDA(ctx, cudaIpcCloseMemHandle(ptr));
    return CCG_OK;
}

/* ---- f1: octreeDownsample (utils.h:16-62) -------------------------------------------------------------- */
int ccg_voxel_downsample(ccg_ctx *ctx, const void *points, size_t stride_bytes, int n_floats, float resolution, int as_written,
                         int32_t *big_to_small, int64_t *n_out) {
    CCG_TRY(begin_call(ctx, true));
    if (!n_out) return set_error(ctx, CCG_ERR_INVALID, "n_out is NULL");
    *n_out = 0;
    ctx->vox_n = 0;
    if (!(resolution > 0.f)) return set_error(ctx, CCG_ERR_INVALID, "resolution must be > 0");
    if (stride_bytes < 12 || (stride_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be >= 12 and a multiple of 4");
    if (n_floats < 1 || n_floats > 16 || (size_t)n_floats * 4 > stride_bytes)
        return set_error(ctx, CCG_ERR_INVALID, "n_floats must be in [1,16] and fit the stride");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    if (!points) return set_error(ctx, CCG_ERR_INVALID, "points is NULL");
    const void *pd = points;
    if (!is_dev(ctx, points)) {
        const size_t bytes = (n - 1) * stride_bytes + (size_t)n_floats * 4;
        CCG_CUDA(ctx, ctx->raw.reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->raw.p, points, bytes, cudaMemcpyHostToDevice, ctx->stream));
        pd = ctx->raw.p;
    }
    OutBuf ob;
    CCG_TRY(ob.open(ctx, 0, big_to_small, n * 4));
    CCG_TRY(query_begin(ctx));
    CCG_TRY(voxel_downsample(ctx, pd, stride_bytes, n_floats, resolution, as_written, ob.dev<int32_t>(), n_out));
    CCG_TRY(query_end(ctx));
    CCG_TRY(ob.close());
    return finish_call(ctx);
}

int ccg_voxel_downsample_fetch(ccg_ctx *ctx, float *out_points) {
    CCG_TRY(begin_call(ctx, false));
    if (ctx->vox_n == 0) return CCG_OK;
    if (!out_points) return set_error(ctx, CCG_ERR_INVALID, "out_points is NULL");
    const size_t bytes = (size_t)ctx->vox_n * ctx->vox_nf * sizeof(float);
    CCG_CUDA(ctx, cudaMemcpyAsync(out_points, ctx->vox_out.p, bytes, cudaMemcpyDefault, ctx->stream));
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return CCG_OK;
}

int ccg_pca_knn(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, float *eig3, float *normal4) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    return pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, 0, ctx->n, 0, eig3, normal4, nullptr, false);
}

int ccg_pca_knn_range(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, size_t begin, size_t end,
                      float *eig3_sorted, float *normal4_sorted) {
    CCG_TRY(begin_call(ctx, true));
    if (k < 1 || k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", k, CCG_MAX_K);
    if (begin % kLeaf) return set_error(ctx, CCG_ERR_INVALID, "begin must be a multiple of 32");
    if (end > ctx->n) end = ctx->n;
    return pca_common(ctx, k, INFINITY, cov_norm, 0, viewpoint3, begin, end, 1, eig3_sorted, normal4_sorted, nullptr, false);
}

int ccg_pca_radius(ccg_ctx *ctx, double r, int max_nn, int cov_norm, int normalise_hist, const float *viewpoint3,
                   float *eig3, float *normal4, int32_t *count) {
    CCG_TRY(begin_call(ctx, true));
    if (!(r >= 0.0)) return set_error(ctx, CCG_ERR_INVALID, "radius must be >= 0");
    if (max_nn < 0) return set_error(ctx, CCG_ERR_INVALID, "max_nn must be >= 0");
    const float rb = radius_bound(r, 1);
    // FLANN caps only when 0 < max_nn < n; otherwise every point inside the radius is used
    bool uncapped = (max_nn == 0) || ((size_t)max_nn >= ctx->n);
    if (!uncapped && max_nn > CCG_MAX_K) {
        // a cap beyond the register lists: per-query counting select of the max_nn-th distance, then one streaming pass over
        // {d2 < T or (d2 == T and index <= I)} — FLANN's "max_nn nearest inside r" for any cap (radsel.cu)
        if (cov_norm < 0 || cov_norm > 2) return set_error(ctx, CCG_ERR_INVALID, "cov_norm must be 0, 1 or 2");
        const size_t n = ctx->n;
        OutBuf oe, on, oc;
        CCG_TRY(oe.open(ctx, 0, eig3, n * 12));
        CCG_TRY(on.open(ctx, 1, normal4, n * 16));
        CCG_TRY(oc.open(ctx, 2, count, n * 4));
        CCG_TRY(query_begin(ctx));
        const float *Tp; const int32_t *Ip;
        CCG_TRY(launch_radius_rule(ctx, rb, max_nn, &Tp, &Ip));
        StreamArgs s{};
        s.cov_norm = cov_norm; s.min_pts = 3; s.normalise_hist = normalise_hist; s.has_vp = viewpoint3 ? 1 : 0;
        if (viewpoint3) for (int i = 0; i < 3; ++i) s.vp[i] = viewpoint3[i];
        s.eig3 = oe.dev<float>(); s.normal4 = on.dev<float>(); s.count = oc.dev<int32_t>();
        CCG_TRY(launch_rule_pca(ctx, rb, Tp, Ip, s));
        CCG_TRY(query_end(ctx));
        CCG_TRY(oe.close()); CCG_TRY(on.close()); CCG_TRY(oc.close());
        return finish_call(ctx);
    }
    return pca_common(ctx, max_nn, rb, cov_norm, normalise_hist, viewpoint3, 0, ctx->n, 0, eig3, normal4, count, uncapped);
}

int ccg_features(ccg_ctx *ctx, const float *eig3, size_t n, int formula_set, float *out) {
    CCG_TRY(begin_call(ctx, false));
    if (n == 0) return CCG_OK;
    if (!eig3 || !out) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    const float *e = eig3;
    if (!is_dev(ctx, eig3)) {
        CCG_CUDA(ctx, ctx->tmp[1].reserve(n * 12));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, eig3, n * 12, cudaMemcpyHostToDevice, ctx->stream));
        e = ctx->tmp[1].as<float>();
    }
    OutBuf o;
    CCG_TRY(o.open(ctx, 0, out, n * 28));
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_features(ctx, e, n, formula_set, o.dev<float>()));
    CCG_TRY(query_end(ctx));
    CCG_TRY(o.close());
    return finish_call(ctx);
}

int ccg_eigen_scalar(ccg_ctx *ctx, const float *eig3, size_t n, int which, float *out) {
    CCG_TRY(begin_call(ctx, false));
    if (n == 0) return CCG_OK;
    if (!eig3 || !out) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (which < 0 || which > 1) return set_error(ctx, CCG_ERR_INVALID, "which must be 0 or 1");
    const float *e = eig3;
    if (!is_dev(ctx, eig3)) {
        CCG_CUDA(ctx, ctx->tmp[1].reserve(n * 12));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, eig3, n * 12, cudaMemcpyHostToDevice, ctx->stream));
        e = ctx->tmp[1].as<float>();
    }
    OutBuf o;
    CCG_TRY(o.open(ctx, 0, out, n * 4));
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_eigen_scalar(ctx, e, n, which, o.dev<float>()));
    CCG_TRY(query_end(ctx));
    CCG_TRY(o.close());
    return finish_call(ctx);
}

int ccg_sor_mean_range(ccg_ctx *ctx, int mean_k, size_t begin, size_t end, float *mean_d_sorted) {
    CCG_TRY(begin_call(ctx, true));
    if (mean_k < 1 || mean_k + 1 > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "mean_k=%d outside [1,%d]", mean_k, CCG_MAX_K - 1);
    if (begin % kLeaf) return set_error(ctx, CCG_ERR_INVALID, "begin must be a multiple of 32");
    if (end > ctx->n) end = ctx->n;
    if (!mean_d_sorted) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    OutBuf om;
    CCG_TRY(om.open(ctx, 0, mean_d_sorted, (end > begin ? end - begin : 0) * 4));
    QueryArgs a = base_args(ctx);
    a.mode = MODE_SOR; a.k = mean_k + 1; a.mean_k = mean_k; a.begin = begin; a.end = end; a.sorted_rows = 1; a.mean_d = om.dev<float>();
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_tile_query(ctx, a));
    CCG_TRY(query_end(ctx));
    CCG_TRY(om.close());
    return finish_call(ctx);
}

int ccg_sor_mask(ccg_ctx *ctx, int mean_k, double stddev_mul, uint8_t *outlier, float *mean_d) {
    CCG_TRY(begin_call(ctx, true));
    if (mean_k < 1 || mean_k + 1 > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "mean_k=%d outside [1,%d]", mean_k, CCG_MAX_K - 1);
    if (!outlier) return set_error(ctx, CCG_ERR_INVALID, "outlier is NULL");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    CCG_CUDA(ctx, ctx->tmp[0].reserve(n * 4));
    float *md = ctx->tmp[0].as<float>();
    QueryArgs a = base_args(ctx);
    a.mode = MODE_SOR; a.k = mean_k + 1; a.mean_k = mean_k; a.mean_d = md;
    CCG_TRY(query_begin(ctx));
    // The threshold is a sum over all points in INDEX ORDER in fp64 (PCL StatisticalOutlierRemoval); a parallel
    // reduction would round differently, so the n floats go to the host and are summed sequentially there.
    // This is the contract's arithmetic, not a fallback: the per-point work stays on the GPU.  On large clouds the
    // 30-40 ms host loop hides behind the query: tiles run in order of their smallest original index, mean distances
    // arrive in row-prefix chunks on the copy stream, and the host adds each chunk (in index order) while the GPU works
    // on the next one.
    CCG_CUDA(ctx, ctx->hpin[0].reserve(n * 4));
    const float *h = ctx->hpin[0].as<float>();
    double sum = 0.0, sq = 0.0;
    if (ctx->pipeline_chunks > 1 && !ctx->collect_stats && n >= (size_t)ctx->pipeline_min_points) {
        OutBuf ob;
        ob.ctx = ctx; ob.user = ctx->hpin[0].p; ob.bytes = n * 4; ob.d = md; ob.staged = true;
        PipeOut po{&ob, 4};
        size_t bound[18];
        int nch = 0;
        CCG_TRY(pipelined_query(ctx, a, &po, 1, ctx->pipeline_chunks, bound, &nch));
        for (int c = 0; c < nch; ++c) {
            CCG_CUDA(ctx, cudaEventSynchronize(ctx->ev_copy[c]));
            for (size_t i = bound[c]; i < bound[c + 1]; ++i) { sum += (double)h[i]; sq += (double)h[i] * (double)h[i]; }
        }
    } else {
        CCG_TRY(launch_tile_query(ctx, a));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->hpin[0].p, md, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (size_t i = 0; i < n; ++i) { sum += (double)h[i]; sq += (double)h[i] * (double)h[i]; }
    }
    const double mu = sum / (double)n;
    const double var = n > 1 ? (sq - sum * sum / (double)n) / ((double)n - 1.0) : 0.0;
    const double thr = mu + stddev_mul * sqrt(var > 0 ? var : 0.0);
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, outlier, n));
    k_sor_mask<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(md, n, thr, oo.dev<uint8_t>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    if (mean_d) {
        if (is_dev(ctx, mean_d)) CCG_CUDA(ctx, cudaMemcpyAsync(mean_d, md, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        else memcpy(mean_d, h, n * 4);
    }
    return finish_call(ctx);
}

/* The second half of the statistical outlier filter for SHARDED runs: the ranks computed their mean distances with
 * ccg_sor_mean_range and gathered them (original order); one rank turns them into the mask with the contract's arithmetic —
 * fp64 sums in index order on the host, threshold mu + stddev_mul * sigma, compare in fp64. */
int ccg_sor_finish(ccg_ctx *ctx, const float *mean_d, size_t n, double stddev_mul, uint8_t *outlier) {
    CCG_TRY(begin_call(ctx, false));
    if (n == 0) return CCG_OK;
    if (!mean_d || !outlier) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    CCG_TRY(query_begin(ctx));
    const float *h = mean_d, *d = mean_d;
    if (is_dev(ctx, mean_d)) {
        CCG_CUDA(ctx, ctx->hpin[0].reserve(n * 4));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->hpin[0].p, mean_d, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        h = ctx->hpin[0].as<float>();
    } else {
        CCG_CUDA(ctx, ctx->tmp[0].reserve(n * 4));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[0].p, mean_d, n * 4, cudaMemcpyHostToDevice, ctx->stream));
        d = ctx->tmp[0].as<float>();
    }
    double sum = 0.0, sq = 0.0;
    for (size_t i = 0; i < n; ++i) { sum += (double)h[i]; sq += (double)h[i] * (double)h[i]; }
    const double mu = sum / (double)n;
    const double var = n > 1 ? (sq - sum * sum / (double)n) / ((double)n - 1.0) : 0.0;
    const double thr = mu + stddev_mul * sqrt(var > 0 ? var : 0.0);
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, outlier, n));
    k_sor_mask<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d, n, thr, oo.dev<uint8_t>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_radius_count_mask(ccg_ctx *ctx, double r, int min_neighbours, int prefilter, uint8_t *outlier) {
    CCG_TRY(begin_call(ctx, true));
    if (!(r >= 0.0) || min_neighbours < 1) return set_error(ctx, CCG_ERR_INVALID, "need r >= 0 and min_neighbours >= 1");
    if (!outlier) return set_error(ctx, CCG_ERR_INVALID, "outlier is NULL");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    cudaStream_t st = ctx->stream;
    CCG_CUDA(ctx, ctx->tmp[0].reserve(n * 4));
    int32_t *cnt = ctx->tmp[0].as<int32_t>();
    uint8_t *cand = nullptr;
    CCG_TRY(query_begin(ctx));
    if (prefilter) {
        CCG_CUDA(ctx, ctx->tmp[1].reserve(n));
        const unsigned long long *ks; const unsigned *vs;
        cand = ctx->tmp[1].as<uint8_t>();
        CCG_TRY(voxel_sort(ctx, (float)(2.0 * r), &ks, &vs));
        k_voxel_candidates<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ks, vs, n, min_neighbours, cand);
        ctx->stats.launches += 1;
    }
    StreamArgs s{};
    s.mode = 0; s.rb = radius_bound(r, 0); s.count_cap = min_neighbours; s.begin = 0; s.end = n; s.count = cnt;
    CCG_TRY(launch_tile_stream(ctx, s));
    OutBuf oo;
    CCG_TRY(oo.open(ctx, 0, outlier, n));
    k_count_mask<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(cnt, cand, n, min_neighbours, oo.dev<uint8_t>());
    ctx->stats.launches++;
    CCG_TRY(query_end(ctx));
    CCG_TRY(oo.close());
    return finish_call(ctx);
}

int ccg_grid_normals(ccg_ctx *ctx, const int32_t *cloud_to_grid, int scan_w, int scan_h, int do_fallback,
                     float *normal4, int32_t *n_missing) {
    CCG_TRY(begin_call(ctx, true));
    if (!cloud_to_grid || !normal4) return set_error(ctx, CCG_ERR_INVALID, "NULL buffer");
    if (scan_w < 1 || scan_h < 1) return set_error(ctx, CCG_ERR_INVALID, "scan grid must be at least 1 x 1");
    const size_t n = ctx->n;
    if (n == 0) { if (n_missing) *n_missing = 0; return CCG_OK; }
    const int32_t *c2g = cloud_to_grid;
    if (!is_dev(ctx, cloud_to_grid)) {
        CCG_CUDA(ctx, ctx->out[3].reserve(n * 4));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->out[3].p, cloud_to_grid, n * 4, cudaMemcpyHostToDevice, ctx->stream));
        c2g = ctx->out[3].as<int32_t>();
    }
    OutBuf on;
    CCG_TRY(on.open(ctx, 1, normal4, n * 16));
    CCG_TRY(query_begin(ctx));
    int32_t still = 0;
    CCG_TRY(launch_grid_normals(ctx, c2g, scan_w, scan_h, do_fallback, on.dev<float>(), &still));
    CCG_TRY(query_end(ctx));
    CCG_TRY(on.close());
    if (n_missing) *n_missing = still;
    return finish_call(ctx);
}

int ccg_unpermute(ccg_ctx *ctx, const void *rows_sorted, size_t row_bytes, void *rows_original) {
    CCG_TRY(begin_call(ctx, true));
    if (!rows_sorted || !rows_original || row_bytes == 0 || (row_bytes & 3)) return set_error(ctx, CCG_ERR_INVALID, "bad rows");
    const size_t n = ctx->n;
    if (n == 0) return CCG_OK;
    const void *in = rows_sorted;
    if (!is_dev(ctx, rows_sorted)) {
        CCG_CUDA(ctx, ctx->tmp[1].reserve(n * row_bytes));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->tmp[1].p, rows_sorted, n * row_bytes, cudaMemcpyHostToDevice, ctx->stream));
        in = ctx->tmp[1].p;
    }
    OutBuf o;
    CCG_TRY(o.open(ctx, 0, rows_original, n * row_bytes));
    CCG_TRY(query_begin(ctx));
    CCG_TRY(launch_unpermute(ctx, in, row_bytes, o.d));
    CCG_TRY(query_end(ctx));
    CCG_TRY(o.close());
    return finish_call(ctx);
}

}  // extern "C"
