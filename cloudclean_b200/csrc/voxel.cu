// voxel.cu — voxel lattice helpers: the sorted voxel keys shared by the outlier filter's candidate step and the
// f1 row of SURVEY.md §8: octreeDownsample (src/utilities/utils.h:16-62), the step in front of every feature call
// of the reference (featureeval.cpp:687,995,1029; markov.cpp:306; flood.cpp:269; outlierfilter.cpp:128).
//
// The reference walks a PCL octree and, per leaf voxel, averages ALL float fields of the points inside
// (utils.h:41-55, accumulator = a default-constructed PointT, so float[3] starts at 1) and records the
// big -> small index map.  Here: voxel key per point (lattice: the first cloud point sits at a voxel centre, the
// box grows by whole edges — the same rule the oracle states), one device radix sort of (key, index) pairs
// (stable, so points of a voxel stay in index order like PCL's leaf index vectors), segment heads by a flagged
// select, and one warp per voxel for the mean.  HBM-streaming: 16 B key/index traffic x 8 sort passes + one gather of
// the caller's AoS (stride bytes per point) + nf floats per voxel out.
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace ccg {

// voxel key of octreeDownsample's lattice (utils.h:29-31 via outlierfilter.cpp:128): first point at a voxel centre
__global__ void __launch_bounds__(256) k_voxel_keys(const float4 *__restrict__ orig, size_t n, float edge,
                                                    unsigned long long *__restrict__ keys, unsigned *__restrict__ vals) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p0 = orig[0], p = orig[i];
    const float h = __fmul_rn(edge, 0.5f);
    long long kx = (long long)floorf(__fdiv_rn(__fsub_rn(p.x, __fsub_rn(p0.x, h)), edge));
    long long ky = (long long)floorf(__fdiv_rn(__fsub_rn(p.y, __fsub_rn(p0.y, h)), edge));
    long long kz = (long long)floorf(__fdiv_rn(__fsub_rn(p.z, __fsub_rn(p0.z, h)), edge));
    keys[i] = ((unsigned long long)(kx & 0x1fffff) << 42) | ((unsigned long long)(ky & 0x1fffff) << 21) | (unsigned long long)(kz & 0x1fffff);
    vals[i] = (unsigned)i;
}

// (key, original index) pairs of the whole cloud sorted by key; buffers live in ctx->tmp[2] / tmp[3]
int voxel_sort(ccg_ctx *ctx, float edge, const unsigned long long **keys_sorted, const unsigned **vals_sorted) {
    const size_t n = ctx->n;
    cudaStream_t st = ctx->stream;
    CCG_CUDA(ctx, ctx->tmp[2].reserve(n * 16 + 64));
    CCG_CUDA(ctx, ctx->tmp[3].reserve(n * 8 + 64));
    unsigned long long *k0 = ctx->tmp[2].as<unsigned long long>(), *k1 = k0 + n;
    unsigned *v0 = ctx->tmp[3].as<unsigned>(), *v1 = v0 + n;
    k_voxel_keys<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ctx->orig.as<float4>(), n, edge, k0, v0);
    cub::DoubleBuffer<unsigned long long> dk(k0, k1);
    cub::DoubleBuffer<unsigned> dv(v0, v1);
    size_t tmp_bytes = 0;
    CCG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, 63, st));
    CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
    CCG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tmp_bytes, dk, dv, (int)n, 0, 63, st));
    ctx->stats.launches += 10;
    *keys_sorted = dk.Current();
    *vals_sorted = dv.Current();
    return CCG_OK;
}

struct IsHead {
    const unsigned long long *keys;
    __host__ __device__ __forceinline__ bool operator()(int i) const { return i == 0 || keys[i] != keys[i - 1]; }
};

// One warp per voxel: lane-strided sums of nf floats per point (index order inside each lane), shuffle tree,
// then (sum + default-constructed accumulator) * (float)(1.0 / count)   [utils.h:41-55].
template <int NF>
__global__ void __launch_bounds__(256) k_voxel_mean(const char *__restrict__ pts, size_t stride, int nf, const unsigned *__restrict__ vals,
                                                    const int *__restrict__ starts, int nvox, int n, int as_written,
                                                    float *__restrict__ out, int32_t *__restrict__ big_to_small) {
    const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (v >= nvox) return;
    const int b = starts[v], e = v + 1 < nvox ? starts[v + 1] : n;
    float acc[NF];
#pragma unroll
    for (int f = 0; f < NF; ++f) acc[f] = 0.f;
    for (int j = b + lane; j < e; j += 32) {
        const unsigned idx = vals[j];
        if (big_to_small) big_to_small[idx] = v;
        const float *p = reinterpret_cast<const float *>(pts + (size_t)idx * stride);
#pragma unroll
        for (int f = 0; f < NF; ++f)
            if (f < nf) acc[f] = __fadd_rn(acc[f], __ldg(p + f));
    }
#pragma unroll
    for (int f = 0; f < NF; ++f)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[f] = __fadd_rn(acc[f], __shfl_xor_sync(kFull, acc[f], o));
    if (lane == 0) {
        const float size_inv = (float)(1.0 / (double)(e - b));
#pragma unroll
        for (int f = 0; f < NF; ++f)
            if (f < nf) {
                float s = acc[f];
                if (as_written && f == 3) s = __fadd_rn(s, 1.0f);       // PCL points are constructed with data[3] = 1
                out[(size_t)v * nf + f] = __fmul_rn(s, size_inv);
            }
    }
}

int voxel_downsample(ccg_ctx *ctx, const void *points, size_t stride, int nf, float resolution, int as_written,
                     int32_t *big_to_small_dev, int64_t *n_out) {
    const size_t n = ctx->n;
    cudaStream_t st = ctx->stream;
    const unsigned long long *ks; const unsigned *vs;
    CCG_TRY(voxel_sort(ctx, resolution, &ks, &vs));
    // segment heads -> starts[]
    CCG_CUDA(ctx, ctx->tmp[1].reserve((n + 2) * sizeof(int)));
    int *starts = ctx->tmp[1].as<int>();
    int *nsel = (int *)((char *)ctx->scratch.p + 3072);
    cub::CountingInputIterator<int> iota(0);
    IsHead pred{ks};
    size_t tmp_bytes = 0;
    CCG_CUDA(ctx, cub::DeviceSelect::If(nullptr, tmp_bytes, iota, starts, nsel, (int)n, pred, st));
    CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
    CCG_CUDA(ctx, cub::DeviceSelect::If(ctx->sort_tmp.p, tmp_bytes, iota, starts, nsel, (int)n, pred, st));
    int nvox = 0;
    CCG_CUDA(ctx, cudaMemcpyAsync(&nvox, nsel, sizeof(int), cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaStreamSynchronize(st));
    ctx->stats.launches += 2;
    CCG_CUDA(ctx, ctx->vox_out.reserve((size_t)nvox * nf * sizeof(float)));
    const int blocks = (nvox + 7) / 8;
    const char *p = (const char *)points;
    float *out = ctx->vox_out.as<float>();
    if (nf <= 4) k_voxel_mean<4><<<blocks, 256, 0, st>>>(p, stride, nf, vs, starts, nvox, (int)n, as_written, out, big_to_small_dev);
    else if (nf <= 8) k_voxel_mean<8><<<blocks, 256, 0, st>>>(p, stride, nf, vs, starts, nvox, (int)n, as_written, out, big_to_small_dev);
    else if (nf <= 12) k_voxel_mean<12><<<blocks, 256, 0, st>>>(p, stride, nf, vs, starts, nvox, (int)n, as_written, out, big_to_small_dev);
    else k_voxel_mean<16><<<blocks, 256, 0, st>>>(p, stride, nf, vs, starts, nvox, (int)n, as_written, out, big_to_small_dev);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    ctx->vox_n = nvox;
    ctx->vox_nf = nf;
    *n_out = nvox;
    return CCG_OK;
}

}  // namespace ccg
