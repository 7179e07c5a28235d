// common.cuh — internals shared by the libccgpu translation units (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include "../../include/ccgpu.h"

namespace ccg {

constexpr int kLeaf = 32;        // points per leaf of the Morton order == one warp-wide float4 load (512 B)
constexpr int kFan = 32;         // children per inner node == one lane per child
constexpr int kRelStride = 33;   // float4 per leaf of the prefilter block array (centre + 16 pair-interleaved pairs, build.cu)
constexpr int kMaxLevels = 8;    // 32^8 leaves is far beyond any cloud
constexpr unsigned kFull = 0xffffffffu;

// Device view of the bounding-box hierarchy over the Morton-sorted cloud (passed by value to kernels).
struct TreeView {
    const float4 *pts;               // sorted points, npad = nleaf*32; w = original index (int bits); pads = +inf
    const float4 *orig;              // original-order points (w unused), for the PCA gather
    const float4 *rel;               // per leaf kRelStride float4: prefilter block (leaf-centre-relative, pair-interleaved)
    const float4 *lo[kMaxLevels];    // per level: box min (w unused)
    const float4 *hi[kMaxLevels];    // per level: box max
    int count[kMaxLevels];           // nodes per level; level 0 = leaves
    int nlevels;                     // count[nlevels-1] == 1
    int n;                           // valid points
};

// grow-only device buffer
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct HostBuf {   // grow-only pinned host buffer
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

}  // namespace ccg

struct ccg_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;       // the stream calls run on (own_stream unless ccg_set_stream)
    cudaStream_t own_stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev[8] = {};
    cudaEvent_t ev_chunk[17] = {};
    cudaEvent_t ev_copy[17] = {};        // D2H of chunk c done (host-side consumers, e.g. the SOR index-order sum)       // output pipelining (created on first use)
    const volatile int *cancel = nullptr;
    int async_mode = 0;              // ccg_set_async: entry points enqueue and return; ccg_sync collects
    int pending_stats = 0;
    int pending_cloud_check = 0;     // an asynchronous build left its bbox / finiteness word in hpin[1]

    bool has_cloud = false;
    size_t n = 0;            // points
    size_t npad = 0;         // nleaf * 32
    int nleaf = 0;
    int nlevels = 0;
    int level_count[ccg::kMaxLevels] = {};
    size_t level_off[ccg::kMaxLevels] = {};   // offset (in float4) of each level inside boxes_lo / boxes_hi

    ccg::DevBuf raw;         // staged AoS input (host clouds only)
    ccg::DevBuf orig;        // float4[n]   original order
    ccg::DevBuf sorted;      // float4[npad] curve order, w = original index
    ccg::DevBuf rel;         // float4[nleaf * kRelStride] prefilter blocks
    ccg::DevBuf keys[2];     // uint64[n] Morton keys (double buffer)
    ccg::DevBuf vals[2];     // uint32[n] original index (double buffer); vals[order_sel] is the final order
    int order_sel = 0;
    ccg::DevBuf sort_tmp;    // radix sort scratch
    ccg::DevBuf boxes_lo, boxes_hi;
    ccg::DevBuf scratch;     // bbox / flags / counters
    ccg::DevBuf out[4];      // staged outputs for host destinations
    ccg::DevBuf tmp[4];      // per-call temporaries
    ccg::DevBuf vox_out;     // result of the last ccg_voxel_downsample: float[vox_n * vox_nf]
    int64_t vox_n = 0; int vox_nf = 0;
    ccg::DevBuf ptx_bytes, ptx_starts, ptx_points, ptx_c2g;   // last ccg_ptx_parse: file bytes, line starts, 32-byte rows, cloud_to_grid_map
    int64_t ptx_n = 0;
    ccg::DevBuf node_mom;         // 3 float4 per box of every level (laid out like boxes_lo): count and moments of its points about the
    bool node_mom_valid = false;  // box centre; filled on the first uncapped radius sum after a build (ensure_node_moments, build.cu)
    int node_moments = 1;         // CCG_NODE_MOMENTS=0: the uncapped radius sums visit every point (A/B switch)
    ccg::DevBuf rule_T, rule_I;   // per sorted position: d2 threshold / tie index threshold of a capped radius search (radsel.cu)
    ccg::DevBuf defer, defer_b;   // [count | query list | output rows] of queries deferred to the exact point kernel (double buffer)
    int defer_sel = 0, side_pending = 0, last_side = 0;
    cudaStream_t side_stream = nullptr;  // the point kernel of launch i runs here, next to tile launch i+1
    cudaEvent_t ev_side[2] = {}, ev_tile = nullptr;
    ccg::HostBuf hpin[2];    // pinned bounce buffers
    float bb_lo[3] = {0, 0, 0}, bb_hi[3] = {0, 0, 0};

    ccg_stats stats = {};
    int collect_stats = 0;
    int pipeline_chunks = 8;         // host-output pipelining of the PCA query (CCG_PIPELINE_CHUNKS; <= 1 disables)
    int pipeline_min_points = 1 << 18;
    int shard_world = 0, shard_rank = 0, shard_chunk = 0;   // set by ccg_*_shard around pca_common
    int route_world = 0; size_t route_rows = 0;             // set by ccg_pca_knn_features_route around pca_common
    float *route_ptr[3][16] = {};
    int knn_use_topk = 0;    // CCG_KNN_TOPK=1: ordered lists from the 64-bit keyed k_tile_topk instead of k_tile_sel<MODE_KNN>
    int key_bits = 16;       // curve-key bits per axis, 4..21 (CCG_KEY_BITS; 13 fits key + index of 33 M points in one word)
    int curve_hilbert = 1;   // leaf order along the Hilbert curve (CCG_CURVE=morton: Z-order, for A/B runs)
    int sel_drain_at = 0;    // k_tile_sel: pending survivors per lane that trigger a drain; 0 = only at the end of the list (CCG_SEL_DRAIN_AT)
    int sel_net = 1;         // k_tile_sel: own leaf seeded by a register sorting network when k == K (CCG_SEL_NET=0: ordered inserts)
    int sel_minb = 0;        // k_tile_sel<20>: 12 = the 80-register build (CCG_SEL_MINB=12), else the default 16-block build
    int sel_stage = 1;       // k_tile_sel leaf staging: 1 = cp.async.bulk + mbarrier ring, 0 = LDG/STS (CCG_SEL_STAGE=ldg)
    int sel_chunks = 4;      // sub-launches of a long k_tile_sel launch (CCG_SEL_CHUNKS)
    int point_overlap = 1;   // deferred exact queries on a side stream (CCG_POINT_OVERLAP=0: same stream)
    int split_sparse = 1;    // move isolated points to the end of the order (CCG_SPLIT_SPARSE=0 disables)
    char err[512] = "";

    ccg::TreeView view() const {
        ccg::TreeView t;
        t.pts = sorted.as<float4>();
        t.orig = orig.as<float4>();
        t.rel = rel.as<float4>();
        for (int l = 0; l < ccg::kMaxLevels; ++l) {
            t.lo[l] = boxes_lo.as<float4>() + level_off[l];
            t.hi[l] = boxes_hi.as<float4>() + level_off[l];
            t.count[l] = l < nlevels ? level_count[l] : 0;
        }
        t.nlevels = nlevels;
        t.n = (int)n;
        return t;
    }
};

namespace ccg {

int set_error(ccg_ctx *ctx, int code, const char *fmt, ...);

#define CCG_CUDA(ctx, call)                                                                     \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess)                                                                 \
            return ccg::set_error((ctx), CCG_ERR_CUDA, "%s failed: %s (%s:%d)", #call,          \
                                  cudaGetErrorString(e__), __FILE__, __LINE__);                 \
    } while (0)

#define CCG_TRY(expr)                                   \
    do {                                                \
        int rc__ = (expr);                              \
        if (rc__ != CCG_OK) return rc__;                \
    } while (0)

// build.cu
int build_cloud(ccg_ctx *ctx, const void *xyz, size_t n, size_t stride_bytes);
int finish_cloud_check(ccg_ctx *ctx);
int ensure_node_moments(ccg_ctx *ctx);

// query.cu
enum TileMode { MODE_KNN = 0, MODE_PCA = 1, MODE_SOR = 2 };
struct QueryArgs {
    int mode;
    int k;                  // list entries wanted (kNN k, or capped max_nn)
    float rb;               // candidates need d2 <= rb (+inf for pure kNN)
    size_t begin, end;      // sorted-position range (begin multiple of 32)
    int sorted_rows;        // 1: output row = pos - begin, 0: output row = original index
    // MODE_KNN
    int32_t *idx; float *d2; int32_t *count;
    // MODE_PCA
    int cov_norm; int min_pts; int normalise_hist; int has_vp; float vp[3];
    float *eig3; float *normal4; float *feat7; int formula_set;
    // MODE_SOR
    int mean_k; float *mean_d;
    // MODE_PCA / MODE_SOR: optional processing order; begin/end then count 32-query SLOTS of this list (x32), not positions
    const int32_t *tile_order;
    // interleaved shard (multi-GPU load balance): this call answers the tiles t with (t / shard_chunk) % shard_world ==
    // shard_rank and ignores begin/end; shard_world <= 1 = off
    int shard_world, shard_rank, shard_chunk;
    // owner-routed output rows (ccg_pca_knn_features_route); route_world <= 1 = off
    int route_world; size_t route_rows; float *const *r_eig; float *const *r_nrm; float *const *r_feat;
    int async_side;         // 1: the caller joins the side stream itself (join_side) — output pipelining
};
int launch_tile_query(ccg_ctx *ctx, const QueryArgs &a);
int join_side(ccg_ctx *ctx);

struct StreamArgs {
    int mode;               // 0: count (early-out at count_cap when > 0), 1: moments -> PCA
    float rb;
    int count_cap;
    size_t begin, end;
    int sorted_rows;
    int32_t *count;
    int cov_norm; int min_pts; int normalise_hist; int has_vp; float vp[3];
    float *eig3; float *normal4; float *feat7; int formula_set;
};
int launch_tile_stream(ccg_ctx *ctx, const StreamArgs &a);

// warp-per-query kernel over an explicit list of query points (original indices, device pointer)
struct PointArgs {
    const int32_t *queries; int nq;
    int mode, k; float rb;
    int list_rows;          // 1: output row = position in the list, 0: row = original index
    int32_t *idx; float *d2; int32_t *count;
    int cov_norm; int min_pts; int normalise_hist; int has_vp; float vp[3];
    float *eig3; float *normal4; float *feat7; int formula_set;
    int mean_k; float *mean_d;
};
int launch_point_query(ccg_ctx *ctx, const PointArgs &a);

// radsel.cu: capped radius neighbourhoods for any cap (rule arrays) and the list-free reductions over them
int launch_radius_rule(ccg_ctx *ctx, float rb, int max_nn, const float **Tpos, const int32_t **Ipos);
int launch_rule_pca(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const StreamArgs &a);
int launch_rule_stdev_dist(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, int use_depth, const float centre[3], float *out);
int launch_rule_normal_stdev(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const void *normals, size_t stride, float *out);
int launch_rule_hist(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, int bins, int which, double radius, const void *attr,
                     size_t stride, float *hist);

int launch_radius_csr_count(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, long long *offsets_dev, long long *total_host);
int launch_radius_csr_fill(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const long long *offsets_dev, long long total,
                           int32_t *idx_dev, float *d2_dev);

int launch_tile_attrsum(ccg_ctx *ctx, float rb, const void *attr_dev, size_t stride, float *sum3, int32_t *count);
int launch_tile_attrsum2(ccg_ctx *ctx, float rb1, float rb2, const void *attr_dev, size_t stride, float *sum1, int32_t *count1,
                         float *sum2, int32_t *count2);
int launch_tile_principal_curv(ccg_ctx *ctx, float rb, const void *attr_dev, size_t stride, float *pc5);
int launch_tile_fpfh(ccg_ctx *ctx, float rb, const void *attr_dev, size_t stride, float *spfh33, float *fpfh33);
int launch_features(ccg_ctx *ctx, const float *eig3, size_t n, int formula_set, float *out);
int launch_eigen_scalar(ccg_ctx *ctx, const float *eig3, size_t n, int which, float *out);
int launch_unpermute(ccg_ctx *ctx, const void *in, size_t row_bytes, void *out);

// voxel.cu
int voxel_sort(ccg_ctx *ctx, float edge, const unsigned long long **keys_sorted, const unsigned **vals_sorted);
int voxel_downsample(ccg_ctx *ctx, const void *points_dev, size_t stride, int nf, float resolution, int as_written,
                     int32_t *big_to_small_dev, int64_t *n_out);

// gridnormals.cu
int launch_grid_normals(ccg_ctx *ctx, const int32_t *c2g_dev, int scan_w, int scan_h, int do_fallback, float *normal4_dev,
                        int32_t *n_missing_host);

}  // namespace ccg
