/*
 * ccgpu.h — C ABI of libccgpu.so: CloudClean's per-point neighbourhood hot path on one B200 (sm_100a).
 *
 * This is the drop-in boundary.  Each entry point replaces one serial
 * `for (i < cloud->size())` loop of the reference, which today calls PCL/FLANN/Eigen in-process; the file:line
 * of the loop it replaces is cited on every declaration (paths relative to the reference tree).
 *
 * Conventions
 *   - Every function returns CCG_OK (0) or a negative CCG_ERR_* code; nothing throws; ccg_last_error()
 *     gives the message.  There is no CPU fallback: without a usable CUDA device ccg_create fails.
 *   - One ccg_ctx drives ONE GPU (one process per GPU; multi-GPU runs replicate the cloud and shard the
 *     query range, see ccg_*_range).  A ctx may be used from any single thread at a time — the reference's
 *     std::async normal-estimation worker (normalestimation.cpp:196) must be able to call it — so every call
 *     sets the device itself and relies on no thread-local CUDA state.
 *   - All buffers are caller-allocated.  Input and output pointers may be HOST pointers (pageable or
 *     pinned) or DEVICE pointers of the ctx's GPU; the library detects which (cudaPointerGetAttributes)
 *     and stages copies only when needed.
 *   - Indices are 32-bit *original* cloud indices (the internal Morton permutation never leaks), except in
 *     the *_range entry points which are explicitly in sorted order.
 *   - Distances are the contract's fp32 d2 = ((dx*dx + dy*dy) + dz*dz), no fma; neighbour order is the
 *     total order (d2, original index); the query point itself is a neighbour (d2 = 0).
 */
#ifndef CCGPU_H
#define CCGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCG_VERSION 100

enum {
    CCG_OK = 0,
    CCG_ERR_INVALID = -1,      /* bad argument */
    CCG_ERR_CUDA = -2,         /* CUDA runtime error (message has the call) */
    CCG_ERR_NO_CLOUD = -3,     /* query before ccg_set_cloud */
    CCG_ERR_NONFINITE = -4,    /* cloud holds NaN/Inf coordinates (PCL's kd-tree rejects those too) */
    CCG_ERR_UNSUPPORTED = -5,  /* parameter outside the compiled range (e.g. k > CCG_MAX_K) */
    CCG_ERR_CANCELLED = -6     /* the caller's cancel flag was raised between stages */
};

#define CCG_MAX_K 64           /* largest k / capped max_nn kept as an ordered list */

/* covariance normalisation of the 3x3 scatter S = sum (p-mu)(p-mu)^T */
enum {
    CCG_COV_SCATTER = 0,       /* S            — pcl::PCA in PCL 1.7.0 (README.md:26-27) */
    CCG_COV_NM1 = 1,           /* S/(n-1)      — pcl::PCA in PCL >= 1.8 (CMakeLists.txt:110)  [default] */
    CCG_COV_N = 2              /* S/n          — pcl::NormalEstimation::computePointNormal */
};

typedef struct ccg_ctx ccg_ctx;

/* Per-call statistics of the most recent query on a ctx (device time from CUDA events on the ctx stream). */
typedef struct ccg_stats {
    float ms_upload;           /* H2D of the cloud (0 for device-resident input) */
    float ms_build;            /* ingest + curve (Hilbert) keys + radix sort + reorder + bounding-box levels */
    float ms_query;            /* the query kernels of the last call */
    float ms_download;         /* D2H of the outputs (0 for device-resident outputs) */
    int32_t launches;          /* kernels of this library launched by the last call (build counted by set_cloud) */
    int32_t launches_build;
    int64_t n_points;
    int64_t n_leaves;          /* 32-point leaves of the curve order */
    int64_t leaves_scanned;    /* leaf scans summed over query tiles (0 unless CCG_COLLECT_STATS=1) */
    int64_t deferred_queries;  /* queries finished by the warp-per-query kernel */
} ccg_stats;

/* ---- lifecycle ------------------------------------------------------------------------------------ */
int ccg_create(ccg_ctx **out, int device_id);
void ccg_destroy(ccg_ctx *ctx);
const char *ccg_last_error(const ccg_ctx *ctx);     /* never NULL; owned by the ctx (static text for ctx==NULL) */
int ccg_version(void);
int ccg_get_stats(const ccg_ctx *ctx, ccg_stats *out);
/* Optional cancellation: checked between stages, honours PointCloud::deleting_ (normalestimation.cpp:305-306). */
int ccg_set_cancel_flag(ccg_ctx *ctx, const volatile int *flag);
/* Run every later call on the caller's CUDA stream (a cudaStream_t, e.g. the stream NCCL collectives are enqueued
 * on, so that broadcast -> build -> query -> gather needs no extra synchronisation); NULL restores the ctx's own. */
int ccg_set_stream(ccg_ctx *ctx, void *cuda_stream);
/* Asynchronous mode (off by default).  A synchronous call returns when its results are in the caller's buffers — what the
 * reference's call sites expect (cv.h:96-113 reads the result in the next statement).  With async on, ccg_set_cloud and
 * the query entry points only ENQUEUE their work on the ctx stream (device or pinned host buffers; pageable host buffers
 * still work but make the copies blocking) and return at once, so that a caller which owns the stream — broadcast ->
 * build -> query -> gather over NCCL on several GPUs — chains the stages without a host round trip in between.  Results,
 * the deferred input checks (non-finite coordinates) and the timing statistics become valid after ccg_sync.  Calls whose
 * contract needs the host in the middle (ccg_sor_mask: index-order fp64 sum; capped radius searches with max_nn > 64:
 * two error bits; CCG_COLLECT_STATS) still synchronise internally. */
int ccg_set_async(ccg_ctx *ctx, int on);
/* Wait for everything enqueued on the ctx's streams; returns the first deferred error (e.g. CCG_ERR_NONFINITE of an
 * asynchronous ccg_set_cloud), CCG_OK otherwise.  Harmless in synchronous mode. */
int ccg_sync(ccg_ctx *ctx);

/* ---- cloud ---------------------------------------------------------------------------------------- */
/* Replaces pcl::KdTreeFLANN::setInputCloud (cv.h:83-84, cv.cpp:254-255, normalestimation.cpp:428-429) and the
 * async octree build (pointcloud.cpp:312-323).  `xyz` points at the first float of an array of structs with
 * `stride_bytes` between points: 32 for pcl::PointXYZI, 48 for pcl::PointXYZINormal, 16 for packed float4,
 * 12 for packed xyz.  Host or device pointer.  The data is copied before return. */
int ccg_set_cloud(ccg_ctx *ctx, const void *xyz, size_t n, size_t stride_bytes);
int64_t ccg_cloud_size(const ccg_ctx *ctx);
/* sorted position -> original index (int32[n]); the space-filling-curve order used to shard queries across GPUs */
int ccg_get_order(ccg_ctx *ctx, int32_t *order);

/* ---- a1: exact kNN -------------------------------------------------------------------------------- */
/* Replaces the per-point pcl::KdTreeFLANN::nearestKSearch loops: normalestimation.cpp:436-439 (k=50),
 * flood.cpp:306-327 (k=8), graphcut/mincut.cpp:462-466 (k=14).  idx, d2: n*k, row i = neighbours of point i,
 * ascending (d2, index), self first.  Rows are padded with idx=-1, d2=+inf when k > n.  d2 may be NULL. */
int ccg_knn(ccg_ctx *ctx, int k, int32_t *idx, float *d2);

/* ---- a2 / a3: fixed-radius search ----------------------------------------------------------------- */
/* Replaces KdTreeFLANN::radiusSearch (cv.h:98, cv.cpp:273, featureeval/utils.h:50) with strict_lt=1
 * [d2 < (float)(r*r)] and OctreePointCloudSearch::radiusSearch (outlierfilter.cpp:169, cv.cpp:194) with
 * strict_lt=0 [(double)d2 <= r*r].  1 <= max_nn <= CCG_MAX_K: the max_nn nearest inside the radius, ascending;
 * idx/d2: n*max_nn padded with -1/+inf; count[i] = entries kept.  idx and d2 may be NULL (count only), and
 * with idx==d2==NULL max_nn may be 0 (count everything inside the radius). */
int ccg_radius(ccg_ctx *ctx, double r, int max_nn, int strict_lt, int32_t *idx, float *d2, int32_t *count);

/* ---- a12: kNN + PCA normals / curvature ----------------------------------------------------------- */
/* The north-star op (PCL NormalEstimation definition): kNN(k, self included) -> 3x3 covariance -> symmetric
 * eigensolve.  eig3: n*3 eigenvalues descending (scaled per cov_norm).  normal4: n*4 = unit eigenvector of the
 * smallest eigenvalue + curvature l2/(l0+l1+l2); flipped towards viewpoint3 (3 floats) when that is non-NULL
 * (pcl::flipNormalTowardsViewpoint, cf. the commented call at normalestimation.cpp:417).  Either output may be NULL. */
int ccg_pca_knn(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, float *eig3, float *normal4);

/* Same with the a9 epilogue fused into the kernel: feat7 (n*7, see ccg_features) is computed from the eigenvalues
 * while they are still in registers, so Markov::mkFeatureVector's inputs (markov.cpp:449-475) cost no second pass. */
int ccg_pca_knn_features(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, float *eig3, float *normal4,
                         float *feat7);

/* ---- a4 / a5 / a6: radius + PCA (getPCA / getHist) ------------------------------------------------ */
/* Replaces the body of getPCA<PointT> (cv.h:96-113) and getHist (cv.cpp:271-295): strict radius search with
 * the max_nn cap (0 = uncapped; ANY positive cap: 20, tests.json's 1000, markov.cpp:54's 100000 — beyond CCG_MAX_K the
 * max_nn-th distance comes from a counting select, radsel.cu), `<3 neighbours -> (0,0,1)` marker (cv.h:105-109), else PCA eigenvalues
 * descending.  normalise_hist!=0 applies getHist's L2 normalisation (cv.cpp:293).  normal4 / count optional.
 * count is exact for every cap; the moments are fp32 sums in traversal order (uncapped: boxes of the hierarchy that lie inside
 * the radius enter as one aggregated term each), eigenvalues within 1e-4 * lambda0 of the fp64 truth. */
int ccg_pca_radius(ccg_ctx *ctx, double r, int max_nn, int cov_norm, int normalise_hist, const float *viewpoint3,
                   float *eig3, float *normal4, int32_t *count);

/* ---- a9: eigen-derived features ------------------------------------------------------------------- */
/* out: n*7 = anisotropy, planarity, sphericity, linearity, omnivariance, eigenentropy, curvature.
 * formula_set 0 = exactly as written in Markov::mkFeatureVector (markov.cpp:449-475, including its slips),
 * 1 = textbook definitions. */
int ccg_features(ccg_ctx *ctx, const float *eig3, size_t n, int formula_set, float *out);
/* featureeval.cpp:1104-1137 (which=0: eigen ratio) and :1201-1220 (which=1: plane cosine similarity); out: n */
int ccg_eigen_scalar(ccg_ctx *ctx, const float *eig3, size_t n, int which, float *out);

/* ---- a11: statistical outlier filter -------------------------------------------------------------- */
/* PCL StatisticalOutlierRemoval definition: mean distance to the mean_k nearest neighbours (self dropped),
 * global mean + stddev_mul * sigma threshold, fp64 sums in index order.  outlier: n bytes (1 = outlier);
 * mean_d: n floats or NULL. */
int ccg_sor_mask(ccg_ctx *ctx, int mean_k, double stddev_mul, uint8_t *outlier, float *mean_d);

/* ---- a10: radius / min-neighbour outlier filter --------------------------------------------------- */
/* Replaces OutlierFilter::filter's loop (outlierfilter.cpp:157-173): outlier[i] = 1 when the inclusive r-ball
 * of point i (self included) holds fewer than min_neighbours points.  prefilter!=0 also applies the
 * octreeDownsample(2r) candidate step (:126-147): only points whose voxel of edge 2r holds fewer than
 * min_neighbours points can be flagged (voxel lattice: first cloud point at a voxel centre). */
int ccg_radius_count_mask(ccg_ctx *ctx, double r, int min_neighbours, int prefilter, uint8_t *outlier);

/* ---- a7: scan-grid triangle-fan normals ----------------------------------------------------------- */
/* Replaces NormalEstimator::estimateNormals (normalestimation.cpp:268-470).  cloud_to_grid: int32[n] grid index
 * of every point (PointCloud::cloudToGridMap, pointcloud.cpp:389-391); scan_w/scan_h as PointCloud::scan_width()/
 * scan_height().  normal4: n*4 = nx,ny,nz,data_n[3]; NaN x4 where the fan has no face.  do_fallback!=0 runs stage
 * (iii): kNN(50) mean of neighbours' normals for the missing ones, in order.  *n_missing (optional) = normals
 * still NaN at the end. */
int ccg_grid_normals(ccg_ctx *ctx, const int32_t *cloud_to_grid, int scan_w, int scan_h, int do_fallback,
                     float *normal4, int32_t *n_missing);

/* ---- scan-grid normals without any allocation ------------------------------------------------------ */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Stages (i) + (ii) of
 * ccg_grid_normals — the result of ccg_grid_normals(do_fallback = 0), bit for bit — with the kernel's scratch
 * (grid_to_cloud, int32 per grid cell) in a file-scope __device__ array of CCG_STATIC_MAX_CELLS cells instead of the
 * context's buffers: 16 MiB that every GPU loading the library holds until exit.  The call allocates nothing (no
 * cudaMalloc / cudaFree / allocating library routine), copies nothing to or from the host, and does NOT synchronise:
 * the kernels are queued on cuda_stream (NULL = the context's stream) and the caller waits on that stream.
 *   xyz_dev           DEVICE pointer, n points, x y z as float32 at the start of every stride_bytes (>= 12, multiple of 4);
 *                     the cloud given to ccg_set_cloud is NOT used
 *   cloud_to_grid_dev DEVICE int32[n]
 *   normal4_dev       DEVICE float32[n * 4]: nx, ny, nz, 1; NaN x4 where the fan has no face
 *   status2_dev       DEVICE int32[2] or NULL: [0] = normals missing (NaN rows), [1] = 1 when cloud_to_grid held an index
 *                     outside the grid (such points are skipped; ccg_grid_normals returns CCG_ERR_INVALID for them)
 * scan_w * scan_h > CCG_STATIC_MAX_CELLS returns CCG_ERR_UNSUPPORTED.  Every call on a GPU shares the one array and
 * clears what it uses first: calls that overlap on the GPU (two streams, two threads, two contexts) race; calls queued on
 * one stream do not. */
#define CCG_STATIC_MAX_CELLS 4194304
int ccg_grid_normals_static(ccg_ctx *ctx, void *cuda_stream, const void *xyz_dev, size_t stride_bytes, size_t n,
                            const int32_t *cloud_to_grid_dev, int scan_w, int scan_h, float *normal4_dev, int32_t *status2_dev);

/* ---- multi-GPU sharding ----------------------------------------------------------- */
/* Every rank holds the whole cloud and builds the identical Morton order; rank g owns sorted positions
 * [begin, end) (multiples of 32 except the last end = n).  Outputs are compact, row (pos - begin), in SORTED
 * order; ccg_unpermute scatters gathered rows back to original order. */
int ccg_pca_knn_range(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, size_t begin, size_t end,
                      float *eig3_sorted, float *normal4_sorted);
int ccg_pca_knn_features_range(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, size_t begin, size_t end,
                               float *eig3_sorted, float *normal4_sorted, float *feat7_sorted);
int ccg_sor_mean_range(ccg_ctx *ctx, int mean_k, size_t begin, size_t end, float *mean_d_sorted);
/* Second half of the filter for sharded runs: mean_d = the n gathered mean distances in ORIGINAL order (host
 * or device); index-order fp64 sums on the host, threshold mu + stddev_mul * sigma, compare in fp64 -> outlier (n bytes, host
 * or device).  ccg_sor_mask == ccg_sor_mean_range over the whole order + un-permute + ccg_sor_finish, bit for bit. */
int ccg_sor_finish(ccg_ctx *ctx, const float *mean_d, size_t n, double stddev_mul, uint8_t *outlier);
/* out[order[i]] = in[i] for rows of row_bytes (multiple of 4) bytes; in/out host or device */
int ccg_unpermute(ccg_ctx *ctx, const void *rows_sorted, size_t row_bytes, void *rows_original);

/* ---- fused query + scatter over NVLink peer memory (SURVEY.md §8e "collect") -------------------------------
 * Same range query, but rows are written in ORIGINAL order at row = original index, straight from the query
 * kernel.  The output pointers may be PEER memory of another GPU on the same NVSwitch domain (a buffer of the
 * collecting rank opened with ccg_ipc_open): every rank then stores its finished rows directly into the
 * collector's arrays through NVLink while it computes — no all-gather, no un-permute pass, no staging copy.
 * Ranks own disjoint rows, so no two GPUs write the same address; the caller orders the collector's reads
 * after all ranks' calls with one barrier-like collective on the stream (bench.py: a 1-element all_reduce). */
int ccg_pca_knn_features_scatter(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, size_t begin, size_t end,
                                 float *eig3, float *normal4, float *feat7);
/* Same, for an INTERLEAVED shard instead of one contiguous range: the Morton order is cut into chunks of chunk_points
 * (a multiple of 32; 8192 is a good value) and rank r answers chunks r, r+world, r+2*world, ...  Query cost per point
 * varies along the order (volumetric vegetation vs flat walls, the isolated points collected at its end), so equal
 * contiguous ranges leave one GPU with several times the work of another; interleaving averages that out
 * (DESIGN.md §5).
 * sorted_rows = 0: row = original index (12/16/28-byte stores scattered over the collector's arrays: fine inside one
 * GPU and for 2 GPUs, but measured at only ~100 GB/s of NVLink ingress when 3-7 peers scatter into one GPU).
 * sorted_rows = 1: row = position in the Morton order (arrays of ceil(n/32)*32 rows): every warp stores 32 consecutive
 * rows, i.e. 384/512/896 contiguous bytes per array, which crosses NVLink at full packet size; the collector then
 * runs ccg_unpermute once per array on its own HBM. */
int ccg_pca_knn_features_shard(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, int rank, int world,
                               int chunk_points, int sorted_rows, float *eig3, float *normal4, float *feat7);
/* Variable-length neighbour lists in CSR form — what KdTreeFLANN::radiusSearch(i, r, idx, d2, max_nn) fills into its
 * std::vector<int> / std::vector<float> (cv.h:98, cv.h:162, cv.cpp:273, featureeval/utils.h:50,145, markov.cpp:360-367) for
 * EVERY cap: max_nn = 0 (all points inside r) or any positive cap (the max_nn nearest, index-ordered ties).  Row i is
 * idx[offsets[i] .. offsets[i+1]) / d2[...], ascending (d2, index), self first.  offsets: int64[n+1].  Call once with
 * idx = d2 = NULL to learn *total (and offsets), then with buffers of `capacity` >= *total entries.  One call handles up to
 * 2^31-1 neighbours in total.  ccg_radius remains the fixed-width form for caps <= 64. */
int ccg_radius_csr(ccg_ctx *ctx, double r, int max_nn, int strict_lt, int64_t *offsets, int32_t *idx, float *d2, int64_t capacity,
                   int64_t *total);

/* Owner-routed variant of ccg_pca_knn_features_shard for end-to-end multi-GPU runs: the output rows are partitioned by
 * ORIGINAL index — rank o owns rows [o * rows_per_rank, (o + 1) * rows_per_rank) — and this rank stores every row it
 * computes (its interleaved chunks of the curve order) at local row (i - o * rows_per_rank) of rank o's arrays,
 * eig3_of_rank[o] / normal4_of_rank[o] / feat7_of_rank[o] (own memory for o == rank, NVLink peer memory opened with
 * ccg_ipc_open otherwise; any of the three tables may be NULL).  After a barrier every rank holds a contiguous,
 * original-order slice of the result, which it can copy to the host through its own PCIe link — no rank funnels
 * everything (the gather step of the north star, spread over all links).  world <= 16. */
int ccg_pca_knn_features_route(ccg_ctx *ctx, int k, int cov_norm, const float *viewpoint3, int formula_set, int rank, int world,
                               int chunk_points, size_t rows_per_rank, float *const *eig3_of_rank, float *const *normal4_of_rank,
                               float *const *feat7_of_rank);

/* Plain cudaMalloc / cudaFree on the ctx's device (IPC handles need an allocation base, which a caching
 * allocator's sub-block is not). */
int ccg_dev_alloc(ccg_ctx *ctx, size_t bytes, void **out);
int ccg_dev_free(ccg_ctx *ctx, void *ptr);
/* handle64: 64 bytes (cudaIpcMemHandle_t) to be sent to the other local ranks by any means. */
int ccg_ipc_export(ccg_ctx *ctx, const void *ptr, void *handle64);
int ccg_ipc_open(ccg_ctx *ctx, const void *handle64, void **out);     /* enables peer access lazily */
int ccg_ipc_close(ccg_ctx *ctx, void *ptr);

/* ---- f1: voxel downsample (octreeDownsample, src/utilities/utils.h:16-62) ----------------------------------
 * The step in front of every feature call of the reference (featureeval.cpp:687,995,1029; markov.cpp:306;
 * flood.cpp:269; outlierfilter.cpp:128).  `points` is the SAME array of structs that was given to ccg_set_cloud
 * (host or device); the first n_floats floats of every struct are averaged per voxel of edge `resolution`
 * (the reference averages sizeof(PointT)/sizeof(float) floats, padding included: 8 for PointXYZI).
 * as_written != 0 reproduces the reference's accumulator, a default-constructed PCL point whose float[3] is 1
 * (utils.h:41-48), so float[3] of a voxel comes out as (count + 1) / count; 0 gives the plain mean.
 * Voxel lattice: the first cloud point sits at a voxel centre, voxel(p) = floor((p - (p0 - res/2)) / res) in fp32
 * (PCL grows its octree box by whole voxels from the first point; its exact lattice depends on insertion history and
 * cannot be reproduced — set-level parity only).  Voxels are numbered in ascending packed (x,y,z) key
 * order; big_to_small[i] (int32[n], optional) = voxel of point i (utils.h:50).  *n_out = number of voxels.
 * ccg_voxel_downsample_fetch then copies the n_out * n_floats packed means to out_points (host or device). */
int ccg_voxel_downsample(ccg_ctx *ctx, const void *points, size_t stride_bytes, int n_floats, float resolution, int as_written,
                         int32_t *big_to_small, int64_t *n_out);
int ccg_voxel_downsample_fetch(ccg_ctx *ctx, float *out_points);

/* ---- f3: kNN graph for the serial graph consumers ---------------------------------------------------------
 * Replaces the per-node tree queries of MinCut::buildGraph (graphcut/mincut.cpp:458-475, k = 14; the markov copy
 * at markov/mincut.cpp:461) and hands Flood::flood (flood.cpp:313-356) a precomputed neighbour table: idx is the
 * ccg_knn table (n*k, self first, ascending (d2, index)), weight (n*k doubles) is calculateBinaryPotential
 * (mincut.cpp:574-585) of every edge (i, idx[i][j]): exp(-inverse_sigma * (dx*dx + dy*dy + dz*dz)) with the three float
 * products summed in double.  The reference adds edges for j >= 1 only (:467); weight[i][0] (self) is 1.  The graph
 * algorithms themselves stay on the CPU. */
int ccg_knn_graph(ccg_ctx *ctx, int k, double inverse_sigma, int32_t *idx, double *weight);

/* ---- f2: radius-neighbourhood reductions (same search as a2/a3, different epilogue) ------------------------
 * Any max_nn the reference accepts: 0 = every point inside the radius, else the max_nn nearest (featureeval's default is
 * 20, featureeval.cpp:67-68; the GUI value reaches the calls at featureeval.cpp:816,999,1032).  With 1 <= max_nn <=
 * CCG_MAX_K the neighbours are visited in ascending (d2, index) order and the fp32 sums are the reference's bit for bit;
 * with 0 or a larger cap the same neighbour SET is streamed in traversal order (sums within 1e-5 of the value range;
 * histograms and counts exact).  Arithmetic and quirks are kept as written.
 * ccg_normal_stdev — normal_stdev<PointT,NormalT> (cv.h:123-206): per point the standard deviation of the normalised
 *   angle acos(clamp(cos,0,1))/pi between its normal and its neighbours' normals (strict kd-tree radius); `normals` is an
 *   array of structs whose first three floats are the normal (pcl::Normal: stride 32), host or device, indexed like the
 *   cloud.  0 for NaN normals, fewer than 3 neighbours and NaN results.  The reference's cosine is
 *   dot / |neighbour| * |current| (operator precedence, cv.h:181-182) and is reproduced.
 * ccg_stdev_dist — stdev_dist (cv.cpp:166-230): per point (sum_sq - sum^2/m)/(m-1) of the distances from its
 *   neighbours (inclusive octree radius) to the point itself, or to centre3 (the sensor origin) when use_depth — a
 *   variance despite the name (cv.cpp:222-223).  The octree's own cap keeps the first max_nn points in traversal
 *   order [EXT]; here the cap keeps the nearest, so results agree with the reference wherever the cap does not bind. */
int ccg_normal_stdev(ccg_ctx *ctx, double r, int max_nn, const void *normals, size_t stride_bytes, float *out);
int ccg_stdev_dist(ccg_ctx *ctx, double r, int max_nn, int use_depth, const float *centre3, float *out);

/* ccg_neighbour_hist — calcDistHist (featureeval/utils.h:18-110, which = 0) and calcIntensityHist (:113-195, which = 1):
 *   hist is n*bins floats; per point, over its strict-radius neighbourhood (any max_nn, 0 = uncapped), every
 *   neighbour adds float(1.0/m) to bin int((sqrt(d2)/r)*(bins-1)) [which = 0] or int(attr*(bins-1)) [which = 1, attr =
 *   one float per point at attr + i*attr_stride_bytes, e.g. &points[0].intensity with stride 32]; rows of points
 *   with fewer than 3 neighbours stay 0.  Bin indices are clamped to [0, bins-1] (the reference would index out of
 *   bounds for intensities outside [0,1]). */
int ccg_neighbour_hist(ccg_ctx *ctx, double r, int max_nn, int bins, int which, const void *attr, size_t attr_stride_bytes, float *hist);

/* ccg_don — don<PointT,NormalT>(cloud, normals, radius1, radius2) (featureeval.cpp:694-739), the difference of normals:
 *   per point, for each of the two (strict, UNCAPPED) radius neighbourhoods the mean of the neighbours' normals with the
 *   point's own normal counted twice (the search returns it and :729 appends it again), then (mean1 - mean2)/2.
 *   normals: array of structs, first three floats = normal (stride 32 for pcl::Normal), host or device.  don4: n*4 floats
 *   (nx, ny, nz, 0).  Sums run in traversal order in fp32 (the reference's order is its tree's, unsorted): 1e-5 agreement.
 *   The density count of markov.cpp:359-370 is ccg_radius(r, 0, 1, NULL, NULL, count). */
int ccg_don(ccg_ctx *ctx, const void *normals, size_t stride_bytes, double r1, double r2, float *don4);

/* ccg_principal_curvatures — pcl::PrincipalCurvaturesEstimation with setRadiusSearch(r) [EXT], as called at
 *   markov.cpp:389-402 and featureeval.cpp:947-956: per point, the normals of its (strict, UNCAPPED) radius neighbourhood
 *   are projected into its tangent plane (M = I - n n^T), their scatter matrix is decomposed, and the row is
 *   pcl::PrincipalCurvatures' five floats {pcx, pcy, pcz (unit eigenvector of the largest eigenvalue, sign free),
 *   pc1 = l_max / N, pc2 = l_mid / N} — the stride-5 / offset-3 layout featureeval.cpp:963-964 reads.
 *   normals: array of structs, first three floats = normal, host or device.  Sums run in traversal order in fp32
 *   (one pass over shifted moments; PCL runs two passes in its tree's order): pc1/pc2 agree to ~1e-5 relative. */
int ccg_principal_curvatures(ccg_ctx *ctx, double r, const void *normals, size_t stride_bytes, float *pc5);

/* ccg_fpfh — pcl::FPFHEstimation<.., pcl::FPFHSignature33> with setRadiusSearch(r) [EXT], as called at
 *   featureeval.cpp:925-932: pass 1, per point the SPFH (3 x 11 bins of the Darboux pair features f1 = atan2(w.n2, n1.n2),
 *   f2 = v.n2, f3 = n1.d/|d| against every other point of its strict, UNCAPPED radius neighbourhood, increments of
 *   100/(N-1)); pass 2, per point the sum of its neighbours' SPFH rows weighted by 1/d2 (d2 = 0 skipped, so the point's own
 *   SPFH does not enter — as PCL), each group of 11 normalised to sum 100.  fpfh33: n*33 floats (pcl::FPFHSignature33);
 *   spfh33 (optional, n*33): the pass-1 rows.  Pair arithmetic in fp32, bin indices in double as PCL writes them; a pair
 *   whose feature lies within rounding of a bin edge may land in the adjacent bin (statistical agreement, see the test). */
int ccg_fpfh(ccg_ctx *ctx, double r, const void *normals, size_t stride_bytes, float *fpfh33, float *spfh33);

/* ---- f4: PTX ingest (PointCloud::load_ptx, src/model/pointcloud.cpp:127-330) -------------------------------
 * The reference reads the Leica PTX ASCII format with one fscanf per line — for 10^7..10^8 lines the user's longest
 * wait.  Here the 10 header lines are read on the host and the body is parsed on the GPU (one thread per line, decimal
 * -> float conversions correctly rounded like strtof; line starts, line numbers and cloud indices by device scans).
 *   bytes/nbytes         the whole file as it lies on disk (host pointer — e.g. an mmap — or device pointer)
 *   decimation_factor    1 or even (:130); lines with (row+1)%d != 0 or (col+1)%d != 0 are skipped (:223-230),
 *                        row = line / file_width, col = line % file_width exactly as written there
 *   result (kept in the ctx until the next parse): n = info->points rows of 32 bytes = pcl::PointXYZI as load_ptx
 *     leaves it {x, y, z, 1.0f, intensity, 0, float(index in the cloud), float(index in the sampled grid)} (:284-286;
 *     the reference leaves data_c[1] uninitialised, here it is 0) and cloud_to_grid_map int32[n] (:287).
 *     (0,0,0) returns are dropped (:263-265); at most file_width*file_height lines are read; a file that ends
 *     early fails with CCG_ERR_INVALID ("File read fail at line", :232-235) exactly when the reference does: when a
 *     sampled line is reached after the stream already reported end-of-file (one missing last sampled line can slip
 *     through as a plain failed fscanf, :240-246 — then the cloud is one point short, as there).
 *   differences, all refusals instead of guesses: a sampled line that is not "x y z intensity" (+ "r g b" when the
 *     first body line has 7 space-separated tokens, :199-216) fails the call with CCG_ERR_INVALID and the line number
 *     (the reference would continue from the middle of that line); numbers with more than 19 significant digits,
 *     magnitude outside [1e-37, 3e38], nan/inf/hex floats fail with CCG_ERR_UNSUPPORTED.  Blank lines are ignored.
 *     Lines may have any length (the reference's fgets(buff, 1024) splits lines beyond 1023 characters when it skips
 *     them).  The bounding box of :268-281 (which writes the maxima into min_bounding_box_) is not produced.
 * ccg_ptx_load = open + mmap + ccg_ptx_parse (what load_ptx(filename, decimation_factor) becomes);
 * ccg_ptx_fetch copies the rows / the map out (either may be NULL; host or device destinations);
 * ccg_ptx_set_cloud makes the parsed rows the ctx's cloud (ccg_set_cloud on the device-resident rows, no host trip). */
typedef struct ccg_ptx_info {
    int32_t file_width, file_height;      /* header lines 1-2 */
    int32_t scan_width, scan_height;      /* file_* / decimation_factor (:157-158) */
    int32_t has_rgb;                      /* first body line had 7 tokens */
    float sensor_origin[3];               /* header line 3 (:161-162) */
    float orientation[9];                 /* header lines 4-6 row-major as read; the reference stores
                                             Quaternionf(orientation^T) (:167-172) */
    int64_t lines;                        /* non-blank body lines found */
    int64_t sampled;                      /* lines that passed the decimation rule (sampled_idx at the end) */
    int64_t points;                       /* rows of the cloud */
    size_t body_offset;                   /* byte offset of the first body line */
} ccg_ptx_info;
int ccg_ptx_parse(ccg_ctx *ctx, const void *bytes, size_t nbytes, int decimation_factor, ccg_ptx_info *info);
int ccg_ptx_load(ccg_ctx *ctx, const char *filename, int decimation_factor, ccg_ptx_info *info);
int ccg_ptx_fetch(ccg_ctx *ctx, void *points32, int32_t *cloud_to_grid);
int ccg_ptx_set_cloud(ccg_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* CCGPU_H */
