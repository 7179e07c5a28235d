// query.cu — the per-point neighbourhood kernels: exact kNN / fixed-radius search over the Morton box
// hierarchy with a fused covariance + eigensolve (normals, curvature, eigenvalues) or SOR epilogue.
//
// Replaces the serial loops  cv.h:96-113 (getPCA), cv.cpp:271-295 (getHist), normalestimation.cpp:436-459,
// outlierfilter.cpp:157-173  and the PCL StatisticalOutlierRemoval / NormalEstimation definitions.
//
// Mapping: one warp = one TILE = the 32 queries of one leaf (Morton-adjacent, so they share
// their candidate set).  The warp walks the box hierarchy bottom-up from its own leaf; at inner nodes one
// lane tests one child box against the tile box; a surviving leaf is re-tested per query, staged once in
// shared memory (a contiguous 512-byte run) and read back as broadcast LDS.128, so every lane scores the
// same candidate against its own query.  Each lane keeps its own bounded top-k as 64-bit (d2 bits, index)
// keys in registers, which makes the (d2, original index) tie-break a single unsigned compare.
//
// Exactness: d2 is evaluated with __fmul_rn/__fadd_rn in the contract's order (no fma); a box is pruned
// only if its fp32 box distance, built from the same monotone operations, is strictly greater than the
// current k-th d2, so ties at the k-th distance are always visited.
#include "tile.cuh"
#include <stdlib.h>

namespace ccg {


// ---------------------------------------------------------------------------------------------------
// bounded top-k in registers: K sorted 64-bit keys (d2 bits << 32 | original index)
// ---------------------------------------------------------------------------------------------------
typedef unsigned long long u64;
constexpr u64 kEmptyKey = ((u64)0x7f800000u << 32) | 0x7fffffffu;   // (+inf, INT_MAX): larger than any real key

template <int K>
struct TopK {
    u64 key[K];
    float bound;     // min(radius bound, current k-th d2); -1 for inactive lanes
    float rb;

    __device__ __forceinline__ void init(int k, float rb_, bool active) {
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = (i < K - k) ? 0ull : kEmptyKey;   // front slots pinned when k < K
        rb = rb_;
        bound = active ? rb_ : -1.f;
    }
    __device__ __forceinline__ void insert(u64 ck) {
        // precondition: ck < key[K-1]; keeps key[] ascending
        bool lt_i = true;
#pragma unroll
        for (int i = K - 1; i > 0; --i) {
            bool lt_p = ck < key[i - 1];
            key[i] = lt_p ? key[i - 1] : (lt_i ? ck : key[i]);
            lt_i = lt_p;
        }
        key[0] = lt_i ? ck : key[0];
    }
    // Score the 32 staged candidates of one leaf.  Phase 1 is branch-free: every lane marks the candidates
    // inside its bound in a 32-bit mask.  Phase 2 drains the masks together, so the (expensive, ~6 K
    // instructions) ordered insert runs max-over-lanes(popcount) times per leaf instead of once per candidate
    // that ANY lane accepts — the lanes' accepted candidates are different ones, and scoring them one candidate
    // at a time left the insert running with a handful of active lanes (ncu r01: 17.5 of 32).
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
        unsigned m = 0;
        const float b = bound;
#pragma unroll
        for (int j = 0; j < kLeaf; ++j) {
            const float d2 = d2_point(q, leaf[j]);
            m |= (d2 <= b) ? (1u << j) : 0u;
        }
        while (__any_sync(kFull, m != 0)) {
            if (m != 0) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const float4 c = leaf[j];
                const float d2 = d2_point(q, c);                 // same expression as phase 1: same bits
                const u64 ck = ((u64)__float_as_uint(d2) << 32) | (unsigned)__float_as_int(c.w);
                if (ck < key[K - 1]) {                           // d2 <= rb already holds (phase 1, b <= rb)
                    insert(ck);
                    bound = fminf(rb, __uint_as_float((unsigned)(key[K - 1] >> 32)));
                }
            }
        }
    }
};

// ---------------------------------------------------------------------------------------------------
// streaming visitor (no list): count and shifted moments of everything inside the radius
// ---------------------------------------------------------------------------------------------------
struct Stream {
    float bound, rb;
    int count, cap;
    bool moments;
    float m[9];   // sx sy sz sxx sxy sxz syy syz szz of (p - q)
    __device__ __forceinline__ void init(float rb_, int cap_, bool moments_, bool active) {
        rb = rb_; cap = cap_; moments = moments_; count = 0;
        bound = active ? rb_ : -1.f;
#pragma unroll
        for (int i = 0; i < 9; ++i) m[i] = 0.f;
    }
    __device__ __forceinline__ void candidate(const float4 &q, const float4 &c) {
        float d2 = d2_point(q, c);
        if (d2 <= bound) {
            ++count;
            if (moments) {
                float x = c.x - q.x, y = c.y - q.y, z = c.z - q.z;
                m[0] += x; m[1] += y; m[2] += z;
                m[3] += x * x; m[4] += x * y; m[5] += x * z; m[6] += y * y; m[7] += y * z; m[8] += z * z;
            }
            if (cap > 0 && count >= cap) bound = -1.f;   // enough neighbours seen: this lane is finished
        }
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 8
        for (int j = 0; j < kLeaf; ++j) candidate(q, leaf[j]);
    }
};

// the same sums with whole boxes taken at once (tile.cuh: visits_nodes, add_box_moments): a box inside the radius of every lane
// that wants it adds its count and its moments, shifted from the box centre to the query (per-box arrays: build.cu)
struct StreamNodes : Stream {
    static constexpr bool kNodes = true;
    const float4 *mom;
    __device__ __forceinline__ bool node_test() const { return true; }
    __device__ __forceinline__ bool whole(float dfar) const { return dfar <= bound; }
    __device__ __forceinline__ void add_node(size_t node, const float4 &lo, const float4 &hi, const float4 &q) {
        count += add_box_moments(mom, node, lo, hi, q, moments, m);
        if (cap > 0 && count >= cap) { count = cap; bound = -1.f; }   // what the per-point count stops at
    }
};

// streaming visitor that sums a 3-float per-point attribute (e.g. normals) over everything inside the radius;
// the attribute is gathered by ORIGINAL index (the candidate's w), so the caller's array is used as it lies
struct AttrSum {
    float bound;
    int count;
    float s[3];
    const char *attr; size_t stride;
    __device__ __forceinline__ void init(float rb_, bool active, const char *attr_, size_t stride_) {
        bound = active ? rb_ : -1.f; count = 0; s[0] = s[1] = s[2] = 0.f; attr = attr_; stride = stride_;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 4
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (d2_point(q, c) <= bound) {
                const float *a = reinterpret_cast<const float *>(attr + (size_t)__float_as_int(c.w) * stride);
                ++count;
                s[0] = __fadd_rn(s[0], __ldg(a)); s[1] = __fadd_rn(s[1], __ldg(a + 1)); s[2] = __fadd_rn(s[2], __ldg(a + 2));
            }
        }
    }
};

// two radii in one traversal (difference of normals): the walk is bounded by the larger one
struct AttrSum2 {
    float bound;                 // max(b1, b2): what the traversal prunes with
    float b1, b2;
    int c1, c2;
    float s1[3], s2[3];
    const char *attr; size_t stride;
    __device__ __forceinline__ void init(float rb1, float rb2, bool active, const char *attr_, size_t stride_) {
        b1 = active ? rb1 : -1.f; b2 = active ? rb2 : -1.f; bound = fmaxf(b1, b2);
        c1 = c2 = 0; attr = attr_; stride = stride_;
#pragma unroll
        for (int i = 0; i < 3; ++i) s1[i] = s2[i] = 0.f;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 4
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            const float d2 = d2_point(q, c);
            if (d2 <= bound) {
                const float *a = reinterpret_cast<const float *>(attr + (size_t)__float_as_int(c.w) * stride);
                const float x = __ldg(a), y = __ldg(a + 1), z = __ldg(a + 2);
                if (d2 <= b1) { ++c1; s1[0] = __fadd_rn(s1[0], x); s1[1] = __fadd_rn(s1[1], y); s1[2] = __fadd_rn(s1[2], z); }
                if (d2 <= b2) { ++c2; s2[0] = __fadd_rn(s2[0], x); s2[1] = __fadd_rn(s2[1], y); s2[2] = __fadd_rn(s2[2], z); }
            }
        }
    }
};

// streaming visitor for f4 (principal curvatures): first and second moments of (attr_j - ref) over everything inside
// the radius, ref = the query's own attribute — the shift keeps the sums small where the attribute varies slowly
struct AttrMoments {
    float bound;
    int count;
    float ref[3];
    float s[3], m[6];
    const char *attr; size_t stride;
    __device__ __forceinline__ void init(float rb_, bool active, const char *attr_, size_t stride_, const float4 &q) {
        bound = active ? rb_ : -1.f; count = 0; attr = attr_; stride = stride_;
#pragma unroll
        for (int i = 0; i < 3; ++i) s[i] = 0.f;
#pragma unroll
        for (int i = 0; i < 6; ++i) m[i] = 0.f;
        ref[0] = ref[1] = ref[2] = 0.f;
        if (active) {
            const float *a = reinterpret_cast<const float *>(attr + (size_t)__float_as_int(q.w) * stride);
            ref[0] = __ldg(a); ref[1] = __ldg(a + 1); ref[2] = __ldg(a + 2);
        }
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (d2_point(q, c) <= bound) {
                const float *a = reinterpret_cast<const float *>(attr + (size_t)__float_as_int(c.w) * stride);
                const float x = __ldg(a) - ref[0], y = __ldg(a + 1) - ref[1], z = __ldg(a + 2) - ref[2];
                ++count;
                s[0] += x; s[1] += y; s[2] += z;
                m[0] += x * x; m[1] += x * y; m[2] += x * z; m[3] += y * y; m[4] += y * z; m[5] += z * z;
            }
        }
    }
};

// f4 (FPFH, pass 1): pcl::computePairFeatures [EXT] for the pair (query p1/n1, candidate p2/n2) -> (f1, f2, f3); false when
// the points coincide or the Darboux frame is undefined
__device__ __forceinline__ bool pair_features(const float4 &p1, const float n1[3], const float4 &p2, const float n2[3],
                                              float &f1, float &f2, float &f3) {
    float dx = p2.x - p1.x, dy = p2.y - p1.y, dz = p2.z - p1.z;
    const float f4 = sqrtf(dx * dx + dy * dy + dz * dz);
    if (f4 == 0.f) return false;
    float ax = n1[0], ay = n1[1], az = n1[2], bx = n2[0], by = n2[1], bz = n2[2];
    const float angle1 = (ax * dx + ay * dy + az * dz) / f4;
    const float angle2 = (bx * dx + by * dy + bz * dz) / f4;
    if (acosf(fabsf(angle1)) > acosf(fabsf(angle2))) {          // the source is the point whose normal is closer to the line
        float t;
        t = ax; ax = bx; bx = t; t = ay; ay = by; by = t; t = az; az = bz; bz = t;
        dx = -dx; dy = -dy; dz = -dz;
        f3 = -angle2;
    } else f3 = angle1;
    // v = dp x n1, w = n1 x v
    float vx = dy * az - dz * ay, vy = dz * ax - dx * az, vz = dx * ay - dy * ax;
    const float vn = sqrtf(vx * vx + vy * vy + vz * vz);
    if (vn == 0.f) return false;
    vx /= vn; vy /= vn; vz /= vn;
    const float wx = ay * vz - az * vy, wy = az * vx - ax * vz, wz = ax * vy - ay * vx;
    f2 = vx * bx + vy * by + vz * bz;
    f1 = atan2f(wx * bx + wy * by + wz * bz, ax * bx + ay * by + az * bz);
    return true;
}

// SPFH of pcl::FPFHEstimation::computeSPFHSignature [EXT]: 3 x 11 bins per query, counts kept per lane in shared memory
// (hist[bin * 32 + lane]: one bank per lane), scaled by 100 / (N - 1) afterwards
struct Spfh {
    float bound;
    int count;
    int self;
    float nrm[3];
    const char *attr; size_t stride;
    unsigned *hist;              // this lane's column of the warp's [33][32] table
    __device__ __forceinline__ void init(float rb_, bool active, const char *attr_, size_t stride_, const float4 &q, unsigned *hist_) {
        bound = active ? rb_ : -1.f; count = 0; attr = attr_; stride = stride_; hist = hist_;
        self = __float_as_int(q.w);
        nrm[0] = nrm[1] = nrm[2] = 0.f;
        if (active) {
            const float *a = reinterpret_cast<const float *>(attr + (size_t)self * stride);
            nrm[0] = __ldg(a); nrm[1] = __ldg(a + 1); nrm[2] = __ldg(a + 2);
        }
#pragma unroll
        for (int b = 0; b < 33; ++b) hist[b * 32] = 0;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (d2_point(q, c) <= bound) {
                ++count;
                const int other = __float_as_int(c.w);
                if (other == self) continue;
                const float *a = reinterpret_cast<const float *>(attr + (size_t)other * stride);
                const float n2[3] = {__ldg(a), __ldg(a + 1), __ldg(a + 2)};
                float f1, f2, f3;
                if (!pair_features(q, nrm, c, n2, f1, f2, f3)) continue;
                // bin arithmetic in double as PCL writes it (M_PI and 1.0 / 0.5 are double literals; d_pi_ is a float)
                const float d_pi = 1.0f / (2.0f * 3.14159274101257324f);
                int b1 = (int)floor(11.0 * (((double)f1 + 3.14159265358979323846) * (double)d_pi));
                int b2 = (int)floor(11.0 * (((double)f2 + 1.0) * 0.5));
                int b3 = (int)floor(11.0 * (((double)f3 + 1.0) * 0.5));
                b1 = b1 < 0 ? 0 : (b1 > 10 ? 10 : b1);
                b2 = b2 < 0 ? 0 : (b2 > 10 ? 10 : b2);
                b3 = b3 < 0 ? 0 : (b3 > 10 ? 10 : b3);
                hist[b1 * 32] += 1; hist[(11 + b2) * 32] += 1; hist[(22 + b3) * 32] += 1;
            }
        }
    }
};

// FPFH pass 2 (weightPointSPFHSignature [EXT]): sum of the neighbours' SPFH rows weighted by 1 / d2, d2 == 0 skipped
struct FpfhSum {
    float bound;
    int count;
    float acc[33];
    const float *spfh;
    __device__ __forceinline__ void init(float rb_, bool active, const float *spfh_) {
        bound = active ? rb_ : -1.f; count = 0; spfh = spfh_;
#pragma unroll
        for (int b = 0; b < 33; ++b) acc[b] = 0.f;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            const float d2 = d2_point(q, c);
            if (d2 <= bound) {
                ++count;
                if (d2 == 0.f) continue;
                const float w = 1.0f / d2;
                const float *row = spfh + (size_t)__float_as_int(c.w) * 33;
#pragma unroll
                for (int b = 0; b < 33; ++b) acc[b] += __ldg(row + b) * w;
            }
        }
    }
};



// ---------------------------------------------------------------------------------------------------
// tile kernel with a top-k list
// ---------------------------------------------------------------------------------------------------
struct PointParams {
    const int32_t *queries;      // original indices; NULL = 0..nq-1
    int nq;
    int mode, k;
    float rb;
    int list_rows;               // 1: output row = position in `queries`, 0: row = original index
    int32_t *idx; float *d2; int32_t *count;
    PcaOut pca;
    int mean_k; float *mean_d;
    const int *nq_dev;           // non-NULL: the list length lives on the device (deferred queries of k_tile_sel)
    const long long *rows;       // non-NULL: explicit output row per list entry
};
__global__ void k_point_query(const __grid_constant__ TreeView T, const __grid_constant__ PointParams P);

struct TileParams {
    int mode, k;
    float rb;
    int tile_begin, tile_end;
    long long row_begin;          // sorted_rows: row = pos - row_begin
    int sorted_rows;
    int32_t *idx; float *d2; int32_t *count;
    PcaOut pca;
    int mean_k; float *mean_d;
    unsigned long long *stat_scanned;
    int *defer_count; int32_t *defer_list; long long *defer_row;   // k_tile_sel: queries handed to k_point_query
    int *slow_count; int32_t *slow_tiles;                          // k_tile_sel: tiles handed to k_tile_sel_stack
    int seed_net;                 // k_tile_sel: own leaf seeded by the sorting network when k == K (0: ordered inserts, for A/B runs)
    int drain_at;                 // k_tile_sel: pending survivors per lane that trigger a pooled drain
    const int32_t *tile_order;    // k_tile_sel: non-NULL = tile_begin/tile_end index this list of tiles (output pipelining)
    int shard_world, shard_rank, shard_chunk;   // k_tile_sel: > 1 = slots enumerate the interleaved chunks of one rank
};

template <int K>
__global__ void __launch_bounds__(128) k_tile_topk(const __grid_constant__ TreeView T, const __grid_constant__ TileParams P) {
    __shared__ WarpScratch s_ws[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = P.tile_begin + blockIdx.x * 4 + warp;
    if (tile >= P.tile_end) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    const int qidx = __float_as_int(q.w);
    TopK<K> v;
    v.init(P.k, P.rb, active);
    int scanned = traverse_tile(T, tile, q, v, s_ws[warp]);
    if (P.stat_scanned && lane == 0) atomicAdd(P.stat_scanned, (unsigned long long)scanned);

    const size_t row = P.sorted_rows ? (size_t)((long long)pos - P.row_begin) : (size_t)qidx;
    const int k = P.k;
    if (P.mode == MODE_KNN) {
        if (!active) return;
        int cnt = 0;
#pragma unroll
        for (int i = 0; i < K; ++i) {
            if (i >= K - k) {
                const int j = i - (K - k);
                const bool real = v.key[i] != kEmptyKey;
                cnt += real ? 1 : 0;
                if (P.idx) P.idx[row * k + j] = real ? (int)(unsigned)(v.key[i] & 0xffffffffu) : -1;
                if (P.d2) P.d2[row * k + j] = __uint_as_float((unsigned)(v.key[i] >> 32));
            }
        }
        if (P.count) P.count[row] = cnt;
    } else if (P.mode == MODE_PCA) {
        // gather the neighbours (original-order array) and accumulate moments of (p - q)
        float m[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) m[i] = 0.f;
        int cnt = 0;
#pragma unroll
        for (int i = 0; i < K; ++i) {
            if (i >= K - k) {
                const bool real = active && v.key[i] != kEmptyKey;
                if (real) {
                    const float4 p = T.orig[(unsigned)(v.key[i] & 0xffffffffu)];
                    float x = p.x - q.x, y = p.y - q.y, z = p.z - q.z;
                    m[0] += x; m[1] += y; m[2] += z;
                    m[3] += x * x; m[4] += x * y; m[5] += x * z; m[6] += y * y; m[7] += y * z; m[8] += z * z;
                    ++cnt;
                }
            }
        }
        if (!active) cnt = 0;
        PcaOut o = P.pca;
        if (!active) { o.eig3 = nullptr; o.normal4 = nullptr; o.feat7 = nullptr; }
        pca_finish(o, q, m, cnt, row);
        if (P.count && active) P.count[row] = cnt;
    } else {   // MODE_SOR: mean distance to the mean_k nearest, self (first entry) dropped
        if (!active) return;
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < K; ++i) {
            if (i > K - k) {
                float d = __uint_as_float((unsigned)(v.key[i] >> 32));
                if (v.key[i] != kEmptyKey) s += (double)__fsqrt_rn(d);
            }
        }
        P.mean_d[row] = (float)(s / (double)P.mean_k);
    }
}

// ---------------------------------------------------------------------------------------------------
// tile kernel, two-pass selection: MODE_PCA (kNN or capped radius -> covariance -> eigen) and MODE_SOR
// ---------------------------------------------------------------------------------------------------
// Candidate leaves of a tile, found ONCE and used by both passes.  After the tile's own leaf and its two
// neighbours in the Morton order have been scanned every lane has a finite bound; a top-down breadth-first walk
// with one lane per child box then lists every leaf whose box is within the warp's largest bound of the tile
// box (ballot + popc compaction into shared memory).  Compared with the nearest-first stack walk this costs a
// few hundred instructions per tile instead of ~100 per visited child, and pass 2 re-uses the list.
constexpr int kListCap = 253;     // leaves listed per tile (overflow: sub-box retry, then the point kernel); list[253..255] hold the home leaves
constexpr int kQueueCap = 128;     // inner nodes per level
struct WarpScratch2 {
    WarpScratch ws;
    int list[256];
    float4 sub[16];
};

// SUB: the tile box alone is a poor description of a tile that straddles a jump of the Morton curve (two far-apart
// clusters in one leaf): everything between the clusters is "inside" it.  Such tiles are described by 8 sub-boxes
// of 4 consecutive lanes each with that group's own largest bound (w.sub[2g], w.sub[2g+1].w = bound).
template <bool SUB>
__device__ __forceinline__ int collect_leaves(const TreeView &T, int tile, const float4 &tlo, const float4 &thi, float rmax,
                                              WarpScratch2 &w, int (*queue)[kQueueCap]) {
    const int lane = threadIdx.x & 31;
    const int top = T.nlevels - 1;
    if (top == 0) return 0;
    if (lane == 0) queue[0][0] = 0;
    __syncwarp();
    int cur = 0, ncur = 1, nlist = 0;
    for (int lv = top; lv >= 1; --lv) {
        int nnext = 0;
        for (int i = 0; i < ncur; ++i) {
            const int node = queue[cur][i];
            const int c = (node << 5) + lane;
            bool pass = false;
            if (c < T.count[lv - 1]) {
                const float4 clo = T.lo[lv - 1][c], chi = T.hi[lv - 1][c];
                pass = d2_box_box(tlo, thi, clo, chi) <= rmax;
                if (lv == 1) pass = pass && (c < tile - 1 || c > tile + 1);     // already scanned
                if (SUB && pass) {
                    pass = false;
#pragma unroll 1
                    for (int g = 0; g < 8; ++g) {
                        const float4 slo = w.sub[2 * g], shi = w.sub[2 * g + 1];
                        pass = pass || d2_box_box(slo, shi, clo, chi) <= shi.w;
                    }
                }
            }
            const unsigned m = __ballot_sync(kFull, pass);
            const int before = __popc(m & ((1u << lane) - 1u));
            if (lv == 1) {
                if (pass && nlist + before < kListCap) w.list[nlist + before] = c;
                nlist += __popc(m);
            } else {
                if (pass && nnext + before < kQueueCap) queue[cur ^ 1][nnext + before] = c;
                nnext += __popc(m);
            }
        }
        if (nnext > kQueueCap || nlist > kListCap) return -1;
        __syncwarp();
        cur ^= 1; ncur = nnext;
    }
    return nlist;
}

// boxes and largest bound of each group of 4 consecutive lanes -> w.sub (inactive lanes: bound -1, box from +inf pads)
__device__ __forceinline__ void make_sub_boxes(const float4 &q, float bound, bool active, WarpScratch2 &w) {
    const int lane = threadIdx.x & 31;
    float lx = active ? q.x : INFINITY, ly = active ? q.y : INFINITY, lz = active ? q.z : INFINITY;
    float hx = active ? q.x : -INFINITY, hy = active ? q.y : -INFINITY, hz = active ? q.z : -INFINITY;
    float b = bound;
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
        lx = fminf(lx, __shfl_xor_sync(kFull, lx, o)); ly = fminf(ly, __shfl_xor_sync(kFull, ly, o));
        lz = fminf(lz, __shfl_xor_sync(kFull, lz, o)); hx = fmaxf(hx, __shfl_xor_sync(kFull, hx, o));
        hy = fmaxf(hy, __shfl_xor_sync(kFull, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(kFull, hz, o));
        b = fmaxf(b, __shfl_xor_sync(kFull, b, o));
    }
    if ((lane & 3) == 0) {
        w.sub[2 * (lane >> 2)] = make_float4(lx, ly, lz, 0.f);
        w.sub[2 * (lane >> 2) + 1] = make_float4(hx, hy, hz, b);
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------------
// k_tile_sel, round 2: queue-based two-pass selection
// ---------------------------------------------------------------------------------------------------
// What r01's ncu profile said (profiles/r01_sel_v13_ncu_lines.txt) and what tools/sim_tiles.cpp adds to it: per tile the
// kernel prefiltered ~28 leaves and then DRAINED each leaf's survivors on the spot — ~70 instructions per round, with
// ~65 rounds per tile for the listed leaves although they hold only ~8 survivors per query (a leaf's survivors sit in the
// 2-3 lanes next to it: 10 of 32 lanes active), plus ~78 rounds for the three home leaves.  This version
//   * scores the home leaves directly (every lane wants nearly every candidate there: no mask, no drain);
//   * keeps, of each listed leaf, only the per-lane 32-bit SURVIVOR MASK of a 4-instruction-per-candidate prefilter
//     (packed FFMA2 on pair-interleaved, leaf-centre-relative blocks that build.cu precomputes, sign bits collected by a
//     funnel shift) in a per-lane queue in shared memory, and scores the queued survivors in pooled rounds — lanes
//     balance across leaves, so the expensive ordered insert runs ~20 times per tile instead of ~65;
//   * refines the candidate list per query ONCE (one lane per listed leaf against the 32 queries) instead of testing
//     a box per leaf and lane inside the loop, which also lets the staging copy only wanted leaves;
//   * stages a leaf's 528-byte block with ONE cp.async.bulk into a 2-stage mbarrier ring (the north star's "TMA bulk
//     copy where the cell run is contiguous"; CCG_SEL_STAGE=ldg selects an LDG/STS ring for the A/B profile);
//   * re-uses the queue as the log of pass 2 (drains clear the bits that fell outside the bound).
constexpr int kSelStages = 2;
constexpr int kQCap = 20;              // queue entries per lane: listed leaves use <= kQCap - 3, pass 2 adds the 3 home leaves
constexpr int kSlotHome = 253;         // list[253] / [254] / [255] = own leaf / previous / next along the curve
constexpr int kListMax = 252;          // listed leaves a tile may keep after the per-query refinement (u8 slots)

// The rare second try of the candidate list (wide or split tiles) stays out of line: k_tile_sel is short of instruction-fetch
// bandwidth, and this copy of the list walk is executed by a few tiles in a thousand.
static __device__ __noinline__ int collect_leaves_split(const TreeView &T, int tile, const float4 &tlo, const float4 &thi, float rmax,
                                                        const float4 &q, float bound, WarpScratch2 &w, int (*queue)[kQueueCap]) {
    make_sub_boxes(q, bound, bound >= 0.f, w);
    return collect_leaves<true>(T, tile, tlo, thi, rmax, w, queue);
}

struct alignas(16) SelScratch {
    WarpScratch2 w;                        // w.ws.leaf: the 32 queries; w.list (+ the home leaves at [253..255]); w.sub
    union {
        float4 blk[kSelStages][kRelStride];    // staged prefilter blocks
        int queue[2][kQueueCap];               // breadth-first queues of collect_leaves (no copy is in flight while it runs)
    };
    struct { unsigned qmask[kQCap * 32]; unsigned char qslot[kQCap * 32]; } q;   // entry e of lane l at [e * 32 + l]
    float sbound[32];
    unsigned long long bar[kSelStages];
};
static_assert(sizeof(int) * 2 * kQueueCap <= sizeof(float4) * kSelStages * kRelStride, "the BFS queues must fit the staging ring");

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

// Prefilter of one staged block against the lane's query: bit (31 - j) of the result is set when candidate j MAY lie within
// `bound` of q.  With C the leaf's box centre, P = p - C (precomputed, W = -|P|^2) and Q = q - C:
//     d2 = |P|^2 - 2 P.Q + |Q|^2,   s_j = thr + W_j + 2 Q.P_j,   thr = bound - |Q|^2 + margin,   pass <=> s_j >= 0.
// Rounding: P, Q carry 2^-24 relative error per coordinate, W and |Q|^2 three fused roundings each, s_j one add and three
// fma; together below 2^-20 (|P| + |Q|)^2.  A true neighbour (contract d2 <= bound) has |P| <= |Q| + sqrt(bound)(1 + 2^-22),
// so (|P| + |Q|)^2 <= 8 |Q|^2 + 2 bound + ..., i.e. an error below 2^-17 |Q|^2 + 2^-19 bound.  margin = 2^-15 (|Q|^2 + bound),
// rounded up, never rejects one.  Survivors are always re-scored with the contract arithmetic before they count.
// Two candidates per instruction: packed fma.rn.f32x2 (FFMA2, sm_100) on the pair-interleaved block layout; the sign
// bits are shifted into the mask with one funnel shift each (reject bits; the caller gets the complement).
__device__ __noinline__ unsigned prefilter_blk(const float4 *blk, const float4 &q, float bound) {
    const float4 c = blk[0];
    const float qx = q.x - c.x, qy = q.y - c.y, qz = q.z - c.z;
    const float qq = fmaf(qx, qx, fmaf(qy, qy, qz * qz));
    const float thr = __fadd_ru(__fsub_ru(bound, qq), __fmul_ru(3.0517578125e-5f, __fadd_ru(qq, bound)));
    const f32x2 AX = pack2(2.f * qx, 2.f * qx), AY = pack2(2.f * qy, 2.f * qy), AZ = pack2(2.f * qz, 2.f * qz), TH = pack2(thr, thr);
    unsigned m = 0;
#ifndef CCG_PF_UNROLL
#define CCG_PF_UNROLL 4        // 16 pairs in 4 rounds: the fully unrolled loop was 4.5 % slower (instruction fetch; profiles/r02_sel_v16_codesize_variants.log)
#endif
    constexpr int kPfUnroll = CCG_PF_UNROLL;
#pragma unroll kPfUnroll
    for (int p = 0; p < kLeaf / 2; ++p) {
        const float4 a = blk[1 + 2 * p], b = blk[2 + 2 * p];      // (X0 X1 Y0 Y1), (Z0 Z1 W0 W1)
        f32x2 s = add2(TH, pack2(b.z, b.w));
        s = fma2(AZ, pack2(b.x, b.y), s);
        s = fma2(AY, pack2(a.z, a.w), s);
        s = fma2(AX, pack2(a.x, a.y), s);
        m = __funnelshift_l((unsigned)s, m, 1);
        m = __funnelshift_l((unsigned)(s >> 32), m, 1);
    }
    return bound >= 0.f ? ~m : 0u;
}

// sorted list of the K smallest d2 VALUES seen (front slots pinned to -1 when k < K); an insert is two independent FMNMX per slot: new[i] = max(old[i-1], min(old[i], c))
template <int K>
struct ValList {
    float key[K];
    float bound, rb;          // bound = min(rb, key[K-1]); -1 for lanes that take no part (pads, isolated points)
    __device__ __forceinline__ void init(int k, float rb_, bool active) {
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = (i < K - k) ? -1.f : INFINITY;
        rb = rb_;
        bound = active ? rb_ : -1.f;
    }
    __device__ __forceinline__ void insert(float c) {      // a no-op for c >= key[K-1]
#pragma unroll
        for (int i = K - 1; i > 0; --i) key[i] = fmaxf(key[i - 1], fminf(key[i], c));
        key[0] = fminf(key[0], c);
    }
    __device__ __forceinline__ void tighten() { bound = bound < 0.f ? -1.f : fminf(rb, key[K - 1]); }
    // Own leaf: the 32 exact distances of the lane are sorted in registers by a
    // bitonic network (240 compare-exchanges = 480 FMNMX; the stages that only feed outputs beyond K are dead code) and the K
    // smallest become the list — against 32 ordered inserts of 2K FMNMX each (query 29.1 -> 28.2 ms on the 30 M scan).
    // Extending the idea to the two curve neighbours (sort + bitonic merge with the list, ~0.95 k instead of 1.7 k
    // instructions per leaf) cut the instruction count by another 5 % and LOST time (28.2 -> 28.7 ms): the straight-line
    // networks are fetched once per tile, stall no_instruction rose 3.4 -> 4.6 (profiles/r02_sel_v17_merge_all_*).
    __device__ __forceinline__ void seed_sorted(const float4 &q, const float4 *pts, int k) {
        float d[kLeaf];
#pragma unroll
        for (int j = 0; j < kLeaf; ++j) {
            const float x = d2_point(q, pts[j]);
            d[j] = x <= rb ? x : INFINITY;                   // NaN (pad lanes) and out-of-radius candidates sort last
        }
#pragma unroll
        for (int k2 = 2; k2 <= kLeaf; k2 <<= 1) {
#pragma unroll
            for (int j = k2 >> 1; j > 0; j >>= 1) {
#pragma unroll
                for (int i = 0; i < kLeaf; ++i) {
                    const int l = i ^ j;
                    if (l > i) {
                        const float a = d[i], b = d[l];
                        const bool up = (i & k2) == 0;
                        d[i] = up ? fminf(a, b) : fmaxf(a, b);
                        d[l] = up ? fmaxf(a, b) : fminf(a, b);
                    }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = i < kLeaf ? d[i < kLeaf ? i : 0] : INFINITY;
        // k < K: the K - k front slots are pinned to -1 and the sorted values start behind them — a right shift by the
        // warp-uniform amount K - k, done as log2(K) conditional shifts by constant distances (register indices stay static)
        const int shift = K - k;
#pragma unroll
        for (int b = 1; b < K; b <<= 1) {
            if (shift & b) {
#pragma unroll
                for (int i = K - 1; i >= 0; --i) key[i] = i >= b ? key[i - b < 0 ? 0 : i - b] : -1.f;
            }
        }
        tighten();
    }
    // every staged point scored exactly and inserted in turn (own leaf with pinned slots)
    __device__ __forceinline__ void score_direct(const float4 &q, const float4 *pts) {
#pragma unroll 1
        for (int j = 0; j < kLeaf; ++j) {
            const float d = d2_point(q, pts[j]);
            const float c = d <= rb ? d : INFINITY;          // NaN (pad lanes) and out-of-radius candidates: no-op insert
            if (__any_sync(kFull, c < key[K - 1])) insert(c);
        }
        tighten();
    }
};

// ---- staging ring: block of `leaf` -> S.blk[seq & 1] --------------------------------------------------------
// TMA flavour: lane 0 arms the stage's mbarrier with the byte count and issues one 528-byte cp.async.bulk; the warp
// waits on the barrier's phase parity when it gets there.  LDG flavour: every lane fetches its float4 of the block into a
// register when the copy is "issued" and stores it to shared memory when it is "awaited".
template <bool TMA>
struct Stager {
    float4 r0, r1;            // LDG flavour: lane's float4 (index lane + 1) and, lane 0, the header
    __device__ __forceinline__ void issue(const TreeView &T, SelScratch &S, int leaf, unsigned seq, int lane) {
        const float4 *src = T.rel + (size_t)leaf * kRelStride;
        if (TMA) {
            if (lane == 0) {
                unsigned long long *bar = &S.bar[seq & (kSelStages - 1)];
                mbar_expect_tx(bar, kRelStride * 16);
                bulk_g2s(S.blk[seq & (kSelStages - 1)], src, kRelStride * 16, bar);
            }
        } else {
            r1 = src[1 + lane];
            if (lane == 0) r0 = src[0];
        }
    }
    __device__ __forceinline__ const float4 *await(SelScratch &S, unsigned seq, int lane) {
        float4 *dst = S.blk[seq & (kSelStages - 1)];
        if (TMA) {
            mbar_wait(&S.bar[seq & (kSelStages - 1)], (seq / kSelStages) & 1u);
        } else {
            dst[1 + lane] = r1;
            if (lane == 0) dst[0] = r0;
            __syncwarp();
        }
        return dst;
    }
};

// keep the listed leaves that at least one query of the tile wants (point-box distance within that query's bound): one
// lane per listed leaf against the 32 staged queries.  Needs S.w.ws.leaf = queries, S.sbound = bounds.
__device__ __forceinline__ int refine_list(const TreeView &T, SelScratch &S, int nlist, int lane) {
    int out = 0;
    for (int base = 0; base < nlist; base += 32) {
        const int i = base + lane;
        int leaf = -1;
        bool want = false;
        if (i < nlist) {
            leaf = S.w.list[i];
            const float4 lo = T.lo[0][leaf], hi = T.hi[0][leaf];
#ifndef CCG_REFINE_UNROLL
#define CCG_REFINE_UNROLL 4
#endif
            constexpr int kRefUnroll = CCG_REFINE_UNROLL;
#pragma unroll kRefUnroll
            for (int j = 0; j < kLeaf; ++j) want |= d2_point_box(S.w.ws.leaf[j], lo, hi) <= S.sbound[j];
        }
        const unsigned m = __ballot_sync(kFull, want);
        __syncwarp();
        if (want) S.w.list[out + __popc(m & ((1u << lane) - 1u))] = leaf;
        out += __popc(m);
        __syncwarp();
    }
    return out;
}

// generator over the set bits of this lane's queue entries [ge, cnt): one candidate per call, its point fetched from the
// sorted array (L1/L2) — the caller issues the fetch of candidate r+1 before it scores candidate r
struct QueueWalk {
    int ge, cnt;
    unsigned gcur;
    __device__ __forceinline__ void start(const SelScratch &S, int first, int cnt_, int lane) {
        ge = first; cnt = cnt_;
        gcur = ge < cnt ? S.q.qmask[ge * 32 + lane] : 0u;
    }
    // every queued mask is non-zero (entries are appended only with survivors, and a drain is followed by a squeeze), so
    // one step to the next entry is enough when the current mask runs out
    __device__ __forceinline__ void next(const TreeView &T, const SelScratch &S, int lane, float4 &c, int &e, unsigned &bit) {
        if (gcur == 0u && ge + 1 < cnt) {
            ++ge;
            gcur = S.q.qmask[ge * 32 + lane];
        }
        e = -1;
        if (gcur) {
            const int j = __clz(gcur);
            bit = 0x80000000u >> j;
            gcur ^= bit;
            e = ge;
            c = T.pts[(unsigned)(S.w.list[S.q.qslot[ge * 32 + lane]] * kLeaf + j)];
        }
    }
};

// score the queued survivors of entries [first, cnt): `rounds` = the largest number of pending bits any lane holds.
// Bits whose candidate is outside the lane's bound are cleared, so that pass 2 does not meet them again.
template <int K>
__device__ __forceinline__ void drain_queue(const TreeView &T, SelScratch &S, int lane, const float4 &q, ValList<K> &v,
                                            int first, int cnt, int rounds) {
    QueueWalk w;
    w.start(S, first, cnt, lane);
    float4 fc = q; int fe; unsigned fbit = 0;
    w.next(T, S, lane, fc, fe, fbit);
    for (int r = 0; r < rounds; ++r) {
        const float4 c = fc; const int e = fe; const unsigned bit = fbit;
        w.next(T, S, lane, fc, fe, fbit);
        if (e >= 0) {
            const float d = d2_point(q, c);
            if (d <= v.bound) {
                if (d < v.key[K - 1]) { v.insert(d); v.bound = fminf(v.rb, v.key[K - 1]); }
            } else {
                S.q.qmask[e * 32 + lane] &= ~bit;
            }
        }
    }
}

// drop the entries a drain emptied; returns the lane's new entry count
__device__ __forceinline__ int squeeze_queue(SelScratch &S, int cnt, int lane) {
    const int maxc = __reduce_max_sync(kFull, cnt);
    int w = 0;
#pragma unroll 1
    for (int e = 0; e < maxc; ++e) {
        if (e < cnt) {
            const unsigned m = S.q.qmask[e * 32 + lane];
            if (m) {
                if (w != e) { S.q.qmask[w * 32 + lane] = m; S.q.qslot[w * 32 + lane] = S.q.qslot[e * 32 + lane]; }
                ++w;
            }
        }
    }
    return w;
}

// Pass 1 for one tile.  Returns leaves prefiltered (>= 0), or < 0 when the whole tile goes to the exact point kernel
// (-3 queue overflow = that many ties, -4 unbounded search, -5 candidate list too long).
// On return: v = value lists and bounds, cnt = queue entries of this lane, lone = lane handed to the point kernel.
template <int K, bool TMA>
__device__ __forceinline__ int sel_pass1(const TreeView &T, const TileParams &P, int tile, const float4 &tlo, const float4 &thi,
                                         const float4 &q, ValList<K> &v, SelScratch &S, Stager<TMA> &stg, unsigned &seq, int &cnt,
                                         int &nlist, bool &lone, unsigned long long *st5) {
    const int lane = threadIdx.x & 31;
    lone = false; cnt = 0; nlist = 0;
    // ---- own leaf, scored directly (every lane wants nearly every candidate of its own leaf: no mask, no drain) -------
    S.w.ws.leaf[lane] = q;
    if (lane < 3) S.w.list[kSlotHome + lane] = tile + (lane == 0 ? 0 : (lane == 1 ? -1 : 1));
    __syncwarp();
    if (P.seed_net) v.seed_sorted(q, S.w.ws.leaf, P.k);
#ifdef CCG_SCORE_TWO_SITES                 // A/B build only: the own leaf's direct scoring as a second inlined copy
    else v.score_direct(q, S.w.ws.leaf);
    const int h0 = 1;
#else
    const int h0 = P.seed_net ? 1 : 0;
#endif
    // ---- the own leaf without the network, and the two neighbours along the curve, scored directly through ONE call site.
    //      (Queueing the neighbours like the listed leaves was measured: their survivors fill half the lanes, the pooled drain
    //      needs as many rounds as the direct loop has candidates, and pass 2 then walks 44 instead of 31 rounds per tile:
    //      28.1 -> 32.6 ms, profiles/r02_sel_v15_*.) ---------------------------------------------------------------------
    int first = 0, pend = 0;
#pragma unroll 1
    for (int h = h0; h < 3; ++h) {
        const int leaf = tile + (h == 0 ? 0 : (h == 1 ? -1 : 1));
        if (leaf < 0 || leaf >= T.count[0]) continue;
        if (h) {
            const float4 p = T.pts[(size_t)leaf * kLeaf + lane];
            __syncwarp();
            S.w.ws.leaf[lane] = p;
            __syncwarp();
        }
        v.score_direct(q, S.w.ws.leaf);
    }
    // ---- isolated lanes -> point kernel (a lane whose bound is far above the warp's median would make the tile's
    //      candidate list hundreds of leaves long) --------------------------------------------------------------
    {
        float b = v.bound >= 0.f ? v.bound : INFINITY;
        // (rolled on purpose, like the two queue loops below: the kernel is short of instruction-fetch bandwidth — 26.40 -> 25.78 ms)
#pragma unroll 1
        for (int kk = 2; kk <= 32; kk <<= 1) {
#pragma unroll 1
            for (int j = kk >> 1; j > 0; j >>= 1) {
                const float o = __shfl_xor_sync(kFull, b, j);
                const bool up = ((lane & kk) == 0) == ((lane & j) == 0);
                b = up ? fminf(b, o) : fmaxf(b, o);
            }
        }
        const int nact = __popc(__ballot_sync(kFull, v.bound >= 0.f));
        const float med = __shfl_sync(kFull, b, nact >> 1);
        lone = v.bound >= 0.f && v.bound > 16.f * med;
        if (lone) v.bound = -1.f;
    }
    const float rmax = warp_max(v.bound);
    if (!(rmax < INFINITY)) return -4;
    __syncwarp();
    S.w.ws.leaf[lane] = q;                 // the refinement reads the queries and their bounds from shared memory
    S.sbound[lane] = v.bound;
    nlist = collect_leaves<false>(T, tile, tlo, thi, rmax, S.w, S.queue);
    if (nlist < 0) {                       // wide or split tile: describe it by 8 sub-boxes and try again
        nlist = collect_leaves_split(T, tile, tlo, thi, rmax, q, v.bound, S.w, S.queue);
        if (nlist < 0) return -5;
    }
    if (TMA) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the BFS queues were written over the staging ring
    __syncwarp();
    const int listed = nlist;
    nlist = refine_list(T, S, nlist, lane);
    if (st5 && lane == 0) { atomicAdd(st5 + 2, (unsigned long long)listed); atomicAdd(st5 + 5, (unsigned long long)nlist); }
    if (nlist > kListMax) return -5;
    // ---- listed leaves: stage, prefilter, queue the survivor masks; pooled drains ----------------------------------
    const int drain_at = P.drain_at;
    if (nlist > 0) stg.issue(T, S, S.w.list[0], seq, lane);
    for (int i = 0; i < nlist; ++i) {
        const float4 *blk = stg.await(S, seq, lane);
        if (i + 1 < nlist) {
            __syncwarp();                                // every lane is done with the stage the next copy overwrites
            stg.issue(T, S, S.w.list[i + 1], seq + 1, lane);
        }
        ++seq;
        const unsigned pm = prefilter_blk(blk, q, v.bound);
        if (pm) {
            S.q.qmask[cnt * 32 + lane] = pm;
            S.q.qslot[cnt * 32 + lane] = (unsigned char)i;
            ++cnt;
            pend += __popc(pm);
        }
        // ONE drain site (code size): pending survivors reached the trigger, the last leaf was filtered, or the queue is
        // full — then score what is pending and drop the entries that emptied.
        const int rounds = __reduce_max_sync(kFull, pend);
        const bool full = __reduce_max_sync(kFull, cnt) >= kQCap - 3;
        if (rounds >= drain_at || ((i + 1 == nlist || full) && rounds > 0)) {
            drain_queue<K>(T, S, lane, q, v, first, cnt, rounds);
            cnt = squeeze_queue(S, cnt, lane);
            first = cnt; pend = 0;
            if (st5 && lane == 0) { atomicAdd(st5 + 6, 1ull); atomicAdd(st5 + 10, (unsigned long long)rounds); }
        }
        if (full && i + 1 < nlist && __reduce_max_sync(kFull, cnt) >= kQCap - 3) {
            stg.await(S, seq, lane);                     // still full after the drain: that many ties.  No copy may be in
            ++seq;                                       // flight when the block retires
            return -3;
        }
    }
    return nlist;
}

template <int K>
__device__ __forceinline__ void sor_finish(const float *key, int k, int mean_k, float *mean_d, size_t row) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < K; ++i)
        if (i > K - k && key[i] != INFINITY) s += (double)__fsqrt_rn(key[i]);
    mean_d[row] = (float)(s / (double)mean_k);
}

// STATS: the counters behind CCG_VERBOSE_STATS live in an instantiation of their own, so that the measured kernel carries none
// of their code (the kernel is short of instruction-fetch bandwidth: every removed instruction shows in the time).
template <int K, int MODE, bool TMA, int MINB, bool STATS>
__global__ void __launch_bounds__(64, MINB) k_tile_sel(const __grid_constant__ TreeView T, const __grid_constant__ TileParams P) {
    __shared__ SelScratch s_ws[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = P.tile_begin + blockIdx.x * 2 + warp;
    if (slot >= P.tile_end) return;
    int tile = slot;
    if (P.tile_order) {
        tile = P.tile_order[slot];
    } else if (P.shard_world > 1) {
        const int c = slot / P.shard_chunk;
        tile = (c * P.shard_world + P.shard_rank) * P.shard_chunk + (slot - c * P.shard_chunk);
        if (tile >= T.count[0]) return;
    }
    SelScratch &S = s_ws[warp];
    unsigned long long *const st5 = STATS ? P.stat_scanned : nullptr;
    if (TMA) {
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < kSelStages; ++s) mbar_init(&S.bar[s], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    const int qidx = __float_as_int(q.w);
    const size_t row = P.sorted_rows ? (size_t)((long long)pos - P.row_begin) : (size_t)qidx;
    const int k = P.k;
    const float4 tlo = T.lo[0][tile], thi = T.hi[0][tile];

    ValList<K> v;
    v.init(k, P.rb, active);
    Stager<TMA> stg;
    unsigned seq = 0;
    int cnt = 0, nlist = 0;
    bool lone = false;
    const int scanned = sel_pass1<K, TMA>(T, P, tile, tlo, thi, q, v, S, stg, seq, cnt, nlist, lone, st5);
    if (scanned >= 0) {                             // isolated lanes of a dense tile -> point kernel
        const unsigned lm = __ballot_sync(kFull, lone);
        if (lm) {
            int base = 0;
            if (lane == 0) base = atomicAdd(P.defer_count, __popc(lm));
            base = __shfl_sync(kFull, base, 0);
            if (lone) {
                const int ds = base + __popc(lm & ((1u << lane) - 1u));
                P.defer_list[ds] = qidx;
                P.defer_row[ds] = (long long)row;
            }
        }
    }
    if (scanned < 0) {          // unbounded / very wide / heavily tied search: the tile's queries go to the exact point kernel
        int base = 0;
        const unsigned am = __ballot_sync(kFull, active);
        if (lane == 0) {
            base = atomicAdd(P.defer_count, __popc(am));
            if (st5) { atomicAdd(st5 + 3, 1ull); atomicAdd(st5 + 4 - scanned, 1ull); }
        }
        base = __shfl_sync(kFull, base, 0);
        if (active) {
            const int ds = base + __popc(am & ((1u << lane) - 1u));
            P.defer_list[ds] = qidx;
            P.defer_row[ds] = (long long)row;
        }
        return;
    }
    if (st5 && lane == 0) atomicAdd(st5, (unsigned long long)scanned);
    if (MODE == MODE_SOR) {
        if (active && !lone) sor_finish<K>(v.key, k, P.mean_k, P.mean_d, row);
        return;
    }
    // ---- pass 2: the neighbour set {d2 <= t} lies in the home leaves and the queue ---------------------------------
    float t = (active && !lone) ? fminf(fminf(P.rb, v.key[K - 1]), FLT_MAX) : -1.f;   // finite: +inf pads never qualify
    // MODE_KNN (ordered lists): the value list is already the sorted d2 row; a neighbour's slot is the number of list
    // values below its d2, which is unique as long as no two list values are equal.  Equal values need the
    // (d2, index) tie-break: those queries go to the exact point kernel, like tie overflow at the k-th value.
    bool dup = false;
    if (MODE == MODE_KNN) {
#pragma unroll
        for (int i = 1; i < K; ++i)
            if (i > K - k) dup = dup || (v.key[i] == v.key[i - 1] && v.key[i] != INFINITY);
        if (dup) t = -1.f;
    }
    // the three home leaves join the queue with their prefilter masks against t (3 entries are reserved for them)
    int tot = 0;
    {
        const int maxc = __reduce_max_sync(kFull, cnt);
#pragma unroll 1
        for (int e = 0; e < maxc; ++e)
            if (e < cnt) tot += __popc(S.q.qmask[e * 32 + lane]);
    }
#pragma unroll 1
    for (int h = 0; h < 3; ++h) {
        const int leaf = tile + (h == 0 ? 0 : (h == 1 ? -1 : 1));
        if (leaf < 0 || leaf >= T.count[0]) continue;
        __syncwarp();
        stg.issue(T, S, leaf, seq, lane);
        const float4 *blk = stg.await(S, seq, lane);
        ++seq;
        const unsigned pm = prefilter_blk(blk, q, t);
        if (pm) {
            S.q.qmask[cnt * 32 + lane] = pm;
            S.q.qslot[cnt * 32 + lane] = (unsigned char)(kSlotHome + h);
            ++cnt;
            tot += __popc(pm);
        }
    }
    int count = 0;
    float m[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) m[i] = 0.f;
    {
        const int rounds = __reduce_max_sync(kFull, tot);
        if (st5 && lane == 0) atomicAdd(st5 + 1, (unsigned long long)rounds);
        QueueWalk w;
        w.start(S, 0, cnt, lane);
        float4 fc = q; int fe; unsigned fbit = 0;
        w.next(T, S, lane, fc, fe, fbit);
        for (int r = 0; r < rounds; ++r) {
            const float4 c = fc; const int e = fe;
            w.next(T, S, lane, fc, fe, fbit);
            if (e >= 0) {
                const float dd = d2_point(q, c);
                if (dd <= t) {
                    ++count;
                    if (MODE == MODE_KNN) {
                        int below = 0;
#pragma unroll
                        for (int i2 = 0; i2 < K; ++i2) below += v.key[i2] < dd ? 1 : 0;
                        if (P.idx) P.idx[row * k + (below - (K - k))] = __float_as_int(c.w);
                    } else {
                        const float x = c.x - q.x, y = c.y - q.y, z = c.z - q.z;
                        m[0] += x; m[1] += y; m[2] += z;
                        m[3] += x * x; m[4] += x * y; m[5] += x * z; m[6] += y * y; m[7] += y * z; m[8] += z * z;
                    }
                }
            }
        }
    }
    if (!active || lone) return;
    if (count > k || dup) {                                // ties at the k-th distance overflow the list: exact kernel decides
        const int ds = atomicAdd(P.defer_count, 1);
        P.defer_list[ds] = qidx;
        P.defer_row[ds] = (long long)row;
        return;
    }
    if (MODE == MODE_KNN) {
#pragma unroll
        for (int i = 0; i < K; ++i) {
            if (i >= K - k) {
                const int j = i - (K - k);
                if (P.d2) P.d2[row * k + j] = j < count ? v.key[i] : INFINITY;
                if (j >= count && P.idx) P.idx[row * k + j] = -1;
            }
        }
        if (P.count) P.count[row] = count;
        return;
    }
    pca_finish(P.pca, q, m, count, row);
    if (P.count) P.count[row] = count;
}

// ---------------------------------------------------------------------------------------------------
// tile kernel without a list (radius count / uncapped radius PCA)
// ---------------------------------------------------------------------------------------------------
struct StreamParams {
    int mode; float rb; int count_cap;
    int tile_begin, tile_end; long long row_begin; int sorted_rows;
    int32_t *count;
    PcaOut pca;
    unsigned long long *stat_scanned;
    const float4 *mom;               // per-box moments (StreamNodes) or NULL
};

template <class V>
__global__ void __launch_bounds__(128) k_tile_stream(const __grid_constant__ TreeView T, const __grid_constant__ StreamParams P) {
    __shared__ WarpScratch s_ws[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = P.tile_begin + blockIdx.x * 4 + warp;
    if (tile >= P.tile_end) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    V v;
    v.init(P.rb, P.count_cap, P.mode == 1, active);
    if constexpr (visits_nodes<V>::value) v.mom = P.mom;
    int scanned = traverse_tile(T, tile, q, v, s_ws[warp]);
    if (P.stat_scanned && lane == 0) atomicAdd(P.stat_scanned, (unsigned long long)scanned);
    const size_t row = P.sorted_rows ? (size_t)((long long)pos - P.row_begin) : (size_t)__float_as_int(q.w);
    if (P.mode == 1) {
        PcaOut o = P.pca;
        if (!active) { o.eig3 = nullptr; o.normal4 = nullptr; o.feat7 = nullptr; }
        pca_finish(o, q, v.m, active ? v.count : 0, row);
    }
    if (P.count && active) P.count[row] = v.count;
}

// radius neighbourhood sum of a per-point 3-vector (f2: difference of normals, featureeval.cpp:694-739)
__global__ void __launch_bounds__(128) k_tile_attrsum(const __grid_constant__ TreeView T, float rb, const char *attr, size_t stride,
                                                      float *sum3, int32_t *count) {
    __shared__ WarpScratch s_ws[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    AttrSum v;
    v.init(rb, active, attr, stride);
    traverse_tile(T, tile, q, v, s_ws[warp]);
    if (!active) return;
    const size_t row = (size_t)__float_as_int(q.w);
    sum3[3 * row] = v.s[0]; sum3[3 * row + 1] = v.s[1]; sum3[3 * row + 2] = v.s[2];
    count[row] = v.count;
}

// f4: pcl::PrincipalCurvaturesEstimation::computePointPrincipalCurvatures [EXT] (called at markov.cpp:389-402,
// featureeval.cpp:947-956) in one pass: neighbour normals projected into the query's tangent plane, M = I - n n^T, their
// scatter matrix C = sum (p_j - mean p)(p_j - mean p)^T = M (Q - S S^T / N) M with S, Q the moments of (n_j - n) (a common
// shift cancels in C), then pc = eigenvector of the largest eigenvalue, pc1 = l_max / N, pc2 = l_mid / N.
__global__ void __launch_bounds__(128) k_tile_principal_curv(const __grid_constant__ TreeView T, float rb, const char *attr, size_t stride,
                                                             float *pc5) {
    __shared__ WarpScratch s_ws[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    AttrMoments v;
    v.init(rb, active, attr, stride, q);
    traverse_tile(T, tile, q, v, s_ws[warp]);
    if (!active) return;
    float *o = pc5 + 5 * (size_t)__float_as_int(q.w);
    if (v.count == 0) {                  // radiusSearch found nothing (cannot happen for a finite query: it finds itself)
        o[0] = o[1] = o[2] = o[3] = o[4] = NAN;
        return;
    }
    const float invn = 1.f / (float)v.count;
    // B = Q - S S^T / N  (symmetric, {xx xy xz yy yz zz})
    float b[6];
    b[0] = v.m[0] - v.s[0] * v.s[0] * invn; b[1] = v.m[1] - v.s[0] * v.s[1] * invn; b[2] = v.m[2] - v.s[0] * v.s[2] * invn;
    b[3] = v.m[3] - v.s[1] * v.s[1] * invn; b[4] = v.m[4] - v.s[1] * v.s[2] * invn; b[5] = v.m[5] - v.s[2] * v.s[2] * invn;
    const float nx = v.ref[0], ny = v.ref[1], nz = v.ref[2];
    // C = M B M, M = I - n n^T:  C = B - n (Bn)^T - (Bn) n^T + (n^T B n) n n^T
    const float bn0 = b[0] * nx + b[1] * ny + b[2] * nz, bn1 = b[1] * nx + b[3] * ny + b[4] * nz, bn2 = b[2] * nx + b[4] * ny + b[5] * nz;
    const float nbn = nx * bn0 + ny * bn1 + nz * bn2;
    float c[6];
    c[0] = -(b[0] - 2.f * nx * bn0 + nbn * nx * nx);
    c[1] = -(b[1] - nx * bn1 - ny * bn0 + nbn * nx * ny);
    c[2] = -(b[2] - nx * bn2 - nz * bn0 + nbn * nx * nz);
    c[3] = -(b[3] - 2.f * ny * bn1 + nbn * ny * ny);
    c[4] = -(b[4] - ny * bn2 - nz * bn1 + nbn * ny * nz);
    c[5] = -(b[5] - 2.f * nz * bn2 + nbn * nz * nz);
    // the solver returns the eigenvector of the SMALLEST eigenvalue: hand it -C
    const Eig3 e = eig3_sym(c);
    o[0] = e.n[0]; o[1] = e.n[1]; o[2] = e.n[2];
    o[3] = -e.w[2] * invn;
    o[4] = -e.w[1] * invn;
}

int launch_tile_principal_curv(ccg_ctx *ctx, float rb, const void *attr_dev, size_t stride, float *pc5) {
    if (ctx->n == 0) return CCG_OK;
    k_tile_principal_curv<<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb, (const char *)attr_dev, stride, pc5);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

// f4: pcl::FPFHEstimation with setRadiusSearch(r) [EXT] (featureeval.cpp:925-932) in two traversals
__global__ void __launch_bounds__(128) k_tile_spfh(const __grid_constant__ TreeView T, float rb, const char *attr, size_t stride, float *spfh33) {
    __shared__ WarpScratch s_ws[4];
    __shared__ unsigned s_hist[4][33 * 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    Spfh v;
    v.init(rb, active, attr, stride, q, &s_hist[warp][lane]);
    traverse_tile(T, tile, q, v, s_ws[warp]);
    if (!active) return;
    float *o = spfh33 + 33 * (size_t)__float_as_int(q.w);
    const float incr = 100.0f / (float)(v.count - 1);            // computeSPFHSignature: 100 / (indices.size() - 1)
#pragma unroll
    for (int b = 0; b < 33; ++b) {
        const unsigned c = v.hist[b * 32];
        o[b] = c ? (float)c * incr : 0.f;
    }
}

__global__ void __launch_bounds__(128) k_tile_fpfh(const __grid_constant__ TreeView T, float rb, const float *spfh33, float *fpfh33) {
    __shared__ WarpScratch s_ws[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    FpfhSum v;
    v.init(rb, active, spfh33);
    traverse_tile(T, tile, q, v, s_ws[warp]);
    if (!active) return;
    float *o = fpfh33 + 33 * (size_t)__float_as_int(q.w);
#pragma unroll
    for (int g = 0; g < 3; ++g) {
        float sum = 0.f;
#pragma unroll
        for (int b = 0; b < 11; ++b) sum += v.acc[g * 11 + b];
        const float sc = sum != 0.f ? 100.0f / sum : 0.f;
#pragma unroll
        for (int b = 0; b < 11; ++b) o[g * 11 + b] = v.acc[g * 11 + b] * sc;
    }
}

int launch_tile_fpfh(ccg_ctx *ctx, float rb, const void *attr_dev, size_t stride, float *spfh33, float *fpfh33) {
    if (ctx->n == 0) return CCG_OK;
    k_tile_spfh<<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb, (const char *)attr_dev, stride, spfh33);
    k_tile_fpfh<<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb, spfh33, fpfh33);
    ctx->stats.launches += 2;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

__global__ void __launch_bounds__(128) k_tile_attrsum2(const __grid_constant__ TreeView T, float rb1, float rb2, const char *attr, size_t stride,
                                                       float *sum1, int32_t *count1, float *sum2, int32_t *count2) {
    __shared__ WarpScratch s_ws[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    AttrSum2 v;
    v.init(rb1, rb2, active, attr, stride);
    traverse_tile(T, tile, q, v, s_ws[warp]);
    if (!active) return;
    const size_t row = (size_t)__float_as_int(q.w);
    sum1[3 * row] = v.s1[0]; sum1[3 * row + 1] = v.s1[1]; sum1[3 * row + 2] = v.s1[2]; count1[row] = v.c1;
    sum2[3 * row] = v.s2[0]; sum2[3 * row + 1] = v.s2[1]; sum2[3 * row + 2] = v.s2[2]; count2[row] = v.c2;
}

int launch_tile_attrsum2(ccg_ctx *ctx, float rb1, float rb2, const void *attr_dev, size_t stride, float *sum1, int32_t *count1,
                         float *sum2, int32_t *count2) {
    if (ctx->n == 0) return CCG_OK;
    k_tile_attrsum2<<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb1, rb2, (const char *)attr_dev, stride, sum1, count1, sum2, count2);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

int launch_tile_attrsum(ccg_ctx *ctx, float rb, const void *attr_dev, size_t stride, float *sum3, int32_t *count) {
    if (ctx->n == 0) return CCG_OK;
    k_tile_attrsum<<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb, (const char *)attr_dev, stride, sum3, count);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

// ---------------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------------
template <int K>
static void launch_topk(const TreeView &T, const TileParams &P, int blocks, cudaStream_t st) {
    k_tile_topk<K><<<blocks, 128, 0, st>>>(T, P);
}

template <int K, bool TMA, int MINB, bool STATS>
static void launch_sel4(const TreeView &T, const TileParams &P, cudaStream_t st) {
    const int ntiles = P.tile_end - P.tile_begin;
    if (P.mode == MODE_PCA) k_tile_sel<K, MODE_PCA, TMA, MINB, STATS><<<(ntiles + 1) / 2, 64, 0, st>>>(T, P);
    else if (P.mode == MODE_KNN) k_tile_sel<K, MODE_KNN, TMA, MINB, STATS><<<(ntiles + 1) / 2, 64, 0, st>>>(T, P);
    else k_tile_sel<K, MODE_SOR, TMA, MINB, STATS><<<(ntiles + 1) / 2, 64, 0, st>>>(T, P);
}
template <int K, bool TMA, int MINB>
static void launch_sel3(const TreeView &T, const TileParams &P, cudaStream_t st) {
    if (P.stat_scanned) {                   // CCG_VERBOSE_STATS runs: the counting build, TMA staging, default occupancy
        if constexpr (TMA) { launch_sel4<K, true, MINB, true>(T, P, st); return; }
    }
    launch_sel4<K, TMA, MINB, false>(T, P, st);
}
// Resident blocks per SM the kernel is compiled for (register cap = 65536 / (64 * MINB)).  Shared memory (12.4 KB per
// block) allows 16; measured on the 30 M scan, K = 20: 12 blocks / 80 registers 28.96 ms, 14 / 72 28.77 ms, 16 / 64 (16 bytes
// of spills) 28.10 ms — occupancy beats registers here.  CCG_SEL_MINB=12 selects the 80-register build of K = 20 for A/B runs.
template <int K>
static void launch_sel(const TreeView &T, const TileParams &P, int tma, int minb, cudaStream_t st) {
#ifndef CCG_MINB20
#define CCG_MINB20 16
#endif
    constexpr int kDef = K > 32 ? 6 : (K > 24 ? 10 : (K == 20 ? CCG_MINB20 : 12));
    if constexpr (K == 20) {
        if (minb == 12) { if (tma) launch_sel3<K, true, 12>(T, P, st); else launch_sel3<K, false, 12>(T, P, st); return; }
    }
    if (tma) launch_sel3<K, true, kDef>(T, P, st);
    else launch_sel3<K, false, kDef>(T, P, st);
}

int launch_tile_query(ccg_ctx *ctx, const QueryArgs &a) {
    if (ctx->n == 0) return CCG_OK;
    if (a.k < 1 || a.k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", a.k, CCG_MAX_K);
    if (a.begin % kLeaf) return set_error(ctx, CCG_ERR_INVALID, "range begin must be a multiple of 32");
    size_t end = a.end > ctx->n ? ctx->n : a.end;
    if (a.begin >= end) return CCG_OK;
    TileParams P{};
    P.mode = a.mode; P.k = a.k; P.rb = a.rb;
    P.tile_begin = (int)(a.begin / kLeaf);
    P.tile_end = (int)((end + kLeaf - 1) / kLeaf);
    P.row_begin = (long long)a.begin; P.sorted_rows = a.sorted_rows;
    P.idx = a.idx; P.d2 = a.d2; P.count = a.count;
    P.pca.cov_norm = a.cov_norm; P.pca.min_pts = a.min_pts; P.pca.normalise_hist = a.normalise_hist; P.pca.has_vp = a.has_vp;
    for (int i = 0; i < 3; ++i) P.pca.vp[i] = a.vp[i];
    P.pca.eig3 = a.eig3; P.pca.normal4 = a.normal4; P.pca.feat7 = a.feat7; P.pca.formula_set = a.formula_set;
    P.pca.route_world = 0;
    if (a.route_world > 1) {
        if (a.route_world > kMaxRoute || a.route_rows == 0) return set_error(ctx, CCG_ERR_INVALID, "route: world must be <= %d and rows_per_rank > 0", kMaxRoute);
        P.pca.route_world = a.route_world; P.pca.route_rows = (unsigned)a.route_rows;
        for (int r = 0; r < a.route_world; ++r) {
            P.pca.r_eig[r] = a.r_eig ? a.r_eig[r] : nullptr; P.pca.r_nrm[r] = a.r_nrm ? a.r_nrm[r] : nullptr;
            P.pca.r_feat[r] = a.r_feat ? a.r_feat[r] : nullptr;
        }
    }
    P.mean_k = a.mean_k; P.mean_d = a.mean_d;
    P.tile_order = a.tile_order;
    P.seed_net = ctx->sel_net;
    P.drain_at = ctx->sel_drain_at > 0 ? ctx->sel_drain_at : 1 << 30;
    P.shard_world = a.shard_world; P.shard_rank = a.shard_rank; P.shard_chunk = a.shard_chunk;
    if (a.shard_world > 1) {
        if (a.mode == MODE_KNN && ctx->knn_use_topk) return set_error(ctx, CCG_ERR_INVALID, "interleaved shards need the selection kernel");
        if (a.shard_chunk < 1 || a.shard_rank < 0 || a.shard_rank >= a.shard_world) return set_error(ctx, CCG_ERR_INVALID, "bad shard");
        const int chunks_total = (ctx->nleaf + a.shard_chunk - 1) / a.shard_chunk;
        const int mine = a.shard_rank < chunks_total ? (chunks_total - a.shard_rank + a.shard_world - 1) / a.shard_world : 0;
        if (mine == 0) return CCG_OK;
        P.tile_begin = 0;
        P.tile_end = mine * a.shard_chunk;
    }
    if (a.tile_order && a.mode == MODE_KNN && ctx->knn_use_topk) return set_error(ctx, CCG_ERR_INVALID, "tile_order is only supported by the selection kernel");
    P.stat_scanned = nullptr;
    if (ctx->collect_stats) {
        CCG_CUDA(ctx, cudaMemsetAsync((char *)ctx->scratch.p + 1024, 0, 96, ctx->stream));
        P.stat_scanned = (unsigned long long *)((char *)ctx->scratch.p + 1024);
    }
    const int ntiles = P.tile_end - P.tile_begin;
    const int blocks = (ntiles + 3) / 4;
    TreeView T = ctx->view();
    cudaStream_t st = ctx->stream;
    const int k = a.k;
    if (a.mode == MODE_KNN && ctx->knn_use_topk) {
        if (k <= 8) launch_topk<8>(T, P, blocks, st);
        else if (k <= 16) launch_topk<16>(T, P, blocks, st);
        else if (k <= 20) launch_topk<20>(T, P, blocks, st);
        else if (k <= 24) launch_topk<24>(T, P, blocks, st);
        else if (k <= 32) launch_topk<32>(T, P, blocks, st);
        else if (k <= 48) launch_topk<48>(T, P, blocks, st);
        else launch_topk<64>(T, P, blocks, st);
    } else {
        const bool overlap = ctx->point_overlap && !ctx->collect_stats;
        // two-pass selection; queries whose k-th distance is tied beyond the list go to the exact point kernel.
        // The point kernel is latency-bound (r01: 7.6 % of GPU time for 0.3 % of the queries, serialised behind every tile
        // launch), so it runs on a side stream next to the NEXT tile launch; the defer list is double-buffered for that and
        // join_side() (every entry point calls it before it touches the outputs) brings the streams together again.
        // A long launch is cut into sel_chunks sub-launches so that chunk c's point kernel overlaps chunk c+1's tiles.
        const int tb0 = P.tile_begin, te0 = P.tile_end;
        int nsub = (overlap && !a.async_side && ntiles >= 65536) ? ctx->sel_chunks : 1;
        if (nsub < 1) nsub = 1;
        const int per = ((ntiles + nsub - 1) / nsub + 1) & ~1;        // even: two tiles per block
        for (int tb = tb0; tb < te0; tb += per) {
        P.tile_begin = tb; P.tile_end = tb + per < te0 ? tb + per : te0;
        const size_t rows = (size_t)(P.tile_end - P.tile_begin) * kLeaf;
        const size_t off_list = 256, off_rows = off_list + ((rows * 4 + 255) & ~(size_t)255);
        const size_t off_end = off_rows + ((rows * 8 + 255) & ~(size_t)255);
        const int sel = overlap ? (ctx->defer_sel ^= 1) : 0;
        DevBuf &db = sel ? ctx->defer_b : ctx->defer;
        if (overlap) {
            for (int i = 0; i < 2; ++i)
                if (!ctx->ev_side[i]) CCG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_side[i], cudaEventDisableTiming));
            if (!ctx->ev_tile) CCG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_tile, cudaEventDisableTiming));
            if (!ctx->side_stream) CCG_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking));
            CCG_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_side[sel], 0));      // the launch before last still reads this buffer
        }
        CCG_CUDA(ctx, db.reserve(off_end));
        CCG_CUDA(ctx, cudaMemsetAsync(db.p, 0, 8, st));
        P.defer_count = db.as<int>();
        P.slow_count = db.as<int>() + 1;
        P.defer_list = (int32_t *)((char *)db.p + off_list);
        P.defer_row = (long long *)((char *)db.p + off_rows);
        P.slow_tiles = nullptr;
        if (k <= 8) launch_sel<8>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        else if (k <= 16) launch_sel<16>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        else if (k <= 20) launch_sel<20>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        else if (k <= 24) launch_sel<24>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        else if (k <= 32) launch_sel<32>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        else if (k <= 48) launch_sel<48>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        else launch_sel<64>(T, P, ctx->sel_stage, ctx->sel_minb, st);
        {
            PointParams Q{};
            Q.queries = P.defer_list; Q.nq = 0; Q.nq_dev = P.defer_count; Q.rows = P.defer_row;
            Q.mode = a.mode; Q.k = k; Q.rb = a.rb; Q.list_rows = 0;
            Q.idx = a.idx; Q.d2 = a.d2;
            Q.count = a.count; Q.pca = P.pca; Q.mean_k = a.mean_k; Q.mean_d = a.mean_d;
            cudaStream_t ps = st;
            if (overlap) {
                CCG_CUDA(ctx, cudaEventRecord(ctx->ev_tile, st));
                CCG_CUDA(ctx, cudaStreamWaitEvent(ctx->side_stream, ctx->ev_tile, 0));
                ps = ctx->side_stream;
            }
            k_point_query<<<ctx->sm_count * 8, 128, 0, ps>>>(T, Q);
            if (overlap) {
                CCG_CUDA(ctx, cudaEventRecord(ctx->ev_side[sel], ps));
                ctx->side_pending = 1;
                ctx->last_side = sel;
            }
            ctx->stats.launches += 2;
        }
        }   // sub-launches
        ctx->stats.launches--;
        if (overlap && !a.async_side) CCG_TRY(join_side(ctx));
    }
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    if (ctx->collect_stats) {
        unsigned long long s4[12] = {};
        CCG_CUDA(ctx, cudaMemcpyAsync(s4, P.stat_scanned, 96, cudaMemcpyDeviceToHost, st));
        int nd = 0;
        if (P.defer_count) CCG_CUDA(ctx, cudaMemcpyAsync(&nd, P.defer_count, 4, cudaMemcpyDeviceToHost, st));
        CCG_CUDA(ctx, cudaStreamSynchronize(st));
        ctx->stats.leaves_scanned = (int64_t)s4[0];
        ctx->stats.deferred_queries = nd;
        if (getenv("CCG_VERBOSE_STATS"))
            fprintf(stderr, "[ccg] tiles=%d leaves listed=%llu kept=%llu prefiltered=%llu | drains=%llu drain rounds=%llu pass-2 rounds=%llu | "
                            "slow tiles=%llu (queue full %llu, unbounded %llu, wide %llu) deferred=%d\n", ntiles,
                    s4[2], s4[5], s4[0], s4[6], s4[10], s4[1], s4[3], s4[7], s4[8], s4[9], nd);
    }
    return CCG_OK;
}

// the main stream waits for the point kernels still running on the side stream
int join_side(ccg_ctx *ctx) {
    if (!ctx->side_pending) return CCG_OK;
    for (int i = 0; i < 2; ++i)
        if (ctx->ev_side[i]) CCG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_side[i], 0));
    ctx->side_pending = 0;
    return CCG_OK;
}

int launch_tile_stream(ccg_ctx *ctx, const StreamArgs &a) {
    if (ctx->n == 0) return CCG_OK;
    if (a.begin % kLeaf) return set_error(ctx, CCG_ERR_INVALID, "range begin must be a multiple of 32");
    size_t end = a.end > ctx->n ? ctx->n : a.end;
    if (a.begin >= end) return CCG_OK;
    StreamParams P{};
    P.mode = a.mode; P.rb = a.rb; P.count_cap = a.count_cap;
    P.tile_begin = (int)(a.begin / kLeaf);
    P.tile_end = (int)((end + kLeaf - 1) / kLeaf);
    P.row_begin = (long long)a.begin; P.sorted_rows = a.sorted_rows;
    P.count = a.count;
    P.pca.cov_norm = a.cov_norm; P.pca.min_pts = a.min_pts; P.pca.normalise_hist = a.normalise_hist; P.pca.has_vp = a.has_vp;
    for (int i = 0; i < 3; ++i) P.pca.vp[i] = a.vp[i];
    P.pca.eig3 = a.eig3; P.pca.normal4 = a.normal4; P.pca.feat7 = a.feat7; P.pca.formula_set = a.formula_set;
    P.stat_scanned = nullptr;
    if (ctx->collect_stats) {
        CCG_CUDA(ctx, cudaMemsetAsync((char *)ctx->scratch.p + 1024, 0, 8, ctx->stream));
        P.stat_scanned = (unsigned long long *)((char *)ctx->scratch.p + 1024);
    }
    const int ntiles = P.tile_end - P.tile_begin;
    P.mom = nullptr;
    if (ctx->node_moments) {
        CCG_TRY(ensure_node_moments(ctx));
        P.mom = ctx->node_mom.as<float4>();
        k_tile_stream<StreamNodes><<<(ntiles + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), P);
    } else {
        k_tile_stream<Stream><<<(ntiles + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), P);
    }
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    if (ctx->collect_stats) {
        unsigned long long s = 0;
        CCG_CUDA(ctx, cudaMemcpyAsync(&s, P.stat_scanned, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->stats.leaves_scanned = (int64_t)s;
    }
    return CCG_OK;
}

// ---------------------------------------------------------------------------------------------------
// a9: eigen-derived features (elementwise; HBM-streaming: 12 B in, 28 B out per point)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_features(const float *__restrict__ eig3, size_t n, int formula_set, float *__restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float f[7];
    features7(eig3[3 * i], eig3[3 * i + 1], eig3[3 * i + 2], formula_set, f);
    float *o = out + 7 * i;
#pragma unroll
    for (int j = 0; j < 7; ++j) o[j] = f[j];
}

__global__ void __launch_bounds__(256) k_eigen_scalar(const float *__restrict__ eig3, size_t n, int which, float *__restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float l0 = eig3[3 * i], l1 = eig3[3 * i + 1], l2 = eig3[3 * i + 2];
    float r;
    if (l1 < l2) r = 0.f;                                    // the (0,0,1) marker: featureeval.cpp:1107-1110
    else if (which == 0) r = l0 > l2 ? __fdiv_rn(l1, l0) : 1.f;   // :1130-1135
    else {                                                   // :1201-1220, cv.h:498-501
        float nn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(l0, l0), __fmul_rn(l1, l1)), __fmul_rn(l2, l2)));
        float c = __fdiv_rn(l0, nn);
        r = c < 0.f ? 0.f : (c > 1.f ? 1.f : c);
    }
    out[i] = r;
}

int launch_features(ccg_ctx *ctx, const float *eig3, size_t n, int formula_set, float *out) {
    if (n == 0) return CCG_OK;
    k_features<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(eig3, n, formula_set, out);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}
int launch_eigen_scalar(ccg_ctx *ctx, const float *eig3, size_t n, int which, float *out) {
    if (n == 0) return CCG_OK;
    k_eigen_scalar<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(eig3, n, which, out);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

// out[order[i]] = in[i], rows of `words` 32-bit words
__global__ void __launch_bounds__(256) k_unpermute(const uint32_t *__restrict__ in, const unsigned *__restrict__ order, size_t n,
                                                   int words, uint32_t *__restrict__ out) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t i = t / words;
    int w = (int)(t % words);
    if (i >= n) return;
    out[(size_t)order[i] * words + w] = in[i * words + w];
}
int launch_unpermute(ccg_ctx *ctx, const void *in, size_t row_bytes, void *out) {
    if (ctx->n == 0) return CCG_OK;
    int words = (int)(row_bytes / 4);
    size_t total = ctx->n * (size_t)words;
    k_unpermute<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>((const uint32_t *)in, ctx->vals[ctx->order_sel].as<unsigned>(),
                                                                          ctx->n, words, (uint32_t *)out);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

// ---------------------------------------------------------------------------------------------------
// warp-per-query kernel: arbitrary query lists (stage (iii) of the grid normals, deferred stragglers).
// The 32 lanes score the 32 candidates of a leaf at once; the bounded top-k (k <= 64) lives one or two
// keys per lane and is updated with warp shuffles.
// ---------------------------------------------------------------------------------------------------

struct WarpStack {
    int base[kMaxLevels];
    unsigned mask[kMaxLevels];
};

__device__ __forceinline__ u64 shfl64(u64 v, int src) {
    unsigned lo = __shfl_sync(kFull, (unsigned)v, src), hi = __shfl_sync(kFull, (unsigned)(v >> 32), src);
    return ((u64)hi << 32) | lo;
}
__device__ __forceinline__ u64 shfl_up64(u64 v) {
    unsigned lo = __shfl_up_sync(kFull, (unsigned)v, 1), hi = __shfl_up_sync(kFull, (unsigned)(v >> 32), 1);
    return ((u64)hi << 32) | lo;
}

__global__ void __launch_bounds__(128) k_point_query(const __grid_constant__ TreeView T, const __grid_constant__ PointParams P) {
    __shared__ WarpStack s_st[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = P.nq_dev ? *P.nq_dev : P.nq;
    WarpStack &ws = s_st[warp];
    for (int qslot = blockIdx.x * 4 + warp; qslot < nq; qslot += gridDim.x * 4) {
    const int qi = P.queries ? P.queries[qslot] : qslot;
    const float4 q = T.orig[qi];
    const int k = P.k;
    u64 key0 = kEmptyKey, key1 = kEmptyKey;        // list slots lane and 32+lane, ascending
    u64 worst = kEmptyKey;                         // key at slot k-1
    float bound = P.rb;

    int sp = 0;
    const int top = T.nlevels - 1;                 // root level; entry d holds children at level top-1-d
    if (top == 0) {                                // single leaf
        if (lane == 0) { ws.base[0] = 0; ws.mask[0] = 1u; }
        sp = 1;
    } else {
        const int c = lane;
        bool pass = false;
        if (c < T.count[top - 1]) pass = d2_point_box(q, T.lo[top - 1][c], T.hi[top - 1][c]) <= bound;
        unsigned m = __ballot_sync(kFull, pass);
        if (lane == 0) { ws.base[0] = 0; ws.mask[0] = m; }
        sp = 1;
    }
    __syncwarp();
    const int lv0 = top == 0 ? 0 : top - 1;
    while (sp > 0) {
        const int lv = lv0 - (sp - 1);
        const int base = ws.base[sp - 1];
        unsigned m = ws.mask[sp - 1];
        __syncwarp();
        if (m == 0) { --sp; continue; }
        const bool mine = (m >> lane) & 1u;
        float bd = INFINITY;
        unsigned bmin = 0xffffffffu;                  // smallest original index below this lane's box
        if (mine) {
            const float4 blo = T.lo[lv][base + lane];
            bmin = (unsigned)__float_as_int(blo.w);
            bd = d2_point_box(q, blo, T.hi[lv][base + lane]);
            // A box exactly AT the k-th distance can only hold candidates tied with it: they win only with a smaller index.
            // blo.w = smallest original index below the box, so a pile of duplicates (PTX files are written with 6
            // decimals: common on real scans) costs one visit of its low-index leaf instead of all of them.
            if (bd == bound && __float_as_uint(bd) == (unsigned)(worst >> 32) && (unsigned)__float_as_int(blo.w) >= (unsigned)(worst & 0xffffffffu)) bd = INFINITY;
        }
        if (lv == 0) {
            // A group of up to 32 leaves: rank them ONCE by box distance (bitonic sort of the distance bits with the lane
            // id in the low 5 bits) and score them in that order until the first one beyond the bound — every later one
            // is at least as far.  Re-ranking the group after every leaf cost more than the scoring itself.
            if (lane == 0) ws.mask[sp - 1] = 0;
            --sp;
            unsigned key = mine ? ((__float_as_uint(bd) & ~31u) | (unsigned)lane) : (0xffffffe0u | (unsigned)lane);
#pragma unroll
            for (int kk = 2; kk <= 32; kk <<= 1) {
#pragma unroll
                for (int j = kk >> 1; j > 0; j >>= 1) {
                    const unsigned o = __shfl_xor_sync(kFull, key, j);
                    const bool up = ((lane & kk) == 0) == ((lane & j) == 0);
                    key = up ? min(key, o) : max(key, o);
                }
            }
            const int nmine = __popc(m);
            for (int r = 0; r < nmine; ++r) {
                const int cbit = (int)(__shfl_sync(kFull, key, r) & 31u);
                const float bdc = __shfl_sync(kFull, bd, cbit);
                if (bdc > bound) {
                    // ranks ignore the low 5 distance bits: a later leaf may still be a hair closer than this one
                    if (__float_as_uint(bdc) - __float_as_uint(bound) > 32u) break;
                    continue;
                }
                // the list tightened since the group was entered: a leaf AT the k-th distance holding only larger indices is moot
                if (__float_as_uint(bdc) == (unsigned)(worst >> 32) && __shfl_sync(kFull, bmin, cbit) >= (unsigned)(worst & 0xffffffffu)) continue;
                const float4 c = T.pts[(size_t)(base + cbit) * kLeaf + lane];
                const float dd = d2_point(q, c);
                const u64 ck = ((u64)__float_as_uint(dd) << 32) | (unsigned)__float_as_int(c.w);
                unsigned passers = __ballot_sync(kFull, dd <= bound && ck < worst);
                while (passers) {
                    const int src = __ffs(passers) - 1;
                    passers &= passers - 1;
                    const u64 nk = shfl64(ck, src);
                    if (!(nk < worst)) continue;                  // the bound moved since the ballot
                    const int pos = __popc(__ballot_sync(kFull, key0 < nk)) + __popc(__ballot_sync(kFull, key1 < nk));
                    const u64 up0 = shfl_up64(key0), up1 = shfl_up64(key1), carry = shfl64(key0, 31);
                    const int s0 = lane, s1 = 32 + lane;
                    key0 = s0 > pos ? up0 : (s0 == pos ? nk : key0);
                    key1 = s1 > pos ? (lane == 0 ? carry : up1) : (s1 == pos ? nk : key1);
                    worst = (k - 1) < 32 ? shfl64(key0, k - 1) : shfl64(key1, k - 1 - 32);
                    bound = fminf(P.rb, __uint_as_float((unsigned)(worst >> 32)));
                }
            }
        } else {
            // inner node: nearest surviving child first (tightens the bound quickly)
            int cbit;
            {
                unsigned kk = __float_as_uint(bd);
                unsigned best = kk;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(kFull, best, o));
                const unsigned who = __ballot_sync(kFull, mine && kk == best);
                cbit = who ? (__ffs(who) - 1) : (__ffs(m) - 1);
                // drop children that no longer pass the current bound
                const unsigned still = __ballot_sync(kFull, mine && bd <= bound);
                m &= still;
                if (!((m >> cbit) & 1u)) { if (lane == 0) ws.mask[sp - 1] = 0; __syncwarp(); --sp; continue; }
            }
            m &= ~(1u << cbit);
            if (lane == 0) ws.mask[sp - 1] = m;
            const int first = (base + cbit) << 5;
            const int c = first + lane;
            bool pass = false;
            if (c < T.count[lv - 1]) pass = d2_point_box(q, T.lo[lv - 1][c], T.hi[lv - 1][c]) <= bound;
            const unsigned m2 = __ballot_sync(kFull, pass);
            if (m2) {
                if (lane == 0) { ws.base[sp] = first; ws.mask[sp] = m2; }
                ++sp;
            }
        }
        __syncwarp();
    }

    const size_t row = P.rows ? (size_t)P.rows[qslot] : (P.list_rows ? (size_t)qslot : (size_t)qi);
    if (P.mode == MODE_KNN) {
        int cnt = 0;
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int s = r * 32 + lane;
            const u64 kk = r ? key1 : key0;
            const bool real = s < k && kk != kEmptyKey;
            cnt += __popc(__ballot_sync(kFull, real));
            if (s < k) {
                if (P.idx) P.idx[row * k + s] = real ? (int)(unsigned)(kk & 0xffffffffu) : -1;
                if (P.d2) P.d2[row * k + s] = real ? __uint_as_float((unsigned)(kk >> 32)) : INFINITY;
            }
        }
        if (P.count && lane == 0) P.count[row] = cnt;
    } else if (P.mode == MODE_PCA) {
        float m[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) m[i] = 0.f;
        int cnt = 0;
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int s = r * 32 + lane;
            const u64 kk = r ? key1 : key0;
            if (s < k && kk != kEmptyKey) {
                const float4 p = T.orig[(unsigned)(kk & 0xffffffffu)];
                float x = p.x - q.x, y = p.y - q.y, z = p.z - q.z;
                m[0] += x; m[1] += y; m[2] += z;
                m[3] += x * x; m[4] += x * y; m[5] += x * z; m[6] += y * y; m[7] += y * z; m[8] += z * z;
                ++cnt;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int i = 0; i < 9; ++i) m[i] += __shfl_xor_sync(kFull, m[i], o);
            cnt += __shfl_xor_sync(kFull, cnt, o);
        }
        if (lane == 0) {
            pca_finish(P.pca, q, m, cnt, row);
            if (P.count) P.count[row] = cnt;
        }
    } else {
        double s = 0.0;
        for (int j = 1; j < k; ++j) {
            const u64 kk = j < 32 ? shfl64(key0, j) : shfl64(key1, j - 32);
            if (kk != kEmptyKey) s += (double)__fsqrt_rn(__uint_as_float((unsigned)(kk >> 32)));
        }
        if (lane == 0) P.mean_d[row] = (float)(s / (double)P.mean_k);
    }
    __syncwarp();
    }   // query loop
}

int launch_point_query(ccg_ctx *ctx, const PointArgs &a) {
    if (ctx->n == 0 || a.nq == 0) return CCG_OK;
    if (a.k < 1 || a.k > CCG_MAX_K) return set_error(ctx, CCG_ERR_UNSUPPORTED, "k=%d outside [1,%d]", a.k, CCG_MAX_K);
    PointParams P{};
    P.queries = a.queries; P.nq = a.nq; P.mode = a.mode; P.k = a.k; P.rb = a.rb; P.list_rows = a.list_rows;
    P.idx = a.idx; P.d2 = a.d2; P.count = a.count;
    P.pca.cov_norm = a.cov_norm; P.pca.min_pts = a.min_pts; P.pca.normalise_hist = a.normalise_hist; P.pca.has_vp = a.has_vp;
    for (int i = 0; i < 3; ++i) P.pca.vp[i] = a.vp[i];
    P.pca.eig3 = a.eig3; P.pca.normal4 = a.normal4; P.pca.feat7 = a.feat7; P.pca.formula_set = a.formula_set;
    P.mean_k = a.mean_k; P.mean_d = a.mean_d;
    k_point_query<<<(a.nq + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), P);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

}  // namespace ccg
