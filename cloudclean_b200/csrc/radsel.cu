// radsel.cu — capped radius neighbourhoods for ANY cap, and the radius-neighbourhood reductions without a list.
//
// FLANN's radiusSearch(i, r, idx, d2, max_nn) (call sites cv.h:98, cv.h:162, cv.cpp:273, featureeval/utils.h:50,145) returns the
// max_nn NEAREST points inside r whenever 0 < max_nn < neighbourhood size.  The list kernels of query.cu keep at most 64
// values per query in registers; the reference runs with pca_max_nn = 1000 (tests.json:7) and max_nn_ = 100000
// (markov.cpp:54), so larger caps are answered here WITHOUT a list:
//
//   k_tile_kth     per query the RULE of its neighbourhood: T = the max_nn-th smallest d2 inside r (r^2 bound itself when the
//                  cap does not bind) and I = the largest original index admitted among candidates tied at T.  T is found by
//                  a counting select over the traversal: 64-bin histograms (uniform in the distance) per lane in shared
//                  memory narrow the bracket level by level until it holds <= 64 candidates, which are then collected and
//                  ranked exactly.  The FIRST pass lowers its bound while it walks (as soon as the bins up to b hold max_nn
//                  candidates nothing above bin b matters), later passes are bounded by the bracket chosen so far: a cap far
//                  below the neighbourhood size never opens the far part of the neighbourhood.  I comes from the same kind
//                  of select over the index bits when the cap admits more than 64 members of a tie class.
//   k_tile_rule<V> one traversal bounded by T that feeds {d2 < T or (d2 == T and index <= I)} — exactly the
//                  (d2, index)-ordered max_nn nearest — to a streaming visitor: PCA moments (getPCA / getHist), stdev_dist
//                  (cv.cpp:166-230), normal_stdev (cv.h:123-206), calcDistHist / calcIntensityHist (featureeval/utils.h).
//
// The fp32 sums of the visitors run in traversal order, not in the reference's ascending-distance order: counts, histograms
// and neighbour SETS are exact, the sums agree to a stated 1e-5 (tests/test_gpu_parity.py).
#include "tile.cuh"
#include <cub/device/device_scan.cuh>
#include <cub/device/device_segmented_sort.cuh>
#include <limits.h>

namespace ccg {

constexpr int kSelBins = 64;          // histogram bins per level and lane
constexpr int kSelKeep = 64;          // candidates a lane can rank exactly
constexpr int kSelLevels = 4;

// bin of d2 at level `lvl` given the bins chosen at the levels above; `match` = the candidate lies inside those bins.
// The same expression runs in every pass, so a candidate always lands in the same bin.  The bins are uniform in the DISTANCE
// (f = 64 d / r, monotone in d2): on a surface the count grows with d2, so bin b of level 0 holds ~(2b + 1) / 4096 of the
// neighbourhood and a cap far below the neighbourhood size (1000 of 24 k, 1000 of 570 k) still lands a few bins up, where the
// edge above it overshoots the cap distance by a few percent instead of several times.
__device__ __forceinline__ int bin_path(float d2, float scale, int lvl, const int sel[kSelLevels], bool &match) {
    float f = __fmul_rn(__fsqrt_rn(d2), scale);
    int b = min(kSelBins - 1, (int)f);
    match = true;
#pragma unroll
    for (int l = 0; l < kSelLevels; ++l) {
        if (l < lvl) {
            match = match && (b == sel[l]);
            f = __fmul_rn(__fsub_rn(f, (float)sel[l]), (float)kSelBins);
            b = min(kSelBins - 1, max(0, (int)f));
        }
    }
    return b;
}

// Upper edge, as a d2 bound, of the bracket chosen at the first `lvl` levels: every candidate whose bins match sel[0..lvl) has
// d2 <= the returned value.  bin_path is exact after its first multiplication (f - sel is exact, * 64 is exact), so the bracket
// is f = sqrt(d2) * scale in [F, F + 64^-(lvl-1)) in units of level-0 bins; the margins cover the roundings of this sum, of the
// division, of the square and of f itself (< 1e-6 together).  The top bins are clamped (f may reach 64 + 2^-18): their edge is
// r^2 itself.
__device__ __forceinline__ float bracket_edge(float rb, float scale, int lvl, const int sel[kSelLevels]) {
    if (lvl == 0) return rb;
    float F = 0.f, w = 1.f, wl = 1.f;
#pragma unroll
    for (int l = 0; l < kSelLevels; ++l) {
        if (l < lvl) { F += (float)sel[l] * w; wl = w; w *= 1.f / (float)kSelBins; }
    }
    const float e = __fmul_rn(__fdiv_rn(F + wl, scale), 1.00001f);
    return fminf(rb, __fmul_rn(__fmul_rn(e, e), 1.00001f));
}

// Inner edge of the same bracket: every candidate with d2 below the returned value has f < F, i.e. lies below the bracket.  A box
// whose farthest corner is nearer than this holds nothing a bracket pass is looking for (tile.cuh: visits_nodes).
__device__ __forceinline__ float bracket_inner(float scale, int lvl, const int sel[kSelLevels]) {
    if (lvl == 0) return 0.f;
    float F = 0.f, w = 1.f;
#pragma unroll
    for (int l = 0; l < kSelLevels; ++l) {
        if (l < lvl) { F += (float)sel[l] * w; w *= 1.f / (float)kSelBins; }
    }
    const float e = __fmul_rn(__fdiv_rn(F, scale), 0.99999f);
    const float in = __fmul_rn(__fmul_rn(e, e), 0.99999f);
    return in == in ? in : 0.f;                       // scale == 0 (r == 0): no inner edge
}

struct HistV {                        // level pass: histogram of the candidates inside the chosen bins
    float bound, scale; int lvl; int sel[kSelLevels];
    unsigned *hist;                   // this lane's column: bin b at hist[b * 32]
    // First pass only (max_nn > 0): the bound FOLLOWS the histogram down.  The traversal visits near boxes first; as soon as the
    // bins up to b hold max_nn candidates, the max_nn-th smallest d2 cannot lie above bin b and the bound drops to that bin's
    // edge, so the far part of a neighbourhood much larger than the cap (24 k points inside r = 0.2 m on the 30 M scan, cap
    // 1000) is never opened.  Bins <= b stay complete: a box is only ever pruned against a bound >= the final one.
    int max_nn, inb, limit;
    static constexpr bool kNodes = true;              // later passes: boxes inside the bracket's inner edge are skipped
    float inner; bool nodes;                          // nodes: WARP-UNIFORM switch (the test votes with a full mask)
    __device__ __forceinline__ bool node_test() const { return nodes; }
    __device__ __forceinline__ bool whole(float dfar) const { return dfar < inner; }
    __device__ __forceinline__ void add_node(size_t, const float4 &, const float4 &, const float4 &) {}
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
        int add = 0;
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float d2 = d2_point(q, leaf[j]);
            if (d2 <= bound) {
                bool match;
                const int b = bin_path(d2, scale, lvl, sel, match);
                if (match) { hist[b * 32] += 1; ++add; }
            }
        }
        if (max_nn > 0) {
            inb += add;
            if (inb > limit) {
                int cum = 0, b = 0;
                for (; b < kSelBins - 1; ++b) {
                    const int hb = (int)hist[b * 32];
                    if (cum + hb >= max_nn) break;
                    cum += hb;
                }
                const int one[kSelLevels] = {b, 0, 0, 0};
                bound = fminf(bound, bracket_edge(bound, scale, 1, one));
                inb = cum + (int)hist[b * 32];
                limit = inb < (1 << 30) ? inb + (max_nn >> 2) + 32 : INT_MAX;
            }
        }
    }
};

struct KeepV {                        // select pass: the d2 values inside the chosen bins (<= kSelKeep kept), their min / max
    float bound, scale; int lvl; int sel[kSelLevels];
    float *vals; int cnt; float vmin, vmax;
    static constexpr bool kNodes = true;
    float inner; bool nodes;                          // nodes: WARP-UNIFORM switch (the test votes with a full mask)
    __device__ __forceinline__ bool node_test() const { return nodes; }
    __device__ __forceinline__ bool whole(float dfar) const { return dfar < inner; }
    __device__ __forceinline__ void add_node(size_t, const float4 &, const float4 &, const float4 &) {}
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float d2 = d2_point(q, leaf[j]);
            if (d2 <= bound) {
                bool match;
                bin_path(d2, scale, lvl, sel, match);
                if (match) {
                    if (cnt < kSelKeep) vals[cnt * 32] = d2;
                    ++cnt;
                    vmin = fminf(vmin, d2); vmax = fmaxf(vmax, d2);
                }
            }
        }
    }
};

struct TieV {                         // tie pass (bound = T): candidates below T, and the smallest indices among those AT T
    float bound; int less, tied;
    int *idx;                         // up to kSelKeep smallest tied indices, unsorted
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            const float d2 = d2_point(q, c);
            if (d2 < bound) ++less;
            else if (d2 == bound) {
                const int ci = __float_as_int(c.w);
                if (tied < kSelKeep) idx[tied * 32] = ci;
                else {                                      // keep the kSelKeep smallest: replace the largest kept one
                    int worst = 0, wv = idx[0];
                    for (int e = 1; e < kSelKeep; ++e) { const int v = idx[e * 32]; if (v > wv) { wv = v; worst = e; } }
                    if (ci < wv) idx[worst * 32] = ci;
                }
                ++tied;
            }
        }
    }
};

// index select (a tie class AT the cap distance of which the cap admits more than kSelKeep members — piles of duplicates,
// lattices): histogram of 6 index bits of the candidates tied at `bound` whose higher bits equal `prefix`
struct TieIdxV {
    float bound; int shift; unsigned prefix;
    unsigned *hist;
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (d2_point(q, c) == bound) {
                const unsigned long long ci = (unsigned)__float_as_int(c.w);
                if ((unsigned)(ci >> (shift + 6)) == prefix) hist[((unsigned)(ci >> shift) & 63u) * 32] += 1;
            }
        }
    }
};

// err bit 1 = more than kSelKeep distinct values left after the last level
__global__ void __launch_bounds__(128) k_tile_kth(const __grid_constant__ TreeView T, float rb, int max_nn, float *__restrict__ Tpos,
                                                  int32_t *__restrict__ Ipos, int *__restrict__ err) {
    __shared__ WarpScratch s_ws[4];
    __shared__ unsigned s_hist[4][kSelBins * 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    unsigned *hist = &s_hist[warp][lane];
    float t_out = rb; int i_out = INT_MAX;

    // The first histogram pass (level 0, every lane) also counts the neighbourhood: count = sum of the bins (a lane whose bound
    // dropped on the way holds more than max_nn candidates below it, so "count > max_nn" reads the same either way).
    const float scale = rb > 0.f ? __fdiv_rn((float)kSelBins, __fsqrt_rn(rb)) : 0.f;
    int sel[kSelLevels] = {0, 0, 0, 0};
    int lvl = 0, kprime = max_nn, nb = 0;                 // lvl: levels THIS lane has refined; nb: candidates left in its bracket
    bool need = active, first = true;
    while (first || __any_sync(kFull, need && nb > kSelKeep && lvl < kSelLevels)) {
        const bool mine = first ? active : (need && nb > kSelKeep && lvl < kSelLevels);
#pragma unroll 8
        for (int b = 0; b < kSelBins; ++b) hist[b * 32] = 0;
        HistV h; h.scale = scale; h.lvl = lvl; h.hist = hist;
#pragma unroll
        for (int l = 0; l < kSelLevels; ++l) h.sel[l] = sel[l];
        // later passes only need the bracket chosen so far: everything in it lies below its upper edge
        h.bound = mine ? bracket_edge(rb, scale, lvl, sel) : -1.f;
        h.inner = bracket_inner(scale, lvl, sel); h.nodes = !first;
        h.max_nn = first ? max_nn : 0; h.inb = 0; h.limit = max_nn < (1 << 30) ? max_nn + (max_nn >> 2) + 32 : INT_MAX;
        traverse_tile(T, tile, q, h, s_ws[warp]);
        if (first) {
            int c = 0;
            for (int b = 0; b < kSelBins; ++b) c += (int)hist[b * 32];
            nb = h.bound < rb ? INT_MAX : c;              // a dropped bound: c undercounts, the bracket must be refined
            need = active && c > max_nn;
            first = false;
            if (!__any_sync(kFull, need)) { Tpos[pos] = t_out; Ipos[pos] = i_out; return; }   // the cap binds nowhere in this tile
        }
        if (mine && need && (lvl > 0 || nb > kSelKeep)) {
            int cum = 0, b = 0;
            for (; b < kSelBins - 1; ++b) {
                const int hb = (int)hist[b * 32];
                if (cum + hb >= kprime) break;
                cum += hb;
            }
            kprime -= cum;
            nb = (int)hist[b * 32];
#pragma unroll
            for (int l = 0; l < kSelLevels; ++l) if (l == lvl) sel[l] = b;
            ++lvl;
        }
    }
    // select pass: the candidates inside the lane's bracket, ranked exactly
    bool tie_pass = false;                                 // does this lane need the index tie-break at T?
    {
        KeepV kv; kv.bound = need ? bracket_edge(rb, scale, lvl, sel) : -1.f; kv.scale = scale; kv.lvl = lvl; kv.vals = reinterpret_cast<float *>(hist);
        kv.cnt = 0; kv.vmin = INFINITY; kv.vmax = -INFINITY; kv.inner = bracket_inner(scale, lvl, sel); kv.nodes = true;
#pragma unroll
        for (int l = 0; l < kSelLevels; ++l) kv.sel[l] = sel[l];
        traverse_tile(T, tile, q, kv, s_ws[warp]);
        if (need) {
            if (kv.cnt <= kSelKeep) {
                // the kprime-th smallest of the kept values: v with #{< v} < kprime <= #{<= v}.  Every candidate AT that value
                // lies in this bracket, so #{<= v} == kprime says the tie class fits the cap and no index decides anything.
                for (int i = 0; i < kv.cnt; ++i) {
                    const float vi = kv.vals[i * 32];
                    int lt = 0, le = 0;
                    for (int j = 0; j < kv.cnt; ++j) { const float vj = kv.vals[j * 32]; lt += vj < vi ? 1 : 0; le += vj <= vi ? 1 : 0; }
                    if (lt < kprime && kprime <= le) { t_out = vi; tie_pass = le > kprime; break; }
                }
            } else if (kv.vmin == kv.vmax) {
                t_out = kv.vmin;                             // a bracket of identical values: the threshold is that value
                tie_pass = true;
            } else {
                atomicOr(err, 1);
            }
        }
    }
    __syncwarp();
    // tie pass (only when some lane's tie class overflows its cap): how many candidates lie strictly below T, which indices tie AT it
    bool wide = false; int mleft = 0;
    if (__any_sync(kFull, tie_pass)) {
        TieV tv; tv.bound = tie_pass ? t_out : -1.f; tv.less = 0; tv.tied = 0; tv.idx = reinterpret_cast<int *>(hist);
        traverse_tile(T, tile, q, tv, s_ws[warp]);
        if (tie_pass) {
            const int m = max_nn - tv.less;                  // tied candidates the cap still admits (>= 1)
            if (tv.tied > m) {
                if (m <= kSelKeep) {
                    const int kept = min(tv.tied, kSelKeep);
                    for (int i = 0; i < kept; ++i) {         // the m-th smallest kept index (indices are distinct)
                        const int vi = tv.idx[i * 32];
                        int lt = 0;
                        for (int j = 0; j < kept; ++j) lt += tv.idx[j * 32] < vi ? 1 : 0;
                        if (lt == m - 1) { i_out = vi; break; }
                    }
                } else {
                    wide = true; mleft = m;
                }
            }
        }
    }
    // the m-th smallest tied index for m > kSelKeep: counting select over the index bits, 6 bits per traversal (bounded by T)
    if (__any_sync(kFull, wide)) {
        unsigned prefix = 0;
        for (int l = 0; l < 6; ++l) {
            const int shift = 30 - 6 * l;
            __syncwarp();
#pragma unroll 8
            for (int b = 0; b < kSelBins; ++b) hist[b * 32] = 0;
            TieIdxV iv; iv.bound = wide ? t_out : -1.f; iv.shift = shift; iv.prefix = prefix; iv.hist = hist;
            traverse_tile(T, tile, q, iv, s_ws[warp]);
            if (wide) {
                int cum = 0, b = 0;
                for (; b < kSelBins - 1; ++b) {
                    const int hb = (int)hist[b * 32];
                    if (cum + hb >= mleft) break;
                    cum += hb;
                }
                mleft -= cum;
                prefix = (prefix << 6) | (unsigned)b;
            }
        }
        if (wide) i_out = (int)prefix;
    }
    Tpos[pos] = t_out; Ipos[pos] = i_out;
}

// ---------------------------------------------------------------------------------------------------
// streaming visitors over the rule {d2 < T or (d2 == T and index <= I)}
// ---------------------------------------------------------------------------------------------------
struct RuleBase {
    float bound; int tie; int count;
    __device__ __forceinline__ bool accept(float d2, int idx) const { return d2 < bound || (d2 == bound && idx <= tie); }
};

struct RuleMoments : RuleBase {       // getPCA / getHist with a cap beyond the list kernels (cv.h:98-113)
    struct Args { PcaOut pca; int32_t *count; const float4 *mom; };
    static constexpr int kSmemWords = 0;
    static constexpr bool kNodes = true;  // a box strictly inside the cap distance is admitted whatever its indices: aggregated term
    const float4 *mom;
    float m[9];
    __device__ __forceinline__ bool node_test() const { return mom != nullptr; }
    __device__ __forceinline__ bool whole(float dfar) const { return dfar < bound; }
    __device__ __forceinline__ void add_node(size_t node, const float4 &lo, const float4 &hi, const float4 &q) {
        count += add_box_moments(mom, node, lo, hi, q, true, m);
    }
    __device__ __forceinline__ void init(const Args &a, const float4 &, unsigned *) {
        mom = a.mom;
#pragma unroll
        for (int i = 0; i < 9; ++i) m[i] = 0.f;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 4
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (accept(d2_point(q, c), __float_as_int(c.w))) {
                ++count;
                const float x = c.x - q.x, y = c.y - q.y, z = c.z - q.z;
                m[0] += x; m[1] += y; m[2] += z;
                m[3] += x * x; m[4] += x * y; m[5] += x * z; m[6] += y * y; m[7] += y * z; m[8] += z * z;
            }
        }
    }
    __device__ __forceinline__ void finish(const Args &a, const float4 &q, size_t row) {
        pca_finish(a.pca, q, m, count, row);
        if (a.count) a.count[row] = count;
    }
};

struct RuleStdevDist : RuleBase {     // stdev_dist (cv.cpp:166-230): fp32 sums of the neighbour distances, variance as written
    struct Args { int use_depth; float cx, cy, cz; float *out; };
    static constexpr int kSmemWords = 0;
    float rx, ry, rz, sum, sumsq;
    __device__ __forceinline__ void init(const Args &a, const float4 &q, unsigned *) {
        rx = a.use_depth ? a.cx : q.x; ry = a.use_depth ? a.cy : q.y; rz = a.use_depth ? a.cz : q.z;
        sum = sumsq = 0.f;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 4
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (accept(d2_point(q, c), __float_as_int(c.w))) {
                ++count;
                const float dx = __fsub_rn(c.x, rx), dy = __fsub_rn(c.y, ry), dz = __fsub_rn(c.z, rz);
                const float dist = __fsqrt_rn(eigen_dot3(dx, dy, dz, dx, dy, dz));
                sum = __fadd_rn(sum, dist);
                sumsq = __fadd_rn(sumsq, __fmul_rn(dist, dist));
            }
        }
    }
    __device__ __forceinline__ void finish(const Args &a, const float4 &, size_t row) {
        const float fm = (float)count;
        a.out[row] = count > 2 ? __fdiv_rn(__fsub_rn(sumsq, __fdiv_rn(__fmul_rn(sum, sum), fm)), __fsub_rn(fm, 1.f)) : 0.f;
    }
};

struct RuleNormalStdev : RuleBase {   // normal_stdev (cv.h:123-206), arithmetic as written there (see k_normal_stdev in api.cu)
    struct Args { const char *normals; size_t stride; float *out; };
    static constexpr int kSmemWords = 0;
    const char *normals; size_t stride;
    float cx, cy, cz, cn, sum, sumsq;
    __device__ __forceinline__ void init(const Args &a, const float4 &q, unsigned *) {
        normals = a.normals; stride = a.stride; sum = sumsq = 0.f; cx = cy = cz = cn = 0.f;
        if (bound >= 0.f) {
            const float *cur = reinterpret_cast<const float *>(normals + (size_t)__float_as_int(q.w) * stride);
            cx = __ldg(cur); cy = __ldg(cur + 1); cz = __ldg(cur + 2);
            cn = __fsqrt_rn(eigen_dot3(cx, cy, cz, cx, cy, cz));
        }
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            if (accept(d2_point(q, c), __float_as_int(c.w))) {
                ++count;
                const float *nb = reinterpret_cast<const float *>(normals + (size_t)__float_as_int(c.w) * stride);
                const float x = __ldg(nb), y = __ldg(nb + 1), z = __ldg(nb + 2);
                const float dot = eigen_dot3(x, y, z, cx, cy, cz);
                const float nn = __fsqrt_rn(eigen_dot3(x, y, z, x, y, z));
                float cosine = __fmul_rn(__fdiv_rn(dot, nn), cn);
                cosine = cosine < 0.f ? 0.f : (cosine > 1.f ? 1.f : cosine);
                const float angle = (float)((double)acosf(cosine) / 3.14159265358979323846);
                sum = __fadd_rn(sum, angle);
                sumsq = __fadd_rn(sumsq, __fmul_rn(angle, angle));
            }
        }
    }
    __device__ __forceinline__ void finish(const Args &a, const float4 &, size_t row) {
        float res = 0.f;
        if (cx == cx && count >= 3) {
            const float fm = (float)count;
            const double mean = (double)__fdiv_rn(sum, fm);
            const float sd = (float)sqrt((double)__fdiv_rn(sumsq, fm) - mean * mean);
            if (sd == sd) res = sd;
        }
        a.out[row] = res;
    }
};

struct RuleHist : RuleBase {          // calcDistHist / calcIntensityHist (featureeval/utils.h:18-195): counts per bin, 64 bins per pass
    struct Args { int bins, bin0, which; double radius; const char *attr; size_t stride; float *hist; };
    static constexpr int kSmemWords = 64 * 32;
    unsigned *cnt; int bins, bin0, which; double radius; const char *attr; size_t stride;
    __device__ __forceinline__ void init(const Args &a, const float4 &, unsigned *smem) {
        cnt = smem; bins = a.bins; bin0 = a.bin0; which = a.which; radius = a.radius; attr = a.attr; stride = a.stride;
#pragma unroll 8
        for (int b = 0; b < 64; ++b) cnt[b * 32] = 0;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 2
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            const float d2 = d2_point(q, c);
            if (accept(d2, __float_as_int(c.w))) {
                ++count;
                int bin;
                if (which == 0) bin = (int)(((double)__fsqrt_rn(d2) / radius) * (double)(bins - 1));
                else bin = (int)__fmul_rn(__ldg(reinterpret_cast<const float *>(attr + (size_t)__float_as_int(c.w) * stride)), (float)(bins - 1));
                bin = bin < 0 ? 0 : (bin > bins - 1 ? bins - 1 : bin);
                bin -= bin0;
                if (bin >= 0 && bin < 64) cnt[bin * 32] += 1;
            }
        }
    }
    __device__ __forceinline__ void finish(const Args &a, const float4 &, size_t row) {
        float *o = a.hist + row * (size_t)bins;
        const float inc = (float)(1.0 / (double)count);
        for (int b = 0; b < 64 && bin0 + b < bins; ++b) {
            float v = 0.f;
            if (count >= 3) {
                const unsigned c = cnt[b * 32];
                for (unsigned i = 0; i < c; ++i) v = __fadd_rn(v, inc);        // the reference adds the increment once per neighbour
            }
            o[bin0 + b] = v;
        }
    }
};

struct RuleCount : RuleBase {         // CSR pass 1: neighbours per point
    struct Args { long long *count64; const float4 *mom; };
    static constexpr int kSmemWords = 0;
    static constexpr bool kNodes = true;
    const float4 *mom;
    __device__ __forceinline__ bool node_test() const { return mom != nullptr; }
    __device__ __forceinline__ bool whole(float dfar) const { return dfar < bound; }
    __device__ __forceinline__ void add_node(size_t node, const float4 &, const float4 &, const float4 &) {
        count += __float_as_int(__ldg(mom + 3 * node).w);
    }
    __device__ __forceinline__ void init(const Args &a, const float4 &, unsigned *) { mom = a.mom; }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 8
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            count += accept(d2_point(q, c), __float_as_int(c.w)) ? 1 : 0;
        }
    }
    __device__ __forceinline__ void finish(const Args &a, const float4 &, size_t row) { a.count64[row] = count; }
};

struct RuleFill : RuleBase {          // CSR pass 2: (d2 bits << 32 | index) keys of the neighbours, in traversal order
    struct Args { const long long *offsets; unsigned long long *keys; };
    static constexpr int kSmemWords = 0;
    unsigned long long *dst;
    __device__ __forceinline__ void init(const Args &a, const float4 &q, unsigned *) {
        dst = bound >= 0.f ? a.keys + a.offsets[__float_as_int(q.w)] : nullptr;
    }
    __device__ __forceinline__ void scan_leaf(const float4 &q, const float4 *leaf) {
#pragma unroll 4
        for (int j = 0; j < kLeaf; ++j) {
            const float4 c = leaf[j];
            const float d2 = d2_point(q, c);
            if (accept(d2, __float_as_int(c.w))) {
                dst[count] = ((unsigned long long)__float_as_uint(d2) << 32) | (unsigned)__float_as_int(c.w);
                ++count;
            }
        }
    }
    __device__ __forceinline__ void finish(const Args &, const float4 &, size_t) {}
};

// sorted keys -> the two output arrays
__global__ void __launch_bounds__(256) k_csr_unpack(const unsigned long long *__restrict__ keys, long long total, int32_t *__restrict__ idx,
                                                    float *__restrict__ d2) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const unsigned long long k = keys[i];
    if (idx) idx[i] = (int32_t)(unsigned)(k & 0xffffffffu);
    if (d2) d2[i] = __uint_as_float((unsigned)(k >> 32));
}

template <class V>
__global__ void __launch_bounds__(128) k_tile_rule(const __grid_constant__ TreeView T, float rb, const float *__restrict__ Tpos,
                                                   const int32_t *__restrict__ Ipos, const __grid_constant__ typename V::Args A) {
    __shared__ WarpScratch s_ws[4];
    __shared__ unsigned s_x[4][V::kSmemWords ? V::kSmemWords : 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * 4 + warp;
    if (tile >= T.count[0]) return;
    const size_t pos = (size_t)tile * kLeaf + lane;
    const float4 q = T.pts[pos];
    const bool active = pos < (size_t)T.n;
    V v;
    v.bound = active ? (Tpos ? Tpos[pos] : rb) : -1.f;
    v.tie = Ipos ? Ipos[pos] : INT_MAX;
    v.count = 0;
    v.init(A, q, &s_x[warp][V::kSmemWords ? lane : 0]);
    traverse_tile(T, tile, q, v, s_ws[warp]);
    if (active) v.finish(A, q, (size_t)__float_as_int(q.w));
}

// ---------------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------------
// The rule arrays of a capped search: *Tpos / *Ipos stay NULL when the cap cannot bind (max_nn == 0 or >= n: FLANN caps only
// when 0 < max_nn < n).  Synchronises once to read the error bit.
int launch_radius_rule(ccg_ctx *ctx, float rb, int max_nn, const float **Tpos, const int32_t **Ipos) {
    *Tpos = nullptr; *Ipos = nullptr;
    if (ctx->n == 0 || max_nn <= 0 || (size_t)max_nn >= ctx->n) return CCG_OK;
    CCG_CUDA(ctx, ctx->rule_T.reserve(ctx->npad * 4));
    CCG_CUDA(ctx, ctx->rule_I.reserve(ctx->npad * 4));
    int *err = (int *)((char *)ctx->scratch.p + 3072);
    CCG_CUDA(ctx, cudaMemsetAsync(err, 0, 4, ctx->stream));
    k_tile_kth<<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb, max_nn, ctx->rule_T.as<float>(), ctx->rule_I.as<int32_t>(), err);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    int h = 0;
    CCG_CUDA(ctx, cudaMemcpyAsync(&h, err, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h & 1) return set_error(ctx, CCG_ERR_UNSUPPORTED, "capped radius search: more than %d distinct distances inside a 2^-24 r^2 bracket", kSelKeep);
    *Tpos = ctx->rule_T.as<float>(); *Ipos = ctx->rule_I.as<int32_t>();
    return CCG_OK;
}

template <class V>
static int launch_rule(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const typename V::Args &A) {
    if (ctx->n == 0) return CCG_OK;
    k_tile_rule<V><<<(ctx->nleaf + 3) / 4, 128, 0, ctx->stream>>>(ctx->view(), rb, Tpos, Ipos, A);
    ctx->stats.launches++;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

int launch_rule_pca(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const StreamArgs &a) {
    RuleMoments::Args A{};
    A.pca.cov_norm = a.cov_norm; A.pca.min_pts = a.min_pts; A.pca.normalise_hist = a.normalise_hist; A.pca.has_vp = a.has_vp;
    for (int i = 0; i < 3; ++i) A.pca.vp[i] = a.vp[i];
    A.pca.eig3 = a.eig3; A.pca.normal4 = a.normal4; A.pca.feat7 = a.feat7; A.pca.formula_set = a.formula_set;
    A.count = a.count;
    A.mom = nullptr;
    if (ctx->node_moments) { CCG_TRY(ensure_node_moments(ctx)); A.mom = ctx->node_mom.as<float4>(); }
    return launch_rule<RuleMoments>(ctx, rb, Tpos, Ipos, A);
}
int launch_rule_stdev_dist(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, int use_depth, const float centre[3], float *out) {
    RuleStdevDist::Args A{use_depth, centre[0], centre[1], centre[2], out};
    return launch_rule<RuleStdevDist>(ctx, rb, Tpos, Ipos, A);
}
int launch_rule_normal_stdev(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const void *normals, size_t stride, float *out) {
    RuleNormalStdev::Args A{(const char *)normals, stride, out};
    return launch_rule<RuleNormalStdev>(ctx, rb, Tpos, Ipos, A);
}
int launch_rule_hist(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, int bins, int which, double radius, const void *attr,
                     size_t stride, float *hist) {
    for (int bin0 = 0; bin0 < bins; bin0 += 64) {
        RuleHist::Args A{bins, bin0, which, radius, (const char *)attr, stride, hist};
        CCG_TRY(launch_rule<RuleHist>(ctx, rb, Tpos, Ipos, A));
    }
    return CCG_OK;
}

// Variable-length neighbour lists (CSR): counts -> offsets (exclusive scan) -> keys in traversal order -> every row sorted
// ascending by (d2, index) with one segmented radix sort -> idx / d2.  offsets_dev: long long[n + 1] (device).
int launch_radius_csr_count(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, long long *offsets_dev, long long *total_host) {
    const size_t n = ctx->n;
    RuleCount::Args A{offsets_dev, nullptr};
    if (ctx->node_moments) { CCG_TRY(ensure_node_moments(ctx)); A.mom = ctx->node_mom.as<float4>(); }
    CCG_TRY(launch_rule<RuleCount>(ctx, rb, Tpos, Ipos, A));
    size_t tmp_bytes = 0;
    CCG_CUDA(ctx, cudaMemsetAsync(offsets_dev + n, 0, sizeof(long long), ctx->stream));
    CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, offsets_dev, offsets_dev, (int)(n + 1), ctx->stream));
    CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
    CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->sort_tmp.p, tmp_bytes, offsets_dev, offsets_dev, (int)(n + 1), ctx->stream));
    ctx->stats.launches += 2;
    CCG_CUDA(ctx, cudaMemcpyAsync(total_host, offsets_dev + n, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return CCG_OK;
}

int launch_radius_csr_fill(ccg_ctx *ctx, float rb, const float *Tpos, const int32_t *Ipos, const long long *offsets_dev, long long total,
                           int32_t *idx_dev, float *d2_dev) {
    if (total <= 0) return CCG_OK;
    if (total > 0x7fffffffLL) return set_error(ctx, CCG_ERR_UNSUPPORTED, "radius lists: %lld neighbours in total exceed 2^31-1 per call", total);
    const size_t n = ctx->n;
    CCG_CUDA(ctx, ctx->tmp[2].reserve((size_t)total * 16 + 64));
    unsigned long long *k0 = ctx->tmp[2].as<unsigned long long>(), *k1 = k0 + total;
    RuleFill::Args A{offsets_dev, k0};
    CCG_TRY(launch_rule<RuleFill>(ctx, rb, Tpos, Ipos, A));
    cub::DoubleBuffer<unsigned long long> dk(k0, k1);
    size_t tmp_bytes = 0;
    CCG_CUDA(ctx, cub::DeviceSegmentedSort::SortKeys(nullptr, tmp_bytes, dk, (int)total, (int)n, offsets_dev, offsets_dev + 1, ctx->stream));
    CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
    CCG_CUDA(ctx, cub::DeviceSegmentedSort::SortKeys(ctx->sort_tmp.p, tmp_bytes, dk, (int)total, (int)n, offsets_dev, offsets_dev + 1, ctx->stream));
    k_csr_unpack<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(dk.Current(), total, idx_dev, d2_dev);
    ctx->stats.launches += 4;
    CCG_CUDA(ctx, cudaGetLastError());
    return CCG_OK;
}

}  // namespace ccg
