// tile.cuh — device building blocks shared by the tile kernels (query.cu, radsel.cu): the contract distance arithmetic,
// the bottom-up tile traversal over the box hierarchy, and the fused covariance + eigensolve epilogue.
#pragma once
#include "common.cuh"
#include <cfloat>
#include <type_traits>

namespace ccg {

// ---------------------------------------------------------------------------------------------------
// distance primitives (contract: ((dx*dx + dy*dy) + dz*dz), round-to-nearest, never fused)
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ float sumsq3(float dx, float dy, float dz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
// Sum of a fixed-size 3-vector expression as Eigen <= 3.2 evaluates it (.dot(), .norm(), .sum() on Vector3f / Map<Vector3f>:
// 12 bytes are not packet-sized, so Redux.h's redux_novec_unroller<Func, Derived, Start, Length> splits Length = 3 into
// HalfLength = 1 and the rest: func(a0, func(a1, a2)) [EXT Eigen/src/Core/Redux.h].  Used wherever the reference reduces
// with Eigen (cv.h:181-182, cv.cpp:212-214, normalestimation.cpp:380-420); the SEARCH distance above is FLANN's
// sequential L2_Simple accumulation instead.
__device__ __forceinline__ float eigen_sum3(float a0, float a1, float a2) { return __fadd_rn(a0, __fadd_rn(a1, a2)); }
__device__ __forceinline__ float eigen_dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return eigen_sum3(__fmul_rn(ax, bx), __fmul_rn(ay, by), __fmul_rn(az, bz));
}
__device__ __forceinline__ float d2_point(const float4 &q, const float4 &p) {
    return sumsq3(__fsub_rn(q.x, p.x), __fsub_rn(q.y, p.y), __fsub_rn(q.z, p.z));
}
// lower bound of d2_point(q, p) over all p inside [lo, hi] (monotone in every rounding step)
__device__ __forceinline__ float d2_point_box(const float4 &q, const float4 &lo, const float4 &hi) {
    float dx = fmaxf(fmaxf(__fsub_rn(lo.x, q.x), __fsub_rn(q.x, hi.x)), 0.f);
    float dy = fmaxf(fmaxf(__fsub_rn(lo.y, q.y), __fsub_rn(q.y, hi.y)), 0.f);
    float dz = fmaxf(fmaxf(__fsub_rn(lo.z, q.z), __fsub_rn(q.z, hi.z)), 0.f);
    return sumsq3(dx, dy, dz);
}
// UPPER bound of d2_point(q, p) over all p inside [lo, hi]: |q.x - p.x| rounds to a value between the rounded differences to the
// two faces (rounding is monotone), and squares and sums of non-negative terms are monotone in every rounding step
__device__ __forceinline__ float d2_point_box_far(const float4 &q, const float4 &lo, const float4 &hi) {
    float dx = fmaxf(fabsf(__fsub_rn(q.x, lo.x)), fabsf(__fsub_rn(q.x, hi.x)));
    float dy = fmaxf(fabsf(__fsub_rn(q.y, lo.y)), fabsf(__fsub_rn(q.y, hi.y)));
    float dz = fmaxf(fabsf(__fsub_rn(q.z, lo.z)), fabsf(__fsub_rn(q.z, hi.z)));
    return sumsq3(dx, dy, dz);
}
// lower bound over all (q in [alo, ahi], p in [blo, bhi])
__device__ __forceinline__ float d2_box_box(const float4 &alo, const float4 &ahi, const float4 &blo, const float4 &bhi) {
    float dx = fmaxf(fmaxf(__fsub_rn(blo.x, ahi.x), __fsub_rn(alo.x, bhi.x)), 0.f);
    float dy = fmaxf(fmaxf(__fsub_rn(blo.y, ahi.y), __fsub_rn(alo.y, bhi.y)), 0.f);
    float dz = fmaxf(fmaxf(__fsub_rn(blo.z, ahi.z), __fsub_rn(alo.z, bhi.z)), 0.f);
    return sumsq3(dx, dy, dz);
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------------
// tile traversal
// ---------------------------------------------------------------------------------------------------
struct WarpScratch {
    float4 leaf[kLeaf];
};

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}

// A visitor with `static constexpr bool kNodes = true` takes a whole SUBTREE at once: when node_test() is on (it must be the same
// in every lane of the warp: the test votes with a full mask), every box chosen by
// the walk is also measured by its FARTHEST corner, and when whole(far distance) holds for every lane that wants the box, the
// lanes call add_node(node, lo, hi, q) instead of descending (node = index into the per-node arrays laid out like the boxes:
// level offset + node number).  Uncapped radius sums add the box's aggregated count and moments (one term per box inside the
// radius instead of a visit per point); the bracket passes of the capped search simply skip boxes that lie inside the
// bracket's inner edge.
template <class V, class = void> struct visits_nodes : std::false_type {};
template <class V> struct visits_nodes<V, std::void_t<decltype(V::kNodes)>> : std::bool_constant<V::kNodes> {};

// count and moments of a box about its own centre (build.cu: k_leaf_moments / k_level_moments), shifted to the query: with
// d = centre - q,  sum (p - q) = S + n d,  sum (p - q)(p - q)^T = M + S d^T + d S^T + n d d^T
__device__ __forceinline__ int add_box_moments(const float4 *__restrict__ mom, size_t node, const float4 &lo, const float4 &hi, const float4 &q,
                                               bool moments, float m[9]) {
    const float4 a = __ldg(mom + 3 * node);
    const int nn = __float_as_int(a.w);
    if (moments) {
        const float4 b = __ldg(mom + 3 * node + 1), c = __ldg(mom + 3 * node + 2);
        const float fn = (float)nn;
        const float dx = __fmul_rn(0.5f, __fadd_rn(lo.x, hi.x)) - q.x, dy = __fmul_rn(0.5f, __fadd_rn(lo.y, hi.y)) - q.y,
                    dz = __fmul_rn(0.5f, __fadd_rn(lo.z, hi.z)) - q.z;
        const float tx = fn * dx, ty = fn * dy, tz = fn * dz;
        m[0] += a.x + tx; m[1] += a.y + ty; m[2] += a.z + tz;
        m[3] += b.x + dx * (2.f * a.x + tx);
        m[4] += b.y + a.x * dy + dx * (a.y + ty);
        m[5] += b.z + a.x * dz + dx * (a.z + tz);
        m[6] += b.w + dy * (2.f * a.y + ty);
        m[7] += c.x + a.y * dz + dy * (a.z + tz);
        m[8] += c.y + dz * (2.f * a.z + tz);
    }
    return nn;
}

// Bottom-up, nearest-first walk of the box hierarchy for the 32 queries of one leaf ("tile").
//   * the tile's own leaf is scanned first (it alone fills a k <= 32 list and gives a finite bound);
//   * then, level by level, the siblings of the tile's ancestor are searched top-down.  The (uniform) stack of
//     {first child, surviving-children mask} lives in registers, entry d in lane d, read back with a shuffle;
//   * children are taken in order of the distance from the tile's centre to their box, so that the bound
//     tightens before the far ones are looked at; before every pop the survivors are re-tested against the
//     warp's largest current bound (box-box distance, one child per lane) and the chosen one against every
//     lane's own bound (point-box distance).
template <class V>
__device__ __forceinline__ int traverse_tile(const TreeView &T, int tile, const float4 &q, V &v, WarpScratch &ws) {
    const int lane = threadIdx.x & 31;
    const float4 tlo = T.lo[0][tile], thi = T.hi[0][tile];
    const float4 tc = make_float4(0.5f * (tlo.x + thi.x), 0.5f * (tlo.y + thi.y), 0.5f * (tlo.z + thi.z), 0.f);
    int scanned = 1;
    ws.leaf[lane] = q;                       // the own leaf is exactly the 32 query points of the warp
    __syncwarp();
    v.scan_leaf(q, ws.leaf);

    int node = tile;                         // ancestor (level L-1) whose subtree is finished
    for (int L = 1; L < T.nlevels; ++L) {
        const int parent = node >> 5;
        float rmax = warp_max(v.bound);
        if (rmax < 0.f) break;               // every lane is finished
        int st_base = 0;                     // lane d holds stack entry d: children at level (L-1-d)
        unsigned st_mask = 0;
        int sp = 0;
        bool mine = false;                   // registers describing the TOP entry: this lane's child survives,
        float dbb = 0.f, dc = 0.f;           // its box-box distance to the tile and its distance to the tile centre
        {
            const int c = (parent << 5) + lane;
            if (c < T.count[L - 1] && c != node) {
                const float4 clo = T.lo[L - 1][c], chi = T.hi[L - 1][c];
                dbb = d2_box_box(tlo, thi, clo, chi);
                dc = d2_point_box(tc, clo, chi);
                mine = dbb <= rmax;
            }
            const unsigned m = __ballot_sync(kFull, mine);
            if (m) { if (lane == 0) { st_base = parent << 5; st_mask = m; } sp = 1; }
        }
        bool top_valid = true;
        while (sp > 0) {
            const int lv = L - sp;
            const int base = __shfl_sync(kFull, st_base, sp - 1);
            unsigned m = __shfl_sync(kFull, st_mask, sp - 1);
            if (!top_valid) {                // back on an outer entry: reload this lane's child of it
                mine = (m >> lane) & 1u;
                if (mine) {
                    const float4 clo = T.lo[lv][base + lane], chi = T.hi[lv][base + lane];
                    dbb = d2_box_box(tlo, thi, clo, chi);
                    dc = d2_point_box(tc, clo, chi);
                }
                top_valid = true;
            }
            rmax = warp_max(v.bound);
            mine = mine && ((m >> lane) & 1u) && dbb <= rmax;
            m = __ballot_sync(kFull, mine);
            if (m == 0) { --sp; top_valid = false; continue; }
            const float best = warp_min(mine ? dc : INFINITY);
            const unsigned who = __ballot_sync(kFull, mine && dc == best);
            const int cbit = who ? __ffs(who) - 1 : __ffs(m) - 1;
            m &= ~(1u << cbit);
            if (lane == cbit) mine = false;
            if (lane == sp - 1) st_mask = m;
            const int child = base + cbit;
            // exact per-query test of the chosen child against each lane's own bound
            const float4 clo = T.lo[lv][child], chi = T.hi[lv][child];
            const bool want = d2_point_box(q, clo, chi) <= v.bound;
            if (!__any_sync(kFull, want)) continue;
            if constexpr (visits_nodes<V>::value) {
                if (v.node_test()) {
                    const bool whole = want && v.whole(d2_point_box_far(q, clo, chi));
                    if (__all_sync(kFull, whole || !want)) {
                        if (whole) v.add_node((size_t)(T.lo[lv] - T.lo[0]) + (size_t)child, clo, chi, q);
                        continue;
                    }
                }
            }
            if (lv == 0) {
                // stage the leaf (a contiguous 512-byte run) and let every lane score its 32 candidates
                const float4 p = T.pts[(size_t)child * kLeaf + lane];
                __syncwarp();
                ws.leaf[lane] = p;
                __syncwarp();
                v.scan_leaf(q, ws.leaf);
                ++scanned;
            } else {
                // expand: one lane per grandchild against the tile box
                const int first = child << 5;
                const int c = first + lane;
                bool pass = false;
                float ndbb = 0.f, ndc = 0.f;
                if (c < T.count[lv - 1]) {
                    const float4 glo = T.lo[lv - 1][c], ghi = T.hi[lv - 1][c];
                    ndbb = d2_box_box(tlo, thi, glo, ghi);
                    ndc = d2_point_box(tc, glo, ghi);
                    pass = ndbb <= rmax;
                }
                const unsigned m2 = __ballot_sync(kFull, pass);
                if (m2) {
                    if (lane == sp) { st_base = first; st_mask = m2; }
                    ++sp;
                    mine = pass; dbb = ndbb; dc = ndc;      // the registers now describe the new top entry
                } 
            }
        }
        node = parent;
    }
    return scanned;
}

// ---------------------------------------------------------------------------------------------------
// fused covariance + symmetric 3x3 eigensolve (fp32 cyclic Jacobi: high relative accuracy on PSD input)
// ---------------------------------------------------------------------------------------------------
struct Eig3 {
    float w[3];      // descending
    float n[3];      // unit eigenvector of w[2]
};

__device__ __forceinline__ void jacobi_rot(float &app, float &aqq, float &apq, float &arp, float &arq,
                                           float &v0p, float &v0q, float &v1p, float &v1q, float &v2p, float &v2q) {
    if (apq != 0.f) {
        float theta = (aqq - app) / (2.f * apq);
        float t = copysignf(1.f, theta) / (fabsf(theta) + sqrtf(fmaf(theta, theta, 1.f)));
        if (!isfinite(theta)) t = 0.f;
        float c = rsqrtf(fmaf(t, t, 1.f));
        float s = t * c;
        float tau = s / (1.f + c);
        app -= t * apq;
        aqq += t * apq;
        apq = 0.f;
        float rp = arp, rq = arq;
        arp = rp - s * (rq + tau * rp);
        arq = rq + s * (rp - tau * rq);
        float a, b;
        a = v0p; b = v0q; v0p = a - s * (b + tau * a); v0q = b + s * (a - tau * b);
        a = v1p; b = v1q; v1p = a - s * (b + tau * a); v1q = b + s * (a - tau * b);
        a = v2p; b = v2q; v2p = a - s * (b + tau * a); v2q = b + s * (a - tau * b);
    }
}

// a = {xx, xy, xz, yy, yz, zz}
__device__ __forceinline__ Eig3 eig3_sym(const float a_in[6]) {
    float a00 = a_in[0], a01 = a_in[1], a02 = a_in[2], a11 = a_in[3], a12 = a_in[4], a22 = a_in[5];
    float scale = fmaxf(fmaxf(fmaxf(fabsf(a00), fabsf(a11)), fabsf(a22)), fmaxf(fmaxf(fabsf(a01), fabsf(a02)), fabsf(a12)));
    float inv = scale > 0.f ? 1.f / scale : 0.f;
    a00 *= inv; a01 *= inv; a02 *= inv; a11 *= inv; a12 *= inv; a22 *= inv;
    float v00 = 1.f, v01 = 0.f, v02 = 0.f, v10 = 0.f, v11 = 1.f, v12 = 0.f, v20 = 0.f, v21 = 0.f, v22 = 1.f;
#pragma unroll 1
    for (int sweep = 0; sweep < 8; ++sweep) {
        float off = fabsf(a01) + fabsf(a02) + fabsf(a12);
        if (off <= 1e-10f) break;                     // entries are scaled to max |a_ij| = 1
        // (p,q,r) = (0,1,2): a_rp = a02, a_rq = a12
        jacobi_rot(a00, a11, a01, a02, a12, v00, v01, v10, v11, v20, v21);
        // (0,2,1): a_rp = a01, a_rq = a12
        jacobi_rot(a00, a22, a02, a01, a12, v00, v02, v10, v12, v20, v22);
        // (1,2,0): a_rp = a01, a_rq = a02
        jacobi_rot(a11, a22, a12, a01, a02, v01, v02, v11, v12, v21, v22);
    }
    Eig3 r;
    float w0 = a00 * scale, w1 = a11 * scale, w2 = a22 * scale;
    // column of the smallest eigenvalue, eigenvalues descending
    float nx, ny, nz, wmin;
    if (w0 <= w1 && w0 <= w2) { wmin = w0; nx = v00; ny = v10; nz = v20; r.w[0] = fmaxf(w1, w2); r.w[1] = fminf(w1, w2); }
    else if (w1 <= w2) { wmin = w1; nx = v01; ny = v11; nz = v21; r.w[0] = fmaxf(w0, w2); r.w[1] = fminf(w0, w2); }
    else { wmin = w2; nx = v02; ny = v12; nz = v22; r.w[0] = fmaxf(w0, w1); r.w[1] = fminf(w0, w1); }
    r.w[2] = wmin;
    float nn = rsqrtf(nx * nx + ny * ny + nz * nz);
    r.n[0] = nx * nn; r.n[1] = ny * nn; r.n[2] = nz * nn;
    return r;
}

constexpr int kMaxRoute = 16;
struct PcaOut {
    int cov_norm, min_pts, normalise_hist, has_vp;
    float vp[3];
    float *eig3, *normal4;
    float *feat7; int formula_set;     // optional fused a9 epilogue
    // owner-routed rows (multi-GPU, ccg_pca_knn_features_route): row r belongs to rank r / route_rows and is stored at local
    // row r % route_rows of THAT rank's arrays (own memory or NVLink peer memory); route_world <= 1 = off
    int route_world; unsigned route_rows;
    float *r_eig[kMaxRoute], *r_nrm[kMaxRoute], *r_feat[kMaxRoute];
};

// a9: the 7 eigen features of Markov::mkFeatureVector (markov.cpp:449-475) from descending (l0, l1, l2).
// formula_set 0 = exactly as written there (including its slips), 1 = textbook definitions.
// out of line on purpose: logf x3 + cbrtf are ~250 instructions that every inlining site would add to a kernel that is short of
// instruction-fetch bandwidth (k_tile_sel: 25.78 -> 25.29 ms, profiles/r02_sel_v16_codesize_variants.log)
static __device__ __noinline__ void features7(float l0, float l1, float l2, int formula_set, float *o) {
    o[0] = __fdiv_rn(__fsub_rn(l0, l2), l0);                 // markov.cpp:451
    o[1] = __fdiv_rn(__fsub_rn(l1, l2), l0);                 // :455
    o[2] = __fdiv_rn(l2, l0);                                // :459
    o[3] = __fdiv_rn(__fsub_rn(l0, l1), l0);                 // :463
    if (formula_set == 0) {
        o[4] = cbrtf(__fmul_rn(__fmul_rn(l0, l1), l0));      // :467 (sic)
        o[5] = -__fadd_rn(__fadd_rn(__fmul_rn(l0, logf(l0)), __fmul_rn(l1, logf(l1))), __fmul_rn(l2, logf(l0)));   // :471 (sic)
        o[6] = __fdiv_rn(l0, __fadd_rn(__fadd_rn(l0, l1), l2));     // :475 (sic)
    } else {
        o[4] = cbrtf(__fmul_rn(__fmul_rn(l0, l1), l2));
        o[5] = -__fadd_rn(__fadd_rn(__fmul_rn(l0, logf(l0)), __fmul_rn(l1, logf(l1))), __fmul_rn(l2, logf(l2)));
        o[6] = __fdiv_rn(l2, __fadd_rn(__fadd_rn(l0, l1), l2));
    }
}

// moments of (p - q): m = {sx sy sz sxx sxy sxz syy syz szz}, cnt points -> outputs for one row
#ifdef CCG_NOINLINE_FIN
#define CCG_FIN_INLINE __noinline__
#else
#define CCG_FIN_INLINE __forceinline__
#endif
__device__ CCG_FIN_INLINE void pca_finish(const PcaOut &o, const float4 &q, const float m[9], int cnt, size_t row) {
    float *eig3 = o.eig3, *normal4 = o.normal4, *feat7 = o.feat7;
    if (o.route_world > 1) {
        const unsigned owner = (unsigned)row / o.route_rows;
        row = (unsigned)row - owner * o.route_rows;
        eig3 = o.r_eig[owner]; normal4 = o.r_nrm[owner]; feat7 = o.r_feat[owner];
    }
    float e0 = 0.f, e1 = 0.f, e2 = 1.f;                       // cv.h:105-109 "isolated point" marker
    float nx = NAN, ny = NAN, nz = NAN, curv = NAN;
    if (cnt >= o.min_pts) {
        float invn = 1.f / (float)cnt;
        float mx = m[0] * invn, my = m[1] * invn, mz = m[2] * invn;
        float s[6];
        s[0] = m[3] - m[0] * mx; s[1] = m[4] - m[0] * my; s[2] = m[5] - m[0] * mz;
        s[3] = m[6] - m[1] * my; s[4] = m[7] - m[1] * mz; s[5] = m[8] - m[2] * mz;
        float sc = o.cov_norm == CCG_COV_NM1 ? 1.f / (float)(cnt - 1) : (o.cov_norm == CCG_COV_N ? invn : 1.f);
#pragma unroll
        for (int i = 0; i < 6; ++i) s[i] *= sc;
        Eig3 e = eig3_sym(s);
        e0 = e.w[0]; e1 = e.w[1]; e2 = e.w[2];
        float tr = e0 + e1 + e2;
        curv = tr > 0.f ? fabsf(e2 / tr) : 0.f;
        nx = e.n[0]; ny = e.n[1]; nz = e.n[2];
        if (o.has_vp) {
            float dot = (o.vp[0] - q.x) * nx + (o.vp[1] - q.y) * ny + (o.vp[2] - q.z) * nz;
            if (dot < 0.f) { nx = -nx; ny = -ny; nz = -nz; }
        }
        if (o.normalise_hist) {                                // cv.cpp:293
            float nn = sqrtf(e0 * e0 + e1 * e1 + e2 * e2);
            if (nn > 0.f) { e0 /= nn; e1 /= nn; e2 /= nn; }
        }
    }
    if (eig3) { float *p = eig3 + row * 3; p[0] = e0; p[1] = e1; p[2] = e2; }
    if (normal4) reinterpret_cast<float4 *>(normal4)[row] = make_float4(nx, ny, nz, curv);
    if (feat7) {
        float f[7];
        features7(e0, e1, e2, o.formula_set, f);
        float *p = feat7 + row * 7;
#pragma unroll
        for (int i = 0; i < 7; ++i) p[i] = f[i];
    }
}

}  // namespace ccg
