// build.cu — cloud ingest and the Morton-ordered bounding-box hierarchy.
//
// Replaces what pcl::KdTreeFLANN::setInputCloud (cv.h:83-84) and the async octree build
// (pointcloud.cpp:312-323) do for the reference: make the cloud searchable.  B200-first design: instead of a
// pointer-chasing tree we sort the points along a 63-bit Morton curve (device radix sort), cut the sorted
// array into 32-point leaves (one warp-wide float4 load each, always contiguous) and put a 32-ary hierarchy
// of axis-aligned boxes on top (one lane per child).  Unlike a uniform grid this adapts to the 1/r^2 density
// of a terrestrial scan with no tuning, and every leaf is a contiguous 512-byte run.
//
// All stages are HBM-streaming kernels; algorithmic bytes per point are listed in DESIGN.md §4.
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_partition.cuh>
#include <cfloat>

namespace ccg {

// order-preserving float <-> uint map for atomicMin/Max
__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(unsigned u) {
    unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}

struct IngestScratch {           // lives in ctx->scratch
    unsigned lo[3], hi[3];
    unsigned nonfinite;
    unsigned pad;
};

// AoS (x,y,z first, arbitrary stride) -> float4 original-order array + bounding box + finiteness check.
// Each thread handles one point; a 16-byte aligned stride lets one LDG.128 fetch xyz(+pad).
template <bool VEC16>
__global__ void __launch_bounds__(256) k_ingest(const char *__restrict__ src, size_t stride, int n,
                                                float4 *__restrict__ orig, IngestScratch *sc) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    unsigned bad = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float x, y, z;
        if (VEC16) {
            float4 v = __ldg(reinterpret_cast<const float4 *>(src + (size_t)i * stride));
            x = v.x; y = v.y; z = v.z;
        } else {
            const float *p = reinterpret_cast<const float *>(src + (size_t)i * stride);
            x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
        }
        orig[i] = make_float4(x, y, z, 0.f);
        bool fin = isfinite(x) && isfinite(y) && isfinite(z);
        bad |= fin ? 0u : 1u;
        if (fin) {
            lo[0] = fminf(lo[0], x); lo[1] = fminf(lo[1], y); lo[2] = fminf(lo[2], z);
            hi[0] = fmaxf(hi[0], x); hi[1] = fmaxf(hi[1], y); hi[2] = fmaxf(hi[2], z);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            lo[d] = fminf(lo[d], __shfl_xor_sync(kFull, lo[d], o));
            hi[d] = fmaxf(hi[d], __shfl_xor_sync(kFull, hi[d], o));
        }
        bad |= __shfl_xor_sync(kFull, bad, o);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            atomicMin(&sc->lo[d], f2ord(lo[d]));
            atomicMax(&sc->hi[d], f2ord(hi[d]));
        }
        if (bad) atomicOr(&sc->nonfinite, 1u);
    }
}

constexpr int kMortonBits = 21;     // most bits per axis a curve key can hold (63-bit key); ctx->key_bits of them are used

__device__ __forceinline__ unsigned long long spread21(unsigned v) {
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// 3-D Hilbert index of the quantised coordinates.  Unlike the Morton curve the Hilbert curve has no jumps: consecutive
// points along it are always spatial neighbours, so a run of 32 consecutive points (a leaf) is a compact blob instead of,
// every so often, two pieces of surface a Z-jump apart.  Measured on the 30 M bench scan: candidate leaves per tile
// 29.2 -> 17.7, query 38.3 -> 29.7 ms (profiles/r02_sel_v14_variants.log).
// The curve is the one of Skilling's transpose algorithm ("Programming the Hilbert curve", 2004), evaluated by its state
// machine: 24 orientations x 8 octants -> (3 output bits, next orientation), one table step per level on the octants of
// the Morton code.  The table was generated from the textbook loop and checked against it (tools/hilbert_table.py);
// the branchy loop itself cost 0.5 ms per 30 M points (lanes of a warp sit in different octants at every level).
__device__ const unsigned char kHilbertTab[192] = {
    8, 17, 27, 2, 39, 46, 52, 5, 56, 71, 73, 86, 91, 20, 10, 13,
    48, 1, 103, 110, 115, 18, 12, 21, 126, 129, 29, 26, 79, 80, 140, 3,
    148, 43, 37, 34, 127, 128, 78, 81, 156, 45, 35, 42, 31, 6, 160, 105,
    72, 87, 139, 4, 57, 70, 50, 53, 0, 171, 111, 76, 49, 58, 102, 61,
    180, 143, 83, 184, 69, 54, 66, 97, 16, 123, 9, 74, 47, 60, 38, 77,
    132, 95, 85, 14, 67, 144, 82, 33, 142, 55, 185, 96, 93, 116, 90, 11,
    188, 107, 175, 176, 101, 98, 62, 65, 164, 109, 119, 22, 99, 106, 152, 41,
    174, 177, 63, 64, 117, 114, 92, 19, 30, 125, 161, 122, 7, 172, 104, 75,
    130, 25, 133, 166, 179, 136, 84, 191, 94, 15, 141, 28, 145, 32, 138, 51,
    146, 155, 149, 36, 137, 24, 190, 167, 154, 157, 147, 44, 169, 182, 120, 135,
    162, 165, 121, 134, 187, 108, 168, 183, 118, 173, 23, 124, 153, 170, 40, 59,
    178, 113, 131, 88, 181, 158, 68, 151, 186, 163, 89, 112, 189, 100, 150, 159,
};
__device__ __forceinline__ unsigned long long hilbert_key(unsigned long long morton, int bits, const unsigned char *tab) {
    unsigned long long key = 0;
    unsigned st = 0;
#pragma unroll 4
    for (int l = bits - 1; l >= 0; --l) {
        const unsigned o = (unsigned)(morton >> (3 * l)) & 7u;            // octant at this level: x y z bits
        const unsigned e = tab[st * 8 + o];
        key = (key << 3) | (e & 7u);
        st = e >> 3;
    }
    return key;
}

// Space-filling-curve key of every point (`bits` per axis, cubic cells over the bounding box).  Points sharing a cell keep
// their original (scan) order — the order is only a permutation, search results do not depend on it — but the QUALITY of
// the leaves does: a cell holding hundreds of points is cut into thin scan-order stripes.  16 bits per axis (1.5 mm cells
// on a 100 m scan, 6 radix passes of (key, index) pairs) is the default; 13 bits fit key and index of a 33 M-point cloud in
// one 64-bit word (5 passes of 8 bytes: build 3.48 -> 2.86 ms at 30 M) but cost the query what the build saves at 30 M
// (27.6 -> 28.3 ms) and 35 % at 100 M, and 12 bits there made it ten times slower (profiles/r02_keybits.log).
// The bounding box is read from the ingest scratch ON THE DEVICE (no host round trip in the middle of the build).
// The point's index rides in the low `ib` bits of the same 64-bit word (key << ib | i), so the sort moves 8 bytes per
// point and pass instead of a 12-byte (key, value) pair, and only the key bits [ib, ib + kb) are sorted.
// vals == NULL: packed words; vals != NULL: (key, index) pairs — the path of clouds whose index bits leave no room for
// enough key bits in one word (see build_cloud).
template <bool HILBERT>
__global__ void __launch_bounds__(256) k_morton(const float4 *__restrict__ orig, int n, const IngestScratch *__restrict__ sc,
                                                int ib, int bits, unsigned long long *__restrict__ keys, unsigned *__restrict__ vals) {
    __shared__ unsigned char s_tab[192];
    if (HILBERT) {
        if (threadIdx.x < 192) s_tab[threadIdx.x] = kHilbertTab[threadIdx.x];
        __syncthreads();
    }
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float lx = ord2f(sc->lo[0]), ly = ord2f(sc->lo[1]), lz = ord2f(sc->lo[2]);
    const float ext = fmaxf(fmaxf(ord2f(sc->hi[0]) - lx, ord2f(sc->hi[1]) - ly), ord2f(sc->hi[2]) - lz);
    const float inv_cell = ext > 0.f ? __fdiv_rn((float)(1u << bits), __fmul_rn(ext, 1.0001f)) : 0.f;
    float4 p = orig[i];
    const float qmax = (float)((1u << bits) - 1u);
    unsigned qx = (unsigned)fminf(fmaxf((p.x - lx) * inv_cell, 0.f), qmax);
    unsigned qy = (unsigned)fminf(fmaxf((p.y - ly) * inv_cell, 0.f), qmax);
    unsigned qz = (unsigned)fminf(fmaxf((p.z - lz) * inv_cell, 0.f), qmax);
    unsigned long long key = spread21(qz) | (spread21(qy) << 1) | (spread21(qx) << 2);
    if (HILBERT) key = hilbert_key(key, bits, s_tab);
    if (vals) { keys[i] = key; vals[i] = (unsigned)i; }
    else keys[i] = (key << ib) | (unsigned long long)(unsigned)i;
}


// Isolated points (laser "mixed pixels", airborne returns: the very points the outlier filters exist for) sit between
// surface points in the Morton order and stretch the box of whatever 32-point leaf they fall into from centimetres to
// metres; a few thousand such boxes make every wide search intersect all of them (r01: 35 k instructions per deferred
// query).  They are therefore moved to the END of the order, where they form leaves of their own — large boxes, but few
// and hardly overlapping — and the remaining leaves stay tight.  Any classification is correct (the order is only a
// permutation); this one calls a point isolated when the distance to its nearer neighbour along the curve is more than
// kIsoFactor times the median of that distance over its aligned 32-point window.  One warp per window.
constexpr float kIsoFactor = 8.f;
// Also unpacks the sorted words: order32[i] = index bits of packed[i].
// packed == NULL: order32 already holds the sorted indices (pairs path).
__global__ void __launch_bounds__(256) k_flag_dense(const float4 *__restrict__ orig, const unsigned long long *__restrict__ packed,
                                                    unsigned long long idx_mask, int n, unsigned *__restrict__ order32,
                                                    unsigned char *__restrict__ dense) {
    const int win = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int i = win * 32 + lane;
    if (win * 32 >= n) return;
    const bool valid = i < n;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
        unsigned o;
        if (packed) { o = (unsigned)(packed[i] & idx_mask); order32[i] = o; }
        else o = order32[i];
        p = orig[o];
    }
    float px = __shfl_up_sync(kFull, p.x, 1), py = __shfl_up_sync(kFull, p.y, 1), pz = __shfl_up_sync(kFull, p.z, 1);
    float nx = __shfl_down_sync(kFull, p.x, 1), ny = __shfl_down_sync(kFull, p.y, 1), nz = __shfl_down_sync(kFull, p.z, 1);
    bool has_prev = valid && i > 0, has_next = valid && i + 1 < n;
    if (lane == 0 && has_prev) { const float4 t = orig[packed ? (unsigned)(packed[i - 1] & idx_mask) : order32[i - 1]]; px = t.x; py = t.y; pz = t.z; }
    if (lane == 31 && has_next) { const float4 t = orig[packed ? (unsigned)(packed[i + 1] & idx_mask) : order32[i + 1]]; nx = t.x; ny = t.y; nz = t.z; }
    float gp = INFINITY, gn = INFINITY;
    if (has_prev) { const float x = p.x - px, y = p.y - py, z = p.z - pz; gp = x * x + y * y + z * z; }
    if (has_next) { const float x = p.x - nx, y = p.y - ny, z = p.z - nz; gn = x * x + y * y + z * z; }
    const float iso = valid ? fminf(gp, gn) : INFINITY;      // squared distance to the nearer curve neighbour
    // bitonic sort of the 32 values across the warp -> median of the valid ones
    float v = iso;
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const float o = __shfl_xor_sync(kFull, v, j);
            const bool up = ((lane & k) == 0) == ((lane & j) == 0);
            v = up ? fminf(v, o) : fmaxf(v, o);
        }
    }
    const int nvalid = min(32, n - win * 32);
    const float med = __shfl_sync(kFull, v, nvalid >> 1);
    if (valid) dense[i] = (iso > kIsoFactor * kIsoFactor * med) ? 0 : 1;
}

// One warp per leaf: gather the 32 points of the leaf in curve order, tag them with their original index, reduce the
// leaf's bounding box, and write the leaf's PREFILTER BLOCK (kRelStride float4 = 528 B, see query.cu):
//   [0]      = (cx, cy, cz, 0)                    box centre c = 0.5 * (lo + hi)
//   [1 + 2p] = (X(2p), X(2p+1), Y(2p), Y(2p+1))   P = p - c of candidates 2p and 2p+1, pair-interleaved so that one
//   [2 + 2p] = (Z(2p), Z(2p+1), W(2p), W(2p+1))   packed FFMA2 scores two candidates; W = -|P|^2
// A candidate leaf is then staged into shared memory by a plain copy (one cp.async.bulk), with no per-tile arithmetic.
// Pad slots of the last leaf are +inf points with index INT_MAX (their block entries are inf / -inf: never a neighbour).
__global__ void __launch_bounds__(256) k_reorder_leaves(const float4 *__restrict__ orig, const unsigned *__restrict__ order,
                                                        int n, int nleaf, float4 *__restrict__ sorted, float4 *__restrict__ rel,
                                                        float4 *__restrict__ lo0, float4 *__restrict__ hi0) {
    int leaf = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (leaf >= nleaf) return;
    int pos = leaf * kLeaf + lane;
    float4 p = make_float4(INFINITY, INFINITY, INFINITY, __int_as_float(0x7fffffff));
    float lx = INFINITY, ly = INFINITY, lz = INFINITY, hx = -INFINITY, hy = -INFINITY, hz = -INFINITY;
    if (pos < n) {
        unsigned o = order[pos];
        float4 v = orig[o];
        p = make_float4(v.x, v.y, v.z, __int_as_float((int)o));
        lx = hx = v.x; ly = hy = v.y; lz = hz = v.z;
    }
    sorted[pos] = p;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lx = fminf(lx, __shfl_xor_sync(kFull, lx, o)); ly = fminf(ly, __shfl_xor_sync(kFull, ly, o));
        lz = fminf(lz, __shfl_xor_sync(kFull, lz, o)); hx = fmaxf(hx, __shfl_xor_sync(kFull, hx, o));
        hy = fmaxf(hy, __shfl_xor_sync(kFull, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(kFull, hz, o));
    }
    // lo.w = smallest original index inside the leaf (int bits; INT_MAX pads never win): lets the exact point kernel skip a
    // box that can only hold candidates TIED with its current k-th distance but with larger indices (duplicate piles)
    const int mi = __reduce_min_sync(kFull, __float_as_int(p.w));
    if (lane == 0) {
        lo0[leaf] = make_float4(lx, ly, lz, __int_as_float(mi));
        hi0[leaf] = make_float4(hx, hy, hz, 0.f);
    }
    const float cx = __fmul_rn(0.5f, __fadd_rn(lx, hx)), cy = __fmul_rn(0.5f, __fadd_rn(ly, hy)), cz = __fmul_rn(0.5f, __fadd_rn(lz, hz));
    const float X = p.x - cx, Y = p.y - cy, Z = p.z - cz;
    const float W = -fmaf(X, X, fmaf(Y, Y, Z * Z));
    // even lane 2p writes [1 + 2p] = (X, X', Y, Y'), odd lane 2p+1 writes [2 + 2p] = (Z', Z, W', W) with ' = the partner's value
    const float oX = __shfl_xor_sync(kFull, X, 1), oY = __shfl_xor_sync(kFull, Y, 1), oZ = __shfl_xor_sync(kFull, Z, 1),
                oW = __shfl_xor_sync(kFull, W, 1);
    float4 *blk = rel + (size_t)leaf * kRelStride;
    blk[1 + lane] = (lane & 1) ? make_float4(oZ, Z, oW, W) : make_float4(X, oX, Y, oY);
    if (lane == 0) blk[0] = make_float4(cx, cy, cz, 0.f);
}

// One warp per parent: lane c reduces child c's box.
__global__ void __launch_bounds__(256) k_build_level(const float4 *__restrict__ clo, const float4 *__restrict__ chi,
                                                     int nchild, float4 *__restrict__ plo, float4 *__restrict__ phi,
                                                     int nparent) {
    int node = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (node >= nparent) return;
    int c = node * kFan + lane;
    float lx = INFINITY, ly = INFINITY, lz = INFINITY, hx = -INFINITY, hy = -INFINITY, hz = -INFINITY;
    int mi = 0x7fffffff;
    if (c < nchild) {
        float4 a = clo[c], b = chi[c];
        lx = a.x; ly = a.y; lz = a.z; hx = b.x; hy = b.y; hz = b.z;
        mi = __float_as_int(a.w);
    }
    mi = __reduce_min_sync(kFull, mi);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lx = fminf(lx, __shfl_xor_sync(kFull, lx, o)); ly = fminf(ly, __shfl_xor_sync(kFull, ly, o));
        lz = fminf(lz, __shfl_xor_sync(kFull, lz, o)); hx = fmaxf(hx, __shfl_xor_sync(kFull, hx, o));
        hy = fmaxf(hy, __shfl_xor_sync(kFull, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(kFull, hz, o));
    }
    if (lane == 0) {
        plo[node] = make_float4(lx, ly, lz, __int_as_float(mi));       // smallest original index below this node
        phi[node] = make_float4(hx, hy, hz, 0.f);
    }
}

// ---------------------------------------------------------------------------------------------------
// per-box moments (uncapped radius sums: a box inside the radius contributes ONE aggregated term)
//   mom[3 b + 0] = (Sx, Sy, Sz, count as int bits)   S = sum (p - c), c = the box centre 0.5 (lo + hi)
//   mom[3 b + 1] = (Mxx, Mxy, Mxz, Myy)              M = sum (p - c)(p - c)^T
//   mom[3 b + 2] = (Myz, Mzz, 0, 0)
// Sums about the box's OWN centre are of the size of the box, so shifting them to a query (tile.cuh / query.cu: StreamNodes)
// loses nothing to cancellation whatever the cloud's coordinates are.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

__global__ void __launch_bounds__(256) k_leaf_moments(const float4 *__restrict__ sorted, int n, int nleaf, const float4 *__restrict__ lo0,
                                                      const float4 *__restrict__ hi0, float4 *__restrict__ mom) {
    const int leaf = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (leaf >= nleaf) return;
    const int pos = leaf * kLeaf + lane;
    const bool valid = pos < n;
    const float4 lo = lo0[leaf], hi = hi0[leaf];
    const float cx = __fmul_rn(0.5f, __fadd_rn(lo.x, hi.x)), cy = __fmul_rn(0.5f, __fadd_rn(lo.y, hi.y)), cz = __fmul_rn(0.5f, __fadd_rn(lo.z, hi.z));
    const float4 p = sorted[pos];
    const float X = valid ? p.x - cx : 0.f, Y = valid ? p.y - cy : 0.f, Z = valid ? p.z - cz : 0.f;
    const float sx = warp_sum(X), sy = warp_sum(Y), sz = warp_sum(Z);
    const float xx = warp_sum(X * X), xy = warp_sum(X * Y), xz = warp_sum(X * Z), yy = warp_sum(Y * Y), yz = warp_sum(Y * Z), zz = warp_sum(Z * Z);
    const int cnt = __popc(__ballot_sync(kFull, valid));
    if (lane == 0) {
        float4 *o = mom + 3 * (size_t)leaf;
        o[0] = make_float4(sx, sy, sz, __int_as_float(cnt));
        o[1] = make_float4(xx, xy, xz, yy);
        o[2] = make_float4(yz, zz, 0.f, 0.f);
    }
}

// one warp per parent: lane c shifts child c's moments to the parent's centre, the warp adds them up
__global__ void __launch_bounds__(256) k_level_moments(const float4 *__restrict__ clo, const float4 *__restrict__ chi, const float4 *__restrict__ cmom,
                                                       int nchild, const float4 *__restrict__ plo, const float4 *__restrict__ phi,
                                                       float4 *__restrict__ pmom, int nparent) {
    const int node = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (node >= nparent) return;
    const int c = node * kFan + lane;
    const float4 PL = plo[node], PH = phi[node];
    const float Cx = __fmul_rn(0.5f, __fadd_rn(PL.x, PH.x)), Cy = __fmul_rn(0.5f, __fadd_rn(PL.y, PH.y)), Cz = __fmul_rn(0.5f, __fadd_rn(PL.z, PH.z));
    float v[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int cnt = 0;
    if (c < nchild) {
        const float4 lo = clo[c], hi = chi[c];
        const float4 a = cmom[3 * (size_t)c], b = cmom[3 * (size_t)c + 1], d = cmom[3 * (size_t)c + 2];
        cnt = __float_as_int(a.w);
        const float fn = (float)cnt;
        const float ex = __fmul_rn(0.5f, __fadd_rn(lo.x, hi.x)) - Cx, ey = __fmul_rn(0.5f, __fadd_rn(lo.y, hi.y)) - Cy,
                    ez = __fmul_rn(0.5f, __fadd_rn(lo.z, hi.z)) - Cz;
        const float tx = fn * ex, ty = fn * ey, tz = fn * ez;
        v[0] = a.x + tx; v[1] = a.y + ty; v[2] = a.z + tz;
        v[3] = b.x + ex * (2.f * a.x + tx);
        v[4] = b.y + a.x * ey + ex * (a.y + ty);
        v[5] = b.z + a.x * ez + ex * (a.z + tz);
        v[6] = b.w + ey * (2.f * a.y + ty);
        v[7] = d.x + a.y * ez + ey * (a.z + tz);
        v[8] = d.y + ez * (2.f * a.z + tz);
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) v[i] = warp_sum(v[i]);
    cnt = __reduce_add_sync(kFull, cnt);
    if (lane == 0) {
        float4 *o = pmom + 3 * (size_t)node;
        o[0] = make_float4(v[0], v[1], v[2], __int_as_float(cnt));
        o[1] = make_float4(v[3], v[4], v[5], v[6]);
        o[2] = make_float4(v[7], v[8], 0.f, 0.f);
    }
}

// Filled lazily (the kNN paths never need it): two tiny passes over the sorted cloud and the boxes, on the ctx stream.
int ensure_node_moments(ccg_ctx *ctx) {
    if (ctx->node_mom_valid || ctx->n == 0) return CCG_OK;
    size_t total = 0;
    for (int l = 0; l < ctx->nlevels; ++l) total = ctx->level_off[l] + (size_t)ctx->level_count[l];
    CCG_CUDA(ctx, ctx->node_mom.reserve(3 * total * sizeof(float4)));
    cudaStream_t st = ctx->stream;
    const float4 *lo = ctx->boxes_lo.as<float4>(), *hi = ctx->boxes_hi.as<float4>();
    float4 *mom = ctx->node_mom.as<float4>();
    k_leaf_moments<<<(ctx->nleaf + 7) / 8, 256, 0, st>>>(ctx->sorted.as<float4>(), (int)ctx->n, ctx->nleaf, lo + ctx->level_off[0],
                                                         hi + ctx->level_off[0], mom + 3 * ctx->level_off[0]);
    ctx->stats.launches++;
    for (int l = 1; l < ctx->nlevels; ++l) {
        const int np = ctx->level_count[l];
        k_level_moments<<<(np + 7) / 8, 256, 0, st>>>(lo + ctx->level_off[l - 1], hi + ctx->level_off[l - 1], mom + 3 * ctx->level_off[l - 1],
                                                      ctx->level_count[l - 1], lo + ctx->level_off[l], hi + ctx->level_off[l],
                                                      mom + 3 * ctx->level_off[l], np);
        ctx->stats.launches++;
    }
    CCG_CUDA(ctx, cudaGetLastError());
    ctx->node_mom_valid = true;
    return CCG_OK;
}

// ccg_sync after an asynchronous build: the deferred finiteness check, bounding box and timings
int finish_cloud_check(ccg_ctx *ctx) {
    if (!ctx->pending_cloud_check) return CCG_OK;
    ctx->pending_cloud_check = 0;
    const IngestScratch *res = ctx->hpin[1].as<IngestScratch>();
    cudaEventElapsedTime(&ctx->stats.ms_upload, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&ctx->stats.ms_build, ctx->ev[1], ctx->ev[2]);
    if (res->nonfinite) {
        ctx->has_cloud = false;
        return set_error(ctx, CCG_ERR_NONFINITE, "cloud holds non-finite coordinates");
    }
    for (int d = 0; d < 3; ++d) { ctx->bb_lo[d] = ord2f(res->lo[d]); ctx->bb_hi[d] = ord2f(res->hi[d]); }
    return CCG_OK;
}

static int is_device_ptr(const void *p, int device) {
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) { cudaGetLastError(); return 0; }
    return (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) && (at.device == device || at.type == cudaMemoryTypeManaged);
}

int build_cloud(ccg_ctx *ctx, const void *xyz, size_t n, size_t stride) {
    cudaStream_t st = ctx->stream;
    ctx->n = 0;
    ctx->has_cloud = false;
    ctx->node_mom_valid = false;
    ctx->stats = ccg_stats{};
    if (n == 0) {
        ctx->npad = 0; ctx->nleaf = 0; ctx->nlevels = 0;
        ctx->has_cloud = true;
        return CCG_OK;
    }
    if (n > (size_t)0x7fffffe0) return set_error(ctx, CCG_ERR_UNSUPPORTED, "clouds above 2^31-32 points are not supported (int indices)");
    if (stride < 12 || (stride & 3)) return set_error(ctx, CCG_ERR_INVALID, "stride_bytes must be >= 12 and a multiple of 4 (got %zu)", stride);
    if (((uintptr_t)xyz) & 3) return set_error(ctx, CCG_ERR_INVALID, "xyz must be 4-byte aligned");

    const int nleaf = (int)((n + kLeaf - 1) / kLeaf);
    const size_t npad = (size_t)nleaf * kLeaf;
    // hierarchy geometry
    int nlevels = 1;
    int counts[kMaxLevels];
    counts[0] = nleaf;
    while (counts[nlevels - 1] > 1) {
        if (nlevels == kMaxLevels) return set_error(ctx, CCG_ERR_UNSUPPORTED, "hierarchy too deep");
        counts[nlevels] = (counts[nlevels - 1] + kFan - 1) / kFan;
        ++nlevels;
    }
    size_t off = 0;
    for (int l = 0; l < kMaxLevels; ++l) {
        ctx->level_off[l] = off;
        ctx->level_count[l] = l < nlevels ? counts[l] : 0;
        if (l < nlevels) off += ((size_t)counts[l] + 31) / 32 * 32;
    }
    CCG_CUDA(ctx, ctx->boxes_lo.reserve(off * sizeof(float4)));
    CCG_CUDA(ctx, ctx->boxes_hi.reserve(off * sizeof(float4)));
    CCG_CUDA(ctx, ctx->orig.reserve(n * sizeof(float4)));
    CCG_CUDA(ctx, ctx->sorted.reserve(npad * sizeof(float4)));
    CCG_CUDA(ctx, ctx->rel.reserve((size_t)nleaf * kRelStride * sizeof(float4)));
    for (int b = 0; b < 2; ++b) {
        CCG_CUDA(ctx, ctx->keys[b].reserve(n * sizeof(unsigned long long)));
        CCG_CUDA(ctx, ctx->vals[b].reserve(n * sizeof(unsigned)));
    }
    CCG_CUDA(ctx, ctx->scratch.reserve(4096));

    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
    // ---- stage the input on the device -----------------------------------------------------------
    const char *src_dev;
    if (is_device_ptr(xyz, ctx->device)) {
        src_dev = (const char *)xyz;
    } else {
        size_t bytes = (n - 1) * stride + 12;
        CCG_CUDA(ctx, ctx->raw.reserve(bytes + 16));
        CCG_CUDA(ctx, cudaMemcpyAsync(ctx->raw.p, xyz, bytes, cudaMemcpyHostToDevice, st));
        src_dev = (const char *)ctx->raw.p;
    }
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));

    // ---- ingest: float4 copy + bbox + finiteness ---------------------------------------------------
    IngestScratch init;
    for (int d = 0; d < 3; ++d) { init.lo[d] = 0xffffffffu; init.hi[d] = 0u; }
    init.nonfinite = 0; init.pad = 0;
    CCG_CUDA(ctx, cudaMemcpyAsync(ctx->scratch.p, &init, sizeof(init), cudaMemcpyHostToDevice, st));
    {
        int blocks = (int)((n + 255) / 256);
        int maxb = ctx->sm_count * 8;
        if (blocks > maxb) blocks = maxb;
        bool vec = (stride % 16 == 0) && (((uintptr_t)src_dev) % 16 == 0);
        // a 16-byte load of the last point must stay inside the staged buffer: true for device AoS of full
        // structs and for our own staging buffer (allocated with slack)
        if (vec) k_ingest<true><<<blocks, 256, 0, st>>>(src_dev, stride, (int)n, ctx->orig.as<float4>(), ctx->scratch.as<IngestScratch>());
        else k_ingest<false><<<blocks, 256, 0, st>>>(src_dev, stride, (int)n, ctx->orig.as<float4>(), ctx->scratch.as<IngestScratch>());
        ctx->stats.launches_build++;
    }
    // ---- curve keys + radix sort ------------------------------------------------------------------------
    // key_bits per axis of the curve index are sorted.  When key and index fit one 64-bit word (3 * key_bits + index bits
    // <= 64) the sort moves packed words (8 B per point and pass, keys only); otherwise (key, index) pairs (12 B).
    // The resolution matters where the scan is dense: 13 bits over a 100 m box are 1.2 cm cells, which hold hundreds of
    // points near the scanner, in scan order — leaves inside such a cell are thin stripes.  Measured: the 100 M-point
    // scan with 12-bit packed keys ran the query 12x slower than with 16 bits (profiles/r02_bench_n8_cfg4_*.json).
    int ib = 1;
    while (ib < 32 && ((n - 1) >> ib)) ++ib;                       // bits of the largest index
    int bits = ctx->key_bits < 4 ? 4 : (ctx->key_bits > kMortonBits ? kMortonBits : ctx->key_bits);
    const int kb = 3 * bits;
    const bool packed_sort = kb + ib <= 64;
    {
        int blocks = (int)((n + 255) / 256);
        unsigned *v0 = packed_sort ? nullptr : ctx->vals[0].as<unsigned>();
        if (ctx->curve_hilbert)
            k_morton<true><<<blocks, 256, 0, st>>>(ctx->orig.as<float4>(), (int)n, ctx->scratch.as<IngestScratch>(), ib, bits,
                                                   ctx->keys[0].as<unsigned long long>(), v0);
        else
            k_morton<false><<<blocks, 256, 0, st>>>(ctx->orig.as<float4>(), (int)n, ctx->scratch.as<IngestScratch>(), ib, bits,
                                                    ctx->keys[0].as<unsigned long long>(), v0);
        ctx->stats.launches_build++;
    }
    int ksel = 0, vsel = 0;
    {
        cub::DoubleBuffer<unsigned long long> dk(ctx->keys[0].as<unsigned long long>(), ctx->keys[1].as<unsigned long long>());
        cub::DoubleBuffer<unsigned> dv(ctx->vals[0].as<unsigned>(), ctx->vals[1].as<unsigned>());
        size_t tmp_bytes = 0;
        if (packed_sort) {
            CCG_CUDA(ctx, cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, dk, (int)n, ib, ib + kb, st));
            CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
            CCG_CUDA(ctx, cub::DeviceRadixSort::SortKeys(ctx->sort_tmp.p, tmp_bytes, dk, (int)n, ib, ib + kb, st));
        } else {
            CCG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, kb, st));
            CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
            CCG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tmp_bytes, dk, dv, (int)n, 0, kb, st));
            vsel = dv.selector;
        }
        ksel = dk.selector;
        ctx->stats.launches_build += 2 + (kb + 7) / 8;            // onesweep: histogram + scan + one pass per 8-bit digit
    }
    // ---- the order (unpacked from the sorted words, or the sorted values); isolated points to the end of it --------
    {
        const unsigned long long *packed = packed_sort ? ctx->keys[ksel].as<unsigned long long>() : nullptr;
        const unsigned long long mask = (1ull << ib) - 1ull;
        unsigned *ord = ctx->vals[packed_sort ? 0 : vsel].as<unsigned>();
        unsigned *ord2 = ctx->vals[packed_sort ? 1 : (vsel ^ 1)].as<unsigned>();
        unsigned char *flags = ctx->keys[ksel ^ 1].as<unsigned char>();    // the other key buffer is free now
        int *nsel = (int *)((char *)ctx->scratch.p + 2048);
        const int nwin = (int)((n + 31) / 32);
        k_flag_dense<<<(nwin + 7) / 8, 256, 0, st>>>(ctx->orig.as<float4>(), packed, mask, (int)n, ord, flags);
        ctx->stats.launches_build++;
        ctx->order_sel = packed_sort ? 0 : vsel;
        if (ctx->split_sparse && n > 64) {
            size_t tmp_bytes = 0;
            CCG_CUDA(ctx, cub::DevicePartition::Flagged(nullptr, tmp_bytes, ord, flags, ord2, nsel, (int)n, st));
            CCG_CUDA(ctx, ctx->sort_tmp.reserve(tmp_bytes));
            CCG_CUDA(ctx, cub::DevicePartition::Flagged(ctx->sort_tmp.p, tmp_bytes, ord, flags, ord2, nsel, (int)n, st));
            ctx->order_sel ^= 1;
            ctx->stats.launches_build += 2;
        }
    }
    // ---- leaves + box levels ----------------------------------------------------------------------------
    {
        float4 *lo = ctx->boxes_lo.as<float4>(), *hi = ctx->boxes_hi.as<float4>();
        int blocks = (nleaf + 7) / 8;
        k_reorder_leaves<<<blocks, 256, 0, st>>>(ctx->orig.as<float4>(), ctx->vals[ctx->order_sel].as<unsigned>(), (int)n, nleaf,
                                                 ctx->sorted.as<float4>(), ctx->rel.as<float4>(), lo + ctx->level_off[0], hi + ctx->level_off[0]);
        ctx->stats.launches_build++;
        for (int l = 1; l < nlevels; ++l) {
            int np = counts[l];
            k_build_level<<<(np + 7) / 8, 256, 0, st>>>(lo + ctx->level_off[l - 1], hi + ctx->level_off[l - 1], counts[l - 1],
                                                        lo + ctx->level_off[l], hi + ctx->level_off[l], np);
            ctx->stats.launches_build++;
        }
    }
    // the bounding box and the finiteness flag come back with the ONE synchronisation at the end of the build
    CCG_CUDA(ctx, ctx->hpin[1].reserve(256));
    IngestScratch *res = ctx->hpin[1].as<IngestScratch>();
    CCG_CUDA(ctx, cudaMemcpyAsync(res, ctx->scratch.p, sizeof(IngestScratch), cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[2], st));
    if (ctx->async_mode) {
        ctx->pending_cloud_check = 1;                 // ccg_sync reads the finiteness word and the bounding box
    } else {
        CCG_CUDA(ctx, cudaStreamSynchronize(st));
        CCG_CUDA(ctx, cudaGetLastError());
        if (res->nonfinite) return set_error(ctx, CCG_ERR_NONFINITE, "cloud holds non-finite coordinates");
        for (int d = 0; d < 3; ++d) { ctx->bb_lo[d] = ord2f(res->lo[d]); ctx->bb_hi[d] = ord2f(res->hi[d]); }
        cudaEventElapsedTime(&ctx->stats.ms_upload, ctx->ev[0], ctx->ev[1]);
        cudaEventElapsedTime(&ctx->stats.ms_build, ctx->ev[1], ctx->ev[2]);
    }
    ctx->n = n; ctx->npad = npad; ctx->nleaf = nleaf; ctx->nlevels = nlevels;
    ctx->has_cloud = true;
    ctx->stats.n_points = (int64_t)n;
    ctx->stats.n_leaves = nleaf;
    return CCG_OK;
}

}  // namespace ccg
