// ptx.cu — f4 of SURVEY.md §8: PointCloud::load_ptx (src/model/pointcloud.cpp:127-330) with the ASCII body parsed on the GPU.
//
// The reference reads a Leica PTX file with fscanf: a 10-line header (columns, rows, scanner position, 3x3 axes,
// 4x4 transform), then file_width*file_height lines "x y z intensity [r g b]".  It keeps every decimation-th row and
// column (:223-230), drops (0,0,0) returns (:263-265) and appends pcl::PointXYZI {x, y, z, 1, intensity, -, float(index
// in the cloud), float(index in the sampled grid)} plus cloud_to_grid_map (:284-287).  That loop is the user's real
// wall-clock wait (10^7..10^8 fscanf calls).
//
// Here the header (10 lines) is read on the host and the body is parsed on the device:
//   1. k_ptx_count + scan    line starts per 16 KB segment of the body -> first line number of every segment
//   2. k_ptx_parse           one block per segment: bytes staged in shared memory, each thread parses the lines that start
//                            in its 64-byte chunk: four correctly rounded decimal -> float conversions per line
//   3. exclusive scan        non-blank lines -> line number (row, column of the scan grid)
//   4. k_ptx_classify        decimation, (0,0,0) test, first malformed sampled line
//   5. exclusive scan        kept lines -> index in the cloud
//   6. k_ptx_emit            32-byte pcl::PointXYZI rows + cloud_to_grid_map, in file order
// Everything streams: the bytes twice from HBM (coalesced), 34 B of temporaries per line, 36 B out.
//
// Decimal conversion: sign, up to 19 significant digits and a decimal exponent give V = m * 10^e exactly; a double
// estimate of V is within a few ulp, so (float)estimate is the correctly rounded result unless the estimate lies within
// 2^-47 (relative) of the midpoint of two adjacent floats; those cases are decided exactly by a 320-bit integer
// comparison of m * 5^a * 2^s against the midpoint.  Results equal glibc's strtof (round to nearest even), which is
// what fscanf("%f") uses.  More than 19 significant digits, magnitudes outside [1e-37, 3e38] (other than 0), "nan",
// "inf" and hex floats are reported as unsupported (the line number is returned) — never guessed.
#include "common.cuh"
#include <locale.h>
#include <limits.h>
#include <cub/block/block_reduce.cuh>
#include <cub/block/block_scan.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

namespace ccg {

enum { kLineBlank = 0, kLineOk = 1, kLineMalformed = 2, kLineUnsupported = 3 };

__constant__ double c_pow10[65] = {
    1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
    1e17, 1e18, 1e19, 1e20, 1e21, 1e22, 1e23, 1e24, 1e25, 1e26, 1e27, 1e28, 1e29, 1e30, 1e31, 1e32, 1e33,
    1e34, 1e35, 1e36, 1e37, 1e38, 1e39, 1e40, 1e41, 1e42, 1e43, 1e44, 1e45, 1e46, 1e47, 1e48, 1e49, 1e50,
    1e51, 1e52, 1e53, 1e54, 1e55, 1e56, 1e57, 1e58, 1e59, 1e60, 1e61, 1e62, 1e63, 1e64};

__constant__ double c_inv_pow10[65] = {
    1e-0,  1e-1,  1e-2,  1e-3,  1e-4,  1e-5,  1e-6,  1e-7,  1e-8,  1e-9,  1e-10, 1e-11, 1e-12, 1e-13, 1e-14, 1e-15, 1e-16,
    1e-17, 1e-18, 1e-19, 1e-20, 1e-21, 1e-22, 1e-23, 1e-24, 1e-25, 1e-26, 1e-27, 1e-28, 1e-29, 1e-30, 1e-31, 1e-32, 1e-33,
    1e-34, 1e-35, 1e-36, 1e-37, 1e-38, 1e-39, 1e-40, 1e-41, 1e-42, 1e-43, 1e-44, 1e-45, 1e-46, 1e-47, 1e-48, 1e-49, 1e-50,
    1e-51, 1e-52, 1e-53, 1e-54, 1e-55, 1e-56, 1e-57, 1e-58, 1e-59, 1e-60, 1e-61, 1e-62, 1e-63, 1e-64};

// ---------------------------------------------------------------------------------------------------
// exact comparison  m * 10^e10  <=>  K * 2^p   (K < 2^26) with 320-bit unsigned integers
// ---------------------------------------------------------------------------------------------------
constexpr int kLimbs = 10;
struct Big {
    unsigned w[kLimbs];
};
__device__ __forceinline__ void big_set(Big &b, unsigned long long v) {
#pragma unroll
    for (int i = 0; i < kLimbs; ++i) b.w[i] = 0;
    b.w[0] = (unsigned)v; b.w[1] = (unsigned)(v >> 32);
}
__device__ __forceinline__ void big_mul_small(Big &b, unsigned f) {
    unsigned long long carry = 0;
#pragma unroll
    for (int i = 0; i < kLimbs; ++i) {
        unsigned long long t = (unsigned long long)b.w[i] * f + carry;
        b.w[i] = (unsigned)t; carry = t >> 32;
    }
}
__device__ __forceinline__ void big_mul_pow5(Big &b, int e) {
    while (e >= 13) { big_mul_small(b, 1220703125u); e -= 13; }        // 5^13 < 2^32
    unsigned f = 1;
    for (int i = 0; i < e; ++i) f *= 5u;
    if (e > 0) big_mul_small(b, f);
}
__device__ __forceinline__ void big_shl(Big &b, int s) {
    const int limbs = s >> 5, bits = s & 31;
    Big r;
#pragma unroll
    for (int i = 0; i < kLimbs; ++i) {
        const int j = i - limbs;
        unsigned lo = (j >= 0 && j < kLimbs) ? b.w[j] : 0u;
        unsigned below = (j - 1 >= 0 && j - 1 < kLimbs) ? b.w[j - 1] : 0u;
        r.w[i] = bits ? ((lo << bits) | (below >> (32 - bits))) : lo;
    }
    b = r;
}
__device__ __forceinline__ int big_cmp(const Big &a, const Big &b) {
    for (int i = kLimbs - 1; i >= 0; --i) {
        if (a.w[i] != b.w[i]) return a.w[i] > b.w[i] ? 1 : -1;
    }
    return 0;
}
// sign of (m * 10^e10  -  K * 2^p);  callers guarantee both sides fit 320 bits (|e10| <= 64, value within float range)
__device__ __noinline__ int cmp_decimal_binary(unsigned long long m, int e10, unsigned K, int p) {
    Big a, b;
    big_set(a, m);
    big_set(b, K);
    if (e10 >= 0) big_mul_pow5(a, e10); else big_mul_pow5(b, -e10);
    const int t = p - e10;              // m * 5^a  <=>  K * 5^b * 2^t
    if (t >= 0) big_shl(b, t); else big_shl(a, -t);
    return big_cmp(a, b);
}

// ---------------------------------------------------------------------------------------------------
// decimal token -> float, correctly rounded (round to nearest even), or a status saying why not
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool is_blank(unsigned char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }
__device__ __forceinline__ bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }

// m * 10^e10 (m > 0) -> float; force_exact routes every number through the integer comparison (test knob)
__device__ __forceinline__ int decimal_to_float(unsigned long long m, int e10, bool force_exact, float &out) {
    if (e10 > 64 || e10 < -64) return kLineUnsupported;
    const double dm = (double)m;                                  // rounds when m >= 2^53: inside the error budget below
    // a multiplication by the rounded reciprocal instead of a division: one more rounding (2^-53), far inside the budget
    const double d = e10 >= 0 ? dm * c_pow10[e10] : dm * c_inv_pow10[-e10];
    if (!(d >= 1e-37 && d <= 3e38)) return kLineUnsupported;
    // d = V (1 + eps), |eps| < 2^-50.  Bits of d below the float mantissa: 29; the midpoint pattern is 1 << 28.
    const unsigned long long bits = (unsigned long long)__double_as_longlong(d);
    const long long low = (long long)(bits & 0x1fffffffULL) - (1LL << 28);
    if (!force_exact && (low > 64 || low < -64)) { out = (float)d; return kLineOk; }
    // candidate pair: fl = k * 2^q (the float below the midpoint region d sits in) and its successor
    int E = (int)((bits >> 52) & 0x7ff) - 1023;
    unsigned long long mant = (bits & 0xfffffffffffffULL) | (1ULL << 52);
    // force_exact may arrive here far from a midpoint, possibly just below a float: take the float at or below d
    unsigned k = (unsigned)(mant >> 29);
    int q = E - 23;
    // exact V against the midpoint (2k+1) * 2^(q-1)
    int c = cmp_decimal_binary(m, e10, 2u * k + 1u, q - 1);
    if (force_exact && c < 0) {
        // V may even be below fl itself when d was rounded up across a float (d's error): compare with fl, step down if so
        if (cmp_decimal_binary(m, e10, k, q) < 0) {
            if (k == (1u << 23)) { k = (1u << 24) - 1u; q -= 1; } else k -= 1u;
            c = cmp_decimal_binary(m, e10, 2u * k + 1u, q - 1);
        }
    }
    // (V cannot lie above the NEXT midpoint: d would have to be off by more than half a float ulp)
    unsigned fb = ((unsigned)(q + 23 + 127) << 23) | (k & 0x7fffffu);
    if (c > 0 || (c == 0 && (k & 1u))) fb += 1u;                 // carries into the exponent when k + 1 = 2^24
    out = __uint_as_float(fb);
    return kLineOk;
}

// one "%f" of fscanf restricted to plain decimal notation; p advances past the token
template <class Bytes, class Idx>
__device__ __forceinline__ int parse_float(const Bytes &b, Idx &p, Idx end, bool force_exact, float &out) {
    while (p < end && is_blank(b[p])) ++p;
    if (p >= end) return kLineBlank;                             // no token left on this line
    bool neg = false;
    if (b[p] == '-' || b[p] == '+') { neg = b[p] == '-'; ++p; }
    unsigned long long m = 0;
    int nd = 0, e10 = 0, any = 0;
    bool dropped = false;
    // integer and fraction digits in ONE loop, so that lanes in different parts of their numbers run the same code:
    // up to 19 significant digits go into m; digits beyond that only move the exponent (integer part) and must be 0
    bool dot = false;
    for (; p < end; ++p) {
        const unsigned ch = b[p];
        const unsigned c = ch - '0';
        if (c <= 9u) {
            any = 1;
            const bool room = nd < 19;
            if (room && (m != 0 || c != 0)) { m = m * 10 + c; ++nd; }
            e10 += room ? (dot ? -1 : 0) : (dot ? 0 : 1);
            dropped |= !room && c != 0;
        } else if (ch == '.' && !dot) dot = true;
        else break;
    }
    if (!any) return kLineMalformed;                             // "nan", "inf", stray text
    if (p < end && (b[p] == 'e' || b[p] == 'E')) {
        Idx s = p + 1;
        bool eneg = false;
        if (s < end && (b[s] == '-' || b[s] == '+')) { eneg = b[s] == '-'; ++s; }
        if (s >= end || !is_digit(b[s])) return kLineMalformed;
        int ex = 0;
        while (s < end && is_digit(b[s])) { if (ex < 100000) ex = ex * 10 + (b[s] - '0'); ++s; }
        e10 += eneg ? -ex : ex;
        p = s;
    }
    if (p < end && !is_blank(b[p])) return kLineMalformed;       // "1.5abc", hex floats ("0x..." stops at 'x')
    if (m == 0) { out = neg ? -0.f : 0.f; return kLineOk; }
    if (dropped) return kLineUnsupported;
    float v;
    const int st = decimal_to_float(m, e10, force_exact, v);
    if (st != kLineOk) return st;
    out = neg ? -v : v;
    return kLineOk;
}

// ---------------------------------------------------------------------------------------------------
// The body is cut into segments of kSeg bytes on a 16-byte aligned grid (base_al = the aligned address at or below the
// first body byte); one block per segment, one 60-byte chunk per thread (15 words: an odd word stride, so threads
// reading their own chunk from shared memory hit distinct banks without padding).
// ---------------------------------------------------------------------------------------------------
constexpr int kThreads = 256;
constexpr int kChunk = 60;
constexpr int kSeg = kThreads * kChunk;      // 15360
constexpr int kLead = 16;                     // bytes staged in front of the segment (the previous byte decides a line start)
constexpr int kTail = 512;                    // bytes staged beyond it: lines that end within kTail are parsed from shared memory
constexpr int kStage = kLead + kSeg + kTail;  // 15888, a multiple of 16
constexpr int kListCap = 2048;                // line starts kept per segment (a valid line has >= 8 bytes: <= 1920)

struct GlobalBytes {
    const unsigned char *b;
    __device__ __forceinline__ unsigned char operator[](size_t g) const { return __ldg(b + g); }
};
struct StagedBytes {                 // the shared-memory image of bytes [org, org + kStage)
    const unsigned char *s;
    size_t org;
    __device__ __forceinline__ unsigned char operator[](size_t g) const { return s[g - org]; }      // absolute offset
};
struct StageLocal {                  // same image, 32-bit offsets relative to org (the parse loops live on these)
    const unsigned char *s;
    __device__ __forceinline__ unsigned char operator[](int o) const { return s[o]; }
};

// bit j of the result: a line starts at byte c0 + j (0 <= j < 60), i.e. byte c0 + j - 1 is '\n' or c0 + j == body; only
// starts inside [lo, hi) count.  chunk = address of byte c0 (4-byte aligned; shared or global), chunk[-1] is readable.
__device__ __forceinline__ unsigned long long chunk_starts(const unsigned char *chunk, size_t c0, size_t body, size_t lo, size_t hi) {
    if (c0 >= hi || c0 + kChunk <= lo) return 0ull;
    const unsigned *w = reinterpret_cast<const unsigned *>(chunk);
    const int nwords = (int)(((hi - c0) < (size_t)kChunk ? (hi - c0) : (size_t)kChunk) + 3) >> 2;
    unsigned long long nl = 0;
#pragma unroll
    for (int i = 0; i < kChunk / 4; ++i) {
        if (i < nwords) {
            const unsigned x = w[i] ^ 0x0a0a0a0au;
            const unsigned m = ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu);      // 0x80 in every byte that was '\n'
            const unsigned nib = (((m >> 7) * 0x00204081u) >> 21) & 0xfu;
            nl |= (unsigned long long)nib << (4 * i);
        }
    }
    unsigned long long st = ((nl << 1) | (chunk[-1] == '\n' ? 1ull : 0ull)) & ((1ull << kChunk) - 1);
    if (body >= c0 && body < c0 + kChunk) st |= 1ull << (int)(body - c0);
    if (c0 < lo) st &= ~0ull << (int)(lo - c0);
    if (c0 + kChunk > hi) st &= (1ull << (int)(hi - c0)) - 1;
    return st;
}

__global__ void __launch_bounds__(kThreads) k_ptx_count(const unsigned char *__restrict__ bytes, size_t base_al, size_t body, size_t nbytes,
                                                        unsigned *__restrict__ counts) {
    const size_t seg_a = base_al + (size_t)blockIdx.x * kSeg;
    const size_t lo = seg_a > body ? seg_a : body;
    const size_t hi = (seg_a + kSeg) < nbytes ? (seg_a + kSeg) : nbytes;
    const size_t c0 = seg_a + (size_t)threadIdx.x * kChunk;
    const unsigned c = (unsigned)__popcll(chunk_starts(bytes + c0, c0, body, lo, hi));
    typedef cub::BlockReduce<unsigned, kThreads> Reduce;
    __shared__ typename Reduce::TempStorage tmp;
    const unsigned total = Reduce(tmp).Sum(c);
    if (threadIdx.x == 0) counts[blockIdx.x] = total;
}

template <class Bytes, class Idx>
__device__ __forceinline__ void parse_line(const Bytes &b, Idx p, Idx end, int has_rgb, bool force_exact, float4 &out, unsigned char &status) {
    while (end > p && (b[end - 1] == '\n' || is_blank(b[end - 1]))) --end;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    int st = kLineOk;
    if (p >= end) st = kLineBlank;
    for (int t = 0; t < 4 && st == kLineOk; ++t) {
        const int r = parse_float(b, p, end, force_exact, v[t]);
        st = r == kLineBlank ? kLineMalformed : r;               // fewer than four numbers
    }
    if (st == kLineOk && !has_rgb) {
        while (p < end && is_blank(b[p])) ++p;
        if (p < end) st = kLineMalformed;                        // the reference's next fscanf would start mid-line
    }
    out = make_float4(v[0], v[1], v[2], v[3]);
    status = (unsigned char)st;
}

// line [p, e) -> vals/status[row]; e = position of the terminating '\n' or nbytes
__device__ __forceinline__ void parse_and_store(const StagedBytes &sb, const GlobalBytes &gb, size_t p, size_t e, size_t stage_end,
                                                int has_rgb, bool force_exact, float4 *vals, unsigned char *status, unsigned long long row) {
    float4 v; unsigned char st;
    if (e <= stage_end) parse_line(StageLocal{sb.s}, (int)(p - sb.org), (int)(e - sb.org), has_rgb, force_exact, v, st);
    else parse_line(gb, p, e, has_rgb, force_exact, v, st);
    vals[row] = v;
    status[row] = st;
}

__device__ __forceinline__ size_t find_newline(const StagedBytes &sb, const GlobalBytes &gb, size_t from, size_t stage_end, size_t nbytes) {
    size_t e = from;
    while (e < stage_end && sb[e] != '\n') ++e;
    if (e == stage_end && stage_end < nbytes) { while (e < nbytes && gb[e] != '\n') ++e; }
    return e;
}

// One block per segment: the bytes are staged in shared memory with 16-byte loads, every thread finds the line starts of its
// chunk with word arithmetic, a block scan ranks them, the starts go to a shared list and the lines are dealt out round
// robin (thread t takes lines t, t + 256, ...), so lanes stay busy whatever the line lengths: "x y z intensity" (+ "r g b"
// ignored when has_rgb, pointcloud.cpp:253-256).  A line that runs past the staged tail is parsed from global memory; a
// segment with more starts than the list holds (a flood of blank lines) keeps the lines with the thread that found them.
__global__ void __launch_bounds__(kThreads) k_ptx_parse(const unsigned char *__restrict__ bytes, size_t base_al, size_t body, size_t nbytes,
                                                        const unsigned long long *__restrict__ seg_base, int has_rgb, int force_exact,
                                                        float4 *__restrict__ vals, unsigned char *__restrict__ status) {
    __shared__ __align__(16) unsigned char s_bytes[kStage];
    __shared__ unsigned short s_starts[kListCap];
    typedef cub::BlockScan<unsigned, kThreads> Scan;
    __shared__ typename Scan::TempStorage s_scan;
    const size_t seg_a = base_al + (size_t)blockIdx.x * kSeg;
    const size_t org = seg_a - kLead;                             // 16-byte aligned address, >= 0 (checked by the host)
    const size_t stage_end = (org + kStage) < nbytes ? (org + kStage) : nbytes;
    {
        // every 16-byte vector that holds at least one valid byte lies inside the allocation (aligned, granule >= 16)
        const uint4 *src = reinterpret_cast<const uint4 *>(bytes + org);
        const int nvec = (int)((stage_end - org + 15) >> 4);
        for (int v = threadIdx.x; v < nvec; v += kThreads) reinterpret_cast<uint4 *>(s_bytes)[v] = __ldg(src + v);
    }
    __syncthreads();
    const StagedBytes sb{s_bytes, org};
    const GlobalBytes gb{bytes};
    const size_t lo = seg_a > body ? seg_a : body;
    const size_t hi = (seg_a + kSeg) < nbytes ? (seg_a + kSeg) : nbytes;
    const size_t c0 = seg_a + (size_t)threadIdx.x * kChunk;
    unsigned long long mask = chunk_starts(s_bytes + kLead + threadIdx.x * kChunk, c0, body, lo, hi);
    unsigned rank = 0, total = 0;
    Scan(s_scan).ExclusiveSum((unsigned)__popcll(mask), rank, total);
    const unsigned long long row0 = seg_base[blockIdx.x];
    const bool fe = force_exact != 0;
    if (total <= (unsigned)kListCap) {
        unsigned r = rank;
        while (mask) {
            const int bit = __ffsll((long long)mask) - 1;
            mask &= mask - 1;
            s_starts[r++] = (unsigned short)(kLead + threadIdx.x * kChunk + bit);
        }
        __syncthreads();
        for (unsigned i = threadIdx.x; i < total; i += kThreads) {
            const size_t p = org + s_starts[i];
            const size_t e = i + 1 < total ? org + s_starts[i + 1] - 1 : find_newline(sb, gb, p, stage_end, nbytes);
            parse_and_store(sb, gb, p, e, stage_end, has_rgb, fe, vals, status, row0 + i);
        }
    } else {
        unsigned long long row = row0 + rank;
        while (mask) {
            const int bit = __ffsll((long long)mask) - 1;
            mask &= mask - 1;
            const size_t p = c0 + (size_t)bit;
            parse_and_store(sb, gb, p, find_newline(sb, gb, p, stage_end, nbytes), stage_end, has_rgb, fe, vals, status, row++);
        }
    }
}

struct ToU64 {
    __host__ __device__ __forceinline__ unsigned long long operator()(unsigned v) const { return (unsigned long long)v; }
};
struct NonBlank {
    const unsigned char *status;
    __host__ __device__ __forceinline__ unsigned long long operator()(unsigned long long i) const { return status[i] != kLineBlank ? 1ull : 0ull; }
};
struct Kept {
    const unsigned char *keep;
    __host__ __device__ __forceinline__ unsigned long long operator()(unsigned long long i) const { return keep[i] ? 1ull : 0ull; }
};

// decimation rule of pointcloud.cpp:219-230 on the line number (blank lines do not count), the (0,0,0) drop of :263-265,
// and the first sampled line that could not be parsed.  STRICTER THAN THE REFERENCE, on purpose: there a failed fscanf without
// a stream error `continue`s (:247-257) with the file position left in the middle of the bad line, so every later row is
// read from wherever the stream happens to stand; this loader refuses the file with the line number instead (documented in
// include/ccgpu.h and INTEGRATION.md).  Only a stream error makes the reference return false (:259-264).
__global__ void __launch_bounds__(256) k_ptx_classify(const float4 *__restrict__ vals, const unsigned char *__restrict__ status,
                                                      const unsigned long long *__restrict__ lineno, size_t nlines, long long file_width,
                                                      long long line_count, int dec, unsigned char *__restrict__ keep,
                                                      long long *__restrict__ grid_index, unsigned long long *__restrict__ first_bad,
                                                      unsigned long long *__restrict__ counters) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlines) return;
    unsigned char k = 0;
    long long gi = -1;
    const int st = status[i];
    const long long ln = (long long)lineno[i];
    if (st != kLineBlank && ln < line_count) {
        const long long row = ln / file_width, col = ln % file_width;
        if ((row + 1) % dec == 0 && (col + 1) % dec == 0) {
            if (st != kLineOk) {
                atomicMin(first_bad, ((unsigned long long)ln << 2) | (unsigned long long)st);
            } else {
                gi = ((row + 1) / dec - 1) * (file_width / dec) + ((col + 1) / dec - 1);     // sampled_idx - 1
                atomicAdd(&counters[0], 1ull);
                const float4 v = vals[i];
                k = !(v.x == 0.f && v.y == 0.f && v.z == 0.f);
            }
        }
    }
    keep[i] = k;
    grid_index[i] = gi;
}

// pcl::PointXYZI rows as load_ptx leaves them (:284-286): data[3] = 1, data_c = {intensity, 0, float(cloud index),
// float(grid index)}; cloud_to_grid_map (:287)
__global__ void __launch_bounds__(256) k_ptx_emit(const float4 *__restrict__ vals, const unsigned char *__restrict__ keep,
                                                  const long long *__restrict__ grid_index, const unsigned long long *__restrict__ pos,
                                                  size_t nlines, float4 *__restrict__ points, int32_t *__restrict__ c2g) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlines || !keep[i]) return;
    const unsigned long long o = pos[i];
    const float4 v = vals[i];
    const long long gi = grid_index[i];
    points[2 * o] = make_float4(v.x, v.y, v.z, 1.f);
    points[2 * o + 1] = make_float4(v.w, 0.f, (float)(unsigned long long)o, (float)(unsigned)gi);
    c2g[o] = (int32_t)gi;
}


// ---------------------------------------------------------------------------------------------------
// host side: the 10 header lines (pointcloud.cpp:148-180) and the rgb test on the first body line (:199-216)
// ---------------------------------------------------------------------------------------------------
static int parse_header(ccg_ctx *ctx, const char *text, size_t len, int dec, ccg_ptx_info *info) {
    // text is NUL-terminated (len bytes + 0)
    // The reference parses these 14 numbers with fscanf under setlocale(LC_NUMERIC, "C") (pointcloud.cpp:132: the Qt host
    // application runs in the user's locale).  A library must not touch the process locale, so the numbers are read with
    // strtol / strtof_l against an explicit "C" locale object: same grammar as fscanf's %d / %f (leading white space
    // skipped, longest valid prefix), independent of LC_NUMERIC.
    static locale_t c_loc = newlocale(LC_ALL_MASK, "C", (locale_t)0);
    const char *cur = text;
    char *end = nullptr;
    const long wl = strtol(cur, &end, 10);
    if (end == cur) return set_error(ctx, CCG_ERR_INVALID, "PTX header: cannot read the grid size");
    cur = end;
    const long hl = strtol(cur, &end, 10);
    if (end == cur) return set_error(ctx, CCG_ERR_INVALID, "PTX header: cannot read the grid size");
    cur = end;
    const int w = (int)wl, h = (int)hl;
    if (wl <= 0 || hl <= 0 || wl > INT_MAX || hl > INT_MAX) return set_error(ctx, CCG_ERR_INVALID, "PTX header: grid %ld x %ld", wl, hl);
    float f[12];
    for (int i = 0; i < 12; ++i) {
        f[i] = c_loc ? strtof_l(cur, &end, c_loc) : strtof(cur, &end);
        if (end == cur) return set_error(ctx, CCG_ERR_INVALID, "PTX header: cannot read scanner pose value %d", i);
        cur = end;
    }
    const int off = (int)(cur - text);
    // five fgets: the rest of the current line, then the 4x4 transform (:175-180)
    size_t p = (size_t)off;
    for (int i = 0; i < 5; ++i) {
        while (p < len && text[p] != '\n') ++p;
        if (p < len) ++p;
    }
    info->file_width = w; info->file_height = h;
    info->scan_width = w / dec; info->scan_height = h / dec;
    for (int i = 0; i < 3; ++i) info->sensor_origin[i] = f[i];
    for (int i = 0; i < 9; ++i) info->orientation[i] = f[3 + i];
    info->body_offset = p;
    // tokens of the first body line, split at spaces only (strtok(buff, " "), :203-211): 7 means x y z i r g b
    int tokens = 0;
    size_t q = p;
    while (q < len && text[q] != '\n') {
        while (q < len && text[q] == ' ') ++q;
        if (q >= len) break;
        if (text[q] == '\n') { ++tokens; break; }                // "... \n": strtok returns the newline as a token of its own
        ++tokens;
        while (q < len && text[q] != ' ' && text[q] != '\n') ++q;
    }
    info->has_rgb = tokens == 7 ? 1 : 0;
    return CCG_OK;
}

static int ptx_begin(ccg_ctx *ctx) {
    if (!ctx) return CCG_ERR_INVALID;
    ctx->err[0] = 0;
    CCG_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->cancel && *ctx->cancel) return set_error(ctx, CCG_ERR_CANCELLED, "cancelled by caller");
    ctx->stats.launches = 0;
    return CCG_OK;
}

static bool on_device(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Host <-> device transfers of pageable memory (an mmap of the file, the caller's std::vector) through pinned staging
// slots filled / drained by a few host threads: a plain cudaMemcpy from pageable memory runs at one core's memcpy speed
// (measured 11 GB/s for the 1.2 GB file), the link takes five times that.  to_device: src (host) -> dst (device);
// otherwise src (device) -> dst (host).
static int staged_copy(ccg_ctx *ctx, void *dst, const void *src, size_t nbytes, bool to_device) {
    const size_t kSlot = (size_t)16 << 20;
    unsigned hw = std::thread::hardware_concurrency();
    int T = (int)(hw / 2);
    if (T < 1) T = 1;
    if (T > 8) T = 8;
    const char *env = getenv("CCG_COPY_THREADS");
    if (env && atoi(env) > 0) T = atoi(env) > 16 ? 16 : atoi(env);
    const size_t nchunks = (nbytes + kSlot - 1) / kSlot;
    if (nchunks < 4 || T == 1) {
        CCG_CUDA(ctx, cudaMemcpyAsync(dst, src, nbytes, to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, ctx->stream));
        return CCG_OK;
    }
    if ((size_t)T > nchunks / 2) T = (int)(nchunks / 2);
    CCG_CUDA(ctx, ctx->hpin[0].reserve((size_t)T * 2 * kSlot));
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));           // device-side producers of src / consumers of dst are done
    std::vector<cudaError_t> errs((size_t)T, cudaSuccess);
    std::vector<std::thread> pool;
    char *pin = ctx->hpin[0].as<char>();
    const int device = ctx->device;
    for (int t = 0; t < T; ++t) {
        pool.emplace_back([=, &errs]() {
            cudaError_t e = cudaSetDevice(device);
            cudaStream_t st = nullptr;
            cudaEvent_t ev[2] = {nullptr, nullptr};
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
            for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
            size_t k = 0;
            size_t pending_off[2] = {0, 0}, pending_len[2] = {0, 0};
            for (size_t c = (size_t)t; c < nchunks && e == cudaSuccess; c += (size_t)T, ++k) {
                const int sl = (int)(k & 1);
                char *slot = pin + ((size_t)t * 2 + sl) * kSlot;
                const size_t off = c * kSlot, len = (nbytes - off) < kSlot ? (nbytes - off) : kSlot;
                if (to_device) {
                    if (k >= 2) e = cudaEventSynchronize(ev[sl]);                     // the slot's previous upload has left
                    if (e != cudaSuccess) break;
                    memcpy(slot, (const char *)src + off, len);
                    e = cudaMemcpyAsync((char *)dst + off, slot, len, cudaMemcpyHostToDevice, st);
                    if (e == cudaSuccess) e = cudaEventRecord(ev[sl], st);
                } else {
                    if (k >= 2) {                                                      // drain what this slot downloaded before
                        e = cudaEventSynchronize(ev[sl]);
                        if (e != cudaSuccess) break;
                        memcpy((char *)dst + pending_off[sl], slot, pending_len[sl]);
                    }
                    e = cudaMemcpyAsync(slot, (const char *)src + off, len, cudaMemcpyDeviceToHost, st);
                    if (e == cudaSuccess) e = cudaEventRecord(ev[sl], st);
                    pending_off[sl] = off; pending_len[sl] = len;
                }
            }
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (!to_device && e == cudaSuccess) {
                const size_t done = k;                                                 // chunks this thread issued
                for (size_t j = done >= 2 ? done - 2 : 0; j < done; ++j) {
                    const int sl = (int)(j & 1);
                    memcpy((char *)dst + pending_off[sl], pin + ((size_t)t * 2 + sl) * kSlot, pending_len[sl]);
                }
            }
            for (int i = 0; i < 2; ++i) if (ev[i]) cudaEventDestroy(ev[i]);
            if (st) cudaStreamDestroy(st);
            errs[(size_t)t] = e;
        });
    }
    for (auto &th : pool) th.join();
    for (int t = 0; t < T; ++t)
        if (errs[(size_t)t] != cudaSuccess) return set_error(ctx, CCG_ERR_CUDA, "staged copy failed: %s", cudaGetErrorString(errs[(size_t)t]));
    return CCG_OK;
}

int ptx_parse(ccg_ctx *ctx, const void *bytes, size_t nbytes, int dec, ccg_ptx_info *info) {
    cudaStream_t st = ctx->stream;
    ctx->ptx_n = 0;
    memset(info, 0, sizeof(*info));
    // header on the host
    const size_t head = nbytes < 16384 ? nbytes : 16384;
    char *text = (char *)malloc(head + 1);
    if (!text) return set_error(ctx, CCG_ERR_INVALID, "out of host memory");
    const bool dev_in = on_device(bytes);
    if (dev_in) {
        cudaError_t e = cudaMemcpy(text, bytes, head, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { free(text); return set_error(ctx, CCG_ERR_CUDA, "header copy failed: %s", cudaGetErrorString(e)); }
    } else memcpy(text, bytes, head);
    text[head] = 0;
    int rc = parse_header(ctx, text, head, dec, info);
    free(text);
    if (rc != CCG_OK) return rc;
    const size_t body = info->body_offset;
    if (body >= nbytes) return CCG_OK;                                  // header only: empty cloud
    // does the last non-blank line end with a newline?  (decides when the reference's stream reports end-of-file)
    bool final_newline = false;
    {
        const size_t tail = (nbytes - body) < 4096 ? (nbytes - body) : 4096;
        char tbuf[4096];
        if (dev_in) {
            cudaError_t e = cudaMemcpy(tbuf, (const char *)bytes + (nbytes - tail), tail, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) return set_error(ctx, CCG_ERR_CUDA, "tail copy failed: %s", cudaGetErrorString(e));
        } else memcpy(tbuf, (const char *)bytes + (nbytes - tail), tail);
        for (size_t i = tail; i-- > 0;) {
            const char c = tbuf[i];
            if (c == '\n') { final_newline = true; break; }
            if (!(c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f')) break;
        }
    }
    // bytes on the device
    const unsigned char *db = (const unsigned char *)bytes;
    float upload_ms = 0.f;
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[3], st));
    if (!dev_in) {
        CCG_CUDA(ctx, ctx->ptx_bytes.reserve(nbytes + 16));
        const auto t0 = std::chrono::steady_clock::now();
        CCG_TRY(staged_copy(ctx, ctx->ptx_bytes.p, bytes, nbytes, true));
        CCG_CUDA(ctx, cudaStreamSynchronize(st));
        upload_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        db = ctx->ptx_bytes.as<unsigned char>();
    }
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[4], st));
    // 1. line starts per 16 KB segment -> first line number of every segment
    CCG_CUDA(ctx, ctx->scratch.reserve(256));
    unsigned long long *d_num = ctx->scratch.as<unsigned long long>();          // [0] unused, [1] first_bad, [2] sampled
    // segment grid on 16-byte aligned addresses (vector loads); the header always leaves room for the 16 staged lead bytes
    const size_t base_al = body - (size_t)((reinterpret_cast<uintptr_t>(db) + body) & 15);
    if (base_al < (size_t)kLead) return set_error(ctx, CCG_ERR_INVALID, "PTX header shorter than %d bytes", 2 * kLead);
    const size_t nseg = (nbytes - base_al + kSeg - 1) / kSeg;
    if (nseg >= ((size_t)1 << 31)) return set_error(ctx, CCG_ERR_UNSUPPORTED, "PTX body too large (%zu bytes)", nbytes - body);
    CCG_CUDA(ctx, ctx->ptx_starts.reserve(nseg * 12 + 64));
    unsigned long long *seg_base = ctx->ptx_starts.as<unsigned long long>();
    unsigned *seg_count = reinterpret_cast<unsigned *>(seg_base + nseg);
    k_ptx_count<<<(unsigned)nseg, kThreads, 0, st>>>(db, base_al, body, nbytes, seg_count);
    {
        cub::TransformInputIterator<unsigned long long, ToU64, const unsigned *> it(seg_count, ToU64());
        size_t tb = 0;
        CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, it, seg_base, (int)nseg, st));
        CCG_CUDA(ctx, ctx->sort_tmp.reserve(tb));
        CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->sort_tmp.p, tb, it, seg_base, (int)nseg, st));
    }
    size_t nlines = 0;
    {
        unsigned long long last_base = 0; unsigned last_count = 0;
        CCG_CUDA(ctx, cudaMemcpyAsync(&last_base, seg_base + (nseg - 1), 8, cudaMemcpyDeviceToHost, st));
        CCG_CUDA(ctx, cudaMemcpyAsync(&last_count, seg_count + (nseg - 1), 4, cudaMemcpyDeviceToHost, st));
        CCG_CUDA(ctx, cudaStreamSynchronize(st));
        nlines = (size_t)(last_base + last_count);
    }
    ctx->stats.launches += 3;
    if (nlines == 0) return CCG_OK;
    if (nlines >= ((size_t)1 << 31)) return set_error(ctx, CCG_ERR_UNSUPPORTED, "PTX body has %zu lines (limit 2^31)", nlines);
    // 2. parse every line (segment by segment, bytes staged in shared memory)
    const unsigned grid = (unsigned)((nlines + 255) / 256);
    CCG_CUDA(ctx, ctx->tmp[0].reserve(nlines * 16 + 64));
    CCG_CUDA(ctx, ctx->tmp[1].reserve(nlines * 2 + 64));
    CCG_CUDA(ctx, ctx->tmp[2].reserve(nlines * 8 + 64));
    CCG_CUDA(ctx, ctx->tmp[3].reserve(nlines * 8 + 64));
    float4 *vals = ctx->tmp[0].as<float4>();
    unsigned char *status = ctx->tmp[1].as<unsigned char>(), *keep = status + nlines;
    unsigned long long *lineno = ctx->tmp[2].as<unsigned long long>();
    long long *grid_index = ctx->tmp[3].as<long long>();
    const char *fe = getenv("CCG_PTX_FORCE_EXACT");
    k_ptx_parse<<<(unsigned)nseg, kThreads, 0, st>>>(db, base_al, body, nbytes, seg_base, info->has_rgb, fe && atoi(fe) ? 1 : 0, vals, status);
    // 3. line numbers (blank lines do not count)
    {
        cub::CountingInputIterator<unsigned long long> cnt(0);
        cub::TransformInputIterator<unsigned long long, NonBlank, cub::CountingInputIterator<unsigned long long>> it(cnt, NonBlank{status});
        size_t tb = 0;
        CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, it, lineno, (int)nlines, st));
        CCG_CUDA(ctx, ctx->sort_tmp.reserve(tb));
        CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->sort_tmp.p, tb, it, lineno, (int)nlines, st));
    }
    // 4. decimation / zero returns / first bad line
    const unsigned long long init[3] = {0ull, ~0ull, 0ull};
    CCG_CUDA(ctx, cudaMemcpyAsync(d_num, init, sizeof(init), cudaMemcpyHostToDevice, st));
    const long long line_count = (long long)info->file_width * (long long)info->file_height;
    k_ptx_classify<<<grid, 256, 0, st>>>(vals, status, lineno, nlines, (long long)info->file_width, line_count, dec, keep, grid_index,
                                         d_num + 1, d_num + 2);
    unsigned long long tail_line[2] = {0, 0};
    unsigned char tail_status = 0;
    CCG_CUDA(ctx, cudaMemcpyAsync(&tail_line[0], lineno + (nlines - 1), 8, cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaMemcpyAsync(&tail_status, status + (nlines - 1), 1, cudaMemcpyDeviceToHost, st));
    // 5. cloud index of every kept line (the scan may overwrite the line numbers)
    unsigned long long *pos = lineno;
    {
        cub::CountingInputIterator<unsigned long long> cnt(0);
        cub::TransformInputIterator<unsigned long long, Kept, cub::CountingInputIterator<unsigned long long>> it(cnt, Kept{keep});
        size_t tb = 0;
        CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, it, pos, (int)nlines, st));
        CCG_CUDA(ctx, ctx->sort_tmp.reserve(tb));
        CCG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->sort_tmp.p, tb, it, pos, (int)nlines, st));
    }
    unsigned long long h_num[3] = {0, 0, 0}, last_pos = 0;
    unsigned char last_keep = 0;
    CCG_CUDA(ctx, cudaMemcpyAsync(h_num, d_num, sizeof(h_num), cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaMemcpyAsync(&last_pos, pos + (nlines - 1), 8, cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaMemcpyAsync(&last_keep, keep + (nlines - 1), 1, cudaMemcpyDeviceToHost, st));
    CCG_CUDA(ctx, cudaStreamSynchronize(st));
    ctx->stats.launches += 6;
    info->lines = (int64_t)(tail_line[0] + (tail_status != kLineBlank ? 1 : 0));
    info->sampled = (int64_t)h_num[2];
    if (h_num[1] != ~0ull) {
        const unsigned long long ln = h_num[1] >> 2;
        return set_error(ctx, (h_num[1] & 3ull) == kLineUnsupported ? CCG_ERR_UNSUPPORTED : CCG_ERR_INVALID,
                         "PTX body line %llu (file line %llu) %s", ln, ln + 11,
                         (h_num[1] & 3ull) == kLineUnsupported ? "holds a number outside the supported decimal range"
                                                               : "is not \"x y z intensity [r g b]\"");
    }
    {
        // pointcloud.cpp:232-235: a sampled line reached with feof() ALREADY set fails the load.  feof gets set by the
        // read that first hits the end of the file: the fscanf of the last existing line when its trailing "\n"
        // directive runs into the end (no rgb), an fgets that ends without a newline, or else the first read attempt
        // past the last line (which then merely `continue`s, :240-246).  So with L lines in the file the load fails iff
        // a sampled line exists beyond t, t = L - 1 or L accordingly.
        const long long L = (long long)info->lines, fw = info->file_width;
        const long long sw = fw / dec, sh = info->file_height / dec;
        const long long need = (sw > 0 && sh > 0) ? (sh * dec - 1) * fw + sw * dec : 0;      // last sampled line + 1
        if (L < need) {
            bool last_sampled = false;
            if (L > 0) { const long long row = (L - 1) / fw, col = (L - 1) % fw; last_sampled = (row + 1) % dec == 0 && (col + 1) % dec == 0; }
            const bool t_is_L = L == 0 ? true : (final_newline && (!last_sampled || info->has_rgb));
            if (L < need - (t_is_L ? 1 : 0))
                return set_error(ctx, CCG_ERR_INVALID, "File read fail at line %lld: the body has %lld lines, the %d x %d grid needs %lld",
                                 L + 1, L, info->file_width, info->file_height, need);
        }
    }
    const size_t npts = (size_t)(last_pos + (last_keep ? 1 : 0));
    info->points = (int64_t)npts;
    // 6. rows in file order
    if (npts) {
        CCG_CUDA(ctx, ctx->ptx_points.reserve(npts * 32 + 64));
        CCG_CUDA(ctx, ctx->ptx_c2g.reserve(npts * 4 + 64));
        k_ptx_emit<<<grid, 256, 0, st>>>(vals, keep, grid_index, pos, nlines, ctx->ptx_points.as<float4>(), ctx->ptx_c2g.as<int32_t>());
        ctx->stats.launches++;
    }
    CCG_CUDA(ctx, cudaEventRecord(ctx->ev[5], st));
    CCG_CUDA(ctx, cudaStreamSynchronize(st));
    CCG_CUDA(ctx, cudaGetLastError());
    ctx->stats.ms_upload = upload_ms;                            // host clock: the upload runs on the copy threads' streams
    cudaEventElapsedTime(&ctx->stats.ms_query, ctx->ev[4], ctx->ev[5]);
    ctx->ptx_n = (int64_t)npts;
    return CCG_OK;
}

}  // namespace ccg

// ---------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

using namespace ccg;

extern "C" int ccg_ptx_parse(ccg_ctx *ctx, const void *bytes, size_t nbytes, int decimation_factor, ccg_ptx_info *info) {
    CCG_TRY(ptx_begin(ctx));
    if (!info) return set_error(ctx, CCG_ERR_INVALID, "info is NULL");
    if (!bytes && nbytes) return set_error(ctx, CCG_ERR_INVALID, "bytes is NULL");
    // pointcloud.cpp:130: assert(decimation_factor%2 == 0 || decimation_factor == 1)
    if (decimation_factor < 1 || (decimation_factor != 1 && decimation_factor % 2 != 0))
        return set_error(ctx, CCG_ERR_INVALID, "decimation_factor must be 1 or even (got %d)", decimation_factor);
    if (nbytes == 0) return set_error(ctx, CCG_ERR_INVALID, "empty PTX buffer");
    return ptx_parse(ctx, bytes, nbytes, decimation_factor, info);
}

extern "C" int ccg_ptx_load(ccg_ctx *ctx, const char *filename, int decimation_factor, ccg_ptx_info *info) {
    CCG_TRY(ptx_begin(ctx));
    if (!filename || !info) return set_error(ctx, CCG_ERR_INVALID, "NULL argument");
    const int fd = open(filename, O_RDONLY);
    if (fd < 0) return set_error(ctx, CCG_ERR_INVALID, "Error opening file %s", filename);      // pointcloud.cpp:139-142
    struct stat sb;
    if (fstat(fd, &sb) != 0 || sb.st_size <= 0) { close(fd); return set_error(ctx, CCG_ERR_INVALID, "cannot stat %s (or empty file)", filename); }
    const size_t nbytes = (size_t)sb.st_size;
    void *map = mmap(nullptr, nbytes, PROT_READ, MAP_PRIVATE, fd, 0);      // pages are touched by the copy threads
    close(fd);
    if (map == MAP_FAILED) return set_error(ctx, CCG_ERR_INVALID, "cannot map %s", filename);
    madvise(map, nbytes, MADV_SEQUENTIAL);
    madvise(map, nbytes, MADV_WILLNEED);
    const int rc = ccg_ptx_parse(ctx, map, nbytes, decimation_factor, info);
    munmap(map, nbytes);
    return rc;
}

extern "C" int ccg_ptx_fetch(ccg_ctx *ctx, void *points32, int32_t *cloud_to_grid) {
    CCG_TRY(ptx_begin(ctx));
    if (ctx->ptx_n == 0) return CCG_OK;
    const size_t n = (size_t)ctx->ptx_n;
    if (points32) {
        if (on_device(points32)) CCG_CUDA(ctx, cudaMemcpyAsync(points32, ctx->ptx_points.p, n * 32, cudaMemcpyDeviceToDevice, ctx->stream));
        else CCG_TRY(staged_copy(ctx, points32, ctx->ptx_points.p, n * 32, false));
    }
    if (cloud_to_grid) {
        if (on_device(cloud_to_grid)) CCG_CUDA(ctx, cudaMemcpyAsync(cloud_to_grid, ctx->ptx_c2g.p, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        else CCG_TRY(staged_copy(ctx, cloud_to_grid, ctx->ptx_c2g.p, n * 4, false));
    }
    CCG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return CCG_OK;
}

extern "C" int ccg_ptx_set_cloud(ccg_ctx *ctx) {
    if (!ctx) return CCG_ERR_INVALID;
    if (ctx->ptx_n == 0) return set_error(ctx, CCG_ERR_NO_CLOUD, "no parsed PTX cloud: call ccg_ptx_parse / ccg_ptx_load first");
    return ccg_set_cloud(ctx, ctx->ptx_points.p, (size_t)ctx->ptx_n, 32);
}
