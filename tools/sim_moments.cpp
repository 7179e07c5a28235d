// Host check of add_box_moments (tile.cuh): the count and moments of a box about its OWN centre, shifted to a query by the
// parallel-axis rule, against fp64 sums of (p - q) taken point by point — also for clouds hundreds of metres from the origin,
// where sums about the origin would lose everything to cancellation.
// Build + run (from the repo root; tests/test_radsel_model.py does the same):
//   python -c "s=open('cloudclean_b200/csrc/tile.cuh').read(); a=s.index('__device__ __forceinline__ int add_box_moments'); open('build/sim/box_moments.inc','w').write(s[a:s.index('// Bottom-up, nearest-first walk')])"
//   g++ -O2 -ffp-contract=off -std=c++17 -Ibuild/sim tools/sim_moments.cpp -o build/sim/sim_moments && ./build/sim/sim_moments
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#define __device__
#define __forceinline__ inline
struct float4 { float x, y, z, w; };
static inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline int __float_as_int(float f) { int i; __builtin_memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; __builtin_memcpy(&f, &i, 4); return f; }
static inline float4 __ldg(const float4 *p) { return *p; }
#include "box_moments.inc"

int main() {
    srand(11);
    double worst = 0;
    long cases = 0, fails = 0;
    for (int rep = 0; rep < 4000; ++rep) {
        const float off = (rep % 4 == 0) ? 0.f : (rep % 4 == 1 ? 30.f : (rep % 4 == 2 ? 250.f : -900.f));
        const float ext = 0.002f + 0.2f * (rand() / (float)RAND_MAX);             // box extent
        const int n = 1 + rand() % 1024;
        std::vector<float4> p(n);
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (auto &v : p) {
            v = {off + ext * (rand() / (float)RAND_MAX), 0.5f * off + ext * (rand() / (float)RAND_MAX), 0.1f * off + 0.3f * ext * (rand() / (float)RAND_MAX), 0.f};
            lo[0] = fminf(lo[0], v.x); hi[0] = fmaxf(hi[0], v.x); lo[1] = fminf(lo[1], v.y); hi[1] = fmaxf(hi[1], v.y);
            lo[2] = fminf(lo[2], v.z); hi[2] = fmaxf(hi[2], v.z);
        }
        const float4 L = {lo[0], lo[1], lo[2], 0.f}, H = {hi[0], hi[1], hi[2], 0.f};
        const float cx = 0.5f * (lo[0] + hi[0]), cy = 0.5f * (lo[1] + hi[1]), cz = 0.5f * (lo[2] + hi[2]);
        // the box's stored moments: fp32 sums about its centre, as k_leaf_moments / k_level_moments produce them
        float s[9] = {0};
        for (auto &v : p) {
            const float X = v.x - cx, Y = v.y - cy, Z = v.z - cz;
            s[0] += X; s[1] += Y; s[2] += Z; s[3] += X * X; s[4] += X * Y; s[5] += X * Z; s[6] += Y * Y; s[7] += Y * Z; s[8] += Z * Z;
        }
        float4 mom[3] = {{s[0], s[1], s[2], __int_as_float(n)}, {s[3], s[4], s[5], s[6]}, {s[7], s[8], 0.f, 0.f}};
        // a query within a radius r of the box
        const float r = ext * (1.f + 3.f * (rand() / (float)RAND_MAX));
        const float4 q = {cx + r * (rand() / (float)RAND_MAX - 0.5f), cy + r * (rand() / (float)RAND_MAX - 0.5f), cz + r * (rand() / (float)RAND_MAX - 0.5f), 0.f};
        float m[9] = {0};
        const int cnt = add_box_moments(mom, 0, L, H, q, true, m);
        double t[9] = {0};
        for (auto &v : p) {
            const double x = (double)v.x - q.x, y = (double)v.y - q.y, z = (double)v.z - q.z;
            t[0] += x; t[1] += y; t[2] += z; t[3] += x * x; t[4] += x * y; t[5] += x * z; t[6] += y * y; t[7] += y * z; t[8] += z * z;
        }
        const double R = std::sqrt(3.0) * (0.5 * ext + 0.5 * r) * 1.5;           // scale of |p - q|
        double err = 0;
        for (int i = 0; i < 3; ++i) err = fmax(err, fabs(m[i] - t[i]) / (n * R));
        for (int i = 3; i < 9; ++i) err = fmax(err, fabs(m[i] - t[i]) / (n * R * R));
        worst = fmax(worst, err);
        ++cases;
        // fp32 sums of n <= 1024 terms + the rounding of the coordinates themselves at |p| up to 900 m (2^-24 * 900 / R per term)
        const double tol = 2e-5 + 3.0 * 6e-8 * fabs(off) / R;
        if (cnt != n || err > tol) { ++fails; printf("FAIL rep %d off %g n %d ext %g r %g err %.3g tol %.3g cnt %d\n", rep, off, n, ext, r, err, tol, cnt); }
    }
    printf("cases %ld fails %ld worst relative error %.3g\n", cases, fails, worst);
    return fails != 0;
}
