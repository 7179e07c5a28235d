// Build + run (from the repo root):
//   mkdir -p build/sim && python -c "s=open('cloudclean_b200/csrc/radsel.cu').read(); open('build/sim/radsel_structs.inc','w').write(s[s.index('constexpr int kSelBins'):s.index('// err bit 1')])"
//   g++ -O2 -ffp-contract=off -std=c++17 -Ibuild/sim tools/sim_radsel.cpp -o build/sim/sim_radsel && ./build/sim/sim_radsel 1
// Host model of k_tile_kth for ONE lane: the device structs of radsel.cu compiled with shims, a nearest-first leaf walk with the
// same pruning rule (a leaf is skipped when its lower-bound distance exceeds the current bound), checked against brute force.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <algorithm>
#include <climits>
#include <type_traits>
#define __device__
#define __forceinline__ inline
struct float4 { float x, y, z, w; };
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline int __float_as_int(float f) { int i; __builtin_memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; __builtin_memcpy(&f, &i, 4); return f; }
using std::min; using std::max;
constexpr int kLeaf = 32;
static inline float d2_point(const float4 &q, const float4 &p) {
    float dx = q.x - p.x, dy = q.y - p.y, dz = q.z - p.z;
    return (dx * dx + dy * dy) + dz * dz;
}
#include "radsel_structs.inc"

struct Leaf { float4 p[kLeaf]; float lb, ub; };
template <class V, class = void> struct visits_nodes : std::false_type {};
template <class V> struct visits_nodes<V, std::void_t<decltype(V::kNodes)>> : std::bool_constant<V::kNodes> {};
static std::vector<Leaf> leaves;
static long skipped = 0;
template <class V> static long walk(const float4 &q, V &v) {
    long scanned = 0;
    for (auto &L : leaves) {                       // sorted by lb: nearest first
        if (!(L.lb <= v.bound)) continue;          // pruned against the CURRENT bound
        if constexpr (visits_nodes<V>::value) {
            if (v.node_test() && v.whole(L.ub)) { ++skipped; continue; }   // the whole leaf lies inside the bracket's inner edge
        }
        v.scan_leaf(q, L.p); ++scanned;
    }
    return scanned;
}

int main(int argc, char **argv) {
    int seed = argc > 1 ? atoi(argv[1]) : 1;
    srand(seed);
    long total_fail = 0, cases = 0, wides = 0; double sc_full = 0, sc_new = 0;
    for (int rep = 0; rep < 400; ++rep) {
        const int n = 200 + rand() % 60000;
        const int mode = rand() % 4;               // 0 surface, 1 volume, 2 lattice ties, 3 duplicates pile
        std::vector<float4> pts(n);
        for (int i = 0; i < n; ++i) {
            float x = rand() / (float)RAND_MAX - 0.5f, y = rand() / (float)RAND_MAX - 0.5f, z = rand() / (float)RAND_MAX - 0.5f;
            if (mode == 0) z *= 0.01f;
            if (mode == 2) { x = roundf(x * 16) / 16; y = roundf(y * 16) / 16; z = roundf(z * 4) / 4; }
            if (mode == 3 && i % 3 == 0) { x = 0.01f; y = 0.02f; z = 0.f; }
            pts[i] = {x, y, z, __int_as_float(i)};
        }
        const float4 q = mode == 3 ? float4{0.01f, 0.02f, 0.f, 0.f} : pts[rand() % n];
        float r = 0.02f + 0.6f * (rand() / (float)RAND_MAX);
        if (mode == 3 && rep % 2) r = 0.f;
        const float rb = r * r;
        int max_nn = 1 + rand() % 3000;
        if (rep % 7 == 0) max_nn = 1 + rand() % 70;
        // leaves: chunks of 32 in a crude spatial order (sort by distance to q with noise) with a valid lower bound
        std::vector<std::pair<float,int>> ord(n);
        for (int i = 0; i < n; ++i) ord[i] = {d2_point(q, pts[i]) * (1.f + 0.5f * (rand() / (float)RAND_MAX)), i};
        std::sort(ord.begin(), ord.end());
        leaves.clear();
        for (int i = 0; i < n; i += kLeaf) {
            Leaf L; L.lb = INFINITY; L.ub = 0.f;
            for (int j = 0; j < kLeaf; ++j) {
                if (i + j < n) { L.p[j] = pts[ord[i + j].second]; L.lb = fminf(L.lb, d2_point(q, L.p[j])); L.ub = fmaxf(L.ub, d2_point(q, L.p[j])); }
                else L.p[j] = {INFINITY, INFINITY, INFINITY, __int_as_float(-1)};
            }
            leaves.push_back(L);
        }
        std::sort(leaves.begin(), leaves.end(), [](const Leaf &a, const Leaf &b) { return a.lb < b.lb; });
        // brute force
        std::vector<float> d(n);
        for (int i = 0; i < n; ++i) d[i] = d2_point(q, pts[i]);
        std::vector<float> in;
        for (int i = 0; i < n; ++i) if (d[i] <= rb) in.push_back(d[i]);
        std::sort(in.begin(), in.end());
        const bool need_ref = (long)in.size() > max_nn;
        const float T_ref = need_ref ? in[max_nn - 1] : rb;
        // ---- the kernel flow for one lane ----
        unsigned hist[kSelBins * 32];
        const float scale = rb > 0.f ? __fdiv_rn((float)kSelBins, __fsqrt_rn(rb)) : 0.f;
        int sel[kSelLevels] = {0, 0, 0, 0};
        int lvl = 0, kprime = max_nn, nb = 0;
        bool need = true, first = true; float t_out = rb; int err = 0; bool tie_pass = false;
        long scanned = 0;
        while (first || (need && nb > kSelKeep && lvl < kSelLevels)) {
            const bool mine = first ? true : (need && nb > kSelKeep && lvl < kSelLevels);
            for (int b = 0; b < kSelBins; ++b) hist[b * 32] = 0;
            HistV h; h.scale = scale; h.lvl = lvl; h.hist = hist;
            for (int l = 0; l < kSelLevels; ++l) h.sel[l] = sel[l];
            h.bound = mine ? bracket_edge(rb, scale, lvl, sel) : -1.f;
            h.inner = bracket_inner(scale, lvl, sel); h.nodes = !first;
            h.max_nn = first ? max_nn : 0; h.inb = 0; h.limit = max_nn < (1 << 30) ? max_nn + (max_nn >> 2) + 32 : INT_MAX;
            scanned += walk(q, h);
            if (first) {
                int c = 0;
                for (int b = 0; b < kSelBins; ++b) c += (int)hist[b * 32];
                nb = h.bound < rb ? INT_MAX : c;
                need = c > max_nn;
                first = false;
                if (!need) break;
            }
            if (mine && need && (lvl > 0 || nb > kSelKeep)) {
                int cum = 0, b = 0;
                for (; b < kSelBins - 1; ++b) { const int hb = (int)hist[b * 32]; if (cum + hb >= kprime) break; cum += hb; }
                kprime -= cum; nb = (int)hist[b * 32];
                for (int l = 0; l < kSelLevels; ++l) if (l == lvl) sel[l] = b;
                ++lvl;
            }
        }
        if (need) {
            KeepV kv; kv.bound = bracket_edge(rb, scale, lvl, sel); kv.scale = scale; kv.lvl = lvl; kv.vals = reinterpret_cast<float *>(hist);
            kv.cnt = 0; kv.vmin = INFINITY; kv.vmax = -INFINITY; kv.inner = bracket_inner(scale, lvl, sel); kv.nodes = true;
            for (int l = 0; l < kSelLevels; ++l) kv.sel[l] = sel[l];
            scanned += walk(q, kv);
            if (kv.cnt <= kSelKeep) {
                bool found = false;
                for (int i = 0; i < kv.cnt; ++i) {
                    const float vi = kv.vals[i * 32]; int lt = 0, le = 0;
                    for (int j = 0; j < kv.cnt; ++j) { const float vj = kv.vals[j * 32]; lt += vj < vi; le += vj <= vi; }
                    if (lt < kprime && kprime <= le) { t_out = vi; tie_pass = le > kprime; found = true; break; }
                }
                if (!found) err |= 4;
            } else if (kv.vmin == kv.vmax) { t_out = kv.vmin; tie_pass = true; }
            else err |= 1;
        }
        int i_out = INT_MAX;
        if (need && !err && tie_pass) {
            TieV tv; tv.bound = t_out; tv.less = 0; tv.tied = 0; tv.idx = reinterpret_cast<int *>(hist);
            scanned += walk(q, tv);
            const int m = max_nn - tv.less;
            bool wide = false; int mleft = 0;
            if (tv.tied > m) {
                if (m <= kSelKeep) {
                    const int kept = min(tv.tied, kSelKeep);
                    for (int i = 0; i < kept; ++i) {
                        const int vi = tv.idx[i * 32]; int lt = 0;
                        for (int j = 0; j < kept; ++j) lt += tv.idx[j * 32] < vi ? 1 : 0;
                        if (lt == m - 1) { i_out = vi; break; }
                    }
                } else { wide = true; mleft = m; }
            }
            if (wide) {
                unsigned prefix = 0;
                for (int l = 0; l < 6; ++l) {
                    const int shift = 30 - 6 * l;
                    for (int b = 0; b < kSelBins; ++b) hist[b * 32] = 0;
                    TieIdxV iv; iv.bound = t_out; iv.shift = shift; iv.prefix = prefix; iv.hist = hist;
                    scanned += walk(q, iv);
                    int cum = 0, b = 0;
                    for (; b < kSelBins - 1; ++b) { const int hb = (int)hist[b * 32]; if (cum + hb >= mleft) break; cum += hb; }
                    mleft -= cum; prefix = (prefix << 6) | (unsigned)b;
                }
                i_out = (int)prefix; ++wides;
            }
        }
        if (need && !err) {                           // the rule must admit exactly the (d2, index)-ordered max_nn nearest
            std::vector<std::pair<float, int>> all;
            for (int i = 0; i < n; ++i) if (d[i] <= rb) all.push_back({d[i], i});
            std::sort(all.begin(), all.end());
            long acc = 0; bool ok = true;
            for (size_t i = 0; i < all.size(); ++i) {
                const bool a = all[i].first < t_out || (all[i].first == t_out && all[i].second <= i_out);
                acc += a; if (a != ((long)i < max_nn)) ok = false;
            }
            if (!ok || acc != max_nn) { ++total_fail; printf("RULE FAIL rep %d mode %d max_nn %d admitted %ld I %d\n", rep, mode, max_nn, acc, i_out); }
        }
        ++cases;
        sc_new += scanned; sc_full += (double)std::count_if(leaves.begin(), leaves.end(), [&](const Leaf &L) { return L.lb <= rb; });
        if (need != need_ref || (err & 4) || (!err && t_out != T_ref)) {
            ++total_fail;
            printf("FAIL rep %d mode %d n %d r %g max_nn %d in %zu: need %d/%d T %.9g/%.9g err %d lvl %d\n", rep, mode, n, r, max_nn, in.size(), need, need_ref, t_out, T_ref, err, lvl);
        } else if (err) printf("note rep %d mode %d: err %d (bracket overflow)\n", rep, mode, err);
        (void)tie_pass;
    }
    printf("wide tie selects %ld; leaves skipped inside the inner edge %ld; ", wides, skipped);
    printf("cases %ld fails %ld; leaves scanned (all passes) %.0f vs one full-radius pass %.0f\n", cases, total_fail, sc_new, sc_full);
    return total_fail != 0;
}
