// sim_tiles.cpp — CPU model of k_tile_sel's pass 1 (developer tool, no product code): for a cloud in a raw float32 [n,3]
// file it builds the leaf order (Morton or Hilbert curve, 16 bits per axis, isolated-point split like build.cu), then for a
// sample of tiles replays the home-leaf bound, the candidate-leaf list, the per-lane survivor masks and the drain rounds.
// Prints leaves listed / scanned per tile, survivor scorings, inserts, and drain rounds when P consecutive leaves are
// pooled before draining.   g++ -O2 -fopenmp -o /tmp/sim_tiles tools/sim_tiles.cpp
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
typedef unsigned long long u64;
struct P3 { float x, y, z; };
static const int K = 20;

static u64 spread(unsigned v) {
    u64 x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull; x = (x | x << 16) & 0x1f0000ff0000ffull; x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull; x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
static u64 morton(unsigned x, unsigned y, unsigned z) { return spread(x) | spread(y) << 1 | spread(z) << 2; }
static u64 hilbert(unsigned x, unsigned y, unsigned z, int b) {
    unsigned X[3] = {x, y, z};
    const unsigned M = 1u << (b - 1);
    for (unsigned Q = M; Q > 1; Q >>= 1) {
        const unsigned Pm = Q - 1;
        for (int i = 0; i < 3; ++i) {
            if (X[i] & Q) X[0] ^= Pm;
            else { const unsigned t = (X[0] ^ X[i]) & Pm; X[0] ^= t; X[i] ^= t; }
        }
    }
    for (int i = 1; i < 3; ++i) X[i] ^= X[i - 1];
    unsigned t = 0;
    for (unsigned Q = M; Q > 1; Q >>= 1) if (X[2] & Q) t ^= Q - 1;
    for (int i = 0; i < 3; ++i) X[i] ^= t;
    return spread(X[2]) | spread(X[1]) << 1 | spread(X[0]) << 2;
}
static inline float d2pp(const P3 &a, const P3 &b) { float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z; return (dx * dx + dy * dy) + dz * dz; }
struct Box { P3 lo, hi; };
static inline float d2pb(const P3 &q, const Box &b) {
    float dx = fmaxf(fmaxf(b.lo.x - q.x, q.x - b.hi.x), 0.f), dy = fmaxf(fmaxf(b.lo.y - q.y, q.y - b.hi.y), 0.f), dz = fmaxf(fmaxf(b.lo.z - q.z, q.z - b.hi.z), 0.f);
    return (dx * dx + dy * dy) + dz * dz;
}
static inline float d2bb(const Box &a, const Box &b) {
    float dx = fmaxf(fmaxf(b.lo.x - a.hi.x, a.lo.x - b.hi.x), 0.f), dy = fmaxf(fmaxf(b.lo.y - a.hi.y, a.lo.y - b.hi.y), 0.f), dz = fmaxf(fmaxf(b.lo.z - a.hi.z, a.lo.z - b.hi.z), 0.f);
    return (dx * dx + dy * dy) + dz * dz;
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: sim_tiles cloud.f32 curve(0 morton,1 hilbert) [sample_stride] [leaf]\n"); return 1; }
    const int curve = atoi(argv[2]);
    const int sstride = argc > 3 ? atoi(argv[3]) : 37;
    const int kbits = argc > 4 ? atoi(argv[4]) : 16;      // bits per axis actually kept in the key (low bits dropped)
    const int kdrop = argc > 5 ? atoi(argv[5]) : 0;       // extra low KEY bits dropped after interleaving
    FILE *f = fopen(argv[1], "rb");
    fseek(f, 0, SEEK_END); const size_t n = ftell(f) / 12; fseek(f, 0, SEEK_SET);
    std::vector<P3> pts(n);
    if (fread(pts.data(), 12, n, f) != n) return 2;
    fclose(f);
    P3 lo = {1e30f, 1e30f, 1e30f}, hi = {-1e30f, -1e30f, -1e30f};
    for (auto &p : pts) { lo.x = fminf(lo.x, p.x); lo.y = fminf(lo.y, p.y); lo.z = fminf(lo.z, p.z); hi.x = fmaxf(hi.x, p.x); hi.y = fmaxf(hi.y, p.y); hi.z = fmaxf(hi.z, p.z); }
    const float ext = fmaxf(fmaxf(hi.x - lo.x, hi.y - lo.y), hi.z - lo.z);
    const float inv = 65536.f / (ext * 1.0001f);
    std::vector<std::pair<u64, unsigned>> kv(n);
#pragma omp parallel for
    for (size_t i = 0; i < n; ++i) {
        unsigned qx = (unsigned)fminf(fmaxf((pts[i].x - lo.x) * inv, 0.f), 65535.f), qy = (unsigned)fminf(fmaxf((pts[i].y - lo.y) * inv, 0.f), 65535.f),
                 qz = (unsigned)fminf(fmaxf((pts[i].z - lo.z) * inv, 0.f), 65535.f);
        qx >>= (16 - kbits); qy >>= (16 - kbits); qz >>= (16 - kbits);
        kv[i] = {(curve ? hilbert(qx, qy, qz, kbits) : morton(qx, qy, qz)) >> kdrop, (unsigned)i};
    }
    std::stable_sort(kv.begin(), kv.end(), [](auto &a, auto &b) { return a.first < b.first; });
    // isolated-point split (k_flag_dense)
    std::vector<unsigned> order; order.reserve(n);
    {
        std::vector<unsigned> sparse;
        for (size_t w = 0; w < n; w += 32) {
            float iso[32]; int nv = (int)std::min<size_t>(32, n - w);
            for (int l = 0; l < nv; ++l) {
                size_t i = w + l; float gp = INFINITY, gn = INFINITY;
                if (i > 0) gp = d2pp(pts[kv[i].second], pts[kv[i - 1].second]);
                if (i + 1 < n) gn = d2pp(pts[kv[i].second], pts[kv[i + 1].second]);
                iso[l] = fminf(gp, gn);
            }
            float s[32]; memcpy(s, iso, sizeof(float) * nv); std::sort(s, s + nv);
            const float med = s[nv >> 1];
            for (int l = 0; l < nv; ++l) (iso[l] > 64.f * med ? sparse : order).push_back(kv[w + l].second);
        }
        printf("isolated points moved: %zu\n", sparse.size());
        order.insert(order.end(), sparse.begin(), sparse.end());
    }
    const int nleaf = (int)((n + 31) / 32);
    std::vector<P3> sp((size_t)nleaf * 32, P3{INFINITY, INFINITY, INFINITY});
    std::vector<Box> box(nleaf);
    for (int L = 0; L < nleaf; ++L) {
        Box b = {{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
        for (int l = 0; l < 32 && (size_t)L * 32 + l < n; ++l) {
            const P3 p = pts[order[(size_t)L * 32 + l]]; sp[(size_t)L * 32 + l] = p;
            b.lo.x = fminf(b.lo.x, p.x); b.lo.y = fminf(b.lo.y, p.y); b.lo.z = fminf(b.lo.z, p.z);
            b.hi.x = fmaxf(b.hi.x, p.x); b.hi.y = fmaxf(b.hi.y, p.y); b.hi.z = fmaxf(b.hi.z, p.z);
        }
        box[L] = b;
    }
    // coarse level for the listing: boxes of 32 leaves
    const int nl1 = (nleaf + 31) / 32;
    std::vector<Box> box1(nl1);
    for (int g = 0; g < nl1; ++g) {
        Box b = {{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
        for (int L = g * 32; L < std::min(nleaf, g * 32 + 32); ++L) {
            b.lo.x = fminf(b.lo.x, box[L].lo.x); b.lo.y = fminf(b.lo.y, box[L].lo.y); b.lo.z = fminf(b.lo.z, box[L].lo.z);
            b.hi.x = fmaxf(b.hi.x, box[L].hi.x); b.hi.y = fmaxf(b.hi.y, box[L].hi.y); b.hi.z = fmaxf(b.hi.z, box[L].hi.z);
        }
        box1[g] = b;
    }
    double s_listed = 0, s_scanned = 0, s_surv = 0, s_ins = 0, s_r1 = 0, s_r2 = 0, s_r4 = 0, s_rall = 0, s_home_r = 0, s_lone = 0, s_logmax = 0;
    long tiles = 0;
    std::vector<int> listed_hist;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : s_listed, s_scanned, s_surv, s_ins, s_r1, s_r2, s_r4, s_rall, s_home_r, s_lone, s_logmax, tiles)
    for (int t = 0; t < nleaf - 1; t += sstride) {
        const P3 *q = &sp[(size_t)t * 32];
        float key[32][K]; float bound[32]; int logc[32];
        for (int l = 0; l < 32; ++l) { for (int i = 0; i < K; ++i) key[l][i] = INFINITY; bound[l] = INFINITY; logc[l] = 0; }
        auto insert = [&](int l, float c) { for (int i = K - 1; i > 0; --i) key[l][i] = fmaxf(key[l][i - 1], fminf(key[l][i], c)); key[l][0] = fminf(key[l][0], c); bound[l] = key[l][K - 1]; };
        int home_rounds = 0;
        for (int h = 0; h < 3; ++h) {
            const int L = t + (h == 0 ? 0 : (h == 1 ? -1 : 1));
            if (L < 0 || L >= nleaf) continue;
            int mx = 0;
            for (int l = 0; l < 32; ++l) {
                int c = 0; const float b0 = bound[l];
                for (int j = 0; j < 32; ++j) { const float d = d2pp(q[l], sp[(size_t)L * 32 + j]); if (d <= b0) { ++c; if (d < key[l][K - 1]) { insert(l, d); ++s_ins; } } }
                mx = std::max(mx, c);
            }
            home_rounds += mx;
        }
        float sb[32]; memcpy(sb, bound, sizeof(sb)); std::sort(sb, sb + 32);
        const float med = sb[16];
        bool act[32];
        for (int l = 0; l < 32; ++l) { act[l] = bound[l] <= 16.f * med; if (!act[l]) { bound[l] = -1.f; ++s_lone; } }
        float rmax = 0.f; for (int l = 0; l < 32; ++l) rmax = fmaxf(rmax, bound[l]);
        std::vector<int> list;
        for (int g = 0; g < nl1; ++g) {
            if (d2bb(box[t], box1[g]) > rmax) continue;
            for (int L = g * 32; L < std::min(nleaf, g * 32 + 32); ++L)
                if ((L < t - 1 || L > t + 1) && d2bb(box[t], box[L]) <= rmax) list.push_back(L);
        }
        // replay with pooling P = 1 (drives the bounds), and count what pooled rounds would be with the same masks (approximation:
        // pooled masks use the bound at group start, so they are recomputed in separate replays below)
        int rounds[3] = {0, 0, 0};
        const int Ps[3] = {1, 2, 4};
        double surv1 = 0; int scanned1 = 0, logmax1 = 0;
        for (int v = 0; v < 3; ++v) {
            float k2[32][K], b2[32]; int lc[32];
            memcpy(k2, key, sizeof(k2)); memcpy(b2, bound, sizeof(b2)); memset(lc, 0, sizeof(lc));
            const int Pn = Ps[v];
            size_t li = 0; int scanned = 0; double surv = 0;
            while (li < list.size()) {
                int grp[4], ng = 0;
                while (li < list.size() && ng < Pn) {
                    const int L = list[li++]; bool any = false;
                    for (int l = 0; l < 32 && !any; ++l) any = b2[l] >= 0.f && d2pb(q[l], box[L]) <= b2[l];
                    if (any) grp[ng++] = L;
                }
                if (!ng) continue;
                scanned += ng;
                int mx = 0;
                for (int l = 0; l < 32; ++l) {
                    if (b2[l] < 0.f) continue;
                    const float b0 = b2[l]; int c = 0;
                    for (int g = 0; g < ng; ++g)
                        for (int j = 0; j < 32; ++j) {
                            const float d = d2pp(q[l], sp[(size_t)grp[g] * 32 + j]);
                            if (d <= b0) {
                                ++c;
                                if (d <= b2[l]) { ++lc[l]; if (d < k2[l][K - 1]) { for (int i = K - 1; i > 0; --i) k2[l][i] = fmaxf(k2[l][i - 1], fminf(k2[l][i], d)); k2[l][0] = fminf(k2[l][0], d); b2[l] = k2[l][K - 1]; if (v == 0) ++s_ins; } }
                            }
                        }
                    surv += c; mx = std::max(mx, c);
                }
                rounds[v] += mx;
            }
            if (v == 0) { surv1 = surv; scanned1 = scanned; for (int l = 0; l < 32; ++l) logmax1 = std::max(logmax1, lc[l]); }
        }
        s_listed += list.size(); s_scanned += scanned1; s_surv += surv1; s_r1 += rounds[0]; s_r2 += rounds[1]; s_r4 += rounds[2];
        s_home_r += home_rounds; s_logmax += logmax1; ++tiles;
    }
    printf("curve=%s n=%zu tiles sampled=%ld\n", curve ? "hilbert" : "morton", n, tiles);
    printf("per tile: listed %.1f  scanned %.1f  listed-leaf survivor scorings %.0f (%.1f per query)  inserts %.0f\n", s_listed / tiles, s_scanned / tiles,
           s_surv / tiles, s_surv / tiles / 32, s_ins / tiles);
    printf("drain rounds per tile: home %.1f | listed P=1 %.1f  P=2 %.1f  P=4 %.1f | ideal (survivors/32) %.1f | lone lanes %.2f | max log %.1f\n", s_home_r / tiles,
           s_r1 / tiles, s_r2 / tiles, s_r4 / tiles, s_surv / tiles / 32, s_lone / tiles, s_logmax / tiles);
    return 0;
}
