// test_adapters.cpp — exercises adapters/cloudclean_gpu.hpp (the reference-named C++ host layer) on cuda:0 and
// checks every adapter against brute-force CPU restatements written in this file (contract arithmetic,
// SURVEY.md §8c).  Built by `make adapters`, run by tests/test_gpu_adapters.py (-m gpu).  Exit code 0 = all green.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <chrono>
#include <thread>
#include <future>
#include <atomic>
#include <cstdlib>
#include <limits>
#include <numeric>

#include "cloudclean_gpu.hpp"

using namespace ccgpu;

static int g_fail = 0;
#define CHECK(cond, ...)                                  \
    do {                                                  \
        if (!(cond)) {                                    \
            ++g_fail;                                     \
            std::printf("FAIL %s:%d: ", __FILE__, __LINE__); \
            std::printf(__VA_ARGS__);                     \
            std::printf("\n");                            \
        }                                                 \
    } while (0)

static uint32_t g_rng = 12345u;
static float urand() {  // LCG, [0,1)
    g_rng = g_rng * 1664525u + 1013904223u;
    return (float)(g_rng >> 8) * (1.0f / 16777216.0f);
}

// contract distance: ((dx*dx + dy*dy) + dz*dz) in fp32, no fma (this file is compiled with -ffp-contract=off)
static float d2f(const PointXYZI &a, const PointXYZI &b) {
    volatile float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
    volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz;
    volatile float s = xx + yy;
    return s + zz;
}

// fp64 cyclic Jacobi of a symmetric 3x3 (a = xx xy xz yy yz zz); eigenvalues descending, v2 = eigenvector of the smallest
static void eig3d(const double a[6], double ev[3], double vmin[3]) {
    double A[3][3] = {{a[0], a[1], a[2]}, {a[1], a[3], a[4]}, {a[2], a[4], a[5]}};
    double V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = std::fabs(A[0][1]) + std::fabs(A[0][2]) + std::fabs(A[1][2]);
        if (off < 1e-300) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (A[p][q] == 0.0) continue;
                double th = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                double t = (th >= 0 ? 1.0 : -1.0) / (std::fabs(th) + std::sqrt(th * th + 1.0));
                double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int r = 0; r < 3; ++r) {
                    double arp = A[r][p], arq = A[r][q];
                    A[r][p] = c * arp - s * arq; A[r][q] = s * arp + c * arq;
                }
                for (int r = 0; r < 3; ++r) {
                    double apr = A[p][r], aqr = A[q][r];
                    A[p][r] = c * apr - s * aqr; A[q][r] = s * apr + c * aqr;
                }
                for (int r = 0; r < 3; ++r) {
                    double vrp = V[r][p], vrq = V[r][q];
                    V[r][p] = c * vrp - s * vrq; V[r][q] = s * vrp + c * vrq;
                }
            }
    }
    int o[3] = {0, 1, 2};
    std::sort(o, o + 3, [&](int i, int j) { return A[i][i] > A[j][j]; });
    for (int i = 0; i < 3; ++i) ev[i] = A[o[i]][o[i]];
    for (int r = 0; r < 3; ++r) vmin[r] = V[r][o[2]];
}

// PCA of a neighbour list in fp64: eigenvalues (desc) of S/(n-1) or S/n, eigenvector of the smallest
static void pca_ref(const std::vector<PointXYZI> &pts, const std::vector<int> &nb, int cov_norm, double ev[3], double vmin[3]) {
    double mu[3] = {0, 0, 0};
    for (int j : nb) { mu[0] += pts[j].x; mu[1] += pts[j].y; mu[2] += pts[j].z; }
    for (double &m : mu) m /= (double)nb.size();
    double a[6] = {0, 0, 0, 0, 0, 0};
    for (int j : nb) {
        double x = pts[j].x - mu[0], y = pts[j].y - mu[1], z = pts[j].z - mu[2];
        a[0] += x * x; a[1] += x * y; a[2] += x * z; a[3] += y * y; a[4] += y * z; a[5] += z * z;
    }
    double div = cov_norm == CCG_COV_NM1 ? (double)nb.size() - 1.0 : (cov_norm == CCG_COV_N ? (double)nb.size() : 1.0);
    for (double &v : a) v /= div;
    eig3d(a, ev, vmin);
}

// neighbours of i ordered by (d2, index)
static std::vector<std::pair<float, int>> sorted_neighbours(const std::vector<PointXYZI> &pts, int i) {
    std::vector<std::pair<float, int>> v(pts.size());
    for (size_t j = 0; j < pts.size(); ++j) v[j] = {d2f(pts[i], pts[j]), (int)j};
    std::sort(v.begin(), v.end());
    return v;
}

int main() {
    // organised patch: W columns x H rows, a floor with a step and noise, 5 % dropouts, a few far isolated points
    const int W = 72, H = 56;
    std::vector<PointXYZI> pts;
    std::vector<int> c2g;
    for (int g = 0; g < W * H; ++g) {
        const int col = g / H, row = g % H;
        if (urand() < 0.05f) continue;
        PointXYZI p{};
        p.x = 0.05f * col + 0.004f * (urand() - 0.5f);
        p.y = 0.05f * row + 0.004f * (urand() - 0.5f);
        p.z = (col > W / 2 ? 0.3f : 0.0f) + 0.002f * (urand() - 0.5f);
        p.w = 1.0f; p.intensity = urand();
        if (urand() < 0.004f) p.z += 2.0f + urand();      // isolated outliers above the floor
        pts.push_back(p); c2g.push_back(g);
    }
    const size_t n = pts.size();
    std::printf("adapters test: n=%zu\n", n);

    Session s(0);
    CHECK(s.ok(), "Session: %s", s.lastError().c_str());
    if (!s.ok()) return 2;

    std::vector<std::vector<std::pair<float, int>>> nbrs(n);
    for (size_t i = 0; i < n; ++i) nbrs[i] = sorted_neighbours(pts, (int)i);

    // ---- getPCA / getHist (cv.h:76-116, cv.cpp:249-308): strict radius, cap = max_nn nearest, <3 -> (0,0,1) ----
    {
        const double radius = 0.12; const unsigned max_nn = 20;
        const float r2 = (float)(radius * radius);
        auto pca = getPCA(s, &pts, radius, max_nn);                  // reference-style call: cloud pointer
        auto pca2 = getPCA(s, pts.data(), n, radius, max_nn);        // raw pointer + size
        auto hist = getHist(s, pts.data(), n, radius, max_nn);
        CHECK(pca->size() == n && pca2->size() == n && hist->size() == n, "getPCA/getHist sizes (%s)", s.lastError().c_str());
        double worst = 0, worst_h = 0; int markers = 0;
        for (size_t i = 0; i < n && pca->size() == n; ++i) {
            std::vector<int> nb;
            for (auto &c : nbrs[i]) { if (!(c.first < r2) || nb.size() == max_nn) break; nb.push_back(c.second); }
            const Vector3f &e = (*pca)[i];
            CHECK(std::memcmp(&e, &(*pca2)[i], 12) == 0, "getPCA overloads differ at %zu", i);
            if (nb.size() < 3) {
                ++markers;
                CHECK(e[0] == 0.f && e[1] == 0.f && e[2] == 1.f, "marker at %zu: %g %g %g", i, e[0], e[1], e[2]);
                continue;
            }
            double ev[3], vm[3];
            pca_ref(pts, nb, CCG_COV_NM1, ev, vm);
            double nrm = std::sqrt(ev[0] * ev[0] + ev[1] * ev[1] + ev[2] * ev[2]);
            for (int d = 0; d < 3; ++d) {
                worst = std::max(worst, std::fabs(e[d] - ev[d]) / ev[0]);
                worst_h = std::max(worst_h, std::fabs((*hist)[i][d] - ev[d] / nrm));
            }
        }
        std::printf("getPCA: max |d lambda|/lambda0 = %.3g, getHist max abs err = %.3g, markers = %d\n", worst, worst_h, markers);
        CHECK(worst <= 1e-4, "getPCA eigenvalues off by %g", worst);
        CHECK(worst_h <= 1e-4, "getHist off by %g", worst_h);
        CHECK(markers > 0, "the test cloud should hold isolated points");

        auto f7 = eigenFeatures(s, *pca, true);
        auto ratio = pcaEigenValueRatio(s, *pca);
        CHECK(f7.size() == 7 * n && ratio.size() == n, "feature sizes");
        double wf = 0;
        for (size_t i = 0; i < n && f7.size() == 7 * n; ++i) {
            const float l0 = (*pca)[i][0], l1 = (*pca)[i][1], l2 = (*pca)[i][2];
            if (!(l0 > 0.f)) continue;
            const float ref[4] = {(l0 - l2) / l0, (l1 - l2) / l0, l2 / l0, (l0 - l1) / l0};   // markov.cpp:451-463
            for (int d = 0; d < 4; ++d) wf = std::max(wf, (double)std::fabs(f7[7 * i + d] - ref[d]));
            if (!(l0 == 0.f && l1 == 0.f && l2 == 1.f)) {
                const float rr = l0 > l2 ? l1 / l0 : 1.f;                                      // featureeval.cpp:1107-1135
                wf = std::max(wf, (double)std::fabs(ratio[i] - rr));
            }
        }
        CHECK(wf <= 1e-6, "eigen features off by %g", wf);
    }

    // ---- OutlierFilter::filter (outlierfilter.cpp:157-173): inclusive radius count incl. self < min_neighbours ----
    {
        const double radius_cm = 20.0; const int min_nb = 10;
        const double r2 = (radius_cm / 100.0) * (radius_cm / 100.0);
        auto got = outlierFilter(s, pts.data(), n, radius_cm, min_nb, /*octree_prefilter=*/false);
        std::vector<int> want;
        for (size_t i = 0; i < n; ++i) {
            int c = 0;
            for (auto &p : nbrs[i]) { if (!((double)p.first <= r2)) break; ++c; }
            if (c < min_nb) want.push_back((int)i);
        }
        std::printf("outlierFilter: %zu flagged (expected %zu)\n", got->size(), want.size());
        CHECK(*got == want, "outlierFilter index list differs");
        CHECK(!want.empty(), "the test cloud should hold radius outliers");
        auto pre = outlierFilter(s, pts.data(), n, radius_cm, min_nb, true);
        CHECK(std::includes(want.begin(), want.end(), pre->begin(), pre->end()), "prefiltered list must be a subset");
    }

    // ---- statistical outliers: mean distance to mean_k neighbours (self dropped), mu + alpha*sigma in fp64 ----
    {
        const int mean_k = 12; const double alpha = 1.0;
        std::vector<float> md(n);
        double sum = 0, sq = 0;
        for (size_t i = 0; i < n; ++i) {
            double acc = 0;
            for (int j = 1; j <= mean_k; ++j) acc += (double)std::sqrt(nbrs[i][j].first);
            md[i] = (float)(acc / mean_k);
            sum += md[i]; sq += (double)md[i] * (double)md[i];
        }
        const double mean = sum / (double)n, var = (sq - sum * sum / (double)n) / ((double)n - 1.0);
        const double thr = mean + alpha * std::sqrt(var);
        std::vector<int> want;
        for (size_t i = 0; i < n; ++i) if ((double)md[i] > thr) want.push_back((int)i);
        auto got = statisticalOutliers(s, pts.data(), n, mean_k, alpha);
        std::printf("statisticalOutliers: %zu flagged (expected %zu)\n", got->size(), want.size());
        CHECK(*got == want, "statistical outlier list differs");
    }

    // ---- knnNormals: kNN(k)+PCA normal = eigenvector of the smallest eigenvalue (S/n), |cos| >= 1-1e-5 ----
    {
        const int k = 16;
        const float origin[3] = {1.8f, 1.4f, 5.0f};
        auto nrm = knnNormals(s, pts.data(), n, k, origin);
        CHECK(nrm.size() == n, "knnNormals size (%s)", s.lastError().c_str());
        double worst = 0, worst_c = 0; int checked = 0, flipped_wrong = 0;
        for (size_t i = 0; i < n && nrm.size() == n; ++i) {
            std::vector<int> nb;
            for (int j = 0; j < k; ++j) nb.push_back(nbrs[i][j].second);
            double ev[3], vm[3];
            pca_ref(pts, nb, CCG_COV_N, ev, vm);
            if (ev[1] - ev[2] <= 1e-3 * ev[0]) continue;          // no unique normal
            ++checked;
            const Normal &o = nrm[i];
            const double c = std::fabs(o.normal_x * vm[0] + o.normal_y * vm[1] + o.normal_z * vm[2]);
            worst = std::max(worst, 1.0 - c);
            worst_c = std::max(worst_c, std::fabs(o.curvature - ev[2] / (ev[0] + ev[1] + ev[2])));
            const double tov = (origin[0] - pts[i].x) * o.normal_x + (origin[1] - pts[i].y) * o.normal_y + (origin[2] - pts[i].z) * o.normal_z;
            if (tov < 0) ++flipped_wrong;
        }
        std::printf("knnNormals: checked %d, max 1-|cos| = %.3g, max curvature err = %.3g\n", checked, worst, worst_c);
        CHECK(checked > (int)n / 2, "too few well-defined normals (%d)", checked);
        CHECK(worst <= 1e-5, "normals off: 1-|cos| = %g", worst);
        CHECK(worst_c <= 1e-4, "curvature off by %g", worst_c);
        CHECK(flipped_wrong == 0, "%d normals point away from the viewpoint", flipped_wrong);
    }

    // ---- estimateNormals (normalestimation.cpp:268-470): organised fan normals + kNN(50) fallback ----
    {
        auto nrm = estimateNormals(s, pts.data(), n, c2g, W, H, true);
        CHECK(nrm.size() == n, "estimateNormals size (%s)", s.lastError().c_str());
        int nan = 0, unit = 0, upright = 0;
        for (size_t i = 0; i < nrm.size(); ++i) {
            const Normal &o = nrm[i];
            if (std::isnan(o.normal_x)) { ++nan; continue; }
            const double len = std::sqrt((double)o.normal_x * o.normal_x + (double)o.normal_y * o.normal_y + (double)o.normal_z * o.normal_z);
            if (std::fabs(len - 1.0) < 1e-4) ++unit;
            if (std::fabs(o.normal_z) > 0.9) ++upright;
            CHECK(o.curvature == 0.f, "curvature must stay 0 (normalestimation.cpp leaves it untouched)");
        }
        std::printf("estimateNormals: %d NaN, %d unit, %d upright of %zu\n", nan, unit, upright, n);
        CHECK(nan == 0, "%d normals left NaN after the fallback", nan);
        CHECK(unit > (int)(0.9 * n), "fan normals must be unit length (%d of %zu)", unit, n);
        CHECK(upright > (int)(0.8 * n), "a floor scan should have mostly vertical normals (%d of %zu)", upright, n);
        std::vector<int> bad_map(n, 0);
        bad_map.pop_back();
        CHECK(estimateNormals(s, pts.data(), n, bad_map, W, H).empty(), "a short cloud_to_grid map must be refused");
    }

    // ---- octreeDownsample (utils.h:16-62): voxel means of all 8 floats + big -> small map ----
    {
        const float res = 0.2f;
        std::vector<int> sub;
        auto small = octreeDownsample(s, pts.data(), n, res, sub);
        CHECK(!small.empty() && sub.size() == n, "octreeDownsample returned nothing (%s)", s.lastError().c_str());
        std::vector<double> sx(small.size(), 0), sy(small.size(), 0), sz(small.size(), 0), si(small.size(), 0);
        std::vector<int> cnt(small.size(), 0);
        bool in_range = true, same_voxel = true;
        const float h = res * 0.5f;
        for (size_t i = 0; i < n && !small.empty(); ++i) {
            const int v = sub[i];
            if (v < 0 || v >= (int)small.size()) { in_range = false; continue; }
            sx[v] += pts[i].x; sy[v] += pts[i].y; sz[v] += pts[i].z; si[v] += pts[i].intensity; cnt[v]++;
        }
        // points of one voxel share the lattice cell floor((p - (p0 - res/2)) / res)
        std::vector<long long> cell(small.size(), -1);
        for (size_t i = 0; i < n && in_range; ++i) {
            long long kx = (long long)std::floor((pts[i].x - (pts[0].x - h)) / res), ky = (long long)std::floor((pts[i].y - (pts[0].y - h)) / res),
                      kz = (long long)std::floor((pts[i].z - (pts[0].z - h)) / res);
            long long key = ((kx & 0x1fffff) << 42) | ((ky & 0x1fffff) << 21) | (kz & 0x1fffff);
            if (cell[sub[i]] == -1) cell[sub[i]] = key;
            else if (cell[sub[i]] != key) same_voxel = false;
        }
        CHECK(in_range, "sub_idxs out of range");
        CHECK(same_voxel, "points of different lattice cells share a voxel");
        for (size_t v = 1; v < cell.size(); ++v) CHECK(cell[v] > cell[v - 1], "voxels must be numbered by ascending key");
        double worst = 0;
        for (size_t v = 0; v < small.size() && in_range; ++v) {
            CHECK(cnt[v] > 0, "empty voxel %zu", v);
            worst = std::max(worst, std::fabs(small[v].x - sx[v] / cnt[v]));
            worst = std::max(worst, std::fabs(small[v].y - sy[v] / cnt[v]));
            worst = std::max(worst, std::fabs(small[v].z - sz[v] / cnt[v]));
            worst = std::max(worst, std::fabs(small[v].intensity - si[v] / cnt[v]));
            worst = std::max(worst, std::fabs(small[v].w - (cnt[v] + 1.0) / cnt[v]));       // as written: float[3] starts at 1
        }
        std::printf("octreeDownsample: %zu voxels from %zu points, max mean error %.3g\n", small.size(), n, worst);
        CHECK(worst <= 1e-5, "voxel means off by %g", worst);
    }

    // ---- knnGraph: MinCut::buildGraph's neighbour loop + calculateBinaryPotential (mincut.cpp:458-475,574-585) ----
    {
        const int k = 14; const double inv_sigma = 16.0;
        std::vector<int> nb; std::vector<double> w;
        CHECK(knnGraph(s, pts.data(), n, k, inv_sigma, nb, w), "knnGraph failed (%s)", s.lastError().c_str());
        int bad_idx = 0; double worst = 0;
        for (size_t i = 0; i < n; ++i)
            for (int j = 0; j < k; ++j) {
                const int t = nb[i * k + j];
                if (t != nbrs[i][j].second) { ++bad_idx; continue; }
                double d = 0.0;
                d += (pts[i].x - pts[t].x) * (pts[i].x - pts[t].x);
                d += (pts[i].y - pts[t].y) * (pts[i].y - pts[t].y);
                d += (pts[i].z - pts[t].z) * (pts[i].z - pts[t].z);
                const double ref = std::exp(-(d * inv_sigma));
                worst = std::max(worst, std::fabs(w[i * k + j] - ref) / ref);
            }
        std::printf("knnGraph: %d neighbour mismatches, max relative weight error %.3g\n", bad_idx, worst);
        CHECK(bad_idx == 0, "%d neighbour entries differ from the (d2, index) order", bad_idx);
        CHECK(worst <= 1e-14, "edge weights off by %g", worst);
    }

    // ---- stdev_dist (cv.cpp:166-230): variance of neighbour distances, inclusive radius, float sums in (d2, index) order ----
    {
        const double radius = 0.15; const int max_nn = 32;
        auto got = stdev_dist(s, pts.data(), n, radius, max_nn);
        const double r2 = radius * radius;
        int bad = 0;
        for (size_t i = 0; i < n && got->size() == n; ++i) {
            float sum = 0.f, sum_sq = 0.f; int m = 0;
            for (auto &c : nbrs[i]) {
                if (!((double)c.first <= r2) || m == max_nn) break;
                const PointXYZI &p = pts[c.second];
                volatile float dx = p.x - pts[i].x, dy = p.y - pts[i].y, dz = p.z - pts[i].z;
                volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz;
                volatile float ss = yy + zz; ss = xx + ss;         // Eigen <= 3.2 Vector3f::norm(): a0 + (a1 + a2) [EXT Redux.h]
                volatile float dist = std::sqrt((float)ss);
                sum = sum + dist;
                volatile float d2v = dist * dist;
                sum_sq = sum_sq + d2v;
                ++m;
            }
            volatile float s2 = sum * sum; volatile float q = s2 / (float)m; volatile float num = sum_sq - q;
            const float want = m > 2 ? num / ((float)m - 1.f) : 0.f;
            if (want != (*got)[i]) ++bad;
        }
        std::printf("stdev_dist: %d of %zu differ\n", bad, n);
        CHECK(bad == 0, "stdev_dist differs at %d points", bad);
    }

    // ---- principalCurvatures (pcl::PrincipalCurvaturesEstimation, markov.cpp:389-402): normals live in PointXYZINormal ----
    {
        const float origin0[3] = {0, 0, 0};
        auto nrm = knnNormals(s, pts.data(), n, 12, origin0);
        std::vector<PointXYZINormal> pn(n);
        for (size_t i = 0; i < n && nrm.size() == n; ++i) {
            std::memset(&pn[i], 0, sizeof(PointXYZINormal));
            pn[i].x = pts[i].x; pn[i].y = pts[i].y; pn[i].z = pts[i].z;
            pn[i].normal_x = nrm[i].normal_x; pn[i].normal_y = nrm[i].normal_y; pn[i].normal_z = nrm[i].normal_z;
        }
        const double radius = 0.3;
        auto pc = principalCurvatures(s, pts.data(), pn.data(), n, radius);
        CHECK(pc.size() == n, "principalCurvatures size (%s)", s.lastError().c_str());
        const float r2 = (float)(radius * radius);
        double worst_tr = 0, worst_det = 0; int checked = 0;
        for (size_t i = 0; i < n && pc.size() == n; i += 7) {
            // the two non-zero eigenvalues of the projected-normal scatter matrix are fixed by its trace and by the sum
            // of its principal 2x2 minors (the third eigenvalue is 0: everything lies in the tangent plane)
            const double nx = pn[i].normal_x, ny = pn[i].normal_y, nz = pn[i].normal_z;
            std::vector<std::array<double, 3>> pr;
            for (auto &c : nbrs[i]) {
                if (!(c.first < r2)) break;
                const PointXYZINormal &o = pn[c.second];
                const double d = nx * o.normal_x + ny * o.normal_y + nz * o.normal_z;
                pr.push_back({o.normal_x - nx * d, o.normal_y - ny * d, o.normal_z - nz * d});
            }
            double cen[3] = {0, 0, 0};
            for (auto &v : pr) for (int a = 0; a < 3; ++a) cen[a] += v[a];
            for (int a = 0; a < 3; ++a) cen[a] /= (double)pr.size();
            double C[3][3] = {{0}};
            for (auto &v : pr) for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) C[a][b] += (v[a] - cen[a]) * (v[b] - cen[b]);
            const double N = (double)pr.size();
            const double tr = (C[0][0] + C[1][1] + C[2][2]) / N;
            const double minors = (C[0][0] * C[1][1] - C[0][1] * C[0][1] + C[0][0] * C[2][2] - C[0][2] * C[0][2] + C[1][1] * C[2][2] - C[1][2] * C[1][2]) / (N * N);
            const double g1 = pc[i].pc1, g2 = pc[i].pc2;
            worst_tr = std::max(worst_tr, std::fabs(g1 + g2 - tr) / (tr + 1e-9));
            worst_det = std::max(worst_det, std::fabs(g1 * g2 - minors) / (tr * tr + 1e-12));
            ++checked;
        }
        std::printf("principalCurvatures: checked %d, max rel trace err = %.3g, max minors err = %.3g\n", checked, worst_tr, worst_det);
        CHECK(checked > 100, "too few rows checked (%d)", checked);
        CHECK(worst_tr <= 2e-4, "pc1 + pc2 off by %g (relative)", worst_tr);
        CHECK(worst_det <= 2e-4, "pc1 * pc2 off by %g (relative to trace^2)", worst_det);
    }

    // ---- fpfh (featureeval.cpp:925-932): structural checks (the numerical comparison lives in tests/test_gpu_parity.py) ----
    {
        const float origin0[3] = {0, 0, 0};
        auto nrm = knnNormals(s, pts.data(), n, 12, origin0);
        auto sig = fpfh(s, pts.data(), nrm.data(), n, 0.3);
        CHECK(sig.size() == n, "fpfh size (%s)", s.lastError().c_str());
        int full = 0, empty = 0, other = 0;
        for (size_t i = 0; i < sig.size(); ++i) {
            double g[3] = {0, 0, 0};
            for (int b = 0; b < 33; ++b) g[b / 11] += sig[i].histogram[b];
            if (std::fabs(g[0] - 100) < 1e-2 && std::fabs(g[1] - 100) < 1e-2 && std::fabs(g[2] - 100) < 1e-2) ++full;
            else if (g[0] == 0 && g[1] == 0 && g[2] == 0) ++empty;
            else ++other;
        }
        std::printf("fpfh: %d rows sum to 100 per group, %d empty (isolated points), %d other\n", full, empty, other);
        CHECK(other == 0 && full > (int)n / 2, "fpfh rows must be normalised to 100 per group or empty (%d / %d / %d)", full, empty, other);
    }

    // ---- load_ptx (pointcloud.cpp:127-330): write the test scan as a PTX file, read it back through the GPU parser ----
    {
        const char *path = "/tmp/ccgpu_adapter_test.ptx";
        FILE *f = std::fopen(path, "w");
        CHECK(f != nullptr, "cannot write %s", path);
        if (f) {
            std::fprintf(f, "%d\n%d\n0.5 0.25 -1.5\n1 0 0\n0 1 0\n0 0 1\n1 0 0 0\n0 1 0 0\n0 0 1 0\n0.5 0.25 -1.5 1\n", W, H);
            size_t next = 0;
            for (int g = 0; g < W * H; ++g) {
                if (next < n && c2g[next] == g) { std::fprintf(f, "%.9g %.9g %.9g %.9g\n", pts[next].x, pts[next].y, pts[next].z, pts[next].intensity); ++next; }
                else std::fprintf(f, "0 0 0 0.5\n");
            }
            std::fclose(f);
            PtxCloud pc = load_ptx(s, path, 1, true);
            CHECK(pc.ok, "load_ptx failed: %s", s.lastError().c_str());
            CHECK(pc.points.size() == n && pc.scan_width == W && pc.scan_height == H, "load_ptx: %zu points, %d x %d", pc.points.size(), pc.scan_width, pc.scan_height);
            CHECK(pc.sensor_origin[0] == 0.5f && pc.sensor_origin[1] == 0.25f && pc.sensor_origin[2] == -1.5f, "sensor origin");
            int bad = 0;
            for (size_t i = 0; i < pc.points.size() && i < n; ++i) {
                // %.9g round-trips a float exactly, so the parsed rows must equal the cloud bit for bit
                if (pc.points[i].x != pts[i].x || pc.points[i].y != pts[i].y || pc.points[i].z != pts[i].z || pc.points[i].intensity != pts[i].intensity) ++bad;
                if (pc.cloud_to_grid_map[i] != c2g[i]) ++bad;
                if (pc.points[i].w != 1.f) ++bad;
            }
            std::printf("load_ptx: %zu points, %d mismatches\n", pc.points.size(), bad);
            CHECK(bad == 0, "load_ptx: %d rows differ", bad);
            PtxCloud half = load_ptx(s, path, 2);
            CHECK(half.ok && half.scan_width == W / 2 && half.scan_height == H / 2, "decimated load_ptx");
            CHECK(!load_ptx(s, "/tmp/ccgpu_no_such_file.ptx").ok && !s.lastError().empty(), "missing file must fail with a message");
            std::remove(path);
        }
    }

    // ---- error behaviour: empty result + message, no exception, session stays usable ----
    {
        std::vector<PointXYZI> bad(pts.begin(), pts.begin() + 100);
        bad[17].y = std::numeric_limits<float>::quiet_NaN();
        auto r = getPCA(s, bad.data(), bad.size(), 0.1, 20);
        CHECK(r->empty() && !s.lastError().empty(), "NaN cloud must be refused with a message");
        std::printf("refused NaN cloud: %s\n", s.lastError().c_str());
        auto ok = getPCA(s, pts.data(), (size_t)64, 0.1, 20);
        CHECK(ok->size() == 64, "session must stay usable after an error");
        std::vector<PointXYZI> none;
        CHECK(getPCA(s, none.data(), (size_t)0, 0.1, 20)->empty(), "empty cloud -> empty result");
        volatile int cancel = 1;
        s.setCancelFlag(&cancel);
        CHECK(estimateNormals(s, pts.data(), n, c2g, W, H).empty(), "a raised deleting_ flag must cancel estimateNormals");
        s.setCancelFlag(nullptr);
    }

    // ---- radiusSearch for every point, variable-length rows, uncapped and with a cap beyond the list kernels ----
    {
        for (unsigned cap : {0u, 100u}) {
            std::vector<int64_t> off; std::vector<int> ri; std::vector<float> rd;
            const double radius = 0.25;
            CHECK(radiusSearchAll(s, pts.data(), n, radius, cap, off, ri, rd), "radiusSearchAll failed (%s)", s.lastError().c_str());
            const float r2 = (float)(radius * radius);
            int bad = 0; size_t widest = 0;
            for (size_t i = 0; i < n && off.size() == n + 1; ++i) {
                size_t m = 0;
                while (m < nbrs[i].size() && nbrs[i][m].first < r2 && (cap == 0 || m < cap)) ++m;
                const size_t got = (size_t)(off[i + 1] - off[i]);
                widest = std::max(widest, got);
                if (got != m) { ++bad; continue; }
                for (size_t j = 0; j < m; ++j)
                    if (ri[off[i] + j] != nbrs[i][j].second || rd[off[i] + j] != nbrs[i][j].first) { ++bad; break; }
            }
            std::printf("radiusSearchAll(max_nn=%u): %zu entries, widest row %zu, %d rows differ\n", cap, ri.size(), widest, bad);
            CHECK(bad == 0, "radiusSearchAll(max_nn=%u) differs at %d points", cap, bad);
        }
    }

    // ---- the threading contract of NormalEstimator (normalestimation.cpp:196, :305-306): estimateNormals runs on a
    //      std::async worker with a Session of its own while the GUI thread keeps using ITS Session, and the GUI thread
    //      raises PointCloud::deleting_ at any moment.  Every outcome must be a full result or an empty one (cancelled),
    //      never a crash, and both Sessions must answer correctly afterwards.
    {
        auto reference = estimateNormals(s, pts.data(), n, c2g, W, H, true);
        CHECK(reference.size() == n, "estimateNormals baseline for the threading check");
        int full = 0, cancelled = 0;
        for (int round = 0; round < 6; ++round) {
            std::atomic<int> deleting(0);
            auto fut = std::async(std::launch::async, [&]() {
                Session worker(0);                                       // the worker thread owns its context
                worker.setCancelFlag(reinterpret_cast<const volatile int *>(&deleting));
                std::vector<Normal> last;
                for (int rep = 0; rep < 4; ++rep) {
                    last = estimateNormals(worker, pts.data(), n, c2g, W, H, true);
                    if (last.empty()) break;
                }
                return last;
            });
            // meanwhile the "GUI thread" works on its own Session ...
            auto pca = getPCA(s, pts.data(), n, 0.1, 20);
            CHECK(pca->size() == n, "GUI-thread getPCA while the worker runs (%s)", s.lastError().c_str());
            // ... and deletes the cloud at some point (rounds 0-1: never, later rounds: earlier and earlier)
            if (round >= 2) {
                std::this_thread::sleep_for(std::chrono::microseconds(4000 >> (round - 2)));
                deleting.store(1);
            }
            std::vector<Normal> got = fut.get();
            if (got.empty()) {
                ++cancelled;
                CHECK(deleting.load() == 1, "estimateNormals returned nothing although deleting_ was never raised");
            } else {
                ++full;
                CHECK(got.size() == n && std::memcmp(got.data(), reference.data(), n * sizeof(Normal)) == 0,
                      "worker-thread normals differ from the GUI-thread result");
            }
        }
        std::printf("threading: %d full results, %d cancelled\n", full, cancelled);
        CHECK(full >= 2, "the rounds without deletion must complete");
        auto again = estimateNormals(s, pts.data(), n, c2g, W, H, true);
        CHECK(again.size() == n && std::memcmp(again.data(), reference.data(), n * sizeof(Normal)) == 0, "session unusable after the threaded rounds");
    }

    std::printf(g_fail ? "adapters: %d check(s) FAILED\n" : "adapters: all checks passed\n", g_fail);
    return g_fail ? 1 : 0;
}
