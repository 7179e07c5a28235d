/*
 * ccoracle.c — CPU oracle for the CloudClean kNN / radius / PCA / outlier hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under cloudclean_b200/ or adapters/ may link, import or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs do.
 *
 * PARITY UNPINNED: every arithmetic step of this path lives in un-vendored third-party code
 * (PCL 1.7/1.8 kdtree+octree+common/pca, FLANN 1.8.4, Eigen <= 3.2.10; README.md:19-27,
 * CMakeLists.txt:110) which is absent from /root/reference and from this image, and the
 * reference's own tests pin no numerical result (test/main.cpp:32-50 only prints).  This file
 * restates the published algorithms and anchors control flow on the reference's call sites:
 *   src/utilities/cv.h:76-116          getPCA   (radius search -> <3 marker -> PCA eigenvalues desc)
 *   src/utilities/cv.cpp:249-308       getHist  (same + L2 normalisation of the eigenvalue vector)
 *   src/plugins/normalestimation/normalestimation.cpp:246-470   grid fan normals + kNN(50) fallback
 *   src/plugins/outlierfilter/outlierfilter.cpp:120-179         radius / min-neighbour filter
 *   src/plugins/markov/markov.cpp:408-485, featureeval.cpp:1104-1137,1201-1220   feature formulas
 *
 * Contract:
 *   d2  = fl(fl(fl(dx*dx)+fl(dy*dy))+fl(dz*dz)), IEEE fp32, NO fma (compile with -ffp-contract=off).
 *   kNN = k smallest under the total order (d2, original index); the query itself is included.
 *   radius (kd-tree flavour, strict)  : d2 <  (float)(r*r)   [r double]; cap = max_nn nearest.
 *   radius (octree flavour, inclusive): (double)d2 <= r*r.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static int cco_nt(int nthreads) {
#ifdef _OPENMP
    return nthreads > 0 ? nthreads : omp_get_max_threads();
#else
    (void)nthreads; return 1;
#endif
}

#define CCO_LEAF 15 /* FLANN KDTreeSingleIndexParams default leaf_max_size used by pcl::KdTreeFLANN */

/* ------------------------------------------------------------------------------------------ */
/* kd-tree (single tree, exact) — stands in for pcl::KdTreeFLANN (cv.h:83-84).                 */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    int left, right;   /* children, -1 for leaf */
    int dim;           /* split dimension */
    float lo_max;      /* max coordinate (dim) of the left subtree  */
    float hi_min;      /* min coordinate (dim) of the right subtree */
    int begin, end;    /* leaf: range in perm */
} cco_node;

typedef struct {
    int n;
    float *xyz;        /* n*3 packed copy (what KdTreeFLANN::setInputCloud does) */
    int *perm;
    cco_node *nodes;
    int n_nodes, cap_nodes;
    float bb_lo[3], bb_hi[3];
} cco_tree;

/* nodes are pre-allocated (median splits down to <= CCO_LEAF points leave >= 8 points per leaf: fewer than n/4 + 2
 * nodes), so that sub-trees can be built by concurrent OpenMP tasks; only the numbering of the nodes depends on the
 * schedule, the tree and every search result do not */
static int cco_new_node(cco_tree *t) {
    int id;
    #pragma omp atomic capture
    id = t->n_nodes++;
    return id;
}

/* quickselect on perm[b..e) by coordinate dim, ties by index so the build is deterministic */
static inline int cco_less(const cco_tree *t, int a, int b, int dim) {
    float va = t->xyz[3 * (size_t)a + dim], vb = t->xyz[3 * (size_t)b + dim];
    return va < vb || (va == vb && a < b);
}
static void cco_select(cco_tree *t, int b, int e, int m, int dim) {
    int *p = t->perm;
    while (e - b > 1) {
        int piv = p[b + (e - b) / 2];
        int i = b, j = e - 1;
        while (i <= j) {
            while (cco_less(t, p[i], piv, dim)) i++;
            while (cco_less(t, piv, p[j], dim)) j--;
            if (i <= j) { int tmp = p[i]; p[i] = p[j]; p[j] = tmp; i++; j--; }
        }
        if (m <= j) e = j + 1;
        else if (m >= i) b = i;
        else return;
    }
}

static int cco_build_rec(cco_tree *t, int b, int e) {
    int id = cco_new_node(t);
    if (e - b <= CCO_LEAF) {
        cco_node *nd = &t->nodes[id];
        nd->left = nd->right = -1; nd->dim = 0; nd->lo_max = nd->hi_min = 0.f;
        nd->begin = b; nd->end = e;
        return id;
    }
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = b; i < e; i++) {
        const float *p = &t->xyz[3 * (size_t)t->perm[i]];
        for (int d = 0; d < 3; d++) { if (p[d] < lo[d]) lo[d] = p[d]; if (p[d] > hi[d]) hi[d] = p[d]; }
    }
    int dim = 0; float ext = hi[0] - lo[0];
    for (int d = 1; d < 3; d++) if (hi[d] - lo[d] > ext) { ext = hi[d] - lo[d]; dim = d; }
    int m = b + (e - b) / 2;
    cco_select(t, b, e, m, dim);
    float lo_max = -FLT_MAX, hi_min = FLT_MAX;
    for (int i = b; i < m; i++) { float v = t->xyz[3 * (size_t)t->perm[i] + dim]; if (v > lo_max) lo_max = v; }
    for (int i = m; i < e; i++) { float v = t->xyz[3 * (size_t)t->perm[i] + dim]; if (v < hi_min) hi_min = v; }
    int l, r;
    if (e - b > 32768) {
        #pragma omp task shared(l) firstprivate(t, b, m)
        l = cco_build_rec(t, b, m);
        r = cco_build_rec(t, m, e);
        #pragma omp taskwait
    } else {
        l = cco_build_rec(t, b, m);
        r = cco_build_rec(t, m, e);
    }
    cco_node *nd = &t->nodes[id];
    nd->left = l; nd->right = r; nd->dim = dim; nd->lo_max = lo_max; nd->hi_min = hi_min;
    nd->begin = b; nd->end = e;
    return id;
}

/* xyz: first float of an AoS with `stride_bytes` between points (pcl::PointXYZI = 32, XYZINormal = 48) */
void *cco_tree_build(const void *xyz, int n, int stride_bytes) {
    cco_tree *t = (cco_tree *)calloc(1, sizeof(cco_tree));
    t->n = n;
    t->xyz = (float *)malloc(sizeof(float) * 3 * (size_t)(n > 0 ? n : 1));
    t->perm = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int d = 0; d < 3; d++) { t->bb_lo[d] = FLT_MAX; t->bb_hi[d] = -FLT_MAX; }
    for (int i = 0; i < n; i++) {
        const float *p = (const float *)((const char *)xyz + (size_t)i * (size_t)stride_bytes);
        for (int d = 0; d < 3; d++) {
            t->xyz[3 * (size_t)i + d] = p[d];
            if (p[d] < t->bb_lo[d]) t->bb_lo[d] = p[d];
            if (p[d] > t->bb_hi[d]) t->bb_hi[d] = p[d];
        }
        t->perm[i] = i;
    }
    t->cap_nodes = n / 4 + 64;
    t->nodes = (cco_node *)malloc(sizeof(cco_node) * (size_t)t->cap_nodes);
    if (n > 0) {
        #pragma omp parallel
        #pragma omp single
        cco_build_rec(t, 0, n);
    }
    return t;
}
void cco_tree_free(void *tv) {
    cco_tree *t = (cco_tree *)tv;
    if (!t) return;
    free(t->xyz); free(t->perm); free(t->nodes); free(t);
}

/* the contract's distance: ((dx*dx + dy*dy) + dz*dz), fp32, no fma (FLANN L2_Simple, [EXT]) */
static inline float cco_d2(const float *a, const float *b) {
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    float xx = dx * dx, yy = dy * dy, zz = dz * dz;
    float s = xx + yy;
    return s + zz;
}

/* bounded result set kept sorted ascending by (d2, idx) */
typedef struct {
    int k, cnt;
    float *d2; int *idx;
    float bound;        /* radius bound (d2 must pass radius test) */
    int use_radius;     /* 0: none, 1: strict d2 < bound (float), 2: inclusive (double)d2 <= bound_d */
    double bound_d;
    long total_in_radius; /* counts everything inside the radius, even beyond the cap */
} cco_rs;

static inline int cco_in_radius(const cco_rs *rs, float d2) {
    if (rs->use_radius == 1) return d2 < rs->bound;
    if (rs->use_radius == 2) return (double)d2 <= rs->bound_d;
    return 1;
}
static inline void cco_rs_insert(cco_rs *rs, float d2, int idx) {
    int k = rs->k, c = rs->cnt;
    if (c == k) {
        float wd = rs->d2[k - 1]; int wi = rs->idx[k - 1];
        if (!(d2 < wd || (d2 == wd && idx < wi))) return;
        c = k - 1;
    }
    int pos = c;
    while (pos > 0 && (d2 < rs->d2[pos - 1] || (d2 == rs->d2[pos - 1] && idx < rs->idx[pos - 1]))) {
        rs->d2[pos] = rs->d2[pos - 1]; rs->idx[pos] = rs->idx[pos - 1]; pos--;
    }
    rs->d2[pos] = d2; rs->idx[pos] = idx;
    rs->cnt = c + 1;
}
/* a cell whose box distance already fails the radius test holds no acceptable point */
static inline int cco_outside_radius(const cco_rs *rs, float boxd2) {
    if (rs->use_radius == 1) return !(boxd2 < rs->bound);
    if (rs->use_radius == 2) return !((double)boxd2 <= rs->bound_d);
    return 0;
}
/* prune iff no point of the cell can enter the result set; ties (boxd2 == worst) are kept */
static inline int cco_prune(const cco_rs *rs, float boxd2) {
    if (cco_outside_radius(rs, boxd2)) return 1;
    if (rs->cnt == rs->k && rs->k > 0 && boxd2 > rs->d2[rs->k - 1]) return 1;
    return 0;
}

static void cco_search(const cco_tree *t, int node, const float *q, float off[3], cco_rs *rs, int count_only) {
    const cco_node *nd = &t->nodes[node];
    if (nd->left < 0) {
        for (int i = nd->begin; i < nd->end; i++) {
            int id = t->perm[i];
            float d2 = cco_d2(q, &t->xyz[3 * (size_t)id]);
            if (!cco_in_radius(rs, d2)) continue;
            rs->total_in_radius++;
            if (!count_only) cco_rs_insert(rs, d2, id);
        }
        return;
    }
    int dim = nd->dim;
    float v = q[dim];
    /* choose the nearer child by the gap midpoint */
    int go_left = (v - nd->lo_max) < (nd->hi_min - v) || v <= nd->lo_max;
    int near = go_left ? nd->left : nd->right;
    int far = go_left ? nd->right : nd->left;
    cco_search(t, near, q, off, rs, count_only);
    float fo = go_left ? (nd->hi_min - v) : (v - nd->lo_max);   /* lower bound on |q[dim]-p[dim]| in far */
    if (fo < 0.f) fo = 0.f;
    float save = off[dim];
    if (fo > save) off[dim] = fo;
    float xx = off[0] * off[0], yy = off[1] * off[1], zz = off[2] * off[2];
    float s = xx + yy; float boxd2 = s + zz;
    if (count_only ? !cco_outside_radius(rs, boxd2) : !cco_prune(rs, boxd2))
        cco_search(t, far, q, off, rs, count_only);
    off[dim] = save;
}

static void cco_root_off(const cco_tree *t, const float *q, float off[3]) {
    for (int d = 0; d < 3; d++) {
        float o = 0.f;
        if (q[d] < t->bb_lo[d]) o = t->bb_lo[d] - q[d];
        else if (q[d] > t->bb_hi[d]) o = q[d] - t->bb_hi[d];
        off[d] = o;
    }
}

/* a1: exact kNN of cloud points [q_begin,q_end) among the cloud, self included, ascending (d2, idx).
 * Rows are padded with idx=-1, d2=+inf when k > n.  out arrays are (q_end-q_begin)*k.            */
void cco_knn(const void *tv, int k, int q_begin, int q_end, int32_t *idx_out, float *d2_out, int nthreads) {
    const cco_tree *t = (const cco_tree *)tv;
    int nt = cco_nt(nthreads);
    #pragma omp parallel num_threads(nt)
    {
        float *d2 = (float *)malloc(sizeof(float) * (size_t)(k + 1));
        int *ix = (int *)malloc(sizeof(int) * (size_t)(k + 1));
        #pragma omp for schedule(dynamic, 1024)
        for (int qi = q_begin; qi < q_end; qi++) {
            cco_rs rs; rs.k = k; rs.cnt = 0; rs.d2 = d2; rs.idx = ix; rs.use_radius = 0; rs.total_in_radius = 0;
            rs.bound = 0; rs.bound_d = 0;
            const float *q = &t->xyz[3 * (size_t)qi];
            float off[3]; cco_root_off(t, q, off);
            if (t->n > 0 && k > 0) cco_search(t, 0, q, off, &rs, 0);
            size_t o = (size_t)(qi - q_begin) * (size_t)k;
            for (int j = 0; j < k; j++) {
                idx_out[o + j] = j < rs.cnt ? ix[j] : -1;
                d2_out[o + j] = j < rs.cnt ? d2[j] : INFINITY;
            }
        }
        free(d2); free(ix);
    }
}

/* brute-force twin of cco_knn (O(n) per query) used to validate the kd-tree itself */
void cco_knn_brute(const void *xyz, int n, int stride_bytes, int k, int q_begin, int q_end,
                   int32_t *idx_out, float *d2_out) {
    #pragma omp parallel
    {
        float *d2 = (float *)malloc(sizeof(float) * (size_t)(k + 1));
        int *ix = (int *)malloc(sizeof(int) * (size_t)(k + 1));
        #pragma omp for schedule(dynamic, 64)
        for (int qi = q_begin; qi < q_end; qi++) {
            cco_rs rs; rs.k = k; rs.cnt = 0; rs.d2 = d2; rs.idx = ix; rs.use_radius = 0; rs.total_in_radius = 0;
            const float *q = (const float *)((const char *)xyz + (size_t)qi * (size_t)stride_bytes);
            for (int i = 0; i < n; i++) {
                const float *p = (const float *)((const char *)xyz + (size_t)i * (size_t)stride_bytes);
                if (k > 0) cco_rs_insert(&rs, cco_d2(q, p), i);
            }
            size_t o = (size_t)(qi - q_begin) * (size_t)k;
            for (int j = 0; j < k; j++) {
                idx_out[o + j] = j < rs.cnt ? ix[j] : -1;
                d2_out[o + j] = j < rs.cnt ? d2[j] : INFINITY;
            }
        }
        free(d2); free(ix);
    }
}

/* a2 / a3: radius search of cloud points [q_begin,q_end).
 *   strict_lt=1: kd-tree flavour  d2 <  (float)(r*r)        (cv.h:98)
 *   strict_lt=0: octree flavour   (double)d2 <= r*r          (outlierfilter.cpp:169)
 * cap>0: keep the `cap` nearest under (d2, idx) — FLANN KNNRadiusResultSet [EXT]; rows padded with -1/inf.
 * count_out = number kept (min(total, cap)); total_out (optional) = all points inside the radius.
 * idx_out/d2_out may be NULL (count only).                                                       */
void cco_radius(const void *tv, double r, int cap, int strict_lt, int q_begin, int q_end,
                int32_t *idx_out, float *d2_out, int32_t *count_out, int32_t *total_out, int nthreads) {
    const cco_tree *t = (const cco_tree *)tv;
    int nt = cco_nt(nthreads);
    int count_only = (idx_out == NULL && d2_out == NULL);
    #pragma omp parallel num_threads(nt)
    {
        int kk = cap > 0 ? cap : 1;
        float *d2 = (float *)malloc(sizeof(float) * (size_t)(kk + 1));
        int *ix = (int *)malloc(sizeof(int) * (size_t)(kk + 1));
        #pragma omp for schedule(dynamic, 1024)
        for (int qi = q_begin; qi < q_end; qi++) {
            cco_rs rs; rs.k = kk; rs.cnt = 0; rs.d2 = d2; rs.idx = ix; rs.total_in_radius = 0;
            rs.use_radius = strict_lt ? 1 : 2;
            rs.bound = (float)(r * r); rs.bound_d = r * r;
            const float *q = &t->xyz[3 * (size_t)qi];
            float off[3]; cco_root_off(t, q, off);
            if (t->n > 0) cco_search(t, 0, q, off, &rs, count_only);
            size_t row = (size_t)(qi - q_begin);
            long kept = count_only ? (cap > 0 && rs.total_in_radius > cap ? cap : rs.total_in_radius) : rs.cnt;
            if (count_out) count_out[row] = (int32_t)kept;
            if (total_out) total_out[row] = (int32_t)rs.total_in_radius;
            if (!count_only) {
                for (int j = 0; j < kk; j++) {
                    if (idx_out) idx_out[row * (size_t)kk + j] = j < rs.cnt ? ix[j] : -1;
                    if (d2_out) d2_out[row * (size_t)kk + j] = j < rs.cnt ? d2[j] : INFINITY;
                }
            }
        }
        free(d2); free(ix);
    }
}

/* ------------------------------------------------------------------------------------------ */
/* 3x3 symmetric eigen                                                                          */
/* ------------------------------------------------------------------------------------------ */

/* fp64 cyclic Jacobi: the adjudicating "truth".  a = {xx,xy,xz,yy,yz,zz}.  w descending, v columns
 * (v[3*c+r] = component r of eigenvector c).                                                    */
static void cco_eig3_f64(const double a_in[6], double w[3], double v[9]) {
    double A[3][3] = {{a_in[0], a_in[1], a_in[2]}, {a_in[1], a_in[3], a_in[4]}, {a_in[2], a_in[4], a_in[5]}};
    double V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 64; sweep++) {
        double offn = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
        double dn = fabs(A[0][0]) + fabs(A[1][1]) + fabs(A[2][2]);
        if (offn <= 1e-300 || offn <= 1e-22 * dn) break;
        for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
            if (A[p][q] == 0.0) continue;
            double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
            double tt = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            double c = 1.0 / sqrt(tt * tt + 1.0), s = tt * c;
            for (int r = 0; r < 3; r++) {   /* A <- A J */
                double arp = A[r][p], arq = A[r][q];
                A[r][p] = c * arp - s * arq; A[r][q] = s * arp + c * arq;
            }
            for (int r = 0; r < 3; r++) {   /* A <- J^T A */
                double apr = A[p][r], aqr = A[q][r];
                A[p][r] = c * apr - s * aqr; A[q][r] = s * apr + c * aqr;
            }
            for (int r = 0; r < 3; r++) {
                double vrp = V[r][p], vrq = V[r][q];
                V[r][p] = c * vrp - s * vrq; V[r][q] = s * vrp + c * vrq;
            }
        }
    }
    int o[3] = {0, 1, 2};
    double d[3] = {A[0][0], A[1][1], A[2][2]};
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2 - i; j++)
        if (d[o[j]] < d[o[j + 1]]) { int tmp = o[j]; o[j] = o[j + 1]; o[j + 1] = tmp; }
    for (int c = 0; c < 3; c++) {
        w[c] = d[o[c]];
        for (int r = 0; r < 3; r++) v[3 * c + r] = V[r][o[c]];
    }
}

/* fp32 "reference-like" solver: Householder tridiagonalisation + implicit-shift QL on the
 * tridiagonal (Golub & Van Loan alg. 8.3.3 / EISPACK tql1), the published algorithm behind
 * Eigen::SelfAdjointEigenSolver<Matrix3f>::compute that pcl::PCA calls [EXT].  Eigenvalues only,
 * descending.  Rounding will not match Eigen bit for bit; it is reported beside the fp64 truth.   */
static void cco_eig3_f32_ql(const float a_in[6], float w[3]) {
    float a00 = a_in[0], a01 = a_in[1], a02 = a_in[2], a11 = a_in[3], a12 = a_in[4], a22 = a_in[5];
    float d[3], e[3];
    /* Householder: zero a02 */
    if (a02 == 0.f) {
        d[0] = a00; d[1] = a11; d[2] = a22; e[0] = a01; e[1] = a12;
    } else {
        float beta = sqrtf(a01 * a01 + a02 * a02);
        float invb = 1.f / beta;
        float m01 = a01 * invb, m02 = a02 * invb;
        float q = 2.f * m01 * a12 + m02 * (a22 - a11);
        d[0] = a00; d[1] = a11 + m02 * q; d[2] = a22 - m02 * q;
        e[0] = beta; e[1] = a12 - m01 * q;
    }
    e[2] = 0.f;
    /* implicit QL with Wilkinson shift */
    for (int l = 0; l < 3; l++) {
        int iter = 0, m;
        do {
            for (m = l; m < 2; m++) {
                float dd = fabsf(d[m]) + fabsf(d[m + 1]);
                if (fabsf(e[m]) <= FLT_EPSILON * dd) break;
            }
            if (m != l) {
                if (iter++ == 60) break;
                float g = (d[l + 1] - d[l]) / (2.f * e[l]);
                float r = hypotf(g, 1.f);
                g = d[m] - d[l] + e[l] / (g + (g >= 0.f ? fabsf(r) : -fabsf(r)));
                float s = 1.f, c = 1.f, p = 0.f;
                int i;
                for (i = m - 1; i >= l; i--) {
                    float f = s * e[i], b = c * e[i];
                    r = hypotf(f, g);
                    e[i + 1] = r;
                    if (r == 0.f) { d[i + 1] -= p; e[m] = 0.f; break; }
                    s = f / r; c = g / r;
                    g = d[i + 1] - p;
                    r = (d[i] - g) * s + 2.f * c * b;
                    p = s * r;
                    d[i + 1] = g + p;
                    g = c * r - b;
                }
                if (r == 0.f && i >= l) continue;
                d[l] -= p; e[l] = g; e[m] = 0.f;
            }
        } while (m != l);
    }
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2 - i; j++)
        if (d[j] < d[j + 1]) { float tmp = d[j]; d[j] = d[j + 1]; d[j + 1] = tmp; }
    w[0] = d[0]; w[1] = d[1]; w[2] = d[2];
}

void cco_eig3_f64_export(const double a[6], double w[3], double v[9]) { cco_eig3_f64(a, w, v); }
void cco_eig3_f32_export(const float a[6], float w[3]) { cco_eig3_f32_ql(a, w); }

/* ------------------------------------------------------------------------------------------ */
/* a4/a5/a6/a12: PCA of given neighbour lists                                                  */
/* ------------------------------------------------------------------------------------------ */
/* nbr: rows*cap int32, count: rows (entries beyond count ignored; count NULL => every row has cap).
 * cov_norm: 0 scatter (PCL 1.7.0), 1 scatter/(n-1) (PCL >= 1.8 pcl::PCA), 2 scatter/n (pcl::NormalEstimation).
 * min_pts: rows with fewer neighbours get the (0,0,1) marker (cv.h:105-109 uses 3).
 * Outputs (any may be NULL):
 *   eig32  rows*3  fp32 reference-like (fp32 centroid in neighbour order, fp32 scatter, fp32 QL) desc
 *   eig64  rows*3  fp64 truth, desc
 *   nrm64  rows*4  fp64 truth: unit eigenvector of the smallest eigenvalue + curvature l2/(l0+l1+l2)
 *   gap64  rows    (l1-l2)/l0 from the truth (conditioning of the normal)                           */
void cco_pca_lists(const void *xyz, int stride_bytes, const int32_t *nbr, const int32_t *count, int rows, int cap,
                   int cov_norm, int min_pts, float *eig32, double *eig64, double *nrm64, double *gap64) {
    #pragma omp parallel for schedule(dynamic, 1024)
    for (int row = 0; row < rows; row++) {
        int n = count ? count[row] : cap;
        if (n > cap) n = cap;
        const int32_t *nb = nbr + (size_t)row * (size_t)cap;
        while (n > 0 && nb[n - 1] < 0) n--;
        if (n < min_pts) {
            if (eig32) { eig32[3 * (size_t)row] = 0.f; eig32[3 * (size_t)row + 1] = 0.f; eig32[3 * (size_t)row + 2] = 1.f; }
            if (eig64) { eig64[3 * (size_t)row] = 0; eig64[3 * (size_t)row + 1] = 0; eig64[3 * (size_t)row + 2] = 1; }
            if (nrm64) { for (int c = 0; c < 4; c++) nrm64[4 * (size_t)row + c] = NAN; }
            if (gap64) gap64[row] = 0;
            continue;
        }
        /* fp32 path */
        float mx = 0.f, my = 0.f, mz = 0.f;
        for (int j = 0; j < n; j++) {
            const float *p = (const float *)((const char *)xyz + (size_t)nb[j] * (size_t)stride_bytes);
            mx += p[0]; my += p[1]; mz += p[2];
        }
        mx /= (float)n; my /= (float)n; mz /= (float)n;
        float s[6] = {0, 0, 0, 0, 0, 0};
        double dm[3] = {0, 0, 0};
        for (int j = 0; j < n; j++) {
            const float *p = (const float *)((const char *)xyz + (size_t)nb[j] * (size_t)stride_bytes);
            float x = p[0] - mx, y = p[1] - my, z = p[2] - mz;
            s[0] += x * x; s[1] += x * y; s[2] += x * z; s[3] += y * y; s[4] += y * z; s[5] += z * z;
            dm[0] += p[0]; dm[1] += p[1]; dm[2] += p[2];
        }
        float sc32 = cov_norm == 1 ? 1.f / (float)(n - 1) : cov_norm == 2 ? 1.f / (float)n : 1.f;
        if (eig32) {
            float a[6]; for (int c = 0; c < 6; c++) a[c] = s[c] * sc32;
            cco_eig3_f32_ql(a, &eig32[3 * (size_t)row]);
        }
        /* fp64 truth */
        dm[0] /= n; dm[1] /= n; dm[2] /= n;
        double S[6] = {0, 0, 0, 0, 0, 0};
        for (int j = 0; j < n; j++) {
            const float *p = (const float *)((const char *)xyz + (size_t)nb[j] * (size_t)stride_bytes);
            double x = p[0] - dm[0], y = p[1] - dm[1], z = p[2] - dm[2];
            S[0] += x * x; S[1] += x * y; S[2] += x * z; S[3] += y * y; S[4] += y * z; S[5] += z * z;
        }
        double sc = cov_norm == 1 ? 1.0 / (double)(n - 1) : cov_norm == 2 ? 1.0 / (double)n : 1.0;
        for (int c = 0; c < 6; c++) S[c] *= sc;
        double w[3], v[9];
        cco_eig3_f64(S, w, v);
        if (eig64) { for (int c = 0; c < 3; c++) eig64[3 * (size_t)row + c] = w[c]; }
        if (nrm64) {
            double tr = w[0] + w[1] + w[2];
            nrm64[4 * (size_t)row + 0] = v[6]; nrm64[4 * (size_t)row + 1] = v[7]; nrm64[4 * (size_t)row + 2] = v[8];
            nrm64[4 * (size_t)row + 3] = tr > 0 ? fabs(w[2] / tr) : 0.0;
        }
        if (gap64) gap64[row] = w[0] > 0 ? (w[1] - w[2]) / w[0] : 0.0;
    }
}

/* a5: getHist's extra step (cv.cpp:293): L2-normalise each eigenvalue triple in place (fp32). */
void cco_hist_normalise(float *eig, int rows) {
    for (int i = 0; i < rows; i++) {
        float *e = &eig[3 * (size_t)i];
        float nn = sqrtf((e[0] * e[0] + e[1] * e[1]) + e[2] * e[2]);
        if (nn > 0.f) { e[0] /= nn; e[1] /= nn; e[2] /= nn; }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* a9: eigen-derived features, 7 per point:                                                    */
/*   0 anisotropy 1 planarity 2 sphericity 3 linearity 4 omnivariance 5 eigenentropy 6 curvature */
/* formula_set 0 = as written in markov.cpp:449-475 (incl. its slips), 1 = textbook.             */
/* ------------------------------------------------------------------------------------------ */
void cco_features(const float *eig3, int n, int formula_set, float *out7) {
    for (int i = 0; i < n; i++) {
        float l0 = eig3[3 * (size_t)i], l1 = eig3[3 * (size_t)i + 1], l2 = eig3[3 * (size_t)i + 2];
        float *o = &out7[7 * (size_t)i];
        o[0] = (l0 - l2) / l0;                       /* markov.cpp:451 */
        o[1] = (l1 - l2) / l0;                       /* :455 */
        o[2] = l2 / l0;                              /* :459 */
        o[3] = (l0 - l1) / l0;                       /* :463 */
        if (formula_set == 0) {
            o[4] = cbrtf(l0 * l1 * l0);              /* :467 (sic: l0 twice) */
            o[5] = -(l0 * logf(l0) + l1 * logf(l1) + l2 * logf(l0));  /* :471 (sic) */
            o[6] = l0 / (l0 + l1 + l2);              /* :475 (sic: largest) */
        } else {
            o[4] = cbrtf(l0 * l1 * l2);
            o[5] = -(l0 * logf(l0) + l1 * logf(l1) + l2 * logf(l2));
            o[6] = l2 / (l0 + l1 + l2);
        }
    }
}

/* featureeval.cpp:1104-1137: eigen ratio  (0 for the <3 marker, l1/l0 if l0>l2 else 1) */
void cco_eigen_ratio(const float *eig3, int n, float *out) {
    for (int i = 0; i < n; i++) {
        float l0 = eig3[3 * (size_t)i], l1 = eig3[3 * (size_t)i + 1], l2 = eig3[3 * (size_t)i + 2];
        if (l1 < l2) { out[i] = 0.f; continue; }
        out[i] = l0 > l2 ? l1 / l0 : 1.f;
    }
}
/* featureeval.cpp:1201-1220 + cv.h:498-501: clamp(cos(eig,(1,0,0)),0,1), 0 for the marker */
void cco_plane_similarity(const float *eig3, int n, float *out) {
    for (int i = 0; i < n; i++) {
        float l0 = eig3[3 * (size_t)i], l1 = eig3[3 * (size_t)i + 1], l2 = eig3[3 * (size_t)i + 2];
        if (l1 < l2) { out[i] = 0.f; continue; }
        float nn = sqrtf((l0 * l0 + l1 * l1) + l2 * l2);
        float c = l0 / (nn * 1.0f);
        out[i] = c < 0.f ? 0.f : (c > 1.f ? 1.f : c);
    }
}

/* ------------------------------------------------------------------------------------------ */
/* a11: statistical outlier filter — PCL StatisticalOutlierRemoval definition [EXT]             */
/* ------------------------------------------------------------------------------------------ */
/* mean_d[i] = (float)( sum_{j=1..k} (double)sqrtf(d2_ij) / k ), neighbours ascending, self (j=0) dropped.
 * mu, sigma over all i in fp64 in index order; outlier <=> (double)mean_d > mu + alpha*sigma.       */
void cco_sor(const void *tv, int mean_k, double stddev_mul, uint8_t *outlier, float *mean_d_out,
             double *mu_out, double *sigma_out, int nthreads) {
    const cco_tree *t = (const cco_tree *)tv;
    int n = t->n, k = mean_k + 1;
    int32_t *idx = (int32_t *)malloc(sizeof(int32_t) * (size_t)n * (size_t)k);
    float *d2 = (float *)malloc(sizeof(float) * (size_t)n * (size_t)k);
    float *md = mean_d_out ? mean_d_out : (float *)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    cco_knn(tv, k, 0, n, idx, d2, nthreads);
    for (int i = 0; i < n; i++) {
        double s = 0.0;
        for (int j = 1; j < k; j++) {
            float dd = d2[(size_t)i * k + j];
            if (isinf(dd)) continue;          /* fewer than k+1 points in the cloud */
            s += (double)sqrtf(dd);
        }
        md[i] = (float)(s / (double)mean_k);
    }
    double sum = 0.0, sq = 0.0;
    for (int i = 0; i < n; i++) { sum += (double)md[i]; sq += (double)md[i] * (double)md[i]; }
    double mu = n > 0 ? sum / (double)n : 0.0;
    double var = n > 1 ? (sq - sum * sum / (double)n) / ((double)n - 1.0) : 0.0;
    double sigma = sqrt(var > 0 ? var : 0.0);
    double thr = mu + stddev_mul * sigma;
    if (outlier) for (int i = 0; i < n; i++) outlier[i] = (double)md[i] > thr ? 1 : 0;
    if (mu_out) *mu_out = mu;
    if (sigma_out) *sigma_out = sigma;
    free(idx); free(d2);
    if (!mean_d_out) free(md);
}

/* ------------------------------------------------------------------------------------------ */
/* a10: radius / min-neighbour outlier filter (outlierfilter.cpp:120-179)                       */
/* ------------------------------------------------------------------------------------------ */
/* prefilter=0: flag every point whose inclusive r-ball (self included) holds < min_neighbours points.
 * prefilter=1: additionally reproduce the octreeDownsample(2r) candidate step (:126-147): only points
 *   whose voxel (edge 2r) holds < min_neighbours points are examined.  The voxel lattice follows PCL's
 *   OctreePointCloud growth rule [EXT, recalled]: the first point sits at the centre of its voxel and the
 *   box only ever grows by whole multiples of the edge, i.e. voxel(p) = floor((p - (p0 - edge/2)) / edge).  */
static inline int64_t cco_voxel_key(const float *p, const float *p0, float edge) {
    /* PCL computes the key in float: (point - min) / resolution, truncated */
    int64_t k[3];
    for (int d = 0; d < 3; d++) {
        float o = p0[d] - edge * 0.5f;
        k[d] = (int64_t)floorf((p[d] - o) / edge);
    }
    return ((k[0] & 0x1FFFFF) << 42) | ((k[1] & 0x1FFFFF) << 21) | (k[2] & 0x1FFFFF);
}
static int cco_cmp_i64(const void *a, const void *b) {
    int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
    return x < y ? -1 : x > y;
}
void cco_radius_outliers(const void *tv, double r, int min_neighbours, int prefilter, uint8_t *outlier, int nthreads) {
    const cco_tree *t = (const cco_tree *)tv;
    int n = t->n;
    uint8_t *cand = (uint8_t *)malloc((size_t)(n > 0 ? n : 1));
    memset(cand, 1, (size_t)n);
    if (prefilter && n > 0) {
        float edge = (float)(2.0 * r);
        int64_t *keys = (int64_t *)malloc(sizeof(int64_t) * 2 * (size_t)n);
        for (int i = 0; i < n; i++) keys[i] = keys[n + i] = cco_voxel_key(&t->xyz[3 * (size_t)i], &t->xyz[0], edge);
        qsort(keys + n, (size_t)n, sizeof(int64_t), cco_cmp_i64);
        for (int i = 0; i < n; i++) {
            /* count of keys[i] in the sorted copy */
            int64_t key = keys[i];
            int lo = 0, hi = n;
            while (lo < hi) { int m = (lo + hi) / 2; if (keys[n + m] < key) lo = m + 1; else hi = m; }
            int first = lo; hi = n;
            while (lo < hi) { int m = (lo + hi) / 2; if (keys[n + m] <= key) lo = m + 1; else hi = m; }
            cand[i] = (lo - first) < min_neighbours;
        }
        free(keys);
    }
    int32_t *total = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    cco_radius(tv, r, 0, 0, 0, n, NULL, NULL, NULL, total, nthreads);
    for (int i = 0; i < n; i++) outlier[i] = (cand[i] && total[i] < min_neighbours) ? 1 : 0;
    free(total); free(cand);
}

/* ------------------------------------------------------------------------------------------ */
/* f1: octreeDownsample (src/utilities/utils.h:16-62)                                           */
/* ------------------------------------------------------------------------------------------ */
/* Per voxel of edge `resolution` (lattice of cco_voxel_key; voxels numbered in ascending key order — PCL's own
 * leaf order is an octree DFS over an insertion-dependent box [EXT] and is not reproduced): the first nf floats
 * of every point's struct are summed in INDEX order in fp32 into an accumulator that starts as a
 * default-constructed PCL point (float[3] = 1 when as_written, utils.h:41-48), then scaled by
 * (float)(1.0 / count) (utils.h:52-54).  big_to_small[i] = voxel of point i (utils.h:50).  Returns the voxel count;
 * out must hold n*nf floats. */
typedef struct { int64_t key; int idx; } cco_kv;
static int cco_cmp_kv(const void *a, const void *b) {
    const cco_kv *x = (const cco_kv *)a, *y = (const cco_kv *)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->idx < y->idx ? -1 : x->idx > y->idx;
}
int cco_voxel_downsample(const void *aos, int n, int stride_bytes, int nf, float resolution, int as_written,
                         float *out, int32_t *big_to_small) {
    if (n <= 0) return 0;
#define CCO_PT(i) ((const float *)((const char *)aos + (size_t)(i) * (size_t)stride_bytes))
    cco_kv *kv = (cco_kv *)malloc(sizeof(cco_kv) * (size_t)n);
    for (int i = 0; i < n; i++) { kv[i].key = cco_voxel_key(CCO_PT(i), CCO_PT(0), resolution); kv[i].idx = i; }
    qsort(kv, (size_t)n, sizeof(cco_kv), cco_cmp_kv);
    int nv = 0;
    for (int b = 0; b < n;) {
        int e = b;
        while (e < n && kv[e].key == kv[b].key) e++;
        float acc[16];
        for (int f = 0; f < nf; f++) acc[f] = (as_written && f == 3) ? 1.0f : 0.0f;
        for (int j = b; j < e; j++) {
            const float *p = CCO_PT(kv[j].idx);
            for (int f = 0; f < nf; f++) acc[f] = acc[f] + p[f];
            if (big_to_small) big_to_small[kv[j].idx] = nv;
        }
        float size_inv = (float)(1.0 / (double)(e - b));
        for (int f = 0; f < nf; f++) out[(size_t)nv * nf + f] = acc[f] * size_inv;
        nv++;
        b = e;
    }
    free(kv);
#undef CCO_PT
    return nv;
}

/* ------------------------------------------------------------------------------------------ */
/* a7: scan-grid triangle-fan normals (normalestimation.cpp:268-470)                            */
/* ------------------------------------------------------------------------------------------ */
static const int CCO_RING[8][2] = {{1, 0}, {1, -1}, {0, -1}, {-1, -1}, {-1, 0}, {-1, 1}, {0, 1}, {1, 1}}; /* :293-302 */

static inline int cco_valid(int dx, int dy, int width, int height, int index, const int *g2c) { /* :246-266 */
    int cx = index % width, cy = index / width;
    int x = cx + dx, y = cy + dy;
    if (x > width - 1 || x < 0) return 0;
    if (y > height - 1 || y < 0) return 0;
    return g2c[y * width + x] != -1;
}

/* normal4: n*4 floats (nx,ny,nz,data_n[3]); data_n[3]=1 for fan normals, NaN x4 where no face, and the
 * kNN(k_fallback) neighbour-normal mean (not re-normalised, data_n[3] left NaN) after stage (iii).
 * do_fallback=0 stops after stage (ii).  Returns the number of normals still missing.              */
int cco_grid_normals(const void *xyz, int n, int stride_bytes, const int32_t *cloud_to_grid, int scan_w, int scan_h,
                     int do_fallback, int k_fallback, float *normal4) {
    int height = scan_w, width = scan_h;            /* sic, :288-289 */
    int length = width * height;
    int *g2c = (int *)malloc(sizeof(int) * (size_t)(length > 0 ? length : 1));
    for (int i = 0; i < length; i++) g2c[i] = -1;
    for (int i = 0; i < n; i++) g2c[cloud_to_grid[i]] = i;                    /* :279-283 */
    int *missing = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    int n_missing = 0;
    for (int i = 0; i < n; i++) { normal4[4 * (size_t)i] = normal4[4 * (size_t)i + 1] = normal4[4 * (size_t)i + 2] = normal4[4 * (size_t)i + 3] = 0.f; }
#define P(ci) ((const float *)((const char *)xyz + (size_t)(ci) * (size_t)stride_bytes))
    for (int i = 0; i < length; i++) {
        if (g2c[i] == -1) continue;
        float ax = 0.f, ay = 0.f, az = 0.f;
        int face_count = 0, next = 0, first = -1;
        while (1) {
            int p1 = -1, p2 = -1;
            while (next < 8) {
                if (cco_valid(CCO_RING[next][0], CCO_RING[next][1], width, height, i, g2c)) { p1 = next++; break; }
                next++;
            }
            while (next < 9) {
                if (next == 8 && face_count == 1) break;
                if (next == 8 && face_count > 1) { p2 = first; break; }
                int id = next % 8;
                if (cco_valid(CCO_RING[id][0], CCO_RING[id][1], width, height, i, g2c)) { p2 = id; break; }
                next++;
            }
            if (p1 == -1 || p2 == -1) break;
            if (first == -1) first = p1;
            int cx = i % width, cy = i / width;
            int idx1 = (cy + CCO_RING[p1][1]) * width + (cx + CCO_RING[p1][0]);
            int idx2 = (cy + CCO_RING[p2][1]) * width + (cx + CCO_RING[p2][0]);
            const float *p0 = P(g2c[i]), *q1 = P(g2c[idx1]), *q2 = P(g2c[idx2]);
            float v1x = q1[0] - p0[0], v1y = q1[1] - p0[1], v1z = q1[2] - p0[2];
            float v2x = q2[0] - p0[0], v2y = q2[1] - p0[1], v2z = q2[2] - p0[2];
            float cxp = v1y * v2z - v1z * v2y, cyp = v1z * v2x - v1x * v2z, czp = v1x * v2y - v1y * v2x;
            float nn = sqrtf(cxp * cxp + (cyp * cyp + czp * czp));   /* Eigen 3.2 fixed-size redux halves the range: a0 + (a1 + a2) [EXT] */
            float tx = cxp / nn, ty = cyp / nn, tz = czp / nn;          /* 0/0 = NaN, :376-382 */
            if (tx != tx || ty != ty || tz != tz) continue;
            ax += tx; ay += ty; az += tz;
            face_count++;
        }
        float *o = &normal4[4 * (size_t)g2c[i]];
        if (face_count == 0) {
            missing[n_missing++] = g2c[i];
            o[0] = o[1] = o[2] = o[3] = NAN;
            continue;
        }
        float nn = sqrtf(ax * ax + (ay * ay + az * az));
        o[0] = ax / nn; o[1] = ay / nn; o[2] = az / nn; o[3] = 1.0f;
    }
#undef P
    int still = n_missing;
    if (do_fallback && n_missing > 0) {                                        /* :428-459 */
        void *tree = cco_tree_build(xyz, n, stride_bytes);
        int k = k_fallback;
        int32_t *nb = (int32_t *)malloc(sizeof(int32_t) * (size_t)k);
        float *d2 = (float *)malloc(sizeof(float) * (size_t)k);
        for (int m = 0; m < n_missing; m++) {
            int idx = missing[m];
            cco_knn(tree, k, idx, idx + 1, nb, d2, 1);
            float sx = 0.f, sy = 0.f, sz = 0.f; int count = 0;
            for (int j = 0; j < k; j++) {
                if (nb[j] < 0) continue;
                const float *nbn = &normal4[4 * (size_t)nb[j]];
                if (nbn[0] != nbn[0]) continue;
                ++count; sx += nbn[0]; sy += nbn[1]; sz += nbn[2];
            }
            if (count == 0) continue;
            normal4[4 * (size_t)idx] = sx / (float)count;
            normal4[4 * (size_t)idx + 1] = sy / (float)count;
            normal4[4 * (size_t)idx + 2] = sz / (float)count;
            --still;
        }
        free(nb); free(d2); cco_tree_free(tree);
    }
    free(g2c); free(missing);
    return still;
}

/* ------------------------------------------------------------------------------------------ */
/* CPU baseline of the bench metric: the reference's per-point loop shape (cv.h:96-113: search,  */
/* then pcl::PCA of the neighbourhood, then the feature formulas of markov.cpp:449-475), fused:  */
/* kNN(k) -> fp32 centroid/covariance -> eigenvalues (fp32 QL) + normal (eigenvector of the      */
/* smallest eigenvalue via the fp64 Jacobi, as SelfAdjointEigenSolver returns vectors too) ->    */
/* 7 features.  No neighbour table is materialised.  OpenMP over queries (nthreads <= 0: all;    */
/* 1 = what the reference does: its loop is serial).  Outputs may be NULL.                       */
/* ------------------------------------------------------------------------------------------ */
void cco_knn_pca_features(const void *tv, int k, int cov_norm, int formula_set, int q_begin, int q_end, float *eig3, float *normal4,
                          float *feat7, int nthreads) {
    const cco_tree *t = (const cco_tree *)tv;
    int nt = cco_nt(nthreads);
    #pragma omp parallel num_threads(nt)
    {
        float *d2 = (float *)malloc(sizeof(float) * (size_t)(k + 1));
        int *ix = (int *)malloc(sizeof(int) * (size_t)(k + 1));
        #pragma omp for schedule(dynamic, 1024)
        for (int qi = q_begin; qi < q_end; qi++) {
            cco_rs rs; rs.k = k; rs.cnt = 0; rs.d2 = d2; rs.idx = ix; rs.use_radius = 0; rs.total_in_radius = 0;
            rs.bound = 0; rs.bound_d = 0;
            const float *q = &t->xyz[3 * (size_t)qi];
            float off[3]; cco_root_off(t, q, off);
            if (t->n > 0 && k > 0) cco_search(t, 0, q, off, &rs, 0);
            const int n = rs.cnt;
            float w[3] = {0.f, 0.f, 1.f}, nrm[4] = {NAN, NAN, NAN, NAN};
            if (n >= 3) {
                float mx = 0.f, my = 0.f, mz = 0.f;
                for (int j = 0; j < n; j++) { const float *p = &t->xyz[3 * (size_t)ix[j]]; mx += p[0]; my += p[1]; mz += p[2]; }
                mx /= (float)n; my /= (float)n; mz /= (float)n;
                float s[6] = {0, 0, 0, 0, 0, 0};
                for (int j = 0; j < n; j++) {
                    const float *p = &t->xyz[3 * (size_t)ix[j]];
                    float x = p[0] - mx, y = p[1] - my, z = p[2] - mz;
                    s[0] += x * x; s[1] += x * y; s[2] += x * z; s[3] += y * y; s[4] += y * z; s[5] += z * z;
                }
                float sc = cov_norm == 1 ? 1.f / (float)(n - 1) : cov_norm == 2 ? 1.f / (float)n : 1.f;
                for (int c = 0; c < 6; c++) s[c] *= sc;
                cco_eig3_f32_ql(s, w);
                if (normal4) {
                    double S[6], wd[3], v[9];
                    for (int c = 0; c < 6; c++) S[c] = s[c];
                    cco_eig3_f64(S, wd, v);
                    double tr = wd[0] + wd[1] + wd[2];
                    nrm[0] = (float)v[6]; nrm[1] = (float)v[7]; nrm[2] = (float)v[8]; nrm[3] = tr > 0 ? (float)fabs(wd[2] / tr) : 0.f;
                }
            }
            if (eig3) { eig3[3 * (size_t)qi] = w[0]; eig3[3 * (size_t)qi + 1] = w[1]; eig3[3 * (size_t)qi + 2] = w[2]; }
            if (normal4) for (int c = 0; c < 4; c++) normal4[4 * (size_t)qi + c] = nrm[c];
            if (feat7) cco_features(w, 1, formula_set, &feat7[7 * (size_t)qi]);
        }
        free(d2); free(ix);
    }
}

int cco_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---------------------------------------------------------------------------------------------------
 * f4 — PointCloud::load_ptx (src/model/pointcloud.cpp:127-330) restated with the same stdio calls in the same order:
 * fscanf("%d %d") (:149), three + nine "%f" (:161-170), five fgets (:175-180), strtok(" ") token count of the first
 * body line (:199-216), then per sample: the decimation test with an fgets skip (:219-230), fscanf("%f %f %f %f\n")
 * (:237), fgets of the rest when rgb (:253-256), the (0,0,0) drop (:263-265) and the row {x,y,z,1, intensity, 0,
 * float(cloud index), float(grid index)} + cloud_to_grid_map (:284-287).  A failed fscanf without a stream error
 * `continue`s (:240-246) exactly as there.  Test infrastructure only.
 * out_points: cap rows of 8 floats; returns the number of rows (or -1 when the file cannot be opened, -2 on the
 * reference's hard failure path); hdr = {file_width, file_height, has_rgb, sampled_idx}, pose = origin[3] + matrix[9]. */
#include <stdio.h>
#include <locale.h>
long long cco_load_ptx(const char *filename, int decimation_factor, float *out_points, int32_t *c2g, long long cap,
                       long long *hdr, float *pose) {
    setlocale(LC_NUMERIC, "C");
    FILE *pfile = fopen(filename, "r");
    if (!pfile) return -1;
    int file_width = 0, file_height = 0;
    if (fscanf(pfile, "%d %d", &file_width, &file_height) != 2) { fclose(pfile); return -2; }
    if (fscanf(pfile, "%f %f %f", &pose[0], &pose[1], &pose[2]) != 3) { fclose(pfile); return -2; }
    for (int i = 0; i < 9; i++)
        if (fscanf(pfile, "%f", &pose[3 + i]) != 1) { fclose(pfile); return -2; }
    char buff[1024];
    for (int i = 0; i < 5; i++)
        if (!fgets(buff, 1024, pfile)) break;
    long file_pos = ftell(pfile);
    int tokens = 0;
    if (fgets(buff, 1024, pfile)) {
        for (char *ch = strtok(buff, " "); ch; ch = strtok(NULL, " ")) ++tokens;
    }
    const int has_rgb = tokens == 7;
    fseek(pfile, file_pos, SEEK_SET);
    const long long line_count = (long long)file_width * (long long)file_height;
    long long file_sample_idx = 0, n = 0;
    unsigned int sampled_idx = 0;
    while (file_sample_idx < line_count) {
        const long long row = file_sample_idx / file_width, col = file_sample_idx % file_width;
        file_sample_idx++;
        if ((row + 1) % decimation_factor != 0 || (col + 1) % decimation_factor != 0) {
            if (!fgets(buff, 1024, pfile)) { /* end of file: the reference keeps looping */ }
            continue;
        }
        if (ferror(pfile) || feof(pfile)) { fclose(pfile); hdr[0] = file_width; hdr[1] = file_height; hdr[2] = has_rgb; hdr[3] = sampled_idx; return -2; }
        float x, y, z, intensity;
        const int filled = fscanf(pfile, "%f %f %f %f\n", &x, &y, &z, &intensity);
        if (filled != 4) {
            if (ferror(pfile)) { fclose(pfile); return -2; }
            continue;
        }
        if (has_rgb)
            if (!fgets(buff, 1024, pfile)) { /* last line without newline */ }
        sampled_idx++;
        if (x == 0 && y == 0 && z == 0) continue;
        if (n < cap) {
            float *o = out_points + 8 * n;
            o[0] = x; o[1] = y; o[2] = z; o[3] = 1.0f;
            o[4] = intensity; o[5] = 0.0f; o[6] = (float)(size_t)n; o[7] = (float)(sampled_idx - 1);
            c2g[n] = (int32_t)(sampled_idx - 1);
        }
        n++;
    }
    fclose(pfile);
    hdr[0] = file_width; hdr[1] = file_height; hdr[2] = has_rgb; hdr[3] = sampled_idx;
    return n;
}

/* strtof on a batch of NUL-separated tokens (the reference's "%f" conversions are glibc's strtof) */
void cco_strtof_batch(const char *tokens, int count, float *out) {
    setlocale(LC_NUMERIC, "C");
    const char *p = tokens;
    for (int i = 0; i < count; i++) {
        out[i] = strtof(p, NULL);
        p += strlen(p) + 1;
    }
}

/* writes a PTX file from rows of 8 floats (x y z - intensity ...) — test/bench infrastructure for full-size files:
 * width*height body lines "%.6f %.6f %.6f %.6f", rows beyond n (or flagged by a NaN x) become "0 0 0 0.500000" */
long long cco_write_ptx(const char *filename, const float *rows8, long long n, int width, int height) {
    FILE *f = fopen(filename, "w");
    if (!f) return -1;
    static char big[1 << 22];
    setvbuf(f, big, _IOFBF, sizeof(big));
    fprintf(f, "%d\n%d\n0.000000 0.000000 0.000000\n1.000000 0.000000 0.000000\n0.000000 1.000000 0.000000\n0.000000 0.000000 1.000000\n", width, height);
    fprintf(f, "1.000000 0.000000 0.000000 0.000000\n0.000000 1.000000 0.000000 0.000000\n0.000000 0.000000 1.000000 0.000000\n0.000000 0.000000 0.000000 1.000000\n");
    const long long total = (long long)width * height;
    for (long long i = 0; i < total; i++) {
        const float *r = rows8 + 8 * i;
        if (i < n && r[0] == r[0]) fprintf(f, "%.6f %.6f %.6f %.6f\n", (double)r[0], (double)r[1], (double)r[2], (double)r[4]);
        else fputs("0 0 0 0.500000\n", f);
    }
    fclose(f);
    return total;
}
