/* CPU neighbour search of the oracle (TEST INFRASTRUCTURE - see oracle/__init__.py).
 *
 * Plain C restatement of the three neighbour queries the reference's hot path issues, accelerated by a uniform cell
 * grid so that the oracle finishes the full-size scans (32.5 k / 114 k points) in seconds.  The RESULT CONTRACT is the
 * one of oracle/tp_ops.py (which stays as the brute-force statement these functions are tested against,
 * tests/test_oracle_numerics.py):
 *
 *   oracle_frnn   third_party/FRNN/frnn/frnn.py:210-347 + csrc/grid/find_nbrs.cu:418-504 as called at
 *                 src/models/utils/point_ops.py:459: up to K points of p2 with d2 <= r*r (non-strict, find_nbrs.cu:490),
 *                 ascending by (d2, index), idx / d2 padded with -1.
 *   oracle_ball   pytorch3d.ops.ball_query as called at point_ops.py:81: the first K points of p2 IN INDEX ORDER with
 *                 d2 < r*r (strict), idx padded with -1.
 *   oracle_knn    pytorch3d.ops.knn_points as called at point_ops.py:91: the K nearest by (d2, index), ascending,
 *                 no radius; idx / d2 padded with -1 when p2 has fewer than K points (the caller applies upstream's 0-fill).
 *
 * d2 = ((dx*dx)+(dy*dy))+(dz*dz) in fp32, every operation rounded separately (compile with -ffp-contract=off, no
 * -ffast-math): the total order (d2, index) that decides every neighbour index is defined on exactly this value.
 * The grid only prunes candidates; every kept candidate goes through the same fp32 test and the same packed-key order,
 * so the results equal the brute-force statement bit for bit.
 *
 * One cloud per call (the python side loops over the clouds of a batch); `threads` worker threads over the queries.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { int64_t key; int32_t idx; } cell_ent;

static int cmp_ent(const void* a, const void* b) {
    const cell_ent *x = (const cell_ent*)a, *y = (const cell_ent*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx);
}

typedef struct {
    double lo[3], h;          /* grid origin and cell edge */
    int64_t n[3];             /* cells per axis */
    int64_t np;               /* points */
    int64_t* keys;            /* sorted cell keys */
    float* pts;               /* points in key order (x, y, z) */
    int32_t* ids;             /* original index of every sorted point */
} grid_t;

static inline int64_t cell_of(const grid_t* g, double v, int ax) { return (int64_t)floor((v - g->lo[ax]) / g->h); }

static int grid_build(grid_t* g, const float* p, int64_t np, double h) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int64_t i = 0; i < np; ++i)
        for (int a = 0; a < 3; ++a) {
            const double v = p[3 * i + a];
            if (v < lo[a]) lo[a] = v;
            if (v > hi[a]) hi[a] = v;
        }
    g->h = h; g->np = np;
    for (int a = 0; a < 3; ++a) {
        g->lo[a] = lo[a];
        g->n[a] = (int64_t)floor((hi[a] - lo[a]) / h) + 1;
        if (g->n[a] < 1) g->n[a] = 1;
    }
    /* keep the linear key inside int64: n0*n1*n2 < 2^62 is guaranteed by the callers' choice of h (>= extent / 2^20) */
    cell_ent* e = (cell_ent*)malloc(sizeof(cell_ent) * (size_t)(np > 0 ? np : 1));
    g->keys = (int64_t*)malloc(sizeof(int64_t) * (size_t)(np > 0 ? np : 1));
    g->pts = (float*)malloc(sizeof(float) * 3 * (size_t)(np > 0 ? np : 1));
    g->ids = (int32_t*)malloc(sizeof(int32_t) * (size_t)(np > 0 ? np : 1));
    if (!e || !g->keys || !g->pts || !g->ids) return -1;
    for (int64_t i = 0; i < np; ++i) {
        const int64_t cx = cell_of(g, p[3 * i], 0), cy = cell_of(g, p[3 * i + 1], 1), cz = cell_of(g, p[3 * i + 2], 2);
        e[i].key = (cz * g->n[1] + cy) * g->n[0] + cx;
        e[i].idx = (int32_t)i;
    }
    qsort(e, (size_t)np, sizeof(cell_ent), cmp_ent);
    for (int64_t i = 0; i < np; ++i) {
        g->keys[i] = e[i].key;
        g->ids[i] = e[i].idx;
        memcpy(g->pts + 3 * i, p + 3 * (int64_t)e[i].idx, 3 * sizeof(float));
    }
    free(e);
    return 0;
}

static void grid_free(grid_t* g) { free(g->keys); free(g->pts); free(g->ids); }

static inline int64_t lower_bound(const int64_t* a, int64_t n, int64_t v) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t m = (lo + hi) >> 1;
        if (a[m] < v) lo = m + 1; else hi = m;
    }
    return lo;
}

static inline float sqdist3(const float* a, const float* b) {
    const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    float s = dx * dx + dy * dy;     /* -ffp-contract=off: two rounded products, one rounded sum */
    s = s + dz * dz;
    return s;
}

static inline uint64_t pack_key(float d2, int32_t idx) {
    uint32_t bits;
    memcpy(&bits, &d2, 4);
    return ((uint64_t)bits << 32) | (uint32_t)idx;
}

/* insert `key` into the ascending array best[0..*cnt) of capacity K (largest dropped when full) */
static inline void topk_insert(uint64_t* best, int* cnt, int K, uint64_t key) {
    int n = *cnt;
    if (n == K) {
        if (key >= best[K - 1]) return;
        n = K - 1;
    }
    int j = n;
    while (j > 0 && best[j - 1] > key) { best[j] = best[j - 1]; --j; }
    best[j] = key;
    *cnt = n + 1;
}

enum { MODE_FRNN = 0, MODE_BALL = 1, MODE_KNN = 2 };

typedef struct {
    const grid_t* g;
    const float* q;
    int64_t nq, q0, q1;
    int K, mode;
    float r2;
    int64_t* idx;
    float* d2;
} job_t;

/* all points of the x-range [x0, x1] of grid row (cy, cz) */
static inline void scan_row(const job_t* J, const float* qp, int64_t x0, int64_t x1, int64_t cy, int64_t cz, uint64_t* best,
                            int* cnt) {
    const grid_t* g = J->g;
    if (cy < 0 || cy >= g->n[1] || cz < 0 || cz >= g->n[2]) return;
    if (x0 < 0) x0 = 0;
    if (x1 >= g->n[0]) x1 = g->n[0] - 1;
    if (x0 > x1) return;
    const int64_t base = (cz * g->n[1] + cy) * g->n[0];
    int64_t i = lower_bound(g->keys, g->np, base + x0);
    const int64_t kend = base + x1;
    for (; i < g->np && g->keys[i] <= kend; ++i) {
        const float d = sqdist3(qp, g->pts + 3 * i);
        if (J->mode == MODE_FRNN) {
            if (d <= J->r2) topk_insert(best, cnt, J->K, pack_key(d, g->ids[i]));
        } else if (J->mode == MODE_BALL) {
            if (d < J->r2) topk_insert(best, cnt, J->K, (uint64_t)(uint32_t)g->ids[i]);
        } else {
            topk_insert(best, cnt, J->K, pack_key(d, g->ids[i]));
        }
    }
}

static void* worker(void* arg) {
    const job_t* J = (const job_t*)arg;
    const grid_t* g = J->g;
    const int K = J->K;
    uint64_t* best = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)K);
    int64_t maxn = g->n[0];
    if (g->n[1] > maxn) maxn = g->n[1];
    if (g->n[2] > maxn) maxn = g->n[2];
    for (int64_t i = J->q0; i < J->q1; ++i) {
        const float* qp = J->q + 3 * i;
        int cnt = 0;
        const int64_t cx = cell_of(g, qp[0], 0), cy = cell_of(g, qp[1], 1), cz = cell_of(g, qp[2], 2);
        if (J->mode != MODE_KNN) {
            for (int64_t dz = -1; dz <= 1; ++dz)
                for (int64_t dy = -1; dy <= 1; ++dy) scan_row(J, qp, cx - 1, cx + 1, cy + dy, cz + dz, best, &cnt);
        } else {
            /* expanding Chebyshev shells: after shell R every unexamined point is farther than R*h from the query */
            /* how far the query's cell is from the grid: shells below that are empty */
            int64_t far = 0;
            const int64_t c[3] = {cx, cy, cz};
            for (int a = 0; a < 3; ++a) {
                if (-c[a] > far) far = -c[a];
                if (c[a] - (g->n[a] - 1) > far) far = c[a] - (g->n[a] - 1);
            }
            for (int64_t R = 0; R <= far + maxn; ++R) {
                if (R >= far) {
                    for (int64_t dz = -R; dz <= R; ++dz)
                        for (int64_t dy = -R; dy <= R; ++dy) {
                            const int64_t m = (dz < 0 ? -dz : dz) > (dy < 0 ? -dy : dy) ? (dz < 0 ? -dz : dz) : (dy < 0 ? -dy : dy);
                            if (m == R) {
                                scan_row(J, qp, cx - R, cx + R, cy + dy, cz + dz, best, &cnt);
                            } else {
                                scan_row(J, qp, cx - R, cx - R, cy + dy, cz + dz, best, &cnt);
                                if (R > 0) scan_row(J, qp, cx + R, cx + R, cy + dy, cz + dz, best, &cnt);
                            }
                        }
                }
                if (cnt == K) {
                    float kth;
                    const uint32_t bits = (uint32_t)(best[K - 1] >> 32);
                    memcpy(&kth, &bits, 4);
                    const double bound = (double)R * g->h;
                    if ((double)kth < bound * bound * (1.0 - 1e-6)) break;
                }
            }
        }
        for (int k = 0; k < K; ++k) {
            if (k < cnt) {
                J->idx[i * K + k] = (int64_t)(uint32_t)(best[k] & 0xFFFFFFFFu);
                if (J->d2) {
                    if (J->mode == MODE_BALL) {
                        J->d2[i * K + k] = 0.f;
                    } else {
                        const uint32_t bits = (uint32_t)(best[k] >> 32);
                        memcpy(&J->d2[i * K + k], &bits, 4);
                    }
                }
            } else {
                J->idx[i * K + k] = -1;
                if (J->d2) J->d2[i * K + k] = -1.f;
            }
        }
    }
    free(best);
    return NULL;
}

static int run(int mode, const float* p1, int64_t n1, const float* p2, int64_t n2, int K, float r, double h, int64_t* idx,
               float* d2, int threads) {
    if (n1 < 0 || n2 < 0 || K <= 0 || !idx) return -1;
    if (n1 == 0) return 0;
    if (n2 == 0) {
        for (int64_t i = 0; i < n1 * K; ++i) { idx[i] = -1; if (d2) d2[i] = -1.f; }
        return 0;
    }
    grid_t g;
    if (grid_build(&g, p2, n2, h)) return -2;
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    if (threads > n1) threads = (int)n1;
    job_t jobs[64];
    pthread_t th[64];
    const float r2 = r * r;       /* fp32 product, like the python statement's r * r on float32 tensors */
    for (int t = 0; t < threads; ++t) {
        jobs[t].g = &g; jobs[t].q = p1; jobs[t].nq = n1; jobs[t].K = K; jobs[t].mode = mode; jobs[t].r2 = r2;
        jobs[t].idx = idx; jobs[t].d2 = d2;
        jobs[t].q0 = n1 * t / threads; jobs[t].q1 = n1 * (t + 1) / threads;
    }
    for (int t = 1; t < threads; ++t) pthread_create(&th[t], NULL, worker, &jobs[t]);
    worker(&jobs[0]);
    for (int t = 1; t < threads; ++t) pthread_join(th[t], NULL);
    grid_free(&g);
    return 0;
}

static double extent_max(const float* p, int64_t n) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, m = 0;
    for (int64_t i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) {
            const double v = p[3 * i + a];
            if (v < lo[a]) lo[a] = v;
            if (v > hi[a]) hi[a] = v;
        }
    for (int a = 0; a < 3; ++a)
        if (n > 0 && hi[a] - lo[a] > m) m = hi[a] - lo[a];
    return m;
}

/* cell edge for a radius query: a little more than r (so that a pair within the fp32-rounded radius test is never
 * more than one cell apart), but never so small that the key overflows */
static double radius_cell(const float* p2, int64_t n2, float r) {
    double h = (double)r * 1.001;
    const double e = extent_max(p2, n2);
    if (h < e / 1048576.0) h = e / 1048576.0;
    if (!(h > 0)) h = 1.0;
    return h;
}

int oracle_frnn(const float* p1, int64_t n1, const float* p2, int64_t n2, int K, float r, int64_t* idx, float* d2,
                int threads) {
    return run(MODE_FRNN, p1, n1, p2, n2, K, r, radius_cell(p2, n2, r), idx, d2, threads);
}

int oracle_ball(const float* p1, int64_t n1, const float* p2, int64_t n2, int K, float r, int64_t* idx, int threads) {
    return run(MODE_BALL, p1, n1, p2, n2, K, r, radius_cell(p2, n2, r), idx, NULL, threads);
}

int oracle_knn(const float* p1, int64_t n1, const float* p2, int64_t n2, int K, int64_t* idx, float* d2, int threads) {
    /* ~2 points per cell if the cloud filled a plane of its largest extent: lidar sweeps and object scans are surfaces */
    double e = extent_max(p2, n2);
    double h = e / fmax(8.0, sqrt((double)(n2 > 0 ? n2 : 1) / 2.0));
    if (!(h > 0)) h = 1.0;
    return run(MODE_KNN, p1, n1, p2, n2, K, 0.f, h, idx, d2, threads);
}
