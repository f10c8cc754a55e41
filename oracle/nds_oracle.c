/*
 * CPU restatement (C11 + OpenMP) of EvoX's environmental-selection hot path.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.
 * The shipped CUDA path never calls into it.
 *
 * It is the fast twin of oracle/nds_oracle.py (which is the line-by-line NumPy
 * restatement pinned by the reference's golden vectors); tests/test_oracle_*.py
 * check the two against each other.  Citations are relative to /root/reference/.
 *
 * Algorithm = the reference's own formulation:
 *   dominance matrix  A[i][j] = all(x_i <= x_j) & any(x_i < x_j)      src/evox/utils/common.py:95-104
 *   dominator count   count[j] = sum_i A[i][j]                        src/evox/operators/selection/non_dominate.py:61
 *   peel loop         rank[front]=r; count -= sum_{i in front} A[i,:]; count -= front; front = count==0
 *                                                                     non_dominate.py:11-27 and :72-111
 *   crowding          per objective: lexsort((cost, ~mask)), neighbour gap / range, +-inf ends, -inf masked
 *                                                                     non_dominate.py:116-156
 *   NSGA-II select    worst_rank = rank[argsort(rank)[pop]]; mask; crowding; lexsort((-crowd, rank))[:pop]
 *                                                                     src/evox/algorithms/mo/nsga2.py:90-96
 * The only liberty is storage: A is kept as one bit per pair instead of one byte,
 * so that N = 262144 fits in host memory (8 GiB instead of 64 GiB).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

/* ---- dominance of one row against a block of columns (SoA), one byte per pair ---- */
#define CLONES __attribute__((target_clones("avx512f", "avx2", "default")))

CLONES static void dom_row_bytes(const float *xi, const float *col, int64_t n, int m, int64_t j0, int64_t j1,
                                 uint8_t *le, uint8_t *lt) {
    /* common.py:98 -- IEEE compares, so NaN gives false on both tests */
    for (int64_t j = j0; j < j1; ++j) {
        le[j - j0] = 1;
        lt[j - j0] = 0;
    }
    for (int k = 0; k < m; ++k) {
        const float a = xi[k];
        const float *c = col + (int64_t)k * n;
        for (int64_t j = j0; j < j1; ++j) {
            le[j - j0] &= (uint8_t)(a <= c[j]);
            lt[j - j0] |= (uint8_t)(a < c[j]);
        }
    }
}

static inline int64_t words_per_row(int64_t n) { return (n + 63) / 64; }

/* Builds the bit-packed dominance matrix (row i, bit j) and the dominator counts. */
static int build_matrix(const float *x, int64_t n, int m, uint64_t *bits, int64_t *count, int nthreads) {
    const int64_t wpr = words_per_row(n);
    float *col = (float *)malloc(sizeof(float) * (size_t)n * (size_t)(m > 0 ? m : 1));
    if (!col) return -1;
    for (int64_t i = 0; i < n; ++i)
        for (int k = 0; k < m; ++k) col[(int64_t)k * n + i] = x[i * m + k];
    memset(count, 0, sizeof(int64_t) * (size_t)n);
    int err = 0;
#pragma omp parallel num_threads(nthreads)
    {
        uint8_t *le = (uint8_t *)malloc((size_t)n + 64);
        uint8_t *lt = (uint8_t *)malloc((size_t)n + 64);
        if (!le || !lt) {
#pragma omp atomic write
            err = -1;
        } else {
#pragma omp for schedule(dynamic, 16)
            for (int64_t i = 0; i < n; ++i) {
                dom_row_bytes(x + i * m, col, n, m, 0, n, le, lt);
                uint64_t *row = bits + i * wpr;
                for (int64_t j = 0; j < n; ++j) le[j] &= lt[j]; /* all(<=) & any(<) */
                for (int64_t j = n; j < wpr * 64; ++j) le[j] = 0;
                for (int64_t w = 0; w < wpr; ++w) {
                    uint64_t acc = 0;
                    for (int q = 0; q < 8; ++q) { /* eight 0/1 bytes -> eight bits */
                        uint64_t v;
                        memcpy(&v, le + w * 64 + q * 8, 8);
                        acc |= ((v * 0x0102040810204080ULL) >> 56) << (8 * q);
                    }
                    row[w] = acc;
                }
            }
        }
        free(le);
        free(lt);
    }
    free(col);
    if (err) return err;
    /* non_dominate.py:61 -- column sums; each thread owns a block of word columns */
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (int64_t w = 0; w < wpr; ++w) {
        for (int64_t i = 0; i < n; ++i) {
            uint64_t v = bits[i * wpr + w];
            while (v) {
                int b = __builtin_ctzll(v);
                count[w * 64 + b] += 1;
                v &= v - 1;
            }
        }
    }
    return 0;
}

/* non_dominate.py:72-111 (same loop as :11-27).  Returns the number of fronts, or <0. */
ORACLE_API int oracle_nds_rank_f32(const float *x, int64_t n, int m, int32_t *rank, int nthreads) {
    if (n < 0 || m < 0) return -2;
    if (n == 0) return 0;
    if (nthreads <= 0) {
#ifdef _OPENMP
        nthreads = omp_get_max_threads();
#else
        nthreads = 1;
#endif
    }
    const int64_t wpr = words_per_row(n);
    uint64_t *bits = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)n * (size_t)wpr);
    int64_t *count = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    int64_t *desc = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    int64_t *members = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    uint8_t *front = (uint8_t *)malloc((size_t)n);
    if (!bits || !count || !desc || !members || !front) {
        free(bits); free(count); free(desc); free(members); free(front);
        return -1;
    }
    int rc = build_matrix(x, n, m, bits, count, nthreads);
    if (rc) {
        free(bits); free(count); free(desc); free(members); free(front);
        return rc;
    }
    memset(rank, 0, sizeof(int32_t) * (size_t)n); /* :72 rank = zeros */
    int32_t current = 0;
    for (;;) {
        int64_t nf = 0;
        for (int64_t j = 0; j < n; ++j) { /* front = count == 0 */
            front[j] = (count[j] == 0);
            if (front[j]) members[nf++] = j;
        }
        if (nf == 0) break;                                        /* :76-78 */
        for (int64_t t = 0; t < nf; ++t) rank[members[t]] = current; /* :82 */
        /* :91-100 count_desc = sum over front members of their rows */
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int64_t w = 0; w < wpr; ++w) {
            int64_t *d = desc + w * 64;
            const int64_t lim = (n - w * 64) < 64 ? (n - w * 64) : 64;
            for (int64_t b = 0; b < lim; ++b) d[b] = 0;
            for (int64_t t = 0; t < nf; ++t) {
                uint64_t v = bits[members[t] * wpr + w];
                while (v) {
                    int b = __builtin_ctzll(v);
                    d[b] += 1;
                    v &= v - 1;
                }
            }
        }
        for (int64_t j = 0; j < n; ++j) count[j] = count[j] - desc[j] - (int64_t)front[j]; /* :102-103 */
        current += 1;                                                                    /* :105 */
    }
    free(bits); free(count); free(desc); free(members); free(front);
    return (int)current;
}

/* Dense boolean matrix for small cases: out[i*ny + j] = x_i dominates y_j (common.py:101-104). */
ORACLE_API int oracle_dominate_relation_f32(const float *x, int64_t nx, const float *y, int64_t ny, int m, uint8_t *out) {
    for (int64_t i = 0; i < nx; ++i)
        for (int64_t j = 0; j < ny; ++j) {
            int le = 1, lt = 0;
            for (int k = 0; k < m; ++k) {
                le &= x[i * m + k] <= y[j * m + k];
                lt |= x[i * m + k] < y[j * m + k];
            }
            out[i * ny + j] = (uint8_t)(le & lt);
        }
    return 0;
}

/* ---- sorting with lax.sort / np.lexsort float order: NaN last, -0 == +0, stable ---- */
typedef struct { uint32_t primary; float key; int64_t idx; } sort_item;

static int cmp_item(const void *pa, const void *pb) {
    const sort_item *a = (const sort_item *)pa, *b = (const sort_item *)pb;
    if (a->primary != b->primary) return a->primary < b->primary ? -1 : 1;
    const int an = isnan(a->key), bn = isnan(b->key);
    if (an != bn) return an ? 1 : -1;
    if (!an) {
        if (a->key < b->key) return -1;
        if (a->key > b->key) return 1;
    }
    return a->idx < b->idx ? -1 : (a->idx > b->idx ? 1 : 0); /* stability */
}

/* non_dominate.py:116-156.  mask may be NULL (all valid). */
ORACLE_API int oracle_crowding_f32(const float *costs, const uint8_t *mask, int64_t n, int m, float *dist) {
    if (n == 0) return 0;
    int64_t nv = 0;
    if (mask) { for (int64_t i = 0; i < n; ++i) nv += mask[i] != 0; } else nv = n;
    sort_item *it = (sort_item *)malloc(sizeof(sort_item) * (size_t)n);
    float *d = (float *)malloc(sizeof(float) * (size_t)n);
    if (!it || !d) { free(it); free(d); return -1; }
    for (int64_t i = 0; i < n; ++i) dist[i] = 0.0f;
    const int64_t last = (nv - 1 + n) % n; /* index num_valid-1 wraps when num_valid == 0 */
    for (int k = 0; k < m; ++k) {
        for (int64_t i = 0; i < n; ++i) {
            it[i].primary = mask ? (mask[i] ? 0u : 1u) : 0u; /* ~mask is the primary key (:145) */
            it[i].key = costs[i * m + k];
            it[i].idx = i;
        }
        qsort(it, (size_t)n, sizeof(sort_item), cmp_item);
        const float range = it[last].key - it[0].key; /* :147 */
        for (int64_t i = 0; i < n; ++i) d[i] = 0.0f;    /* :148 jnp.empty == zeros */
        for (int64_t p = 1; p + 1 < n; ++p) d[it[p].idx] = (it[p + 1].key - it[p - 1].key) / range; /* :149 */
        d[it[0].idx] = INFINITY;    /* :150 */
        d[it[last].idx] = INFINITY; /* :151 */
        for (int64_t i = 0; i < n; ++i) {
            const float v = (!mask || mask[i]) ? d[i] : -INFINITY; /* :153 */
            dist[i] = dist[i] + v;                                 /* :156 */
        }
    }
    free(it); free(d);
    return 0;
}

/* nsga2.py:90-96.  order_out has pop entries; rank and crowd (n each) may be NULL. */
ORACLE_API int oracle_nsga2_select_f32(const float *fit, int64_t n, int m, int64_t pop, int32_t *order_out,
                                       int32_t *rank_out, float *crowd_out, int nthreads) {
    if (pop < 0 || pop >= n) return -2; /* rank[order[pop]] must exist */
    int32_t *rank = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    float *crowd = (float *)malloc(sizeof(float) * (size_t)n);
    uint8_t *mask = (uint8_t *)malloc((size_t)n);
    sort_item *it = (sort_item *)malloc(sizeof(sort_item) * (size_t)n);
    if (!rank || !crowd || !mask || !it) { free(rank); free(crowd); free(mask); free(it); return -1; }
    int rc = oracle_nds_rank_f32(fit, n, m, rank, nthreads);
    if (rc < 0) { free(rank); free(crowd); free(mask); free(it); return rc; }
    /* worst_rank = rank[argsort(rank)[pop]] : the value of the pop-th smallest rank (:91-92) */
    int64_t *hist = (int64_t *)calloc((size_t)rc + 1, sizeof(int64_t));
    for (int64_t i = 0; i < n; ++i) hist[rank[i]]++;
    int32_t worst = 0;
    int64_t acc = 0;
    for (int32_t r = 0; r <= rc; ++r) { acc += hist[r]; if (acc > pop) { worst = r; break; } }
    free(hist);
    for (int64_t i = 0; i < n; ++i) mask[i] = rank[i] == worst; /* :93 */
    oracle_crowding_f32(fit, mask, n, m, crowd);                /* :94 */
    for (int64_t i = 0; i < n; ++i) {                           /* :96 lexsort((-crowd, rank)) */
        it[i].primary = (uint32_t)rank[i];
        it[i].key = -crowd[i];
        it[i].idx = i;
    }
    qsort(it, (size_t)n, sizeof(sort_item), cmp_item);
    for (int64_t p = 0; p < pop; ++p) order_out[p] = (int32_t)it[p].idx;
    if (rank_out) memcpy(rank_out, rank, sizeof(int32_t) * (size_t)n);
    if (crowd_out) memcpy(crowd_out, crowd, sizeof(float) * (size_t)n);
    free(rank); free(crowd); free(mask); free(it);
    return 0;
}

ORACLE_API int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
