// NSGA-III environmental selection after the ranking: normalisation, association with the reference
// directions, niche-preserving selection.
//
// Replaces src/evox/algorithms/mo/nsga3.py:97-211 of the reference (reference points:
// src/evox/operators/sampling/uniform.py:10-60, cos_dist: src/evox/utils/common.py:82-87).  Same arithmetic
// in float32, different schedule:
//   * ideal point, extreme points and the worst point are three atomic reductions over the rows kept
//     (rank <= last_rank); keys are the order-preserving integer image of the floats, the row index rides in
//     the low word so that ties resolve to the first row like nanargmin;
//   * the M x M hyperplane system is solved by one thread (LU with partial pivoting in float32; the rank test
//     of jnp.linalg.matrix_rank is a one-sided Jacobi SVD with the same tolerance);
//   * association is the N x R x M hot loop: one thread per row, reference directions staged in shared memory;
//   * the reference's select loop takes ONE survivor per while_loop iteration (:156-201), each an argmin
//     over all niches plus a sort of all rows.  Its order is reproduced exactly by rounds: all niches that
//     share the current minimum count are served in index order within one round, the t-th of them using
//     key number selected + t of jax.random.split(key, pop + 1) -- so a round is parallel over niches and the
//     number of rounds is the number of distinct count levels visited, not pop_size.
// The sampler (threefry-2x32 split / randint) follows JAX 0.4.x; see oracle/jax_prng.py for the restatement
// it is tested against.
#pragma once

#include <float.h>

#include "common.cuh"
#include "sort.cuh"

namespace evoxb {

__host__ __device__ __forceinline__ float key_to_float(uint32_t k) {
    const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } cv; cv.u = u; return cv.f;
#endif
}

// ---- threefry-2x32-20 and the two JAX samplers built on it ---------------------------------------------
__host__ __device__ inline void threefry2x32(uint32_t k0, uint32_t k1, uint32_t x0, uint32_t x1, uint32_t &o0, uint32_t &o1) {
    const uint32_t ks[3] = {k0, k1, k0 ^ k1 ^ 0x1BD11BDAu};
    const int rot[2][4] = {{13, 15, 26, 6}, {17, 29, 16, 24}};
    x0 += ks[0];
    x1 += ks[1];
    for (int g = 0; g < 5; ++g) {
        for (int q = 0; q < 4; ++q) {
            const int r = rot[g & 1][q];
            x0 += x1;
            x1 = (x1 << r) | (x1 >> (32 - r));
            x1 ^= x0;
        }
        x0 += ks[(g + 1) % 3];
        x1 += ks[(g + 2) % 3] + (uint32_t)(g + 1);
    }
    o0 = x0;
    o1 = x1;
}

// word q of the 2 * num words jax.random.split(key, num) produces (counters 0 .. 2 num - 1 hashed in two halves)
__host__ __device__ inline uint32_t jax_split_word(uint32_t k0, uint32_t k1, uint32_t num, uint32_t q) {
    uint32_t a, b;
    if (q < num) {
        threefry2x32(k0, k1, q, num + q, a, b);
        return a;
    }
    threefry2x32(k0, k1, q - num, q, a, b);
    return b;
}

// jax.random.randint(key, (), 0, span) for int32, span >= 1
__host__ __device__ inline uint32_t jax_randint(uint32_t k0, uint32_t k1, uint32_t span) {
    // k1', k2' = split(key): words 0..3 of the split with num = 2
    const uint32_t a0 = jax_split_word(k0, k1, 2u, 0u), a1 = jax_split_word(k0, k1, 2u, 1u);
    const uint32_t b0 = jax_split_word(k0, k1, 2u, 2u), b1 = jax_split_word(k0, k1, 2u, 3u);
    uint32_t hi, lo, unused;
    threefry2x32(a0, a1, 0u, 0u, hi, unused);  // random_bits of shape (): one counter, padded with a zero
    threefry2x32(b0, b1, 0u, 0u, lo, unused);
    uint32_t mult = 65536u % span;
    mult = (mult * mult) % span;
    const uint32_t off = (hi % span) * mult + (lo % span);
    return off % span;
}

struct Nsga3Layout {
    int *hist;            // [n] rank histogram
    int *kept_list;       // [n] rows with rank <= last_rank, any order
    int *info;            // [64]: [0] last_rank, [1] full-rank flag, [2] rounds of the niche loop, [3] rows kept
    uint32_t *ideal_key;  // [m]
    uint32_t *worst_key;  // [m]
    unsigned long long *ext_key;  // [m] (key of the scalarised value << 32) | row
    float *norm;          // [3 m]: ideal, nadir, extreme rows (as floats)
    float *xu;            // [n][m] unit vectors of the normalised rows
    float *xnorm;         // [n] their norms (NaN: row not kept)
    float *yu;            // [r][m] unit reference directions
    int *gid;             // [n]
    float *gdist;         // [n]
    int *rho;             // [r]
    int *rho_last;        // [r]
    int *off;             // [r + 1] CSR offsets of the last front's rows by niche
    int *cursor;          // [r]
    int *members;         // [n] rows of the last front, niche-major, ascending inside a niche
    uint8_t *taken;       // [n] parallel to members
    uint8_t *keep;        // [n] 1 = survivor
    size_t bytes;
};

inline Nsga3Layout nsga3_layout(void *ws, int64_t n, int m, int64_t r) {
    Nsga3Layout L{};
    Carver c(ws);
    const size_t n1 = (size_t)(n > 0 ? n : 1), r1 = (size_t)(r > 0 ? r : 1);
    L.hist = c.take<int>(n1);
    L.kept_list = c.take<int>(n1);
    L.info = c.take<int>(64);
    L.ideal_key = c.take<uint32_t>((size_t)m);
    L.worst_key = c.take<uint32_t>((size_t)m);
    L.ext_key = c.take<unsigned long long>((size_t)m);
    L.norm = c.take<float>((size_t)(3 * m));
    L.xu = c.take<float>(n1 * (size_t)m);
    L.xnorm = c.take<float>(n1);
    L.yu = c.take<float>(r1 * (size_t)m);
    L.gid = c.take<int>(n1);
    L.gdist = c.take<float>(n1);
    L.rho = c.take<int>(r1);
    L.rho_last = c.take<int>(r1);
    L.off = c.take<int>(r1 + 1);
    L.cursor = c.take<int>(r1);
    L.members = c.take<int>(n1);
    L.taken = c.take<uint8_t>(n1);
    L.keep = c.take<uint8_t>(n1);
    L.bytes = c.total();
    return L;
}

// ---- nsga3.py:99-105: rows kept (rank <= last_rank), ideal point = nanmin over them ---------------------
__global__ void __launch_bounds__(256)
n3_ideal_kernel(const float *__restrict__ f, const int *__restrict__ rank, int n, int m, const int *__restrict__ info,
                uint32_t *__restrict__ ideal_key) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int last = info[0];
    const bool kept = i < n && rank[i] <= last;
    for (int k = 0; k < m; ++k) {
        uint32_t key = 0xffffffffu;  // NaN and rows not kept do not take part
        if (kept) key = float_to_key(f[(int64_t)i * m + k]);
        key = __reduce_min_sync(0xffffffffu, key);
        if ((threadIdx.x & 31) == 0 && key != 0xffffffffu) atomicMin(ideal_key + k, key);
    }
}

// ---- nsga3.py:106-112 and :123-125: extreme rows (argmin of the scalarised maximum) and the worst point ---
__global__ void __launch_bounds__(256)
n3_extreme_kernel(const float *__restrict__ f, const int *__restrict__ rank, int n, int m, const int *__restrict__ info,
                  const uint32_t *__restrict__ ideal_key, unsigned long long *__restrict__ ext_key,
                  uint32_t *__restrict__ worst_key) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int last = info[0];
    const bool kept = i < n && rank[i] <= last;
    float t[EVOXB_MAX_OBJECTIVES];
#pragma unroll 1
    for (int k = 0; k < m; ++k) {
        t[k] = kept ? f[(int64_t)i * m + k] - key_to_float(ideal_key[k]) : __int_as_float(0x7fc00000);
        uint32_t wk = (t[k] == t[k]) ? float_to_key(t[k]) : 0u;  // nanmax: NaN does not take part
        wk = __reduce_max_sync(0xffffffffu, wk);
        if ((threadIdx.x & 31) == 0 && wk) atomicMax(worst_key + k, wk);
    }
    const float w_diag = 1.0f + 1e-6f, w_off = 1e-6f;  // jnp.eye(M) + 1e-6 in float32
#pragma unroll 1
    for (int a = 0; a < m; ++a) {
        float best = __int_as_float(0x7fc00000);  // nanmax over the objectives; all NaN stays NaN
#pragma unroll 1
        for (int k = 0; k < m; ++k) {
            const float v = t[k] / (k == a ? w_diag : w_off);
            if (v == v && !(best >= v)) best = v;
        }
        // nanargmin = argmin with NaN read as +inf: rows that are not kept (or all NaN) still take part, and tie
        // with genuine +inf values by index -- observable on populations holding infinities
        unsigned long long key = ~0ull;
        if (i < n) key = ((unsigned long long)float_to_key(best == best ? best : __int_as_float(0x7f800000)) << 32) | (unsigned)i;
        // warp minimum of a 64-bit key: high word first, then the low word among the lanes that tie
        const uint32_t hi = __reduce_min_sync(0xffffffffu, (uint32_t)(key >> 32));
        const uint32_t lo = __reduce_min_sync(0xffffffffu, ((uint32_t)(key >> 32) == hi) ? (uint32_t)key : 0xffffffffu);
        if ((threadIdx.x & 31) == 0 && !(hi == 0xffffffffu && lo == 0xffffffffu))
            atomicMin(ext_key + a, ((unsigned long long)hi << 32) | lo);
    }
}

// ---- nsga3.py:113-131: intercepts of the hyperplane through the extreme points, or the worst point --------
__global__ void n3_nadir_kernel(const float *__restrict__ f, int m, const uint32_t *__restrict__ ideal_key,
                                const unsigned long long *__restrict__ ext_key, const uint32_t *__restrict__ worst_key,
                                float *__restrict__ norm, int *__restrict__ info) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float e[EVOXB_MAX_OBJECTIVES][EVOXB_MAX_OBJECTIVES];
    double u[EVOXB_MAX_OBJECTIVES][EVOXB_MAX_OBJECTIVES];
    for (int k = 0; k < m; ++k) norm[k] = key_to_float(ideal_key[k]);
    for (int a = 0; a < m; ++a) {
        const unsigned row = (unsigned)(ext_key[a] & 0xffffffffull);
        norm[2 * m + a] = (ext_key[a] == ~0ull) ? -1.0f : (float)row;
        for (int k = 0; k < m; ++k) {
            e[a][k] = (ext_key[a] == ~0ull) ? __int_as_float(0x7fc00000) : f[(int64_t)row * m + k] - norm[k];
            u[a][k] = (double)e[a][k];
        }
    }
    // matrix_rank: singular values by one-sided Jacobi (columns of u are rotated until orthogonal)
    bool finite = true;
    for (int a = 0; a < m; ++a)
        for (int k = 0; k < m; ++k) finite = finite && isfinite(u[a][k]);
    int rk = 0;
    if (finite) {
        for (int sweep = 0; sweep < 40; ++sweep) {
            double offd = 0.0;
            for (int p = 0; p < m - 1; ++p)
                for (int q = p + 1; q < m; ++q) {
                    double al = 0.0, be = 0.0, ga = 0.0;
                    for (int i = 0; i < m; ++i) {
                        al += u[i][p] * u[i][p];
                        be += u[i][q] * u[i][q];
                        ga += u[i][p] * u[i][q];
                    }
                    if (ga == 0.0) continue;
                    offd = fmax(offd, fabs(ga) / sqrt(fmax(al * be, 1e-300)));
                    const double zeta = (be - al) / (2.0 * ga);
                    const double tt = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                    const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
                    for (int i = 0; i < m; ++i) {
                        const double up = u[i][p], uq = u[i][q];
                        u[i][p] = cs * up - sn * uq;
                        u[i][q] = sn * up + cs * uq;
                    }
                }
            if (offd < 1e-15) break;
        }
        double smax = 0.0, sv[EVOXB_MAX_OBJECTIVES];
        for (int q = 0; q < m; ++q) {
            double s2 = 0.0;
            for (int i = 0; i < m; ++i) s2 += u[i][q] * u[i][q];
            sv[q] = sqrt(s2);
            smax = fmax(smax, sv[q]);
        }
        const double tol = (double)(float)smax * (double)m * (double)FLT_EPSILON;  // S.max() * max(M, N) * eps(float32)
        for (int q = 0; q < m; ++q) rk += (float)sv[q] > (float)tol ? 1 : 0;
    }
    info[1] = (rk == m) ? 1 : 0;
    if (rk == m) {
        // jnp.linalg.solve(extreme, ones): LU with partial pivoting, float32
        float b[EVOXB_MAX_OBJECTIVES];
        for (int i = 0; i < m; ++i) b[i] = 1.0f;
        for (int c = 0; c < m; ++c) {
            int piv = c;
            for (int i = c + 1; i < m; ++i)
                if (fabsf(e[i][c]) > fabsf(e[piv][c])) piv = i;
            if (piv != c) {
                for (int k = 0; k < m; ++k) { const float tmp = e[c][k]; e[c][k] = e[piv][k]; e[piv][k] = tmp; }
                const float tb = b[c]; b[c] = b[piv]; b[piv] = tb;
            }
            for (int i = c + 1; i < m; ++i) {
                const float l = e[i][c] / e[c][c];
                for (int k = c + 1; k < m; ++k) e[i][k] -= l * e[c][k];
                b[i] -= l * b[c];
            }
        }
        for (int i = m - 1; i >= 0; --i) {
            float acc = b[i];
            for (int k = i + 1; k < m; ++k) acc -= e[i][k] * b[k];
            b[i] = acc / e[i][i];
        }
        for (int k = 0; k < m; ++k) norm[m + k] = 1.0f / b[k];
    } else {
        for (int k = 0; k < m; ++k) norm[m + k] = worst_key[k] ? key_to_float(worst_key[k]) : __int_as_float(0x7fc00000);
    }
}

// ---- nsga3.py:133-134 (row side): normalised rows as unit vectors + their norms; unit reference directions ----
__global__ void __launch_bounds__(256)
n3_unit_rows_kernel(const float *__restrict__ f, const int *__restrict__ rank, int n, int m, int *__restrict__ info,
                    const float *__restrict__ norm, float *__restrict__ xu, float *__restrict__ xnorm, int *__restrict__ kept_list,
                    int *__restrict__ gid, float *__restrict__ gdist, int r) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool kept = i < n && rank[i] <= info[0];
    // rows that are kept go to a dense list (any order): the association kernel only runs over those
    const uint32_t bal = __ballot_sync(0xffffffffu, kept);
    int base = 0;
    if ((threadIdx.x & 31) == 0 && bal) base = atomicAdd(info + 3, __popc(bal));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (i >= n) return;
    if (!kept) {
        xnorm[i] = __int_as_float(0x7fc00000);
        gid[i] = r;  // no direction (nsga3.py:141), distance NaN
        gdist[i] = __int_as_float(0x7fc00000);
        return;
    }
    kept_list[base + __popc(bal & ((1u << (threadIdx.x & 31)) - 1u))] = i;
    float v[EVOXB_MAX_OBJECTIVES];
    float s = 0.0f;
#pragma unroll 1
    for (int k = 0; k < m; ++k) {
        v[k] = (f[(int64_t)i * m + k] - norm[k]) / norm[m + k];
        s += v[k] * v[k];
    }
    const float nn = sqrtf(s);
    xnorm[i] = nn;
#pragma unroll 1
    for (int k = 0; k < m; ++k) xu[(int64_t)i * m + k] = v[k] / nn;
}

__global__ void __launch_bounds__(256)
n3_unit_refs_kernel(const float *__restrict__ ref, int r, int m, float *__restrict__ yu) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= r) return;
    float s = 0.0f;
    for (int k = 0; k < m; ++k) s += ref[(int64_t)j * m + k] * ref[(int64_t)j * m + k];
    const float nn = sqrtf(s);
    for (int k = 0; k < m; ++k) yu[(int64_t)j * m + k] = ref[(int64_t)j * m + k] / nn;
}

// ---- nsga3.py:134-142: perpendicular distance to every direction, nearest one (NaN skipped, first on ties) ----
// Eight threads share two kept rows: thread s of the eight takes the directions j = s (mod 8) of every tile (a
// direction read from shared memory serves both rows), and the eight partial (distance, index) minima are merged
// by shuffles, the smaller index winning among equal distances.  Directions are staged as float4-padded records
// (a record is ceil(M / 4) 16-byte loads); the dot
// product runs k = 0 .. M-1 in one FMA chain.
template <int M>
struct N3Tile {
    static constexpr int kPad = (M + 3) / 4 * 4;
    static constexpr int value = 8192 / kPad;  // directions per shared-memory tile (32 KB of floats)
};
constexpr int kN3Rows = 2;                            // rows per thread group (4 measured slower: 95 registers)
constexpr int kN3Split = 8;                           // threads per group (slices of the directions)
constexpr int kN3RowsPerCta = 256 / kN3Split * kN3Rows;  // 64

template <int M>
__global__ void __launch_bounds__(256)
n3_associate_kernel(const float *__restrict__ xu, const float *__restrict__ xnorm, const int *__restrict__ kept_list,
                    const int *__restrict__ info, const float *__restrict__ yu, int r, int *__restrict__ gid,
                    float *__restrict__ gdist) {
    constexpr int kPad = N3Tile<M>::kPad, kTile = N3Tile<M>::value;
    __shared__ __align__(16) float ys[kTile * kPad];
    const int nkept = info[3];
    if ((int)(blockIdx.x * kN3RowsPerCta) >= nkept) return;  // uniform over the CTA
    const int slice = threadIdx.x & (kN3Split - 1), group = threadIdx.x / kN3Split;
    int row[kN3Rows];
    float nn[kN3Rows], x[kN3Rows][kPad], best[kN3Rows];
    int arg[kN3Rows];
    bool some[kN3Rows], ok[kN3Rows];
    float sthr[kN3Rows], smax[kN3Rows];
#pragma unroll
    for (int q = 0; q < kN3Rows; ++q) {
        const int slot = blockIdx.x * kN3RowsPerCta + group * kN3Rows + q;
        row[q] = slot < nkept ? kept_list[slot] : -1;
        nn[q] = row[q] >= 0 ? xnorm[row[q]] : __int_as_float(0x7fc00000);
#pragma unroll
        for (int k = 0; k < kPad; ++k) x[q][k] = (row[q] >= 0 && k < M) ? xu[(int64_t)row[q] * M + k] : 0.0f;
        best[q] = __int_as_float(0x7f800000);
        arg[q] = -1;
        some[q] = false;
        ok[q] = nn[q] > 0.0f && nn[q] < __int_as_float(0x7f800000);
        sthr[q] = __int_as_float(0x7f800000);
        smax[q] = __int_as_float(0xff800000);
    }
    for (int j0 = 0; j0 < r; j0 += kTile) {
        const int jn = min(kTile, r - j0);
        __syncthreads();
        for (int q = threadIdx.x; q < jn * kPad; q += 256) {
            const int j = q / kPad, k = q - j * kPad;
            ys[q] = k < M ? yu[(int64_t)(j0 + j) * M + k] : 0.0f;
        }
        __syncthreads();
#pragma unroll 2
        for (int j = slice; j < jn; j += kN3Split) {
            float y[kPad];
#pragma unroll
            for (int k4 = 0; k4 < kPad / 4; ++k4) {
                const float4 t = *reinterpret_cast<const float4 *>(ys + j * kPad + 4 * k4);
                y[4 * k4] = t.x; y[4 * k4 + 1] = t.y; y[4 * k4 + 2] = t.z; y[4 * k4 + 3] = t.w;
            }
#pragma unroll
            for (int q = 0; q < kN3Rows; ++q) {
                float c = 0.0f;
#pragma unroll
                for (int k = 0; k < M; ++k) c = fmaf(x[q][k], y[k], c);
                // d = |x| sqrt(1 - cos^2), compared exactly like the reference -- but the square root is only taken
                // when 1 - cos^2 is small enough to matter: sthr bounds from above every value whose distance could
                // still be <= the best one (two roundings of <= 2^-24 each separate d from |x| sqrt(s))
                const float sq = __fsub_rn(1.0f, __fmul_rn(c, c));
                if (ok[q]) {
                    // `seen` (a distance that is not NaN exists) only matters for rows that end without a direction;
                    // it is recovered after the loop from the largest sq (fmaxf skips NaN)
                    smax[q] = fmaxf(smax[q], sq);
                    if (sq <= sthr[q]) {
                        const float d = nn[q] * sqrtf(sq);
                        if (d < best[q]) {  // NaN compares false: skipped; strict: the first of equal distances stays
                            best[q] = d;
                            arg[q] = j0 + j;
                            sthr[q] = sq * (1.0f + 1e-6f);
                        }
                    }
                } else {  // |x| is 0, inf or NaN: no short cut
                    const float d = nn[q] * sqrtf(sq);
                    some[q] = some[q] || (d == d);
                    if (d < best[q]) {
                        best[q] = d;
                        arg[q] = j0 + j;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < kN3Rows; ++q) {
        if (ok[q]) some[q] = smax[q] >= 0.0f;  // finite |x| > 0: d is a number exactly when 1 - cos^2 >= 0
        // merge the slices: smaller distance, then smaller direction index (NaN never entered `best`)
#pragma unroll
        for (int o = 1; o < kN3Split; o <<= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best[q], o);
            const int oa = __shfl_xor_sync(0xffffffffu, arg[q], o);
            const bool os = __shfl_xor_sync(0xffffffffu, some[q] ? 1 : 0, o) != 0;
            if (oa >= 0 && (arg[q] < 0 || ob < best[q] || (ob == best[q] && oa < arg[q]))) {
                best[q] = ob;
                arg[q] = oa;
            }
            some[q] = some[q] || os;
        }
        if (row[q] < 0 || slice != 0) continue;
        // nanargmin reads NaN as +inf: a row whose distances are +inf or NaN (but not all NaN) gets direction 0 and
        // distance +inf; all NaN -> -1 -> len(ref) (nsga3.py:140-142)
        if (arg[q] < 0 && some[q]) arg[q] = 0;
        gid[row[q]] = arg[q] >= 0 ? arg[q] : r;
        gdist[row[q]] = arg[q] >= 0 ? best[q] : __int_as_float(0x7fc00000);
    }
}

// ---- nsga3.py:143-151: niche counts of the rows below / in the last front ---------------------------------
__global__ void __launch_bounds__(256)
n3_count_kernel(const int *__restrict__ rank, const int *__restrict__ gid, int n, int r, const int *__restrict__ info,
                int *__restrict__ rho, int *__restrict__ rho_last, uint8_t *__restrict__ keep) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int last = info[0], g = gid[i];
    keep[i] = rank[i] < last ? 1 : 0;
    if (g >= r) return;  // bincount(..., length=len(ref)) drops the rows without a direction
    if (rank[i] < last) atomicAdd(rho + g, 1);
    else if (rank[i] == last) atomicAdd(rho_last + g, 1);
}

// exclusive scan of rho_last -> off (one CTA)
__global__ void __launch_bounds__(1024)
n3_offsets_kernel(const int *__restrict__ rho_last, int r, int *__restrict__ off) {
    __shared__ int wsum[32];
    __shared__ int carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < r; base += 1024) {
        const int j = base + threadIdx.x;
        const int v = j < r ? rho_last[j] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int before = carry;
        for (int q = 0; q < warp; ++q) before += wsum[q];
        if (j < r) off[j] = before + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[r] = carry;
}

__global__ void __launch_bounds__(256)
n3_fill_kernel(const int *__restrict__ rank, const int *__restrict__ gid, int n, int r, const int *__restrict__ info,
               const int *__restrict__ off, int *__restrict__ cursor, int *__restrict__ members) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || rank[i] != info[0]) return;
    const int g = gid[i];
    if (g >= r) return;
    members[off[g] + atomicAdd(cursor + g, 1)] = i;
}

// ---- nsga3.py:152-211: the select loop in rounds, then the survivors in index order ------------------------
struct NicheArgs {
    int n, r, pop;
    uint32_t key0, key1;
    int *rho, *rho_last;
    const int *off;
    int *members;
    uint8_t *taken, *keep;
    const float *gdist;
    int *info;
    int *selected_idx;  // out [pop]
};

__global__ void __launch_bounds__(1024)
n3_niche_kernel(NicheArgs a) {
    __shared__ int wsum[32];
    __shared__ int sh_min, sh_selected, sh_carry;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int kInf = 0x7fffffff;
    // rows of a niche in ascending index order (they were appended through an atomic cursor)
    for (int g = t; g < a.r; g += 1024) {
        const int lo = a.off[g], hi = a.off[g + 1];
        for (int p = lo + 1; p < hi; ++p) {
            const int v = a.members[p];
            int q = p;
            while (q > lo && a.members[q - 1] > v) {
                a.members[q] = a.members[q - 1];
                --q;
            }
            a.members[q] = v;
        }
        for (int p = lo; p < hi; ++p) a.taken[p] = 0;
    }
    // selected_number = sum(rho); rho = inf where the last front offers nobody (nsga3.py:152-153)
    int mine = 0;
    for (int g = t; g < a.r; g += 1024) {
        mine += a.rho[g];
        if (a.rho_last[g] == 0) a.rho[g] = kInf;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if (lane == 0) wsum[warp] = mine;
    __syncthreads();
    if (t == 0) {
        int s = 0;
        for (int q = 0; q < 32; ++q) s += wsum[q];
        sh_selected = s;
    }
    __syncthreads();
    int rounds = 0;
    while (sh_selected < a.pop) {
        // the smallest count among the niches that still have candidates
        int v = kInf;
        for (int g = t; g < a.r; g += 1024) v = min(v, a.rho[g]);
        v = __reduce_min_sync(0xffffffffu, v);
        __syncthreads();
        if (lane == 0) wsum[warp] = v;
        if (t == 0) sh_carry = 0;
        __syncthreads();
        if (t == 0) {
            int mn = kInf;
            for (int q = 0; q < 32; ++q) mn = min(mn, wsum[q]);
            sh_min = mn;
        }
        __syncthreads();
        const int level = sh_min;
        if (level == kInf) break;  // nobody left to take (cannot happen when the last front is complete)
        const int selected0 = sh_selected, room = a.pop - selected0;
        // niches at this level, in index order: number tidx of a niche = how many such niches precede it
        for (int base = 0; base < a.r; base += 1024) {
            const int g = base + t;
            const bool at = g < a.r && a.rho[g] == level;
            const uint32_t bal = __ballot_sync(0xffffffffu, at);
            __syncthreads();
            if (lane == 0) wsum[warp] = __popc(bal);
            __syncthreads();
            int before = sh_carry;
            for (int q = 0; q < warp; ++q) before += wsum[q];
            int total = 0;
            for (int q = 0; q < 32; ++q) total += wsum[q];
            const int tidx = before + __popc(bal & ((1u << lane) - 1u));
            if (at && tidx < room) {
                const int lo = a.off[g], hi = a.off[g + 1];
                const int left = a.rho_last[g];
                int pick = -1;
                if (level == 0 || left == 1) {  // nsga3.py:179: the closest candidate (first on ties)
                    float bd = __int_as_float(0x7f800000);
                    for (int p = lo; p < hi; ++p) {
                        if (a.taken[p]) continue;
                        const float d = a.gdist[a.members[p]];
                        if (pick < 0 || d < bd) {
                            bd = d;
                            pick = p;
                        }
                    }
                } else {  // a uniformly drawn candidate: the j-th remaining row in index order
                    const uint32_t num = (uint32_t)a.pop + 1u, ki = (uint32_t)(selected0 + tidx);
                    const uint32_t kw0 = jax_split_word(a.key0, a.key1, num, 2u * ki);
                    const uint32_t kw1 = jax_split_word(a.key0, a.key1, num, 2u * ki + 1u);
                    int j = (int)jax_randint(kw0, kw1, (uint32_t)left);
                    for (int p = lo; p < hi; ++p) {
                        if (a.taken[p]) continue;
                        if (j-- == 0) {
                            pick = p;
                            break;
                        }
                    }
                }
                if (pick >= 0) {
                    a.taken[pick] = 1;
                    a.keep[a.members[pick]] = 1;  // rank[candidate] = last_rank - 1
                }
                a.rho_last[g] = left - 1;
                a.rho[g] = (left - 1 == 0) ? kInf : level + 1;
            }
            __syncthreads();
            if (t == 0) sh_carry += total;
            __syncthreads();
        }
        if (t == 0) sh_selected = selected0 + min(room, sh_carry);
        ++rounds;
        __syncthreads();
    }
    // survivors in index order (nsga3.py:203-205)
    if (t == 0) sh_carry = 0;
    __syncthreads();
    for (int base = 0; base < a.n; base += 1024) {
        const int i = base + t;
        const bool k = i < a.n && a.keep[i];
        const uint32_t bal = __ballot_sync(0xffffffffu, k);
        if (lane == 0) wsum[warp] = __popc(bal);
        __syncthreads();
        int before = sh_carry;
        for (int q = 0; q < warp; ++q) before += wsum[q];
        int total = 0;
        for (int q = 0; q < 32; ++q) total += wsum[q];
        const int pos = before + __popc(bal & ((1u << lane) - 1u));
        if (k && pos < a.pop) a.selected_idx[pos] = i;
        __syncthreads();
        if (t == 0) sh_carry += total;
        __syncthreads();
    }
    if (t == 0) a.info[2] = rounds;
}

}  // namespace evoxb
