/*
 * evoxb -- B200-native Pareto ranking and crowding distance (C ABI).
 *
 * This is the drop-in boundary for ONE hot path of EvoX (reference: MinyangChen/evox-myc,
 * EvoX 0.8.1): multi-objective environmental selection.  The reference has no FFI layer; its
 * boundary is two Python callables (src/evox/operators/__init__.py:1-5).  Each entry point
 * below names the reference function it replaces (paths relative to the reference tree).
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name says host; arrays are row-major,
 *     dense, and never modified unless documented as outputs;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); nothing in here
 *     calls cudaDeviceSynchronize and nothing allocates: scratch comes from the caller
 *     (`ws`, at least *_workspace_bytes() bytes, 256-byte aligned);
 *   - return value 0 = success, negative = error (EVOXB_E_*); evoxb_last_error() returns a
 *     thread-local message for the last failing call on this thread;
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef EVOXB_H_
#define EVOXB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EVOXB_VERSION 200 /* 0.2.0 */

#if defined(__GNUC__)
#define EVOXB_API __attribute__((visibility("default")))
#else
#define EVOXB_API
#endif

#define EVOXB_OK 0
#define EVOXB_E_INVALID (-1)   /* bad argument (null pointer, negative size, m out of range) */
#define EVOXB_E_WORKSPACE (-2) /* workspace missing, misaligned or too small */
#define EVOXB_E_CUDA (-3)      /* a CUDA runtime call failed; message has the CUDA error string */
#define EVOXB_E_UNSUPPORTED (-4)
#define EVOXB_E_NCCL (-5)

#define EVOXB_MAX_OBJECTIVES 32

typedef void *evoxb_stream_t; /* cudaStream_t */

EVOXB_API int evoxb_version(void);
EVOXB_API const char *evoxb_last_error(void);

/* ---- a2: dominate_relation(x, y)  -- src/evox/utils/common.py:95-104 --------------------
 * out[i * ny + j] = 1 iff x_i dominates y_j: all_k(x_ik <= y_jk) && any_k(x_ik < y_jk),
 * IEEE compares (NaN => 0, -0 == +0).  x is (nx, m), y is (ny, m), out is (nx, ny) bytes.   */
EVOXB_API int evoxb_dominate_relation_f32(const float *x, int64_t nx, const float *y, int64_t ny, int m,
                                uint8_t *out, evoxb_stream_t stream);

/* ---- a2+a3+a4: non_dominated_sort(x) -> rank ---------------------------------------------
 * src/evox/operators/selection/non_dominate.py:30-113 (dominance matrix :60, dominator count
 * :61, front-peeling while_loop :72-111).  fitness is (n, m) float32, rank is (n) int32,
 * 0-based.  Bit-exact with the reference for every input, NaN / +-inf / duplicates included.
 * The dominance matrix lives in `ws` as one bit per ordered pair (n * roundup(n,4096) / 8 bytes). */
EVOXB_API int evoxb_nds_workspace_bytes(int64_t n, int m, size_t *bytes);
EVOXB_API int evoxb_nds_rank_f32(const float *fitness, int64_t n, int m, int32_t *rank, void *ws,
                       size_t ws_bytes, evoxb_stream_t stream);
/* How the dominance matrix of the last evoxb_nds_rank_f32 / evoxb_nds_rank_sharded_f32 call on `ws` was built
 * (no counterpart in the reference; for tests and benchmark labels): 0 = float compares, 1 = 15-bit levels +
 * fix-up (m >= 4 from 65536 rows on; environment EVOXB_BUILD_MODE=levels|exact overrides), 2 = levels were
 * prepared but judged too coarse for this population on the device, and the float kernel ran.  The result of
 * the call is the same in all three.  Synchronises `stream` (one 4-byte copy).  With ws == NULL nothing is
 * read from the device: path = 1 if the levels would be used for this shape, else 0.                        */
EVOXB_API int evoxb_nds_build_path(const void *ws, int64_t n, int m, evoxb_stream_t stream, int *path);

/* ---- a5: crowding_distance(costs, mask) -> dist ---------------------------------------------
 * src/evox/operators/selection/non_dominate.py:116-156.  fitness is (n, m) float32; mask is (n)
 * bytes (non-zero = participates) or NULL for "all rows"; dist is (n) float32: +inf at both ends
 * of every objective's sorted order, -inf for masked-out rows, NaN where the reference gives NaN
 * (range == 0).  Stable per-objective sort (NaN last, -0 == +0) as jnp.lexsort((cost, ~mask)).   */
EVOXB_API int evoxb_crowding_workspace_bytes(int64_t n, int m, size_t *bytes);
EVOXB_API int evoxb_crowding_f32(const float *fitness, const uint8_t *mask, int64_t n, int m, float *dist,
                                 void *ws, size_t ws_bytes, evoxb_stream_t stream);

/* ---- crowding distance without any allocation (no counterpart in the reference) ---------------
 * Some callers of this library cannot obtain device memory at the moment they need a result: a
 * custom-call handler running inside a compiled step only owns the buffers its runtime handed it
 * (a cudaMalloc there would stall every stream of the device and bypass the runtime's memory
 * accounting), and a host loop that has committed all of its memory budget has nothing left to
 * pass as a workspace.  For them this variant of evoxb_crowding_f32 keeps its sort buffers in an
 * 8 MiB array that is part of the library itself; it takes no workspace and calls no cudaMalloc,
 * cudaFree or other allocating routine.  Inputs, outputs and results are those of
 * evoxb_crowding_f32.
 *   - Limit: sizes whose evoxb_crowding_workspace_bytes() exceed EVOXB_CROWDING_NOALLOC_MAX_BYTES
 *     (8 MiB: e.g. n = 65536 with m <= 6, n = 32768 with m <= 12) are rejected with
 *     EVOXB_E_WORKSPACE and a message naming both numbers; nothing is launched.
 *   - The first call on each device looks up the array's device address (with that device
 *     current); make that first call before starting any stream capture on the device.
 *   - Every GPU that loads the library holds the 8 MiB until the process exits.  The driver reserves
 *     them when the library's module is first loaded on a device (with CUDA's lazy loading that is the
 *     first call into the library on that device, possibly this one), so on a device whose memory is
 *     already exhausted that first call can itself fail with EVOXB_E_CUDA: make one warm-up call per
 *     device while memory is still available.
 *   - All calls on a device share the one array and it keeps the previous call's contents; each
 *     call writes everything it later reads, so nothing needs clearing -- but two calls that
 *     overlap on the GPU (different streams, or different host threads on the same device)
 *     race on it.  Serialise them, or use evoxb_crowding_f32 with separate workspaces.          */
#define EVOXB_CROWDING_NOALLOC_MAX_BYTES (8u << 20)
EVOXB_API int evoxb_crowding_noalloc_f32(const float *fitness, const uint8_t *mask, int64_t n, int m,
                                         float *dist, evoxb_stream_t stream);

/* ---- a7: NSGA-II survivor selection ------------------------------------------------------------
 * src/evox/algorithms/mo/nsga2.py:90-96 (same block: gde3.py:88-97, im_moea.py:131-141):
 * rank = non_dominated_sort(fitness); worst_rank = rank[argsort(rank)[pop_size]];
 * crowd = crowding_distance(fitness, rank == worst_rank);
 * order = lexsort((-crowd, rank))[:pop_size].  Requires 0 <= pop_size < n like the reference's
 * indexing.  order is (pop_size) int32; rank_out (n) and crowd_out (n) are optional (may be NULL). */
EVOXB_API int evoxb_nsga2_select_workspace_bytes(int64_t n, int m, size_t *bytes);
EVOXB_API int evoxb_nsga2_select_f32(const float *fitness, int64_t n, int m, int64_t pop_size, int32_t *order,
                                     int32_t *rank_out, float *crowd_out, void *ws, size_t ws_bytes,
                                     evoxb_stream_t stream);

/* ---- a8: the cut-off rank, `rank[argsort(rank)[k]]` -------------------------------------------------
 * src/evox/algorithms/mo/nsga3.py:97-98 (k = pop_size) and nsga2.py:91-92: the value of the k-th
 * smallest rank (0-based), from a rank histogram instead of an argsort.  out is one device int32.   */
EVOXB_API int evoxb_kth_rank_workspace_bytes(int64_t n, size_t *bytes);
EVOXB_API int evoxb_kth_rank_i32(const int32_t *rank, int64_t n, int64_t k, int32_t *out, void *ws,
                                 size_t ws_bytes, evoxb_stream_t stream);

/* ---- a6: operator form non_dominate(population, fitness, topk) ------------------------------------
 * src/evox/operators/selection/non_dominate.py:181-195, including its quirks: the cut-off rank is
 * rank[argsort(rank)[topk-1]] (:190) and the secondary sort key is -argsort(crowding distance)
 * used positionally (:191-194).  Requires 1 <= topk <= n.  order is (topk) int32.  Workspace size
 * is evoxb_nsga2_select_workspace_bytes().                                                         */
EVOXB_API int evoxb_non_dominate_select_f32(const float *fitness, int64_t n, int m, int64_t topk, int32_t *order,
                                            void *ws, size_t ws_bytes, evoxb_stream_t stream);

/* ---- a6: crowding_distance_sort = jnp.argsort(distance)  (non_dominate.py:159-178) -------------------
 * Stable ascending argsort of n floats with lax.sort ordering (NaN last, -0 == +0).                  */
EVOXB_API int evoxb_argsort_workspace_bytes(int64_t n, size_t *bytes);
EVOXB_API int evoxb_argsort_f32(const float *x, int64_t n, int32_t *idx, void *ws, size_t ws_bytes,
                                evoxb_stream_t stream);

/* survivor = merged_pop[order]  (nsga2.py:97-98): dst[r, :] = src[idx[r], :], src is (n, d), dst (k, d). */
EVOXB_API int evoxb_gather_rows_f32(const float *src, int64_t n, int64_t d, const int32_t *idx, int64_t k,
                                    float *dst, evoxb_stream_t stream);

/* ---- next row 2: what NSGA-II does between two selections -----------------------------------------
 * evoxb_sbx_pm_clip_f32 = NSGA2.ask (src/evox/algorithms/mo/nsga2.py:76-84): simulated binary
 * crossover of rows [0, n/2) with rows [n/2, 2*(n/2)) (operators/crossover/sbx.py:8-41, type 1),
 * polynomial mutation (operators/mutation/pm_mutation.py:6-54) and clip to [lb, ub].  pop, out are
 * (n, d); lb, ub are (d).  Random numbers are Philox-4x32-10 keyed by (seed, generation): the same
 * arguments give the same offspring; JAX's threefry streams are not reproduced (statistical parity).
 * evoxb_dtlz2_f32 = DTLZ2.evaluate (src/evox/problems/numerical/dtlz.py:84-103): x (n, d) -> f (n, m). */
EVOXB_API int evoxb_sbx_pm_clip_f32(const float *pop, int64_t n, int d, const float *lb, const float *ub,
                                    float pro_c, float dis_c, float pro_m, float dis_m, uint32_t seed,
                                    uint32_t generation, float *out, evoxb_stream_t stream);
EVOXB_API int evoxb_dtlz2_f32(const float *x, int64_t n, int d, int m, float *f, evoxb_stream_t stream);
/* evoxb_lsmop1_f32 = LSMOP1.evaluate (src/evox/problems/numerical/lsmop.py:127-141 with _calc_g :88-111 and the
 * sphere inner function, problems/numerical/sphere.py:5-6): x (n, d) -> f (n, m), 2 <= m <= 16.
 * sublen[i] / start[i] (host arrays of m ints) are the reference's `self.sublen[i]` and `self.len[i]`
 * (lsmop.py:44-54: logistic-map weights, floor(c / sum(c) * d / 5), cumulative sum of 5 * sublen).        */
EVOXB_API int evoxb_lsmop1_f32(const float *x, int64_t n, int d, int m, const int32_t *sublen, const int32_t *start,
                               float *f, evoxb_stream_t stream);

/* ---- f3: NSGA-III environmental selection: all of NSGA3.tell after the concatenation
 * (src/evox/algorithms/mo/nsga3.py:96-211): ranking, last_rank, normalisation by the hyperplane intercepts (or
 * the worst point when the extreme points are rank deficient), association of every kept row with its nearest
 * reference direction (ref: (r, m), e.g. UniformSampling, src/evox/operators/sampling/uniform.py:10-60) and the
 * niche-preserving select loop.  key = the two words of the JAX PRNG key in state.key; the loop draws with
 * jax.random.randint(keys[selected_number], (), 0, rho_last[group]) of keys = jax.random.split(key, pop_size + 1)
 * (threefry, JAX 0.4.x layout) and next_key (host, optional) receives keys[0] (:209).
 *   selected_idx   : out, device (pop_size) int32, ascending row indices of the survivors (:203-205)
 *   rank_out       : optional out, device (n) int32
 *   group_id_out   : optional out, device (n) int32, nearest direction (r = none, :141)
 *   group_dist_out : optional out, device (n) float32 (NaN = none)
 *   norm_out       : optional out, device (3 m) float32: ideal point, nadir point, extreme row indices
 * m <= 16.  Stream-ordered; nothing is allocated; host arguments (key, next_key) are read / written at call time. */
EVOXB_API int evoxb_nsga3_select_workspace_bytes(int64_t n, int m, int64_t r, size_t *bytes);
EVOXB_API int evoxb_nsga3_select_f32(const float *fitness, int64_t n, int m, int64_t pop_size, const float *ref, int64_t r,
                                     const uint32_t key[2], int32_t *selected_idx, uint32_t next_key[2], int32_t *rank_out,
                                     int32_t *group_id_out, float *group_dist_out, float *norm_out, void *ws,
                                     size_t ws_bytes, evoxb_stream_t stream);

/* ---- e: multi-GPU ranking, one process per GPU (no counterpart in the reference, whose tell() is
 * replicated on every device: src/evox/workflows/std_workflow.py:194, :341) ---------------------------
 * The bit-packed dominance matrix is sharded by COLUMN chunk (4096 dominated individuals per chunk,
 * chunks dealt cyclically to the ranks); every rank keeps all rows of its chunks and the dominator
 * counts of its columns, so counts never cross GPUs.  Per front the ranks exchange one bit per
 * individual (N/8 bytes in total).  Every rank receives the full rank vector.  Three exchanges exist:
 *   2  data-flow kernel (peel_flow.cuh): one persistent kernel per GPU, tagged 8-byte words stored
 *      straight into the peers' memory (CUDA IPC over NVLink), no grid-wide barrier;
 *   1  grid-wide peer-memory kernel (peel.cuh): same transport, two grid syncs per front;
 *   0  one ncclAllGather per front, host check every 4 fronts (when peers cannot be mapped).
 * The library picks 2 or 1 from the population size per GPU (both stream-ordered, no host sync) and
 * falls back to 0.  NCCL is loaded at run time (libnccl.so.2 of the process, or $EVOXB_NCCL_LIB).
 *   evoxb_comm_unique_id : rank 0 makes the 128-byte NCCL id; the caller broadcasts it (any transport)
 *   evoxb_comm_create    : collective; call with the GPU of this rank current
 * The sharded entry points must be called by all ranks with the same arguments.  Supports m <= 16.
 * Peer time-out: a kernel that waits longer than the time-out (default 4 s, evoxb_comm_set_timeout_ms)
 * for a peer gives up: the call's ranks are all -1, the error is sticky (evoxb_comm_poll_error != 0)
 * and every later sharded call on the communicator returns EVOXB_E_NCCL without launching; destroy the
 * communicator on every rank and create a new one.                                                   */
typedef struct evoxb_comm evoxb_comm_t;
EVOXB_API int evoxb_comm_unique_id(void *id128);
EVOXB_API int evoxb_comm_create(const void *id128, int rank, int nranks, evoxb_comm_t **out);
EVOXB_API int evoxb_comm_destroy(evoxb_comm_t *comm);
/* p2p_enabled: 1 when the ranks exchange the front bitmask through peer memory inside one persistent
 * kernel (CUDA IPC + NVLink; the default when every rank can map every peer), 0 when they fall back
 * to one ncclAllGather per front.  error: 1 if a peer did not answer within ~4 s during the last
 * call.  fronts: number of fronts of the last peer-memory call.  Synchronous (small D2H copies).    */
EVOXB_API int evoxb_comm_status(evoxb_comm_t *comm, int *p2p_enabled, int *error, int *fronts);
/* 1 after a peer time-out on this communicator, else 0.  Reads a host word the kernels raise: no device
 * synchronisation, so it is meaningful for calls whose work has completed (after a stream sync).      */
EVOXB_API int evoxb_comm_poll_error(evoxb_comm_t *comm);
/* How long a kernel waits for a peer GPU before giving up (first calls can be skewed by seconds of
 * lazy initialisation on another rank).  Same value on every rank; takes effect from the next call.  */
EVOXB_API int evoxb_comm_set_timeout_ms(evoxb_comm_t *comm, int64_t ms);
/* Exchange used by the last sharded call on this communicator: 0, 1 or 2 as listed above.           */
EVOXB_API int evoxb_comm_last_path(evoxb_comm_t *comm);
EVOXB_API int evoxb_shard_plan(int64_t n, int rank, int nranks, int *first_chunk, int *chunk_stride,
                               int *local_chunks, int *total_chunks);
/* Column ownership of the data-flow peel kernel on `rank` of `nranks` with at most max_ctas CTAs (host arithmetic
 * only; no counterpart in the reference).  A matrix row is cut in units of 256 columns, a chunk in 1..16 pieces;
 * pieces (optional) receives 4 ints per piece: owning CTA, global chunk, first unit, number of units.          */
EVOXB_API int evoxb_flow_plan(int64_t n, int rank, int nranks, int tri, int max_ctas, int *grid, int *npieces,
                              int *max_words, int32_t *pieces, int pieces_cap);
EVOXB_API int evoxb_nds_sharded_workspace_bytes(int64_t n, int m, int rank, int nranks, size_t *bytes);
EVOXB_API int evoxb_nds_rank_sharded_f32(const float *fitness, int64_t n, int m, int32_t *rank, void *ws,
                                         size_t ws_bytes, evoxb_comm_t *comm, evoxb_stream_t stream);
/* evoxb_nsga2_select_f32 with the ranking sharded; the crowding / selection tail is replicated. */
EVOXB_API int evoxb_nsga2_select_sharded_f32(const float *fitness, int64_t n, int m, int64_t pop_size,
                                             int32_t *order, int32_t *rank_out, float *crowd_out, void *ws,
                                             size_t ws_bytes, evoxb_comm_t *comm, evoxb_stream_t stream);

/* ---- profiling hook (no counterpart in the reference) --------------------------------------
 * When enabled, every compute entry point brackets its stages with CUDA events on the caller's
 * stream.  evoxb_profile_read() synchronises those events and returns, for the LAST call made
 * on this thread, up to `cap` stage times in milliseconds and their names (static strings).
 * Returns the number of stages.                                                              */
EVOXB_API int evoxb_profile_enable(int on);
EVOXB_API int evoxb_profile_read(float *ms, const char **names, int cap);
/* Number of kernels this library has launched since it was loaded (a statistic for bench.py). */
EVOXB_API unsigned long long evoxb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* EVOXB_H_ */
