/* b200approx.h -- C ABI of libb200approx.so: the B200 (sm_100a) implementation of
 * util.approximate's distance-based predict/fit hot path.
 *
 * The reference crosses Python -> ctypes -> Fortran once PER QUERY POINT
 * (fmodpy-generated `c_<routine>` BIND(C) symbols, every argument by pointer:
 * util/approximate/shepard/weight/weight_bind_c.f90:7-28, weight_from_py.py:54-85).
 * That per-point loop is what this library replaces, so each entry point below
 * takes the whole batch of query points.  Conventions:
 *   - plain C, no exceptions; every function returns 0 on success, non-zero on
 *     error (text via b200_last_error()).
 *   - the caller owns every host buffer; the library owns device memory behind
 *     the model handle; one handle = one fitted data set on one GPU; calls on a
 *     handle are serialised by the caller.
 *   - matrices are row-major fp64 (numpy C order: points x dimensions); indices
 *     are 0-based int64 (numpy's default), status arrays int32.
 *   - there is no CPU fallback: without a CUDA device every call fails.
 */
#ifndef B200APPROX_H
#define B200APPROX_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_model b200_model;

/* sub-model weighting used by NearestNeighbor(method=...) */
enum { B200_AVERAGE = 0, B200_SHEPARD = 1, B200_VORONOI = 2 };
/* knn path selection for b200_set_option("knn_path") */
enum { B200_PATH_AUTO = 0, B200_PATH_EXACT = 1, B200_PATH_TENSOR = 2 };

int b200_version(void);
const char* b200_last_error(void);
int b200_device_count(void);

/* Upload the fitted points x (n,d) to `device`.  Replaces the state the reference
 * models keep after _fit: NearestNeighbor.points + BallTree
 * (nearest_neighbor/__init__.py:29-35), ShepMod.x (modified_shepard/__init__.py:30),
 * LSHEP.x (linear_shepard/__init__.py:38), Voronoi.points (voronoi/__init__.py:33). */
int b200_create(const double* x, int64_t n, int32_t d, int32_t device, b200_model** out);
int b200_destroy(b200_model* m);

/* Options (b200_set_option):
 *   "knn_path"      B200_PATH_AUTO / _EXACT / _TENSOR: which kNN kernel answers (AUTO: by problem size)
 *   "margin"        extra candidates the tensor filter keeps beyond k (-1: max(4, min(k/2, 16)))
 *   "tail_split"    0: do not spread the last, partly filled round of query blocks over idle SMs
 *   "fallback"      1: uncertified queries go to the split brute-force scan instead of the rescan
 *   "tc_halves"     1 / 2: force 128 / 256 query rows per CTA of the tensor kernel (0: automatic)
 *   "tc_mode"       measurement only: 1 = no selection, 2 = no epilogue work (results meaningless)
 *   "rows_per_call" queries per internal batch of the dense scans
 *   "reset_stats"   clear the counters
 * Counters (b200_get_stat): "uncertified", "ms_prep", "ms_stage1", "ms_rerank", "ms_weights",
 * "ms_exact", "ms_collective", "launches", "tc_queries", "exact_queries", and the tensor
 * kernel's shape "tc_K", "tc_cap", "tc_keep", "tc_slots", "tc_halves", "tc_scale", "tc_ymax". */
int b200_set_option(b200_model* m, const char* key, double value);
int b200_get_stat(b200_model* m, const char* key, double* value);

/* Response values y (n,r) of the fitted points, kept on the device: what WeightedApproximator.fit
 * stores as self.y (base.py:127-143).  The predict entry points below then form the weighted sums
 * (base.py:159-169) when called with y == NULL and out != NULL, without sending y again; a call that
 * passes y uses (and keeps) that y instead.  y == NULL here forgets the stored values. */
int b200_set_values(b200_model* m, const double* y, int32_t r);

/* k nearest fitted points of each query, ascending by (squared distance, index):
 * the batched form of `self.tree.query(pt[None,:], k=..., return_distance=False)`
 * (nearest_neighbor/__init__.py:75-76).  idx (q,k); d2 (q,k) squared distances
 * (may be NULL).  self_offset >= 0 excludes data row self_offset+p for query p
 * (the "other nodes" scans of SHEPMOD / LSHEP: modified_shepard.f95:45-53).
 * A NaN or infinite query coordinate fails the call ("query points contain NaN or infinity"),
 * here and in every entry point that searches neighbours, as BallTree.query raises for such input. */
int b200_knn(b200_model* m, const double* z, int64_t q, int32_t k, int64_t self_offset,
             int64_t* idx, double* d2);

/* NearestNeighbor._predict + WeightedApproximator.predict
 * (nearest_neighbor/__init__.py:60-89, base.py:146-173): kNN, sub-model weights
 * on the k neighbours (Average :7-14, Shepard = WEIGHT shepard/weight.f90:1-34,
 * Voronoi = PREDICT voronoi/voronoi.f90:57-155), and, when y != NULL, the
 * weighted sums.  ids/wts (q,k) padded rows with cnt (q) valid entries (any may
 * be NULL); err (q) sub-model error codes (Voronoi ERROR 0/1/2); y (n,r),
 * out (q,r). */
int b200_nn_predict(b200_model* m, const double* z, int64_t q, int32_t k, int32_t method,
                    const double* y, int32_t r, int64_t* ids, double* wts, int32_t* cnt,
                    int32_t* err, double* out);

/* Stand-alone Shepard(): WEIGHT (shepard/weight.f90:1-34) of every query against ALL fitted points,
 * the batched form of Shepard._predict's loop (shepard/__init__.py:35-47).  wts (q,n) dense normalised
 * weights in point order (may be NULL); y (n,r) -> out (q,r) weighted sums (y may be NULL when wts is
 * given). */
int b200_shepard_predict(b200_model* m, const double* z, int64_t q, const double* y, int32_t r, double* wts,
                         double* out);

/* SHEPMOD (modified_shepard.f95:10-60): radius of influence rw (n); *ier 0/1. */
int b200_shepmod_fit(b200_model* m, double* rw, int32_t* ier);
/* EVAL_MESH (voronoi/voronoi.f90:5-44, fmodpy symbol c_eval_mesh): naive full
 * Voronoi mesh from precomputed inner products.  Arrays column-major as the
 * Fortran takes them: int_prods (n,n), app_int_prods (q,n) -> mesh (q,n), rows
 * normalised to sum 1.  Host pointers; no model handle. */
int b200_eval_mesh(int32_t device, const double* int_prods, int32_t n, const double* app_int_prods, int64_t q,
                   double* mesh);

/* Stand-alone Voronoi() over ALL fitted points (Voronoi._predict,
 * voronoi/__init__.py:41-62: PREDICT, voronoi/voronoi.f90:57-155, once per query
 * against the n x n inner-product cache of _fit, :32-38).  The cache is built on
 * the first call (kept in HBM while it fits 40 GB; larger sets recompute the products).  Per query the points with positive weight in
 * ascending index order: ids/wts (q,cap), cnt (q) = their number (> cap: the
 * row was truncated, call again with a larger cap), err (q) 0 / 1 overflow /
 * 2 duplicated fit points; optional y (n,r) -> out (q,r). */
int b200_voronoi_predict(b200_model* m, const double* z, int64_t q, int32_t cap, int64_t* ids,
                         double* wts, int32_t* cnt, int32_t* err, const double* y, int32_t r,
                         double* out);

/* SHEPMODVAL over a batch + ShepMod._predict's `wts > 0` filter
 * (modified_shepard.f95:83-176, modified_shepard/__init__.py:39-58).
 * ids/wts (q,cap) ascending index, cnt (q) true counts (rows with cnt > cap are
 * truncated: call again with a larger cap), ier (q) 0/1; optional y (n,r) ->
 * out (q,r). */
int b200_shepmod_predict(b200_model* m, const double* z, int64_t q, int32_t cap, int64_t* ids,
                         double* wts, int32_t* cnt, int32_t* ier, const double* y, int32_t r,
                         double* out);

/* LSHEP non-robust fit (linear_shepard.f95:23-134,177; DGELSS replaced by a batched Householder QR,
 * and by a Jacobi SVD with DGELSS's singular-value cut -- minimum-norm solution, IER=2 -- for nodes
 * at or near rank deficiency): f (n,r) response columns; a (r,n,d) gradients, rw (n) radii BEFORE
 * the wrapper's `radius` multiplier (linear_shepard/__init__.py:55); *ier 0/1/2. */
int b200_lshep_fit(b200_model* m, const double* f, int32_t r, double* a, double* rw,
                   int32_t* ier);
/* LSHEP with RLSHEP present -- the robust fit the Fortran module exports but the shipped wrapper never
 * reaches (linear_shepard.f95:136-175 with IRLS :181-282, MAD :284-308, QUANTILE :310-335): unit weights,
 * then Huber and Tukey-bisquare re-weighted least squares (five iterations each), RW cut at the first
 * neighbour whose final weight is below 0.8.  One response column f (n); a (n,d), rw (n); *ier 0/1/2/3.
 * b200_lshep_predict then evaluates this fit. */
int b200_lshep_fit_robust(b200_model* m, const double* f, double* a, double* rw, int32_t* ier);
/* RIPPLE (linear_shepard.f95:474-665; exported by the Fortran module, not reachable from the shipped
 * wrapper): least-median-style start from the best (M+1)-point subset of the neighbour chains, then the
 * two IRLS stages.  One response column f (n), d <= 16; a (n,d), rw (n); *ier 0/1/2.
 * b200_lshep_predict then evaluates this fit. */
int b200_ripple_fit(b200_model* m, const double* f, double* a, double* rw, int32_t* ier);
/* LSHEPVAL over a batch (linear_shepard.f95:365-472) with the radii scaled by
 * `radius`: out (q,r), ier (q,r) 0/1/2. */
int b200_lshep_predict(b200_model* m, double radius, const double* z, int64_t q, double* out,
                       int32_t* ier);

/* ---- device-pointer forms (inputs/outputs already in HBM on the model's GPU;
 * `stream` is a cudaStream_t, NULL = default).  No counterpart in the reference
 * (it has no device memory); used for the HBM-resident throughput figure and
 * the multi-GPU path.  idx is int32 here. */
int b200_knn_dev(b200_model* m, const double* z_dev, int64_t q, int32_t k, int64_t self_offset,
                 int32_t* idx_dev, double* d2_dev, void* stream);
int b200_nn_weights_dev(b200_model* m, const double* z_dev, int64_t q, int32_t k, int32_t method,
                        const int32_t* idx_dev, const double* d2_dev, const double* y_dev, int32_t r,
                        double* wts_dev, int32_t* cnt_dev, int32_t* err_dev, double* out_dev,
                        void* stream);
/* Data-sharded kNN merge: `parts` lists of (q,k) sorted (d2, global idx) -> the k
 * best per query by (d2, idx).  d2_parts/idx_parts are (parts,q,k) contiguous. */
int b200_merge_topk_dev(int32_t device, const double* d2_parts, const int32_t* idx_parts,
                        int32_t parts, int64_t q, int32_t k, double* d2_out, int32_t* idx_out,
                        void* stream);

/* ---- one fitted model over several GPUs of one box, driven from one process.  No counterpart in the
 * reference (single-threaded, no devices); this is BASELINE.json configs[4] behind the C-ABI.
 * shard_data = 0: every device holds all fitted points and takes a contiguous slice of the queries (no
 * exchange; the choice when the data set fits one GPU).  shard_data = 1: device g holds rows
 * [g n/G, (g+1) n/G), finds its exact local top-k of all queries, the partial lists travel in one
 * all-to-all (grouped ncclSend/ncclRecv; libnccl.so.2 is opened with dlopen at create time), and each
 * device merges and weights its slice of the queries -- bit-identical to the single-GPU result.
 * Host buffers in and out as for b200_nn_predict; Average / Shepard weighting under data sharding.
 * Stats: the devices' counters summed, plus "ms_collective", "ms_merge", "ms_local" (max over devices). */
typedef struct b200_multi b200_multi;
int b200_multi_create(const double* x, int64_t n, int32_t d, int32_t n_dev, const int32_t* devs, int32_t shard_data,
                      b200_multi** out);
int b200_multi_destroy(b200_multi* mm);
int b200_multi_set_option(b200_multi* mm, const char* key, double value);
int b200_multi_get_stat(b200_multi* mm, const char* key, double* value);
int b200_multi_set_values(b200_multi* mm, const double* y, int32_t r);
int b200_multi_nn_predict(b200_multi* mm, const double* z, int64_t q, int32_t k, int32_t method, int64_t* ids,
                          double* wts, double* out);

/* Exact kNN that allocates nothing.  It has no counterpart in the reference, which owns
 * no device memory.  Meant for callers that cannot let a library allocate while they
 * run: serving loops for which the implicit device-wide synchronisation of
 * cudaMalloc/cudaFree is a stall, and processes whose device memory is already handed
 * out in full (a pool, another framework) or audited, where a late cudaMalloc fails.  All pointers are device pointers on the CURRENT device (the
 * caller owns inputs and outputs): x_dev (n,d) fitted points, z_dev (q,d) queries,
 * idx_dev (q,k) int32 and d2_dev (q,k) results, ordered by (d2, index) like b200_knn_dev.
 * The partial top-k lists the split scan needs are kept in a 12 MiB array compiled into
 * the library (present on every GPU that loads it until the process exits), which bounds
 * the batch: q*k <= 262144 and k <= 128; larger requests are refused with an error, so
 * split the query batch.  The first call on each device looks up that array's device
 * address; make that first call outside any stream capture.  All calls share the array
 * and nothing in it has to start at zero (every slot a call reads it has written first),
 * but two calls whose kernels overlap on the GPU -- different streams or threads --
 * race on it: serialise them. */
int b200_knn_exact_noalloc(const double* x_dev, int64_t n, int32_t d, const double* z_dev,
                           int64_t q, int32_t k, int32_t* idx_dev, double* d2_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif
