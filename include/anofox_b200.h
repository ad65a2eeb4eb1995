/*
 * anofox_b200.h -- C ABI of the B200-native per-group OLS / Ridge / WLS fit path.
 *
 * Two layers live in this one header:
 *
 *  (1) DROP-IN LAYER.  The seven symbols the reference's DuckDB aggregate code
 *      (src/aggregate_functions/{ols,ridge,wls}_aggregate.cpp and
 *      *_predict_aggregate.cpp) binds from its Rust static library.  Names, POD
 *      layouts, by-value passing, malloc/free ownership and error behaviour follow
 *      the reference's hand-kept header src/include/anofox_stats_ffi.h
 *      (types :18-141, :351-364, :446-457, :660-667; functions :155, :161, :166,
 *      :378, :472, :655, :686) and its Rust side
 *      crates/anofox-stats-ffi/src/lib.rs (:98, :272, :289, :984, :1384, :2217, :2266).
 *      A reference build that links libanofox_stats_b200.so instead of
 *      libanofox_stats_ffi.a resolves these seven symbols unchanged.
 *
 *  (2) BATCH LAYER (new; the reference has no equivalent).  DuckDB hands Finalize up
 *      to 2048 group states per call (ols_aggregate.cpp:249-257); a per-group launch
 *      cannot feed a B200, so the product entry point takes ALL groups of a query at
 *      once as a group-sorted CSR: offsets[G+1], y[N], X[N*p] row-major (the layout of
 *      DuckDB's LIST(DOUBLE) child vector, ols_aggregate.cpp:181-183), optional w[N].
 *      Results come back as struct-of-arrays -- one array per field of the reference's
 *      result STRUCT (ols_aggregate.cpp:74-96) plus a per-group status that carries
 *      the NULL-struct decision (ols_aggregate.cpp:263-267, :298-301).
 *
 * No torch types, no C++ types: plain pointers and sizes only.
 * There is no CPU fallback behind any of these calls: without a CUDA device they fail.
 */
#ifndef ANOFOX_B200_H
#define ANOFOX_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * (1) Drop-in layer: POD types shared with the reference (anofox_stats_ffi.h)
 * ---------------------------------------------------------------------------------------- */

/* anofox_stats_ffi.h:18-31 / types.rs:6-21 */
typedef enum {
	ANOFOX_ERROR_SUCCESS = 0,
	ANOFOX_ERROR_INVALID_INPUT = 1,
	ANOFOX_ERROR_SINGULAR_MATRIX = 2,
	ANOFOX_ERROR_CONVERGENCE_FAILURE = 3,
	ANOFOX_ERROR_INVALID_ALPHA = 4,
	ANOFOX_ERROR_INVALID_L1_RATIO = 5,
	ANOFOX_ERROR_INSUFFICIENT_DATA = 6,
	ANOFOX_ERROR_ALLOCATION_FAILURE = 7,
	ANOFOX_ERROR_SERIALIZATION_ERROR = 8,
	ANOFOX_ERROR_DIMENSION_MISMATCH = 9,
	ANOFOX_ERROR_NO_VALID_DATA = 10,
	ANOFOX_ERROR_INTERNAL = 99
} AnofoxErrorCode;

/* anofox_stats_ffi.h:36-39: code + NUL-terminated message (255 chars max) */
typedef struct {
	AnofoxErrorCode code;
	char message[256];
} AnofoxError;

/* anofox_stats_ffi.h:44-51: one f64 column. validity bit i (LSB-first within each byte)
 * = 1 when data[i] is non-NULL; validity == NULL means "all valid". NULL reads as NaN
 * (types.rs:79-89) and such rows are then dropped by the finite-row filter. */
typedef struct {
	const double *data;
	const uint8_t *validity;
	size_t len;
} AnofoxDataArray;

/* anofox_stats_ffi.h:56-73. coefficients has n_features entries; aliased / constant
 * columns hold NaN. intercept is NaN when fit_intercept == false. */
typedef struct {
	double *coefficients;
	size_t coefficients_len;
	double intercept;
	double r_squared;
	double adj_r_squared;
	double residual_std_error;
	size_t n_observations;
	size_t n_features;
} AnofoxFitResultCore;

/* anofox_stats_ffi.h:78-97. Five arrays of `len` (= n_features, no intercept entry). */
typedef struct {
	double *std_errors;
	double *t_values;
	double *p_values;
	double *ci_lower;
	double *ci_upper;
	size_t len;
	double confidence_level;
	double f_statistic;
	double f_pvalue;
} AnofoxFitResultInference;

/* anofox_stats_ffi.h:102-106 */
typedef enum { ANOFOX_SOLVER_QR = 0, ANOFOX_SOLVER_SVD = 1, ANOFOX_SOLVER_CHOLESKY = 2 } AnofoxSolverType;
/* anofox_stats_ffi.h:111-114 */
typedef enum { ANOFOX_LAMBDA_SCALING_RAW = 0, ANOFOX_LAMBDA_SCALING_GLMNET = 1 } AnofoxLambdaScaling;
/* anofox_stats_ffi.h:119-125 */
typedef enum {
	ANOFOX_HC_NONE = 0,
	ANOFOX_HC_HC0 = 1,
	ANOFOX_HC_HC1 = 2,
	ANOFOX_HC_HC2 = 3,
	ANOFOX_HC_HC3 = 4
} AnofoxHcType;

/* anofox_stats_ffi.h:130-141 */
typedef struct {
	bool fit_intercept;
	bool compute_inference;
	double confidence_level;
	AnofoxSolverType solver;
	AnofoxHcType hc_type;
} AnofoxOlsOptions;

/* anofox_stats_ffi.h:351-364 */
typedef struct {
	double alpha;
	bool fit_intercept;
	bool compute_inference;
	double confidence_level;
	AnofoxSolverType solver;
	AnofoxLambdaScaling lambda_scaling;
} AnofoxRidgeOptions;

/* anofox_stats_ffi.h:446-457 (same layout as the OLS options) */
typedef struct {
	bool fit_intercept;
	bool compute_inference;
	double confidence_level;
	AnofoxSolverType solver;
	AnofoxHcType hc_type;
} AnofoxWlsOptions;

/* anofox_stats_ffi.h:660-667 */
typedef struct {
	double yhat;
	double yhat_lower;
	double yhat_upper;
} AnofoxPredictionResult;

/* One fit of one group, column-major input (x = array of x_count columns of length y.len).
 * Replaces lib.rs:98-265 / :984-1155 / :1384-1555. Returns true on success; on failure
 * *out_error (if non-NULL) holds the code and nothing is left allocated. Arrays in the
 * outputs are malloc()ed here and released with the two free functions below.
 * out_inference may be NULL. When inference was requested but the model is the
 * intercept-only shortcut (all columns constant) the inference struct is zero-length with
 * NaN F statistics (lib.rs:251-253). hc_type HC0 .. HC3 (OLS / WLS, with compute_inference) replaces the
 * standard errors, t, p and confidence bounds by their heteroscedasticity-consistent versions (ols.rs:209-245,
 * wls.rs:230-266; csrc/hc.cuh); the F statistic stays classical, as in the reference. */
bool anofox_ols_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxOlsOptions options,
                    AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference, AnofoxError *out_error);
bool anofox_ridge_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxRidgeOptions options,
                      AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference, AnofoxError *out_error);
bool anofox_wls_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxDataArray weights,
                    AnofoxWlsOptions options, AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference,
                    AnofoxError *out_error);

/* lib.rs:272-311: free() the arrays and null the pointers; NULL-safe. */
void anofox_free_result_core(AnofoxFitResultCore *result);
void anofox_free_result_inference(AnofoxFitResultInference *result);

/* lib.rs:2217-2231: two-sided Student-t critical value t_{(1+level)/2, df}; NaN when df == 0
 * or level outside (0,1). Evaluated on the device (same routine the fit kernels use). */
double anofox_t_critical(double confidence_level, size_t df);

/* lib.rs:2266-2350: yhat = intercept (0 if NaN) + sum over non-NaN coefficients; interval
 * yhat +- t_crit(df) * rse * sqrt(1 + 1/n), df = n - len - (intercept given ? 1 : 0); the
 * interval collapses to yhat when rse is NaN or <= 0, n <= len + 1, or df == 0. */
bool anofox_predict_with_interval(const double *coefficients, size_t coefficients_len, double intercept,
                                  const double *x_new, size_t x_len, double residual_std_error,
                                  size_t n_observations, double confidence_level,
                                  AnofoxPredictionResult *out_result);

/* lib.rs:1932-2010 (information_criteria.rs:15-87): n ln(RSS/n) + 2k, n ln(RSS/n) + k ln n,
 * RSS == 0 -> -inf; n == 0 or RSS < 0 -> INVALID_INPUT. Host arithmetic (scalar helpers);
 * the batch layer computes the same formulas on the device per group. */
bool anofox_compute_aic(double rss, size_t n, size_t k, double *out_aic, AnofoxError *out_error);
bool anofox_compute_bic(double rss, size_t n, size_t k, double *out_bic, AnofoxError *out_error);

/* ------------------------------------------------------------------------------------------
 * (2) Batch layer
 * ---------------------------------------------------------------------------------------- */

typedef enum { ANOFOX_B200_OLS = 0, ANOFOX_B200_RIDGE = 1, ANOFOX_B200_WLS = 2 } AnofoxB200Method;

/* Options of one batched call = the bind data of one aggregate (ols_aggregate.cpp:47-69). */
typedef struct {
	int32_t method;            /* AnofoxB200Method */
	int32_t fit_intercept;     /* 0 / 1 */
	int32_t compute_inference; /* 0 / 1 */
	int32_t solver;            /* AnofoxSolverType: accepted, all three must agree to 1e-10
	                              (test_map_options.test); one device solver serves them */
	int32_t lambda_scaling;    /* AnofoxLambdaScaling (ridge) */
	int32_t hc_type;           /* AnofoxHcType: robust standard errors for OLS / WLS (ignored by ridge, which
	                              has no such option: ridge.rs); outside 0..4 -> INVALID_INPUT */
	double confidence_level;
	double alpha;              /* ridge penalty; < 0 -> INVALID_ALPHA for the whole call */
} AnofoxB200FitOptions;

/* Per-group status values written to AnofoxB200FitOutputs.status: 0 = a result struct
 * exists; anything else = the reference yields a NULL struct for this group and the value
 * says why (an AnofoxErrorCode, or ANOFOX_B200_STATUS_TOO_FEW_ROWS for the aggregate's own
 * "fewer than 2 rows in the state" rule, ols_aggregate.cpp:263-267). */
#define ANOFOX_B200_STATUS_OK 0
#define ANOFOX_B200_STATUS_TOO_FEW_ROWS 1000

/* Struct-of-arrays outputs, caller-allocated. G = n_groups, p = n_features.
 * Required: coefficients .. status. The inference block is required iff
 * compute_inference; aic/bic/rss are optional (NULL = skip).
 * For the *_host entry points these are host pointers, for *_device device pointers. */
typedef struct {
	double *coefficients;       /* [G*p] row-major; NaN = aliased / constant column      */
	double *intercept;          /* [G]   NaN when fit_intercept == 0                      */
	double *r_squared;          /* [G]                                                    */
	double *adj_r_squared;      /* [G]                                                    */
	double *residual_std_error; /* [G]                                                    */
	int64_t *n_observations;    /* [G]   rows that survived the NULL / NaN / weight filter */
	int32_t *status;            /* [G]   see above                                         */
	/* inference (ols_aggregate.cpp:86-94) */
	double *std_errors;         /* [G*p] */
	double *t_values;           /* [G*p] */
	double *p_values;           /* [G*p] */
	double *ci_lower;           /* [G*p] */
	double *ci_upper;           /* [G*p] */
	double *f_statistic;        /* [G]   */
	double *f_pvalue;           /* [G]   */
	uint8_t *has_inference;     /* [G]   0 = intercept-only shortcut: empty lists, F = NaN */
	/* optional extras named by the north star (AIC/BIC on device) */
	double *aic;                /* [G] k = estimated parameters incl. intercept */
	double *bic;                /* [G] */
	double *rss;                /* [G] weighted residual sum of squares */
} AnofoxB200FitOutputs;

/* Outputs of the fit+predict path (*_fit_predict_agg, ols_predict_aggregate.cpp:322-426):
 * one entry per input row, in input order. */
typedef struct {
	double *yhat;         /* [N] NaN = NULL triple (non-finite prediction or NULL group)   */
	double *yhat_lower;   /* [N] */
	double *yhat_upper;   /* [N] */
	uint8_t *is_training; /* [N] */
} AnofoxB200PredictOutputs;

/* Number of bytes of device scratch a *_device call needs for (n_groups, n_rows, p). */
size_t anofox_b200_workspace_bytes(int64_t n_groups, int64_t n_rows, int32_t n_features);

/* Device-resident batch fit. All data pointers are device pointers on the current device:
 *   offsets   [G+1] int64, offsets[0] == 0 .. offsets[G] == n_rows, non-decreasing
 *   y         [N]    f64   (NaN = NULL / invalid -> row dropped)
 *   x         [N*p]  f64   row-major
 *   w         [N]    f64 or NULL (required for WLS; rows with w <= 0 or non-finite dropped)
 *   workspace >= anofox_b200_workspace_bytes(..) bytes, 256-byte aligned
 *   stream    a cudaStream_t (NULL = legacy default stream)
 * The call only enqueues work on `stream`; it returns 0 on success or an AnofoxErrorCode
 * (message in *err) for call-level errors (bad arguments, alpha < 0, hc_type, CUDA error). */
int32_t anofox_b200_fit_batch_device(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                     const double *w, int64_t n_rows, int32_t n_features,
                                     const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out,
                                     void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err);

/* Same contract with HOST pointers for inputs and outputs: the library owns device
 * buffers, copies in (chunked, overlapped with compute), runs the same kernels and copies
 * the result arrays back. Blocks until the outputs are complete. This is the call a DuckDB
 * Finalize makes and the one bench.py's `e2e` times. */
int32_t anofox_b200_fit_batch_host(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                   const double *w, int64_t n_rows, int32_t n_features,
                                   const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out,
                                   AnofoxError *err);

/* Fit + predict (engine of the *_fit_predict_by macros, fit_predict_macros.cpp:29-44).
 * Every row is an output row; `train` [N] (u8, may be NULL = "y finite") marks training
 * rows as decided by the aggregate's Update (ols_predict_aggregate.cpp:216-249); rows with
 * NaN y or NaN features never train. Fit runs without inference (…:354-364); each row gets
 * yhat and the simplified interval of anofox_predict_with_interval. `fit` may be NULL or
 * receive the per-group core results as well. */
int32_t anofox_b200_fit_predict_batch_device(const int64_t *offsets, int64_t n_groups, const double *y,
                                             const double *x, const double *w, const uint8_t *train,
                                             int64_t n_rows, int32_t n_features,
                                             const AnofoxB200FitOptions *options,
                                             const AnofoxB200PredictOutputs *pred, const AnofoxB200FitOutputs *fit,
                                             void *workspace, size_t workspace_bytes, void *stream,
                                             AnofoxError *err);
int32_t anofox_b200_fit_predict_batch_host(const int64_t *offsets, int64_t n_groups, const double *y,
                                           const double *x, const double *w, const uint8_t *train, int64_t n_rows,
                                           int32_t n_features, const AnofoxB200FitOptions *options,
                                           const AnofoxB200PredictOutputs *pred, const AnofoxB200FitOutputs *fit,
                                           AnofoxError *err);

/* Ingest: unsorted rows -> the group-sorted CSR the batch entry points take. Replaces what *AggUpdate / *AggCombine
 * do per row on the host (src/aggregate_functions/ols_aggregate.cpp:120-234): the rows of a query stay in scan order,
 *   group_ids [N] int64 in [0, n_groups)  (e.g. the dense ids of DuckDB's group hash table)
 *   y [N], x [N*p] row-major, w [N] or NULL
 * and a STABLE radix partition on the device writes offsets_out [G+1], y_out, x_out, w_out in group order; inside a
 * group the rows keep their arrival order (the "state order" of the reference: first valid row of the constant-column
 * rule, order of the fit_predict output list). row_index_out [N] (optional) = the source row of every sorted row, for
 * scattering per-row outputs back. At most 2^32 - 1 rows and groups per call. Only enqueues work on `stream`. */
size_t anofox_b200_ingest_workspace_bytes(int64_t n_rows, int64_t n_groups);
int32_t anofox_b200_ingest_device(const int64_t *group_ids, const double *y, const double *x, const double *w, int64_t n_rows,
                                  int32_t n_features, int64_t n_groups, int64_t *offsets_out, double *y_out, double *x_out,
                                  double *w_out, int64_t *row_index_out, void *workspace, size_t workspace_bytes, void *stream,
                                  AnofoxError *err);
/* 1 when the last ingest on this workspace met an id outside [0, n_groups) (those rows were filed under group 0);
 * synchronising read of one flag. */
int32_t anofox_b200_ingest_bad_ids(const void *workspace, int64_t n_rows, int64_t n_groups);

/* Ingest WITHOUT any allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and result (bit for bit) as anofox_b200_ingest_device, but the scratch of the radix partition is
 * a file-scope device array inside the library (ANOFOX_B200_INGEST_STATIC_MAX_BYTES = 16 MiB, held by every GPU that
 * loads the library until the process exits) instead of a caller workspace. The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine) and only enqueues work on `stream`.
 *   limit   anofox_b200_ingest_workspace_bytes(n_rows, n_groups) <= ANOFOX_B200_INGEST_STATIC_MAX_BYTES -- about
 *           16.5 bytes per row + 4 per group, e.g. 900 000 rows x 100 000 groups; beyond it the call returns
 *           ANOFOX_ERROR_INVALID_INPUT with a message that names both sizes.
 *   sharing every call on a device uses the SAME array: calls that overlap on the GPU (two streams, two threads) race;
 *           serialise them. The array keeps the last call's contents; the call clears what it needs at zero itself.
 *   bad_ids_out  optional DEVICE int32: set to 1 when an id was outside [0, n_groups) (what anofox_b200_ingest_bad_ids
 *           reports for the workspace form), else 0; written on `stream`. */
#define ANOFOX_B200_INGEST_STATIC_MAX_BYTES (16u * 1024u * 1024u)
int32_t anofox_b200_ingest_device_static(const int64_t *group_ids, const double *y, const double *x, const double *w,
                                         int64_t n_rows, int32_t n_features, int64_t n_groups, int64_t *offsets_out,
                                         double *y_out, double *x_out, double *w_out, int64_t *row_index_out,
                                         int32_t *bad_ids_out, void *stream, AnofoxError *err);

/* Scalar array fits, batched: {ols,ridge,wls}_fit(y LIST, x LIST(LIST) [, w LIST]) computes ONE fit per SQL row from
 * column arrays (src/table_functions/ols_fit.cpp:120-173 makes one anofox_ols_fit call per row of the DataChunk).
 * This entry point takes all rows of a chunk (or of a query) at once:
 *   offsets   [F+1]  rows of fit i are [offsets[i], offsets[i+1])  (device int64)
 *   y         [N]    the concatenated y lists (the LIST(DOUBLE) child vector)
 *   x_columns [N*p]  per fit a block of p columns of n_i values each, blocks in fit order -- the layout of the
 *                    LIST(LIST(DOUBLE)) child when each row's inner lists are stored back to back
 *   w         [N]    concatenated weight lists (WLS) or NULL
 * The columns are turned into rows on the device and fitted by the same kernels as the aggregates. These are direct
 * anofox_*_fit calls: the aggregate's "fewer than 2 rows" rule does not apply, and where the reference throws
 * ("OLS fit failed: ...", ols_fit.cpp:175-177) status[i] carries the error code instead. */
size_t anofox_b200_fit_columns_workspace_bytes(int64_t n_fits, int64_t n_rows, int32_t n_features);
int32_t anofox_b200_fit_columns_batch_device(const int64_t *offsets, int64_t n_fits, const double *y, const double *x_columns,
                                             const double *w, int64_t n_rows, int32_t n_features,
                                             const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out,
                                             void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err);

/* Window form of the fit-predict aggregates:
 *   {ols,ridge,wls}_fit_predict(y, x [, w], opts) OVER (PARTITION BY g ORDER BY t ROWS BETWEEN <start> AND <end>)
 * (src/window_functions/ols_fit_predict.cpp:111-326, ridge_fit_predict.cpp, wls_fit_predict.cpp). DuckDB runs
 * the aggregate once per output row over that row's frame; here the partitions arrive as a group-sorted CSR in
 * ORDER BY order (offsets[G+1]) and every row gets its answer in one call:
 *   frame of row i = rows [i - frame_preceding, i - frame_end_offset] of its partition;
 *   frame_preceding < 0 = UNBOUNDED PRECEDING (expanding frame, prefix scan of the moment sums);
 *   frame_end_offset    = 0 (CURRENT ROW) or k (k PRECEDING; 1 PRECEDING is the frame of the reference's headline
 *                         benchmark, examples/performance_1m_groups/benchmark_ols.sql:17-19).
 * A row trains when y (and w for WLS) is non-NaN; the answer is the NaN triple (= NULL struct) when the frame is
 * empty, holds at most n_features + intercept training rows (ols_fit_predict.cpp:258-262) or the fit fails
 * (:289-292); otherwise the model is fitted without inference and evaluated, with the simplified interval of
 * anofox_predict_with_interval, at the x of the frame's LAST row (:157-163). pred->is_training is not written.
 * n_features <= 8. null_policy 'drop_y_zero_x' is the caller's business (set y of those rows to NaN).
 * The per-frame fits stay on chip: the workspace is a token 256 bytes (the argument is kept for ABI stability). */
size_t anofox_b200_window_workspace_bytes(int64_t n_rows, int32_t n_features);
int32_t anofox_b200_window_fit_predict_device(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                              const double *w, int64_t n_rows, int32_t n_features,
                                              const AnofoxB200FitOptions *options, int64_t frame_preceding,
                                              int64_t frame_end_offset, const AnofoxB200PredictOutputs *pred,
                                              void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err);
int32_t anofox_b200_window_fit_predict_host(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                            const double *w, int64_t n_rows, int32_t n_features,
                                            const AnofoxB200FitOptions *options, int64_t frame_preceding,
                                            int64_t frame_end_offset, const AnofoxB200PredictOutputs *pred,
                                            AnofoxError *err);

/* Row-sharded single group (config "1 B rows x 64 features over 8 GPUs").
 * Step 0, per rank: anofox_b200_first_valid_row_device finds the first row of this rank's shard that
 *   passes the reference's row filter (ols.rs:59-66, wls.rs:76-86) and writes it to shift_out[p+1]
 *   (x_1..x_p, y; zeros when none) and *found_out (0/1), both device memory. The caller all-gathers
 *   (found, shift) and uses the row of the LOWEST rank that found one: that is the first valid row of
 *   the whole group, the row the reference's constant-column test compares against (ols.rs:76-87).
 * Step 1, per rank: accumulate the shifted moment block of this rank's rows.
 *   shift   [p+1] device f64: common origin (x_1..x_p, y) -- every rank must pass the
 *           same values (rank 0's first valid row, broadcast by the caller)
 *   moments device f64 [anofox_b200_moment_block_len(p)]: additive under SUM. Layout: [0] valid rows,
 *           [1] sum of weights, [2] rows seen, [3] spare, then S1[p+1] = sum w d, then the packed upper
 *           triangle (row-major, i <= j) of S2 = sum w d d^T, d = (x_1..x_p, y) - shift
 *   cdev    device f64 [p]: max |x_j - shift_j| over valid rows: additive under MAX
 * Step 2, caller: ncclAllReduce(moments, SUM), ncclAllReduce(cdev, MAX).
 * Step 3, every rank: solve from the reduced block (one group) into `out` (G = 1). */
size_t anofox_b200_moment_block_len(int32_t n_features);
int32_t anofox_b200_first_valid_row_device(const double *y, const double *x, const double *w, int64_t n_rows,
                                           int32_t n_features, double *shift_out, int32_t *found_out, void *workspace,
                                           size_t workspace_bytes, void *stream, AnofoxError *err);
int32_t anofox_b200_moments_device(const double *y, const double *x, const double *w, int64_t n_rows,
                                   int32_t n_features, const double *shift, double *moments, double *cdev,
                                   void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err);
int32_t anofox_b200_solve_moments_device(const double *moments, const double *cdev, const double *shift,
                                         int32_t n_features, const AnofoxB200FitOptions *options,
                                         const AnofoxB200FitOutputs *out, void *stream, AnofoxError *err);

/* Exact-residual refinement of the row-sharded path. R^2, the residual standard error and everything derived
 * from them come from RSS = TSS - ESS of the moment block; for (nearly) exact fits that difference cancels, and
 * the batched calls then recompute the residuals from the rows (refine_kernel). Across ranks the same takes one
 * more tiny collective, with no host synchronisation anywhere:
 *   a. anofox_b200_solve_moments_refine_device(..., needs_refine, rss_exact = NULL): first solve; sets
 *      *needs_refine (device int32) to 1 when rss <= 1e-5 tss.
 *   b. every rank: anofox_b200_shard_residual_device with the coefficients / intercept just written (device) and
 *      run_if = needs_refine: its rows' sum w (y - b0 - x'b)^2 (evaluated about `shift`) -> rss_partial (device f64; 0 when not flagged,
 *      the kernel then returns at once).
 *   c. ncclAllReduce(rss_partial, SUM).
 *   d. anofox_b200_solve_moments_refine_device(..., needs_refine, rss_exact = reduced value): solves again with
 *      the exact RSS when flagged, does nothing otherwise.
 * A moment block carries no rows, so the QR fallback of the batched calls (ill-conditioned groups) is not
 * available behind these entry points; hc_type must be NONE. */
int32_t anofox_b200_solve_moments_refine_device(const double *moments, const double *cdev, const double *shift,
                                                int32_t n_features, const AnofoxB200FitOptions *options,
                                                const AnofoxB200FitOutputs *out, int32_t *needs_refine,
                                                const double *rss_exact, void *stream, AnofoxError *err);
int32_t anofox_b200_shard_residual_device(const double *y, const double *x, const double *w, int64_t n_rows,
                                          int32_t n_features, const double *coefficients, const double *intercept,
                                          const double *shift /* the common shift row [p+1], may be NULL */,
                                          const int32_t *run_if, double *rss_partial, void *workspace,
                                          size_t workspace_bytes, void *stream, AnofoxError *err);

/* Library / device probe: 0 when a CUDA device of compute capability 10.x is usable. */
int32_t anofox_b200_device_check(AnofoxError *err);
/* How many kernels this library has launched since load (bench.py's gpu_launches). */
int64_t anofox_b200_launch_count(void);
const char *anofox_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ANOFOX_B200_H */
