// groupfit.cu -- kernels' launch side and the C ABI of libanofox_stats_b200.so (include/anofox_b200.h).
//
// Host-side mirror of the reference interface for this path:
//   anofox_{ols,ridge,wls}_fit ........ crates/anofox-stats-ffi/src/lib.rs:98-265, :984-1155, :1384-1555
//   anofox_free_result_* .............. lib.rs:272-311
//   anofox_t_critical / predict ....... lib.rs:2217-2231, :2266-2350
//   anofox_compute_aic / _bic ......... lib.rs:1932-2010
// plus the batch layer the DuckDB Finalize (ols_aggregate.cpp:249-338) should call instead of one
// FFI call per group. There is NO CPU implementation behind these entry points: every fit runs the
// CUDA kernels below, and a missing / unsuitable device is an error.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "acc_launch.h"
#include "solve_mid.cuh"
#include "solve_wide.cuh"
#include "accumulate.cuh"
#include "common.cuh"
#include "fused_launch.h"
#include "solve.cuh"
#include "specfun.cuh"
#include "qr.cuh"
#include "hc.cuh"
#include "hc_launch.h"
#include "resident.cuh"
#include "resident_launch.h"
#include "window.cuh"
#include "window_launch.h"
#include "ingest.cuh"

namespace b200 {

static std::atomic<long long> g_launches{0};
#define LAUNCHED() (g_launches.fetch_add(1, std::memory_order_relaxed))

// Optional per-stage timing (bench.py's roofline leg): CUDA events recorded on the caller's stream
// around the accumulation kernel and around the solve + refine kernels of every batched call.
struct StageTimer {
	std::mutex mu;
	bool enabled = false;
	std::vector<cudaEvent_t> ev; // triples: before accumulate, after accumulate, after solve/refine
	void clear() {
		for (auto e : ev) cudaEventDestroy(e);
		ev.clear();
	}
};
static StageTimer &stage_timer() {
	static StageTimer t;
	return t;
}
static cudaEvent_t timer_mark(cudaStream_t st) {
	StageTimer &t = stage_timer();
	if (!t.enabled) return nullptr;
	std::lock_guard<std::mutex> lock(t.mu);
	cudaEvent_t e = nullptr;
	if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
	cudaEventRecord(e, st);
	t.ev.push_back(e);
	return e;
}

// ---------------------------------------------------------------------------------------------
// Small kernels
// ---------------------------------------------------------------------------------------------
template <int PMAX>
__global__ void __launch_bounds__(128) solve_kernel(const double *__restrict__ records, int64_t n_groups, FitParams P,
                                                    AnofoxB200FitOutputs o, int *refine_count, int64_t *refine_list) {
	const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= n_groups) return;
	solve_group<PMAX>(records + g * rec_len(P.p), 1, g, P, o, -1.0, nullptr, refine_count, refine_list);
}

// Same, for the groups named in a list (the big groups of a batch whose other groups went through the
// fused kernel).
template <int PMAX>
__global__ void __launch_bounds__(128) solve_list_kernel(const double *__restrict__ records, const int64_t *__restrict__ list,
                                                         const int *__restrict__ n_list, FitParams P, AnofoxB200FitOutputs o,
                                                         int *refine_count, int64_t *refine_list) {
	const int n = *n_list; // decided on the device: the host never reads the count
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const int64_t g = list[i];
		solve_group<PMAX>(records + g * rec_len(P.p), 1, g, P, o, -1.0, nullptr, refine_count, refine_list);
	}
}

// Exact residual pass for the groups the solve stage queued (residual sum of squares too small
// relative to the total for the moment form to keep 1e-9): one warp per queued group re-reads the
// group's rows, forms sum w (y - b0 - x'b)^2 with the coefficients just written (NaN = aliased,
// contributes nothing -- lib.rs:2296-2303) and re-runs the statistics with that value.
struct RefineArgs {
	const int64_t *offsets;
	const double *y, *x, *w;
	const uint8_t *train;
	const double *records;
	int64_t rec_gs, rec_es; // record element (g, i) at records[g * rec_gs + i * rec_es]
	const int *refine_count;
	const int64_t *refine_list;
};
template <int PMAX>
__global__ void __launch_bounds__(256) refine_kernel(RefineArgs a, FitParams P, AnofoxB200FitOutputs o) {
	const int lane = threadIdx.x & 31;
	const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
	const int count = *a.refine_count;
	const int p = P.p;
	for (int q = wid; q < count; q += nw) {
		const int64_t g = a.refine_list[q];
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		const double icpt = o.intercept[g];
		const bool ic = !isnan(icpt);
		const double *coef = o.coefficients + g * p;
		// With an intercept the residuals are evaluated about the group's shift row (the record's origin):
		// e = (y - sh_y) - b0' - sum beta (x - sh), b0' = (S1_y - sum beta S1_j) / sum w. Every term then has the
		// size of the data's variation instead of its offset.
		const double *rec = a.records + g * a.rec_gs;
		const double *shf = rec + (int64_t)rec_off_shift(p) * a.rec_es, *s1 = rec + (int64_t)rec_off_s1(p) * a.rec_es;
		double b0 = 0.0, shy = 0.0;
		if (ic) {
			shy = shf[(int64_t)p * a.rec_es];
			double t = s1[(int64_t)p * a.rec_es];
			for (int j = 0; j < p; ++j) {
				const double c = coef[j];
				if (!isnan(c)) t -= c * s1[(int64_t)j * a.rec_es];
			}
			b0 = t / rec[a.rec_es];
		}
		double rss = 0.0;
		for (int64_t r = r0 + lane; r < r1; r += 32) {
			const double yv = a.y[r];
			const double wv = a.w ? a.w[r] : 1.0;
			bool ok = row_ok_yw(yv, wv, a.train ? (int)a.train[r] : 1, a.w != nullptr);
			double fit = b0;
			for (int j = 0; j < p; ++j) {
				const double xv = a.x[r * p + j];
				ok = ok && finite_f64(xv);
				const double c = coef[j];
				if (!isnan(c)) fit = fma(c, ic ? xv - shf[(int64_t)j * a.rec_es] : xv, fit);
			}
			if (ok) {
				const double e = (yv - shy) - fit;
				rss = fma(wv * e, e, rss);
			}
		}
		for (int off = 16; off; off >>= 1) rss += __shfl_xor_sync(0xffffffffu, rss, off);
		if (lane == 0) solve_group<PMAX>(a.records + g * a.rec_gs, a.rec_es, g, P, o, rss, nullptr, nullptr, nullptr);
	}
}

// mode 0: two-sided t p-value (a = t, b = df); 1: t critical (a = level, b = df); 2: F upper tail (a = f, b = d1, c = d2)
// 3 / 4: the table-driven fast paths the solve kernel uses.
__global__ void specfun_eval_kernel(int mode, int n, const double *a, const double *b, const double *c, double *out,
                                    DfTables T) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (mode == 0)
		out[i] = t_two_sided_p(a[i], b[i]);
	else if (mode == 1)
		out[i] = t_critical_dev(a[i], b[i]);
	else if (mode == 2)
		out[i] = f_upper_p(a[i], b[i], c[i]);
	else if (mode == 3)
		out[i] = t_two_sided_p_fast(a[i], b[i], T);
	else
		out[i] = f_upper_p_fast(a[i], b[i], c[i], T);
}

// Default training rule of ols_predict_aggregate.cpp:216-239 when the caller passes no flags:
// y present (NaN stands for NULL) and no missing feature.
__global__ void default_train_kernel(const double *__restrict__ y, const double *__restrict__ x, int64_t n_rows, int p,
                                     uint8_t *__restrict__ flag) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= n_rows) return;
	bool tr = !isnan(y[r]);
	for (int j = 0; tr && j < p; ++j) tr = !isnan(x[r * p + j]);
	flag[r] = tr ? 1 : 0;
}

// The same rule without the features: "y present". Used with the ring kernel for unweighted fits, where a
// training row with a NaN feature poisons its group's sums and sends the group through the checked path
// (redo_kernel, flags_from_y), which applies the feature half of the rule; the is_training OUTPUT gets the full
// rule from the predict kernels, which read every feature anyway. Saves one full pass over X.
__global__ void default_train_y_kernel(const double *__restrict__ y, int64_t n_rows, uint8_t *__restrict__ flag) {
	// eight rows per thread: the flags leave as one 8-byte store when the slice is complete and aligned
	const int64_t r0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
	if (r0 >= n_rows) return;
	if (r0 + 8 <= n_rows && (((uintptr_t)(flag + r0)) & 7) == 0) {
		unsigned long long bits = 0ull;
#pragma unroll
		for (int i = 0; i < 8; ++i) bits |= (unsigned long long)(isnan(y[r0 + i]) ? 0 : 1) << (8 * i);
		*reinterpret_cast<unsigned long long *>(flag + r0) = bits;
	} else {
		for (int64_t r = r0; r < n_rows && r < r0 + 8; ++r) flag[r] = isnan(y[r]) ? 0 : 1;
	}
}

// One warp per group: yhat and the simplified interval of anofox_predict_with_interval
// (lib.rs:2292-2347) for every row of the group, in input order.
// flags_y != null: also write is_training[r] = flags_y[r] && no NaN feature in row r.
// list != null: only the groups list[0 .. *count) (after the resident kernel: the big groups and the groups the QR
// fallback re-fitted).
__global__ void predict_kernel(const int64_t *__restrict__ offsets, int64_t n_groups, const double *__restrict__ x,
                               int p, FitParams P, AnofoxB200FitOutputs fit, AnofoxB200PredictOutputs out,
                               const uint8_t *__restrict__ flags_y, const int64_t *__restrict__ list = nullptr,
                               const int *__restrict__ count = nullptr) {
	int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (list) {
		if (g >= (int64_t)*count) return;
		g = list[g];
	}
	if (g >= n_groups) return;
	const int64_t r0 = offsets[g], r1 = offsets[g + 1];
	const double nanv = nan_f64();
	const bool ok = fit.status[g] == 0;
	double b0 = 0.0, margin = 0.0;
	const double *coef = fit.coefficients + g * p;
	if (ok) {
		const double icpt = fit.intercept[g];
		const bool has_i = !isnan(icpt);
		b0 = has_i ? icpt : 0.0;
		const double rse = fit.residual_std_error[g];
		const int64_t n = fit.n_observations[g];
		if (!(isnan(rse) || rse <= 0.0 || n <= (int64_t)p + 1)) {
			const int64_t kk = (int64_t)p + (has_i ? 1 : 0);
			const int64_t df = n > kk ? n - kk : 0;
			if (df > 0) {
				const double tc = (P.tcrit && df <= kTcritTable) ? P.tcrit[df] : t_critical_dev(P.level, (double)df);
				if (!isnan(tc)) margin = tc * (rse * sqrt(1.0 + 1.0 / (double)n));
			}
		}
	}
	for (int64_t r = r0 + lane; r < r1; r += 32) {
		double yh = nanv, lo = nanv, hi = nanv;
		if (flags_y) {
			bool tr = flags_y[r] != 0;
			for (int j = 0; tr && j < p; ++j) tr = !isnan(x[r * p + j]);
			out.is_training[r] = tr ? 1 : 0;
		}
		if (ok) {
			double s = b0;
			for (int j = 0; j < p; ++j) {
				const double c = coef[j];
				if (!isnan(c)) s += c * x[r * p + j];
			}
			if (finite_f64(s)) { // ols_predict_aggregate.cpp:409-417
				yh = s;
				lo = s - margin;
				hi = s + margin;
			}
		}
		out.yhat[r] = yh;
		out.yhat_lower[r] = lo;
		out.yhat_upper[r] = hi;
	}
}

// Same predictions for p = 2, 4, 8, 16 with fully coalesced reads: P/2 lanes share a row (one 16-byte load
// each, a warp-wide load instruction covers 512 contiguous bytes), the row's dot product meets in a shuffle
// tree and the row's first lane writes the three outputs. x must be 16-byte aligned.
#ifndef ANOFOX_PREDICT_MINB
#define ANOFOX_PREDICT_MINB 3
#endif
#ifndef ANOFOX_PREDICT_U
#define ANOFOX_PREDICT_U 8
#endif
template <int P>
__global__ void __launch_bounds__(256, ANOFOX_PREDICT_MINB) predict_flat_kernel(const int64_t *__restrict__ offsets, int64_t n_groups,
                                                           const double *__restrict__ x, FitParams Pm,
                                                           AnofoxB200FitOutputs fit, AnofoxB200PredictOutputs out,
                                                           const uint8_t *__restrict__ flags_y) {
	constexpr int LPR = P / 2, RPW = 32 / LPR; // lanes per row, rows per warp-wide load
	const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (g >= n_groups) return;
	const int64_t r0 = offsets[g], r1 = offsets[g + 1];
	const double nanv = nan_f64();
	const bool ok = fit.status[g] == 0;
	const int sub = lane % LPR, rw = lane / LPR;
	double b0 = 0.0, margin = 0.0, c0 = 0.0, c1 = 0.0;
	bool n0 = true, n1 = true;
	if (ok) {
		const double icpt = fit.intercept[g];
		const bool has_i = !isnan(icpt);
		b0 = has_i ? icpt : 0.0;
		const double rse = fit.residual_std_error[g];
		const int64_t n = fit.n_observations[g];
		if (!(isnan(rse) || rse <= 0.0 || n <= (int64_t)P + 1)) {
			const int64_t kk = (int64_t)P + (has_i ? 1 : 0);
			const int64_t df = n > kk ? n - kk : 0;
			if (df > 0) {
				const double tc = (Pm.tcrit && df <= kTcritTable) ? Pm.tcrit[df] : t_critical_dev(Pm.level, (double)df);
				if (!isnan(tc)) margin = tc * (rse * sqrt(1.0 + 1.0 / (double)n));
			}
		}
		c0 = fit.coefficients[g * P + 2 * sub];
		c1 = fit.coefficients[g * P + 2 * sub + 1];
		n0 = isnan(c0); // aliased / constant columns contribute nothing (lib.rs:2292-2300)
		n1 = isnan(c1);
	}
	const double2 *xv = reinterpret_cast<const double2 *>(x);
	constexpr int U = ANOFOX_PREDICT_U; // independent loads in flight per lane
	for (int64_t rb = r0; rb < r1; rb += (int64_t)RPW * U) {
		double2 v[U];
		unsigned nanb[U];
		constexpr int NF = (RPW * U + 31) / 32; // flag bytes per lane and pass, loaded with the rows (not after them)
		unsigned fl[NF];
#pragma unroll
		for (int i = 0; i < NF; ++i) {
			const int64_t r = rb + lane + 32 * i;
			fl[i] = (flags_y && lane + 32 * i < RPW * U && r < r1) ? (unsigned)__ldg(flags_y + r) : 0u;
		}
#pragma unroll
		for (int u = 0; u < U; ++u) {
			const int64_t r = rb + u * RPW + rw;
			v[u] = r < r1 ? __ldg(xv + (r * P) / 2 + sub) : make_double2(0.0, 0.0);
			nanb[u] = 0u;
		}
#pragma unroll
		for (int u = 0; u < U; ++u) {
			const int64_t r = rb + u * RPW + rw;
			double s = (n0 ? 0.0 : c0 * v[u].x) + (n1 ? 0.0 : c1 * v[u].y);
#pragma unroll
			for (int m = 1; m < LPR; m <<= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
			if (flags_y) nanb[u] = __ballot_sync(0xffffffffu, isnan(v[u].x) || isnan(v[u].y)); // the lanes of a row vote
			if (sub == 0 && r < r1) {
				double yh = nanv, lo = nanv, hi = nanv;
				s += b0;
				if (ok && finite_f64(s)) { // ols_predict_aggregate.cpp:409-417
					yh = s;
					lo = s - margin;
					hi = s + margin;
				}
				out.yhat[r] = yh;
				out.yhat_lower[r] = lo;
				out.yhat_upper[r] = hi;
			}
		}
		if (flags_y) {
			// is_training = y present and no NaN feature, for the RPW * U consecutive rows of this pass: one
			// coalesced byte per lane instead of scattered single-byte stores from the rows' first lanes
#pragma unroll
			for (int i = 0; i < NF; ++i) {
				const int j = lane + 32 * i;
				const int64_t r = rb + j;
				if (j < RPW * U && r < r1) {
					const unsigned rowmask = (LPR == 32 ? 0xffffffffu : ((1u << LPR) - 1u)) << ((j % RPW) * LPR);
					unsigned nb = nanb[0];
#pragma unroll
					for (int u = 1; u < U; ++u) nb = (j / RPW == u) ? nanb[u] : nb;
					out.is_training[r] = (fl[i] != 0u && (nb & rowmask) == 0u) ? 1 : 0;
				}
			}
		}
	}
}

// Shifted moment block of the row-sharded path -> standard record. The block layout is the
// record without header rewrite: we only need to splice shift and cdev in.
__global__ void assemble_record_kernel(const double *__restrict__ moments, const double *__restrict__ cdev,
                                       const double *__restrict__ shift, int p, double *__restrict__ rec) {
	const int L = rec_len(p);
	for (int i = threadIdx.x; i < L; i += blockDim.x) {
		double v;
		if (i < 4)
			v = moments[i];
		else if (i < rec_off_cdev(p))
			v = shift[i - rec_off_shift(p)];
		else if (i < rec_off_cdev(p) + p)
			v = cdev[i - rec_off_cdev(p)];
		else if (i < rec_off_s1(p))
			v = 0.0;
		else
			v = moments[4 + (i - rec_off_s1(p))];
		rec[i] = v;
	}
}
__global__ void split_record_kernel(const double *__restrict__ rec, int p, double *__restrict__ moments,
                                    double *__restrict__ cdev) {
	const int L = rec_len(p);
	for (int i = threadIdx.x; i < L; i += blockDim.x) {
		const double v = rec[i];
		if (i < 4)
			moments[i] = (i == 3) ? 0.0 : v;
		else if (i < rec_off_cdev(p)) {
		} else if (i < rec_off_cdev(p) + p)
			cdev[i - rec_off_cdev(p)] = v;
		else if (i < rec_off_s1(p)) {
		}
		else
			moments[4 + (i - rec_off_s1(p))] = v;
	}
}


// Residual pass of one row shard (row-sharded single group): sum w (y - b0 - x'b)^2 over the shard's valid rows
// with the coefficients of the first solve (NaN = aliased, contributes nothing: lib.rs:2296-2303). One warp per
// row, lane j holds columns j and j + 32; block partials are added in block order by the second kernel. Runs
// only when *run_if != 0 (the first solve found rss <= 1e-5 tss: the moment form of RSS has cancelled).
__global__ void __launch_bounds__(256) shard_residual_kernel(const double *__restrict__ y, const double *__restrict__ x,
                                                             const double *__restrict__ w, int64_t n_rows, int p,
                                                             const double *__restrict__ coef, const double *__restrict__ icpt,
                                                             const double *__restrict__ shift, const int *__restrict__ run_if,
                                                             double *__restrict__ partials) {
	__shared__ double red[8];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	double acc = 0.0;
	if (*run_if != 0) {
		double b0 = icpt[0];
		const bool ic = !isnan(b0);
		if (!ic) b0 = 0.0;
		double c0 = lane < p ? coef[lane] : 0.0, c1 = lane + 32 < p ? coef[lane + 32] : 0.0;
		if (isnan(c0)) c0 = 0.0;
		if (isnan(c1)) c1 = 0.0;
		// residuals about the common shift row when an intercept is fitted (see refine_kernel): the constant
		// b0 - sh_y + sum beta sh is formed once, its rounding is common to all rows
		const bool sf = ic && shift != nullptr;
		const double sh0 = (sf && lane < p) ? shift[lane] : 0.0, sh1 = (sf && lane + 32 < p) ? shift[lane + 32] : 0.0;
		const double shy = sf ? shift[p] : 0.0;
		if (sf) {
			double t = fma(c0, sh0, c1 * sh1);
#pragma unroll
			for (int off = 16; off; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
			b0 = (b0 - shy) + t;
		}
		const int64_t nw = (int64_t)gridDim.x * 8;
		for (int64_t r = (int64_t)blockIdx.x * 8 + warp; r < n_rows; r += nw) {
			const double x0 = lane < p ? x[r * p + lane] : 0.0, x1 = lane + 32 < p ? x[r * p + lane + 32] : 0.0;
			const double yv = y[r], wv = w ? w[r] : 1.0;
			const bool bad = !finite_f64(x0) || !finite_f64(x1);
			const bool ok = row_ok_yw(yv, wv, 1, w != nullptr) && !__any_sync(0xffffffffu, bad);
			double dot = fma(c0, x0 - sh0, c1 * (x1 - sh1));
#pragma unroll
			for (int off = 16; off; off >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, off);
			if (ok) {
				const double e = (yv - shy) - (b0 + dot);
				acc = fma(wv * e, e, acc);
			}
		}
	}
	if (lane == 0) red[warp] = acc;
	__syncthreads();
	if (threadIdx.x == 0) {
		double t = 0.0;
		for (int i = 0; i < 8; ++i) t += red[i];
		partials[blockIdx.x] = t;
	}
}
__global__ void shard_residual_sum_kernel(const double *__restrict__ partials, int n, double *__restrict__ out) {
	if (threadIdx.x == 0 && blockIdx.x == 0) {
		double t = 0.0;
		for (int i = 0; i < n; ++i) t += partials[i];
		*out = t;
	}
}
// Second solve of the row-sharded group with the all-reduced exact residual sum of squares (only when flagged).
__global__ void __launch_bounds__(kWideThreads) solve_wide_override_kernel(const double *__restrict__ rec, FitParams P,
                                                                           AnofoxB200FitOutputs o, const int *__restrict__ flag,
                                                                           const double *__restrict__ rss) {
	extern __shared__ __align__(16) unsigned char wide_smem_raw2[];
	if (*flag == 0) return;
	WideSmem &S = *reinterpret_cast<WideSmem *>(wide_smem_raw2);
	solve_group_wide(S, rec, 1, 0, P, o, fmax(*rss, 0.0), nullptr, nullptr);
}

// First valid row of one row shard (row-sharded single group, step 0): blocks scan disjoint chunks in
// order and stop as soon as an earlier chunk has reported a hit; `best` ends as the smallest valid row.
__global__ void shard_first_valid_kernel(const double *__restrict__ y, const double *__restrict__ x,
                                         const double *__restrict__ w, int64_t n_rows, int p,
                                         unsigned long long *best) {
	const int64_t chunk = ((n_rows + gridDim.x - 1) / gridDim.x + blockDim.x - 1) / blockDim.x * blockDim.x;
	const int64_t b0 = (int64_t)blockIdx.x * chunk;
	int64_t b1 = b0 + chunk;
	if (b1 > n_rows) b1 = n_rows;
	for (int64_t base = b0; base < b1; base += blockDim.x) {
		// decided by an earlier chunk? (one thread reads, the vote makes the exit uniform)
		const bool stop = threadIdx.x == 0 && *(volatile unsigned long long *)best < (unsigned long long)base;
		if (__syncthreads_or(stop ? 1 : 0)) return;
		const int64_t r = base + threadIdx.x;
		bool ok = false;
		if (r < b1) {
			ok = row_ok_yw(y[r], w ? w[r] : 1.0, 1, w != nullptr);
			for (int c = 0; ok && c < p; ++c) ok = finite_f64(x[r * p + c]);
		}
		if (ok) atomicMin(best, (unsigned long long)r);
		if (__syncthreads_or(ok ? 1 : 0)) return;
	}
}
__global__ void shard_shift_kernel(const double *__restrict__ y, const double *__restrict__ x, int p,
                                   const unsigned long long *best, int64_t n_rows, double *__restrict__ shift,
                                   int32_t *__restrict__ found) {
	const unsigned long long f = *best;
	const bool ok = f < (unsigned long long)n_rows;
	for (int c = threadIdx.x; c <= p; c += blockDim.x) shift[c] = ok ? (c < p ? x[(int64_t)f * p + c] : y[f]) : 0.0;
	if (threadIdx.x == 0) *found = ok ? 1 : 0;
}
// Row bounds of `units` equal slices of a shard (multiples of 16 rows) and the one-slice-per-CTA map the ring
// kernel's slice mode reads.
__global__ void slice_bounds_kernel(int64_t n_rows, int units, int64_t *__restrict__ bounds, int64_t *__restrict__ cta_map) {
	int64_t per = (n_rows + units - 1) / units;
	per = (per + 15) & ~(int64_t)15;
	for (int i = threadIdx.x; i <= units; i += blockDim.x) {
		const int64_t b = (int64_t)i * per;
		bounds[i] = b < n_rows ? b : n_rows;
		cta_map[i] = i;
	}
}

// ---------------------------------------------------------------------------------------------
// Dispatch on the feature count
// ---------------------------------------------------------------------------------------------
struct Shape {
	int nt;
	bool side;
};
static Shape shape_of(int p) {
	const int r = p % 8;
	const bool side = (r == 0 || r == 7);
	const int nt = side ? (p + 7) / 8 : (p + 2 + 7) / 8;
	return {nt, side};
}

static cudaError_t launch_resident(const AccArgs &a, int grid, cudaStream_t st) {
	const Shape s = shape_of(a.p);
	const bool hw = a.w != nullptr, ht = a.train != nullptr;
	cudaError_t e = cudaErrorInvalidValue;
	switch (s.nt) {
	case 1: e = launch_resident_nt1(s.side, hw, ht, a, grid, st); break;
	case 2: e = launch_resident_nt2(s.side, hw, ht, a, grid, st); break;
	case 3: e = launch_resident_nt3(s.side, hw, ht, a, grid, st); break;
	case 4: e = launch_resident_nt4(s.side, hw, ht, a, grid, st); break;
	case 5: e = launch_resident_nt5(s.side, hw, ht, a, grid, st); break;
	case 6: e = launch_resident_nt6(s.side, hw, ht, a, grid, st); break;
	case 7: e = launch_resident_nt7(s.side, hw, ht, a, grid, st); break;
	case 8: e = launch_resident_nt8(s.side, hw, ht, a, grid, st); break;
	case 9: e = launch_resident_nt9(s.side, hw, ht, a, grid, st); break;
	}
	LAUNCHED();
	return e;
}

// Big groups, one CTA per group (device-side list)
static cudaError_t launch_big(const StreamArgs &a, int grid, cudaStream_t st) {
	const Shape s = shape_of(a.p);
	cudaError_t e = cudaErrorInvalidValue;
	switch (s.nt) {
	case 1: e = launch_big_nt1(s.side, a, grid, st); break;
	case 2: e = launch_big_nt2(s.side, a, grid, st); break;
	case 3: e = launch_big_nt3(s.side, a, grid, st); break;
	case 4: e = launch_big_nt4(s.side, a, grid, st); break;
	case 5: e = launch_big_nt5(s.side, a, grid, st); break;
	case 6: e = launch_big_nt6(s.side, a, grid, st); break;
	case 7: e = launch_big_nt7(s.side, a, grid, st); break;
	case 8: e = launch_big_nt8(s.side, a, grid, st); break;
	case 9: e = launch_big_nt9(s.side, a, grid, st); break;
	}
	LAUNCHED();
	return e;
}

// Huge-group streaming: each warp owns an equal slice of every listed group (partials = grid * warps).
static cudaError_t launch_stream(const StreamArgs &a, int grid, cudaStream_t st) {
	const Shape s = shape_of(a.p);
	cudaError_t e = cudaErrorInvalidValue;
	switch (s.nt) {
	case 1: e = launch_stream_nt1(s.side, a, grid, st); break;
	case 2: e = launch_stream_nt2(s.side, a, grid, st); break;
	case 3: e = launch_stream_nt3(s.side, a, grid, st); break;
	case 4: e = launch_stream_nt4(s.side, a, grid, st); break;
	case 5: e = launch_stream_nt5(s.side, a, grid, st); break;
	case 6: e = launch_stream_nt6(s.side, a, grid, st); break;
	case 7: e = launch_stream_nt7(s.side, a, grid, st); break;
	case 8: e = launch_stream_nt8(s.side, a, grid, st); break;
	case 9: e = launch_stream_nt9(s.side, a, grid, st); break;
	}
	LAUNCHED();
	return e;
}

static cudaError_t launch_ring(const RingArgs &a, bool hw, bool ht, int grid, size_t dyn, cudaStream_t st) {
	const Shape s = shape_of(a.p);
	cudaError_t e = cudaErrorInvalidValue;
	switch (s.nt) {
	case 1: e = launch_ring_nt1(s.side, hw, ht, a, grid, dyn, st); break;
	case 2: e = launch_ring_nt2(s.side, hw, ht, a, grid, dyn, st); break;
	case 3: e = launch_ring_nt3(s.side, hw, ht, a, grid, dyn, st); break;
	case 4: e = launch_ring_nt4(s.side, hw, ht, a, grid, dyn, st); break;
	case 5: e = launch_ring_nt5(s.side, hw, ht, a, grid, dyn, st); break;
	case 6: e = launch_ring_nt6(s.side, hw, ht, a, grid, dyn, st); break;
	case 7: e = launch_ring_nt7(s.side, hw, ht, a, grid, dyn, st); break;
	case 8: e = launch_ring_nt8(s.side, hw, ht, a, grid, dyn, st); break;
	case 9: e = launch_ring_nt9(s.side, hw, ht, a, grid, dyn, st); break;
	}
	LAUNCHED();
	return e;
}
static cudaError_t launch_slice(const SliceArgs &a, bool hw, int grid, size_t dyn, cudaStream_t st) {
	const Shape s = shape_of(a.p);
	cudaError_t e = cudaErrorInvalidValue;
	switch (s.nt) {
	case 5: e = launch_slice_nt5(s.side, hw, a, grid, dyn, st); break;
	case 6: e = launch_slice_nt6(s.side, hw, a, grid, dyn, st); break;
	case 7: e = launch_slice_nt7(s.side, hw, a, grid, dyn, st); break;
	case 8: e = launch_slice_nt8(s.side, hw, a, grid, dyn, st); break;
	case 9: e = launch_slice_nt9(s.side, hw, a, grid, dyn, st); break;
	}
	LAUNCHED();
	return e;
}
static cudaError_t launch_redo(const RingArgs &a, int grid, cudaStream_t st) {
	const Shape s = shape_of(a.p);
	cudaError_t e = cudaErrorInvalidValue;
	switch (s.nt) {
	case 1: e = launch_redo_nt1(s.side, a, grid, st); break;
	case 2: e = launch_redo_nt2(s.side, a, grid, st); break;
	case 3: e = launch_redo_nt3(s.side, a, grid, st); break;
	case 4: e = launch_redo_nt4(s.side, a, grid, st); break;
	case 5: e = launch_redo_nt5(s.side, a, grid, st); break;
	case 6: e = launch_redo_nt6(s.side, a, grid, st); break;
	case 7: e = launch_redo_nt7(s.side, a, grid, st); break;
	case 8: e = launch_redo_nt8(s.side, a, grid, st); break;
	case 9: e = launch_redo_nt9(s.side, a, grid, st); break;
	}
	LAUNCHED();
	return e;
}

static cudaError_t launch_solve(const double *records, int64_t G, const FitParams &P, const AnofoxB200FitOutputs &o,
                                int *refine_count, int64_t *refine_list, cudaStream_t st) {
	if (G <= 0) return cudaSuccess;
	if (P.p > 8 && P.p <= 16) {
		// 9 <= p <= 16: fixed-layout, fully unrolled shared-memory solve (solve_mid.cuh)
		const size_t dyn = (size_t)FixLayout<16>::LEN * kFixTS * 8;
		cudaError_t e = cudaFuncSetAttribute(solve_fixed_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
		if (e != cudaSuccess) return e;
		solve_fixed_kernel<16><<<(unsigned)((G + kFixT - 1) / kFixT), kFixT, dyn, st>>>(records, G, P, o, refine_count, refine_list);
		LAUNCHED();
		return cudaGetLastError();
	}
	const int mid_threads = P.p > 8 && P.p <= 32 ? solve_mid_threads(P.p) : 0;
	if (mid_threads > 0) {
		// 9 <= p <= 32: per-thread arrays in shared memory (solve_mid.cuh)
		const size_t dyn = (size_t)rec_len(P.p) * (mid_threads + 1) * 8;
		cudaError_t e = cudaFuncSetAttribute(solve_mid_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
		if (e != cudaSuccess) return e;
		solve_mid_kernel<0><<<(unsigned)((G + mid_threads - 1) / mid_threads), mid_threads, dyn, st>>>(records, G, P, o, refine_count,
		                                                                                              refine_list);
		LAUNCHED();
		return cudaGetLastError();
	}
	if (P.p > 32) {
		// 33 <= p <= 64: one thread block per group, factor in shared memory (solve_wide.cuh)
		const size_t dyn = sizeof(WideSmem);
		cudaError_t e = cudaFuncSetAttribute(solve_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
		if (e != cudaSuccess) return e;
		solve_wide_kernel<<<(unsigned)std::min<int64_t>(G, 148 * 12), kWideThreads, dyn, st>>>(records, nullptr, G, P, o, refine_count,
		                                                                                     refine_list);
		LAUNCHED();
		return cudaGetLastError();
	}
	const int threads = 128;
	const unsigned blocks = (unsigned)((G + threads - 1) / threads);
	if (P.p <= 8)
		solve_kernel<8><<<blocks, threads, 0, st>>>(records, G, P, o, refine_count, refine_list);
	else if (P.p <= 16)
		solve_kernel<16><<<blocks, threads, 0, st>>>(records, G, P, o, refine_count, refine_list);
	else if (P.p <= 32)
		solve_kernel<32><<<blocks, threads, 0, st>>>(records, G, P, o, refine_count, refine_list);
	else
		solve_kernel<64><<<blocks, threads, 0, st>>>(records, G, P, o, refine_count, refine_list);
	LAUNCHED();
	return cudaGetLastError();
}

static cudaError_t launch_solve_list(const double *records, const int64_t *list, const int *n_list, int64_t n_max,
                                     const FitParams &P, const AnofoxB200FitOutputs &o, int *refine_count,
                                     int64_t *refine_list, cudaStream_t st) {
	if (n_max <= 0) return cudaSuccess;
	const int threads = 128;
	const unsigned blocks = (unsigned)std::min<int64_t>((n_max + threads - 1) / threads, 592);
	if (P.p <= 8)
		solve_list_kernel<8><<<blocks, threads, 0, st>>>(records, list, n_list, P, o, refine_count, refine_list);
	else if (P.p <= 16)
		solve_list_kernel<16><<<blocks, threads, 0, st>>>(records, list, n_list, P, o, refine_count, refine_list);
	else if (P.p <= 32)
		solve_list_kernel<32><<<blocks, threads, 0, st>>>(records, list, n_list, P, o, refine_count, refine_list);
	else
		solve_list_kernel<64><<<blocks, threads, 0, st>>>(records, list, n_list, P, o, refine_count, refine_list);
	LAUNCHED();
	return cudaGetLastError();
}

static cudaError_t launch_small(const SmallArgs &a, int mode, bool has_w, bool has_t, int grid, cudaStream_t st) {
	cudaError_t e = cudaErrorInvalidValue;
	switch (a.acc.p) {
	case 1: e = launch_small_p1(mode, has_w, has_t, a, grid, st); break;
	case 2: e = launch_small_p2(mode, has_w, has_t, a, grid, st); break;
	case 3: e = launch_small_p3(mode, has_w, has_t, a, grid, st); break;
	case 4: e = launch_small_p4(mode, has_w, has_t, a, grid, st); break;
	case 5: e = launch_small_p5(mode, has_w, has_t, a, grid, st); break;
	case 6: e = launch_small_p6(mode, has_w, has_t, a, grid, st); break;
	case 7: e = launch_small_p7(mode, has_w, has_t, a, grid, st); break;
	case 8: e = launch_small_p8(mode, has_w, has_t, a, grid, st); break;
	}
	LAUNCHED();
	return e;
}
// The p-values the thread-per-group solve left out (SmallGroupView::defer_p): one thread per (group, statistic), the P
// t-tests and the F test of a group side by side. The same pvalue_f on the same arguments as the in-line form, so the
// results are bit-identical; the point is the instruction stream: the distribution code is more than half of the
// solve kernel's instructions and most of its code size, and it diverges per statistic -- here it runs at full
// occupancy (few registers) instead of inside a 128-register thread that holds a whole fit.
__global__ void __launch_bounds__(256) pvalues_deferred_kernel(int64_t n_groups, int p, int ic, FitParams P, AnofoxB200FitOutputs o) {
	// threads [0, G p): the t-tests in output order; threads [G p, G p + G): the F tests -- a warp never mixes the two
	const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const int64_t nt = n_groups * p;
	if (e >= nt + n_groups) return;
	const DfTables T{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
	if (e < nt) {
		const double dfr = o.p_values[e];
		if (isnan(dfr)) return; // dead column or NULL struct: stays NaN
		const int64_t g = e / p;
		if (o.status[g] != 0) return;
		const double t = o.t_values[e];
		o.p_values[e] = pvalue_f(t * t, 1.0, dfr, T);
	} else {
		const int64_t g = e - nt;
		if (o.status[g] != 0) return;                           // NULL struct: every statistic is NaN already
		if (o.has_inference && o.has_inference[g] == 0) return; // intercept-only fit: no inference (lib.rs:251-253)
		const double dfm = o.f_pvalue[g];
		if (isnan(dfm)) return;
		const double dfr = (double)o.n_observations[g] - dfm - (double)ic;
		o.f_pvalue[g] = pvalue_f(o.f_statistic[g], dfm, dfr, T);
	}
}

static cudaError_t launch_solve_small(const double *records, int64_t G, const FitParams &P, const AnofoxB200FitOutputs &o,
                                      int *refine_count, int64_t *refine_list, cudaStream_t st) {
	if (G <= 0) return cudaSuccess;
	cudaError_t e = cudaErrorInvalidValue;
	// A/B (ANOFOX_B200_DEFER_PVALUES=1): p-values in a kernel of their own. Measured on C2: solve stage 0.433 ms against
	// 0.387 ms with the p-values in line (bit-identical results) -- the distribution code costs its ~0.2 ms either way
	// and overlaps better with the rest of the solve inside one thread. Off by default.
	static const int env_defer = getenv("ANOFOX_B200_DEFER_PVALUES") ? atoi(getenv("ANOFOX_B200_DEFER_PVALUES")) : 0;
	const int defer_p = (P.compute_inference && o.std_errors && o.p_values && o.f_pvalue && o.t_values && env_defer) ? 1 : 0;
	switch (P.p) {
	case 1: e = launch_solve_small_p1(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 2: e = launch_solve_small_p2(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 3: e = launch_solve_small_p3(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 4: e = launch_solve_small_p4(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 5: e = launch_solve_small_p5(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 6: e = launch_solve_small_p6(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 7: e = launch_solve_small_p7(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	case 8: e = launch_solve_small_p8(records, G, P, o, refine_count, refine_list, defer_p, st); break;
	}
	LAUNCHED();
	if (e == cudaSuccess && defer_p) {
		const int64_t n = G * (P.p + 1);
		pvalues_deferred_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(G, P.p, P.fit_intercept ? 1 : 0, P, o);
		LAUNCHED();
		e = cudaGetLastError();
	}
	return e;
}

static cudaError_t launch_refine(const RefineArgs &a, const FitParams &P, const AnofoxB200FitOutputs &o, int sms,
                                 cudaStream_t st) {
	const int blocks = sms * 2;
	if (P.p <= 8)
		refine_kernel<8><<<blocks, 256, 0, st>>>(a, P, o);
	else if (P.p <= 16)
		refine_kernel<16><<<blocks, 256, 0, st>>>(a, P, o);
	else if (P.p <= 32)
		refine_kernel<32><<<blocks, 256, 0, st>>>(a, P, o);
	else
		refine_kernel<64><<<blocks, 256, 0, st>>>(a, P, o);
	LAUNCHED();
	return cudaGetLastError();
}

static cudaError_t launch_qr(QrArgs &a, const FitParams &P, const AnofoxB200FitOutputs &o, int grid, cudaStream_t st) {
	size_t dyn = 0;
	qr_geometry(P.p, a.B, a.LD, dyn);
	if (grid < 1) grid = 1;
	cudaError_t e = cudaFuncSetAttribute(qr_fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
	if (e != cudaSuccess) return e;
	qr_fit_kernel<<<grid, kQrThreads, dyn, st>>>(a, P, o);
	LAUNCHED();
	return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Device properties / stage geometry
// ---------------------------------------------------------------------------------------------
struct DeviceInfo {
	int sms = 0;
	int cc_major = 0;
	size_t smem_optin = 0;
	bool ok = false;
	char why[200] = {0};
};
// One entry per CUDA device, filled on first use from a thread whose current device it is (the SM count sizes
// every grid and workspace, so it must be the calling thread's device, not whichever device came first).
static DeviceInfo &device_info() {
	static std::mutex mu;
	static DeviceInfo infos[64];
	static bool filled[64];
	static DeviceInfo none;
	std::lock_guard<std::mutex> lock(mu);
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) {
		snprintf(none.why, sizeof none.why, "no CUDA device: %s", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
		none.ok = false;
		return none;
	}
	int dev = 0;
	cudaGetDevice(&dev);
	if (dev < 0 || dev >= 64) dev = 0;
	DeviceInfo &info = infos[dev];
	if (filled[dev]) return info;
	filled[dev] = true;
	cudaDeviceProp pr;
	if ((e = cudaGetDeviceProperties(&pr, dev)) != cudaSuccess) {
		snprintf(info.why, sizeof info.why, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
		return info;
	}
	info.sms = pr.multiProcessorCount;
	info.cc_major = pr.major;
	info.smem_optin = pr.sharedMemPerBlockOptin;
	if (pr.major != 10) {
		snprintf(info.why, sizeof info.why, "device %s is sm_%d%d; this library is built for sm_100a only", pr.name,
		         pr.major, pr.minor);
		return info;
	}
	info.ok = true;
	return info;
}

struct StageGeom {
	int cap_rows, xcap, ycap, wcap, tcap, stage_bytes, n_stages;
	int64_t window_rows;
};
static inline int align16(int64_t v) { return (int)((v + 15) & ~(int64_t)15); }

// Choose the stage size from the mean group size: a stage should hold several groups, and the ring
// should keep >= ~96 KB in flight per SM (microbenchmark: 4 x 32 KB stages reach 7.3 TB/s).
static StageGeom stage_geometry(int p, bool has_w, bool has_t, int64_t n_rows, int64_t n_groups) {
	const int64_t row_bytes = 8 * (int64_t)(p + 1) + (has_w ? 8 : 0) + (has_t ? 1 : 0);
	const int64_t budget = 196 * 1024; // dynamic shared memory for the ring
	const double mean_rows = n_groups > 0 ? (double)n_rows / (double)n_groups : 1.0;
	int64_t want = (int64_t)(4.0 * mean_rows * (double)row_bytes);
	int64_t stage = 32 * 1024;
	while (stage < want && stage < budget / 2) stage += 16 * 1024;
	if (stage > budget / 2) stage = budget / 2;
	StageGeom g;
	g.n_stages = (int)std::min<int64_t>(6, budget / stage);
	const int64_t overhead = 4 * 32; // alignment slack of the regions
	g.cap_rows = (int)((stage - overhead) / row_bytes);
	if (g.cap_rows < 4) g.cap_rows = 4;
	g.xcap = align16((int64_t)g.cap_rows * p * 8 + 32);
	g.ycap = align16((int64_t)g.cap_rows * 8 + 32);
	g.wcap = has_w ? align16((int64_t)g.cap_rows * 8 + 32) : 0;
	g.tcap = has_t ? align16((int64_t)g.cap_rows + 32) : 0;
	g.stage_bytes = (g.xcap + g.ycap + g.wcap + g.tcap + 127) & ~127;
	while ((int64_t)g.n_stages * g.stage_bytes > budget + 8 * 1024 && g.n_stages > 2) --g.n_stages;
	g.window_rows = (int64_t)g.cap_rows * 8;
	return g;
}

// Ring path (ring.cuh): window size, ring depth, team size and the shared-memory map.
struct RingGeom {
	int win_rows, n_win, team, ncw;
	uint32_t x_off, y_off, w_off, t_off, slot_off;
	size_t dyn_bytes;
	int64_t big_rows;
};
static RingGeom ring_geometry(int p, bool has_w, bool has_t, int64_t n_rows, int64_t n_groups, int grid, int team_override) {
	const Shape sh = shape_of(p);
	RingGeom g;
	g.ncw = ring_warps(sh.nt);
	const int qr = 4 * ring_sets(sh.nt);
	const int nv = sh.nt * (sh.nt + 1) + (sh.side ? 2 * sh.nt + 3 : 0) + 3;
	const int64_t row_bytes = 8 * (int64_t)(p + 1) + (has_w ? 8 : 0) + (has_t ? 1 : 0);
	static const int env_budget_kb = getenv("ANOFOX_B200_RING_KB") ? atoi(getenv("ANOFOX_B200_RING_KB")) : 0;
	const int64_t budget = (env_budget_kb > 0 ? env_budget_kb : 220) * 1024; // of 227 KB per CTA; measured: 200 -> 222 KB is 2-3 % on C3
	// tuning hooks (benchmarks only): window size in KB, team size
	static const int env_win_kb = getenv("ANOFOX_B200_RING_WIN_KB") ? atoi(getenv("ANOFOX_B200_RING_WIN_KB")) : 0;
	static const int env_team = getenv("ANOFOX_B200_RING_TEAM") ? atoi(getenv("ANOFOX_B200_RING_TEAM")) : 0;
	// measured (tools/sweep.sh): 8 KB windows lose a third of the bandwidth to per-copy overhead; one-tile shapes
	// (small p: y and the flags are tiny copies) want 32 KB, wider rows are best at 16 KB; two-tile shapes (9 <= p <= 15
	// and p = 16), re-swept at the end of round 2 on C3: 24 KB windows, ridge accumulate 5.99 -> 5.81 ms, wls 7.44 -> 7.21 ms
	// (28 / 32 KB are better still for one of the two and 10 % worse for the other)
	const int64_t win_bytes = (env_win_kb > 0 ? env_win_kb : (sh.nt == 1 ? 32 : (sh.nt == 2 ? 24 : 16))) * 1024;
	int sr = (int)((win_bytes / row_bytes + 15) / 16 * 16);
	if (sr < 32) sr = 32;
	const double mean_rows = n_groups > 0 ? (double)n_rows / (double)n_groups : 1.0;
	int tw = 1;
	if (team_override > 0) {
		tw = team_override;
	} else {
		// measured: a warp wants >= ~100 rows of a group (C3: 200 rows -> 2 warps, C5: 1000 rows -> 4 warps)
		while (g.ncw % (tw * 2) == 0 && mean_rows >= 50.0 * (tw * 2) * (tw * 2)) tw *= 2;
		if (env_team > 0 && g.ncw % env_team == 0) tw = env_team;
	}
	for (;;) {
		const int64_t slot_bytes = (int64_t)(g.ncw / tw) * (tw - 1) * nv * 256;
		const int64_t apron = (int64_t)kRingApron * row_bytes + 4 * 16 + 4 * 128;
		int64_t ns = (budget - (team_override > 0 ? 0 : slot_bytes) - apron) / ((int64_t)sr * row_bytes);
		if (ns > kRingMaxWin) ns = kRingMaxWin;
		const bool ok = ns >= 4 && (int64_t)ns * sr >= 2 * (int64_t)tw * qr + 2 * sr && (team_override > 0 || slot_bytes <= 64 * 1024);
		if (ok || (tw == 1 && team_override <= 0)) {
			if (ns < 4) { // very wide rows cannot happen for p <= 64, but never hand the kernel fewer than 4 windows
				sr = 32;
				ns = 4;
			}
			g.n_win = (int)ns;
			g.team = tw;
			break;
		}
		if (team_override > 0) { // slice mode: shrink the windows instead of the team
			sr = std::max(16, sr / 2 / 16 * 16);
			continue;
		}
		tw /= 2;
	}
	g.win_rows = sr;
	const int64_t ring_rows = (int64_t)g.win_rows * g.n_win;
	auto a128 = [](int64_t v) { return (uint32_t)((v + 127) & ~(int64_t)127); };
	uint32_t off = 0;
	g.x_off = off;
	off += a128((ring_rows + kRingApron) * p * 8 + 16);
	g.y_off = off;
	off += a128((ring_rows + kRingApron) * 8 + 16);
	g.w_off = off;
	if (has_w) off += a128((ring_rows + kRingApron) * 8 + 16);
	g.t_off = off;
	if (has_t) off += a128(ring_rows + kRingApron + 16);
	g.slot_off = off;
	if (team_override <= 0) off += a128((int64_t)(g.ncw / g.team) * (g.team - 1) * nv * 256);
	g.dyn_bytes = off;
	int64_t big = n_rows / (2 * (int64_t)std::max(grid, 1));
	if (big < 8 * ring_rows) big = 8 * ring_rows;
	if (big > (int64_t)1 << 30) big = (int64_t)1 << 30;
	g.big_rows = big;
	return g;
}

// Fused small-group path (fused_small.cuh): used when p <= 8 and the mean group is small enough for
// many groups to be resident per SM. T lanes per group keep >= ~8 warps' worth of groups on chip.
// 0 = auto (small groups: accumulate-to-record kernel + register solve kernel), 1 = always the DMMA record
// path, 2 = small groups through the single fused kernel (tests / A-B runs)
static std::atomic<int> g_force_path{0};
struct SmallGeom {
	bool use, stash;
	int T, max_groups;
	StageGeom sg;
};
// which small-group kernels: accumulate-to-record + thread-per-group solve kernel by default; 2 = the redundant-solve
// fused kernel, 6 = the stash kernel (rows -> results in one launch, fused_small.cuh) -- both A/B paths: measured
// slower than the two-kernel path on C2
static bool small_use_stash() { return g_force_path.load() == 6; }
static SmallGeom small_geometry(int p, bool has_w, bool has_t, int64_t n_rows, int64_t n_groups) {
	SmallGeom r;
	memset(&r, 0, sizeof r);
	const int64_t row_bytes = 8 * (int64_t)(p + 1) + (has_w ? 8 : 0) + (has_t ? 1 : 0);
	const double mean_rows = n_groups > 0 ? (double)n_rows / (double)n_groups : 0.0;
	const double mean_bytes = mean_rows * (double)row_bytes;
	r.use = g_force_path.load() != 1 && g_force_path.load() != 3 && p <= kSmallMaxFeatures && n_groups > 0 && mean_bytes <= 24.0 * 1024.0;
	if (!r.use) return r;
	// the stash kernel's ring is what the stashes leave of the shared memory (>= ~110 KB): it takes the batches
	// whose typical group is at most a ninth of that, the accumulate-to-record kernel the rest (where the solve
	// stage is a small share of the work anyway)
	const int64_t stash_ring = (int64_t)206 * 1024 - (int64_t)stash_pool_bytes(p, has_w);
	r.stash = small_use_stash() && mean_bytes * 9.0 <= (double)stash_ring;
	// tuning hooks (tools/sweep_small.sh): rounds of 32 lanes per stage and the shared-memory budget of the ring
	static const int env_rounds = getenv("ANOFOX_B200_SMALL_ROUNDS") ? atoi(getenv("ANOFOX_B200_SMALL_ROUNDS")) : 0;
	static const int env_small_kb = getenv("ANOFOX_B200_SMALL_KB") ? atoi(getenv("ANOFOX_B200_SMALL_KB")) : 0;
	static const int env_stage_kb = getenv("ANOFOX_B200_STASH_STAGE_KB") ? atoi(getenv("ANOFOX_B200_STASH_STAGE_KB")) : 0;
	static const int env_T = getenv("ANOFOX_B200_STASH_T") ? atoi(getenv("ANOFOX_B200_STASH_T")) : 0;
	int64_t budget;
	if (r.stash) {
		// a stage is one warp's unit of work: 32 / T groups, T lanes per group, about 40 KB of rows (a power of two
		// of groups, so that two, four, ... stages fill the 32 slots of a warp's stash exactly)
		const double stage_target = (env_stage_kb > 0 ? env_stage_kb : 40) * 1024.0;
		int mg = 32;
		while (mg > 1 && (double)mg * mean_bytes > stage_target) mg >>= 1;
		int T = 32 / mg;
		if (env_T > 0 && env_T <= 32 && (env_T & (env_T - 1)) == 0 && env_T >= T) T = env_T; // more lanes per group, same stage
		r.T = T;
		r.max_groups = mg;
		// the ring gets what the stashes of the worker warps leave of the CTA's shared memory
		budget = stash_ring;
		if (env_small_kb > 0) budget = std::min<int64_t>(budget, (int64_t)env_small_kb * 1024);
	} else {
		// lanes per group: up to 900 bytes of rows per lane up to p = 5 (C2: 600 B = 12.5 rows per lane; C1: 800 B = 25
		// rows, T = 4 measured 10 % faster in the accumulate kernel than T = 8, which a 700-byte rule chose); from p = 6 on the
		// per-group butterfly over the 28 .. 54 sums weighs more than the rows, so a lane takes up to 16 rows
		// (measured at p = 8: 1 M x 50 rows T = 8 -> 4: 1.51 -> 1.23 ms; 250 k x 200 rows T = 32 -> 16: 2.27 -> 1.22 ms)
		int T = 4;
		if (p >= 6)
			while (T < 32 && mean_rows > 16.0 * T) T *= 2;
		else
			while (T < 32 && mean_bytes > 900.0 * T) T *= 2;
		static const int env_small_T = getenv("ANOFOX_B200_SMALL_T") ? atoi(getenv("ANOFOX_B200_SMALL_T")) : 0; // tuning hook
		if (env_small_T >= 4 && env_small_T <= 32 && (env_small_T & (env_small_T - 1)) == 0) T = env_small_T;
		r.T = T;
		const int gpw = 32 / T;
		r.max_groups = std::min(kSmallStageGroups, (env_rounds > 0 ? env_rounds : 2) * gpw); // two rounds per stage
		budget = (env_small_kb > 0 ? env_small_kb : 196) * 1024;
	}
	int64_t cap = (int64_t)std::ceil(std::max((double)r.max_groups, 4.0) * std::max(mean_rows, 1.0)) + 8;
	const int64_t cap_bytes = r.stash ? std::min<int64_t>(96 * 1024, budget / 3 - 256) : 96 * 1024;
	if (cap * row_bytes > cap_bytes) cap = cap_bytes / row_bytes;
	if (cap * row_bytes < 8 * 1024) cap = 8 * 1024 / row_bytes;
	StageGeom &g = r.sg;
	g.cap_rows = (int)cap;
	g.xcap = align16((int64_t)g.cap_rows * p * 8 + 32);
	g.ycap = align16((int64_t)g.cap_rows * 8 + 32);
	g.wcap = has_w ? align16((int64_t)g.cap_rows * 8 + 32) : 0;
	g.tcap = has_t ? align16((int64_t)g.cap_rows + 32) : 0;
	g.stage_bytes = (g.xcap + g.ycap + g.wcap + g.tcap + 127) & ~127;
	g.n_stages = (int)std::min<int64_t>(kSmallMaxStages, budget / g.stage_bytes);
	if (g.n_stages < 2) {
		// (cannot happen with the 96 KB stage cap and >= 100 KB of ring; keep the kernel's minimum anyway)
		g.n_stages = 2;
	}
	g.window_rows = (int64_t)g.cap_rows * 8;
	return r;
}

// ---- resident fit-predict kernel (resident.cuh): groups between the small-group regime and one stage ----------
static cudaError_t launch_resident_fp(const ResidentArgs &a, int p, bool has_w, int grid, cudaStream_t st) {
	cudaError_t e = cudaErrorInvalidValue;
	switch (p) {
	case 1: e = launch_resident_fp_p1(has_w, a, grid, st); break;
	case 2: e = launch_resident_fp_p2(has_w, a, grid, st); break;
	case 3: e = launch_resident_fp_p3(has_w, a, grid, st); break;
	case 4: e = launch_resident_fp_p4(has_w, a, grid, st); break;
	case 5: e = launch_resident_fp_p5(has_w, a, grid, st); break;
	case 6: e = launch_resident_fp_p6(has_w, a, grid, st); break;
	case 7: e = launch_resident_fp_p7(has_w, a, grid, st); break;
	case 8: e = launch_resident_fp_p8(has_w, a, grid, st); break;
	}
	LAUNCHED();
	return e;
}
struct ResidentGeom {
	bool use;
	int max_groups;
	StageGeom sg;
};
// For the fit-predict entry points when p <= 8 and three typical groups fit the CTA's shared memory (C5: 1 000 rows x
// 8 features = 73 KB per group). Opt-in (force_path 8): measured on C5 the one-pass kernel moves 1.0x instead of 1.7x
// the algorithmic bytes but is issue-bound (about 10 k warp instructions per group, 8 warps per SM at 255 registers)
// and ends at 3.85 ms against 3.62 ms for accumulate -> solve -> predict.
static ResidentGeom resident_geometry(int p, bool has_w, int64_t n_rows, int64_t n_groups, int fp_override = -1) {
	ResidentGeom r;
	memset(&r, 0, sizeof r);
	const int fp = fp_override >= 0 ? fp_override : g_force_path.load();
	if (fp != 8 || p > kResidentMaxFeatures || n_groups <= 0 || n_rows <= 0) return r;
	const int64_t row_bytes = 8 * (int64_t)(p + 1) + (has_w ? 8 : 0) + 1;
	const double mean_rows = (double)n_rows / (double)n_groups;
	const int64_t budget = (int64_t)227 * 1024 - kResidentStaticBytes;
	const int64_t slack = 4 * 32 + 4 * 16; // the regions' lead bytes and their 16-byte rounding
	const int64_t cap_max = (budget / 3 - 128 - slack) / row_bytes;
	if (mean_rows > (double)cap_max) return r;
	int64_t cap = (int64_t)std::ceil(mean_rows * 1.25) + 8;
	cap = std::max<int64_t>(cap, 8 * 1024 / row_bytes);
	cap = std::min(cap, cap_max);
	StageGeom &g = r.sg;
	g.cap_rows = (int)cap;
	g.xcap = align16((int64_t)g.cap_rows * p * 8 + 32);
	g.ycap = align16((int64_t)g.cap_rows * 8 + 32);
	g.wcap = has_w ? align16((int64_t)g.cap_rows * 8 + 32) : 0;
	g.tcap = align16((int64_t)g.cap_rows + 32);
	g.stage_bytes = (g.xcap + g.ycap + g.wcap + g.tcap + 127) & ~127;
	g.n_stages = (int)std::min<int64_t>(kResidentMaxStages, budget / g.stage_bytes);
	if (g.n_stages < 3) return r;
	g.window_rows = (int64_t)g.cap_rows * 8;
	r.max_groups = (int)std::max<int64_t>(1, std::min<int64_t>(8, (int64_t)((double)cap / std::max(mean_rows, 1.0))));
	r.use = true;
	return r;
}

static void set_error(AnofoxError *err, int code, const char *msg) {
	if (!err) return;
	err->code = (AnofoxErrorCode)code;
	strncpy(err->message, msg ? msg : "", 255);
	err->message[255] = 0;
}
static void clear_error(AnofoxError *err) {
	if (!err) return;
	err->code = ANOFOX_ERROR_SUCCESS;
	memset(err->message, 0, sizeof err->message);
}
static int cuda_fail(AnofoxError *err, cudaError_t e, const char *where) {
	char buf[256];
	snprintf(buf, sizeof buf, "CUDA error in %s: %s", where, cudaGetErrorString(e));
	set_error(err, ANOFOX_ERROR_INTERNAL, buf);
	return ANOFOX_ERROR_INTERNAL;
}
#define CU(call, where)                                                                                              \
	do {                                                                                                             \
		cudaError_t e__ = (call);                                                                                    \
		if (e__ != cudaSuccess) return cuda_fail(err, e__, where);                                                   \
	} while (0)

// ---------------------------------------------------------------------------------------------
// Per-(device, confidence level) distribution tables, built once and kept for the process lifetime.
// ---------------------------------------------------------------------------------------------
struct TableEntry {
	int device;
	double level;
	double *buf; // [tcrit | gratio | rcp | ce | co]
};
struct TableCache {
	std::mutex mu;
	std::vector<TableEntry> entries;
};
static TableCache &table_cache() {
	static TableCache c;
	return c;
}
// Returns the tables for `level` on the current device. A new (device, level) pair is filled on a private stream
// and waited for on the host, once per process: afterwards a call adds no dependency to the caller's stream at
// all (so that batched calls can be captured in a CUDA graph).
static int get_tables(double level, cudaStream_t, DfTables &T, AnofoxError *err) {
	TableCache &c = table_cache();
	std::lock_guard<std::mutex> lock(c.mu);
	int dev = 0;
	cudaGetDevice(&dev);
	const size_t n1 = kTcritTable + 1, n3 = kRcpTable + 1, n4 = kSeriesTable;
	auto point = [&](double *buf) {
		T.tcrit = buf;
		T.gratio = buf + n1;
		T.rcp = buf + 2 * n1;
		T.ce = buf + 2 * n1 + n3;
		T.co = buf + 2 * n1 + n3 + n4;
	};
	for (auto &e : c.entries)
		if (e.device == dev && (e.level == level || (std::isnan(e.level) && std::isnan(level)))) {
			point(e.buf);
			return 0;
		}
	// unbounded growth guard: the cache keeps at most 64 levels per process; beyond that the oldest tables stay
	// allocated (a running kernel may still read them) and new levels are evaluated directly (T stays null)
	if (c.entries.size() >= 64) {
		T.tcrit = T.gratio = T.rcp = T.ce = T.co = nullptr;
		return 0;
	}
	TableEntry e;
	e.device = dev;
	e.level = level;
	e.buf = nullptr;
	CU(cudaMalloc(&e.buf, (2 * n1 + n3 + 2 * n4) * sizeof(double)), "cudaMalloc tables");
	cudaStream_t ts = nullptr;
	CU(cudaStreamCreateWithFlags(&ts, cudaStreamNonBlocking), "cudaStreamCreate");
	df_tables_kernel<<<(unsigned)((n3 + 127) / 128), 128, 0, ts>>>(level, e.buf, e.buf + n1, e.buf + 2 * n1,
	                                                                e.buf + 2 * n1 + n3, e.buf + 2 * n1 + n3 + n4);
	LAUNCHED();
	cudaError_t ke = cudaGetLastError();
	if (ke == cudaSuccess) ke = cudaStreamSynchronize(ts);
	cudaStreamDestroy(ts);
	if (ke != cudaSuccess) {
		cudaFree(e.buf);
		return cuda_fail(err, ke, "df_tables_kernel");
	}
	c.entries.push_back(e);
	point(e.buf);
	return 0;
}

// ---------------------------------------------------------------------------------------------
// Workspace carving
// ---------------------------------------------------------------------------------------------
struct Workspace {
	double *records;
	int64_t *tile_g0;
	int64_t *big_list;
	int *big_count;
	double *shift;
	int *first_ok;
	double *partials;
	double *ring_partials; // row-shard mode: partial records of the ring kernel's slices
	int64_t *slice_off;    // row-shard mode: row bounds of the slices, then the CTA -> slice map
	int64_t *refine_list;
	int *refine_count;
	int64_t *redo_list;
	int *redo_count;
	int64_t *huge_list; // indices into big_list of the groups with more than kHugeRows rows
	int *huge_count;
	int64_t *qr_list;   // groups queued for the QR fallback (qr.cuh)
	int *qr_count;
	uint8_t *train_tmp;
	// fit outputs used internally by the predict path when the caller passes none
	double *f_coef, *f_icpt, *f_r2, *f_adj, *f_rse;
	int64_t *f_nobs;
	int32_t *f_status;
	size_t bytes;
	int64_t nbig_max, nhuge_max;
	int n_part;
};

// Big groups above this many rows are sliced over the whole grid (acc_stream_kernel, rounds of kHugeRound groups);
// up to it one CTA streams the group (acc_big_kernel). Both decided on the device.
constexpr int64_t kHugeRows = (int64_t)1 << 20;
constexpr int kHugeRound = 64;

static size_t carve(Workspace &w, unsigned char *base, int64_t G, int64_t N, int p, int sms) {
	size_t off = 0;
	auto take = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return base ? base + o : (unsigned char *)nullptr;
	};
	const int L = rec_len(p);
	// bounds that hold for whichever path / (w, train) combination the call ends up using: the smallest
	// stage capacity gives the most row windows and the most groups that can exceed a stage
	int min_cap = 1 << 30;
	for (int c = 0; c < 4; ++c) {
		min_cap = std::min(min_cap, stage_geometry(p, c & 1, c & 2, N, G).cap_rows);
		const SmallGeom sm = small_geometry(p, c & 1, c & 2, N, G);
		if (sm.use) min_cap = std::min(min_cap, sm.sg.cap_rows);
		const ResidentGeom rg = resident_geometry(p, c & 1, N, G, 8); // whichever way the path is forced later
		if (rg.use) min_cap = std::min(min_cap, rg.sg.cap_rows);
	}
	min_cap = std::max(min_cap, 1);
	const int64_t n_tiles = (N + (int64_t)min_cap * 8 - 1) / ((int64_t)min_cap * 8) + 1;
	w.nbig_max = std::min<int64_t>(G, N / min_cap + 1);
	w.n_part = sms * 2 * kStreamWarps;
	w.records = (double *)take((size_t)std::max<int64_t>(G, 1) * L * 8);
	w.tile_g0 = (int64_t *)take((size_t)(std::max<int64_t>(n_tiles, sms) + 2) * 8);
	w.big_list = (int64_t *)take((size_t)(w.nbig_max + 1) * 8);
	w.big_count = (int *)take(768); // the counters of a call sit together: one memset clears them
	w.refine_count = w.big_count ? w.big_count + 64 : nullptr;
	w.redo_count = w.big_count ? w.big_count + 128 : nullptr;
	w.huge_count = w.big_count ? w.big_count + 144 : nullptr;
	w.qr_count = w.big_count ? w.big_count + 160 : nullptr;
	w.nhuge_max = std::min<int64_t>(w.nbig_max, N / kHugeRows + 1);
	w.huge_list = (int64_t *)take((size_t)(w.nhuge_max + 1) * 8);
	w.qr_list = (int64_t *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.shift = (double *)take((size_t)(w.nbig_max + 1) * (p + 1) * 8);
	w.first_ok = (int *)take((size_t)(w.nbig_max + 1) * 4);
	// partial records of the streaming path: one per (big group, warp). Bounded by capping how many
	// big groups one call may hold (see fit_device); sized here for that cap.
	const int64_t part_groups = std::max<int64_t>(1, std::min<int64_t>(w.nhuge_max, kHugeRound));
	w.partials = (double *)take((size_t)part_groups * w.n_part * L * 8);
	w.ring_partials = (double *)take(G == 1 ? (size_t)sms * 16 * L * 8 : 0);
	w.slice_off = (int64_t *)take((size_t)(2 * sms + 4) * 8);
	w.refine_list = (int64_t *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.redo_list = (int64_t *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.train_tmp = (uint8_t *)take((size_t)std::max<int64_t>(N, 1));
	w.f_coef = (double *)take((size_t)std::max<int64_t>(G, 1) * p * 8);
	w.f_icpt = (double *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.f_r2 = (double *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.f_adj = (double *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.f_rse = (double *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.f_nobs = (int64_t *)take((size_t)std::max<int64_t>(G, 1) * 8);
	w.f_status = (int32_t *)take((size_t)std::max<int64_t>(G, 1) * 4);
	w.bytes = off;
	return off;
}

static int validate_common(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w,
                           int64_t N, int p, const AnofoxB200FitOptions *opt, AnofoxError *err) {
	clear_error(err);
	if (!opt) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "options is NULL");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (G < 0 || N < 0 || p <= 0 || !offsets || (N > 0 && (!y || !x))) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "x is NULL or empty"); // lib.rs:120-125
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (p > kMaxFeatures) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "more than 64 features are not supported by the B200 path");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (opt->method < ANOFOX_B200_OLS || opt->method > ANOFOX_B200_WLS) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "unknown method");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (opt->method == ANOFOX_B200_RIDGE && opt->alpha < 0.0) { // ridge.rs:38-40
		set_error(err, ANOFOX_ERROR_INVALID_ALPHA, "Invalid alpha: must be >= 0");
		return ANOFOX_ERROR_INVALID_ALPHA;
	}
	if (opt->method == ANOFOX_B200_WLS && !w && N > 0) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "weights is NULL");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (opt->hc_type < ANOFOX_HC_NONE || opt->hc_type > ANOFOX_HC_HC3) { // map_options_parser.cpp: "Invalid hc_type"
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "Invalid hc_type");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	return 0;
}

// The seven core arrays are written unconditionally by the solve kernels, the inference block when inference
// is on: a partially filled struct would be a device-side NULL dereference.
static int check_outputs(const AnofoxB200FitOutputs *out, bool inference, AnofoxError *err) {
	if (!out->coefficients || !out->intercept || !out->r_squared || !out->adj_r_squared || !out->residual_std_error ||
	    !out->n_observations || !out->status) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "a required output array is NULL (coefficients .. status)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (inference && (!out->std_errors || !out->t_values || !out->p_values || !out->ci_lower || !out->ci_upper ||
	                  !out->f_statistic || !out->f_pvalue)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "compute_inference needs the inference arrays (std_errors .. f_pvalue)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	return 0;
}

// No C++ exception may cross the C boundary into DuckDB (the reference wraps every fit in catch_unwind and
// reports "Internal panic", lib.rs:144-156): every extern "C" body below runs inside this guard.
template <class F> static int32_t guarded(AnofoxError *err, const char *where, F &&f) noexcept {
	try {
		return f();
	} catch (const std::bad_alloc &) {
		set_error(err, ANOFOX_ERROR_ALLOCATION_FAILURE, "allocation failure");
		return ANOFOX_ERROR_ALLOCATION_FAILURE;
	} catch (...) {
		char buf[160];
		snprintf(buf, sizeof buf, "Internal exception in %s", where);
		set_error(err, ANOFOX_ERROR_INTERNAL, buf);
		return ANOFOX_ERROR_INTERNAL;
	}
}

// The whole device pipeline for one batch. `train` non-null selects predict semantics.
static int fit_device(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w,
                      const uint8_t *train, int64_t N, int p, const AnofoxB200FitOptions *opt,
                      const AnofoxB200FitOutputs &out, const AnofoxB200PredictOutputs *pred, void *workspace,
                      size_t workspace_bytes, cudaStream_t st, AnofoxError *err, int min_state_rows = 2) {
	DeviceInfo &di = device_info();
	Workspace ws;
	const size_t need = carve(ws, (unsigned char *)workspace, G, N, p, di.sms);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned (anofox_b200_workspace_bytes)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (G == 0) return 0;
	const bool wls = opt->method == ANOFOX_B200_WLS;
	const double *wuse = wls ? w : nullptr; // weights only matter for WLS (ols.rs / ridge.rs take none)

	FitParams P;
	P.p = p;
	P.method = opt->method;
	P.fit_intercept = opt->fit_intercept != 0;
	P.compute_inference = pred ? 0 : (opt->compute_inference != 0); // ols_predict_aggregate.cpp:354-364
	P.lambda_scaling = opt->lambda_scaling;
	P.level = opt->confidence_level;
	P.alpha = opt->alpha;
	P.tcrit = P.gratio = P.rcp = P.ce = P.co = nullptr;
	P.predict_mode = pred ? 1 : 0;
	P.min_state_rows = min_state_rows;
	P.qr_count = ws.qr_count;
	P.qr_list = ws.qr_list;

	if (P.compute_inference || pred) {
		DfTables T;
		const int trc = get_tables(P.level, st, T, err);
		if (trc) return trc;
		P.tcrit = T.tcrit;
		P.gratio = T.gratio;
		P.rcp = T.rcp;
		P.ce = T.ce;
		P.co = T.co;
	}

	const uint8_t *train_use = nullptr;
	bool y_only_flags = false;
	ResidentGeom resident;
	memset(&resident, 0, sizeof resident);
	if (pred) {
		const ResidentGeom rg = resident_geometry(p, wuse != nullptr, N, G);
		if (rg.use && (g_force_path.load() == 8 || !small_geometry(p, wuse != nullptr, true, N, G).use)) resident = rg;
	}
	if (pred) {
		if (train) {
			train_use = train;
			if (pred->is_training) CU(cudaMemcpyAsync(pred->is_training, train, (size_t)N, cudaMemcpyDeviceToDevice, st), "copy train");
		} else {
			// unweighted fit through the ring kernel: flags from y alone (default_train_y_kernel), the feature half
			// of the rule is applied by the checked path and by the predict kernels
			const SmallGeom sm0 = small_geometry(p, wuse != nullptr, true, N, G);
			y_only_flags = !wuse && !sm0.use && g_force_path.load() != 3 && pred->is_training != nullptr;
			// the resident fit-predict kernel has every row in shared memory: it takes the y half of the rule as its
			// flags, applies the feature half itself and writes is_training
			if (resident.use) y_only_flags = true;
			uint8_t *flags = (pred->is_training && !y_only_flags) ? pred->is_training : ws.train_tmp;
			if (N > 0) {
				if (y_only_flags)
					default_train_y_kernel<<<(unsigned)((N + 2047) / 2048), 256, 0, st>>>(y, N, flags);
				else
					default_train_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(y, x, N, p, flags);
				LAUNCHED();
				CU(cudaGetLastError(), "default_train_kernel");
			}
			train_use = flags;
		}
	}

	AnofoxB200FitOutputs fo = out;
	if (!fo.coefficients) { // predict path without fit outputs: use workspace arrays
		memset(&fo, 0, sizeof fo);
		fo.coefficients = ws.f_coef;
		fo.intercept = ws.f_icpt;
		fo.r_squared = ws.f_r2;
		fo.adj_r_squared = ws.f_adj;
		fo.residual_std_error = ws.f_rse;
		fo.n_observations = ws.f_nobs;
		fo.status = ws.f_status;
	}

	SmallGeom small = small_geometry(p, wuse != nullptr, train_use != nullptr, N, G);
	if (resident.use) small.use = false;
	// 0 = auto, 1 = never the small-group kernels, 2 = fused small kernel, 3 = whole-group stages (the
	// pre-ring DMMA kernel, kept for A/B runs)
	const bool use_ring = !small.use && !resident.use && g_force_path.load() != 3;
	const StageGeom sg = small.use ? small.sg : (resident.use ? resident.sg : stage_geometry(p, wuse != nullptr, train_use != nullptr, N, G));
	CU(cudaMemsetAsync(ws.big_count, 0, 768, st), "memset"); // big_count | refine_count | redo_count
	// small-group kernels: 2 = stash kernel (rows -> results, one launch), 1 = the redundant-solve fused kernel
	// (A/B), 0 = accumulate to transposed records + thread-per-group solve kernel
	// a handful of groups (the per-group anofox_*_fit calls of DuckDB's Finalize: G = 1): the single-launch fused
	// kernel -- its redundant solve costs nothing at this size and two launches (solve, refine) drop out of a call
	// whose cost is launch latency
	const bool tiny_batch = g_force_path.load() == 0 && G <= 32;
	const int small_mode = !small.use ? 0 : (small.stash ? 2 : ((g_force_path.load() == 2 || tiny_batch) ? 1 : 0));
	const bool fused = small_mode != 0 || resident.use; // the small groups leave the first kernel solved
	const bool transposed = small.use && !fused; // record element (g, i) at records[g + i * G]
	const int64_t rec_gs = transposed ? 1 : rec_len(p), rec_es = transposed ? G : 1;
	int64_t big_threshold = sg.cap_rows;
	if (use_ring) {
		// one contiguous range of groups per CTA, balanced by rows
		const int grid = (int)std::min<int64_t>(di.sms, std::max<int64_t>(1, (N + 4095) / 4096));
		const int64_t window_rows = std::max<int64_t>(1, (N + grid - 1) / grid);
		tile_bounds_kernel<<<1, 256, 0, st>>>(offsets, G, window_rows, grid, ws.tile_g0);
		LAUNCHED();
		CU(cudaGetLastError(), "tile_bounds_kernel");
		const RingGeom rg = ring_geometry(p, wuse != nullptr, train_use != nullptr, N, G, grid, 0);
		RingArgs r;
		r.offsets = offsets;
		r.y = y;
		r.x = x;
		r.w = wuse;
		r.train = train_use;
		r.p = p;
		r.tile_g0 = ws.tile_g0;
		r.records = ws.records;
		r.shift_common = nullptr;
		r.flags_from_y = y_only_flags ? 1 : 0;
		r.big_list = ws.big_list;
		r.big_count = ws.big_count;
		r.big_rows = rg.big_rows;
		r.redo_list = ws.redo_list;
		r.redo_count = ws.redo_count;
		r.win_rows = rg.win_rows;
		r.n_win = rg.n_win;
		r.team = rg.team;
		r.x_off = rg.x_off;
		r.y_off = rg.y_off;
		r.w_off = rg.w_off;
		r.t_off = rg.t_off;
		r.slot_off = rg.slot_off;
		big_threshold = rg.big_rows;
		timer_mark(st);
		CU(launch_ring(r, wuse != nullptr, train_use != nullptr, grid, rg.dyn_bytes, st), "acc_ring_kernel");
		CU(launch_redo(r, di.sms * 2, st), "redo_kernel");
		timer_mark(st);
	} else {
		const int64_t n_tiles = (N + sg.window_rows - 1) / sg.window_rows + 1; // +1: groups that start at row N
		tile_bounds_kernel<<<(unsigned)((n_tiles + 1 + 255) / 256), 256, 0, st>>>(offsets, G, sg.window_rows, n_tiles, ws.tile_g0);
		LAUNCHED();
		CU(cudaGetLastError(), "tile_bounds_kernel");

		AccArgs a;
		a.offsets = offsets;
		a.y = y;
		a.x = x;
		a.w = wuse;
		a.train = train_use;
		a.n_groups = G;
		a.p = p;
		a.tile_g0 = ws.tile_g0;
		a.n_tiles = n_tiles;
		a.records = ws.records;
		a.big_list = ws.big_list;
		a.big_count = ws.big_count;
		a.cap_rows = sg.cap_rows;
		a.stage_bytes = sg.stage_bytes;
		a.xcap = sg.xcap;
		a.ycap = sg.ycap;
		a.wcap = sg.wcap;
		a.n_stages = sg.n_stages;
		const int grid = (int)std::min<int64_t>(di.sms, n_tiles);
		timer_mark(st);
		if (resident.use) {
			// fit-predict, groups of up to one stage: rows -> fit results and predictions in one kernel
			ResidentArgs ra;
			ra.acc = a;
			ra.params = P;
			ra.out = fo;
			ra.pred = *pred;
			ra.write_is_training = (y_only_flags && pred->is_training) ? 1 : 0;
			ra.max_stage_groups = resident.max_groups;
			// experiments: ANOFOX_B200_RESIDENT_DEBUG=1 -> per-phase clock totals in the spare slots of the counter block
			static const int env_rdbg = getenv("ANOFOX_B200_RESIDENT_DEBUG") ? atoi(getenv("ANOFOX_B200_RESIDENT_DEBUG")) : 0;
			ra.dbg = env_rdbg ? (unsigned long long *)(ws.big_count + 64 * 0 + 96 * 0) + 12 : nullptr; // ints 24 .. 55 of the block
			CU(launch_resident_fp(ra, p, wuse != nullptr, grid, st), "fit_predict_resident_kernel");
		} else if (small.use) {
			// fused: rows -> results in one kernel; otherwise rows -> transposed moment records
			SmallArgs sa;
			sa.acc = a;
			sa.params = P;
			sa.out = fo;
			sa.lanes_per_group = small.T;
			sa.max_stage_groups = small.max_groups;
			static const int env_dbg = getenv("ANOFOX_B200_STASH_DEBUG") ? atoi(getenv("ANOFOX_B200_STASH_DEBUG")) : 0;
			sa.debug = env_dbg;
			static const int env_hint = getenv("ANOFOX_B200_WAIT_HINT_NS") ? atoi(getenv("ANOFOX_B200_WAIT_HINT_NS")) : 20000;
			sa.wait_hint = (uint32_t)env_hint;
			sa.refine_count = ws.refine_count;
			sa.refine_list = ws.refine_list;
			CU(launch_small(sa, small_mode, wuse != nullptr, train_use != nullptr, grid, st), "fit_small_kernel");
		} else {
			CU(launch_resident(a, grid, st), "acc_resident_kernel");
		}
		timer_mark(st);
	}

	// Big groups (rare): everything is decided on the device -- the call only enqueues work (no host read of
	// the count). The kernels below look at *big_count themselves and return at once when it is zero.
	if (N > big_threshold) {
		StreamArgs s;
		s.offsets = offsets;
		s.big_list = ws.big_list;
		s.big_count = ws.big_count;
		s.fixed_rows = 0;
		s.y = y;
		s.x = x;
		s.w = wuse;
		s.train = train_use;
		s.p = p;
		s.shift = ws.shift;
		s.first_ok = ws.first_ok;
		s.partials = ws.partials;
		s.n_part = ws.n_part;
		s.records = ws.records;
		s.rec_gs = rec_gs;
		s.rec_es = rec_es;
		s.run_if = nullptr;
		s.ring_partials = nullptr;
		s.ring_n_part = 0;
		s.flags_from_y = y_only_flags ? 1 : 0;
		s.huge_list = ws.huge_list;
		s.huge_count = ws.huge_count;
		s.huge_rows = kHugeRows;
		s.b_begin = 0;
		s.b_cap = kHugeRound;
		const int64_t nbig_max = std::min<int64_t>(G, N / std::max<int64_t>(big_threshold, 1) + 1);
		first_valid_kernel<<<(unsigned)std::min<int64_t>(nbig_max, (int64_t)di.sms * 8), 32, 0, st>>>(s);
		LAUNCHED();
		CU(cudaGetLastError(), "first_valid_kernel");
		CU(launch_big(s, (int)std::min<int64_t>(nbig_max, ws.n_part / kStreamWarps), st), "acc_big_kernel");
		// groups above kHugeRows rows: at most N / kHugeRows of them, sliced over the whole grid in rounds
		const int64_t nhuge_max = std::min<int64_t>(nbig_max, N / kHugeRows);
		for (int64_t b0 = 0; b0 < nhuge_max; b0 += kHugeRound) {
			s.b_begin = (int)b0;
			CU(launch_stream(s, ws.n_part / kStreamWarps, st), "acc_stream_kernel");
			dim3 rg((unsigned)((rec_len(p) + 127) / 128), (unsigned)kHugeRound);
			reduce_partials_kernel<<<rg, 128, 0, st>>>(s);
			LAUNCHED();
			CU(cudaGetLastError(), "reduce_partials_kernel");
		}
	}

	if (fused) {
		const int64_t nbig_max = N > big_threshold ? std::min<int64_t>(G, N / std::max<int64_t>(big_threshold, 1) + 1) : 0;
		CU(launch_solve_list(ws.records, ws.big_list, ws.big_count, nbig_max, P, fo, ws.refine_count, ws.refine_list, st),
		   "solve_list_kernel");
	} else if (transposed) {
		CU(launch_solve_small(ws.records, G, P, fo, ws.refine_count, ws.refine_list, st), "solve_small_kernel");
	} else {
		CU(launch_solve(ws.records, G, P, fo, ws.refine_count, ws.refine_list, st), "solve_kernel");
	}
	if ((small_mode != 1 && !resident.use) || N > big_threshold) {
		RefineArgs ra;
		ra.offsets = offsets;
		ra.y = y;
		ra.x = x;
		ra.w = wuse;
		ra.train = train_use;
		ra.records = ws.records;
		ra.rec_gs = rec_gs;
		ra.rec_es = rec_es;
		ra.refine_count = ws.refine_count;
		ra.refine_list = ws.refine_list;
		CU(launch_refine(ra, P, fo, di.sms, st), "refine_kernel");
	}
	{
		// QR fallback for the groups whose Cholesky met a small pivot (qr.cuh); force_path 5: every group
		QrArgs qa;
		qa.offsets = offsets;
		qa.y = y;
		qa.x = x;
		qa.w = wuse;
		qa.train = train_use;
		qa.flags_from_y = y_only_flags ? 1 : 0;
		const bool all = g_force_path.load() == 5;
		qa.count = all ? nullptr : ws.qr_count;
		qa.list = all ? nullptr : ws.qr_list;
		qa.n_all = G;
		CU(launch_qr(qa, P, fo, (int)std::min<int64_t>(G, (int64_t)di.sms * 4), st), "qr_fit_kernel");
	}
	if (!pred && P.compute_inference && opt->hc_type != ANOFOX_HC_NONE && opt->method != ANOFOX_B200_RIDGE) {
		// robust (sandwich) standard errors: a second pass over the rows with the coefficients just written
		// (ols.rs:209-245, wls.rs:230-266; ridge has no hc_type, ridge.rs)
		HcArgs ha;
		ha.offsets = offsets;
		ha.y = y;
		ha.x = x;
		ha.w = wuse;
		ha.train = train_use;
		ha.flags_from_y = y_only_flags ? 1 : 0;
		ha.n_groups = G;
		ha.hc_type = opt->hc_type;
		ha.list = nullptr;
		ha.count = nullptr;
		bool rest = true; // anything left for the CTA-per-group kernel?
		if (p <= kHcSmallMaxFeatures) {
			// p <= 8: four lanes per group, everything in registers (hc_small.cuh); groups above kHcSmallMaxRows
			// rows are listed for hc_kernel. The QR queue's list is free by now (qr_fit_kernel is enqueued) and is
			// reused; the counter is a slot of its own in the block the call cleared at its start.
			int *hc_count = ws.big_count + 176;
			HcSmallArgs hs;
			hs.offsets = offsets;
			hs.y = y;
			hs.x = x;
			hs.w = wuse;
			hs.train = train_use;
			hs.n_groups = G;
			hs.hc_type = opt->hc_type;
			hs.big_list = ws.qr_list;
			hs.big_count = hc_count;
			const int64_t per_cta = 128 / 4;
			const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((G + per_cta - 1) / per_cta, (int64_t)di.sms * 16));
			CU(launch_hc_small(p, hs, P, fo, grid, st), "hc_small_kernel");
			LAUNCHED();
			ha.list = ws.qr_list;
			ha.count = hc_count;
			rest = N > kHcSmallMaxRows;
		}
		if (rest) {
			const size_t dyn = hc_smem_bytes(p);
			CU(cudaFuncSetAttribute(hc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn), "hc_kernel attr");
			const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)200 * 1024 / (dyn + 1024)));
			const int64_t most = ha.list ? std::min<int64_t>(G, N / kHcSmallMaxRows) : G;
			hc_kernel<<<(unsigned)std::max<int64_t>(1, std::min<int64_t>(most, (int64_t)di.sms * per_sm)), kHcThreads, dyn, st>>>(ha, P, fo);
			LAUNCHED();
			CU(cudaGetLastError(), "hc_kernel");
		}
	}
	timer_mark(st);

	if (pred && N > 0 && resident.use && g_force_path.load() != 5) {
		// the resident kernel has predicted every group it fitted; what is left: the big groups (fitted by the
		// streaming path) and the groups the QR fallback re-fitted after their first prediction
		const int threads = 256;
		const uint8_t *fy = (y_only_flags && pred->is_training) ? train_use : nullptr;
		if (N > big_threshold) {
			const int64_t nbig_max = std::min<int64_t>(G, N / std::max<int64_t>(big_threshold, 1) + 1);
			predict_kernel<<<(unsigned)((nbig_max * 32 + threads - 1) / threads), threads, 0, st>>>(offsets, G, x, p, P, fo, *pred, fy,
			                                                                                      ws.big_list, ws.big_count);
			LAUNCHED();
		}
		predict_kernel<<<(unsigned)((G * 32 + threads - 1) / threads), threads, 0, st>>>(offsets, G, x, p, P, fo, *pred, fy, ws.qr_list,
		                                                                               ws.qr_count);
		LAUNCHED();
		CU(cudaGetLastError(), "predict_kernel (lists)");
	} else if (pred && N > 0) {
		const int threads = 256;
		const unsigned blocks = (unsigned)((G * 32 + threads - 1) / threads);
		const bool flat = (((uintptr_t)x) & 15) == 0;
		const uint8_t *fy = (y_only_flags && pred->is_training) ? train_use : nullptr;
		if (flat && p == 2)
			predict_flat_kernel<2><<<blocks, threads, 0, st>>>(offsets, G, x, P, fo, *pred, fy);
		else if (flat && p == 4)
			predict_flat_kernel<4><<<blocks, threads, 0, st>>>(offsets, G, x, P, fo, *pred, fy);
		else if (flat && p == 8)
			predict_flat_kernel<8><<<blocks, threads, 0, st>>>(offsets, G, x, P, fo, *pred, fy);
		else if (flat && p == 16)
			predict_flat_kernel<16><<<blocks, threads, 0, st>>>(offsets, G, x, P, fo, *pred, fy);
		else
			predict_kernel<<<blocks, threads, 0, st>>>(offsets, G, x, p, P, fo, *pred, fy);
		LAUNCHED();
		CU(cudaGetLastError(), "predict_kernel");
	}
	return 0;
}

// ---------------------------------------------------------------------------------------------
// Host-pointer path: library-owned device buffers
// ---------------------------------------------------------------------------------------------
// Library-owned device state of the host-pointer entry points: three streams, the chunk events and one arena.
// DuckDB finalises from many threads at once (radix_partitioned_hashtable.cpp:990), so there is a small POOL of
// these per device: a caller leases a free context (creating one while fewer than kCtxPerDevice exist, waiting
// otherwise) and concurrent calls overlap their copies and kernels instead of queueing on one mutex.
struct HostContext {
	int device = -1;
	bool busy = false;
	cudaStream_t stream = nullptr;           // compute
	cudaStream_t in_stream = nullptr;        // host -> device copies of the chunks
	cudaStream_t out_stream = nullptr;       // device -> host copies of finished chunks
	static constexpr int kMaxChunks = 16;
	cudaEvent_t in_ev[kMaxChunks], done_ev[kMaxChunks];
	bool events = false;
	unsigned char *dev = nullptr;
	size_t dev_bytes = 0;
	// Small calls (DuckDB's per-group anofox_*_fit, a few KB): inputs are packed into one pinned staging buffer and
	// cross the bus as ONE copy each way instead of ~4 + ~19 driver calls.
	static constexpr size_t kStageHalf = (size_t)256 << 10;
	unsigned char *hstage = nullptr;
	bool stage_ready() {
		if (!hstage && cudaHostAlloc((void **)&hstage, 2 * kStageHalf, cudaHostAllocDefault) != cudaSuccess) {
			hstage = nullptr;
			cudaGetLastError();
		}
		return hstage != nullptr;
	}
	// Medium calls (DuckDB's Finalize hands over up to 2 048 groups: ~0.7 MB of results in 19 arrays): the results
	// come back as ONE copy into this pinned buffer and are scattered by the host, instead of 19 small copies.
	static constexpr size_t kOutStage = (size_t)2 << 20;
	unsigned char *hstage_out = nullptr;
	bool out_stage_ready() {
		if (!hstage_out && cudaHostAlloc((void **)&hstage_out, kOutStage, cudaHostAllocDefault) != cudaSuccess) {
			hstage_out = nullptr;
			cudaGetLastError();
		}
		return hstage_out != nullptr;
	}
	int ensure(size_t bytes, AnofoxError *err) {
		if (!stream) CU(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking), "cudaStreamCreate");
		if (!in_stream) CU(cudaStreamCreateWithFlags(&in_stream, cudaStreamNonBlocking), "cudaStreamCreate");
		if (!out_stream) CU(cudaStreamCreateWithFlags(&out_stream, cudaStreamNonBlocking), "cudaStreamCreate");
		if (!events) {
			for (int i = 0; i < kMaxChunks; ++i) {
				CU(cudaEventCreateWithFlags(&in_ev[i], cudaEventDisableTiming), "cudaEventCreate");
				CU(cudaEventCreateWithFlags(&done_ev[i], cudaEventDisableTiming), "cudaEventCreate");
			}
			events = true;
		}
		if (bytes > dev_bytes) {
			if (dev) cudaFree(dev);
			dev = nullptr;
			dev_bytes = 0;
			size_t want = bytes + bytes / 8;
			cudaError_t e = cudaMalloc(&dev, want);
			if (e != cudaSuccess) {
				set_error(err, ANOFOX_ERROR_ALLOCATION_FAILURE, cudaGetErrorString(e));
				return ANOFOX_ERROR_ALLOCATION_FAILURE;
			}
			dev_bytes = want;
		}
		return 0;
	}
	// every copy and kernel this context has enqueued is complete (error exits: the caller's host buffers must
	// not be written or read after the call returns)
	void drain() {
		if (in_stream) cudaStreamSynchronize(in_stream);
		if (stream) cudaStreamSynchronize(stream);
		if (out_stream) cudaStreamSynchronize(out_stream);
	}
};
constexpr int kCtxPerDevice = 8;
struct HostPool {
	std::mutex mu;
	std::condition_variable cv;
	std::vector<HostContext *> all; // never freed: process lifetime
};
static HostPool &host_pool() {
	static HostPool p;
	return p;
}
struct HostLease {
	HostContext *c = nullptr;
	HostLease() {
		HostPool &p = host_pool();
		int dev = 0;
		cudaGetDevice(&dev);
		std::unique_lock<std::mutex> lock(p.mu);
		for (;;) {
			int have = 0;
			for (HostContext *x : p.all)
				if (x->device == dev) {
					++have;
					if (!x->busy) {
						c = x;
						break;
					}
				}
			if (c) break;
			if (have < kCtxPerDevice) {
				c = new HostContext();
				c->device = dev;
				p.all.push_back(c);
				break;
			}
			p.cv.wait(lock);
		}
		c->busy = true;
	}
	~HostLease() {
		HostPool &p = host_pool();
		{
			std::lock_guard<std::mutex> lock(p.mu);
			c->busy = false;
		}
		p.cv.notify_one();
	}
	HostLease(const HostLease &) = delete;
	HostLease &operator=(const HostLease &) = delete;
};
struct DrainGuard { // drains the context's streams on every exit path of a host-pointer call
	HostContext &c;
	~DrainGuard() { c.drain(); }
};

struct DevLayout { // device mirror of one host batch
	size_t off = 0;
	size_t take(size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	}
};

// offsets of a chunk of groups, relative to the chunk's first row
__global__ void rebase_offsets_kernel(const int64_t *__restrict__ in, int64_t n, int64_t r0, int64_t *__restrict__ out) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = in[i] - r0;
}

// Column-major-per-fit -> row-major rows (the scalar array fits: ols_fit(y LIST, x LIST(LIST)), one fit per SQL row,
// src/table_functions/ols_fit.cpp:120-173). Fit i owns rows [offsets[i], offsets[i+1]); its block of the input holds
// column 0, column 1, ... each of n_i values. One warp per fit; the lanes walk the OUTPUT elements (coalesced
// stores, the loads fall on p short runs of the fit's own block, which the L1 holds), index arithmetic in 32 bits
// whenever the fit has fewer than 2^31 elements.
__global__ void columns_to_rows_kernel(const int64_t *__restrict__ offsets, int64_t n_fits, const double *__restrict__ xc, int p,
                                       double *__restrict__ xr) {
	const int lane = threadIdx.x & 31;
	const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
	for (int64_t f = wid; f < n_fits; f += nw) {
		const int64_t r0 = offsets[f], n = offsets[f + 1] - r0;
		const double *__restrict__ src = xc + r0 * p;
		double *__restrict__ dst = xr + r0 * p;
		const int64_t ne = n * p;
		if (ne < ((int64_t)1 << 31)) {
			const uint32_t un = (uint32_t)n, up = (uint32_t)p, une = (uint32_t)ne;
			for (uint32_t o = (uint32_t)lane; o < une; o += 32u) {
				const uint32_t k = o / up, j = o - k * up;
				dst[o] = src[j * un + k];
			}
		} else {
			for (int64_t o = lane; o < ne; o += 32) {
				const int64_t k = o / p, j = o - k * p;
				dst[o] = src[j * n + k];
			}
		}
	}
}

static int fit_host(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w,
                    const uint8_t *train, int64_t N, int p, const AnofoxB200FitOptions *opt,
                    const AnofoxB200FitOutputs *out, const AnofoxB200PredictOutputs *pred, AnofoxError *err,
                    int min_state_rows = 2) {
	const bool inf = !pred && opt->compute_inference;
	if (out) {
		const int orc = check_outputs(out, inf, err);
		if (orc) return orc;
	}
	// the CSR must be well formed before any pointer arithmetic is done with it
	if (G > 0) {
		bool ok = offsets[0] == 0 && offsets[G] == N;
		for (int64_t g = 0; ok && g < G; ++g) ok = offsets[g] <= offsets[g + 1];
		if (!ok) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "offsets must start at 0, be non-decreasing and end at n_rows");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
	}
	HostLease lease;
	HostContext &c = *lease.c;
	DrainGuard drain{c};
	DeviceInfo &di = device_info();
	// Chunks of whole groups, balanced by rows: the copy of chunk i + 1 to the device, the kernels of chunk i and
	// the copy of chunk i - 1's results back to the host run on three streams (PCIe is full duplex and the
	// kernels are ~2 % of a host-to-host call, so the call approaches the one-way transfer time of its inputs).
	int nch = 1;
	const size_t in_bytes = (size_t)N * ((size_t)p + 1 + (w ? 1 : 0)) * 8;
	if (G >= 1024 && in_bytes >= ((size_t)64 << 20)) nch = (int)std::min<size_t>(8, in_bytes >> 25);
	int64_t gb[HostContext::kMaxChunks + 1];
	gb[0] = 0;
	for (int i = 1; i < nch; ++i) {
		const int64_t target = (N / nch) * i;
		const int64_t g = std::lower_bound(offsets, offsets + G + 1, target) - offsets;
		gb[i] = std::min<int64_t>(G, std::max<int64_t>(g, gb[i - 1]));
	}
	gb[nch] = G;
	Workspace wsz;
	size_t ws_bytes = carve(wsz, nullptr, G, N, p, di.sms);
	if (nch > 1)
		for (int i = 0; i < nch; ++i)
			ws_bytes = std::max(ws_bytes, carve(wsz, nullptr, gb[i + 1] - gb[i], offsets[gb[i + 1]] - offsets[gb[i]], p, di.sms));
	DevLayout L;
	const int64_t G1 = std::max<int64_t>(G, 1), N1 = std::max<int64_t>(N, 1);
	const size_t o_off = L.take((size_t)(G + 1) * 8), o_y = L.take((size_t)N1 * 8), o_x = L.take((size_t)N1 * p * 8);
	const size_t o_off2 = L.take((size_t)(G + 1 + nch) * 8);
	const size_t o_w = w ? L.take((size_t)N1 * 8) : 0, o_t = train ? L.take((size_t)N1) : 0;
	const size_t o_gp = L.take((size_t)G1 * p * 8 * (inf ? 6 : 1));
	const size_t o_g = L.take((size_t)G1 * 8 * 12);
	const size_t o_pred = pred ? L.take((size_t)N1 * 8 * 3 + (size_t)N1) : 0;
	const size_t o_ws = L.take(ws_bytes);
	int rc = c.ensure(L.off, err);
	if (rc) return rc;
	unsigned char *d = c.dev;
	cudaStream_t st = c.stream, sin = nch > 1 ? c.in_stream : c.stream, sout = nch > 1 ? c.out_stream : c.stream;
	// ---- small call: one packed copy each way through the context's pinned staging buffer ----
	const size_t in_span = o_gp - o_off, out_span = o_g + (size_t)G1 * 8 * 12 - o_gp;
	const bool small = nch == 1 && !pred && in_span <= HostContext::kStageHalf && out_span <= HostContext::kStageHalf &&
	                   c.stage_ready();
	const bool pack_out = !small && nch == 1 && !pred && out && out_span <= HostContext::kOutStage && c.out_stage_ready();
	struct Pending {
		void *dst;
		size_t off, bytes;
	};
	std::vector<Pending> pending;
	if (small) {
		unsigned char *hs = c.hstage;
		memcpy(hs, offsets, (size_t)(G + 1) * 8);
		if (N > 0) {
			memcpy(hs + (o_y - o_off), y, (size_t)N * 8);
			memcpy(hs + (o_x - o_off), x, (size_t)N * p * 8);
			if (w) memcpy(hs + (o_w - o_off), w, (size_t)N * 8);
			if (train) memcpy(hs + (o_t - o_off), train, (size_t)N);
		}
		CU(cudaMemcpyAsync(d + o_off, hs, in_span, cudaMemcpyHostToDevice, sin), "H2D packed");
	} else {
		// ---- host -> device: offsets first, then the chunks in order ----
		CU(cudaMemcpyAsync(d + o_off, offsets, (size_t)(G + 1) * 8, cudaMemcpyHostToDevice, sin), "H2D offsets");
	}
	auto push = [&](int i) -> int { // chunk i's rows to the device
		if (small) return 0;
		const int64_t r0 = offsets[gb[i]], nr = offsets[gb[i + 1]] - r0;
		if (nr > 0) {
			CU(cudaMemcpyAsync(d + o_y + (size_t)r0 * 8, y + r0, (size_t)nr * 8, cudaMemcpyHostToDevice, sin), "H2D y");
			CU(cudaMemcpyAsync(d + o_x + (size_t)r0 * p * 8, x + r0 * p, (size_t)nr * p * 8, cudaMemcpyHostToDevice, sin), "H2D x");
			if (w) CU(cudaMemcpyAsync(d + o_w + (size_t)r0 * 8, w + r0, (size_t)nr * 8, cudaMemcpyHostToDevice, sin), "H2D w");
			if (train) CU(cudaMemcpyAsync(d + o_t + (size_t)r0, train + r0, (size_t)nr, cudaMemcpyHostToDevice, sin), "H2D train");
		}
		if (nch > 1) CU(cudaEventRecord(c.in_ev[i], sin), "event");
		return 0;
	};
	// two chunks ahead of the kernels (pinned host memory: the calls return at once; pageable memory: the staged
	// copy of chunk i + 2 overlaps the kernels of chunk i)
	for (int i = 0; i < std::min(nch, 2); ++i)
		if ((rc = push(i))) return rc;
	double *gp = (double *)(d + o_gp), *gs = (double *)(d + o_g);
	auto back = [&](void *dst, const void *src, size_t bytes) {
		if (small || pack_out) { // scattered from the staging buffer after the one packed copy below
			if (dst && bytes) pending.push_back({dst, (size_t)((const unsigned char *)src - (d + o_gp)), bytes});
			return cudaSuccess;
		}
		return (dst && bytes) ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, sout) : cudaSuccess;
	};
	for (int i = 0; i < nch; ++i) {
		const int64_t g0 = gb[i], Gc = gb[i + 1] - g0;
		const int64_t r0 = offsets[g0], Nc = offsets[gb[i + 1]] - r0;
		if (Gc <= 0) {
			if (i + 2 < nch && (rc = push(i + 2))) return rc;
			continue;
		}
		// device views of this chunk: every per-group array starts at group g0, every per-row array at row r0
		AnofoxB200FitOutputs dout;
		memset(&dout, 0, sizeof dout);
		dout.coefficients = gp + (size_t)g0 * p;
		if (inf) {
			dout.std_errors = gp + (size_t)G1 * p * 1 + (size_t)g0 * p;
			dout.t_values = gp + (size_t)G1 * p * 2 + (size_t)g0 * p;
			dout.p_values = gp + (size_t)G1 * p * 3 + (size_t)g0 * p;
			dout.ci_lower = gp + (size_t)G1 * p * 4 + (size_t)g0 * p;
			dout.ci_upper = gp + (size_t)G1 * p * 5 + (size_t)g0 * p;
		}
		dout.intercept = gs + G1 * 0 + g0;
		dout.r_squared = gs + G1 * 1 + g0;
		dout.adj_r_squared = gs + G1 * 2 + g0;
		dout.residual_std_error = gs + G1 * 3 + g0;
		dout.n_observations = (int64_t *)(gs + G1 * 4) + g0;
		dout.status = (int32_t *)(gs + G1 * 5) + g0;
		dout.f_statistic = gs + G1 * 6 + g0;
		dout.f_pvalue = gs + G1 * 7 + g0;
		dout.has_inference = (uint8_t *)(gs + G1 * 8) + g0;
		dout.aic = gs + G1 * 9 + g0;
		dout.bic = gs + G1 * 10 + g0;
		dout.rss = gs + G1 * 11 + g0;
		AnofoxB200PredictOutputs dpred;
		memset(&dpred, 0, sizeof dpred);
		if (pred) {
			double *pp = (double *)(d + o_pred);
			dpred.yhat = pp + r0;
			dpred.yhat_lower = pp + N1 + r0;
			dpred.yhat_upper = pp + 2 * N1 + r0;
			dpred.is_training = (uint8_t *)(pp + 3 * N1) + r0;
		}
		const int64_t *doff = (const int64_t *)(d + o_off) + g0;
		if (nch > 1) {
			CU(cudaStreamWaitEvent(st, c.in_ev[i], 0), "wait");
			int64_t *reb = (int64_t *)(d + o_off2) + g0 + i;
			rebase_offsets_kernel<<<(unsigned)((Gc + 1 + 255) / 256), 256, 0, st>>>(doff, Gc + 1, r0, reb);
			LAUNCHED();
			CU(cudaGetLastError(), "rebase_offsets_kernel");
			doff = reb;
		}
		rc = fit_device(doff, Gc, (const double *)(d + o_y) + r0, (const double *)(d + o_x) + r0 * p,
		                w ? (const double *)(d + o_w) + r0 : nullptr, train ? (const uint8_t *)(d + o_t) + r0 : nullptr, Nc, p,
		                opt, dout, pred ? &dpred : nullptr, d + o_ws, ws_bytes, st, err, min_state_rows);
		if (rc) return rc; // DrainGuard waits for the copies already in flight
		if (i + 2 < nch && (rc = push(i + 2))) return rc;
		if (nch > 1) {
			CU(cudaEventRecord(c.done_ev[i], st), "event");
			CU(cudaStreamWaitEvent(sout, c.done_ev[i], 0), "wait");
		}
		// ---- device -> host: this chunk's slice of every output array ----
		if (out) {
			const size_t gpn = (size_t)Gc * p * 8, gn = (size_t)Gc * 8;
			auto at = [&](double *h, size_t k) { return h ? h + (size_t)g0 * k : nullptr; };
			CU(back(at(out->coefficients, p), dout.coefficients, gpn), "D2H");
			CU(back(at(out->intercept, 1), dout.intercept, gn), "D2H");
			CU(back(at(out->r_squared, 1), dout.r_squared, gn), "D2H");
			CU(back(at(out->adj_r_squared, 1), dout.adj_r_squared, gn), "D2H");
			CU(back(at(out->residual_std_error, 1), dout.residual_std_error, gn), "D2H");
			CU(back(out->n_observations ? out->n_observations + g0 : nullptr, dout.n_observations, gn), "D2H");
			CU(back(out->status ? out->status + g0 : nullptr, dout.status, (size_t)Gc * 4), "D2H");
			if (inf) {
				CU(back(at(out->std_errors, p), dout.std_errors, gpn), "D2H");
				CU(back(at(out->t_values, p), dout.t_values, gpn), "D2H");
				CU(back(at(out->p_values, p), dout.p_values, gpn), "D2H");
				CU(back(at(out->ci_lower, p), dout.ci_lower, gpn), "D2H");
				CU(back(at(out->ci_upper, p), dout.ci_upper, gpn), "D2H");
				CU(back(at(out->f_statistic, 1), dout.f_statistic, gn), "D2H");
				CU(back(at(out->f_pvalue, 1), dout.f_pvalue, gn), "D2H");
				CU(back(out->has_inference ? out->has_inference + g0 : nullptr, dout.has_inference, (size_t)Gc), "D2H");
			}
			CU(back(at(out->aic, 1), dout.aic, gn), "D2H");
			CU(back(at(out->bic, 1), dout.bic, gn), "D2H");
			CU(back(at(out->rss, 1), dout.rss, gn), "D2H");
		}
		if (pred && Nc > 0) {
			CU(back(pred->yhat + r0, dpred.yhat, (size_t)Nc * 8), "D2H");
			CU(back(pred->yhat_lower + r0, dpred.yhat_lower, (size_t)Nc * 8), "D2H");
			CU(back(pred->yhat_upper + r0, dpred.yhat_upper, (size_t)Nc * 8), "D2H");
			CU(back(pred->is_training + r0, dpred.is_training, (size_t)Nc), "D2H");
		}
	}
	unsigned char *hback = small ? c.hstage + HostContext::kStageHalf : c.hstage_out;
	if (small || pack_out) CU(cudaMemcpyAsync(hback, d + o_gp, out_span, cudaMemcpyDeviceToHost, sout), "D2H packed");
	CU(cudaStreamSynchronize(sin), "sync");
	CU(cudaStreamSynchronize(st), "sync");
	if (sout != st) CU(cudaStreamSynchronize(sout), "sync");
	for (const Pending &q : pending) memcpy(q.dst, hback + q.off, q.bytes);
	return 0;
}

// ---------------------------------------------------------------------------------------------
// One fit through the reference's own signature (drop-in layer)
// ---------------------------------------------------------------------------------------------
// Error texts = the Display strings of the reference's StatsError (crates/anofox-stats-core/src/errors.rs:7-37), which
// anofox_*_fit copies into AnofoxError.message (lib.rs:258-262) and the scalar SQL functions surface as
// "OLS fit failed: <message>" (src/table_functions/ols_fit.cpp:175-177; test/sql/error_handling.test matches on them).
static std::string rust_f64(double v) { // Rust's Display for f64: shortest digits that round-trip, no exponent
	char buf[64];
	for (int prec = 1; prec <= 17; ++prec) {
		snprintf(buf, sizeof buf, "%.*g", prec, v);
		if (strtod(buf, nullptr) == v) break;
	}
	std::string t(buf);
	if (t.find('e') != std::string::npos) {
		snprintf(buf, sizeof buf, "%.17f", v);
		t = buf;
		while (!t.empty() && t.back() == '0') t.pop_back();
		if (!t.empty() && t.back() == '.') t.pop_back();
	}
	return t;
}
static std::string msg_dimension(size_t y_len, size_t x_rows) {
	return "Dimension mismatch: y has " + std::to_string(y_len) + " elements, X has " + std::to_string(x_rows) + " rows";
}

static bool single_fit(int method, AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count,
                       const AnofoxDataArray *weights, bool fit_intercept, bool compute_inference, double level,
                       double alpha, int lambda_scaling, int solver, int hc_type, AnofoxFitResultCore *out_core,
                       AnofoxFitResultInference *out_inf, AnofoxError *err) {
	clear_error(err); // lib.rs:108-110
	if (!out_core) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "out_core is NULL");
		return false;
	}
	if (!x || x_count == 0) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "x is NULL or empty");
		return false;
	}
	if (method == ANOFOX_B200_RIDGE && alpha < 0.0) { // ridge.rs:38-40 comes first
		set_error(err, ANOFOX_ERROR_INVALID_ALPHA, ("Invalid alpha parameter: " + rust_f64(alpha) + " (must be >= 0)").c_str());
		return false;
	}
	if (y.len == 0 || !y.data) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "Empty input: y cannot be empty"); // ols.rs:38-40
		return false;
	}
	if (weights && weights->len == 0) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "Empty input: weights cannot be empty"); // wls.rs:50-52
		return false;
	}
	if (weights && weights->len != y.len) { // wls.rs:58-63
		set_error(err, ANOFOX_ERROR_DIMENSION_MISMATCH, msg_dimension(y.len, weights->len).c_str());
		return false;
	}
	for (size_t j = 0; j < x_count; ++j)
		if (x[j].len != y.len) { // ols.rs:49-56
			set_error(err, ANOFOX_ERROR_DIMENSION_MISMATCH, msg_dimension(y.len, x[j].len).c_str());
			return false;
		}
	const size_t n = y.len, p = x_count;
	auto valid = [](const uint8_t *v, size_t i) { return v == nullptr || ((v[i >> 3] >> (i & 7)) & 1); };
	// column arrays -> one group of row-major rows; NULL -> NaN (types.rs:79-89)
	std::vector<double> xr(n * p), yv(n), wv;
	for (size_t i = 0; i < n; ++i) yv[i] = valid(y.validity, i) ? y.data[i] : NAN;
	for (size_t j = 0; j < p; ++j)
		for (size_t i = 0; i < n; ++i) xr[i * p + j] = valid(x[j].validity, i) ? x[j].data[i] : NAN;
	if (weights) {
		wv.resize(n);
		for (size_t i = 0; i < n; ++i) wv[i] = valid(weights->validity, i) ? weights->data[i] : NAN;
	}
	AnofoxB200FitOptions o;
	o.method = method;
	o.fit_intercept = fit_intercept;
	o.compute_inference = compute_inference;
	o.solver = solver;
	o.lambda_scaling = lambda_scaling;
	o.hc_type = hc_type;
	o.confidence_level = level;
	o.alpha = alpha;
	int64_t offsets[2] = {0, (int64_t)n};
	// The aggregate's "fewer than 2 rows" rule belongs to the caller (ols_aggregate.cpp:263-267), not
	// to anofox_*_fit: a one-row fit must reach the core rules, hence min_state_rows = 0 below.
	std::vector<double> coef(p), se(p), tv(p), pv(p), lo(p), hi(p);
	double icpt, r2, adj, rse, fs, fp;
	int64_t nobs;
	int32_t status;
	uint8_t has_inf = 0;
	AnofoxB200FitOutputs out;
	memset(&out, 0, sizeof out);
	out.coefficients = coef.data();
	out.intercept = &icpt;
	out.r_squared = &r2;
	out.adj_r_squared = &adj;
	out.residual_std_error = &rse;
	out.n_observations = &nobs;
	out.status = &status;
	out.std_errors = se.data();
	out.t_values = tv.data();
	out.p_values = pv.data();
	out.ci_lower = lo.data();
	out.ci_upper = hi.data();
	out.f_statistic = &fs;
	out.f_pvalue = &fp;
	out.has_inference = &has_inf;
	int rc = validate_common(offsets, 1, yv.data(), xr.data(), weights ? wv.data() : nullptr, (int64_t)n, (int)p, &o, err);
	if (rc) return false;
	rc = fit_host(offsets, 1, yv.data(), xr.data(), weights ? wv.data() : nullptr, nullptr, (int64_t)n, (int)p, &o, &out,
	              nullptr, err, 0);
	if (rc) return false;
	if (status != 0) {
		std::string msg = "fit failed";
		if (status == ANOFOX_ERROR_NO_VALID_DATA) {
			msg = "All rows filtered due to NULL/NaN values"; // errors.rs:16
		} else if (status == ANOFOX_ERROR_INSUFFICIENT_DATA) {
			// errors.rs:13 names the rows that survived the filter: count them for the text (the decision itself
			// was made on the device)
			size_t rows = 0;
			for (size_t i = 0; i < n; ++i) {
				bool ok = std::isfinite(yv[i]);
				if (ok && weights) ok = std::isfinite(wv[i]) && wv[i] > 0.0;
				for (size_t j = 0; ok && j < p; ++j) ok = std::isfinite(xr[i * p + j]);
				rows += ok ? 1 : 0;
			}
			msg = "Insufficient data: " + std::to_string(rows) + " rows, " + std::to_string(p) + " features (need rows > features)";
		}
		set_error(err, status, msg.c_str());
		return false;
	}
	double *c = (double *)malloc(std::max<size_t>(p, 1) * sizeof(double));
	if (!c) {
		set_error(err, ANOFOX_ERROR_ALLOCATION_FAILURE, "Failed to allocate coefficients");
		return false;
	}
	memcpy(c, coef.data(), p * sizeof(double));
	out_core->coefficients = c;
	out_core->coefficients_len = p;
	out_core->intercept = icpt;
	out_core->r_squared = r2;
	out_core->adj_r_squared = adj;
	out_core->residual_std_error = rse;
	out_core->n_observations = (size_t)nobs;
	out_core->n_features = p;
	if (out_inf) {
		if (compute_inference && has_inf) {
			double *arrs[5];
			const std::vector<double> *src[5] = {&se, &tv, &pv, &lo, &hi};
			bool okm = true;
			for (int i = 0; i < 5; ++i) {
				arrs[i] = (double *)malloc(std::max<size_t>(p, 1) * sizeof(double));
				okm = okm && arrs[i];
			}
			if (!okm) {
				for (int i = 0; i < 5; ++i) free(arrs[i]);
				free(c);
				out_core->coefficients = nullptr;
				set_error(err, ANOFOX_ERROR_ALLOCATION_FAILURE, "Failed to allocate inference arrays");
				return false;
			}
			for (int i = 0; i < 5; ++i) memcpy(arrs[i], src[i]->data(), p * sizeof(double));
			out_inf->std_errors = arrs[0];
			out_inf->t_values = arrs[1];
			out_inf->p_values = arrs[2];
			out_inf->ci_lower = arrs[3];
			out_inf->ci_upper = arrs[4];
			out_inf->len = p;
			out_inf->confidence_level = level;
			out_inf->f_statistic = fs;
			out_inf->f_pvalue = fp;
		} else { // lib.rs:251-253
			out_inf->std_errors = out_inf->t_values = out_inf->p_values = out_inf->ci_lower = out_inf->ci_upper = nullptr;
			out_inf->len = 0;
			out_inf->confidence_level = 0.95;
			out_inf->f_statistic = NAN;
			out_inf->f_pvalue = NAN;
		}
	}
	return true;
}

// ---------------------------------------------------------------------------------------------
// Scalar array fits, batched over the SQL rows of a DataChunk: {ols,ridge,wls}_fit(y LIST, x LIST(LIST) [, w LIST])
// (src/table_functions/{ols,ridge,wls}_fit.cpp:120-230: one anofox_*_fit call per row there)
// ---------------------------------------------------------------------------------------------
static int fit_columns_device(const int64_t *offsets, int64_t F, const double *y, const double *x_cm, const double *w,
                              int64_t N, int p, const AnofoxB200FitOptions *opt, const AnofoxB200FitOutputs &out,
                              void *workspace, size_t workspace_bytes, cudaStream_t st, AnofoxError *err) {
	DeviceInfo &di = device_info();
	const size_t xbytes = (((size_t)std::max<int64_t>(N, 1) * p * 8) + 255) & ~(size_t)255;
	Workspace wsz;
	const size_t need = xbytes + carve(wsz, nullptr, F, N, p, di.sms);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned (anofox_b200_fit_columns_workspace_bytes)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (F == 0) return 0;
	double *xr = (double *)workspace;
	if (N > 0) {
		const int64_t warps = std::min<int64_t>(F, (int64_t)di.sms * 32);
		columns_to_rows_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(offsets, F, x_cm, p, xr);
		LAUNCHED();
		CU(cudaGetLastError(), "columns_to_rows_kernel");
	}
	// min_state_rows = 0: these are direct anofox_*_fit calls, the aggregate's "fewer than 2 rows" rule does not apply
	return fit_device(offsets, F, y, xr, w, nullptr, N, p, opt, out, nullptr, (unsigned char *)workspace + xbytes,
	                  workspace_bytes - xbytes, st, err, 0);
}

// ---------------------------------------------------------------------------------------------
// Ingest: unsorted rows -> group-sorted CSR (ingest.cuh)
// ---------------------------------------------------------------------------------------------
struct IngestWs {
	uint32_t *keys[2], *vals[2], *hist, *sums, *counts;
	int *bad;
	int64_t n_tiles;
	size_t bytes;
};
static size_t ingest_carve(IngestWs &w, unsigned char *base, int64_t N, int64_t G) {
	size_t off = 0;
	auto take = [&](size_t bytes) {
		size_t at = off;
		off += (bytes + 255) & ~(size_t)255;
		return base ? base + at : (unsigned char *)nullptr;
	};
	const size_t n1 = (size_t)std::max<int64_t>(N, 1), g1 = (size_t)std::max<int64_t>(G, 1);
	w.n_tiles = (int64_t)((n1 + kRadixTile - 1) / kRadixTile);
	for (int i = 0; i < 2; ++i) {
		w.keys[i] = (uint32_t *)take(n1 * 4);
		w.vals[i] = (uint32_t *)take(n1 * 4);
	}
	const size_t nh = (size_t)256 * w.n_tiles;
	w.hist = (uint32_t *)take(nh * 4);
	w.sums = (uint32_t *)take((std::max(nh, g1) / kScanBlock + 2) * 4);
	w.counts = (uint32_t *)take((g1 + 1) * 4);
	w.bad = (int *)take(256);
	w.bytes = off;
	return off;
}
static int scan_u32(uint32_t *a, int64_t n, uint32_t *sums, cudaStream_t st, AnofoxError *err) {
	if (n <= 0) return 0;
	const int64_t nb = (n + kScanBlock - 1) / kScanBlock;
	scan_local_kernel<<<(unsigned)nb, 256, 0, st>>>(a, n, sums);
	LAUNCHED();
	if (nb > 1) {
		scan_sums_kernel<<<1, 256, 0, st>>>(sums, nb);
		LAUNCHED();
		scan_add_kernel<<<(unsigned)nb, 256, 0, st>>>(a, n, sums);
		LAUNCHED();
	}
	CU(cudaGetLastError(), "scan kernels");
	return 0;
}
static int ingest_device(const int64_t *group_ids, const double *y, const double *x, const double *w, int64_t N, int p,
                         int64_t G, int64_t *offsets_out, double *y_out, double *x_out, double *w_out, int64_t *row_index_out,
                         void *workspace, size_t workspace_bytes, cudaStream_t st, AnofoxError *err) {
	clear_error(err);
	if (N < 0 || G <= 0 || p <= 0 || p > kMaxFeatures || !offsets_out || (N > 0 && (!group_ids || !y || !x || !y_out || !x_out)) ||
	    (w && !w_out)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "ingest: bad arguments");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (N >= ((int64_t)1 << 32) || G >= ((int64_t)1 << 32)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "ingest: at most 2^32 - 1 rows and groups per call");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	IngestWs ws;
	const size_t need = ingest_carve(ws, (unsigned char *)workspace, N, G);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned (anofox_b200_ingest_workspace_bytes)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	CU(cudaMemsetAsync(ws.counts, 0, (size_t)(G + 1) * 4, st), "memset");
	CU(cudaMemsetAsync(ws.bad, 0, 4, st), "memset");
	if (N > 0) {
		ingest_keys_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(group_ids, N, G, ws.keys[0], ws.counts, ws.bad);
		LAUNCHED();
		CU(cudaGetLastError(), "ingest_keys_kernel");
	}
	int rc = scan_u32(ws.counts, G, ws.sums, st, err);
	if (rc) return rc;
	ingest_offsets_kernel<<<(unsigned)((G + 1 + 255) / 256), 256, 0, st>>>(ws.counts, G, N, offsets_out);
	LAUNCHED();
	CU(cudaGetLastError(), "ingest_offsets_kernel");
	if (N == 0) return 0;
	int bits = 1;
	while (bits < 32 && ((int64_t)1 << bits) < G) ++bits;
	const int passes = (bits + 7) / 8;
	int cur = 0;
	for (int ps = 0; ps < passes; ++ps) {
		radix_hist_kernel<<<(unsigned)ws.n_tiles, kRadixThreads, 0, st>>>(ws.keys[cur], N, 8 * ps, ws.n_tiles, ws.hist);
		LAUNCHED();
		rc = scan_u32(ws.hist, (int64_t)256 * ws.n_tiles, ws.sums, st, err);
		if (rc) return rc;
		radix_scatter_kernel<<<(unsigned)ws.n_tiles, kRadixThreads, 0, st>>>(ws.keys[cur], ws.vals[cur], N, 8 * ps, ws.n_tiles, ws.hist,
		                                                                    ws.keys[cur ^ 1], ws.vals[cur ^ 1], ps == 0 ? 1 : 0);
		LAUNCHED();
		CU(cudaGetLastError(), "radix pass");
		cur ^= 1;
	}
	ingest_gather_kernel<<<(unsigned)((N + kGatherRows - 1) / kGatherRows), 256, 0, st>>>(ws.vals[cur], N, p, y, x, w, y_out, x_out, w_out,
	                                                                                       row_index_out);
	LAUNCHED();
	CU(cudaGetLastError(), "ingest_gather_kernel");
	return 0;
}

// ---------------------------------------------------------------------------------------------
// Window form of the fit-predict aggregates (window.cuh)
// ---------------------------------------------------------------------------------------------
static size_t window_scratch(AnofoxB200FitOutputs &o, unsigned char *base, int64_t N, int p) {
	size_t off = 0;
	auto take = [&](size_t bytes) {
		size_t at = off;
		off += (bytes + 255) & ~(size_t)255;
		return base ? base + at : (unsigned char *)nullptr;
	};
	// The per-frame fits stay on chip (SmallGroupView::fit_x): no per-row scratch arrays any more. The workspace
	// argument of the entry points stays (ABI), a token 256 bytes.
	(void)N;
	(void)p;
	memset(&o, 0, sizeof o);
	take(256);
	return off;
}

static int window_device(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w, int64_t N,
                         int p, const AnofoxB200FitOptions *opt, int64_t preceding, int64_t end_offset,
                         const AnofoxB200PredictOutputs *pred, void *workspace, size_t workspace_bytes, cudaStream_t st,
                         AnofoxError *err) {
	if (p > kWindowMaxFeatures) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "the window form supports up to 8 features");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (end_offset < 0) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "frame end must be CURRENT ROW (0) or k PRECEDING (k > 0)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	// the scan serves expanding frames that end at the current or the previous row; an expanding frame that ends
	// further back is summed directly like a bounded one
	if (preceding < 0 && end_offset > 1) preceding = (int64_t)1 << 60;
	if (!pred || !pred->yhat || !pred->yhat_lower || !pred->yhat_upper) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "prediction outputs are NULL");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	AnofoxB200FitOutputs scratch;
	const size_t need = window_scratch(scratch, (unsigned char *)workspace, N, p);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned (anofox_b200_window_workspace_bytes)");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (G == 0 || N == 0) return 0;
	DeviceInfo &di = device_info();
	const bool wls = opt->method == ANOFOX_B200_WLS;
	FitParams P;
	P.p = p;
	P.method = opt->method;
	P.fit_intercept = opt->fit_intercept != 0;
	P.compute_inference = 0; // ols_fit_predict.cpp:281
	P.lambda_scaling = opt->lambda_scaling;
	P.level = opt->confidence_level;
	P.alpha = opt->alpha;
	P.predict_mode = 1;
	P.min_state_rows = 0; // the window form has its own "more than p + intercept training rows" rule
	P.qr_count = nullptr;
	P.qr_list = nullptr;
	DfTables T;
	const int trc = get_tables(P.level, st, T, err);
	if (trc) return trc;
	P.tcrit = T.tcrit;
	P.gratio = T.gratio;
	P.rcp = T.rcp;
	P.ce = T.ce;
	P.co = T.co;
	WindowArgs a;
	a.offsets = offsets;
	a.y = y;
	a.x = x;
	a.w = wls ? w : nullptr;
	a.n_groups = G;
	a.preceding = preceding;
	a.end_offset = end_offset;
	a.yhat = pred->yhat;
	a.lower = pred->yhat_lower;
	a.upper = pred->yhat_upper;
	const int64_t warps = std::min<int64_t>(G, (int64_t)di.sms * 16);
	CU(launch_window(p, wls, a, P, scratch, (int)((warps + 3) / 4), st), "window_fit_predict_kernel");
	LAUNCHED();
	return 0;
}

static int window_host(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w, int64_t N, int p,
                       const AnofoxB200FitOptions *opt, int64_t preceding, int64_t end_offset,
                       const AnofoxB200PredictOutputs *pred, AnofoxError *err) {
	if (G > 0) {
		bool ok = offsets[0] == 0 && offsets[G] == N;
		for (int64_t g = 0; ok && g < G; ++g) ok = offsets[g] <= offsets[g + 1];
		if (!ok) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "offsets must start at 0, be non-decreasing and end at n_rows");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
	}
	if (!pred || !pred->yhat || !pred->yhat_lower || !pred->yhat_upper) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "prediction outputs are NULL");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	HostLease lease;
	HostContext &c = *lease.c;
	DrainGuard drain{c};
	AnofoxB200FitOutputs tmp;
	const size_t ws_bytes = window_scratch(tmp, nullptr, N, p);
	DevLayout L;
	const int64_t N1 = std::max<int64_t>(N, 1);
	const size_t o_off = L.take((size_t)(G + 1) * 8), o_y = L.take((size_t)N1 * 8), o_x = L.take((size_t)N1 * p * 8);
	const size_t o_w = w ? L.take((size_t)N1 * 8) : 0;
	const size_t o_pred = L.take((size_t)N1 * 8 * 3);
	const size_t o_ws = L.take(ws_bytes);
	int rc = c.ensure(L.off, err);
	if (rc) return rc;
	unsigned char *d = c.dev;
	cudaStream_t st = c.stream;
	CU(cudaMemcpyAsync(d + o_off, offsets, (size_t)(G + 1) * 8, cudaMemcpyHostToDevice, st), "H2D offsets");
	if (N > 0) {
		CU(cudaMemcpyAsync(d + o_y, y, (size_t)N * 8, cudaMemcpyHostToDevice, st), "H2D y");
		CU(cudaMemcpyAsync(d + o_x, x, (size_t)N * p * 8, cudaMemcpyHostToDevice, st), "H2D x");
		if (w) CU(cudaMemcpyAsync(d + o_w, w, (size_t)N * 8, cudaMemcpyHostToDevice, st), "H2D w");
	}
	AnofoxB200PredictOutputs dp;
	memset(&dp, 0, sizeof dp);
	dp.yhat = (double *)(d + o_pred);
	dp.yhat_lower = dp.yhat + N1;
	dp.yhat_upper = dp.yhat + 2 * N1;
	rc = window_device((const int64_t *)(d + o_off), G, (const double *)(d + o_y), (const double *)(d + o_x),
	                   w ? (const double *)(d + o_w) : nullptr, N, p, opt, preceding, end_offset, &dp, d + o_ws, ws_bytes, st,
	                   err);
	if (rc) return rc;
	if (N > 0) {
		CU(cudaMemcpyAsync(pred->yhat, dp.yhat, (size_t)N * 8, cudaMemcpyDeviceToHost, st), "D2H");
		CU(cudaMemcpyAsync(pred->yhat_lower, dp.yhat_lower, (size_t)N * 8, cudaMemcpyDeviceToHost, st), "D2H");
		CU(cudaMemcpyAsync(pred->yhat_upper, dp.yhat_upper, (size_t)N * 8, cudaMemcpyDeviceToHost, st), "D2H");
	}
	CU(cudaStreamSynchronize(st), "sync");
	return 0;
}

static int eval_specfun_host(int mode, int n, const double *a, const double *b, const double *c, double *out,
                             AnofoxError *err) {
	clear_error(err);
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	if (n <= 0) return 0;
	HostLease lease;
	HostContext &ctx = *lease.c;
	DrainGuard drain{ctx};
	const size_t bytes = (size_t)n * 8;
	int rc = ctx.ensure(4 * ((bytes + 255) & ~(size_t)255), err);
	if (rc) return rc;
	const size_t step = (bytes + 255) & ~(size_t)255;
	double *da = (double *)ctx.dev, *db = (double *)(ctx.dev + step), *dc = (double *)(ctx.dev + 2 * step),
	       *dout = (double *)(ctx.dev + 3 * step);
	cudaStream_t st = ctx.stream;
	CU(cudaMemcpyAsync(da, a, bytes, cudaMemcpyHostToDevice, st), "H2D");
	CU(cudaMemcpyAsync(db, b, bytes, cudaMemcpyHostToDevice, st), "H2D");
	if (c) CU(cudaMemcpyAsync(dc, c, bytes, cudaMemcpyHostToDevice, st), "H2D");
	DfTables T{nullptr, nullptr, nullptr, nullptr, nullptr};
	if (mode >= 3) {
		rc = get_tables(0.95, st, T, err);
		if (rc) return rc;
	}
	specfun_eval_kernel<<<(n + 127) / 128, 128, 0, st>>>(mode, n, da, db, dc, dout, T);
	LAUNCHED();
	CU(cudaGetLastError(), "specfun_eval_kernel");
	CU(cudaMemcpyAsync(out, dout, bytes, cudaMemcpyDeviceToHost, st), "D2H");
	CU(cudaStreamSynchronize(st), "sync");
	return 0;
}

} // namespace b200

// =============================================================================================
// C ABI
// =============================================================================================
using namespace b200;

extern "C" {

bool anofox_ols_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxOlsOptions options,
                    AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference, AnofoxError *out_error) {
	return guarded(out_error, "OLS fit", [&] {
		       return single_fit(ANOFOX_B200_OLS, y, x, x_count, nullptr, options.fit_intercept, options.compute_inference,
		                         options.confidence_level, 0.0, 0, (int)options.solver, (int)options.hc_type, out_core,
		                         out_inference, out_error)
		                  ? 0
		                  : 1;
	       }) == 0;
}

bool anofox_ridge_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxRidgeOptions options,
                      AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference, AnofoxError *out_error) {
	return guarded(out_error, "Ridge fit", [&] {
		       return single_fit(ANOFOX_B200_RIDGE, y, x, x_count, nullptr, options.fit_intercept, options.compute_inference,
		                         options.confidence_level, options.alpha, (int)options.lambda_scaling, (int)options.solver,
		                         ANOFOX_HC_NONE, out_core, out_inference, out_error)
		                  ? 0
		                  : 1;
	       }) == 0;
}

bool anofox_wls_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxDataArray weights,
                    AnofoxWlsOptions options, AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference,
                    AnofoxError *out_error) {
	return guarded(out_error, "WLS fit", [&] {
		       return single_fit(ANOFOX_B200_WLS, y, x, x_count, &weights, options.fit_intercept, options.compute_inference,
		                         options.confidence_level, 0.0, 0, (int)options.solver, (int)options.hc_type, out_core,
		                         out_inference, out_error)
		                  ? 0
		                  : 1;
	       }) == 0;
}

void anofox_free_result_core(AnofoxFitResultCore *result) {
	if (!result) return;
	free(result->coefficients);
	result->coefficients = nullptr;
}

void anofox_free_result_inference(AnofoxFitResultInference *result) {
	if (!result) return;
	free(result->std_errors);
	free(result->t_values);
	free(result->p_values);
	free(result->ci_lower);
	free(result->ci_upper);
	result->std_errors = result->t_values = result->p_values = result->ci_lower = result->ci_upper = nullptr;
}

double anofox_t_critical(double confidence_level, size_t df) {
	if (df == 0 || !(confidence_level > 0.0) || !(confidence_level < 1.0)) return NAN; // lib.rs:2218-2220
	double a = confidence_level, b = (double)df, out = NAN;
	AnofoxError err;
	if (guarded(&err, "t_critical", [&] { return eval_specfun_host(1, 1, &a, &b, nullptr, &out, &err); }) != 0) return NAN;
	return out;
}

bool anofox_predict_with_interval(const double *coefficients, size_t coefficients_len, double intercept,
                                  const double *x_new, size_t x_len, double residual_std_error,
                                  size_t n_observations, double confidence_level, AnofoxPredictionResult *out) {
	if (!out) return false;
	out->yhat = out->yhat_lower = out->yhat_upper = NAN;
	if (!coefficients || coefficients_len == 0) return false;
	if (!x_new || x_len != coefficients_len) return false;
	// lib.rs:2292-2347. The dot product of one row is scalar host work; the t quantile is the
	// device routine (anofox_t_critical above), as in the batched predict kernel.
	double yhat = std::isnan(intercept) ? 0.0 : intercept;
	for (size_t j = 0; j < coefficients_len; ++j)
		if (!std::isnan(coefficients[j])) yhat += coefficients[j] * x_new[j];
	out->yhat = out->yhat_lower = out->yhat_upper = yhat;
	if (std::isnan(residual_std_error) || residual_std_error <= 0.0 || n_observations <= coefficients_len + 1) return true;
	const size_t k = coefficients_len + (std::isnan(intercept) ? 0 : 1);
	const size_t df = n_observations > k ? n_observations - k : 0;
	if (df == 0) return true;
	const double tc = anofox_t_critical(confidence_level, df);
	if (std::isnan(tc)) return true;
	const double margin = tc * (residual_std_error * std::sqrt(1.0 + 1.0 / (double)n_observations));
	out->yhat_lower = yhat - margin;
	out->yhat_upper = yhat + margin;
	return true;
}

static bool info_crit(double rss, size_t n, size_t k, double *out, bool bic, AnofoxError *err) {
	clear_error(err);
	if (!out) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, bic ? "out_bic is NULL" : "out_aic is NULL");
		return false;
	}
	if (n == 0) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "n must be > 0");
		return false;
	}
	if (rss < 0.0) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "RSS must be non-negative");
		return false;
	}
	if (rss == 0.0) {
		*out = -INFINITY;
		return true;
	}
	const double nf = (double)n, kf = (double)k;
	*out = nf * std::log(rss / nf) + (bic ? kf * std::log(nf) : 2.0 * kf);
	return true;
}
bool anofox_compute_aic(double rss, size_t n, size_t k, double *out_aic, AnofoxError *err) {
	return info_crit(rss, n, k, out_aic, false, err);
}
bool anofox_compute_bic(double rss, size_t n, size_t k, double *out_bic, AnofoxError *err) {
	return info_crit(rss, n, k, out_bic, true, err);
}

size_t anofox_b200_workspace_bytes(int64_t n_groups, int64_t n_rows, int32_t n_features) {
	if (n_groups < 0 || n_rows < 0 || n_features <= 0 || n_features > kMaxFeatures) return 0;
	DeviceInfo &di = device_info();
	Workspace w;
	return carve(w, nullptr, n_groups, n_rows, n_features, di.sms > 0 ? di.sms : 148);
}

int32_t anofox_b200_fit_batch_device(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                     const double *w, int64_t n_rows, int32_t n_features,
                                     const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out,
                                     void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err) {
	return guarded(err, "fit_batch_device", [&]() -> int32_t {
		int rc = validate_common(offsets, n_groups, y, x, w, n_rows, n_features, options, err);
		if (rc) return rc;
		if (!out) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "outputs is NULL");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		if ((rc = check_outputs(out, options->compute_inference != 0, err))) return rc;
		return fit_device(offsets, n_groups, y, x, w, nullptr, n_rows, n_features, options, *out, nullptr, workspace,
		                  workspace_bytes, (cudaStream_t)stream, err);
	});
}

int32_t anofox_b200_fit_batch_host(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                   const double *w, int64_t n_rows, int32_t n_features,
                                   const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out,
                                   AnofoxError *err) {
	return guarded(err, "fit_batch_host", [&]() -> int32_t {
		int rc = validate_common(offsets, n_groups, y, x, w, n_rows, n_features, options, err);
		if (rc) return rc;
		if (!out) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "outputs is NULL");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		return fit_host(offsets, n_groups, y, x, w, nullptr, n_rows, n_features, options, out, nullptr, err);
	});
}

int32_t anofox_b200_fit_predict_batch_device(const int64_t *offsets, int64_t n_groups, const double *y,
                                             const double *x, const double *w, const uint8_t *train,
                                             int64_t n_rows, int32_t n_features,
                                             const AnofoxB200FitOptions *options,
                                             const AnofoxB200PredictOutputs *pred, const AnofoxB200FitOutputs *fit,
                                             void *workspace, size_t workspace_bytes, void *stream,
                                             AnofoxError *err) {
	return guarded(err, "fit_predict_batch_device", [&]() -> int32_t {
		int rc = validate_common(offsets, n_groups, y, x, w, n_rows, n_features, options, err);
		if (rc) return rc;
		if (!pred || !pred->yhat || !pred->yhat_lower || !pred->yhat_upper) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "prediction outputs are NULL");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		if (fit && fit->coefficients && (rc = check_outputs(fit, false, err))) return rc; // all seven, or none
		AnofoxB200FitOutputs none;
		memset(&none, 0, sizeof none);
		return fit_device(offsets, n_groups, y, x, w, train, n_rows, n_features, options, fit ? *fit : none, pred,
		                  workspace, workspace_bytes, (cudaStream_t)stream, err);
	});
}

int32_t anofox_b200_fit_predict_batch_host(const int64_t *offsets, int64_t n_groups, const double *y,
                                           const double *x, const double *w, const uint8_t *train, int64_t n_rows,
                                           int32_t n_features, const AnofoxB200FitOptions *options,
                                           const AnofoxB200PredictOutputs *pred, const AnofoxB200FitOutputs *fit,
                                           AnofoxError *err) {
	return guarded(err, "fit_predict_batch_host", [&]() -> int32_t {
		int rc = validate_common(offsets, n_groups, y, x, w, n_rows, n_features, options, err);
		if (rc) return rc;
		if (!pred || !pred->yhat || !pred->yhat_lower || !pred->yhat_upper || !pred->is_training) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "prediction outputs are NULL");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		return fit_host(offsets, n_groups, y, x, w, train, n_rows, n_features, options, fit, pred, err);
	});
}

size_t anofox_b200_moment_block_len(int32_t n_features) {
	if (n_features <= 0 || n_features > kMaxFeatures) return 0;
	const int k = n_features + 1;
	return (size_t)(4 + k + k * (k + 1) / 2);
}

size_t anofox_b200_ingest_workspace_bytes(int64_t n_rows, int64_t n_groups) {
	IngestWs w;
	return ingest_carve(w, nullptr, n_rows, n_groups);
}

int32_t anofox_b200_ingest_device(const int64_t *group_ids, const double *y, const double *x, const double *w, int64_t n_rows,
                                  int32_t n_features, int64_t n_groups, int64_t *offsets_out, double *y_out, double *x_out,
                                  double *w_out, int64_t *row_index_out, void *workspace, size_t workspace_bytes, void *stream,
                                  AnofoxError *err) {
	return guarded(err, "ingest_device", [&]() -> int32_t {
		DeviceInfo &di = device_info();
		if (!di.ok) {
			set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
			return ANOFOX_ERROR_INTERNAL;
		}
		return ingest_device(group_ids, y, x, w, n_rows, n_features, n_groups, offsets_out, y_out, x_out, w_out, row_index_out,
		                     workspace, workspace_bytes, (cudaStream_t)stream, err);
	});
}

// 1 when the last ingest on this workspace met a group id outside [0, n_groups) (such rows were filed under group 0)
int32_t anofox_b200_ingest_bad_ids(const void *workspace, int64_t n_rows, int64_t n_groups) {
	IngestWs w;
	ingest_carve(w, (unsigned char *)workspace, n_rows, n_groups);
	int v = 0;
	if (cudaMemcpy(&v, w.bad, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
	return v;
}

// ---- ingest without any allocation: scratch in a file-scope device array ----------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it). No counterpart in the reference. The scratch of the radix
// partition (keys, row indices, histograms, counters: ingest_carve) lives in g_ingest_static_scratch, which every
// GPU that loads the library holds until the process exits. In HOST code the array's name is only a stand-in for
// the symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device, with that device
// current, and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_ingest_static_scratch[ANOFOX_B200_INGEST_STATIC_MAX_BYTES];

static int ingest_static_scratch(unsigned char **out, AnofoxError *err) {
	static std::mutex mu;
	static unsigned char *addr[64];
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess || dev < 0 || dev >= 64) {
		set_error(err, ANOFOX_ERROR_INTERNAL, e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range");
		return ANOFOX_ERROR_INTERNAL;
	}
	std::lock_guard<std::mutex> lock(mu);
	if (!addr[dev]) {
		void *ptr = nullptr;
		e = cudaGetSymbolAddress(&ptr, g_ingest_static_scratch);
		if (e != cudaSuccess || !ptr) {
			char msg[200];
			snprintf(msg, sizeof msg, "cudaGetSymbolAddress(g_ingest_static_scratch): %s", cudaGetErrorString(e));
			set_error(err, ANOFOX_ERROR_INTERNAL, msg);
			return ANOFOX_ERROR_INTERNAL;
		}
		addr[dev] = (unsigned char *)ptr;
	}
	*out = addr[dev];
	return 0;
}

int32_t anofox_b200_ingest_device_static(const int64_t *group_ids, const double *y, const double *x, const double *w,
                                         int64_t n_rows, int32_t n_features, int64_t n_groups, int64_t *offsets_out,
                                         double *y_out, double *x_out, double *w_out, int64_t *row_index_out,
                                         int32_t *bad_ids_out, void *stream, AnofoxError *err) {
	return guarded(err, "ingest_device_static", [&]() -> int32_t {
		clear_error(err);
		DeviceInfo &di = device_info();
		if (!di.ok) {
			set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
			return ANOFOX_ERROR_INTERNAL;
		}
		if (n_rows < 0 || n_groups <= 0) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "ingest: bad arguments");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		IngestWs ws;
		const size_t need = ingest_carve(ws, nullptr, n_rows, n_groups);
		if (need > (size_t)ANOFOX_B200_INGEST_STATIC_MAX_BYTES) {
			char msg[240];
			snprintf(msg, sizeof msg,
			         "ingest_device_static: %lld rows x %lld groups need %zu bytes of scratch, the built-in array holds %zu "
			         "(ANOFOX_B200_INGEST_STATIC_MAX_BYTES); use anofox_b200_ingest_device with a workspace",
			         (long long)n_rows, (long long)n_groups, need, (size_t)ANOFOX_B200_INGEST_STATIC_MAX_BYTES);
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, msg);
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		unsigned char *scratch = nullptr;
		int rc = ingest_static_scratch(&scratch, err);
		if (rc) return rc;
		// ingest_device clears the counters and the flag it needs at zero itself; everything else is overwritten
		rc = ingest_device(group_ids, y, x, w, n_rows, n_features, n_groups, offsets_out, y_out, x_out, w_out, row_index_out,
		                   scratch, (size_t)ANOFOX_B200_INGEST_STATIC_MAX_BYTES, (cudaStream_t)stream, err);
		if (rc) return rc;
		if (bad_ids_out) {
			ingest_carve(ws, scratch, n_rows, n_groups);
			CU(cudaMemcpyAsync(bad_ids_out, ws.bad, sizeof(int32_t), cudaMemcpyDeviceToDevice, (cudaStream_t)stream), "bad flag copy");
		}
		return 0;
	});
}

size_t anofox_b200_fit_columns_workspace_bytes(int64_t n_fits, int64_t n_rows, int32_t n_features) {
	if (n_fits < 0 || n_rows < 0 || n_features <= 0 || n_features > kMaxFeatures) return 0;
	DeviceInfo &di = device_info();
	Workspace w;
	return ((((size_t)std::max<int64_t>(n_rows, 1) * n_features * 8) + 255) & ~(size_t)255) +
	       carve(w, nullptr, n_fits, n_rows, n_features, di.ok ? di.sms : 148);
}

int32_t anofox_b200_fit_columns_batch_device(const int64_t *offsets, int64_t n_fits, const double *y, const double *x_columns,
                                             const double *w, int64_t n_rows, int32_t n_features,
                                             const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out,
                                             void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err) {
	return guarded(err, "fit_columns_batch_device", [&]() -> int32_t {
		int rc = validate_common(offsets, n_fits, y, x_columns, w, n_rows, n_features, options, err);
		if (rc) return rc;
		if (!out) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "outputs is NULL");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		if ((rc = check_outputs(out, options->compute_inference != 0, err))) return rc;
		return fit_columns_device(offsets, n_fits, y, x_columns, w, n_rows, n_features, options, *out, workspace,
		                          workspace_bytes, (cudaStream_t)stream, err);
	});
}

size_t anofox_b200_window_workspace_bytes(int64_t n_rows, int32_t n_features) {
	AnofoxB200FitOutputs tmp;
	return window_scratch(tmp, nullptr, n_rows, n_features);
}

int32_t anofox_b200_window_fit_predict_device(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                              const double *w, int64_t n_rows, int32_t n_features,
                                              const AnofoxB200FitOptions *options, int64_t frame_preceding,
                                              int64_t frame_end_offset, const AnofoxB200PredictOutputs *pred,
                                              void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err) {
	return guarded(err, "window_fit_predict_device", [&]() -> int32_t {
		int rc = validate_common(offsets, n_groups, y, x, w, n_rows, n_features, options, err);
		if (rc) return rc;
		return window_device(offsets, n_groups, y, x, w, n_rows, n_features, options, frame_preceding, frame_end_offset, pred,
		                     workspace, workspace_bytes, (cudaStream_t)stream, err);
	});
}

int32_t anofox_b200_window_fit_predict_host(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                            const double *w, int64_t n_rows, int32_t n_features,
                                            const AnofoxB200FitOptions *options, int64_t frame_preceding,
                                            int64_t frame_end_offset, const AnofoxB200PredictOutputs *pred,
                                            AnofoxError *err) {
	return guarded(err, "window_fit_predict_host", [&]() -> int32_t {
		int rc = validate_common(offsets, n_groups, y, x, w, n_rows, n_features, options, err);
		if (rc) return rc;
		if (n_features > kWindowMaxFeatures) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "the window form supports up to 8 features");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		return window_host(offsets, n_groups, y, x, w, n_rows, n_features, options, frame_preceding, frame_end_offset, pred,
		                   err);
	});
}

int32_t anofox_b200_first_valid_row_device(const double *y, const double *x, const double *w, int64_t n_rows,
                                           int32_t n_features, double *shift_out, int32_t *found_out, void *workspace,
                                           size_t workspace_bytes, void *stream, AnofoxError *err) {
	clear_error(err);
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	const int p = n_features;
	if (p <= 0 || p > kMaxFeatures || n_rows < 0 || !shift_out || !found_out || (n_rows > 0 && (!x || !y))) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "invalid argument");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	Workspace ws;
	const size_t need = carve(ws, (unsigned char *)workspace, 1, n_rows, p, di.sms);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	cudaStream_t st = (cudaStream_t)stream;
	unsigned long long *best = (unsigned long long *)ws.big_list;
	CU(cudaMemsetAsync(best, 0xff, sizeof(unsigned long long), st), "memset");
	if (n_rows > 0) {
		const int blocks = (int)std::min<int64_t>(di.sms * 4, (n_rows + 255) / 256);
		shard_first_valid_kernel<<<blocks, 256, 0, st>>>(y, x, w, n_rows, p, best);
		LAUNCHED();
		CU(cudaGetLastError(), "shard_first_valid_kernel");
	}
	shard_shift_kernel<<<1, 128, 0, st>>>(y, x, p, best, n_rows, shift_out, found_out);
	LAUNCHED();
	CU(cudaGetLastError(), "shard_shift_kernel");
	return 0;
}

int32_t anofox_b200_moments_device(const double *y, const double *x, const double *w, int64_t n_rows,
                                   int32_t n_features, const double *shift, double *moments, double *cdev,
                                   void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err) {
	clear_error(err);
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	const int p = n_features;
	if (p <= 0 || p > kMaxFeatures || n_rows < 0 || !shift || !moments || !cdev || (n_rows > 0 && (!x || !y))) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "invalid argument");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	Workspace ws;
	const size_t need = carve(ws, (unsigned char *)workspace, 1, n_rows, p, di.sms);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	cudaStream_t st = (cudaStream_t)stream;
	// Fast pass: the ring kernel in slice mode (ring.cuh) -- one slice of the shard per CTA, the CTA's warps
	// interleave its quads, rows flow through the shared-memory ring. It checks nothing per element; a
	// non-finite value or an undecided constant column raises redo_count, and only then the checked
	// streaming pass below runs (decided on the device: no host round trip).
	const Shape shp = shape_of(p);
	const int ncw = ring_warps(shp.nt);
	const int units = (int)std::max<int64_t>(1, std::min<int64_t>(di.sms, (n_rows + 4095) / 4096));
	const RingGeom rg_ = ring_geometry(p, w != nullptr, false, n_rows, units, units, ncw);
	CU(cudaMemsetAsync(ws.redo_count, 0, 256, st), "memset");
	slice_bounds_kernel<<<1, 256, 0, st>>>(n_rows, units, ws.slice_off, ws.slice_off + units + 1);
	LAUNCHED();
	CU(cudaGetLastError(), "slice_bounds_kernel");
	RingArgs r;
	r.offsets = ws.slice_off;
	r.y = y;
	r.x = x;
	r.w = w;
	r.train = nullptr;
	r.p = p;
	r.tile_g0 = ws.slice_off + units + 1;
	r.records = ws.ring_partials;
	r.shift_common = shift;
	r.flags_from_y = 0;
	r.big_list = ws.big_list;
	r.big_count = ws.big_count;
	r.big_rows = (int64_t)1 << 40;
	r.redo_list = ws.redo_list;
	r.redo_count = ws.redo_count;
	r.win_rows = rg_.win_rows;
	r.n_win = rg_.n_win;
	r.team = rg_.team;
	r.x_off = rg_.x_off;
	r.y_off = rg_.y_off;
	r.w_off = rg_.w_off;
	r.t_off = rg_.t_off;
	r.slot_off = rg_.slot_off;
	StreamArgs s;
	s.offsets = nullptr;
	s.big_list = nullptr;
	s.big_count = nullptr;
	s.fixed_rows = n_rows;
	s.y = y;
	s.x = x;
	s.w = w;
	s.train = nullptr;
	s.p = p;
	s.shift = const_cast<double *>(shift);
	s.first_ok = nullptr;
	s.partials = ws.partials;
	s.n_part = ws.n_part;
	s.records = ws.records;
	s.rec_gs = rec_len(p);
	s.rec_es = 1;
	s.run_if = ws.redo_count;
	s.flags_from_y = 0;
	s.huge_list = nullptr;
	s.huge_count = nullptr;
	s.huge_rows = 0;
	s.b_begin = 0;
	s.b_cap = 1;
	s.ring_partials = ws.ring_partials;
	s.ring_n_part = units * ncw;
	timer_mark(st);
	if (n_rows > 0 && g_force_path.load() != 3 && shp.nt >= kSliceMinTiles && g_force_path.load() != 4) {
		// wide Gram: the split-role slice kernel (slice.cuh), 8 partial records per slice
		SliceArgs sa;
		sa.bounds = ws.slice_off;
		sa.y = y;
		sa.x = x;
		sa.w = w;
		sa.p = p;
		sa.shift = shift;
		sa.partials = ws.ring_partials;
		sa.redo_count = ws.redo_count;
		const int64_t row_bytes = 8 * (int64_t)(p + 1) + (w ? 8 : 0);
		const int sr = kSliceRows;
		sa.win_rows = sr;
		sa.x_slot = (uint32_t)(sr * p * 8 + 16);
		sa.y_slot = (uint32_t)(sr * 8 + 16);
		const int64_t slot_bytes = (int64_t)sa.x_slot + sa.y_slot * (w ? 2 : 1);
		sa.n_win = (int)std::min<int64_t>(kSliceMaxWin, (216 * 1024) / slot_bytes);
		auto a128 = [](int64_t v) { return (uint32_t)((v + 127) & ~(int64_t)127); };
		sa.x_off = 0;
		sa.y_off = a128((int64_t)sa.n_win * sa.x_slot);
		sa.w_off = sa.y_off + a128((int64_t)sa.n_win * sa.y_slot);
		const size_t dyn = sa.w_off + (w ? a128((int64_t)sa.n_win * sa.y_slot) : 0);
		s.ring_n_part = units * kSliceWarps;
		CU(launch_slice(sa, w != nullptr, units, dyn, st), "acc_slice_kernel");
	} else if (n_rows > 0 && g_force_path.load() != 3) {
		CU(launch_ring(r, w != nullptr, false, units, rg_.dyn_bytes, st), "acc_ring_kernel (slices)");
	} else {
		s.run_if = nullptr; // empty shard or A/B run: the checked streaming pass alone
	}
	CU(launch_stream(s, ws.n_part / kStreamWarps, st), "acc_stream_kernel");
	timer_mark(st);
	dim3 rg((unsigned)((rec_len(p) + 127) / 128), 1);
	reduce_partials_kernel<<<rg, 128, 0, st>>>(s);
	LAUNCHED();
	CU(cudaGetLastError(), "reduce_partials_kernel");
	split_record_kernel<<<1, 256, 0, st>>>(ws.records, p, moments, cdev);
	LAUNCHED();
	CU(cudaGetLastError(), "split_record_kernel");
	timer_mark(st);
	return 0;
}

// Common body of the moment-block solves. needs_refine (device int32, may be NULL): set to 1 by the first solve
// (rss_exact == NULL) when the residual sum of squares of the moment form has cancelled (rss <= 1e-5 tss); with
// rss_exact (device double) the group is solved again with that value, only when *needs_refine != 0.
static int solve_moments_impl(const double *moments, const double *cdev, const double *shift, int32_t n_features,
                              const AnofoxB200FitOptions *options, const AnofoxB200FitOutputs *out, int32_t *needs_refine,
                              const double *rss_exact, void *stream, AnofoxError *err) {
	clear_error(err);
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	const int p = n_features;
	if (p <= 0 || p > kMaxFeatures || !moments || !cdev || !shift || !options || !out || (rss_exact && !needs_refine)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "invalid argument");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	// the same option checks as the batched calls (validate_common)
	if (options->method < ANOFOX_B200_OLS || options->method > ANOFOX_B200_WLS) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "unknown method");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (options->method == ANOFOX_B200_RIDGE && options->alpha < 0.0) {
		set_error(err, ANOFOX_ERROR_INVALID_ALPHA, "Invalid alpha: must be >= 0");
		return ANOFOX_ERROR_INVALID_ALPHA;
	}
	if (options->hc_type < ANOFOX_HC_NONE || options->hc_type > ANOFOX_HC_HC3) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "unknown hc_type");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	if (options->hc_type != ANOFOX_HC_NONE) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "hc_type: robust standard errors need the rows; not available from a moment block");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	const int orc = check_outputs(out, options->compute_inference != 0, err);
	if (orc) return orc;
	cudaStream_t st = (cudaStream_t)stream;
	double *rec = nullptr;
	const int L = rec_len(p);
	CU(cudaMallocAsync(&rec, (size_t)(L + 2) * 8, st), "cudaMallocAsync");
	assemble_record_kernel<<<1, 256, 0, st>>>(moments, cdev, shift, p, rec);
	LAUNCHED();
	FitParams P;
	P.p = p;
	P.method = options->method;
	P.fit_intercept = options->fit_intercept != 0;
	P.compute_inference = options->compute_inference != 0;
	P.lambda_scaling = options->lambda_scaling;
	P.level = options->confidence_level;
	P.alpha = options->alpha;
	P.tcrit = P.gratio = P.rcp = P.ce = P.co = nullptr; // one group: evaluate the distribution functions directly
	P.predict_mode = 0;
	P.min_state_rows = 2;
	P.qr_count = nullptr; // no rows at hand here (the header says what a caller with rows can do)
	P.qr_list = nullptr;
	if (!rss_exact) {
		if (needs_refine) CU(cudaMemsetAsync(needs_refine, 0, sizeof(int32_t), st), "memset");
		CU(launch_solve(rec, 1, P, *out, needs_refine, needs_refine ? (int64_t *)(rec + L) : nullptr, st), "solve_kernel");
	} else {
		const size_t dyn = sizeof(WideSmem);
		CU(cudaFuncSetAttribute(solve_wide_override_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn), "attr");
		solve_wide_override_kernel<<<1, kWideThreads, dyn, st>>>(rec, P, *out, needs_refine, rss_exact);
		LAUNCHED();
		CU(cudaGetLastError(), "solve_wide_override_kernel");
	}
	CU(cudaFreeAsync(rec, st), "cudaFreeAsync");
	return 0;
}

int32_t anofox_b200_solve_moments_device(const double *moments, const double *cdev, const double *shift,
                                         int32_t n_features, const AnofoxB200FitOptions *options,
                                         const AnofoxB200FitOutputs *out, void *stream, AnofoxError *err) {
	return guarded(err, "solve_moments_device", [&]() -> int32_t {
		return solve_moments_impl(moments, cdev, shift, n_features, options, out, nullptr, nullptr, stream, err);
	});
}

int32_t anofox_b200_solve_moments_refine_device(const double *moments, const double *cdev, const double *shift,
                                                int32_t n_features, const AnofoxB200FitOptions *options,
                                                const AnofoxB200FitOutputs *out, int32_t *needs_refine,
                                                const double *rss_exact, void *stream, AnofoxError *err) {
	return guarded(err, "solve_moments_refine_device", [&]() -> int32_t {
		if (!needs_refine) {
			set_error(err, ANOFOX_ERROR_INVALID_INPUT, "needs_refine is NULL");
			return ANOFOX_ERROR_INVALID_INPUT;
		}
		return solve_moments_impl(moments, cdev, shift, n_features, options, out, needs_refine, rss_exact, stream, err);
	});
}

int32_t anofox_b200_shard_residual_device(const double *y, const double *x, const double *w, int64_t n_rows,
                                          int32_t n_features, const double *coefficients, const double *intercept,
                                          const double *shift, const int32_t *run_if, double *rss_partial,
                                          void *workspace, size_t workspace_bytes, void *stream, AnofoxError *err) {
	clear_error(err);
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	const int p = n_features;
	if (p <= 0 || p > kMaxFeatures || n_rows < 0 || !coefficients || !intercept || !run_if || !rss_partial ||
	    (n_rows > 0 && (!x || !y))) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "invalid argument");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	Workspace ws;
	const size_t need = carve(ws, (unsigned char *)workspace, 1, n_rows, p, di.sms);
	if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
		set_error(err, ANOFOX_ERROR_INVALID_INPUT, "workspace too small or misaligned");
		return ANOFOX_ERROR_INVALID_INPUT;
	}
	cudaStream_t st = (cudaStream_t)stream;
	const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)di.sms * 8, (n_rows + 7) / 8));
	shard_residual_kernel<<<blocks, 256, 0, st>>>(y, x, w, n_rows, p, coefficients, intercept, shift, run_if, ws.partials);
	LAUNCHED();
	CU(cudaGetLastError(), "shard_residual_kernel");
	shard_residual_sum_kernel<<<1, 32, 0, st>>>(ws.partials, blocks, rss_partial);
	LAUNCHED();
	CU(cudaGetLastError(), "shard_residual_sum_kernel");
	return 0;
}

int32_t anofox_b200_device_check(AnofoxError *err) {
	clear_error(err);
	DeviceInfo &di = device_info();
	if (!di.ok) {
		set_error(err, ANOFOX_ERROR_INTERNAL, di.why);
		return ANOFOX_ERROR_INTERNAL;
	}
	return 0;
}

int64_t anofox_b200_launch_count(void) { return (int64_t)g_launches.load(); }

const char *anofox_b200_version(void) { return "anofox-statistics_b200 0.1 (sm_100a)"; }

// Bench hooks: per-stage device times of the batched calls made while enabled.
void anofox_b200_stage_timing_enable(int32_t on) {
	StageTimer &t = stage_timer();
	std::lock_guard<std::mutex> lock(t.mu);
	t.clear();
	t.enabled = on != 0;
}
// Sums over all recorded calls (synchronises the device). Returns the number of calls.
int64_t anofox_b200_stage_timing_read(double *accumulate_ms, double *solve_ms) {
	StageTimer &t = stage_timer();
	std::lock_guard<std::mutex> lock(t.mu);
	cudaDeviceSynchronize();
	double a = 0.0, b = 0.0;
	int64_t n = 0;
	for (size_t i = 0; i + 3 <= t.ev.size(); i += 3) {
		float m1 = 0.f, m2 = 0.f;
		if (cudaEventElapsedTime(&m1, t.ev[i], t.ev[i + 1]) == cudaSuccess &&
		    cudaEventElapsedTime(&m2, t.ev[i + 1], t.ev[i + 2]) == cudaSuccess) {
			a += m1;
			b += m2;
			++n;
		}
	}
	if (accumulate_ms) *accumulate_ms = a;
	if (solve_ms) *solve_ms = b;
	t.clear();
	return n;
}

// Test / A-B hook: 0 = choose the kernel path automatically, 1 = always the record path (DMMA accumulation
// kernel + solve kernel), even where the fused small-group kernel would be chosen, 2 = fused small kernel,
// 3 = whole-group stages, 4 = ring slices for wide shards, 5 = auto + every group re-fitted by the QR kernel.
void anofox_b200_force_path(int32_t path) { g_force_path.store(path); }

// Test / bench hook: the device-side counters of the last call that used `workspace` (synchronises the device):
// out[0] big groups, [1] groups sent to the exact residual pass, [2] groups redone by the checked accumulation,
// [3] groups above kHugeRows, [4] groups re-fitted by the QR fallback.
int32_t anofox_b200_debug_counters(const void *workspace, int64_t n_groups, int64_t n_rows, int32_t n_features,
                                   int32_t *out) {
	if (!workspace || !out || n_features <= 0 || n_features > kMaxFeatures) return ANOFOX_ERROR_INVALID_INPUT;
	DeviceInfo &di = device_info();
	Workspace ws;
	carve(ws, (unsigned char *)const_cast<void *>(workspace), n_groups, n_rows, n_features, di.sms > 0 ? di.sms : 148);
	int h[192];
	if (cudaDeviceSynchronize() != cudaSuccess) return ANOFOX_ERROR_INTERNAL;
	if (cudaMemcpy(h, ws.big_count, sizeof h, cudaMemcpyDeviceToHost) != cudaSuccess) return ANOFOX_ERROR_INTERNAL;
	out[0] = h[0];
	out[1] = h[64];
	out[2] = h[128];
	out[3] = h[144];
	out[4] = h[160];
	return 0;
}
// Experiments: the whole counter block (192 ints) of the last call that used `workspace`.
int32_t anofox_b200_debug_raw(const void *workspace, int64_t n_groups, int64_t n_rows, int32_t n_features, int32_t *out192) {
	if (!workspace || !out192 || n_features <= 0 || n_features > kMaxFeatures) return ANOFOX_ERROR_INVALID_INPUT;
	DeviceInfo &di = device_info();
	Workspace ws;
	carve(ws, (unsigned char *)const_cast<void *>(workspace), n_groups, n_rows, n_features, di.sms > 0 ? di.sms : 148);
	if (cudaDeviceSynchronize() != cudaSuccess) return ANOFOX_ERROR_INTERNAL;
	if (cudaMemcpy(out192, ws.big_count, 768, cudaMemcpyDeviceToHost) != cudaSuccess) return ANOFOX_ERROR_INTERNAL;
	return 0;
}
// Which path a batch of this shape takes: 1 = fused small-group kernel (lanes_per_group set), 0 = record path.
int32_t anofox_b200_plan(int64_t n_groups, int64_t n_rows, int32_t n_features, int32_t has_w, int32_t has_train,
                         int32_t *lanes_per_group) {
	const SmallGeom sm = small_geometry(n_features, has_w != 0, has_train != 0, n_rows, n_groups);
	if (lanes_per_group) *lanes_per_group = sm.use ? sm.T : 0;
	// has_train == 2: a fit-predict call; 3 = the resident one-pass kernel takes it
	if (has_train == 2 && !sm.use && resident_geometry(n_features, has_w != 0, n_rows, n_groups).use) return 3;
	return sm.use ? (sm.stash ? 2 : 1) : 0;
}

// Test hook: evaluate the device special functions on arrays (host pointers).
int32_t anofox_b200_eval_specfun(int32_t mode, int32_t n, const double *a, const double *b, const double *c,
                                 double *out, AnofoxError *err) {
	return eval_specfun_host(mode, n, a, b, c, out, err);
}

} // extern "C"
