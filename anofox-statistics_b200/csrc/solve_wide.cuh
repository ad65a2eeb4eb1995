// solve_wide.cuh -- solve + inference of ONE group by one thread block, for 33 <= p <= 64 (and for any p
// when only a handful of groups exist: the row-sharded single group of BASELINE config 4).
//
// Same arithmetic and the same decisions as solve_group (solve.cuh) -- fit_ols / fit_ridge / fit_wls of the
// reference (crates/anofox-stats-core/src/models/{ols.rs:36-268, ridge.rs:36-229, wls.rs:37-289}), the FFI
// field conventions (lib.rs:164-253) and compute_aic / compute_bic (information_criteria.rs:15-87) -- but
// thread a of the block owns kept column a: the Gram rows, the Cholesky columns, the two triangular solves
// and the p unit right-hand sides of diag((L L^T)^-1) all run kWideThreads wide out of shared memory instead
// of serially out of local memory (p = 64: 33 ms -> tens of microseconds).
#pragma once
#include "solve.cuh"

namespace b200 {

constexpr int kWideThreads = 64; // = kMaxFeatures: one thread per kept column
constexpr int kWideLd = kWideThreads + 1; // row stride of the factor in shared memory (bank-conflict free)

struct WideSmem {
	double L[kWideThreads * kWideLd]; // Gram, then the Cholesky factor (lower triangle, row-major)
	double X[kWideThreads * kWideThreads]; // X[c * 64 + j]: row c of the solution of L x = e_j
	double b[kWideThreads], xm[kWideThreads], s1[kWideThreads], orig[kWideThreads], v[kWideThreads];
	double red[kWideThreads];
	int idx[kWideThreads];
	int dead[kWideThreads];
	int pe;
	int suspicious; // a Cholesky pivot fell below the QR-fallback threshold (common.cuh)
};

__device__ __forceinline__ double wide_sum(WideSmem &S, double mine, int tid) {
	// fixed-order sum over the block (thread 0 adds in index order: deterministic, p <= 64)
	S.red[tid] = mine;
	__syncthreads();
	double t = 0.0;
	for (int i = 0; i < kWideThreads; ++i) t += S.red[i];
	__syncthreads();
	return t;
}

// Must be called by all kWideThreads threads of the block with identical arguments.
__device__ void solve_group_wide(WideSmem &S, const double *__restrict__ rec, int64_t rs, int64_t g, const FitParams &P,
                                 const AnofoxB200FitOutputs &o, double rss_override, int *refine_count,
                                 int64_t *refine_list) {
	const int tid = threadIdx.x;
	const int p = P.p, k = p + 1;
	const double nanv = nan_f64();
	const double n_valid = rec[0], sw = rec[rs], n_state = rec[2 * rs];
	const double *shift = rec + rec_off_shift(p) * rs;
	const double *cdev = rec + rec_off_cdev(p) * rs;
	const double *S1 = rec + rec_off_s1(p) * rs;
	const double *S2 = rec + rec_off_s2(p) * rs;
	const int ic = P.fit_intercept ? 1 : 0;
	const bool wls = P.method == ANOFOX_B200_WLS;
	int status = 0;
	if (n_state < (double)P.min_state_rows)
		status = ANOFOX_B200_STATUS_TOO_FEW_ROWS; // ols_aggregate.cpp:263-267
	else if (!(n_valid > 0.0))
		status = ANOFOX_ERROR_NO_VALID_DATA; // ols.rs:68-70
	if (tid == 0) {
		int pe = 0;
		if (status == 0)
			for (int j = 0; j < p; ++j)
				if (!(cdev[j * rs] < 1e-10)) S.idx[pe++] = j; // ols.rs:76-87
		S.pe = pe;
	}
	__syncthreads();
	const int pe = S.pe;
	if (status == 0) {
		if (pe == 0 && !ic) status = ANOFOX_ERROR_INSUFFICIENT_DATA;                              // ols.rs:102-107
		else if (pe > 0 && n_valid < (double)(pe + ic)) status = ANOFOX_ERROR_INSUFFICIENT_DATA; // ols.rs:134-139
	}
	if (status != 0) {
		if (tid == 0) write_null_group(o, P, g, status);
		return;
	}
	const double inv_sw = 1.0 / sw;
	const double s1y = S1[p * rs];
	const double ybar = shift[p * rs] + s1y * inv_sw;
	const double syy_c = S2[tri_idx(p, p, k) * rs] - s1y * s1y * inv_sw;
	const double tss = ic ? syy_c : syy_c + sw * ybar * ybar; // R's summary.lm; no_intercept.json
	const bool inf_arrays = P.compute_inference && o.std_errors;
	for (int j = tid; j < p; j += kWideThreads) {
		o.coefficients[g * p + j] = nanv;
		if (inf_arrays) {
			o.std_errors[g * p + j] = nanv;
			o.t_values[g * p + j] = nanv;
			o.p_values[g * p + j] = nanv;
			o.ci_lower[g * p + j] = nanv;
			o.ci_upper[g * p + j] = nanv;
		}
	}
	if (tid == 0) {
		o.status[g] = 0;
		o.n_observations[g] = (int64_t)n_valid;
		if (inf_arrays) {
			o.f_statistic[g] = nanv;
			o.f_pvalue[g] = nanv;
		}
	}
	if (pe == 0) {
		// all columns constant: intercept-only shortcut (ols.rs:108-130, wls.rs:126-150)
		if (tid == 0) {
			const double ss = fmax(syy_c, 0.0);
			const double var = wls ? ss * inv_sw : ss / (n_valid - 1.0);
			o.intercept[g] = ybar;
			o.r_squared[g] = 0.0;
			o.adj_r_squared[g] = 0.0;
			o.residual_std_error[g] = sqrt(var);
			if (inf_arrays && o.has_inference) o.has_inference[g] = 0; // lib.rs:251-253
			double aic, bic;
			info_criteria_dev(ss, n_valid, 1.0, aic, bic);
			if (o.aic) o.aic[g] = aic;
			if (o.bic) o.bic[g] = bic;
			if (o.rss) o.rss[g] = ss;
		}
		return;
	}

	// ---- Gram matrix of the kept columns: thread a builds row a ------------------------------
	const int a = tid;
	const bool act = a < pe;
	const int ja = act ? S.idx[a] : 0;
	double s1a = 0.0, ma = 0.0;
	if (act) {
		s1a = S1[ja * rs];
		ma = shift[ja * rs] + s1a * inv_sw;
		S.s1[a] = s1a;
		S.xm[a] = ma;
	}
	__syncthreads();
	double lambda = 0.0;
	if (P.method == ANOFOX_B200_RIDGE) {
		lambda = P.alpha;
		if (P.lambda_scaling == ANOFOX_LAMBDA_SCALING_GLMNET) {
			const double sdy = sqrt(fmax(tss, 0.0) / n_valid);
			lambda = P.alpha * n_valid;
			if (sdy > 0.0) lambda /= sdy;
		}
	}
	const bool ridge = lambda > 0.0;
	if (act) {
		for (int c = 0; c <= a; ++c) {
			const int jc = S.idx[c];
			double v = S2[tri_idx(jc, ja, k) * rs] - s1a * S.s1[c] * inv_sw; // centred
			if (!ic) v += sw * ma * S.xm[c];                                  // raw second moment
			if (ridge && c == a) v += lambda;
			S.L[a * kWideLd + c] = v;
		}
		double cxy = S2[tri_idx(ja, p, k) * rs] - s1a * s1y * inv_sw;
		if (!ic) cxy += sw * ma * ybar;
		S.b[a] = cxy;
		const double caa = S2[tri_idx(ja, ja, k) * rs] - s1a * s1a * inv_sw;
		S.orig[a] = caa + sw * ma * ma;
	}
	__syncthreads();

	// ---- Cholesky, column by column, limited pivoting (same sums, in the same order, as solve_group) ----
	if (tid == 0) S.suspicious = 0;
	for (int j = 0; j < pe; ++j) {
		double s = 0.0;
		if (act && a >= j) {
			s = S.L[a * kWideLd + j];
			for (int c = 0; c < j; ++c) s -= S.L[a * kWideLd + c] * S.L[j * kWideLd + c];
		}
		__syncthreads(); // row j (read by all) is about to be rewritten by thread j
		if (a == j) {
			S.red[0] = s;
			// thread j owns both the remaining pivot and the diagonal it started from: the ill-conditioning probe
			if (rss_override < 0.0 && chol_pivot_suspicious(s, S.L[j * kWideLd + j], S.orig[j], ridge)) S.suspicious = 1;
		}
		__syncthreads();
		const double d = S.red[0];
		const bool alias = ridge ? !(d > 0.0) : !(d >= 1e-14 * S.orig[j]) || !(d > 0.0);
		if (alias) {
			if (act && a >= j) S.L[a * kWideLd + j] = 0.0; // column j
			if (a == j) {
				for (int c = 0; c < j; ++c) S.L[j * kWideLd + c] = 0.0; // row j
				S.dead[j] = 1;
			}
		} else {
			const double dj = sqrt(d);
			if (a == j) {
				S.L[j * kWideLd + j] = dj;
				S.dead[j] = 0;
			} else if (act && a > j) {
				S.L[a * kWideLd + j] = s / dj;
			}
		}
		__syncthreads();
	}
	if (tid == 0 && S.suspicious) queue_qr(P, g);
	const bool live = act && !S.dead[a];
	int rank = 0;
	for (int i = 0; i < pe; ++i) rank += S.dead[i] ? 0 : 1;

	// ---- forward solve L v = b (column sweep: row i subtracts c = 0 .. i-1 in order) ----------
	double bi = act ? S.b[a] : 0.0;
	for (int j = 0; j < pe; ++j) {
		if (a == j) S.v[j] = live ? bi / S.L[j * kWideLd + j] : 0.0;
		__syncthreads();
		if (act && a > j) bi -= S.L[a * kWideLd + j] * S.v[j];
	}
	__syncthreads();
	const double va = act ? S.v[a] : 0.0;
	const double ess = wide_sum(S, va * va, tid); // explained sum of squares |v|^2
	// ---- back solve L^T beta = v ---------------------------------------------------------------
	double sa = va;
	for (int i = pe - 1; i >= 0; --i) {
		if (a == i) S.b[i] = live ? sa / S.L[i * kWideLd + i] : 0.0;
		__syncthreads();
		if (act && a < i) sa -= S.L[i * kWideLd + a] * S.b[i];
	}
	__syncthreads();
	const double beta = act ? S.b[a] : 0.0;
	const double bnorm2 = wide_sum(S, live ? beta * beta : 0.0, tid);
	const double b0 = ybar - wide_sum(S, live ? S.xm[a] * beta : 0.0, tid);
	double rss = tss - ess - (ridge ? lambda * bnorm2 : 0.0);
	if (refine_count && rss_override < 0.0 && tid == 0) {
		if (!(rss > 1e-5 * tss)) { // cancellation guard, as in solve_group
			const int slot = atomicAdd(refine_count, 1);
			refine_list[slot] = g;
		}
	}
	if (rss_override >= 0.0) rss = rss_override;
	if (!(rss > 0.0)) rss = 0.0;

	const int rank_total = rank + ic;
	const double dfr = n_valid - (double)rank_total;
	const double dfm = (double)rank;
	const double r2 = 1.0 - rss / tss;
	const double adj = dfr > 0.0 ? 1.0 - (1.0 - r2) * (n_valid - (double)ic) / dfr : nanv;
	const double sigma2 = dfr > 0.0 ? rss / dfr : nanv;
	const double rse = sqrt(sigma2);
	if (live) o.coefficients[g * p + ja] = beta;
	const DfTables T{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
	if (tid == 0) {
		o.intercept[g] = ic ? b0 : nanv; // ols.rs:172-176
		o.r_squared[g] = r2;
		o.adj_r_squared[g] = adj;
		o.residual_std_error[g] = rse;
		double aic, bic;
		info_criteria_dev(rss, n_valid, (double)rank_total, aic, bic);
		if (o.aic) o.aic[g] = aic;
		if (o.bic) o.bic[g] = bic;
		if (o.rss) o.rss[g] = rss;
		if (inf_arrays) {
			if (o.has_inference) o.has_inference[g] = 1;
			if (dfm > 0.0 && dfr > 0.0) {
				const double f = ((tss - rss) / dfm) / sigma2;
				o.f_statistic[g] = f;
				o.f_pvalue[g] = f_upper_p_fast(f, dfm, dfr, T);
			}
		}
	}
	if (!inf_arrays) return;

	// ---- diag((L L^T)^-1)_j = sum_i (L^-1)_ij^2: thread j solves L x = e_j ----------------------
	// All threads walk the rows i together (L[i][c] is a broadcast read); x lives in X[c * 64 + j].
	double acc = 0.0;
	for (int i = 0; i < pe; ++i) {
		double xi = 0.0;
		if (live && i >= a && !S.dead[i]) {
			double s = (i == a) ? 1.0 : 0.0;
			for (int c = a; c < i; ++c) s -= S.L[i * kWideLd + c] * S.X[c * kWideThreads + a];
			xi = s / S.L[i * kWideLd + i];
			acc += xi * xi;
		}
		S.X[i * kWideThreads + a] = xi;
	}
	if (live) {
		double tcrit = nanv;
		if (dfr > 0.0) {
			const int idf = (int)dfr;
			tcrit = (P.tcrit && idf <= kTcritTable) ? P.tcrit[idf] : t_critical_dev(P.level, dfr);
		}
		const double se = sqrt(sigma2 * acc);
		const double t = beta / se;
		const int64_t oj = g * p + ja;
		o.std_errors[oj] = se;
		o.t_values[oj] = t;
		o.p_values[oj] = dfr > 0.0 ? t_two_sided_p_fast(t, dfr, T) : nanv;
		o.ci_lower[oj] = beta - tcrit * se;
		o.ci_upper[oj] = beta + tcrit * se;
	}
}

// One block of kWideThreads threads per group; list != null: the groups list[0 .. n_list).
static __global__ void __launch_bounds__(kWideThreads) solve_wide_kernel(const double *__restrict__ records, const int64_t *__restrict__ list,
                                                                         int64_t n_groups, FitParams P, AnofoxB200FitOutputs o,
                                                                         int *refine_count, int64_t *refine_list) {
	extern __shared__ __align__(16) unsigned char wide_smem_raw[];
	WideSmem &S = *reinterpret_cast<WideSmem *>(wide_smem_raw);
	for (int64_t i = blockIdx.x; i < n_groups; i += gridDim.x) {
		const int64_t g = list ? list[i] : i;
		solve_group_wide(S, records + g * rec_len(P.p), 1, g, P, o, -1.0, refine_count, refine_list);
		__syncthreads();
	}
}

} // namespace b200
