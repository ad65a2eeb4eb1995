// solve.cuh -- per-group solve + inference from a moment record (see common.cuh).
//
// Replaces, for one group, everything the reference does after its row buffers are full:
// fit_ols / fit_ridge / fit_wls (crates/anofox-stats-core/src/models/{ols.rs:36-268,
// ridge.rs:36-229,wls.rs:37-289}) including the external regressor call (ols.rs:155-161), the
// NaN scatter-back (ols.rs:166-171) and the FFI field conventions (lib.rs:164-253), plus
// compute_aic / compute_bic (information_criteria.rs:15-87).
//
// Algorithm: the centred (or, without intercept, raw) weighted Gram matrix of the kept columns is
// factorised by a left-to-right Cholesky with LIMITED pivoting: a column whose remaining pivot
// falls below (1e-7)^2 times its original squared norm is aliased (NaN coefficient) and skipped,
// which is the Gram-form of LINPACK dqrdc2 as used by R's lm() -- the convention the reference's
// golden vectors pin (perfect_collinearity.json, rank_deficient.json: the LATER column is NA).
// Ridge adds lambda to the diagonal and never aliases.
#pragma once
#include "common.cuh"
#include "specfun.cuh"

namespace b200 {

__device__ __forceinline__ int lt_idx(int i, int j) { return (i * (i + 1)) / 2 + j; } // j <= i

struct GroupStats { // scalar results of one solve, kept for the predict path
	double intercept, rse;
	int64_t n_obs;
	int status;
};

__device__ inline void write_null_group(const AnofoxB200FitOutputs &o, const FitParams &P, int64_t g, int status) {
	const double nanv = nan_f64();
	const int p = P.p;
	for (int j = 0; j < p; ++j) o.coefficients[g * p + j] = nanv;
	o.intercept[g] = nanv;
	o.r_squared[g] = nanv;
	o.adj_r_squared[g] = nanv;
	o.residual_std_error[g] = nanv;
	o.n_observations[g] = 0;
	o.status[g] = status;
	if (P.compute_inference && o.std_errors) {
		for (int j = 0; j < p; ++j) {
			o.std_errors[g * p + j] = nanv;
			o.t_values[g * p + j] = nanv;
			o.p_values[g * p + j] = nanv;
			o.ci_lower[g * p + j] = nanv;
			o.ci_upper[g * p + j] = nanv;
		}
		o.f_statistic[g] = nanv;
		o.f_pvalue[g] = nanv;
		if (o.has_inference) o.has_inference[g] = 0;
	}
	if (o.aic) o.aic[g] = nanv;
	if (o.bic) o.bic[g] = nanv;
	if (o.rss) o.rss[g] = nanv;
}

__device__ inline void info_criteria_dev(double rss, double n, double k, double &aic, double &bic) {
	if (!(n > 0.0) || !(rss >= 0.0)) {
		aic = bic = nan_f64();
	} else if (rss == 0.0) {
		aic = bic = -__longlong_as_double(0x7ff0000000000000LL);
	} else {
		double l = n * log(rss / n);
		aic = l + 2.0 * k;
		bic = l + k * log(n);
	}
}

// rec: the group's record; rs: stride (in doubles) between consecutive record elements
// (1 for an array-of-records layout, G for a transposed one).
// rss_override: >= 0 replaces the Gram-derived residual sum of squares (refinement pass).
template <int PMAX>
__device__ void solve_group(const double *__restrict__ rec, int64_t rs, int64_t g, const FitParams &P,
                            const AnofoxB200FitOutputs &o, double rss_override, GroupStats *stats_out,
                            int *refine_count, int64_t *refine_list) {
	const int p = P.p, k = p + 1;
	const double nanv = nan_f64();
	const double n_valid = rec[0], sw = rec[rs], n_state = rec[2 * rs];
	const double *shift = rec + rec_off_shift(p) * rs;
	const double *cdev = rec + rec_off_cdev(p) * rs;
	const double *S1 = rec + rec_off_s1(p) * rs;
	const double *S2 = rec + rec_off_s2(p) * rs;
	const int ic = P.fit_intercept ? 1 : 0;
	const bool wls = P.method == ANOFOX_B200_WLS;
	if (stats_out) {
		stats_out->status = -1;
		stats_out->intercept = nanv;
		stats_out->rse = nanv;
		stats_out->n_obs = 0;
	}
	int status = 0;
	if (n_state < (double)P.min_state_rows)
		status = ANOFOX_B200_STATUS_TOO_FEW_ROWS; // ols_aggregate.cpp:263-267
	else if (!(n_valid > 0.0))
		status = ANOFOX_ERROR_NO_VALID_DATA; // ols.rs:68-70

	int idx[PMAX];
	int pe = 0;
	if (status == 0) {
		for (int j = 0; j < p; ++j)
			if (!(cdev[j * rs] < 1e-10)) idx[pe++] = j; // ols.rs:76-87
		if (pe == 0 && !ic) status = ANOFOX_ERROR_INSUFFICIENT_DATA;          // ols.rs:102-107
		else if (pe > 0 && n_valid < (double)(pe + ic)) status = ANOFOX_ERROR_INSUFFICIENT_DATA; // ols.rs:134-139
	}
	if (status != 0) {
		write_null_group(o, P, g, status);
		if (stats_out) stats_out->status = status;
		return;
	}

	const double inv_sw = 1.0 / sw;
	const double s1y = S1[p * rs];
	const double ybar = shift[p * rs] + s1y * inv_sw;
	const double syy_c = S2[tri_idx(p, p, k) * rs] - s1y * s1y * inv_sw; // centred sum of squares of y
	// total sum of squares: centred with intercept, raw without (R's summary.lm; no_intercept.json)
	double tss;
	if (ic)
		tss = syy_c;
	else
		tss = syy_c + sw * ybar * ybar;

	o.status[g] = 0;
	o.n_observations[g] = (int64_t)n_valid;
	for (int j = 0; j < p; ++j) o.coefficients[g * p + j] = nanv;
	const bool inf_arrays = P.compute_inference && o.std_errors;
	if (inf_arrays) {
		for (int j = 0; j < p; ++j) {
			o.std_errors[g * p + j] = nanv;
			o.t_values[g * p + j] = nanv;
			o.p_values[g * p + j] = nanv;
			o.ci_lower[g * p + j] = nanv;
			o.ci_upper[g * p + j] = nanv;
		}
		o.f_statistic[g] = nanv;
		o.f_pvalue[g] = nanv;
	}

	if (pe == 0) {
		// all columns constant: intercept-only shortcut (ols.rs:108-130, wls.rs:126-150)
		double ss = fmax(syy_c, 0.0);
		double var = wls ? ss * inv_sw : ss / (n_valid - 1.0);
		double rse = sqrt(var);
		o.intercept[g] = ybar;
		o.r_squared[g] = 0.0;
		o.adj_r_squared[g] = 0.0;
		o.residual_std_error[g] = rse;
		if (inf_arrays && o.has_inference) o.has_inference[g] = 0; // lib.rs:251-253
		double aic, bic;
		info_criteria_dev(ss, n_valid, 1.0, aic, bic);
		if (o.aic) o.aic[g] = aic;
		if (o.bic) o.bic[g] = bic;
		if (o.rss) o.rss[g] = ss;
		if (stats_out) {
			stats_out->status = 0;
			stats_out->intercept = ybar;
			stats_out->rse = rse;
			stats_out->n_obs = (int64_t)n_valid;
		}
		return;
	}

	// ---- Gram matrix of the kept columns -------------------------------------------------
	double A[PMAX * (PMAX + 1) / 2]; // lower triangle, row-major; overwritten by the Cholesky factor
	double b[PMAX];                  // X'Wy -> forward solve -> coefficients
	double orig[PMAX];               // uncentred squared column norms (aliasing reference)
	double xm[PMAX];                 // column means
	for (int a = 0; a < pe; ++a) {
		int ja = idx[a];
		double s1a = S1[ja * rs];
		double sha = shift[ja * rs];
		double ma = sha + s1a * inv_sw;
		xm[a] = ma;
		for (int c = 0; c <= a; ++c) {
			int jc = idx[c];
			double s1c = S1[jc * rs];
			double v = S2[tri_idx(jc, ja, k) * rs] - s1a * s1c * inv_sw; // centred
			if (!ic) v += sw * ma * xm[c];                               // raw second moment
			A[lt_idx(a, c)] = v;
		}
		double cxy = S2[tri_idx(ja, p, k) * rs] - s1a * s1y * inv_sw;
		if (!ic) cxy += sw * ma * ybar;
		b[a] = cxy;
		double caa = S2[tri_idx(ja, ja, k) * rs] - s1a * s1a * inv_sw;
		orig[a] = caa + sw * ma * ma;
	}

	double lambda = 0.0;
	if (P.method == ANOFOX_B200_RIDGE) {
		lambda = P.alpha;
		if (P.lambda_scaling == ANOFOX_LAMBDA_SCALING_GLMNET) {
			// glmnet's penalty on its unit-variance response == raw lambda * n / sd_y (oracle header)
			double sdy = sqrt(fmax(tss, 0.0) / n_valid);
			lambda = P.alpha * n_valid;
			if (sdy > 0.0) lambda /= sdy;
		}
	}
	const bool ridge = lambda > 0.0;
	if (ridge)
		for (int a = 0; a < pe; ++a) A[lt_idx(a, a)] += lambda;

	// ---- Cholesky, left to right, limited pivoting -----------------------------------------
	bool dead[PMAX];
	int rank = 0;
	bool suspicious = false;
	for (int i = 0; i < pe; ++i) {
		const double a0 = A[lt_idx(i, i)];
		for (int j = 0; j < i; ++j) {
			if (dead[j]) {
				A[lt_idx(i, j)] = 0.0;
				continue;
			}
			double s = A[lt_idx(i, j)];
			for (int c = 0; c < j; ++c) s -= A[lt_idx(i, c)] * A[lt_idx(j, c)];
			A[lt_idx(i, j)] = s / A[lt_idx(j, j)];
		}
		double d = A[lt_idx(i, i)];
		for (int c = 0; c < i; ++c) d -= A[lt_idx(i, c)] * A[lt_idx(i, c)];
		bool alias = ridge ? !(d > 0.0) : !(d >= 1e-14 * orig[i]) || !(d > 0.0);
		suspicious = suspicious || chol_pivot_suspicious(d, a0, orig[i], ridge);
		dead[i] = alias;
		if (alias) {
			for (int c = 0; c <= i; ++c) A[lt_idx(i, c)] = 0.0;
		} else {
			A[lt_idx(i, i)] = sqrt(d);
			++rank;
		}
	}
	if (suspicious && rss_override < 0.0) queue_qr(P, g);
	// forward solve L v = b
	double ess = 0.0; // explained sum of squares |v|^2
	for (int i = 0; i < pe; ++i) {
		if (dead[i]) {
			b[i] = 0.0;
			continue;
		}
		double s = b[i];
		for (int c = 0; c < i; ++c) s -= A[lt_idx(i, c)] * b[c];
		s /= A[lt_idx(i, i)];
		b[i] = s;
		ess += s * s;
	}
	// back solve L^T beta = v
	for (int i = pe - 1; i >= 0; --i) {
		if (dead[i]) continue;
		double s = b[i];
		for (int c = i + 1; c < pe; ++c)
			if (!dead[c]) s -= A[lt_idx(c, i)] * b[c];
		b[i] = s / A[lt_idx(i, i)];
	}
	double bnorm2 = 0.0, b0 = ybar;
	for (int i = 0; i < pe; ++i)
		if (!dead[i]) {
			bnorm2 += b[i] * b[i];
			b0 -= xm[i] * b[i];
		}
	// residual sum of squares from the moments:  tss - |v|^2  (- lambda |beta|^2 for ridge, where
	// beta' C beta = beta' b - lambda |beta|^2).
	double rss = tss - ess - (ridge ? lambda * bnorm2 : 0.0);
	if (refine_count && rss_override < 0.0) {
		// cancellation guard: tss - ess keeps ~16 - log10(tss/rss) digits. Queue the group for the
		// exact residual pass (refine_kernel) when fewer than ~10 would survive.
		if (!(rss > 1e-5 * tss)) {
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
	o.intercept[g] = ic ? b0 : nanv; // ols.rs:172-176
	o.r_squared[g] = r2;
	o.adj_r_squared[g] = adj;
	o.residual_std_error[g] = rse;
	for (int i = 0; i < pe; ++i)
		if (!dead[i]) o.coefficients[g * p + idx[i]] = b[i];
	double aic, bic;
	info_criteria_dev(rss, n_valid, (double)rank_total, aic, bic);
	if (o.aic) o.aic[g] = aic;
	if (o.bic) o.bic[g] = bic;
	if (o.rss) o.rss[g] = rss;
	if (stats_out) {
		stats_out->status = 0;
		stats_out->intercept = ic ? b0 : nanv;
		stats_out->rse = rse;
		stats_out->n_obs = (int64_t)n_valid;
	}

	if (inf_arrays) {
		if (o.has_inference) o.has_inference[g] = 1;
		const DfTables T{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
		double tcrit = nanv;
		if (dfr > 0.0) {
			int idf = (int)dfr;
			tcrit = (P.tcrit && idf <= kTcritTable) ? P.tcrit[idf] : t_critical_dev(P.level, dfr);
		}
		// diag((L L^T)^-1)_j = sum_i (L^-1)_ij^2 : one unit right-hand side per kept column
		double x[PMAX];
		for (int j = 0; j < pe; ++j) {
			if (dead[j]) continue;
			double acc = 0.0;
			for (int i = j; i < pe; ++i) {
				if (dead[i]) {
					x[i] = 0.0;
					continue;
				}
				double s = (i == j) ? 1.0 : 0.0;
				for (int c = j; c < i; ++c) s -= A[lt_idx(i, c)] * x[c];
				s /= A[lt_idx(i, i)];
				x[i] = s;
				acc += s * s;
			}
			double se = sqrt(sigma2 * acc);
			double t = b[j] / se;
			int64_t oj = g * p + idx[j];
			o.std_errors[oj] = se;
			o.t_values[oj] = t;
			o.p_values[oj] = dfr > 0.0 ? t_two_sided_p_fast(t, dfr, T) : nanv;
			o.ci_lower[oj] = b[j] - tcrit * se;
			o.ci_upper[oj] = b[j] + tcrit * se;
		}
		if (dfm > 0.0 && dfr > 0.0) {
			double f = ((tss - rss) / dfm) / sigma2;
			o.f_statistic[g] = f;
			o.f_pvalue[g] = f_upper_p_fast(f, dfm, dfr, T);
		}
	}
}

} // namespace b200
