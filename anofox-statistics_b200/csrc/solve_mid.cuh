// solve_mid.cuh -- solve + inference stage for 9 <= p <= 32: one thread per group, every per-group array in
// SHARED memory instead of local memory.
//
// Same reference code as solve.cuh (fit_ols / fit_ridge / fit_wls after the row buffers are full:
// crates/anofox-stats-core/src/models/ols.rs:76-268, ridge.rs:36-229, wls.rs:37-289, the external regressor's
// statistics, lib.rs:164-253 field conventions, information_criteria.rs:15-87) and the same formulas, line
// for line, as solve_group. What changes is where the numbers live:
//   * a CTA of T threads takes T consecutive groups. Their row-wise moment records are one contiguous block
//     of T * L doubles in HBM: the CTA streams it with fully coalesced loads and transposes it into shared
//     memory as ws[field * (T + 1) + thread] (odd stride: conflict-free both for the transposing stores and
//     for the per-thread accesses afterwards).
//   * the solve then runs IN PLACE in the thread's column of that block: the centred Gram overwrites S2, the
//     Cholesky factor overwrites the Gram, X'y / the coefficients live in the (., y) column of S2, the column
//     means overwrite the shift row, the squared norms overwrite cdev, the unit right-hand sides of the
//     inference pass reuse S1. Constant columns are not compacted away, they start dead (as in small_solve).
//   * two CTAs fit an SM (p = 16: 2 x 109 KB), so one streams its records while the other factorises.
#pragma once
#include "solve.cuh"

namespace b200 {

template <int DUMMY = 0>
__global__ void __launch_bounds__(64) solve_mid_kernel(const double *__restrict__ records, int64_t n_groups, FitParams P,
                                                       AnofoxB200FitOutputs o, int *refine_count, int64_t *refine_list) {
	extern __shared__ __align__(16) double ws_all[];
	const int T = blockDim.x, TS = T + 1;
	const int p = P.p, k = p + 1, L = rec_len(p);
	const int64_t g0 = (int64_t)blockIdx.x * T;
	const int ng = (int)(n_groups - g0 < T ? n_groups - g0 : T);
	// ---- stage the CTA's records: element e of the block is field e % L of group e / L -------------------
	{
		// a warp stages one group at a time, lane l the fields l, l + 32, ...: coalesced reads of the record,
		// conflict-free transposed writes (field stride TS is odd); hundreds of 8-byte copies in flight per thread
		const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = (T + 31) >> 5;
		for (int grp = warp; grp < ng; grp += nwarps) {
			const double *src = records + (g0 + grp) * L;
			for (int f = lane; f < L; f += 32)
				asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(&ws_all[f * TS + grp])), "l"(src + f) : "memory");
		}
	}
	asm volatile("cp.async.wait_all;" ::: "memory");
	__syncthreads();
	if ((int)threadIdx.x >= ng) return;
	const int64_t g = g0 + threadIdx.x;
	double *ws = ws_all + threadIdx.x;
#define W(f) ws[(f) * TS]
	const int o_shift = rec_off_shift(p), o_cdev = rec_off_cdev(p), o_s1 = rec_off_s1(p), o_s2 = rec_off_s2(p);
#define SHIFT(j) W(o_shift + (j))
#define ORIG(j) W(o_cdev + (j))
#define S1(j) W(o_s1 + (j))
#define S2(i, j) W(o_s2 + tri_idx((i), (j), k)) /* i <= j */
#define LL(i, j) S2((j), (i))                    /* Cholesky factor, j <= i */
#define BB(i) S2((i), p)
	const double nanv = nan_f64();
	const double n_valid = W(0), sw = W(1), n_state = W(2);
	const int ic = P.fit_intercept ? 1 : 0;
	const bool wls = P.method == ANOFOX_B200_WLS;
	int status = 0;
	if (n_state < (double)P.min_state_rows)
		status = ANOFOX_B200_STATUS_TOO_FEW_ROWS; // ols_aggregate.cpp:263-267
	else if (!(n_valid > 0.0))
		status = ANOFOX_ERROR_NO_VALID_DATA; // ols.rs:68-70
	unsigned dead = 0u; // bit j: column j carries no coefficient (constant, or aliased below)
	int pe = 0;
	if (status == 0) {
		for (int j = 0; j < p; ++j) {
			if (!(W(o_cdev + j) < 1e-10))
				++pe; // ols.rs:76-87
			else
				dead |= 1u << j;
		}
		if (pe == 0 && !ic) status = ANOFOX_ERROR_INSUFFICIENT_DATA;                                // ols.rs:102-107
		else if (pe > 0 && n_valid < (double)(pe + ic)) status = ANOFOX_ERROR_INSUFFICIENT_DATA; // ols.rs:134-139
	}
	if (status != 0) {
		write_null_group(o, P, g, status);
		return;
	}
	const double inv_sw = 1.0 / sw;
	const double s1y = S1(p);
	const double ybar = SHIFT(p) + s1y * inv_sw;
	const double syy_c = S2(p, p) - s1y * s1y * inv_sw;
	const double tss = ic ? syy_c : syy_c + sw * ybar * ybar; // R's summary.lm; no_intercept.json
	o.status[g] = 0;
	o.n_observations[g] = (int64_t)n_valid;
	const bool inf_arrays = P.compute_inference && o.std_errors;
	if (pe == 0) {
		// all columns constant: intercept-only shortcut (ols.rs:108-130, wls.rs:126-150)
		const double ss = fmax(syy_c, 0.0);
		const double var = wls ? ss * inv_sw : ss / (n_valid - 1.0);
		for (int j = 0; j < p; ++j) o.coefficients[g * p + j] = nanv;
		o.intercept[g] = ybar;
		o.r_squared[g] = 0.0;
		o.adj_r_squared[g] = 0.0;
		o.residual_std_error[g] = sqrt(var);
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
			if (o.has_inference) o.has_inference[g] = 0; // lib.rs:251-253
		}
		double aic, bic;
		info_criteria_dev(ss, n_valid, 1.0, aic, bic);
		if (o.aic) o.aic[g] = aic;
		if (o.bic) o.bic[g] = bic;
		if (o.rss) o.rss[g] = ss;
		return;
	}

	// ---- centred (without intercept: raw) Gram, in place ------------------------------------------------
	for (int a = 0; a < p; ++a) {
		const double s1a = S1(a);
		const double ma = SHIFT(a) + s1a * inv_sw;
		SHIFT(a) = ma; // column mean from here on
		const double caa = S2(a, a) - s1a * s1a * inv_sw;
		ORIG(a) = caa + sw * ma * ma; // uncentred squared norm: the aliasing reference
		for (int c = 0; c <= a; ++c) {
			double v = S2(c, a) - s1a * S1(c) * inv_sw;
			if (!ic) v += sw * ma * SHIFT(c);
			S2(c, a) = v;
		}
		double cxy = S2(a, p) - s1a * s1y * inv_sw;
		if (!ic) cxy += sw * ma * ybar;
		S2(a, p) = cxy;
	}
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
	if (ridge)
		for (int a = 0; a < p; ++a) S2(a, a) += lambda;

	// ---- Cholesky, left to right, limited pivoting (solve.cuh) ---------------------------------------------
	int rank = 0;
	bool suspicious = false;
	for (int i = 0; i < p; ++i) {
		const bool dead_i = (dead >> i) & 1u;
		const double a0 = LL(i, i);
		for (int j = 0; j < i; ++j) {
			if (dead_i || ((dead >> j) & 1u)) {
				LL(i, j) = 0.0;
				continue;
			}
			double s = LL(i, j);
#pragma unroll 4
			for (int c = 0; c < j; ++c) s -= LL(i, c) * LL(j, c);
			LL(i, j) = s / LL(j, j);
		}
		if (dead_i) {
			LL(i, i) = 0.0;
			continue;
		}
		double d = LL(i, i);
#pragma unroll 4
		for (int c = 0; c < i; ++c) d -= LL(i, c) * LL(i, c);
		const bool alias = ridge ? !(d > 0.0) : !(d >= 1e-14 * ORIG(i)) || !(d > 0.0);
		suspicious = suspicious || chol_pivot_suspicious(d, a0, ORIG(i), ridge);
		if (alias) {
			dead |= 1u << i;
			for (int c = 0; c <= i; ++c) LL(i, c) = 0.0;
		} else {
			LL(i, i) = sqrt(d);
			++rank;
		}
	}
	if (suspicious) queue_qr(P, g);
	// forward solve L v = b, back solve L^T beta = v
	double ess = 0.0;
	for (int i = 0; i < p; ++i) {
		if ((dead >> i) & 1u) {
			BB(i) = 0.0;
			continue;
		}
		double s = BB(i);
#pragma unroll 4
		for (int c = 0; c < i; ++c) s -= LL(i, c) * BB(c);
		s /= LL(i, i);
		BB(i) = s;
		ess += s * s;
	}
	for (int i = p - 1; i >= 0; --i) {
		if ((dead >> i) & 1u) continue;
		double s = BB(i);
#pragma unroll 4
		for (int c = i + 1; c < p; ++c) s -= LL(c, i) * BB(c); // dead c: both factors are 0
		BB(i) = s / LL(i, i);
	}
	double bnorm2 = 0.0, b0 = ybar;
	for (int i = 0; i < p; ++i)
		if (!((dead >> i) & 1u)) {
			bnorm2 += BB(i) * BB(i);
			b0 -= SHIFT(i) * BB(i);
		}
	double rss = tss - ess - (ridge ? lambda * bnorm2 : 0.0);
	if (refine_count && !(rss > 1e-5 * tss)) { // cancellation guard: queue for the exact residual pass
		const int slot = atomicAdd(refine_count, 1);
		refine_list[slot] = g;
	}
	if (!(rss > 0.0)) rss = 0.0;

	const int rank_total = rank + ic;
	const double dfr = n_valid - (double)rank_total;
	const double dfm = (double)rank;
	const double r2 = 1.0 - rss / tss;
	const double adj = dfr > 0.0 ? 1.0 - (1.0 - r2) * (n_valid - (double)ic) / dfr : nanv;
	const double sigma2 = dfr > 0.0 ? rss / dfr : nanv;
	o.intercept[g] = ic ? b0 : nanv; // ols.rs:172-176
	o.r_squared[g] = r2;
	o.adj_r_squared[g] = adj;
	o.residual_std_error[g] = sqrt(sigma2);
	for (int i = 0; i < p; ++i) o.coefficients[g * p + i] = ((dead >> i) & 1u) ? nanv : BB(i);
	double aic, bic;
	info_criteria_dev(rss, n_valid, (double)rank_total, aic, bic);
	if (o.aic) o.aic[g] = aic;
	if (o.bic) o.bic[g] = bic;
	if (o.rss) o.rss[g] = rss;
	if (!inf_arrays) return;

	if (o.has_inference) o.has_inference[g] = 1;
	const DfTables Tb{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
	double tcrit = nanv;
	if (dfr > 0.0) {
		const int idf = (int)dfr;
		tcrit = (P.tcrit && idf <= kTcritTable) ? P.tcrit[idf] : t_critical_dev(P.level, dfr);
	}
	// diag((L L^T)^-1)_j = sum_i (L^-1)_ij^2 : one unit right-hand side per live column, kept in the S1 slots
	for (int j = 0; j < p; ++j) {
		const int64_t oj = g * p + j;
		if ((dead >> j) & 1u) {
			o.std_errors[oj] = nanv;
			o.t_values[oj] = nanv;
			o.p_values[oj] = nanv;
			o.ci_lower[oj] = nanv;
			o.ci_upper[oj] = nanv;
			continue;
		}
		double acc = 0.0;
		for (int i = j; i < p; ++i) {
			if ((dead >> i) & 1u) {
				S1(i) = 0.0;
				continue;
			}
			double s = (i == j) ? 1.0 : 0.0;
#pragma unroll 4
			for (int c = j; c < i; ++c) s -= LL(i, c) * S1(c);
			s /= LL(i, i);
			S1(i) = s;
			acc += s * s;
		}
		const double se = sqrt(sigma2 * acc);
		const double bj = BB(j);
		const double t = bj / se;
		o.std_errors[oj] = se;
		o.t_values[oj] = t;
		o.p_values[oj] = dfr > 0.0 ? pvalue_f(t * t, 1.0, dfr, Tb) : nanv;
		o.ci_lower[oj] = bj - tcrit * se;
		o.ci_upper[oj] = bj + tcrit * se;
	}
	double fs = nanv, fp = nanv;
	if (dfm > 0.0 && dfr > 0.0) {
		fs = ((tss - rss) / dfm) / sigma2;
		fp = pvalue_f(fs, dfm, dfr, Tb);
	}
	o.f_statistic[g] = fs;
	o.f_pvalue[g] = fp;
#undef W
#undef SHIFT
#undef ORIG
#undef S1
#undef S2
#undef LL
#undef BB
}

// ---------------------------------------------------------------------------------------------------------
// Fixed-layout variant for p <= PM (PM = 16): the staging pass re-packs every record into a layout that does
// not depend on p -- header | shift[PM+1] | cdev/orig[PM] | S1[PM+1] | S2 packed for K = PM + 1, with y always
// at index PM and the unused columns zero (a zero cdev makes them constant = dead from the start) -- so that
// the whole solve is unrolled with compile-time shared-memory offsets: two loads and one DFMA per inner
// iteration instead of a dozen instructions of index arithmetic. One thread per group, 64 groups per CTA.
// ---------------------------------------------------------------------------------------------------------
constexpr int kFixT = 64, kFixTS = kFixT + 1;
__host__ __device__ constexpr int fx_utri(int i, int j, int k) { return i * k - (i * (i - 1)) / 2 + (j - i); } // i <= j
template <int PM> struct FixLayout {
	static constexpr int K = PM + 1;
	static constexpr int SHIFT = 4, ORIG = SHIFT + K, S1 = ORIG + PM, S2 = S1 + K, LEN = S2 + K * (K + 1) / 2;
};

template <int PM>
__global__ void __launch_bounds__(kFixT) solve_fixed_kernel(const double *__restrict__ records, int64_t n_groups, FitParams P,
                                                            AnofoxB200FitOutputs o, int *refine_count, int64_t *refine_list) {
	using FL = FixLayout<PM>;
	constexpr int K = FL::K, TS = kFixTS;
	extern __shared__ __align__(16) double ws_all[];
	__shared__ short fmap[4 + 3 * (PM + 1) + 4 + (PM + 1) * (PM + 2) / 2];
	const int p = P.p, k = p + 1, L = rec_len(p);
	const int64_t g0 = (int64_t)blockIdx.x * kFixT;
	const int ng = (int)(n_groups - g0 < kFixT ? n_groups - g0 : kFixT);
	// ---- field map: record field (layout of common.cuh for this p) -> fixed field; -1 = padding ------------
	for (int f = threadIdx.x; f < rec_off_s2(p); f += kFixT) {
		int m = -1;
		auto col = [&](int c) { return c < p ? c : PM; };
		if (f < 4)
			m = f;
		else if (f < rec_off_cdev(p))
			m = FL::SHIFT + col(f - rec_off_shift(p));
		else if (f < rec_off_cdev(p) + p)
			m = FL::ORIG + (f - rec_off_cdev(p));
		else if (f >= rec_off_s1(p) && f < rec_off_s2(p))
			m = FL::S1 + col(f - rec_off_s1(p));
		fmap[f] = (short)m;
	}
	for (int i = threadIdx.x; i <= p; i += kFixT) {
		const int mi = i < p ? i : PM;
		for (int j = i; j <= p; ++j) fmap[rec_off_s2(p) + tri_idx(i, j, k)] = (short)(FL::S2 + fx_utri(mi, j < p ? j : PM, K));
	}
	for (int e = threadIdx.x; e < FL::LEN * TS; e += kFixT) ws_all[e] = 0.0;
	__syncthreads();
	{
		// a warp stages one group at a time, lane l the fields l, l + 32, ...: coalesced 256-byte reads of the
		// record, conflict-free transposed writes (field stride TS = 65 doubles), three address updates per element
		const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
		for (int grp = warp; grp < ng; grp += kFixT / 32) {
			const double *src = records + (g0 + grp) * L;
			for (int f = lane; f < L; f += 32) {
				const int m = fmap[f];
				if (m >= 0)
					asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(&ws_all[m * TS + grp])), "l"(src + f) : "memory");
			}
		}
	}
	asm volatile("cp.async.wait_all;" ::: "memory");
	__syncthreads();
	if ((int)threadIdx.x >= ng) return;
	const int64_t g = g0 + threadIdx.x;
	double *ws = ws_all + threadIdx.x;
#define FX(f) ws[(f) * TS]
#define FSHIFT(j) FX(FL::SHIFT + (j))
#define FORIG(j) FX(FL::ORIG + (j))
#define FS1(j) FX(FL::S1 + (j))
#define FU(i, j) FX(FL::S2 + fx_utri((i), (j), K)) /* i <= j */
	const double nanv = nan_f64();
	const double n_valid = FX(0), sw = FX(1), n_state = FX(2);
	const int ic = P.fit_intercept ? 1 : 0;
	const bool wls = P.method == ANOFOX_B200_WLS;
	int status = 0;
	if (n_state < (double)P.min_state_rows)
		status = ANOFOX_B200_STATUS_TOO_FEW_ROWS; // ols_aggregate.cpp:263-267
	else if (!(n_valid > 0.0))
		status = ANOFOX_ERROR_NO_VALID_DATA; // ols.rs:68-70
	unsigned dead = 0u;
#pragma unroll
	for (int c = 0; c < PM; ++c)
		if (FORIG(c) < 1e-10) dead |= 1u << c; // ols.rs:76-87 (and the unused columns c >= p)
	const int pe = PM - __popc(dead);
	if (status == 0) {
		if (pe == 0 && !ic) status = ANOFOX_ERROR_INSUFFICIENT_DATA;                                // ols.rs:102-107
		else if (pe > 0 && n_valid < (double)(pe + ic)) status = ANOFOX_ERROR_INSUFFICIENT_DATA; // ols.rs:134-139
	}
	if (status != 0) {
		write_null_group(o, P, g, status);
		return;
	}
	const double inv_sw = 1.0 / sw;
	const double s1y = FS1(PM);
	const double ybar = FSHIFT(PM) + s1y * inv_sw;
	const double syy_c = FU(PM, PM) - s1y * s1y * inv_sw;
	const double tss = ic ? syy_c : syy_c + sw * ybar * ybar;
	o.status[g] = 0;
	o.n_observations[g] = (int64_t)n_valid;
	const bool inf_arrays = P.compute_inference && o.std_errors;
	if (pe == 0) {
		// all columns constant: intercept-only shortcut (ols.rs:108-130, wls.rs:126-150)
		const double ss = fmax(syy_c, 0.0);
		const double var = wls ? ss * inv_sw : ss / (n_valid - 1.0);
		for (int j = 0; j < p; ++j) o.coefficients[g * p + j] = nanv;
		o.intercept[g] = ybar;
		o.r_squared[g] = 0.0;
		o.adj_r_squared[g] = 0.0;
		o.residual_std_error[g] = sqrt(var);
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
			if (o.has_inference) o.has_inference[g] = 0; // lib.rs:251-253
		}
		double aic, bic;
		info_criteria_dev(ss, n_valid, 1.0, aic, bic);
		if (o.aic) o.aic[g] = aic;
		if (o.bic) o.bic[g] = bic;
		if (o.rss) o.rss[g] = ss;
		return;
	}
	// ---- centred (without intercept: raw) Gram, in place; unused columns (a >= p) stay zero ---------------
#pragma unroll
	for (int a = 0; a < PM; ++a) {
		if (a < p) {
			const double s1a = FS1(a);
			const double ma = FSHIFT(a) + s1a * inv_sw;
			FSHIFT(a) = ma; // column mean from here on
			const double caa = FU(a, a) - s1a * s1a * inv_sw;
			FORIG(a) = caa + sw * ma * ma; // uncentred squared norm: the aliasing reference
#pragma unroll
			for (int c = 0; c <= a; ++c) {
				double v = FU(c, a) - s1a * FS1(c) * inv_sw;
				if (!ic) v += sw * ma * FSHIFT(c);
				FU(c, a) = v;
			}
			double cxy = FU(a, PM) - s1a * s1y * inv_sw;
			if (!ic) cxy += sw * ma * ybar;
			FU(a, PM) = cxy;
		}
	}
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
	if (ridge) {
#pragma unroll
		for (int a = 0; a < PM; ++a) FU(a, a) += lambda;
	}
	// ---- Cholesky, left to right, limited pivoting. L(i, c) lives in FU(c, i); 1 / L(i, i) replaces orig(i) --
	int rank = 0;
	bool suspicious = false;
#pragma unroll
	for (int i = 0; i < PM; ++i) {
		if (i < p) {
			const double a0 = FU(i, i);
#pragma unroll
			for (int j = 0; j < i; ++j) {
				double s = FU(j, i);
#pragma unroll
				for (int c = 0; c < j; ++c) s -= FU(c, i) * FU(c, j);
				FU(j, i) = s * FORIG(j); // dead j: reciprocal 0
			}
			double d = FU(i, i);
#pragma unroll
			for (int c = 0; c < i; ++c) d -= FU(c, i) * FU(c, i);
			bool alias = ridge ? !(d > 0.0) : (!(d >= 1e-14 * FORIG(i)) || !(d > 0.0));
			if (!((dead >> i) & 1u)) suspicious = suspicious || chol_pivot_suspicious(d, a0, FORIG(i), ridge);
			alias = alias || ((dead >> i) & 1u);
			if (alias) {
				dead |= 1u << i;
#pragma unroll
				for (int c = 0; c < i; ++c) FU(c, i) = 0.0;
				FU(i, i) = 0.0;
				FORIG(i) = 0.0;
			} else {
				const double r = rsqrt(d);
				FU(i, i) = d * r;
				FORIG(i) = r;
				++rank;
			}
		} else {
			FORIG(i) = 0.0;
		}
	}
	if (suspicious) queue_qr(P, g);
	// forward solve L v = b, back solve L^T beta = v (dead columns: zero rows / columns, reciprocal 0)
	double ess = 0.0;
#pragma unroll
	for (int i = 0; i < PM; ++i) {
		if (i < p) {
			double s = FU(i, PM);
#pragma unroll
			for (int c = 0; c < i; ++c) s -= FU(c, i) * FU(c, PM);
			s *= FORIG(i);
			FU(i, PM) = s;
			ess = fma(s, s, ess);
		}
	}
#pragma unroll
	for (int i = PM - 1; i >= 0; --i) {
		if (i < p) {
			double s = FU(i, PM);
#pragma unroll
			for (int c = i + 1; c < PM; ++c) s -= FU(i, c) * FU(c, PM);
			FU(i, PM) = s * FORIG(i);
		}
	}
	double bnorm2 = 0.0, b0 = ybar;
#pragma unroll
	for (int i = 0; i < PM; ++i) {
		const double bi = FU(i, PM);
		bnorm2 = fma(bi, bi, bnorm2);
		b0 -= FSHIFT(i) * bi;
	}
	double rss = tss - ess - (ridge ? lambda * bnorm2 : 0.0);
	if (refine_count && !(rss > 1e-5 * tss)) { // cancellation guard: queue for the exact residual pass
		const int slot = atomicAdd(refine_count, 1);
		refine_list[slot] = g;
	}
	if (!(rss > 0.0)) rss = 0.0;
	const int rank_total = rank + ic;
	const double dfr = n_valid - (double)rank_total;
	const double dfm = (double)rank;
	const double r2 = 1.0 - rss / tss;
	const double adj = dfr > 0.0 ? 1.0 - (1.0 - r2) * (n_valid - (double)ic) / dfr : nanv;
	const double sigma2 = dfr > 0.0 ? rss / dfr : nanv;
	o.intercept[g] = ic ? b0 : nanv; // ols.rs:172-176
	o.r_squared[g] = r2;
	o.adj_r_squared[g] = adj;
	o.residual_std_error[g] = sqrt(sigma2);
#pragma unroll
	for (int i = 0; i < PM; ++i)
		if (i < p) o.coefficients[g * p + i] = ((dead >> i) & 1u) ? nanv : FU(i, PM);
	double aic, bic;
	info_criteria_dev(rss, n_valid, (double)rank_total, aic, bic);
	if (o.aic) o.aic[g] = aic;
	if (o.bic) o.bic[g] = bic;
	if (o.rss) o.rss[g] = rss;
	if (!inf_arrays) return;

	if (o.has_inference) o.has_inference[g] = 1;
	const DfTables Tb{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
	double tcrit = nanv;
	if (dfr > 0.0) {
		const int idf = (int)dfr;
		tcrit = (P.tcrit && idf <= kTcritTable) ? P.tcrit[idf] : t_critical_dev(P.level, dfr);
	}
	// diag((L L^T)^-1)_j = sum_i (L^-1)_ij^2: unit right-hand sides, the running solution kept in the S1 slots
#pragma unroll
	for (int j = 0; j < PM; ++j) {
		if (j < p) {
			const int64_t oj = g * p + j;
			if ((dead >> j) & 1u) {
				o.std_errors[oj] = nanv;
				o.t_values[oj] = nanv;
				o.p_values[oj] = nanv;
				o.ci_lower[oj] = nanv;
				o.ci_upper[oj] = nanv;
			} else {
				double acc = 0.0;
#pragma unroll
				for (int i = j; i < PM; ++i) {
					double s2 = (i == j) ? 1.0 : 0.0;
#pragma unroll
					for (int c = j; c < i; ++c) s2 -= FU(c, i) * FS1(c);
					s2 *= FORIG(i);
					FS1(i) = s2;
					acc = fma(s2, s2, acc);
				}
				const double se = sqrt(sigma2 * acc);
				const double bj = FU(j, PM);
				const double t = bj / se;
				o.std_errors[oj] = se;
				o.t_values[oj] = t;
				o.p_values[oj] = dfr > 0.0 ? pvalue_f(t * t, 1.0, dfr, Tb) : nanv;
				o.ci_lower[oj] = bj - tcrit * se;
				o.ci_upper[oj] = bj + tcrit * se;
			}
		}
	}
	double fs = nanv, fp = nanv;
	if (dfm > 0.0 && dfr > 0.0) {
		fs = ((tss - rss) / dfm) / sigma2;
		fp = pvalue_f(fs, dfm, dfr, Tb);
	}
	o.f_statistic[g] = fs;
	o.f_pvalue[g] = fp;
#undef FX
#undef FSHIFT
#undef FORIG
#undef FS1
#undef FU
}

// threads per CTA (= groups per CTA) that fit two CTAs per SM; 0: the records are too wide for this kernel
static inline int solve_mid_threads(int p) {
	const int L = rec_len(p);
	for (int T = 64; T >= 32; T -= 32)
		if ((size_t)L * (T + 1) * 8 <= (size_t)110 * 1024) return T;
	if ((size_t)L * 33 * 8 <= (size_t)220 * 1024) return 32;
	return 0;
}

} // namespace b200
