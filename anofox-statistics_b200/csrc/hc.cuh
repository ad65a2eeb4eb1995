// hc.cuh -- heteroscedasticity-consistent (sandwich) standard errors HC0 .. HC3 for OLS / WLS fits.
//
// Replaces the `hc_type` branch of fit_ols / fit_wls (crates/anofox-stats-core/src/models/ols.rs:209-245,
// wls.rs:230-266), which hands the design, the coefficients and the residuals to the upstream
// anofox_regression::inference::compute_hc_inference (not vendored). Restated from the published estimators
// (MacKinnon & White 1985; R sandwich::vcovHC for weighted fits):
//   V = (X'WX)^-1 [ sum_i omega_i x~_i x~_i' ] (X'WX)^-1,   x~ = sqrt(w) x,  e~ = sqrt(w) e,
//   h_i = x~_i' (X'WX)^-1 x~_i,   omega_i = e~_i^2 (HC0) | e~_i^2 n/(n-k) (HC1) | e~_i^2/(1-h_i) (HC2) | e~_i^2/(1-h_i)^2 (HC3)
// over the ESTIMABLE columns (constant and aliased columns hold NaN coefficients and stay NaN), intercept column
// included. Standard errors, t, p (same residual df) and the confidence bounds of the slopes are overwritten;
// the F statistic stays the classical one, as in the reference (it passes result.f_statistic through).
//
// This is a second streaming pass over the rows. One CTA per group:
//   pass 1  A = sum w z z', z = (1, x_kept - shift): the Gram matrix about the group's first valid row (the shift
//           lies in the span of the intercept column: leverages and slope variances do not depend on it), in
//           shared memory; Cholesky A = L L', L^-1 column by column;
//   pass 2  per row u = L^-1 z~ (one thread per row), h = |u|^2, omega from the residual of the stored
//           coefficients, S = sum omega u u';
//   finish  V_jj = sum_ab (L^-1)_aj S_ab (L^-1)_bj, se = sqrt(V_jj), t, p, bounds.
// Without an intercept the shift is zero (raw design).
#pragma once
#include "common.cuh"
#include "specfun.cuh"

namespace b200 {

constexpr int kHcThreads = 128;
constexpr int kHcChunk = 64; // rows per shared-memory chunk

struct HcArgs {
	const int64_t *offsets;
	const double *y, *x, *w;
	const uint8_t *train;
	int flags_from_y;
	int64_t n_groups;
	int hc_type; // AnofoxHcType 1 .. 4
	// list mode (after hc_small_kernel): only the groups list[0 .. *count); null = every group
	const int64_t *list;
	const int *count;
};

// dynamic shared memory: L (k x k) | Linv (k x k) | S (k x k) | zbuf (chunk x k) | om (chunk) | shift (k) | beta (k) | idx
static inline size_t hc_smem_bytes(int p) {
	const size_t k = (size_t)p + 1;
	return (3 * k * k + (size_t)kHcChunk * k + kHcChunk + 2 * k) * sizeof(double) + (k + 4) * sizeof(int);
}

__device__ __forceinline__ bool hc_row_ok(const HcArgs &a, int64_t r, int p) {
	bool ok = row_ok_yw(a.y[r], a.w ? a.w[r] : 1.0, a.train ? (int)a.train[r] : 1, a.w != nullptr);
	for (int c = 0; ok && c < p; ++c) ok = finite_f64(a.x[r * p + c]);
	return ok;
}

__global__ void __launch_bounds__(kHcThreads) hc_kernel(HcArgs a, FitParams P, AnofoxB200FitOutputs o) {
	extern __shared__ __align__(16) unsigned char hc_raw[];
	const int p = P.p, tid = threadIdx.x;
	const int ic = P.fit_intercept ? 1 : 0;
	const int kmax = p + 1;
	double *L = reinterpret_cast<double *>(hc_raw);
	double *Li = L + kmax * kmax;
	double *S = Li + kmax * kmax;
	double *zb = S + kmax * kmax;
	double *om = zb + kHcChunk * kmax;
	double *sh = om + kHcChunk;
	double *beta = sh + kmax;
	int *idx = reinterpret_cast<int *>(beta + kmax);
	__shared__ int s_first, s_k, s_bad;
	const double nanv = nan_f64();
	const DfTables Tb{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
	const int64_t n_todo = a.list ? (int64_t)*a.count : a.n_groups;
	for (int64_t it = blockIdx.x; it < n_todo; it += gridDim.x) {
		const int64_t g = a.list ? a.list[it] : it;
		__syncthreads();
		if (o.status[g] != 0 || (o.has_inference && o.has_inference[g] == 0)) continue;
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		const int64_t nobs = o.n_observations[g];
		// estimable columns = those with a coefficient
		if (tid == 0) {
			int k = 0;
			if (ic) idx[k++] = -1; // the intercept column
			for (int j = 0; j < p; ++j)
				if (!isnan(o.coefficients[g * p + j])) idx[k++] = j;
			s_k = k;
			s_first = 0x7fffffff;
			s_bad = 0;
		}
		__syncthreads();
		const int k = s_k;
		if (k - ic <= 0) continue;
		// first valid row (state order) = the shift
		for (int64_t r = r0 + tid; r < r1; r += kHcThreads)
			if (hc_row_ok(a, r, p)) {
				atomicMin(&s_first, (int)(r - r0));
				break;
			}
		__syncthreads();
		if (s_first == 0x7fffffff) continue;
		const int64_t rf = r0 + s_first;
		for (int c = tid; c < k; c += kHcThreads) {
			const int j = idx[c];
			sh[c] = (j < 0 || !ic) ? 0.0 : a.x[rf * p + j];
			beta[c] = j < 0 ? 0.0 : o.coefficients[g * p + j];
		}
		for (int e = tid; e < k * k; e += kHcThreads) L[e] = S[e] = 0.0;
		__syncthreads();
		const double shy = ic ? a.y[rf] : 0.0;
		// residual in the shifted frame: e = (y - shy) - c0 - sum beta (x - sh), c0 = b0 + sum beta sh - shy
		double c0 = 0.0;
		if (ic) {
			c0 = o.intercept[g] - shy;
			for (int c = 1; c < k; ++c) c0 += beta[c] * sh[c];
		}
		// ---- pass 1: Gram about the shift ----
		for (int64_t rb = r0; rb < r1; rb += kHcChunk) {
			const int nr = (int)min((int64_t)kHcChunk, r1 - rb);
			for (int e = tid; e < nr * k; e += kHcThreads) {
				const int i = e / k, c = e - i * k;
				const int64_t r = rb + i;
				const int j = idx[c];
				zb[i * k + c] = j < 0 ? 1.0 : a.x[r * p + j] - sh[c];
			}
			for (int i = tid; i < nr; i += kHcThreads) {
				const int64_t r = rb + i;
				bool ok = hc_row_ok(a, r, p);
				om[i] = ok ? (a.w ? a.w[r] : 1.0) : 0.0; // weight; 0 removes the row
			}
			__syncthreads();
			for (int e = tid; e < k * k; e += kHcThreads) {
				const int ra = e / k, cb = e - ra * k;
				if (cb > ra) continue;
				double acc = 0.0;
				for (int i = 0; i < nr; ++i) {
					const double wv = om[i];
					if (wv != 0.0) acc = fma(wv * zb[i * k + ra], zb[i * k + cb], acc);
				}
				L[ra * k + cb] += acc;
			}
			__syncthreads();
		}
		// ---- Cholesky in place (lower), column by column ----
		for (int c = 0; c < k; ++c) {
			if (tid == 0) {
				double d = L[c * k + c];
				for (int q = 0; q < c; ++q) d -= L[c * k + q] * L[c * k + q];
				if (!(d > 0.0)) s_bad = 1;
				L[c * k + c] = sqrt(d);
			}
			__syncthreads();
			if (s_bad) break;
			const double dinv = 1.0 / L[c * k + c];
			for (int r = c + 1 + tid; r < k; r += kHcThreads) {
				double v = L[r * k + c];
				for (int q = 0; q < c; ++q) v -= L[r * k + q] * L[c * k + q];
				L[r * k + c] = v * dinv;
			}
			__syncthreads();
		}
		if (s_bad) continue; // numerically singular Gram: keep the classical statistics (ols.rs:232-244 falls back too)
		// ---- L^-1: thread j solves L x = e_j ----
		for (int j = tid; j < k; j += kHcThreads) {
			for (int r = 0; r < k; ++r) {
				double v = r == j ? 1.0 : 0.0;
				if (r < j) {
					Li[r * k + j] = 0.0;
					continue;
				}
				for (int q = j; q < r; ++q) v -= L[r * k + q] * Li[q * k + j];
				Li[r * k + j] = v / L[r * k + r];
			}
		}
		__syncthreads();
		// ---- pass 2: leverages, weights omega, S = sum omega u u' ----
		const double hc1 = (double)nobs / ((double)nobs - (double)k);
		for (int64_t rb = r0; rb < r1; rb += kHcChunk) {
			const int nr = (int)min((int64_t)kHcChunk, r1 - rb);
			for (int i = tid; i < nr; i += kHcThreads) {
				const int64_t r = rb + i;
				double *u = zb + i * k;
				if (!hc_row_ok(a, r, p)) {
					om[i] = 0.0;
					for (int c = 0; c < k; ++c) u[c] = 0.0;
					continue;
				}
				const double wv = a.w ? a.w[r] : 1.0, sq = sqrt(wv);
				double fit = c0;
				for (int c = 0; c < k; ++c) {
					const int j = idx[c];
					const double zc = j < 0 ? 1.0 : a.x[r * p + j] - sh[c];
					if (j >= 0) fit = fma(beta[c], zc, fit);
					u[c] = sq * zc;
				}
				const double et = sq * ((a.y[r] - shy) - fit);
				double h = 0.0;
				for (int c = 0; c < k; ++c) { // forward solve in place: u := L^-1 z~
					double v = u[c];
					for (int q = 0; q < c; ++q) v -= L[c * k + q] * u[q];
					v /= L[c * k + c];
					u[c] = v;
					h = fma(v, v, h);
				}
				double w2 = et * et;
				if (a.hc_type == ANOFOX_HC_HC1) w2 *= hc1;
				else if (a.hc_type == ANOFOX_HC_HC2) w2 /= (1.0 - h);
				else if (a.hc_type == ANOFOX_HC_HC3) w2 /= (1.0 - h) * (1.0 - h);
				om[i] = w2;
			}
			__syncthreads();
			for (int e = tid; e < k * k; e += kHcThreads) {
				const int ra = e / k, cb = e - ra * k;
				if (cb > ra) continue;
				double acc = 0.0;
				for (int i = 0; i < nr; ++i) acc = fma(om[i] * zb[i * k + ra], zb[i * k + cb], acc);
				S[ra * k + cb] += acc;
			}
			__syncthreads();
		}
		// ---- V = L^-T S L^-1: diagonal entries of the slopes ----
		// u = L^-1 z, (X'WX)^-1 z = L^-T u  =>  V = L^-T S L^-1,  V_jj = sum_ab Li[a][j] S[a][b] Li[b][j]
		const double dfr = (double)nobs - (double)k;
		double tcrit = nanv;
		if (dfr > 0.0) {
			const int idf = (int)dfr;
			tcrit = (P.tcrit && idf <= kTcritTable) ? P.tcrit[idf] : t_critical_dev(P.level, dfr);
		}
		for (int c = ic + tid; c < k; c += kHcThreads) {
			double v = 0.0;
			for (int ra = c; ra < k; ++ra) {
				double t = 0.0;
				for (int cb = c; cb < k; ++cb) t = fma(cb <= ra ? S[ra * k + cb] : S[cb * k + ra], Li[cb * k + c], t);
				v = fma(Li[ra * k + c], t, v);
			}
			const int j = idx[c];
			const int64_t oj = g * p + j;
			const double se = sqrt(v), b = beta[c], t = b / se;
			o.std_errors[oj] = se;
			o.t_values[oj] = t;
			o.p_values[oj] = dfr > 0.0 ? pvalue_f(t * t, 1.0, dfr, Tb) : nanv;
			o.ci_lower[oj] = b - tcrit * se;
			o.ci_upper[oj] = b + tcrit * se;
		}
	}
}

} // namespace b200
