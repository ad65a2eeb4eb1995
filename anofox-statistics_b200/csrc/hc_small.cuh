// hc_small.cuh -- HC0 .. HC3 sandwich standard errors for MANY SMALL groups (p <= 8): four lanes per group.
//
// Same estimator and same reference lines as hc.cuh (ols.rs:209-245, wls.rs:230-266; MacKinnon & White 1985), other
// work mapping: hc_kernel spends one CTA and a dozen barriers on every group, which is right for groups of thousands
// of rows and 35x too slow for C2 (10^6 groups of 50 rows: 32 ms next to a 0.9 ms fit). Here a group belongs to the
// four lanes of a quad; lane q takes rows q, q+4, ...; every matrix lives in registers (P is a template parameter,
// all loops unrolled), the two reductions per group are xor-butterflies inside the quad (all four lanes end with
// bitwise-identical sums, so every later decision is uniform over the quad), and the finish is spread over the
// lanes so that the distribution function is evaluated once per coefficient.
//   pass 1  A = sum w z z',  z = (1, x - shift) with a ZERO in the place of every column without a coefficient
//           (constant / aliased columns, and the intercept place of a fit without one); their diagonal is set to 1,
//           so that one fixed-size Cholesky serves every pattern of dropped columns: L = diag(1) there, u = 0,
//           nothing leaks into the kept columns
//   pass 2  u = L^-1 z~ per row, h = |u|^2, omega from the residual, S = sum omega u u'   (rows come from L1 / L2:
//           a group is a few KB and was read by pass 1 microseconds earlier)
//   finish  V_cc = v' S v, v = column c of L^-1;  se, t, p, bounds
// Groups with more than kHcSmallMaxRows rows are appended to a list and left to hc_kernel (list mode).
#pragma once
#include "accumulate.cuh"
#include "common.cuh"
#include "hc_launch.h"
#include "specfun.cuh"

namespace b200 {

constexpr int kHcSmallThreads = 128;
constexpr int kHcSmallLanes = 4;

template <int P> __device__ __forceinline__ bool hcs_row(const HcSmallArgs &a, int64_t r, double (&xr)[P], double &yv, double &wv) {
	yv = a.y[r];
	wv = a.w ? a.w[r] : 1.0;
	bool ok = row_ok_yw(yv, wv, a.train ? (int)a.train[r] : 1, a.w != nullptr);
#pragma unroll
	for (int j = 0; j < P; ++j) {
		xr[j] = a.x[r * P + j];
		ok = ok && finite_f64(xr[j]);
	}
	return ok;
}

__device__ __forceinline__ double hcs_quad_sum(double v, unsigned gmask) {
	v += __shfl_xor_sync(gmask, v, 1);
	v += __shfl_xor_sync(gmask, v, 2);
	return v;
}

// p <= 5: four CTAs per SM (128 registers, some spills) measured 12 % faster than three at 168 registers (C2 + HC1:
// HC pass 2.29 -> 2.02 ms): the kernel is bound by the latency of its dependent FP64 chains, not by issue slots
template <int P>
__global__ void __launch_bounds__(kHcSmallThreads, (P <= 5 ? 4 : 1)) hc_small_kernel(HcSmallArgs a, FitParams Pm, AnofoxB200FitOutputs o) {
	constexpr int K = P + 1;
	constexpr int TRI = K * (K + 1) / 2;
#define HCS_T(r, c) ((r) * ((r) + 1) / 2 + (c)) /* lower triangle, c <= r */
	const int lane = threadIdx.x & 31, q = lane & 3;
	const unsigned gmask = 0xFu << (lane & ~3);
	const int ic = Pm.fit_intercept ? 1 : 0;
	const double nanv = nan_f64();
	const DfTables Tb{Pm.tcrit, Pm.gratio, Pm.rcp, Pm.ce, Pm.co};
	const int64_t stride = (int64_t)gridDim.x * (kHcSmallThreads / kHcSmallLanes);
	for (int64_t g = (int64_t)blockIdx.x * (kHcSmallThreads / kHcSmallLanes) + (threadIdx.x >> 2); g < a.n_groups; g += stride) {
		if (o.status[g] != 0 || (o.has_inference && o.has_inference[g] == 0)) continue;
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		if (r1 - r0 > kHcSmallMaxRows) {
			if (q == 0) a.big_list[atomicAdd(a.big_count, 1)] = g;
			continue;
		}
		double beta[P];
		unsigned keep = 0;
#pragma unroll
		for (int j = 0; j < P; ++j) {
			const double b = o.coefficients[g * P + j];
			const bool kept = !isnan(b);
			beta[j] = kept ? b : 0.0;
			keep |= kept ? (1u << j) : 0u;
		}
		const int nkept = __popc(keep);
		if (nkept == 0) continue;
		const int k = ic + nkept;
		const int64_t nobs = o.n_observations[g];
		// ---- the shift: the group's first valid row in state order ----
		double xr[P], yv, wv;
		int64_t rf = INT64_MAX;
		for (int64_t r = r0 + q; r < r1; r += kHcSmallLanes)
			if (hcs_row<P>(a, r, xr, yv, wv)) {
				rf = r;
				break;
			}
		{
			int64_t t = __shfl_xor_sync(gmask, rf, 1);
			rf = t < rf ? t : rf;
			t = __shfl_xor_sync(gmask, rf, 2);
			rf = t < rf ? t : rf;
		}
		if (rf == INT64_MAX) continue;
		double sh[P], shy = 0.0;
#pragma unroll
		for (int j = 0; j < P; ++j) sh[j] = ic ? a.x[rf * P + j] : 0.0;
		double c0 = 0.0; // residual in the shifted frame: e = (y - shy) - c0 - sum beta (x - sh)
		if (ic) {
			shy = a.y[rf];
			c0 = o.intercept[g] - shy;
#pragma unroll
			for (int j = 0; j < P; ++j) c0 = fma(beta[j], sh[j], c0);
		}
		// ---- pass 1: Gram about the shift ----
		double L[TRI];
#pragma unroll
		for (int e = 0; e < TRI; ++e) L[e] = 0.0;
		for (int64_t r = r0 + q; r < r1; r += kHcSmallLanes) {
			if (!hcs_row<P>(a, r, xr, yv, wv)) continue;
			double z[K];
			z[0] = ic ? 1.0 : 0.0;
#pragma unroll
			for (int j = 0; j < P; ++j) z[1 + j] = ((keep >> j) & 1u) ? xr[j] - sh[j] : 0.0;
#pragma unroll
			for (int rr = 0; rr < K; ++rr) {
				const double wz = wv * z[rr];
#pragma unroll
				for (int cc = 0; cc <= rr; ++cc) L[HCS_T(rr, cc)] = fma(wz, z[cc], L[HCS_T(rr, cc)]);
			}
		}
#pragma unroll
		for (int e = 0; e < TRI; ++e) L[e] = hcs_quad_sum(L[e], gmask);
		if (!ic) L[HCS_T(0, 0)] = 1.0;
#pragma unroll
		for (int j = 0; j < P; ++j)
			if (!((keep >> j) & 1u)) L[HCS_T(1 + j, 1 + j)] = 1.0;
		// ---- Cholesky in place; dinv = 1 / diagonal ----
		double dinv[K];
		bool bad = false;
#pragma unroll
		for (int c = 0; c < K; ++c) {
			double d = L[HCS_T(c, c)];
#pragma unroll
			for (int s = 0; s < c; ++s) d = fma(-L[HCS_T(c, s)], L[HCS_T(c, s)], d);
			bad = bad || !(d > 0.0);
			const double sd = sqrt(d);
			L[HCS_T(c, c)] = sd;
			dinv[c] = 1.0 / sd;
#pragma unroll
			for (int r = c + 1; r < K; ++r) {
				double v = L[HCS_T(r, c)];
#pragma unroll
				for (int s = 0; s < c; ++s) v = fma(-L[HCS_T(r, s)], L[HCS_T(c, s)], v);
				L[HCS_T(r, c)] = v * dinv[c];
			}
		}
		if (bad) continue; // numerically singular Gram: the classical statistics stay (as hc_kernel, ols.rs:232-244)
		// ---- pass 2: leverages, omega, S = sum omega u u' ----
		const double hc1 = (double)nobs / ((double)nobs - (double)k);
		double S[TRI];
#pragma unroll
		for (int e = 0; e < TRI; ++e) S[e] = 0.0;
		for (int64_t r = r0 + q; r < r1; r += kHcSmallLanes) {
			if (!hcs_row<P>(a, r, xr, yv, wv)) continue;
			const double sq = sqrt(wv);
			double u[K], fit = c0;
			u[0] = ic ? sq : 0.0;
#pragma unroll
			for (int j = 0; j < P; ++j) {
				const double zc = ((keep >> j) & 1u) ? xr[j] - sh[j] : 0.0;
				fit = fma(beta[j], zc, fit);
				u[1 + j] = sq * zc;
			}
			const double et = sq * ((yv - shy) - fit);
			double h = 0.0;
#pragma unroll
			for (int c = 0; c < K; ++c) { // forward solve in place: u := L^-1 z~
				double v = u[c];
#pragma unroll
				for (int s = 0; s < c; ++s) v = fma(-L[HCS_T(c, s)], u[s], v);
				v *= dinv[c];
				u[c] = v;
				h = fma(v, v, h);
			}
			double w2 = et * et;
			if (a.hc_type == ANOFOX_HC_HC1) w2 *= hc1;
			else if (a.hc_type == ANOFOX_HC_HC2) w2 /= (1.0 - h);
			else if (a.hc_type == ANOFOX_HC_HC3) w2 /= (1.0 - h) * (1.0 - h);
#pragma unroll
			for (int rr = 0; rr < K; ++rr) {
				const double ou = w2 * u[rr];
#pragma unroll
				for (int cc = 0; cc <= rr; ++cc) S[HCS_T(rr, cc)] = fma(ou, u[cc], S[HCS_T(rr, cc)]);
			}
		}
#pragma unroll
		for (int e = 0; e < TRI; ++e) S[e] = hcs_quad_sum(S[e], gmask);
		// ---- V_cc = v' S v with v = L^-1 e_c (entries c .. K-1), slopes only ----
		double V[P];
#pragma unroll
		for (int c = 1; c < K; ++c) {
			double v[K];
#pragma unroll
			for (int r = c; r < K; ++r) {
				double t = r == c ? 1.0 : 0.0;
#pragma unroll
				for (int s = c; s < r; ++s) t = fma(-L[HCS_T(r, s)], v[s], t);
				v[r] = t * dinv[r];
			}
			double acc = 0.0;
#pragma unroll
			for (int ra = c; ra < K; ++ra) {
				double t = 0.0;
#pragma unroll
				for (int cb = c; cb < K; ++cb) t = fma(cb <= ra ? S[HCS_T(ra, cb)] : S[HCS_T(cb, ra)], v[cb], t);
				acc = fma(v[ra], t, acc);
			}
			V[c - 1] = acc;
		}
		// ---- finish: lane q takes the slopes q, q+4 ----
		const double dfr = (double)nobs - (double)k;
		double tcrit = nanv;
		if (dfr > 0.0) {
			const int idf = (int)dfr;
			tcrit = (Pm.tcrit && idf <= kTcritTable) ? Pm.tcrit[idf] : t_critical_dev(Pm.level, dfr);
		}
#pragma unroll
		for (int rd = 0; rd < (P + kHcSmallLanes - 1) / kHcSmallLanes; ++rd) {
			const int j = rd * kHcSmallLanes + q;
			if (j >= P || !((keep >> j) & 1u)) continue;
			double vj = 0.0, b = 0.0;
#pragma unroll
			for (int jj = 0; jj < P; ++jj)
				if (jj == j) {
					vj = V[jj];
					b = beta[jj];
				}
			const int64_t oj = g * P + j;
			const double se = sqrt(vj), t = b / se;
			o.std_errors[oj] = se;
			o.t_values[oj] = t;
			o.p_values[oj] = dfr > 0.0 ? pvalue_f(t * t, 1.0, dfr, Tb) : nanv;
			o.ci_lower[oj] = b - tcrit * se;
			o.ci_upper[oj] = b + tcrit * se;
		}
	}
#undef HCS_T
}

} // namespace b200
