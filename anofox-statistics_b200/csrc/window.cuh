// window.cuh -- the window form of the fit-predict aggregates:
//   {ols,ridge,wls}_fit_predict(y, x [, w], opts) OVER (PARTITION BY g ORDER BY t ROWS BETWEEN .. AND ..)
// (src/window_functions/ols_fit_predict.cpp:111-326, ridge_fit_predict.cpp, wls_fit_predict.cpp; the reference's
// headline benchmark examples/performance_1m_groups/benchmark_ols.sql:6-22 is the expanding frame
// "UNBOUNDED PRECEDING AND 1 PRECEDING"). The reference re-runs the whole aggregate for every output row:
// it copies the frame's rows into a fresh state and fits them from scratch -- O(n^2 p^2) per partition.
//
// Here the moment record is additive, so an expanding frame is an inclusive SCAN: one warp walks a partition 32
// rows at a time, every lane takes one row, the shifted sums (n, sum w, S1, S2 about the partition's first training
// row, and the per-column "differs from the first row by >= 1e-10" flags of the constant-column rule, ols.rs:76-87, which for a frame that starts
// at the partition's first row is exact) are prefix-summed across the lanes with shuffles and carried from chunk
// to chunk; lane i then holds the sums of its own frame in registers and runs the thread-per-group solve of the
// small-group kernels (small_solve, no inference: ols_fit_predict.cpp:281) followed by the prediction rule of
// anofox_predict_with_interval (lib.rs:2292-2347) at the x of the frame's last row (:157-163).
// A bounded frame ("k PRECEDING") is summed directly by its lane about the frame's own first training row.
// Work per partition: O(n p^2) accumulation + n solves.
//
// NULL is NaN: a row trains when y (and w for WLS) is present; a NaN feature drops the row inside the fit
// (ols.rs:59-66) but the row still counts for the "more than p + intercept training rows" rule (:258-262).
// Output NaN triple = NULL struct (no x on the last frame row -> its features are NaN -> yhat NaN as well).
#pragma once
#include "fused_small.cuh"

namespace b200 {

struct WindowArgs {
	const int64_t *offsets;
	const double *y, *x, *w;
	int64_t n_groups;
	int64_t preceding;  // < 0: unbounded preceding (expanding frame), else rows before the current one
	int64_t end_offset; // frame ends at row i - end_offset (0 = CURRENT ROW, 1 = 1 PRECEDING)
	double *yhat, *lower, *upper;
};

template <int P> struct WinSums {
	SmallMoments<P> M;  // nv = rows that enter the fit (finite y, x; w > 0 finite)
	unsigned nc;        // bit c: some fitting row has |x_c - first_c| >= 1e-10 (the column is not constant, ols.rs:76-87)
	int n_train;        // rows with y (and w) present: the aggregate's y_values.size()
};

template <int P, bool HASW> __device__ __forceinline__ void win_clear(WinSums<P> &s) {
	s.M.clear();
	s.nc = 0u;
	s.n_train = 0;
}

// a += b (the constant-column flags combine with OR)
template <int P, bool HASW> __device__ __forceinline__ void win_add(WinSums<P> &a, const WinSums<P> &b) {
#pragma unroll
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) a.M.s2[i] += b.M.s2[i];
#pragma unroll
	for (int i = 0; i < P + 1; ++i) a.M.s1[i] += b.M.s1[i];
	if (HASW) a.M.sw += b.M.sw;
	a.nc |= b.nc;
	a.M.nv += b.M.nv;
	a.n_train += b.n_train;
}

template <int P, bool HASW> __device__ __forceinline__ WinSums<P> win_shfl_up(const WinSums<P> &s, int delta) {
	WinSums<P> r;
#pragma unroll
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) r.M.s2[i] = __shfl_up_sync(0xffffffffu, s.M.s2[i], delta);
#pragma unroll
	for (int i = 0; i < P + 1; ++i) r.M.s1[i] = __shfl_up_sync(0xffffffffu, s.M.s1[i], delta);
	r.M.sw = HASW ? __shfl_up_sync(0xffffffffu, s.M.sw, delta) : 0.0;
#pragma unroll
	for (int c = 0; c < P; ++c) r.M.q2[c] = 0.0;
	r.nc = __shfl_up_sync(0xffffffffu, s.nc, delta);
	r.M.nv = __shfl_up_sync(0xffffffffu, s.M.nv, delta);
	r.M.ns = 0;
	r.n_train = __shfl_up_sync(0xffffffffu, s.n_train, delta);
	return r;
}
template <int P, bool HASW> __device__ __forceinline__ WinSums<P> win_bcast(const WinSums<P> &s, int src) {
	WinSums<P> r;
#pragma unroll
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) r.M.s2[i] = __shfl_sync(0xffffffffu, s.M.s2[i], src);
#pragma unroll
	for (int i = 0; i < P + 1; ++i) r.M.s1[i] = __shfl_sync(0xffffffffu, s.M.s1[i], src);
	r.M.sw = HASW ? __shfl_sync(0xffffffffu, s.M.sw, src) : 0.0;
#pragma unroll
	for (int c = 0; c < P; ++c) r.M.q2[c] = 0.0;
	r.nc = __shfl_sync(0xffffffffu, s.nc, src);
	r.M.nv = __shfl_sync(0xffffffffu, s.M.nv, src);
	r.M.ns = 0;
	r.n_train = __shfl_sync(0xffffffffu, s.n_train, src);
	return r;
}

// One row's contribution about `sh` (zero when the row does not enter the fit).
template <int P, bool HASW>
__device__ __forceinline__ void win_row(WinSums<P> &s, const double (&xr)[P], double yv, double wv, const double (&sh)[P + 1]) {
	constexpr int K = P + 1;
	bool train = !isnan(yv);
	if (HASW) train = train && !isnan(wv);
	bool ok = train && finite_f64(yv);
	if (HASW) ok = ok && wv > 0.0 && finite_f64(wv);
#pragma unroll
	for (int c = 0; c < P; ++c) ok = ok && finite_f64(xr[c]);
	s.n_train += train ? 1 : 0;
	if (!ok) return;
	double d[K], a[K];
#pragma unroll
	for (int c = 0; c < P; ++c) d[c] = xr[c] - sh[c];
	d[P] = yv - sh[P];
#pragma unroll
	for (int c = 0; c < K; ++c) a[c] = HASW ? wv * d[c] : d[c];
#pragma unroll
	for (int i = 0; i < K; ++i)
#pragma unroll
		for (int j = i; j < K; ++j) s.M.s2[utri(i, j, K)] = fma(a[i], d[j], s.M.s2[utri(i, j, K)]);
#pragma unroll
	for (int c = 0; c < K; ++c) s.M.s1[c] += a[c];
	if (HASW) s.M.sw += wv;
#pragma unroll
	for (int c = 0; c < P; ++c) s.nc |= !(fabs(d[c]) < 1e-10) ? (1u << c) : 0u;
	s.M.nv += 1;
}

template <int P, bool HASW> __device__ __forceinline__ bool win_row_fits(const WindowArgs &a, int64_t r) {
	const double yv = a.y[r];
	bool ok = finite_f64(yv);
	if (HASW) {
		const double wv = a.w[r];
		ok = ok && wv > 0.0 && finite_f64(wv);
	}
#pragma unroll
	for (int c = 0; c < P; ++c) ok = ok && finite_f64(a.x[r * P + c]);
	return ok;
}

// Fit the frame whose sums are in F (about sh) and predict at xc. `slot` indexes the per-row scratch outputs.
template <int P, bool HASW>
__device__ __forceinline__ void win_fit_predict(const WindowArgs &a, WinSums<P> &F, const double (&sh)[P + 1],
                                                const double (&xc)[P], int64_t slot, int64_t f_lo, int64_t f_hi,
                                                const FitParams &Pm, const AnofoxB200FitOutputs &o, double *fx) {
	const double nanv = nan_f64();
	double yh = nanv, lo = nanv, hi = nanv;
	const int ic = Pm.fit_intercept ? 1 : 0;
	if (F.n_train > P + ic) { // ols_fit_predict.cpp:258-262: y_values.size() <= min_obs -> NULL
		const unsigned nonconst = F.nc; // ols.rs:76-87
		if (!HASW) F.M.sw = (double)F.M.nv;
		F.M.ns = F.n_train;
		SmallGroupView v;
		v.n = 0;
		v.q = 0;
		v.T = 1;
		v.mask = 0u;
		v.pv = nullptr;
		v.refine_count = nullptr;
		v.refine_list = nullptr;
		v.rec_out = nullptr;
		v.stash = nullptr;
		v.fit_x = fx; // the fit's results stay on chip: no per-row scratch in HBM
		small_solve<P, false, false, false>(F.M, sh, nonconst, F.n_train, slot, Pm, o, v);
		if (fx[0] == 0.0) {
			double b[P];
#pragma unroll
			for (int c = 0; c < P; ++c) b[c] = fx[4 + c];
			const double icpt = fx[1];
			double rse = fx[2];
			const int64_t n = (int64_t)fx[3];
			const double r2 = fx[4 + P];
			if (!(1.0 - r2 > 1e-5)) {
				// (nearly) exact fit: tss - ess has cancelled; take the residuals from the frame's rows themselves
				int rank = ic;
#pragma unroll
				for (int c = 0; c < P; ++c) rank += isnan(b[c]) ? 0 : 1;
				double c0 = 0.0;
				if (ic) {
					c0 = icpt - sh[P];
#pragma unroll
					for (int c = 0; c < P; ++c)
						if (!isnan(b[c])) c0 = fma(b[c], sh[c], c0);
				}
				double rss = 0.0;
				for (int64_t r = f_lo; r <= f_hi; ++r) {
					if (!win_row_fits<P, HASW>(a, r)) continue;
					double fit = c0;
#pragma unroll
					for (int c = 0; c < P; ++c)
						if (!isnan(b[c])) fit = fma(b[c], ic ? a.x[r * P + c] - sh[c] : a.x[r * P + c], fit);
					const double e = (ic ? a.y[r] - sh[P] : a.y[r]) - fit;
					rss = fma(HASW ? a.w[r] * e : e, e, rss);
				}
				const double dfr = (double)n - (double)rank;
				rse = dfr > 0.0 ? sqrt(rss / dfr) : nanv;
			}
			// lib.rs:2292-2347
			double s = isnan(icpt) ? 0.0 : icpt;
#pragma unroll
			for (int c = 0; c < P; ++c)
				if (!isnan(b[c])) s = fma(b[c], xc[c], s);
			yh = lo = hi = s;
			if (!(isnan(rse) || rse <= 0.0 || n <= (int64_t)P + 1)) {
				const int64_t kk = (int64_t)P + (isnan(icpt) ? 0 : 1);
				const int64_t df = n > kk ? n - kk : 0;
				if (df > 0) {
					const double tc = (Pm.tcrit && df <= kTcritTable) ? Pm.tcrit[df] : t_critical_dev(Pm.level, (double)df);
					if (!isnan(tc)) {
						const double margin = tc * (rse * sqrt(1.0 + 1.0 / (double)n));
						lo = s - margin;
						hi = s + margin;
					}
				}
			}
		}
	}
	a.yhat[slot] = yh;
	a.lower[slot] = lo;
	a.upper[slot] = hi;
}

// p <= 3: four CTAs per SM (128 registers, some spills) measured 4 % faster on W1 than two CTAs at 218 registers;
// three CTAs (168 registers) measured slower than either
#ifndef ANOFOX_WINDOW_MINB
#define ANOFOX_WINDOW_MINB 4
#endif
template <int P, bool HASW>
__global__ void __launch_bounds__(128, (P <= 3 ? ANOFOX_WINDOW_MINB : 1)) window_fit_predict_kernel(const WindowArgs a, FitParams Pm, AnofoxB200FitOutputs o) {
	constexpr int K = P + 1;
	// per-thread hand-over of a frame's fit (status, intercept, rse, n, coefficients, r2: SmallGroupView::fit_x);
	// 13 doubles apart: odd multiple of two banks
	__shared__ double fit_slots[128][13];
	double *fx = fit_slots[threadIdx.x];
	const int lane = threadIdx.x & 31;
	const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
	const double nanv = nan_f64();
	for (int64_t g = wid; g < a.n_groups; g += nw) {
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		if (a.preceding >= 0) {
			// ---- bounded frame: every lane sums its own frame about the frame's first fitting row ----
			for (int64_t rb = r0; rb < r1; rb += 32) {
				const int64_t r = rb + lane;
				if (r >= r1) continue;
				const int64_t hi = r - a.end_offset;
				int64_t lo = r - a.preceding;
				if (lo < r0) lo = r0;
				if (hi < lo) {
					a.yhat[r] = a.lower[r] = a.upper[r] = nanv;
					continue;
				}
				double sh[K], xr[P], xc[P];
#pragma unroll
				for (int c = 0; c < K; ++c) sh[c] = 0.0;
				int64_t f = -1;
				for (int64_t q = lo; q <= hi && f < 0; ++q)
					if (win_row_fits<P, HASW>(a, q)) f = q;
				if (f >= 0) {
#pragma unroll
					for (int c = 0; c < P; ++c) sh[c] = a.x[f * P + c];
					sh[P] = a.y[f];
				}
				WinSums<P> F;
				win_clear<P, HASW>(F);
				for (int64_t q = lo; q <= hi; ++q) {
#pragma unroll
					for (int c = 0; c < P; ++c) xr[c] = a.x[q * P + c];
					win_row<P, HASW>(F, xr, a.y[q], HASW ? a.w[q] : 1.0, sh);
				}
#pragma unroll
				for (int c = 0; c < P; ++c) xc[c] = a.x[hi * P + c];
				win_fit_predict<P, HASW>(a, F, sh, xc, r, lo, hi, Pm, o, fx);
			}
			continue;
		}
		// ---- expanding frame: scan ----
		// the partition's first fitting row = the shift of every frame (and the row of the constant-column rule)
		int64_t f = -1;
		for (int64_t rb = r0; rb < r1 && f < 0; rb += 32) {
			const int64_t r = rb + lane;
			const bool ok = r < r1 && win_row_fits<P, HASW>(a, r);
			const unsigned m = __ballot_sync(0xffffffffu, ok);
			if (m) f = rb + __ffs(m) - 1;
		}
		double sh[K];
#pragma unroll
		for (int c = 0; c < K; ++c) sh[c] = 0.0;
		if (f >= 0) {
#pragma unroll
			for (int c = 0; c < P; ++c) sh[c] = a.x[f * P + c];
			sh[P] = a.y[f];
		}
		WinSums<P> carry; // sums of all rows before this chunk
		win_clear<P, HASW>(carry);
		double xprev[P]; // x of the row before this chunk (end_offset = 1)
#pragma unroll
		for (int c = 0; c < P; ++c) xprev[c] = nanv;
		for (int64_t rb = r0; rb < r1; rb += 32) {
			const int64_t r = rb + lane;
			double xr[P];
			double yv = nanv, wv = nanv;
			if (r < r1) {
#pragma unroll
				for (int c = 0; c < P; ++c) xr[c] = a.x[r * P + c];
				yv = a.y[r];
				if (HASW) wv = a.w[r];
			} else {
#pragma unroll
				for (int c = 0; c < P; ++c) xr[c] = nanv;
			}
			WinSums<P> S;
			win_clear<P, HASW>(S);
			win_row<P, HASW>(S, xr, yv, wv, sh);
			// inclusive scan over the lanes
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				WinSums<P> t = win_shfl_up<P, HASW>(S, d);
				if (lane >= d) win_add<P, HASW>(S, t);
			}
			win_add<P, HASW>(S, carry); // prefix over the partition, including the own row
			// the frame of row r ends at r - end_offset
			WinSums<P> F = S;
			double xc[P];
#pragma unroll
			for (int c = 0; c < P; ++c) xc[c] = xr[c];
			if (a.end_offset == 1) {
				WinSums<P> t = win_shfl_up<P, HASW>(S, 1);
				F = lane == 0 ? carry : t;
#pragma unroll
				for (int c = 0; c < P; ++c) {
					const double xp = __shfl_up_sync(0xffffffffu, xr[c], 1);
					xc[c] = lane == 0 ? xprev[c] : xp;
				}
			}
			const int64_t hi = r - a.end_offset;
			if (r < r1) {
				if (hi < r0)
					a.yhat[r] = a.lower[r] = a.upper[r] = nanv; // empty frame
				else
					win_fit_predict<P, HASW>(a, F, sh, xc, r, r0, hi, Pm, o, fx);
			}
			__syncwarp();
			carry = win_bcast<P, HASW>(S, 31);
#pragma unroll
			for (int c = 0; c < P; ++c) xprev[c] = __shfl_sync(0xffffffffu, xr[c], 31);
		}
	}
}

} // namespace b200
