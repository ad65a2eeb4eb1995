// qr.cuh -- QR fallback: one group fitted from its ROWS by Householder QR, for the groups whose Gram-Cholesky
// solve met a small pivot (chol_pivot_suspicious, common.cuh).
//
// Reference behaviour this restores: the reference never forms X'X. Its default solver is a decomposition
// of the design matrix itself (crates/anofox-stats-core/src/types.rs:121-130: SVD by default, "Cholesky ...
// falls back to QR"; call site models/ols.rs:155-161), so its error grows with cond(X); a Gram solve grows
// with cond(X)^2. The solve kernels therefore hand every group with a suspicious pivot to this kernel, which
// repeats the whole per-group path of fit_ols / fit_ridge / fit_wls (models/ols.rs:36-268, ridge.rs:36-229,
// wls.rs:37-289) on the rows: row filter, first valid row, constant-column rule, min-observation rules,
// decomposition, NaN scatter-back, statistics. The decomposition is the one the reference's golden vectors
// were produced with (R lm(): LINPACK dqrdc2, Householder with limited column pivoting, tol = 1e-7).
//
// Shape: one CTA of 128 threads per queued group, matrix M column-major in shared memory.
//   pass 0  first valid row, valid / state row counts                    (warp per row, lane per column)
//   pass 1  constant columns, sum w, weighted sums of y about the shift
//   pass 2  streaming QR: rows enter in blocks of B below the running k x k triangle R (columns = intercept,
//           kept features minus the shift row, y; each scaled by sqrt(w)); a Householder sweep folds the block
//           into R. Ridge starts from R = sqrt(lambda) I on the feature columns (the augmented system).
//           Then dqrdc2's limited pivoting is replayed on the small R (QR of R P == QR of A P).
//   pass 3  residual sum of squares from the rows with the coefficients just found (NaN = aliased: no term)
// Every reduction has a fixed order: results do not depend on the grid.
#pragma once
#include "accumulate.cuh"
#include "solve.cuh"

namespace b200 {

constexpr int kQrThreads = 128;
constexpr int kQrWarps = kQrThreads / 32;
constexpr int kQrMaxK = kMaxFeatures + 2;

struct QrArgs {
	const int64_t *offsets;
	const double *y, *x, *w;
	const uint8_t *train;
	int flags_from_y;    // train says "y present" only: a NaN feature also removes the row from the state
	const int *count;    // number of listed groups (device); null: all groups 0 .. n_all
	const int64_t *list;
	int64_t n_all;
	int B;               // rows per streamed block
	int LD;              // leading dimension of M (odd, >= p + 2 + B)
};

// Host side: geometry for p features.
static inline void qr_geometry(int p, int &B, int &LD, size_t &dyn) {
	const int K = p + 2;
	B = K <= 16 ? 160 : (K <= 34 ? 96 : 72);
	LD = (K + B) | 1;
	dyn = (size_t)LD * K * sizeof(double);
}

__device__ __forceinline__ double qr_warp_sum(double v) {
#pragma unroll
	for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
	return v;
}
// Sum over the CTA in a fixed order; every thread gets the total. `red` holds kQrWarps doubles.
__device__ __forceinline__ double qr_cta_sum(double v, double *red) {
	v = qr_warp_sum(v);
	__syncthreads();
	if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
	__syncthreads();
	double t = 0.0;
#pragma unroll
	for (int i = 0; i < kQrWarps; ++i) t += red[i];
	return t;
}

// One Householder step on the column-major matrix M (leading dimension LD): the vector is element (pr, pc)
// plus rows [i0, i1) of column pc; it is reflected onto e_pr and applied to the columns cols[0 .. nc).
// Returns the new diagonal element (to be stored by the caller after the closing barrier).
__device__ __forceinline__ double qr_reflect(double *M, int LD, int pr, int pc, int i0, int i1, const int *cols, int nc,
                                             double *red) {
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const double *v = M + (size_t)pc * LD;
	const double piv = v[pr];
	double part = 0.0;
	for (int i = i0 + tid; i < i1; i += kQrThreads) part = fma(v[i], v[i], part);
	const double tail2 = qr_cta_sum(part, red);
	const double nrm = sqrt(fma(piv, piv, tail2));
	if (nrm == 0.0) return 0.0;
	const double alpha = piv >= 0.0 ? -nrm : nrm;
	const double v0 = piv - alpha;
	const double vtv = fma(v0, v0, tail2);
	if (vtv > 0.0) {
		const double beta = 2.0 / vtv;
		for (int c = warp; c < nc; c += kQrWarps) {
			double *t = M + (size_t)cols[c] * LD;
			double dot = 0.0;
			for (int i = i0 + lane; i < i1; i += 32) dot = fma(v[i], t[i], dot);
			dot = qr_warp_sum(dot);
			dot = fma(v0, t[pr], dot) * beta;
			for (int i = i0 + lane; i < i1; i += 32) t[i] = fma(-dot, v[i], t[i]);
			__syncwarp();
			if (lane == 0) t[pr] = fma(-dot, v0, t[pr]);
		}
	}
	__syncthreads();
	return alpha;
}

struct QrShared {
	double shift[kQrMaxK];   // first valid row: x_1 .. x_p, then y at [p]
	double orig2[kQrMaxK];   // squared norms of the unshifted sqrt(w)-scaled columns (dqrdc2's reference norms)
	double sol[kQrMaxK];     // solution per physical column
	double vd[kQrMaxK];      // diag((R^T R)^-1) per physical column
	double cmaxw[kQrWarps][kMaxFeatures];
	double red[kQrWarps], red2[kQrWarps], red3[kQrWarps];
	int colmap[kMaxFeatures]; // feature -> column of M, -1 = constant column
	int feat[kQrMaxK];        // column of M -> feature (-1: intercept / y)
	int perm[kQrMaxK];
	int cols[kQrMaxK];
	unsigned long long first;
	unsigned long long nv, ns;
	int status, pe, rank;
};

// validity of row r as the reference's filters see it (ols.rs:59-66, wls.rs:76-86), evaluated by a whole warp:
// lane j holds x[r][j] (and x[r][j + 32]); returns through the references.
__device__ __forceinline__ void qr_row_load(const QrArgs &a, int p, int64_t r, int lane, double &x0, double &x1, double &yv,
                                            double &wv, bool &valid, bool &counted) {
	const double *xr = a.x + r * p;
	x0 = lane < p ? xr[lane] : 0.0;
	x1 = lane + 32 < p ? xr[lane + 32] : 0.0;
	yv = a.y[r];
	wv = a.w ? a.w[r] : 1.0;
	const int tr = a.train ? (int)a.train[r] : 1;
	const bool bad = !finite_f64(x0) || !finite_f64(x1);
	const bool nanx = isnan(x0) || isnan(x1);
	const bool anybad = __any_sync(0xffffffffu, bad);
	valid = row_ok_yw(yv, wv, tr, a.w != nullptr) && !anybad;
	counted = tr != 0;
	if (a.flags_from_y) counted = counted && !__any_sync(0xffffffffu, nanx);
}

static __global__ void __launch_bounds__(kQrThreads) qr_fit_kernel(const QrArgs a, const FitParams P, const AnofoxB200FitOutputs o) {
	extern __shared__ __align__(16) double M[];
	__shared__ QrShared S;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int p = P.p, LD = a.LD, B = a.B;
	const int ic = P.fit_intercept ? 1 : 0;
	const bool wls = P.method == ANOFOX_B200_WLS;
	const bool has_w = a.w != nullptr;
	const double nanv = nan_f64();
	const int64_t count = a.count ? (int64_t)*a.count : a.n_all;
	for (int64_t q = blockIdx.x; q < count; q += gridDim.x) {
		const int64_t g = a.list ? a.list[q] : q;
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		__syncthreads(); // previous group's shared state is dead
		if (tid == 0) {
			S.first = ~0ull;
			S.nv = S.ns = 0ull;
			S.status = 0;
		}
		__syncthreads();
		// ---------------- pass 0: first valid row, row counts -------------------------------------------
		{
			unsigned long long first = ~0ull, nv = 0, ns = 0;
			for (int64_t r = r0 + warp; r < r1; r += kQrWarps) {
				double x0, x1, yv, wv;
				bool valid, counted;
				qr_row_load(a, p, r, lane, x0, x1, yv, wv, valid, counted);
				if (valid) {
					++nv;
					if (first == ~0ull) first = (unsigned long long)r;
				}
				if (counted) ++ns;
			}
			if (lane == 0) {
				atomicMin(&S.first, first);
				atomicAdd(&S.nv, nv);
				atomicAdd(&S.ns, ns);
			}
		}
		__syncthreads();
		const int64_t n_valid = (int64_t)S.nv;
		const int64_t n_state = a.train ? (int64_t)S.ns : (a.flags_from_y ? (int64_t)S.ns : (r1 - r0));
		if (n_state < (int64_t)P.min_state_rows || n_valid == 0) {
			if (tid == 0)
				write_null_group(o, P, g, n_state < (int64_t)P.min_state_rows ? ANOFOX_B200_STATUS_TOO_FEW_ROWS : ANOFOX_ERROR_NO_VALID_DATA);
			continue;
		}
		const int64_t rf = (int64_t)S.first;
		for (int c = tid; c <= p; c += kQrThreads) S.shift[c] = c < p ? a.x[rf * p + c] : a.y[rf];
		__syncthreads();
		// ---------------- pass 1: constant columns (ols.rs:76-87), sum w, sums of y about the shift -------
		double sw, swdy, swdy2;
		{
			const double sh0 = lane < p ? S.shift[lane] : 0.0, sh1 = lane + 32 < p ? S.shift[lane + 32] : 0.0;
			const double shy = S.shift[p];
			double c0 = 0.0, c1 = 0.0, a_sw = 0.0, a_dy = 0.0, a_dy2 = 0.0;
			for (int64_t r = r0 + warp; r < r1; r += kQrWarps) {
				double x0, x1, yv, wv;
				bool valid, counted;
				qr_row_load(a, p, r, lane, x0, x1, yv, wv, valid, counted);
				if (valid) {
					c0 = fmax(c0, fabs(x0 - sh0));
					c1 = fmax(c1, fabs(x1 - sh1));
					const double ww = has_w ? wv : 1.0, dy = yv - shy;
					a_sw += ww;
					a_dy = fma(ww, dy, a_dy);
					a_dy2 = fma(ww * dy, dy, a_dy2);
				}
			}
			if (lane < p) S.cmaxw[warp][lane] = c0;
			if (lane + 32 < p) S.cmaxw[warp][lane + 32] = c1;
			if (lane == 0) {
				S.red[warp] = a_sw;
				S.red2[warp] = a_dy;
				S.red3[warp] = a_dy2;
			}
			__syncthreads();
			sw = swdy = swdy2 = 0.0;
#pragma unroll
			for (int i = 0; i < kQrWarps; ++i) {
				sw += S.red[i];
				swdy += S.red2[i];
				swdy2 += S.red3[i];
			}
			if (tid == 0) {
				int pe = 0;
				for (int j = 0; j < p; ++j) {
					double m = 0.0;
					for (int i = 0; i < kQrWarps; ++i) m = fmax(m, S.cmaxw[i][j]);
					if (!(m < 1e-10)) {
						S.colmap[j] = ic + pe;
						S.feat[ic + pe] = j;
						++pe;
					} else {
						S.colmap[j] = -1;
					}
				}
				if (ic) S.feat[0] = -1;
				S.feat[ic + pe] = -1;
				S.pe = pe;
				int st = 0;
				if (pe == 0 && !ic) st = ANOFOX_ERROR_INSUFFICIENT_DATA;                              // ols.rs:102-107
				else if (pe > 0 && n_valid < (int64_t)(pe + ic)) st = ANOFOX_ERROR_INSUFFICIENT_DATA; // ols.rs:134-139
				S.status = st;
			}
			__syncthreads();
		}
		if (S.status != 0) {
			if (tid == 0) write_null_group(o, P, g, S.status);
			continue;
		}
		const int pe = S.pe, m = pe + ic, k = m + 1, ycol = m;
		const double inv_sw = 1.0 / sw;
		const double ybar = S.shift[p] + swdy * inv_sw;
		const double syy_c = swdy2 - swdy * swdy * inv_sw;
		const double tss = ic ? syy_c : syy_c + sw * ybar * ybar;
		const bool inf_arrays = P.compute_inference && o.std_errors;
		if (tid == 0) {
			o.status[g] = 0;
			o.n_observations[g] = n_valid;
		}
		for (int j = tid; j < p; j += kQrThreads) {
			o.coefficients[g * p + j] = nanv;
			if (inf_arrays) {
				o.std_errors[g * p + j] = nanv;
				o.t_values[g * p + j] = nanv;
				o.p_values[g * p + j] = nanv;
				o.ci_lower[g * p + j] = nanv;
				o.ci_upper[g * p + j] = nanv;
			}
		}
		if (tid == 0 && inf_arrays) {
			o.f_statistic[g] = nanv;
			o.f_pvalue[g] = nanv;
		}
		if (pe == 0) {
			// all columns constant: intercept-only shortcut (ols.rs:108-130, wls.rs:126-150)
			if (tid == 0) {
				const double ss = fmax(syy_c, 0.0);
				const double var = wls ? ss * inv_sw : ss / ((double)n_valid - 1.0);
				o.intercept[g] = ybar;
				o.r_squared[g] = 0.0;
				o.adj_r_squared[g] = 0.0;
				o.residual_std_error[g] = sqrt(var);
				if (inf_arrays && o.has_inference) o.has_inference[g] = 0; // lib.rs:251-253
				double aic, bic;
				info_criteria_dev(ss, (double)n_valid, 1.0, aic, bic);
				if (o.aic) o.aic[g] = aic;
				if (o.bic) o.bic[g] = bic;
				if (o.rss) o.rss[g] = ss;
			}
			continue;
		}
		double lambda = 0.0;
		if (P.method == ANOFOX_B200_RIDGE) {
			lambda = P.alpha;
			if (P.lambda_scaling == ANOFOX_LAMBDA_SCALING_GLMNET) {
				const double sdy = sqrt(fmax(tss, 0.0) / (double)n_valid);
				lambda = P.alpha * (double)n_valid;
				if (sdy > 0.0) lambda /= sdy;
			}
		}
		const bool ridge = lambda > 0.0;

		// ---------------- pass 2: streaming Householder QR --------------------------------------------------
		// columns: [0] sqrt(w) (intercept), [ic + c] sqrt(w) (x - shift) of kept feature c, [m] sqrt(w) (y - shift_y);
		// without an intercept nothing is shifted (the fit is not shift-invariant then)
		for (int e = tid; e < k * k; e += kQrThreads) {
			const int c = e / k, r = e - c * k;
			M[(size_t)c * LD + r] = (ridge && r == c && c >= ic && c < m) ? sqrt(lambda) : 0.0;
		}
		for (int c = tid; c < k; c += kQrThreads) S.orig2[c] = 0.0;
		__syncthreads();
		{
			const int j0 = lane, j1 = lane + 32;
			const int m0 = j0 < p ? S.colmap[j0] : -1, m1 = j1 < p ? S.colmap[j1] : -1;
			const double sh0 = (ic && j0 < p) ? S.shift[j0] : 0.0, sh1 = (ic && j1 < p) ? S.shift[j1] : 0.0;
			const double shy = ic ? S.shift[p] : 0.0;
			double o0 = 0.0, o1 = 0.0, oi = 0.0; // squared norms of the UNSHIFTED columns, this warp's rows
			for (int64_t rb = r0; rb < r1; rb += B) {
				const int nb = (int)(r1 - rb < (int64_t)B ? r1 - rb : (int64_t)B);
				for (int i = warp; i < nb; i += kQrWarps) {
					double x0, x1, yv, wv;
					bool valid, counted;
					qr_row_load(a, p, rb + i, lane, x0, x1, yv, wv, valid, counted);
					const double sq = valid ? (has_w ? sqrt(wv) : 1.0) : 0.0;
					const int row = k + i;
					if (m0 >= 0) {
						const double u = valid ? sq * x0 : 0.0;
						o0 = fma(u, u, o0);
						M[(size_t)m0 * LD + row] = valid ? sq * (x0 - sh0) : 0.0;
					}
					if (m1 >= 0) {
						const double u = valid ? sq * x1 : 0.0;
						o1 = fma(u, u, o1);
						M[(size_t)m1 * LD + row] = valid ? sq * (x1 - sh1) : 0.0;
					}
					if (lane == 0) {
						if (ic) M[row] = sq;
						oi = fma(sq, sq, oi);
						M[(size_t)ycol * LD + row] = valid ? sq * (yv - shy) : 0.0;
					}
				}
				__syncthreads();
				// fold the block into R: column l's vector is R[l][l] and the block rows
				for (int l = 0; l < k; ++l) {
					for (int c = tid; c < k - l - 1; c += kQrThreads) S.cols[c] = l + 1 + c;
					__syncthreads();
					const double al = qr_reflect(M, LD, l, l, k, k + nb, S.cols, k - l - 1, S.red);
					if (tid == 0 && al != 0.0) M[(size_t)l * LD + l] = al;
					__syncthreads();
				}
			}
			// unshifted column norms: per-warp partial sums, added in warp order
			if (m0 >= 0) S.cmaxw[warp][j0] = o0;
			if (m1 >= 0) S.cmaxw[warp][j1] = o1;
			if (lane == 0) S.red[warp] = oi;
			__syncthreads();
			if (tid < k - 1) {
				const int f = S.feat[tid];
				double t = 0.0;
				for (int i = 0; i < kQrWarps; ++i) t += f >= 0 ? S.cmaxw[i][f] : S.red[i];
				S.orig2[tid] = t;
			}
			for (int c = tid; c < k; c += kQrThreads) {
				S.perm[c] = c;
				S.sol[c] = nanv;
				S.vd[c] = nanv;
			}
			__syncthreads();
		}
		// ---------------- dqrdc2's limited pivoting replayed on R (k x k, rows 0 .. k-1) -----------------------
		// rows the oracle's matrix has: valid rows (+ pe penalty rows for ridge); QR of R P equals QR of A P.
		int rank;
		{
			const int64_t rows_total = n_valid + (ridge ? pe : 0);
			const int lmax = (int)(rows_total < (int64_t)m ? rows_total : (int64_t)m);
			int lup = m, l = 0;
			while (l < lmax && l < lup) {
				const int col = S.perm[l];
				if (!ridge) {
					double part = 0.0;
					for (int i = l + tid; i < k; i += kQrThreads) {
						const double t = M[(size_t)col * LD + i];
						part = fma(t, t, part);
					}
					const double nrm = sqrt(qr_cta_sum(part, S.red));
					double on = sqrt(S.orig2[col]);
					if (on == 0.0) on = 1.0;
					if (nrm < on * 1e-7) { // negligible: cycle the column to the end, keep l (dqrdc2)
						__syncthreads();
						if (tid == 0) {
							for (int c = l; c + 1 < m; ++c) S.perm[c] = S.perm[c + 1];
							S.perm[m - 1] = col;
						}
						--lup;
						__syncthreads();
						continue;
					}
				}
				if (l + 1 < k) {
					for (int c = tid; c < m - l - 1; c += kQrThreads) S.cols[c] = S.perm[l + 1 + c];
					if (tid == 0) S.cols[m - l - 1] = ycol;
					__syncthreads();
					const double al = qr_reflect(M, LD, l, col, l + 1, k, S.cols, m - l, S.red);
					if (tid == 0 && al != 0.0) M[(size_t)col * LD + l] = al;
					__syncthreads();
				}
				++l;
			}
			rank = lup < lmax ? lup : lmax;
		}
		// ---------------- back substitution, diag((R^T R)^-1) ---------------------------------------------------
		if (tid == 0) {
			for (int qq = rank - 1; qq >= 0; --qq) {
				double s = M[(size_t)ycol * LD + qq];
				for (int c = qq + 1; c < rank; ++c) s -= M[(size_t)S.perm[c] * LD + qq] * S.sol[S.perm[c]];
				S.sol[S.perm[qq]] = s / M[(size_t)S.perm[qq] * LD + qq];
			}
		}
		// thread c: column c of R^-1, kept in the free block rows of physical column c  (rinv(c, r) = M[c LD + k + r])
		if (tid < rank) {
			const int c = tid;
			double *ri = M + (size_t)c * LD + k;
			ri[c] = 1.0 / M[(size_t)S.perm[c] * LD + c];
			for (int r = c - 1; r >= 0; --r) {
				double s = 0.0;
				for (int kk = r + 1; kk <= c; ++kk) s = fma(M[(size_t)S.perm[kk] * LD + r], ri[kk], s);
				ri[r] = -s / M[(size_t)S.perm[r] * LD + r];
			}
		}
		__syncthreads();
		if (tid < rank) {
			const int r = tid;
			double s = 0.0;
			for (int c = r; c < rank; ++c) {
				const double t = M[(size_t)c * LD + k + r];
				s = fma(t, t, s);
			}
			S.vd[S.perm[r]] = s;
		}
		__syncthreads();
		// intercept back on the unshifted scale: y - shy = b0' + sum beta (x - sh)
		double b0 = 0.0;
		if (ic) {
			b0 = S.sol[0] + S.shift[p];
			for (int c = 0; c < pe; ++c) {
				const double bc = S.sol[ic + c];
				if (!isnan(bc)) b0 -= bc * S.shift[S.feat[ic + c]];
			}
		}
		// ---------------- pass 3: residuals from the rows -------------------------------------------------------
		// evaluated in the shifted frame when an intercept is fitted: e = (y - shy) - b0' - sum beta (x - sh). The
		// terms then have the size of the data's variation, not of its offset, so a residual of 1e-6 next to values
		// of 1e6 keeps its digits.
		double rss;
		{
			const int m0 = lane < p ? S.colmap[lane] : -1, m1 = lane + 32 < p ? S.colmap[lane + 32] : -1;
			double be0 = m0 >= 0 ? S.sol[m0] : nanv, be1 = m1 >= 0 ? S.sol[m1] : nanv;
			if (isnan(be0)) be0 = 0.0; // aliased / constant: contributes nothing (lib.rs:2296-2303)
			if (isnan(be1)) be1 = 0.0;
			const double sh0 = (ic && lane < p) ? S.shift[lane] : 0.0, sh1 = (ic && lane + 32 < p) ? S.shift[lane + 32] : 0.0;
			const double shy = ic ? S.shift[p] : 0.0, b0s = ic ? S.sol[0] : 0.0;
			double acc = 0.0;
			for (int64_t r = r0 + warp; r < r1; r += kQrWarps) {
				double x0, x1, yv, wv;
				bool valid, counted;
				qr_row_load(a, p, r, lane, x0, x1, yv, wv, valid, counted);
				if (valid) { // warp-uniform
					const double dot = qr_warp_sum(fma(be0, x0 - sh0, be1 * (x1 - sh1)));
					const double e = (yv - shy) - (b0s + dot);
					acc = fma((has_w ? wv : 1.0) * e, e, acc);
				}
			}
			__syncthreads();
			if (lane == 0) S.red[warp] = acc;
			__syncthreads();
			rss = 0.0;
#pragma unroll
			for (int i = 0; i < kQrWarps; ++i) rss += S.red[i];
		}
		// ---------------- statistics (same formulas as solve_group, solve.cuh) --------------------------------------
		const int rank_total = ridge ? pe + ic : rank;
		const double nvd = (double)n_valid;
		const double dfr = nvd - (double)rank_total;
		const double dfm = (double)(rank_total - ic);
		const double r2 = 1.0 - rss / tss;
		const double sigma2 = dfr > 0.0 ? rss / dfr : nanv;
		if (tid == 0) {
			o.intercept[g] = ic ? b0 : nanv; // ols.rs:172-176
			o.r_squared[g] = r2;
			o.adj_r_squared[g] = dfr > 0.0 ? 1.0 - (1.0 - r2) * (nvd - (double)ic) / dfr : nanv;
			o.residual_std_error[g] = sqrt(sigma2);
			double aic, bic;
			info_criteria_dev(rss, nvd, (double)rank_total, aic, bic);
			if (o.aic) o.aic[g] = aic;
			if (o.bic) o.bic[g] = bic;
			if (o.rss) o.rss[g] = rss;
		}
		const DfTables T{P.tcrit, P.gratio, P.rcp, P.ce, P.co};
		double tcrit = nanv;
		if (inf_arrays && dfr > 0.0) {
			const int idf = (int)dfr;
			tcrit = (P.tcrit && idf <= kTcritTable) ? P.tcrit[idf] : t_critical_dev(P.level, dfr);
		}
		for (int c = tid; c < pe; c += kQrThreads) {
			const double bc = S.sol[ic + c];
			if (isnan(bc)) continue;
			const int64_t oj = g * p + S.feat[ic + c];
			o.coefficients[oj] = bc;
			if (inf_arrays) {
				const double se = sqrt(sigma2 * S.vd[ic + c]);
				const double t = bc / se;
				o.std_errors[oj] = se;
				o.t_values[oj] = t;
				o.p_values[oj] = dfr > 0.0 ? t_two_sided_p_fast(t, dfr, T) : nanv;
				o.ci_lower[oj] = bc - tcrit * se;
				o.ci_upper[oj] = bc + tcrit * se;
			}
		}
		if (tid == 32 && inf_arrays) {
			if (o.has_inference) o.has_inference[g] = 1;
			if (dfm > 0.0 && dfr > 0.0) {
				const double f = ((tss - rss) / dfm) / sigma2;
				o.f_statistic[g] = f;
				o.f_pvalue[g] = f_upper_p_fast(f, dfm, dfr, T);
			}
		}
	}
}

} // namespace b200
