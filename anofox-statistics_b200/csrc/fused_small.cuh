// fused_small.cuh -- ONE kernel for many small groups with few features (p <= 8): rows -> results.
//
// Covers, for the groups that fit a shared-memory stage, the whole reference path
//   *AggUpdate / *AggCombine / *AggFinalize   src/aggregate_functions/ols_aggregate.cpp:120-338
//   anofox_{ols,ridge,wls}_fit                crates/anofox-stats-ffi/src/lib.rs:98-265, :984, :1384
//   fit_ols / fit_ridge / fit_wls             crates/anofox-stats-core/src/models/{ols,ridge,wls}.rs
//   (+ the external regressor's statistics and compute_aic / compute_bic)
// without ever writing a per-group record to HBM: each input row is read from HBM once (TMA bulk
// copy into a shared-memory stage, producer_loop in accumulate.cuh) and only the result struct is
// written back.
//
// Work mapping. A group is owned by T consecutive lanes of a warp (T = 4, 8, 16 or 32, chosen by the
// host from the mean group size so that enough groups are resident per SM). Lane q of the T takes
// rows q, q+T, ... and keeps the full upper triangle of sum w d d^T (d = row - first valid row) in
// registers -- plain DFMA, every lane busy, no per-group prologue at warp granularity -- then a
// log2(T)-step butterfly leaves the sums in all T lanes. The T lanes run the tiny Cholesky solve
// redundantly (it is the same instruction stream) and split the per-coefficient distribution
// functions (the expensive part: t p-values) between them. Rows are still in shared memory at that
// point, so the exact residual pass that the moment form needs for (nearly) exact fits happens in
// place instead of through a refinement queue.
//
// Rounds of 32/T groups are claimed dynamically from the stage by the consumer warps.
#pragma once
#include "accumulate.cuh"
#include "solve.cuh"

namespace b200 {

template <int P> struct SmallMoments {
	static constexpr int K = P + 1;
	static constexpr int NS2 = K * (K + 1) / 2;
	double s2[NS2]; // sum w d d^T, upper triangle row-major (i <= j), index K-1 = y
	double s1[K];   // sum w d
	double sw;      // sum w over valid rows (weighted fits only; otherwise nv)
	double q2[P];   // unweighted sum d^2 per feature (weighted fits only; otherwise the diagonal of s2)
	int nv;         // valid rows
	int ns;         // rows that entered the aggregate state (training flag set)
	__device__ __forceinline__ void clear() {
#pragma unroll
		for (int i = 0; i < NS2; ++i) s2[i] = 0.0;
#pragma unroll
		for (int i = 0; i < K; ++i) s1[i] = 0.0;
#pragma unroll
		for (int i = 0; i < P; ++i) q2[i] = 0.0;
		sw = 0.0;
		nv = ns = 0;
	}
};

__host__ __device__ constexpr int utri(int i, int j, int k) { return i * k - (i * (i - 1)) / 2 + (j - i); } // i <= j
__host__ __device__ constexpr int ltri(int i, int j) { return (i * (i + 1)) / 2 + j; }                       // j <= i

__device__ __forceinline__ double shfl_xor_f64(unsigned mask, double v, int off) {
	return __shfl_xor_sync(mask, v, off);
}

// Rows q, q+T, ... of one group (shared-memory addresses in sg) folded into M.
//   CHECK = false: fast path. Rows are masked only by w > 0 / the training flag; non-finite values
//                  poison the sums and the caller re-runs with CHECK = true.
//   CHECK = true : the full row filter of ols.rs:59-66 / wls.rs:76-86.
//   ROT (P a power of two): slot c of the lane holds column (c + rot) mod P, rot = small_rot<P>(q). A row of
//                  P = 8 doubles is 16 banks wide, so with every lane on column c the 16 lanes of a half-warp
//                  meet in 2 bank pairs (8-way conflict); walking the columns in an order shifted by q / 2 spreads
//                  them over all 32 banks. sh is in slot order too; the caller undoes the rotation of the sums
//                  (small_unrotate) before anything combines the lanes.
template <int P> __device__ __forceinline__ int small_rot(int q) { return ((q * 2 * P) >> 5) & (P - 1); }
template <int P, bool HASW, bool HAST, bool CHECK, bool ROT = false>
__device__ __forceinline__ void small_accumulate(SmallMoments<P> &M, const SmemGroup &sg, int n, int q, int T,
                                                 const double (&sh)[P + 1]) {
	constexpr int K = P + 1;
	static_assert(!ROT || (P & (P - 1)) == 0, "rotated column order: P must be a power of two");
	uint32_t px = sg.x + (uint32_t)(q * P) * 8u, py = sg.y + (uint32_t)q * 8u;
	uint32_t pw = sg.w + (uint32_t)q * 8u, pt = sg.t + (uint32_t)q;
	const uint32_t stepx = (uint32_t)(T * P) * 8u, stepy = (uint32_t)T * 8u;
	constexpr bool MASK = HASW || HAST || CHECK;
	uint32_t co[P]; // byte offset of slot c's column inside the row
#pragma unroll
	for (int c = 0; c < P; ++c) co[c] = ROT ? (uint32_t)((c + small_rot<P>(q)) & (P - 1)) * 8u : (uint32_t)c * 8u;
#pragma unroll 2
	for (int i = q; i < n; i += T) {
		double d[K];
#pragma unroll
		for (int c = 0; c < P; ++c) d[c] = ROT ? lds_f64(px + co[c]) : lds_f64(px + (uint32_t)c * 8u);
		d[P] = lds_f64(py);
		px += stepx;
		py += stepy;
		bool ok = true;
		double wv = 1.0;
		if (HASW) {
			wv = lds_f64(pw);
			pw += stepy;
			ok = wv > 0.0;
			if (CHECK) ok = ok && finite_f64(wv);
		}
		if (HAST) {
			const uint32_t tr = lds_u8(pt);
			pt += (uint32_t)T;
			ok = ok && tr != 0u;
			M.ns += tr != 0u ? 1 : 0;
		}
		if (CHECK) {
#pragma unroll
			for (int c = 0; c < K; ++c) ok = ok && finite_f64(d[c]);
		}
#pragma unroll
		for (int c = 0; c < K; ++c) d[c] -= sh[c];
		if (!MASK || ok) {
			if (MASK) M.nv += 1;
			double a[K];
#pragma unroll
			for (int c = 0; c < K; ++c) a[c] = HASW ? wv * d[c] : d[c];
#pragma unroll
			for (int i2 = 0; i2 < K; ++i2)
#pragma unroll
				for (int j2 = i2; j2 < K; ++j2) M.s2[utri(i2, j2, K)] = fma(a[i2], d[j2], M.s2[utri(i2, j2, K)]);
#pragma unroll
			for (int c = 0; c < K; ++c) M.s1[c] += a[c];
			if (HASW) {
				M.sw += wv;
#pragma unroll
				for (int c = 0; c < P; ++c) M.q2[c] = fma(d[c], d[c], M.q2[c]);
			}
		}
	}
}

// Undo the slot order of a ROT accumulation: slot c held column (c + rot) mod P, so column t sits in slot (t - rot) mod P.
// rot = sum of its bits s: one conditional, compile-time permutation per bit (new[t] = old[(t - s) mod P]).
template <int P, int S, bool HASW> __device__ __forceinline__ void small_unrotate_step(SmallMoments<P> &M, bool on) {
	constexpr int K = P + 1;
	SmallMoments<P> R;
#pragma unroll
	for (int i = 0; i < P; ++i) {
		const int si = (i - S) & (P - 1);
#pragma unroll
		for (int j = i; j < P; ++j) {
			const int sj = (j - S) & (P - 1);
			R.s2[utri(i, j, K)] = M.s2[utri(si < sj ? si : sj, si < sj ? sj : si, K)];
		}
		R.s2[utri(i, P, K)] = M.s2[utri(si, P, K)];
		R.s1[i] = M.s1[si];
		R.q2[i] = M.q2[si];
	}
#pragma unroll
	for (int i = 0; i < P; ++i) {
#pragma unroll
		for (int j = i; j <= P; ++j) M.s2[utri(i, j, K)] = on ? R.s2[utri(i, j, K)] : M.s2[utri(i, j, K)];
		M.s1[i] = on ? R.s1[i] : M.s1[i];
		if (HASW) M.q2[i] = on ? R.q2[i] : M.q2[i];
	}
}
template <int P, bool HASW> __device__ __forceinline__ void small_unrotate(SmallMoments<P> &M, int rot) {
	if (P >= 2) small_unrotate_step<P, 1 % (P > 1 ? P : 2), HASW>(M, (rot & 1) != 0);
	if (P >= 4) small_unrotate_step<P, 2 % (P > 2 ? P : 4), HASW>(M, (rot & 2) != 0);
	if (P >= 8) small_unrotate_step<P, 4 % (P > 4 ? P : 8), HASW>(M, (rot & 4) != 0);
}

// Butterfly over the T lanes of a group: afterwards every lane holds the group's sums.
template <int P, bool HASW, bool HAST> __device__ __forceinline__ void small_reduce(SmallMoments<P> &M, int T, unsigned mask) {
	constexpr int K = P + 1;
	for (int off = 1; off < T; off <<= 1) {
#pragma unroll
		for (int i = 0; i < SmallMoments<P>::NS2; ++i) M.s2[i] += shfl_xor_f64(mask, M.s2[i], off);
#pragma unroll
		for (int i = 0; i < K; ++i) M.s1[i] += shfl_xor_f64(mask, M.s1[i], off);
		if (HASW) {
			M.sw += shfl_xor_f64(mask, M.sw, off);
#pragma unroll
			for (int i = 0; i < P; ++i) M.q2[i] += shfl_xor_f64(mask, M.q2[i], off);
		}
		M.nv += __shfl_xor_sync(mask, M.nv, off);
		if (HAST) M.ns += __shfl_xor_sync(mask, M.ns, off);
	}
}

// Split reduction for the accumulate-to-record path. The record's sums (S1 | S2, NSUM of them) are not
// needed in every lane, only stored once: after the butterfly steps that cross the 4-lane quads
// (offsets >= 4, T > 4 only), the last two steps exchange HALVES, so that lane q ends up with the
// complete sums j = 4k + (q & 3) in keep[k] -- the same shuffle count as a full butterfly but a
// quarter of the stores afterwards, each with a compile-time field index.
template <int P> struct SmallSplit {
	static constexpr int K = P + 1;
	static constexpr int NSUM = K + K * (K + 1) / 2;
	static constexpr int NQ = (NSUM + 3) / 4;
};
template <int P> __device__ __forceinline__ double small_sum_at(const SmallMoments<P> &M, int j) {
	return j < P + 1 ? M.s1[j] : M.s2[j - (P + 1)];
}
template <int P, bool HASW, bool HAST>
__device__ __forceinline__ void small_reduce_split(SmallMoments<P> &M, int T, int q, unsigned mask,
                                                   double (&keep)[SmallSplit<P>::NQ]) {
	constexpr int K = P + 1;
	constexpr int NSUM = SmallSplit<P>::NSUM, NQ = SmallSplit<P>::NQ;
	for (int off = T >> 1; off >= 4; off >>= 1) {
#pragma unroll
		for (int i = 0; i < SmallMoments<P>::NS2; ++i) M.s2[i] += shfl_xor_f64(mask, M.s2[i], off);
#pragma unroll
		for (int i = 0; i < K; ++i) M.s1[i] += shfl_xor_f64(mask, M.s1[i], off);
	}
	for (int off = 1; off < T; off <<= 1) {
		if (HASW) {
			M.sw += shfl_xor_f64(mask, M.sw, off);
#pragma unroll
			for (int i = 0; i < P; ++i) M.q2[i] += shfl_xor_f64(mask, M.q2[i], off);
		}
		M.nv += __shfl_xor_sync(mask, M.nv, off);
		if (HAST) M.ns += __shfl_xor_sync(mask, M.ns, off);
	}
	const bool b1 = (q & 2) != 0, b0 = (q & 1) != 0;
#pragma unroll
	for (int k = 0; k < NQ; ++k) {
		const double v0 = small_sum_at<P>(M, 4 * k);
		const double v1 = 4 * k + 1 < NSUM ? small_sum_at<P>(M, 4 * k + 1) : 0.0;
		const double v2 = 4 * k + 2 < NSUM ? small_sum_at<P>(M, 4 * k + 2) : 0.0;
		const double v3 = 4 * k + 3 < NSUM ? small_sum_at<P>(M, 4 * k + 3) : 0.0;
		const double u0 = (b1 ? v2 : v0) + shfl_xor_f64(mask, b1 ? v0 : v2, 2);
		const double u1 = (b1 ? v3 : v1) + shfl_xor_f64(mask, b1 ? v1 : v3, 2);
		keep[k] = (b0 ? u1 : u0) + shfl_xor_f64(mask, b0 ? u0 : u1, 1);
	}
}

// Shared-memory view of one round's group for this lane.
struct SmallGroupView {
	SmemGroup sg;
	int n;     // rows of the group
	int q, T;  // lane within the group, lanes per group
	unsigned mask; // the T lanes of the group
	double *pv;    // this group's slot of the warp's p-value scratch: f[P+1], d1, d2 (see fit_small_kernel)
	// thread-per-group mode (solve_small_kernel): no rows at hand -- groups that need the exact residual
	// pass are queued for refine_kernel instead
	int *refine_count;
	int64_t *refine_list;
	// stash kernel: no record of the group exists in HBM -- a group queued for refine_kernel gets its standard
	// record written here (array-of-records layout, rec_len(P) doubles per group) at the moment it is queued
	double *rec_out;
	double *stash; // stash kernel: this group's slot (field f at stash[f * 32])
	// resident kernel (small_group<..., TEAM = 2>): two warps share the accumulation of one group. Warp team_w = 1 folds
	// rows 32 + lane, 96 + lane, ..., leaves its warp-reduced sums in team_x (shared memory, kSmallTeamDoubles) and
	// returns; warp 0 adds them after the named barrier team_bar (64 threads) and carries on alone.
	int team_w, team_bar;
	double *team_x;
	// resident kernel: where small_solve leaves what the prediction pass needs (shared memory; null = nowhere):
	// [0] status, [1] intercept (NaN = none), [2] residual standard error, [3] n_observations, [4 .. 4+P) coefficients,
	// [4+P] r_squared. With fit_x set and o.status null (window kernel) nothing is written to the output arrays.
	double *fit_x = nullptr;
	// thread-per-group solve: leave the p-values to pvalues_deferred_kernel (groupfit.cu). The slots then carry what it
	// needs: p_values[g*P+j] = residual degrees of freedom (NaN for a dead column), f_pvalue[g] = model degrees of freedom.
	bool defer_p = false;
};
constexpr int kSmallTeamDoubles = 72; // >= NS2 + K + 1 + P + 2 for P = 8
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
	asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// First row (in state order) that passes `pred(row)`, searched by the T lanes together; -1 if none.
template <class Pred> __device__ __forceinline__ int small_first_row(const SmallGroupView &v, Pred pred) {
	int f = 0x7fffffff;
	for (int i = v.q; i < v.n; i += v.T)
		if (pred(i)) {
			f = i;
			break;
		}
	for (int off = 1; off < v.T; off <<= 1) f = min(f, __shfl_xor_sync(v.mask, f, off));
	return f == 0x7fffffff ? -1 : f;
}

template <int P, bool HASW, bool HAST> __device__ __forceinline__ bool small_row_valid_full(const SmallGroupView &v, int i) {
	bool ok = finite_f64(lds_f64(v.sg.y + (uint32_t)i * 8u));
	if (HASW) {
		const double wv = lds_f64(v.sg.w + (uint32_t)i * 8u);
		ok = ok && wv > 0.0 && finite_f64(wv);
	}
	if (HAST) ok = ok && lds_u8(v.sg.t + (uint32_t)i) != 0u;
#pragma unroll
	for (int c = 0; c < P; ++c) ok = ok && finite_f64(lds_f64(v.sg.x + (uint32_t)(i * P + c) * 8u));
	return ok;
}

// Exact constant-column statistic for one feature: max |x - shift| over the valid rows (ols.rs:76-87).
template <int P, bool HASW, bool HAST>
__device__ __forceinline__ double small_exact_cdev(const SmallGroupView &v, int c, double shc) {
	double m = 0.0;
	for (int i = v.q; i < v.n; i += v.T)
		if (small_row_valid_full<P, HASW, HAST>(v, i))
			m = fmax(m, fabs(lds_f64(v.sg.x + (uint32_t)(i * P + c) * 8u) - shc));
	for (int off = 1; off < v.T; off <<= 1) m = fmax(m, shfl_xor_f64(v.mask, m, off));
	return m;
}

// sum w (y - b0 - x'beta)^2 over the valid rows with the coefficients just computed (NaN = aliased:
// contributes nothing, lib.rs:2296-2303). With an intercept (shifted = true) the residual is evaluated about the
// shift row: e = (y - sh_y) - b0s - sum beta (x - sh), every term of the size of the data's variation.
template <int P, bool HASW, bool HAST>
__device__ __forceinline__ double small_exact_rss(const SmallGroupView &v, const double (&beta)[P], unsigned live,
                                                  double b0s, const double (&sh)[P + 1], bool shifted) {
	double rss = 0.0;
	for (int i = v.q; i < v.n; i += v.T) {
		if (!small_row_valid_full<P, HASW, HAST>(v, i)) continue;
		double fit = b0s;
#pragma unroll
		for (int c = 0; c < P; ++c)
			if ((live >> c) & 1u) {
				const double xv = lds_f64(v.sg.x + (uint32_t)(i * P + c) * 8u);
				fit = fma(beta[c], shifted ? xv - sh[c] : xv, fit);
			}
		const double yv = lds_f64(v.sg.y + (uint32_t)i * 8u);
		const double e = (shifted ? yv - sh[P] : yv) - fit;
		const double wv = HASW ? lds_f64(v.sg.w + (uint32_t)i * 8u) : 1.0;
		rss = fma(wv * e, e, rss);
	}
	for (int off = 1; off < v.T; off <<= 1) rss += shfl_xor_f64(v.mask, rss, off);
	return rss;
}

// The group's sums as a standard moment record (common.cuh layout, element stride 1): what refine_kernel and
// solve_group read. Only the stash kernel needs it, and only for the rare groups it queues.
template <int P>
__device__ __noinline__ void small_write_std_record(const SmallMoments<P> &M, const double (&sh)[P + 1], unsigned nonconst,
                                                    int n_state, double sw, double *__restrict__ rec) {
	constexpr int K = P + 1;
	rec[0] = (double)M.nv;
	rec[1] = sw;
	rec[2] = (double)n_state;
	rec[3] = 0.0;
	for (int c = 0; c < K; ++c) rec[rec_off_shift(P) + c] = sh[c];
	for (int c = 0; c < P; ++c) rec[rec_off_cdev(P) + c] = ((nonconst >> c) & 1u) ? 1.0 : 0.0;
	for (int c = rec_off_cdev(P) + P; c < rec_off_s1(P); ++c) rec[c] = 0.0;
	for (int c = 0; c < K; ++c) rec[rec_off_s1(P) + c] = M.s1[c];
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) rec[rec_off_s2(P) + i] = M.s2[i];
}

// Solve + statistics of one group from its sums, executed by the T lanes of the group together.
// Mirrors solve_group (solve.cuh) formula for formula; differences are structural only: no column
// compaction (constant columns are simply dead from the start), reciprocals of the Cholesky diagonal,
// distribution functions split across the lanes.
template <int P, bool HASW, bool HAST, bool ROWS>
__device__ __forceinline__ void small_solve(const SmallMoments<P> &M, const double (&sh)[P + 1], unsigned nonconst,
                                            int n_state, int64_t g, const FitParams &Pm, const AnofoxB200FitOutputs &o,
                                            const SmallGroupView &v) {
	constexpr int K = P + 1;
	const double nanv = nan_f64();
	const int ic = Pm.fit_intercept ? 1 : 0;
	const bool wls = Pm.method == ANOFOX_B200_WLS;
	const int q = v.q, T = v.T;
	const double n_valid = (double)M.nv;
	const double sw = (HASW || !ROWS) ? M.sw : n_valid;
	int status = 0;
	if (n_state < Pm.min_state_rows)
		status = ANOFOX_B200_STATUS_TOO_FEW_ROWS; // ols_aggregate.cpp:263-267
	else if (M.nv <= 0)
		status = ANOFOX_ERROR_NO_VALID_DATA; // ols.rs:68-70
	const int pe = __popc(nonconst);
	if (status == 0) {
		if (pe == 0 && !ic) status = ANOFOX_ERROR_INSUFFICIENT_DATA;                // ols.rs:102-107
		else if (pe > 0 && M.nv < pe + ic) status = ANOFOX_ERROR_INSUFFICIENT_DATA; // ols.rs:134-139
	}
	const bool wr = o.status != nullptr; // false: results go to v.fit_x only
	if (status != 0) {
		if (q == 0 && wr) write_null_group(o, Pm, g, status);
		if (q == 0 && v.fit_x) v.fit_x[0] = (double)status;
		return;
	}
	const double inv_sw = 1.0 / sw;
	const double s1y = M.s1[P];
	const double ybar = sh[P] + s1y * inv_sw;
	const double syy_c = M.s2[utri(P, P, K)] - s1y * s1y * inv_sw;
	const double tss = ic ? syy_c : syy_c + sw * ybar * ybar;
	const bool inf_arrays = Pm.compute_inference && o.std_errors;

	if (pe == 0) {
		// all columns constant: intercept-only shortcut (ols.rs:108-130, wls.rs:126-150)
		if (q == 0 && !wr) {
			const double ss = fmax(syy_c, 0.0);
			const double var = wls ? ss * inv_sw : ss / (n_valid - 1.0);
			if (v.fit_x) {
				v.fit_x[0] = 0.0;
				v.fit_x[1] = ybar;
				v.fit_x[2] = sqrt(var);
				v.fit_x[3] = (double)M.nv;
				for (int j = 0; j < P; ++j) v.fit_x[4 + j] = nanv;
				v.fit_x[4 + P] = 0.0;
			}
		}
		if (q == 0 && wr) {
			const double ss = fmax(syy_c, 0.0);
			const double var = wls ? ss * inv_sw : ss / (n_valid - 1.0);
			o.status[g] = 0;
			o.n_observations[g] = (int64_t)M.nv;
			for (int j = 0; j < P; ++j) o.coefficients[g * P + j] = nanv;
			o.intercept[g] = ybar;
			o.r_squared[g] = 0.0;
			o.adj_r_squared[g] = 0.0;
			o.residual_std_error[g] = sqrt(var);
			if (inf_arrays) {
				for (int j = 0; j < P; ++j) {
					o.std_errors[g * P + j] = nanv;
					o.t_values[g * P + j] = nanv;
					o.p_values[g * P + j] = nanv;
					o.ci_lower[g * P + j] = nanv;
					o.ci_upper[g * P + j] = nanv;
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
			if (v.fit_x) {
				v.fit_x[0] = 0.0;
				v.fit_x[1] = ybar;
				v.fit_x[2] = sqrt(var);
				v.fit_x[3] = (double)M.nv;
				for (int j = 0; j < P; ++j) v.fit_x[4 + j] = nanv;
				v.fit_x[4 + P] = 0.0;
			}
		}
		return;
	}

	// ---- centred (or raw, without intercept) Gram of all P columns; constant ones start dead ----
	double A[P * (P + 1) / 2], b[P], orig[P], xm[P];
#pragma unroll
	for (int a = 0; a < P; ++a) {
		const double s1a = M.s1[a];
		const double ma = sh[a] + s1a * inv_sw;
		xm[a] = ma;
#pragma unroll
		for (int c = 0; c <= a; ++c) {
			double val = M.s2[utri(c, a, K)] - s1a * M.s1[c] * inv_sw;
			if (!ic) val += sw * ma * xm[c];
			A[ltri(a, c)] = val;
		}
		double cxy = M.s2[utri(a, P, K)] - s1a * s1y * inv_sw;
		if (!ic) cxy += sw * ma * ybar;
		b[a] = cxy;
		const double caa = M.s2[utri(a, a, K)] - s1a * s1a * inv_sw;
		orig[a] = caa + sw * ma * ma;
	}
	double lambda = 0.0;
	if (Pm.method == ANOFOX_B200_RIDGE) {
		lambda = Pm.alpha;
		if (Pm.lambda_scaling == ANOFOX_LAMBDA_SCALING_GLMNET) {
			const double sdy = sqrt(fmax(tss, 0.0) / n_valid);
			lambda = Pm.alpha * n_valid;
			if (sdy > 0.0) lambda /= sdy;
		}
	}
	const bool ridge = lambda > 0.0;
	if (ridge) {
#pragma unroll
		for (int a = 0; a < P; ++a) A[ltri(a, a)] += lambda;
	}

	// ---- Cholesky, left to right, limited pivoting (solve.cuh) --------------------------------
	unsigned live = 0u;  // columns with a coefficient
	const bool want_qr = Pm.qr_count != nullptr; // no queue (window form, moment-only entry points): skip the probe
	double rinv[P];      // 1 / L_ii
	int rank = 0;
	bool suspicious = false;
#pragma unroll
	for (int i = 0; i < P; ++i) {
		const double a0 = A[ltri(i, i)];
#pragma unroll
		for (int j = 0; j < i; ++j) {
			double s = A[ltri(i, j)];
#pragma unroll
			for (int c = 0; c < j; ++c) s -= A[ltri(i, c)] * A[ltri(j, c)];
			A[ltri(i, j)] = ((live >> j) & 1u) ? s * rinv[j] : 0.0;
		}
		double d = A[ltri(i, i)];
#pragma unroll
		for (int c = 0; c < i; ++c) d -= A[ltri(i, c)] * A[ltri(i, c)];
		bool alias = ridge ? !(d > 0.0) : (!(d >= 1e-14 * orig[i]) || !(d > 0.0));
		if (want_qr && ((nonconst >> i) & 1u)) suspicious = suspicious || chol_pivot_suspicious(d, a0, orig[i], ridge);
		alias = alias || !((nonconst >> i) & 1u);
		if (alias) {
#pragma unroll
			for (int c = 0; c < i; ++c) A[ltri(i, c)] = 0.0;
			A[ltri(i, i)] = 0.0;
			rinv[i] = 0.0;
		} else {
			const double r = rsqrt(d);
			A[ltri(i, i)] = d * r;
			rinv[i] = r;
			live |= 1u << i;
			++rank;
		}
	}
	if (suspicious && q == 0) queue_qr(Pm, g);
	// forward solve L v = b, back solve L^T beta = v (dead columns have zero rows/columns and rinv = 0)
	double ess = 0.0;
#pragma unroll
	for (int i = 0; i < P; ++i) {
		double s = b[i];
#pragma unroll
		for (int c = 0; c < i; ++c) s -= A[ltri(i, c)] * b[c];
		s *= rinv[i];
		b[i] = s;
		ess = fma(s, s, ess);
	}
#pragma unroll
	for (int i = P - 1; i >= 0; --i) {
		double s = b[i];
#pragma unroll
		for (int c = i + 1; c < P; ++c) s -= A[ltri(c, i)] * b[c];
		b[i] = s * rinv[i];
	}
	double bnorm2 = 0.0, b0 = ybar;
#pragma unroll
	for (int i = 0; i < P; ++i) {
		bnorm2 = fma(b[i], b[i], bnorm2);
		b0 -= xm[i] * b[i];
	}
	double rss = tss - ess - (ridge ? lambda * bnorm2 : 0.0);
	if (!(rss > 1e-5 * tss)) {
		// tss - ess keeps ~16 - log10(tss/rss) digits: take the residuals from the rows themselves
		if (ROWS) {
			double b0s = 0.0; // intercept in the shifted frame: (S1_y - sum beta S1_j) / sum w
			if (ic) {
				b0s = s1y;
#pragma unroll
				for (int i = 0; i < P; ++i) b0s -= b[i] * M.s1[i];
				b0s *= inv_sw;
			}
			rss = small_exact_rss<P, HASW, HAST>(v, b, live, b0s, sh, ic != 0);
		} else if (v.refine_count) {
			const int slot = atomicAdd(v.refine_count, 1);
			v.refine_list[slot] = g;
			if (v.rec_out) small_write_std_record<P>(M, sh, nonconst, n_state, sw, v.rec_out + g * rec_len(P));
		}
	}
	if (!(rss > 0.0)) rss = 0.0;

	const int rank_total = rank + ic;
	const double dfr = n_valid - (double)rank_total;
	const double dfm = (double)rank;
	const double r2 = 1.0 - rss / tss;
	const double sigma2 = dfr > 0.0 ? rss / dfr : nanv;
	if (q == 0) {
		if (wr) {
			o.status[g] = 0;
			o.n_observations[g] = (int64_t)M.nv;
			o.intercept[g] = ic ? b0 : nanv; // ols.rs:172-176
			o.r_squared[g] = r2;
			o.adj_r_squared[g] = dfr > 0.0 ? 1.0 - (1.0 - r2) * (n_valid - (double)ic) / dfr : nanv;
			o.residual_std_error[g] = sqrt(sigma2);
			if (o.rss) o.rss[g] = rss;
		}
		if (v.fit_x) {
			v.fit_x[4 + P] = r2;
			v.fit_x[0] = 0.0;
			v.fit_x[1] = ic ? b0 : nanv;
			v.fit_x[2] = sqrt(sigma2);
			v.fit_x[3] = (double)M.nv;
#pragma unroll
			for (int j = 0; j < P; ++j) v.fit_x[4 + j] = ((live >> j) & 1u) ? b[j] : nanv;
		}
	}
	if (q == (1 & (T - 1)) && (o.aic || o.bic)) {
		double aic, bic;
		info_criteria_dev(rss, n_valid, (double)rank_total, aic, bic);
		if (o.aic) o.aic[g] = aic;
		if (o.bic) o.bic[g] = bic;
	}
	if (!inf_arrays) {
		if (wr) {
#pragma unroll
			for (int j = 0; j < P; ++j)
				if ((j & (T - 1)) == q) o.coefficients[g * P + j] = ((live >> j) & 1u) ? b[j] : nanv;
		}
		return;
	}

	// ---- inference ------------------------------------------------------------------------------
	// diag((L L^T)^-1)_j = sum_i (L^-1)_ij^2, all columns in every lane (P^3/6 multiply-adds) ...
	double vdiag[P];
#pragma unroll
	for (int j = 0; j < P; ++j) {
		double x[P];
		double acc = 0.0;
#pragma unroll
		for (int i = j; i < P; ++i) {
			double s = (i == j) ? 1.0 : 0.0;
#pragma unroll
			for (int c = j; c < i; ++c) s -= A[ltri(i, c)] * x[c];
			s *= rinv[i];
			x[i] = s;
			acc = fma(s, s, acc);
		}
		vdiag[j] = acc;
	}
	double tcrit = nanv;
	if (dfr > 0.0) {
		const int idf = (int)dfr;
		tcrit = (Pm.tcrit && idf <= kTcritTable) ? Pm.tcrit[idf] : t_critical_dev(Pm.level, dfr);
	}
	// ... then lane q takes coefficients q, q+T, ...: the same instruction stream on different data.
	// The p-values themselves (the expensive part) are evaluated after the round by all 32 lanes of the
	// warp from the statistics deposited in v.pv: t^2 per coefficient and the F statistic.
	for (int j0 = 0; j0 < P; j0 += T) {
		const int j = j0 + q;
		if (j < P) {
			double bj = 0.0, vj = 0.0;
#pragma unroll
			for (int c = 0; c < P; ++c)
				if (c == j) {
					bj = b[c];
					vj = vdiag[c];
				}
			const int64_t oj = g * P + j;
			if ((live >> j) & 1u) {
				const double se = sqrt(sigma2 * vj);
				const double t = bj / se;
				o.coefficients[oj] = bj;
				o.std_errors[oj] = se;
				o.t_values[oj] = t;
				o.ci_lower[oj] = bj - tcrit * se;
				o.ci_upper[oj] = bj + tcrit * se;
				if (ROWS)
					v.pv[j] = t * t;
				else
					o.p_values[oj] = v.defer_p ? dfr : pvalue_f(t * t, 1.0, dfr, DfTables{Pm.tcrit, Pm.gratio, Pm.rcp, Pm.ce, Pm.co});
			} else {
				o.coefficients[oj] = nanv;
				o.std_errors[oj] = nanv;
				o.t_values[oj] = nanv;
				o.ci_lower[oj] = nanv;
				o.ci_upper[oj] = nanv;
				if (ROWS)
					v.pv[j] = nanv;
				else
					o.p_values[oj] = nanv;
			}
		}
	}
	if (q == (2 & (T - 1))) {
		double fs = nanv;
		if (dfm > 0.0 && dfr > 0.0) fs = ((tss - rss) / dfm) / sigma2;
		o.f_statistic[g] = fs;
		if (o.has_inference) o.has_inference[g] = 1;
		if (ROWS) {
			v.pv[P] = fs;
			v.pv[P + 1] = dfm; // > 0 marks "p-values wanted" for this slot
			v.pv[P + 2] = dfr;
		} else {
			o.f_pvalue[g] = v.defer_p ? dfm : pvalue_f(fs, dfm, dfr, DfTables{Pm.tcrit, Pm.gratio, Pm.rcp, Pm.ce, Pm.co});
		}
	}
}

// One round: lanes [sub*T, sub*T+T) own group `gi` of the stage (valid = false: idle lanes).
// Transposed moment record (element stride rs): same field order as the record of common.cuh, so that the
// generic solve_group / refine_kernel can read it too. Header fields (replicated in the lanes of the group)
// are stored by lane f % 4; the sums come from small_reduce_split: lane q holds sum 4k + q in keep[k].
template <int P>
__device__ __forceinline__ void small_store_record(const double (&keep)[SmallSplit<P>::NQ], int nv, double sw,
                                                   const double (&sh)[P + 1], unsigned nonconst, int n_state,
                                                   double *__restrict__ rec, int64_t rs, int q) {
	constexpr int K = P + 1;
	constexpr int NSUM = SmallSplit<P>::NSUM, NQ = SmallSplit<P>::NQ;
	if (q >= 4) return; // lanes 4.. of a wider group hold copies
	int f = 0;
	auto put = [&](double val) {
		if ((f & 3) == q) rec[(int64_t)f * rs] = val;
		++f;
	};
	put((double)nv);
	put(sw);
	put((double)n_state);
	put(0.0);
#pragma unroll
	for (int c = 0; c < K; ++c) put(sh[c]);
#pragma unroll
	for (int c = 0; c < P; ++c) put(((nonconst >> c) & 1u) ? 1.0 : 0.0); // cdev: only "< 1e-10" is ever asked
	double *recq = rec + (int64_t)(((4 + K + P + 3) & ~3) + q) * rs; // rec_off_s1(P) + q
#pragma unroll
	for (int k = 0; k < NQ; ++k)
		if (4 * k + 3 < NSUM || 4 * k + q < NSUM) recq[(int64_t)(4 * k) * rs] = keep[k];
}

// ---- stash (fit_small_stash_kernel) -----------------------------------------------------------------------
// A warp parks the sums of the groups it has accumulated in a private shared-memory table, field-major with
// 32 slots per field (slot = the lane that will solve the group), until 32 groups are waiting; then every lane
// solves one group with all arrays in registers (small_solve, thread per group). Fields of a slot:
//   g | nv, n_state, nonconst packed | [sum w] | shift[K] | S1[K] | S2[K(K+1)/2]
template <int P, bool HASW> struct SmallStash {
	static constexpr int K = P + 1;
	static constexpr int NS2 = K * (K + 1) / 2;
	static constexpr int NF = 2 + (HASW ? 1 : 0) + 2 * K + NS2;
	static constexpr int BYTES = NF * 32 * 8;
};
// Every lane of the group holds all the sums (full butterfly); lane q stores the fields f = q (mod T).
template <int P, bool HASW>
__device__ __forceinline__ void small_stash_store(const SmallMoments<P> &M, const double (&sh)[P + 1], unsigned nonconst,
                                                  int n_state, int64_t g, double *__restrict__ st, int q, int T) {
	constexpr int K = P + 1;
	int f = 0;
	auto put = [&](double val) {
		if ((f & (T - 1)) == q) st[f * 32] = val;
		++f;
	};
	put(__longlong_as_double((long long)g));
	put(__longlong_as_double((long long)M.nv | ((long long)n_state << 21) | ((long long)nonconst << 42)));
	if (HASW) put(M.sw);
#pragma unroll
	for (int c = 0; c < K; ++c) put(sh[c]);
#pragma unroll
	for (int c = 0; c < K; ++c) put(M.s1[c]);
#pragma unroll
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) put(M.s2[i]);
}
// The sums parked in one slot (sl = table + slot) back into registers.
template <int P, bool HASW>
__device__ __forceinline__ void small_stash_load(const double *__restrict__ sl, SmallMoments<P> &M, double (&sh)[P + 1],
                                                 unsigned &nonconst, int &n_state, int64_t &g) {
	constexpr int K = P + 1;
	int f = 0;
	auto get = [&]() { return sl[(f++) * 32]; };
	g = (int64_t)__double_as_longlong(get());
	const long long pk = __double_as_longlong(get());
	M.nv = (int)(pk & 0x1fffff);
	n_state = (int)((pk >> 21) & 0x1fffff);
	nonconst = (unsigned)(pk >> 42) & 0xffu;
	M.ns = n_state;
	M.sw = HASW ? get() : (double)M.nv;
#pragma unroll
	for (int c = 0; c < K; ++c) sh[c] = get();
#pragma unroll
	for (int c = 0; c < K; ++c) M.s1[c] = get();
#pragma unroll
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) M.s2[i] = get();
#pragma unroll
	for (int c = 0; c < P; ++c) M.q2[c] = 0.0;
}

template <int P, bool HASW, bool HAST, int MODE, bool ROT = false, int TEAM = 1>
__device__ __forceinline__ void small_group(const SmallGroupView &v, int64_t g, const FitParams &Pm,
                                            const AnofoxB200FitOutputs &o, double *records, int64_t n_groups) {
	constexpr int K = P + 1;
	constexpr int NQ = SmallSplit<P>::NQ;
	constexpr bool FUSED = MODE != 0; // 1 and 2: every lane of the group ends up with all the sums
	SmallMoments<P> M;
	double keep[NQ];
	double sh[K];
#pragma unroll
	for (int c = 0; c < K; ++c) sh[c] = 0.0;
#pragma unroll
	for (int k = 0; k < NQ; ++k) keep[k] = 0.0;
	const int n = v.n;
	int n_state = n;
	bool have = n > 0;
	// reduce over the lanes of the group; true when a non-finite value poisoned the sums
	auto reduce = [&]() {
		bool bad = false;
		if (FUSED) {
			small_reduce<P, HASW, HAST>(M, v.T, v.mask);
#pragma unroll
			for (int c = 0; c < K; ++c) bad = bad || !finite_f64(M.s2[utri(c, c, K)]);
		} else {
			small_reduce_split<P, HASW, HAST>(M, v.T, v.q, v.mask, keep);
#pragma unroll
			for (int k = 0; k < NQ; ++k) bad = bad || !finite_f64(keep[k]);
			bad = __any_sync(v.mask, bad);
		}
		if (HASW) bad = bad || !finite_f64(M.sw);
		return bad;
	};
	// fast path: the shift row is the first row that passes the cheap mask (row 0 without masks)
	int f = 0;
	if (have && (HASW || HAST)) {
		f = small_first_row(v, [&](int i) {
			bool ok = true;
			if (HASW) ok = lds_f64(v.sg.w + (uint32_t)i * 8u) > 0.0;
			if (HAST) ok = ok && lds_u8(v.sg.t + (uint32_t)i) != 0u;
			return ok;
		});
	}
	bool redo = false;
	M.clear();
	// ROT: the lane's accumulation runs in slot order (small_accumulate) with the shift in the same order
	auto load_shift = [&](int frow, double (&dst)[K], bool slots) {
#pragma unroll
		for (int c = 0; c < P; ++c) {
			const int col = (ROT && slots) ? ((c + small_rot<P>(v.q)) & (P - 1)) : c;
			dst[c] = lds_f64(v.sg.x + (uint32_t)(frow * P + col) * 8u);
		}
		dst[P] = lds_f64(v.sg.y + (uint32_t)frow * 8u);
	};
	if (have && f >= 0) {
		load_shift(f, sh, true);
		small_accumulate<P, HASW, HAST, false, ROT>(M, v.sg, n, TEAM > 1 ? v.q + v.T * v.team_w : v.q, TEAM * v.T, sh);
		if (ROT) {
			small_unrotate<P, HASW>(M, small_rot<P>(v.q));
			load_shift(f, sh, false);
		}
		redo = reduce();
		if (!(HASW || HAST)) M.nv = n;
	} else if (have) {
		redo = true; // no row passes the mask: the checked path settles n_state / NO_VALID_DATA
	}
	if (TEAM > 1) {
		static_assert(TEAM <= 2 && (TEAM == 1 || (MODE == 1 && HAST)), "teams: fused mode with training flags, two warps");
		double *tx = v.team_x;
		if (v.team_w != 0) {
			if (v.q == 0) {
#pragma unroll
				for (int i = 0; i < SmallMoments<P>::NS2; ++i) tx[i] = M.s2[i];
#pragma unroll
				for (int i = 0; i < K; ++i) tx[SmallMoments<P>::NS2 + i] = M.s1[i];
				tx[SmallMoments<P>::NS2 + K] = M.sw;
#pragma unroll
				for (int i = 0; i < P; ++i) tx[SmallMoments<P>::NS2 + K + 1 + i] = M.q2[i];
				tx[SmallMoments<P>::NS2 + K + 1 + P] = (double)M.nv;
				tx[SmallMoments<P>::NS2 + K + 2 + P] = (double)M.ns;
			}
			named_bar_sync(v.team_bar, 64);
			return;
		}
		named_bar_sync(v.team_bar, 64);
#pragma unroll
		for (int i = 0; i < SmallMoments<P>::NS2; ++i) M.s2[i] += tx[i];
#pragma unroll
		for (int i = 0; i < K; ++i) M.s1[i] += tx[SmallMoments<P>::NS2 + i];
		M.sw += tx[SmallMoments<P>::NS2 + K];
#pragma unroll
		for (int i = 0; i < P; ++i) M.q2[i] += tx[SmallMoments<P>::NS2 + K + 1 + i];
		M.nv += (int)tx[SmallMoments<P>::NS2 + K + 1 + P];
		M.ns += (int)tx[SmallMoments<P>::NS2 + K + 2 + P];
#pragma unroll
		for (int c = 0; c < K; ++c) redo = redo || !finite_f64(M.s2[utri(c, c, K)]);
		if (HASW) redo = redo || !finite_f64(M.sw);
	}
	if (redo) {
		// a non-finite value somewhere: the full row filter of ols.rs:59-66
		f = small_first_row(v, [&](int i) { return small_row_valid_full<P, HASW, HAST>(v, i); });
		M.clear();
		if (f >= 0) {
			load_shift(f, sh, true);
			small_accumulate<P, HASW, HAST, true, ROT>(M, v.sg, n, v.q, v.T, sh);
			if (ROT) {
				small_unrotate<P, HASW>(M, small_rot<P>(v.q));
				load_shift(f, sh, false);
			}
		} else if (HAST) {
			for (int i = v.q; i < n; i += v.T) M.ns += lds_u8(v.sg.t + (uint32_t)i) != 0u ? 1 : 0;
		}
		reduce();
		if (f < 0) {
#pragma unroll
			for (int c = 0; c < K; ++c) sh[c] = 0.0;
		}
	}
	if (HAST) n_state = M.ns;
	// constant columns: all |x - shift| < 1e-10  <=>  max |d| < 1e-10. sum d^2 brackets max d^2
	// (max d^2 <= sum d^2 <= n max d^2); only the band in between needs the exact statistic.
	unsigned nonconst = 0u;
	if (M.nv > 0) {
		unsigned bits = 0u; // low half: surely not constant; high half: in the band
#pragma unroll
		for (int c = 0; c < P; ++c) {
			constexpr int dummy = 0;
			(void)dummy;
			const int jd = K + utri(c, c, K); // position of the diagonal sum in S1 | S2
			double s;
			bool mine = true;
			if (HASW)
				s = M.q2[c];
			else if (FUSED)
				s = M.s2[utri(c, c, K)];
			else {
				s = keep[jd >> 2];
				mine = (v.q & 3) == (jd & 3);
			}
			if (mine) {
				if (s > 1.01e-20 * (double)M.nv)
					bits |= 1u << c;
				else if (!(s < 0.99e-20))
					bits |= 0x10000u << c;
			}
		}
		if (!FUSED && !HASW) {
			bits |= __shfl_xor_sync(v.mask, bits, 1);
			bits |= __shfl_xor_sync(v.mask, bits, 2);
		}
		nonconst = bits & 0xffffu;
		if (bits >> 16) {
#pragma unroll
			for (int c = 0; c < P; ++c)
				if ((bits >> (16 + c)) & 1u)
					nonconst |= !(small_exact_cdev<P, HASW, HAST>(v, c, sh[c]) < 1e-10) ? (1u << c) : 0u;
		}
	}
	if (MODE == 2)
		small_stash_store<P, HASW>(M, sh, nonconst, n_state, g, v.stash, v.q, v.T);
	else if (MODE == 1)
		small_solve<P, HASW, HAST, true>(M, sh, nonconst, n_state, g, Pm, o, v);
	else
		small_store_record<P>(keep, M.nv, HASW ? M.sw : (double)M.nv, sh, nonconst, n_state, records + g, n_groups, v.q);
}

// One thread per group: moments from the transposed record (coalesced: consecutive lanes read consecutive
// groups of the same field), then the same solve as the fused kernel with every array in registers.
// register budget of the thread-per-group solve: (512, 1) = 128 registers; A/B builds override the pair
#ifndef ANOFOX_SOLVE_SMALL_T
#define ANOFOX_SOLVE_SMALL_T 512
#endif
#ifndef ANOFOX_SOLVE_SMALL_MINB
#define ANOFOX_SOLVE_SMALL_MINB 1
#endif
template <int P>
__global__ void __launch_bounds__(ANOFOX_SOLVE_SMALL_T, ANOFOX_SOLVE_SMALL_MINB) solve_small_kernel(const double *__restrict__ records, int64_t n_groups, FitParams Pm,
                                                          AnofoxB200FitOutputs o, int *refine_count,
                                                          int64_t *refine_list, int defer_p) {
	constexpr int K = P + 1;
	const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= n_groups) return;
	const double *rec = records + g;
	const int64_t rs = n_groups;
	SmallMoments<P> M;
	double sh[K];
	int f = 0;
	auto get = [&]() { return __ldcs(rec + (int64_t)(f++) * rs); };
	M.nv = (int)get();
	M.sw = get();
	const int n_state = (int)get();
	get();
#pragma unroll
	for (int c = 0; c < K; ++c) sh[c] = get();
	unsigned nonconst = 0u;
#pragma unroll
	for (int c = 0; c < P; ++c) nonconst |= !(get() < 1e-10) ? (1u << c) : 0u;
	f = rec_off_s1(P); // skip the pad fields
#pragma unroll
	for (int c = 0; c < K; ++c) M.s1[c] = get();
#pragma unroll
	for (int i = 0; i < SmallMoments<P>::NS2; ++i) M.s2[i] = get();
	M.ns = n_state;
	SmallGroupView v;
	v.n = 0;
	v.q = 0;
	v.T = 1;
	v.mask = 0u;
	v.pv = nullptr;
	v.refine_count = refine_count;
	v.refine_list = refine_list;
	v.rec_out = nullptr;
	v.stash = nullptr;
	v.defer_p = defer_p != 0;
	small_solve<P, false, false, false>(M, sh, nonconst, n_state, g, Pm, o, v);
}

// rows of 64 bytes (p = 8): the lanes of a group walk the columns in rotated order (small_accumulate ROT), which takes
// the 8-way shared-memory bank conflict out of the row loads: 1 M groups x 50 rows x 8 features, accumulate kernel
// 1.69 -> 1.51 ms (x 200 rows: 2.38 -> 2.27 ms). p = 4 (32-byte rows, 4-way conflict) measured 2 % slower with the
// rotation and its un-rotation than without, so it stays plain. A/B: -DANOFOX_SMALL_ROT=0.
#ifndef ANOFOX_SMALL_ROT
#define ANOFOX_SMALL_ROT 1
#endif
template <int P> constexpr bool kSmallRot = ANOFOX_SMALL_ROT != 0 && P == 8;
constexpr int kSmallStageGroups = 64; // groups per stage of the fused kernel at most (two rounds of 32)
constexpr int kSmallMaxStages = 8;
using SmallStageHeader = StageHeaderT<kSmallStageGroups>;

struct SmallArgs {
	AccArgs acc;
	FitParams params;
	AnofoxB200FitOutputs out;
	int lanes_per_group; // T: 4, 8, 16 or 32 (stash kernel: 1 .. 32)
	int max_stage_groups;
	int *refine_count;     // stash kernel: queue of the groups that need the exact residual pass
	int64_t *refine_list;
	int debug;             // experiments only: 1 = solve warps drop the tables without solving
	uint32_t wait_hint;    // suspend hint (ns) of the stage barriers' waits; 0 = spin
};

template <int P, int NCW, bool HASW, bool HAST, bool FUSED>
__global__ void __launch_bounds__(32 * (NCW + 1), 1) fit_small_kernel(const SmallArgs args) {
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ uint64_t full_bar[kSmallMaxStages], empty_bar[kSmallMaxStages];
	__shared__ SmallStageHeader hdr[kSmallMaxStages];
	constexpr int PVW = P + 3;                   // per slot: f[P] (t^2), F, d1 (0 = nothing to do), d2
	__shared__ double pv_scratch[FUSED ? NCW : 1][8 * PVW]; // at most 8 groups per warp round (T >= 4)
	__shared__ int64_t obuf[2 * kOffsetTile];               // the producer's prefetched group offsets
	const AccArgs &a = args.acc;
	const int NS = a.n_stages;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NS; ++s) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], NCW);
		}
		mbar_fence_init();
	}
	__syncthreads();
	if (warp == 0) {
		producer_loop_pf<HASW, HAST>(a, smem, full_bar, empty_bar, hdr, NS, lane, args.max_stage_groups, obuf, 1, 0, args.wait_hint);
		return;
	}
	const int T = args.lanes_per_group;
	const int lt = 31 - __clz(T);      // log2 T
	const int gpw = 32 >> lt, lg = 5 - lt; // groups per warp round and its log2
	const int sub = lane >> lt, q = lane & (T - 1);
	const unsigned gmask = (T == 32 ? 0xffffffffu : ((1u << T) - 1u)) << (sub * T);
	double *pvw = pv_scratch[FUSED ? warp - 1 : 0];
	const bool want_p = FUSED && args.params.compute_inference && args.out.std_errors;
	const DfTables Tb{args.params.tcrit, args.params.gratio, args.params.rcp, args.params.ce, args.params.co};
	int s = 0;
	uint32_t ph = 0;
	for (;;) {
		mbar_wait_hint(&full_bar[s], ph, args.wait_hint);
		const int cnt = hdr[s].count;
		if (cnt < 0) break;
		unsigned char *base = smem + (size_t)s * a.stage_bytes;
		const uint32_t ax = smem_u32(base + hdr[s].lead_x);
		const uint32_t ay = smem_u32(base + a.xcap + hdr[s].lead_y);
		const uint32_t aw = HASW ? smem_u32(base + a.xcap + a.ycap + hdr[s].lead_w) : 0u;
		const uint32_t at = HAST ? smem_u32(base + a.xcap + a.ycap + a.wcap + hdr[s].lead_t) : 0u;
		const int64_t g0 = hdr[s].g_first;
		const int nrounds = (cnt + gpw - 1) >> lg;
		for (;;) {
			if (*(volatile int *)&hdr[s].next >= nrounds) break; // nothing left to claim: skip the atomic
			int r = 0;
			if (lane == 0) r = atomicAdd(&hdr[s].next, 1);
			r = __shfl_sync(0xffffffffu, r, 0);
			if (r >= nrounds) break;
			const int gi = (r << lg) + sub;
			if (want_p && q == 0) pvw[sub * PVW + P + 1] = 0.0;
			__syncwarp();
			if (gi < cnt) {
				const int rb = (int)hdr[s].rowstart[gi];
				SmallGroupView v;
				v.n = (int)hdr[s].rowstart[gi + 1] - rb;
				v.q = q;
				v.T = T;
				v.mask = gmask;
				v.pv = pvw + sub * PVW;
				v.refine_count = nullptr;
				v.refine_list = nullptr;
				v.rec_out = nullptr;
				v.stash = nullptr;
				v.sg.x = ax + (uint32_t)rb * (uint32_t)(P * 8);
				v.sg.y = ay + (uint32_t)rb * 8u;
				v.sg.w = aw + (uint32_t)rb * 8u;
				v.sg.t = at + (uint32_t)rb;
				v.sg.kone = v.sg.kzero = 0u;
				small_group<P, HASW, HAST, FUSED ? 1 : 0, kSmallRot<P>>(v, g0 + gi, args.params, args.out, a.records, a.n_groups);
			}
			__syncwarp();
			if (want_p) {
				// the round's p-values: (P + 1) per group, spread over all 32 lanes. The t tests are the F(1, df)
				// upper tail of t^2, so every lane runs the same function.
				const int total = gpw * (P + 1);
				for (int e = lane; e < total; e += 32) {
					const int slot = e / (P + 1), item = e - slot * (P + 1);
					const double *pv = pvw + slot * PVW;
					const double d1 = pv[P + 1];
					if (d1 > 0.0) {
						const double pval = pvalue_f(pv[item], item < P ? 1.0 : d1, pv[P + 2], Tb);
						const int64_t g = g0 + (r << lg) + slot;
						if (item < P)
							args.out.p_values[g * P + item] = pval;
						else
							args.out.f_pvalue[g] = pval;
					}
				}
				__syncwarp();
			}
		}
		__syncwarp();
		if (lane == 0) mbar_arrive(&empty_bar[s]);
		if (++s == NS) {
			s = 0;
			ph ^= 1u;
		}
	}
}

// ---------------------------------------------------------------------------------------------------------
// Stash kernel: rows -> results in ONE launch, with no redundant solve and no record round trip through HBM.
// Warp roles inside the persistent CTA (one per SM):
//   warp 0                 producer: TMA bulk copies of whole groups into the ring of stages (producer_loop)
//   warps 1 .. NACC        accumulate: take the next stage by ticket (a stage belongs to exactly one warp), fold
//                          its groups T lanes per group (small_group<MODE 2>) and park each group's sums in one of
//                          the warp's two private 32-slot tables in shared memory; a full table is posted
//   warps NACC+1 ..        solve: take the next posted table by ticket, lane k loads the sums of slot k into
//                          registers, the table is handed back at once, and every lane runs the thread-per-group
//                          solve + inference (the instruction stream of solve_small_kernel) and writes the struct
// The accumulate warps are bound by the TMA stream from HBM, the solve warps by FP64 latency (a dependent chain
// per group: it takes ~16 of them per SM to keep the issue slots busy); inside one CTA the two overlap instead
// of running back to back as two launches, and the 0.7 GB record round trip of the two-kernel path is gone.
#ifndef ANOFOX_STASH_SEG
#define ANOFOX_STASH_SEG 0
#endif
template <int P, int NACC, int NSOL, bool HASW, bool HAST>
__global__ void __launch_bounds__(32 * (1 + NACC + NSOL), 1) fit_small_stash_kernel(const SmallArgs args) {
	extern __shared__ __align__(128) unsigned char smem[];
	constexpr int NB = 32; // full barriers of the ring / of the posted tables: more than the warps that can wait on them
	constexpr int NTAB = 2 * NACC;
	constexpr int TAB = SmallStash<P, HASW>::NF * 32; // doubles per table
	static_assert(NB > kSmallMaxStages + NACC && NB > NSOL + NTAB, "a ticket holder must never be a whole barrier phase ahead");
	static_assert(!ANOFOX_STASH_SEG || 1 + NACC + NSOL >= 4 * NACC + 1, "the accumulate warps are warps 4, 8, .. 4 NACC");
	__shared__ uint64_t full_bar[NB], empty_bar[kSmallMaxStages];
	__shared__ uint64_t tab_full[NB], tab_empty[NTAB];
	__shared__ int tab_ent[NB]; // posted table number k: physical table | (groups + 1) << 8; groups + 1 == 0: no more work
	__shared__ SmallStageHeader hdr[kSmallMaxStages];
	__shared__ int ticket, tab_prod, tab_cons, acc_done;
	__shared__ int64_t obuf[2 * kOffsetTile];
	const AccArgs &a = args.acc;
	const int NS = a.n_stages;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NB; ++s) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&tab_full[s], 1);
		}
		for (int s = 0; s < NS; ++s) mbar_init(&empty_bar[s], 1); // the one warp that took the stage
		for (int s = 0; s < NTAB; ++s) mbar_init(&tab_empty[s], 1);
		mbar_fence_init();
		ticket = tab_prod = tab_cons = acc_done = 0;
	}
	__syncthreads();
	if (warp == 0) {
		producer_loop_pf<HASW, HAST>(a, smem, full_bar, empty_bar, hdr, NS, lane, args.max_stage_groups, obuf, NACC, NB, args.wait_hint);
		return;
	}
	double *pool = reinterpret_cast<double *>(smem + (size_t)NS * a.stage_bytes);
	// Roles by scheduler: warp w issues from sub-partition w % 4. The producer and the accumulate warps (a small
	// loop) all sit on sub-partition 0; the solve warps (40 KB of straight-line code per pass, which streams
	// through the instruction cache of the sub-partition) take the other three and cannot evict that loop.
#if ANOFOX_STASH_SEG
	const bool is_acc = (warp & 3) == 0 && (warp >> 2) <= NACC;
	const int aw_i = (warp >> 2) - 1;
#else
	const bool is_acc = warp <= NACC;
	const int aw_i = warp - 1;
#endif
	if (is_acc) {
		// ------------------------------------------------ accumulate warps ------------------------------------
		const int T = args.lanes_per_group;
		const int lt = 31 - __clz(T);
		const int gpw = 32 >> lt, lg = 5 - lt;
		const int sub = lane >> lt, q = lane & (T - 1);
		const unsigned gmask = (T == 32 ? 0xffffffffu : ((1u << T) - 1u)) << (sub * T);
		int filled = 0, use = 0; // groups parked in the current table; tables this warp has started so far
		auto post = [&](int phys, int groups) {
			__syncwarp();
			if (lane == 0) {
				const int k = atomicAdd(&tab_prod, 1);
				tab_ent[k & (NB - 1)] = phys | ((groups + 1) << 8);
				mbar_arrive(&tab_full[k & (NB - 1)]); // release: the table's stores are visible to the waiter
			}
		};
		for (;;) {
			int tk = 0;
			if (lane == 0) tk = atomicAdd(&ticket, 1);
			tk = __shfl_sync(0xffffffffu, tk, 0);
			const int s = tk % NS;
			mbar_wait_hint(&full_bar[tk & (NB - 1)], (uint32_t)((tk / NB) & 1), args.wait_hint);
			const int cnt = hdr[s].count;
			if (cnt < 0) {
				if (lane == 0) mbar_arrive(&empty_bar[s]);
				break;
			}
			const int phys = 2 * aw_i + (use & 1);
			if (filled == 0) mbar_wait(&tab_empty[phys], (uint32_t)(((use >> 1) & 1) ^ 1)); // its previous use has been read
			double *st = pool + (size_t)phys * TAB;
			unsigned char *base = smem + (size_t)s * a.stage_bytes;
			const uint32_t ax = smem_u32(base + hdr[s].lead_x);
			const uint32_t ay = smem_u32(base + a.xcap + hdr[s].lead_y);
			const uint32_t aw = HASW ? smem_u32(base + a.xcap + a.ycap + hdr[s].lead_w) : 0u;
			const uint32_t at = HAST ? smem_u32(base + a.xcap + a.ycap + a.wcap + hdr[s].lead_t) : 0u;
			const int64_t g0 = hdr[s].g_first;
			const int nrounds = (args.debug & 2) ? 0 : (cnt + gpw - 1) >> lg;
			for (int r = 0; r < nrounds; ++r) {
				const int gi = (r << lg) + sub;
				if (gi < cnt) {
					const int rb = (int)hdr[s].rowstart[gi];
					SmallGroupView v;
					v.n = (int)hdr[s].rowstart[gi + 1] - rb;
					v.q = q;
					v.T = T;
					v.mask = gmask;
					v.pv = nullptr;
					v.refine_count = nullptr;
					v.refine_list = nullptr;
					v.rec_out = nullptr;
					v.stash = st + filled + gi;
					v.sg.x = ax + (uint32_t)rb * (uint32_t)(P * 8);
					v.sg.y = ay + (uint32_t)rb * 8u;
					v.sg.w = aw + (uint32_t)rb * 8u;
					v.sg.t = at + (uint32_t)rb;
					v.sg.kone = v.sg.kzero = 0u;
					small_group<P, HASW, HAST, 2>(v, g0 + gi, args.params, args.out, a.records, a.n_groups);
				}
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(&empty_bar[s]); // the rows are no longer needed: only the sums are kept
			filled += cnt;
			if (filled + args.max_stage_groups > 32) {
				post(phys, filled);
				filled = 0;
				++use;
			}
		}
		if (filled > 0) post(2 * aw_i + (use & 1), filled);
		// the last accumulate warp to finish tells every solve warp that no more tables will come
		int last = 0;
		if (lane == 0) last = atomicAdd(&acc_done, 1) == NACC - 1 ? 1 : 0;
		last = __shfl_sync(0xffffffffu, last, 0);
		if (last)
			for (int i = 0; i < NSOL; ++i) post(255, -1);
		return;
	}
	// ---------------------------------------------------- solve warps -----------------------------------------
	for (;;) {
		int c = 0;
		if (lane == 0) c = atomicAdd(&tab_cons, 1);
		c = __shfl_sync(0xffffffffu, c, 0);
		mbar_wait(&tab_full[c & (NB - 1)], (uint32_t)((c / NB) & 1));
		const int ent = tab_ent[c & (NB - 1)];
		const int n = (ent >> 8) - 1;
		if (n < 0) break;
		const int phys = ent & 255;
		SmallMoments<P> M;
		double sh[P + 1];
		unsigned nonconst = 0u;
		int n_state = 0;
		int64_t g = 0;
		if (lane < n) small_stash_load<P, HASW>(pool + (size_t)phys * TAB + lane, M, sh, nonconst, n_state, g);
		__syncwarp();
		if (lane == 0) mbar_arrive(&tab_empty[phys]); // the sums are in registers: the table can be refilled
		if (args.debug & 1) {
			if (lane < n && !(args.debug & 2)) args.out.intercept[g] = M.s1[0] + sh[0];
			continue;
		}
		if (lane < n) {
			SmallGroupView v;
			v.n = 0;
			v.q = 0;
			v.T = 1;
			v.mask = 0u;
			v.pv = nullptr;
			v.refine_count = args.refine_count;
			v.refine_list = args.refine_list;
			v.rec_out = a.records;
			v.stash = nullptr;
			small_solve<P, false, false, false>(M, sh, nonconst, n_state, g, args.params, args.out, v);
		}
		__syncwarp();
	}
}

} // namespace b200
