// groupfit_oracle.cpp -- CPU ORACLE. TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library. The product (anofox-statistics_b200/) never links, imports or calls it.
//
// What it is: a plain-C++ restatement, rule by rule, of the reference's per-group OLS / Ridge /
// WLS fit path and its prediction helper:
//   * row filter, constant-column rule, intercept-only shortcut, min-observation rule, NaN
//     scatter-back:  crates/anofox-stats-core/src/models/ols.rs:36-268, ridge.rs:36-229,
//     wls.rs:37-289
//   * FFI conventions (error codes, NULL->NaN, intercept NaN when absent, inference default):
//     crates/anofox-stats-ffi/src/lib.rs:98-265, types.rs:24-165
//   * aggregate NULL-struct rule (< 2 rows in the state): src/aggregate_functions/
//     ols_aggregate.cpp:263-267, :298-301
//   * prediction + simplified interval: lib.rs:2217-2231, :2266-2350;
//     fit_predict row semantics: src/aggregate_functions/ols_predict_aggregate.cpp:135-265,322-426
//   * AIC / BIC: crates/anofox-stats-core/src/diagnostics/information_criteria.rs:15-87
//
// The decomposition itself lives in crates that are NOT under /root/reference
// (anofox-regression 0.5.13, faer 0.23.2, statrs 0.18.0 -- Cargo.lock:6-9,347-350,1121-1124), so
// the least-squares core here restates the *published* algorithm the reference's golden vectors
// were generated with (R lm(): LINPACK dqrdc2 Householder QR with limited column pivoting,
// tol = 1e-7; summary.lm formulas) and is PINNED against every golden vector / known-answer
// test the reference holds for this path (tests/test_oracle_golden.py, tests/golden/*).
// It works on RAW ROWS (no Gram matrix), which makes it an independent check of the device
// path's sufficient-statistics algorithm.
//
// PARITY UNPINNED items (no reference test or readable source fixes them):
// which of several exactly collinear columns is aliased (we follow R: the later one), Ridge
// standard errors / adj R^2 / F, the glmnet lambda scaling factor, statistics at zero residual
// degrees of freedom.
//
// Build: see oracle/Makefile (g++ -O2 -fopenmp -shared).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/anofox_b200.h"

namespace {

const double kNaN = std::numeric_limits<double>::quiet_NaN();
const double kInf = std::numeric_limits<double>::infinity();

// ---------------------------------------------------------------------------------------------
// Special functions (Student t, F). statrs 0.18 is what the reference calls (lib.rs:26,2223-2230);
// these are independent implementations checked against scipy/mpmath in tests/.
// ---------------------------------------------------------------------------------------------

template <class R> R lgam(R x);
template <> long double lgam<long double>(long double x) { return lgammal(x); }
template <> double lgam<double>(double x) { return std::lgamma(x); }

// Continued fraction of the incomplete beta function (modified Lentz).
template <class R> R beta_cf(R a, R b, R x) {
	const R tiny = std::numeric_limits<R>::min() * 1e10;
	const R eps = std::numeric_limits<R>::epsilon();
	R qab = a + b, qap = a + 1, qam = a - 1;
	R c = 1, d = 1 - qab * x / qap;
	if (std::fabs(d) < tiny) d = tiny;
	d = 1 / d;
	R h = d;
	for (int m = 1; m <= 100000; ++m) {
		R m2 = 2 * (R)m;
		R aa = (R)m * (b - m) * x / ((qam + m2) * (a + m2));
		d = 1 + aa * d;
		if (std::fabs(d) < tiny) d = tiny;
		c = 1 + aa / c;
		if (std::fabs(c) < tiny) c = tiny;
		d = 1 / d;
		h *= d * c;
		aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
		d = 1 + aa * d;
		if (std::fabs(d) < tiny) d = tiny;
		c = 1 + aa / c;
		if (std::fabs(c) < tiny) c = tiny;
		d = 1 / d;
		R del = d * c;
		h *= del;
		if (std::fabs(del - 1) < 2 * eps) break;
	}
	return h;
}

// Regularised incomplete beta I_x(a,b); xc = 1 - x supplied by the caller without cancellation.
template <class R> R beta_inc(R a, R b, R x, R xc) {
	if (!(x > 0)) return 0;
	if (!(xc > 0)) return 1;
	R lnpre = lgam<R>(a + b) - lgam<R>(a) - lgam<R>(b) + a * std::log(x) + b * std::log(xc);
	if (x < (a + 1) / (a + b + 2)) return std::exp(lnpre) * beta_cf<R>(a, b, x) / a;
	return 1 - std::exp(lnpre) * beta_cf<R>(b, a, xc) / b;
}

// Two-sided p-value 2*(1 - T_df(|t|)) = I_{df/(df+t^2)}(df/2, 1/2).
template <class R> R t_two_sided_p(R t, R df) {
	if (std::isnan(t) || !(df > 0)) return std::numeric_limits<R>::quiet_NaN();
	if (std::isinf(t)) return 0;
	R t2 = t * t;
	R x = df / (df + t2), xc = t2 / (df + t2);
	return beta_inc<R>(df / 2, (R)0.5, x, xc);
}

// Quantile t_q(df) for q in (0.5, 1): bisection on the two-sided tail (monotone), then polish.
template <class R> R t_quantile_upper(R q, R df) {
	R target = 2 * (1 - q); // two-sided tail mass
	R lo = 0, hi = 1;
	while (t_two_sided_p<R>(hi, df) > target) {
		hi *= 2;
		if (hi > 1e300) return std::numeric_limits<R>::infinity();
	}
	for (int it = 0; it < 400; ++it) {
		R mid = (lo + hi) / 2;
		if (mid == lo || mid == hi) break;
		if (t_two_sided_p<R>(mid, df) > target)
			lo = mid;
		else
			hi = mid;
	}
	return (lo + hi) / 2;
}

// anofox_t_critical (lib.rs:2217-2231)
template <class R> double t_critical(double level, size_t df) {
	if (df == 0 || !(level > 0.0) || !(level < 1.0)) return kNaN;
	R q = ((R)1 + (R)level) / 2;
	return (double)t_quantile_upper<R>(q, (R)df);
}

// Upper tail of F(d1, d2) at f.
template <class R> R f_upper_p(R f, R d1, R d2) {
	if (std::isnan(f) || !(d1 > 0) || !(d2 > 0)) return std::numeric_limits<R>::quiet_NaN();
	if (f <= 0) return 1;
	if (std::isinf(f)) return 0;
	R den = d2 + d1 * f;
	return beta_inc<R>(d2 / 2, d1 / 2, d2 / den, d1 * f / den);
}

// ---------------------------------------------------------------------------------------------
// One group
// ---------------------------------------------------------------------------------------------

struct Opts {
	int method;
	bool fit_intercept;
	bool compute_inference;
	double level;
	double alpha;
	int lambda_scaling;
	int hc_type = 0; // AnofoxHcType: robust (sandwich) standard errors for OLS / WLS, ols.rs:209-245, wls.rs:230-266
};

struct GroupResult {
	int code = 0; // AnofoxErrorCode
	std::vector<double> coef, se, tv, pv, lo, hi;
	double intercept = kNaN, r2 = kNaN, adj = kNaN, rse = kNaN;
	double fstat = kNaN, fp = kNaN, rss = kNaN, aic = kNaN, bic = kNaN;
	size_t n_obs = 0;
	bool has_inf = false;
};

struct RowView {
	// Either column arrays (reference ABI) or row-major rows (batch ABI).
	const double *y = nullptr;
	const double *xrm = nullptr;          // row-major [n*p]
	const AnofoxDataArray *xcols = nullptr; // p columns
	const uint8_t *yvalid = nullptr;
	const double *w = nullptr;
	const uint8_t *wvalid = nullptr;
	size_t n = 0, p = 0;
	static bool bit(const uint8_t *v, size_t i) { return v == nullptr || ((v[i >> 3] >> (i & 7)) & 1); }
	double Y(size_t i) const { return bit(yvalid, i) ? y[i] : kNaN; }
	double X(size_t i, size_t j) const {
		if (xrm) return xrm[i * p + j];
		return bit(xcols[j].validity, i) ? xcols[j].data[i] : kNaN;
	}
	double W(size_t i) const { return w ? (bit(wvalid, i) ? w[i] : kNaN) : 1.0; }
};

void info_criteria(double rss, size_t n, size_t k, double &aic, double &bic) {
	// information_criteria.rs:15-33, :68-87
	if (n == 0 || rss < 0 || std::isnan(rss)) {
		aic = bic = kNaN;
		return;
	}
	if (rss == 0.0) {
		aic = bic = -kInf;
		return;
	}
	double nf = (double)n, kf = (double)k;
	aic = nf * std::log(rss / nf) + 2.0 * kf;
	bic = nf * std::log(rss / nf) + kf * std::log(nf);
}

// Householder QR with LINPACK-dqrdc2 style limited pivoting on A (rows x m, column-major in `a`),
// right-hand side b. On return: perm = column order, rank, R in the upper triangle of the permuted
// columns, b overwritten with Q^T b. tol < 0 disables pivoting (ridge).
// orig_given: the reference norms of the negligibility test when the columns of `a` are not the original ones.
template <class R>
void qr_limited(std::vector<R> &a, std::vector<R> &b, size_t rows, size_t m, R tol, std::vector<size_t> &perm,
                size_t &rank, const std::vector<R> *orig_given = nullptr) {
	perm.resize(m);
	for (size_t j = 0; j < m; ++j) perm[j] = j;
	std::vector<R> orig(m);
	for (size_t j = 0; j < m; ++j) {
		R s = 0;
		for (size_t i = 0; i < rows; ++i) s += a[j * rows + i] * a[j * rows + i];
		orig[j] = orig_given ? (*orig_given)[j] : std::sqrt(s);
		if (orig[j] == 0) orig[j] = 1;
	}
	size_t lup = m, l = 0;
	size_t lmax = std::min(rows, m);
	while (l < lmax && l < lup) {
		size_t col = perm[l];
		R s = 0;
		for (size_t i = l; i < rows; ++i) s += a[col * rows + i] * a[col * rows + i];
		R nrm = std::sqrt(s);
		if (tol >= 0 && nrm < orig[col] * tol) {
			// negligible: cycle this column to the end, keep l
			for (size_t k = l; k + 1 < m; ++k) perm[k] = perm[k + 1];
			perm[m - 1] = col;
			--lup;
			continue;
		}
		if (l + 1 < rows && nrm != 0) {
			R *v = &a[col * rows];
			R alpha = (v[l] >= 0) ? -nrm : nrm; // R_ll
			// v := x - alpha e_l (stored in place below the diagonal), beta = 2 / (v^T v)
			R v0 = v[l] - alpha;
			R vtv = v0 * v0;
			for (size_t i = l + 1; i < rows; ++i) vtv += v[i] * v[i];
			if (vtv > 0) {
				R beta = 2 / vtv;
				auto reflect = [&](R *t) {
					R dot = v0 * t[l];
					for (size_t i = l + 1; i < rows; ++i) dot += v[i] * t[i];
					dot *= beta;
					t[l] -= dot * v0;
					for (size_t i = l + 1; i < rows; ++i) t[i] -= dot * v[i];
				};
				for (size_t k = l + 1; k < m; ++k) reflect(&a[perm[k] * rows]);
				reflect(b.data());
			}
			v[l] = alpha;
			for (size_t i = l + 1; i < rows; ++i) v[i] = 0;
		}
		++l;
	}
	rank = std::min(lup, lmax);
}

// per-thread memo of t quantiles: the level is fixed within a batch, df repeats across groups
template <class R> struct TcritMemo {
	double level = -1;
	std::vector<std::pair<double, R>> items;
	R get(double lvl, double df) {
		if (lvl != level) {
			items.clear();
			level = lvl;
		}
		for (auto &it : items)
			if (it.first == df) return it.second;
		R v = t_quantile_upper<R>(((R)1 + (R)lvl) / 2, (R)df);
		if (items.size() < 4096) items.emplace_back(df, v);
		return v;
	}
};

template <class R> void fit_group(const Opts &o, const RowView &d, GroupResult &out, TcritMemo<R> *memo = nullptr) {
	const size_t n = d.n, p = d.p;
	out = GroupResult();
	// ridge.rs:38-40: alpha is validated before anything else
	if (o.method == ANOFOX_B200_RIDGE && o.alpha < 0.0) {
		out.code = ANOFOX_ERROR_INVALID_ALPHA;
		return;
	}
	// ols.rs:38-43 (EmptyInput -> InvalidInput, lib.rs:74)
	if (n == 0 || p == 0) {
		out.code = ANOFOX_ERROR_INVALID_INPUT;
		return;
	}
	const bool wls = o.method == ANOFOX_B200_WLS;
	// ols.rs:59-66 / wls.rs:76-86: finite-row filter
	std::vector<size_t> valid;
	valid.reserve(n);
	for (size_t i = 0; i < n; ++i) {
		if (!std::isfinite(d.Y(i))) continue;
		if (wls) {
			double w = d.W(i);
			if (!(w > 0.0) || !std::isfinite(w)) continue;
		}
		bool ok = true;
		for (size_t j = 0; j < p && ok; ++j) ok = std::isfinite(d.X(i, j));
		if (ok) valid.push_back(i);
	}
	if (valid.empty()) {
		out.code = ANOFOX_ERROR_NO_VALID_DATA;
		return;
	}
	const size_t nv = valid.size();
	out.n_obs = nv;
	// ols.rs:76-87: constant column <=> all |x_i - x_first| < 1e-10
	std::vector<size_t> keep;
	for (size_t j = 0; j < p; ++j) {
		double first = d.X(valid[0], j);
		bool cst = true;
		for (size_t k = 0; k < nv && cst; ++k) cst = std::fabs(d.X(valid[k], j) - first) < 1e-10;
		if (!cst) keep.push_back(j);
	}
	const size_t pe = keep.size();
	const size_t ic = o.fit_intercept ? 1 : 0;
	out.coef.assign(p, kNaN);
	if (pe == 0) {
		// ols.rs:101-130 / wls.rs:119-150
		if (!o.fit_intercept) {
			out.code = ANOFOX_ERROR_INSUFFICIENT_DATA;
			return;
		}
		R sw = 0, swy = 0;
		for (size_t k = 0; k < nv; ++k) {
			R w = wls ? (R)d.W(valid[k]) : (R)1;
			sw += w;
			swy += w * (R)d.Y(valid[k]);
		}
		R mean = swy / sw, ss = 0;
		for (size_t k = 0; k < nv; ++k) {
			R w = wls ? (R)d.W(valid[k]) : (R)1;
			R e = (R)d.Y(valid[k]) - mean;
			ss += w * e * e;
		}
		R var = wls ? ss / sw : ss / (R)((double)nv - 1.0);
		out.intercept = (double)mean;
		out.r2 = 0.0;
		out.adj = 0.0;
		out.rse = (double)std::sqrt(var);
		out.rss = (double)ss;
		info_criteria(out.rss, nv, 1, out.aic, out.bic);
		out.has_inf = false;
		return;
	}
	// ols.rs:134-139
	if (nv < pe + ic) {
		out.code = ANOFOX_ERROR_INSUFFICIENT_DATA;
		return;
	}

	// Weighted means and total sum of squares
	R sw = 0, swy = 0;
	std::vector<R> sqw(nv), xbar(pe, 0);
	for (size_t k = 0; k < nv; ++k) {
		R w = wls ? (R)d.W(valid[k]) : (R)1;
		sqw[k] = std::sqrt(w);
		sw += w;
		swy += w * (R)d.Y(valid[k]);
		for (size_t c = 0; c < pe; ++c) xbar[c] += w * (R)d.X(valid[k], keep[c]);
	}
	R ybar = swy / sw;
	for (size_t c = 0; c < pe; ++c) xbar[c] /= sw;
	R tss = 0;
	for (size_t k = 0; k < nv; ++k) {
		R e = (R)d.Y(valid[k]) - (ic ? ybar : (R)0);
		tss += sqw[k] * sqw[k] * e * e;
	}

	double lambda = 0.0;
	if (o.method == ANOFOX_B200_RIDGE) {
		// Raw: penalty as given. Glmnet: glmnet minimises RSS/(2n) + (lambda/2)|b|^2 on a response it
		// first scales to unit (1/n) variance, which on the original scale is the raw penalty
		// lambda * n / sd_y.  This reproduces the reference's glmnet golden vectors
		// (test/data/ridge_tests) to 5e-7; the upstream crate's own factor is unpinned (iii).
		lambda = o.alpha;
		if (o.lambda_scaling == ANOFOX_LAMBDA_SCALING_GLMNET) {
			double sdy = std::sqrt((double)tss / (double)nv);
			lambda = o.alpha * (double)nv;
			if (sdy > 0.0) lambda /= sdy;
		}
	}

	std::vector<R> beta(pe, std::numeric_limits<R>::quiet_NaN()); // per kept column
	std::vector<R> vdiag(pe, std::numeric_limits<R>::quiet_NaN()); // diag of (X'X [+ lambda I])^-1 per kept column
	std::vector<bool> aliased(pe, false);
	R b0 = ic ? ybar : (R)0;
	size_t rank_total; // estimated parameters incl. intercept
	// frame the residuals are evaluated in: e = (y - cy) - b0c - sum beta (x - cx). With an intercept the centre is
	// the origin the factorisation used, so that a residual of 1e-6 next to features of 1e6 keeps its digits.
	std::vector<R> cx(pe, 0);
	R cy = 0, b0c = 0;
	// kept for the sandwich estimator: R^-1 of the factorised (shifted) design, its column order and rank
	std::vector<R> hc_rinv, hc_shift(pe, 0);
	std::vector<size_t> hc_perm;
	size_t hc_rank = 0;

	if (lambda > 0.0) {
		// Ridge (guides/02_technical_guide.md:129-138; validation/r_seesion_output.txt:668-671):
		// centre when an intercept is fitted, solve the augmented least-squares problem
		// [Xc; sqrt(lambda) I] b = [yc; 0] by unpivoted Householder QR.
		size_t rows = nv + pe;
		std::vector<R> a(rows * pe, 0), b(rows, 0);
		R sl = std::sqrt((R)lambda);
		for (size_t c = 0; c < pe; ++c) {
			for (size_t k = 0; k < nv; ++k)
				a[c * rows + k] = (R)d.X(valid[k], keep[c]) - (ic ? xbar[c] : (R)0);
			a[c * rows + nv + c] = sl;
		}
		for (size_t k = 0; k < nv; ++k) b[k] = (R)d.Y(valid[k]) - (ic ? ybar : (R)0);
		std::vector<size_t> perm;
		size_t rank;
		qr_limited<R>(a, b, rows, pe, (R)-1, perm, rank);
		// back substitution
		for (size_t q = pe; q-- > 0;) {
			R s = b[q];
			for (size_t k = q + 1; k < pe; ++k) s -= a[k * rows + q] * beta[k];
			beta[q] = s / a[q * rows + q];
		}
		// R^-1 rows -> diag((R^T R)^-1)
		std::vector<R> rinv(pe * pe, 0);
		for (size_t c = 0; c < pe; ++c) {
			rinv[c * pe + c] = 1 / a[c * rows + c];
			for (size_t r = c; r-- > 0;) {
				R s = 0;
				for (size_t k = r + 1; k <= c; ++k) s += a[k * rows + r] * rinv[c * pe + k];
				rinv[c * pe + r] = -s / a[r * rows + r];
			}
		}
		for (size_t r = 0; r < pe; ++r) {
			R s = 0;
			for (size_t c = r; c < pe; ++c) s += rinv[c * pe + r] * rinv[c * pe + r];
			vdiag[r] = s;
		}
		if (ic) {
			b0 = ybar;
			for (size_t c = 0; c < pe; ++c) b0 -= xbar[c] * beta[c];
			cx = xbar;
			cy = ybar;
			b0c = 0;
		}
		rank_total = pe + ic;
	} else {
		// OLS / WLS: design [sqrt(w) | sqrt(w) x], dqrdc2 limited pivoting, tol 1e-7 (R lm.fit).
		// With an intercept the feature columns and y enter the factorisation minus the first valid row (a shift lies
		// in the span of the intercept column: same fit, same remaining norms) -- the working precision is then spent
		// on the variation of the data instead of on a common offset. dqrdc2's reference norms stay those of the
		// ORIGINAL columns, so the aliasing rule is the one lm.fit applies to the raw design.
		size_t m = pe + ic;
		std::vector<R> a(nv * m), b(nv), shiftx(pe, 0), orig_norm(m, 0);
		R shifty = 0;
		if (ic) {
			for (size_t c = 0; c < pe; ++c) shiftx[c] = (R)d.X(valid[0], keep[c]);
			shifty = (R)d.Y(valid[0]);
		}
		for (size_t k = 0; k < nv; ++k) {
			if (ic) {
				a[k] = sqw[k];
				orig_norm[0] += sqw[k] * sqw[k];
			}
			for (size_t c = 0; c < pe; ++c) {
				R xv = (R)d.X(valid[k], keep[c]);
				a[(c + ic) * nv + k] = sqw[k] * (xv - shiftx[c]);
				orig_norm[c + ic] += (sqw[k] * xv) * (sqw[k] * xv);
			}
			b[k] = sqw[k] * ((R)d.Y(valid[k]) - shifty);
		}
		for (size_t j = 0; j < m; ++j) orig_norm[j] = std::sqrt(orig_norm[j]);
		std::vector<size_t> perm;
		size_t rank;
		qr_limited<R>(a, b, nv, m, (R)1e-7, perm, rank, &orig_norm);
		std::vector<R> sol(m, std::numeric_limits<R>::quiet_NaN());
		for (size_t q = rank; q-- > 0;) {
			R s = b[q];
			for (size_t k = q + 1; k < rank; ++k) s -= a[perm[k] * nv + q] * sol[perm[k]];
			sol[perm[q]] = s / a[perm[q] * nv + q];
		}
		// inverse of the leading rank x rank triangle
		std::vector<R> rinv(rank * rank, 0), vd(m, std::numeric_limits<R>::quiet_NaN());
		for (size_t c = 0; c < rank; ++c) {
			rinv[c * rank + c] = 1 / a[perm[c] * nv + c];
			for (size_t r = c; r-- > 0;) {
				R s = 0;
				for (size_t k = r + 1; k <= c; ++k) s += a[perm[k] * nv + r] * rinv[c * rank + k];
				rinv[c * rank + r] = -s / a[perm[r] * nv + r];
			}
		}
		for (size_t r = 0; r < rank; ++r) {
			R s = 0;
			for (size_t c = r; c < rank; ++c) s += rinv[c * rank + r] * rinv[c * rank + r];
			vd[perm[r]] = s;
		}
		if (ic) {
			cx = shiftx;
			cy = shifty;
			b0c = sol[0];
			// back on the unshifted scale: y - shifty = b0' + sum beta (x - shiftx)
			R acc = sol[0] + shifty;
			for (size_t c = 0; c < pe; ++c)
				if (!std::isnan((double)sol[c + ic])) acc -= sol[c + ic] * shiftx[c];
			sol[0] = acc;
		}
		if (ic) b0 = sol[0]; // the intercept column is never negligible
		hc_rinv = rinv;
		hc_perm = perm;
		hc_rank = rank;
		hc_shift = shiftx;
		for (size_t c = 0; c < pe; ++c) {
			beta[c] = sol[c + ic];
			vdiag[c] = vd[c + ic];
			aliased[c] = std::isnan((double)sol[c + ic]);
		}
		rank_total = rank;
	}

	// Residuals from the raw rows (NaN coefficients contribute nothing)
	R rss = 0;
	std::vector<R> resid(nv);
	for (size_t k = 0; k < nv; ++k) {
		R fit = b0c;
		for (size_t c = 0; c < pe; ++c)
			if (!aliased[c]) fit += beta[c] * ((R)d.X(valid[k], keep[c]) - cx[c]);
		R e = ((R)d.Y(valid[k]) - cy) - fit;
		resid[k] = e;
		rss += sqw[k] * sqw[k] * e * e;
	}

	const double dfr = (double)nv - (double)rank_total;
	const double dfm = (double)rank_total - (double)ic;
	R r2 = 1 - rss / tss;
	out.r2 = (double)r2;
	// summary.lm: 1 - (1 - r2) * (n - int) / df_resid. Zero residual df: NaN (unpinned iv).
	out.adj = dfr > 0 ? (double)(1 - (1 - r2) * (R)((double)nv - (double)ic) / (R)dfr) : kNaN;
	R sigma2 = dfr > 0 ? rss / (R)dfr : std::numeric_limits<R>::quiet_NaN();
	out.rse = (double)std::sqrt(sigma2);
	out.rss = (double)rss;
	out.intercept = ic ? (double)b0 : kNaN; // ols.rs:172-176, lib.rs:179
	for (size_t c = 0; c < pe; ++c) out.coef[keep[c]] = aliased[c] ? kNaN : (double)beta[c];
	info_criteria(out.rss, nv, rank_total, out.aic, out.bic);

	if (o.compute_inference) {
		out.has_inf = true;
		out.se.assign(p, kNaN);
		out.tv.assign(p, kNaN);
		out.pv.assign(p, kNaN);
		out.lo.assign(p, kNaN);
		out.hi.assign(p, kNaN);
		R tcrit = std::numeric_limits<R>::quiet_NaN();
		if (dfr > 0 && o.level > 0.0 && o.level < 1.0)
			tcrit = memo ? memo->get(o.level, dfr) : t_quantile_upper<R>(((R)1 + (R)o.level) / 2, (R)dfr);
		// Heteroscedasticity-consistent covariance (ols.rs:209-245 / wls.rs:230-266 -> upstream compute_hc_inference,
		// source not vendored). Restated from the published estimators (MacKinnon & White 1985; R sandwich::vcovHC
		// for weighted fits): V = (X'WX)^-1 [sum omega_i x~_i x~_i'] (X'WX)^-1 over the estimable columns incl. the
		// intercept, x~ = sqrt(w) x, e~ = sqrt(w) e, h_i = x~_i'(X'WX)^-1 x~_i, omega_i = e~^2 (HC0), e~^2 n/(n-k)
		// (HC1), e~^2/(1-h_i) (HC2), e~^2/(1-h_i)^2 (HC3). t, p and the interval use the same residual df; the F
		// statistic stays the classical one (the reference passes result.f_statistic through). Ridge has no HC.
		std::vector<R> vhc;
		if (o.hc_type != 0 && lambda == 0.0 && hc_rank > 0) {
			const size_t rk = hc_rank;
			std::vector<R> S(rk * rk, 0), u(rk), z(rk);
			for (size_t k = 0; k < nv; ++k) {
				for (size_t r = 0; r < rk; ++r) {
					const size_t j = hc_perm[r];
					z[r] = (ic && j == 0) ? sqw[k] : sqw[k] * ((R)d.X(valid[k], keep[j - ic]) - hc_shift[j - ic]);
				}
				R h = 0;
				for (size_t c = 0; c < rk; ++c) { // u = R^-T z
					R t = 0;
					for (size_t r = 0; r <= c; ++r) t += hc_rinv[c * rk + r] * z[r];
					u[c] = t;
					h += t * t;
				}
				R et = sqw[k] * resid[k], om = et * et;
				if (o.hc_type == ANOFOX_HC_HC1) om *= (R)nv / (R)((double)nv - (double)rk);
				else if (o.hc_type == ANOFOX_HC_HC2) om /= (1 - h);
				else if (o.hc_type == ANOFOX_HC_HC3) om /= (1 - h) * (1 - h);
				for (size_t a = 0; a < rk; ++a)
					for (size_t b2 = 0; b2 < rk; ++b2) S[a * rk + b2] += om * u[a] * u[b2];
			}
			vhc.assign(pe, std::numeric_limits<R>::quiet_NaN());
			for (size_t r = 0; r < rk; ++r) { // V_rr = sum_ab Rinv[r][a] S[a][b] Rinv[r][b], a, b >= r
				const size_t j = hc_perm[r];
				if (ic && j == 0) continue;
				R v = 0;
				for (size_t a = r; a < rk; ++a)
					for (size_t b2 = r; b2 < rk; ++b2) v += hc_rinv[a * rk + r] * S[a * rk + b2] * hc_rinv[b2 * rk + r];
				vhc[j - ic] = v;
			}
		}
		for (size_t c = 0; c < pe; ++c) {
			if (aliased[c]) continue;
			R se = vhc.empty() ? std::sqrt(sigma2 * vdiag[c]) : std::sqrt(vhc[c]);
			R t = beta[c] / se;
			size_t j = keep[c];
			out.se[j] = (double)se;
			out.tv[j] = (double)t;
			out.pv[j] = dfr > 0 ? (double)t_two_sided_p<R>(t, (R)dfr) : kNaN;
			out.lo[j] = (double)(beta[c] - tcrit * se);
			out.hi[j] = (double)(beta[c] + tcrit * se);
		}
		if (dfm > 0 && dfr > 0) {
			R f = ((tss - rss) / (R)dfm) / sigma2;
			out.fstat = (double)f;
			out.fp = (double)f_upper_p<R>(f, (R)dfm, (R)dfr);
		}
	}
}

void set_err(AnofoxError *e, int code, const char *msg) {
	if (!e) return;
	e->code = (AnofoxErrorCode)code;
	std::strncpy(e->message, msg, 255);
	e->message[255] = 0;
}

// Error texts = the Display strings of the reference's StatsError (crates/anofox-stats-core/src/errors.rs:7-37), which
// anofox_*_fit copies into AnofoxError.message (lib.rs:258-262) and the scalar SQL functions surface as
// "OLS fit failed: <message>" (src/table_functions/ols_fit.cpp:175-177; test/sql/error_handling.test matches on them).
std::string rust_f64(double v) { // Rust's Display for f64: shortest digits that round-trip, no exponent
	char buf[64];
	for (int prec = 1; prec <= 17; ++prec) {
		std::snprintf(buf, sizeof buf, "%.*g", prec, v);
		if (std::strtod(buf, nullptr) == v) break;
	}
	std::string t(buf);
	if (t.find('e') != std::string::npos) {
		std::snprintf(buf, sizeof buf, "%.17f", v);
		t = buf;
		while (!t.empty() && t.back() == '0') t.pop_back();
		if (!t.empty() && t.back() == '.') t.pop_back();
	}
	return t;
}
std::string msg_insufficient(size_t rows, size_t cols) {
	return "Insufficient data: " + std::to_string(rows) + " rows, " + std::to_string(cols) + " features (need rows > features)";
}
std::string msg_dimension(size_t y_len, size_t x_rows) {
	return "Dimension mismatch: y has " + std::to_string(y_len) + " elements, X has " + std::to_string(x_rows) + " rows";
}
std::string msg_empty(const char *field) { return std::string("Empty input: ") + field + " cannot be empty"; }
std::string msg_alpha(double a) { return "Invalid alpha parameter: " + rust_f64(a) + " (must be >= 0)"; }
const char *kMsgNoValidData = "All rows filtered due to NULL/NaN values";

double *dup_array(const std::vector<double> &v) {
	double *p = (double *)std::malloc(std::max<size_t>(v.size(), 1) * sizeof(double));
	if (p && !v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(double));
	return p;
}

bool single_fit(const Opts &o, AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, const AnofoxDataArray *w,
                AnofoxFitResultCore *core, AnofoxFitResultInference *inf, AnofoxError *err) {
	if (err) {
		err->code = ANOFOX_ERROR_SUCCESS;
		std::memset(err->message, 0, sizeof(err->message));
	}
	if (!core) {
		set_err(err, ANOFOX_ERROR_INVALID_INPUT, "out_core is NULL");
		return false;
	}
	if (!x || x_count == 0) {
		set_err(err, ANOFOX_ERROR_INVALID_INPUT, "x is NULL or empty");
		return false;
	}
	if (o.method == ANOFOX_B200_RIDGE && o.alpha < 0.0) {
		set_err(err, ANOFOX_ERROR_INVALID_ALPHA, msg_alpha(o.alpha).c_str());
		return false;
	}
	if (y.len == 0) {
		set_err(err, ANOFOX_ERROR_INVALID_INPUT, msg_empty("y").c_str());
		return false;
	}
	if (w && w->len == 0) {
		set_err(err, ANOFOX_ERROR_INVALID_INPUT, msg_empty("weights").c_str());
		return false;
	}
	if (w && w->len != y.len) {
		set_err(err, ANOFOX_ERROR_DIMENSION_MISMATCH, msg_dimension(y.len, w->len).c_str());
		return false;
	}
	for (size_t j = 0; j < x_count; ++j)
		if (x[j].len != y.len) {
			set_err(err, ANOFOX_ERROR_DIMENSION_MISMATCH, msg_dimension(y.len, x[j].len).c_str());
			return false;
		}
	RowView d;
	d.y = y.data;
	d.yvalid = y.validity;
	d.xcols = x;
	d.n = y.len;
	d.p = x_count;
	if (w) {
		d.w = w->data;
		d.wvalid = w->validity;
	}
	GroupResult r;
	fit_group<long double>(o, d, r);
	if (r.code != 0) {
		std::string m = r.code == ANOFOX_ERROR_INSUFFICIENT_DATA ? msg_insufficient(r.n_obs, x_count)
		                : r.code == ANOFOX_ERROR_NO_VALID_DATA   ? std::string(kMsgNoValidData)
		                                                         : std::string("fit failed");
		set_err(err, r.code, m.c_str());
		return false;
	}
	core->coefficients = dup_array(r.coef);
	core->coefficients_len = r.coef.size();
	core->intercept = r.intercept;
	core->r_squared = r.r2;
	core->adj_r_squared = r.adj;
	core->residual_std_error = r.rse;
	core->n_observations = r.n_obs;
	core->n_features = x_count;
	if (inf) {
		if (r.has_inf) {
			inf->std_errors = dup_array(r.se);
			inf->t_values = dup_array(r.tv);
			inf->p_values = dup_array(r.pv);
			inf->ci_lower = dup_array(r.lo);
			inf->ci_upper = dup_array(r.hi);
			inf->len = r.se.size();
			inf->confidence_level = o.level;
			inf->f_statistic = r.fstat;
			inf->f_pvalue = r.fp;
		} else {
			// lib.rs:251-253 -> FitResultInference::default() (types.rs:151-165)
			inf->std_errors = inf->t_values = inf->p_values = inf->ci_lower = inf->ci_upper = nullptr;
			inf->len = 0;
			inf->confidence_level = 0.95;
			inf->f_statistic = kNaN;
			inf->f_pvalue = kNaN;
		}
	}
	return true;
}

template <class R>
void predict_row(const double *coef, size_t len, double intercept, const double *x, double rse, size_t n_obs,
                 double level, double tcrit_cached, AnofoxPredictionResult *out) {
	// lib.rs:2292-2347
	double yhat = std::isnan(intercept) ? 0.0 : intercept;
	for (size_t j = 0; j < len; ++j)
		if (!std::isnan(coef[j])) yhat += coef[j] * x[j];
	out->yhat = out->yhat_lower = out->yhat_upper = yhat;
	if (std::isnan(rse) || rse <= 0.0 || n_obs <= len + 1) return;
	size_t k = len + (std::isnan(intercept) ? 0 : 1);
	size_t df = n_obs > k ? n_obs - k : 0;
	if (df == 0) return;
	double tc = std::isnan(tcrit_cached) ? t_critical<R>(level, df) : tcrit_cached;
	if (std::isnan(tc)) return;
	double margin = tc * (rse * std::sqrt(1.0 + 1.0 / (double)n_obs));
	out->yhat_lower = yhat - margin;
	out->yhat_upper = yhat + margin;
}

template <class R>
void fit_batch(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w, int32_t p,
               const AnofoxB200FitOptions *fo, const AnofoxB200FitOutputs *out, const uint8_t *train,
               const AnofoxB200PredictOutputs *pred, int threads) {
	Opts o{fo->method, fo->fit_intercept != 0, fo->compute_inference != 0, fo->confidence_level, fo->alpha,
	       fo->lambda_scaling, fo->method == ANOFOX_B200_RIDGE ? 0 : fo->hc_type};
	if (pred) o.compute_inference = false; // ols_predict_aggregate.cpp:354-364
#ifdef _OPENMP
	if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
	{
	TcritMemo<R> memo;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 64)
#endif
	for (int64_t g = 0; g < G; ++g) {
		int64_t r0 = offsets[g], r1 = offsets[g + 1];
		size_t n = (size_t)(r1 - r0);
		GroupResult r;
		std::vector<double> ty, tx, tw;
		size_t n_state = n;
		RowView d;
		d.p = (size_t)p;
		if (pred) {
			// training subset (ols_predict_aggregate.cpp:216-263)
			for (size_t i = 0; i < n; ++i) {
				size_t row = (size_t)r0 + i;
				// `train` is the aggregate Update's decision (split column, null_policy, per-method
				// NULL-feature rule); without it the OLS rule applies: y present and no NULL feature.
				bool tr;
				if (train) {
					tr = train[row] != 0;
				} else {
					tr = !std::isnan(y[row]);
					for (int32_t j = 0; j < p && tr; ++j) tr = !std::isnan(x[row * p + j]);
				}
				if (pred->is_training) pred->is_training[row] = tr ? 1 : 0;
				if (!tr) continue;
				ty.push_back(y[row]);
				for (int32_t j = 0; j < p; ++j) tx.push_back(x[row * p + j]);
				if (w) tw.push_back(w[row]);
			}
			n_state = ty.size();
			d.y = ty.data();
			d.xrm = tx.data();
			d.w = w ? tw.data() : nullptr;
			d.n = n_state;
		} else {
			d.y = y + r0;
			d.xrm = x + (size_t)r0 * p;
			d.w = w ? w + r0 : nullptr;
			d.n = n;
		}
		if (n_state < 2) {
			r.code = ANOFOX_B200_STATUS_TOO_FEW_ROWS; // ols_aggregate.cpp:263-267
			r.coef.assign((size_t)p, kNaN);
			// alpha < 0 still wins: the whole call is rejected upstream; keep per-group rule simple
		} else {
			fit_group<R>(o, d, r, &memo);
		}
		if (out) {
			out->status[g] = r.code;
			bool ok = r.code == 0;
			for (int32_t j = 0; j < p; ++j) out->coefficients[g * p + j] = ok ? r.coef[(size_t)j] : kNaN;
			out->intercept[g] = ok ? r.intercept : kNaN;
			out->r_squared[g] = ok ? r.r2 : kNaN;
			out->adj_r_squared[g] = ok ? r.adj : kNaN;
			out->residual_std_error[g] = ok ? r.rse : kNaN;
			out->n_observations[g] = ok ? (int64_t)r.n_obs : 0;
			if (o.compute_inference && out->std_errors) {
				bool hi = ok && r.has_inf;
				for (int32_t j = 0; j < p; ++j) {
					out->std_errors[g * p + j] = hi ? r.se[(size_t)j] : kNaN;
					out->t_values[g * p + j] = hi ? r.tv[(size_t)j] : kNaN;
					out->p_values[g * p + j] = hi ? r.pv[(size_t)j] : kNaN;
					out->ci_lower[g * p + j] = hi ? r.lo[(size_t)j] : kNaN;
					out->ci_upper[g * p + j] = hi ? r.hi[(size_t)j] : kNaN;
				}
				out->f_statistic[g] = hi ? r.fstat : kNaN;
				out->f_pvalue[g] = hi ? r.fp : kNaN;
				if (out->has_inference) out->has_inference[g] = hi ? 1 : 0;
			}
			if (out->aic) out->aic[g] = ok ? r.aic : kNaN;
			if (out->bic) out->bic[g] = ok ? r.bic : kNaN;
			if (out->rss) out->rss[g] = ok ? r.rss : kNaN;
		}
		if (pred) {
			double tc = kNaN;
			if (r.code == 0) {
				size_t k = (size_t)p + (std::isnan(r.intercept) ? 0 : 1);
				if (r.n_obs > k) tc = t_critical<R>(o.level, r.n_obs - k);
			}
			for (size_t i = 0; i < n; ++i) {
				size_t row = (size_t)r0 + i;
				AnofoxPredictionResult pr{kNaN, kNaN, kNaN};
				if (r.code == 0) {
					predict_row<R>(r.coef.data(), (size_t)p, r.intercept, x + row * p, r.rse, r.n_obs, o.level, tc,
					               &pr);
					if (!std::isfinite(pr.yhat)) pr = AnofoxPredictionResult{kNaN, kNaN, kNaN}; // :409-417
				}
				pred->yhat[row] = pr.yhat;
				pred->yhat_lower[row] = pr.yhat_lower;
				pred->yhat_upper[row] = pr.yhat_upper;
			}
		}
	}
	}
}

} // namespace

// ---------------------------------------------------------------------------------------------
// Window form: {ols,ridge,wls}_fit_predict(y, x, ...) OVER (PARTITION BY g ORDER BY t ROWS BETWEEN ...)
// (src/window_functions/ols_fit_predict.cpp:111-326; ridge / wls analogous). For every output row DuckDB runs the
// aggregate over the row's frame: Update keeps the rows with a non-NULL y as training data (:165-190) and
// remembers the x of the LAST frame row (:157-163); Finalize (:243-326) answers NULL when no frame row carried an
// x list or when the training rows are <= p + intercept, fits WITHOUT inference, and predicts at the remembered
// x with the simplified interval of anofox_predict_with_interval. Restated literally: one raw-row fit per output
// row. frame: rows [i - preceding, i - end_offset] of the partition (preceding < 0: unbounded).
// ---------------------------------------------------------------------------------------------
template <class R>
void window_fit_predict(const int64_t *offsets, int64_t G, const double *y, const double *x, const double *w, int32_t p,
                        const AnofoxB200FitOptions *fo, int64_t preceding, int64_t end_offset,
                        const AnofoxB200PredictOutputs *pred, int threads) {
	Opts o{fo->method, fo->fit_intercept != 0, false, fo->confidence_level, fo->alpha, fo->lambda_scaling, 0};
	const bool wls = fo->method == ANOFOX_B200_WLS;
#ifdef _OPENMP
	if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 4)
#endif
	for (int64_t g = 0; g < G; ++g) {
		const int64_t r0 = offsets[g], r1 = offsets[g + 1];
		for (int64_t r = r0; r < r1; ++r) {
			pred->yhat[r] = pred->yhat_lower[r] = pred->yhat_upper[r] = kNaN;
			const int64_t hi = r - end_offset;
			int64_t lo = preceding < 0 ? r0 : r - preceding;
			if (lo < r0) lo = r0;
			if (hi < lo) continue; // empty frame: the state was never initialised
			std::vector<double> ty, tx, tw;
			for (int64_t q = lo; q <= hi; ++q) {
				bool tr = !std::isnan(y[q]);               // y IS NOT NULL (:170-172)
				if (wls && w) tr = tr && !std::isnan(w[q]); // wls_fit_predict.cpp: weight must be present
				if (!tr) continue;
				ty.push_back(y[q]);
				for (int32_t j = 0; j < p; ++j) tx.push_back(x[q * p + j]);
				if (w) tw.push_back(w[q]);
			}
			const size_t min_obs = (size_t)p + (o.fit_intercept ? 1 : 0);
			if (ty.size() <= min_obs) continue; // :258-262
			RowView d;
			d.p = (size_t)p;
			d.y = ty.data();
			d.xrm = tx.data();
			d.w = w ? tw.data() : nullptr;
			d.n = ty.size();
			GroupResult res;
			fit_group<R>(o, d, res, nullptr);
			if (res.code != 0) continue; // :289-292
			AnofoxPredictionResult pr;
			predict_row<R>(res.coef.data(), (size_t)p, res.intercept, x + hi * p, res.rse, res.n_obs, o.level, kNaN, &pr);
			pred->yhat[r] = pr.yhat;
			pred->yhat_lower[r] = pr.yhat_lower;
			pred->yhat_upper[r] = pr.yhat_upper;
		}
	}
}


extern "C" {

bool oracle_ols_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxOlsOptions options,
                    AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference, AnofoxError *out_error) {
	Opts o{ANOFOX_B200_OLS, options.fit_intercept, options.compute_inference, options.confidence_level, 0.0, 0,
	       (int)options.hc_type};
	return single_fit(o, y, x, x_count, nullptr, out_core, out_inference, out_error);
}

bool oracle_ridge_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxRidgeOptions options,
                      AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference, AnofoxError *out_error) {
	Opts o{ANOFOX_B200_RIDGE, options.fit_intercept, options.compute_inference, options.confidence_level,
	       options.alpha,   (int)options.lambda_scaling};
	return single_fit(o, y, x, x_count, nullptr, out_core, out_inference, out_error);
}

bool oracle_wls_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t x_count, AnofoxDataArray weights,
                    AnofoxWlsOptions options, AnofoxFitResultCore *out_core, AnofoxFitResultInference *out_inference,
                    AnofoxError *out_error) {
	Opts o{ANOFOX_B200_WLS, options.fit_intercept, options.compute_inference, options.confidence_level, 0.0, 0,
	       (int)options.hc_type};
	return single_fit(o, y, x, x_count, &weights, out_core, out_inference, out_error);
}

void oracle_free_result_core(AnofoxFitResultCore *r) {
	if (!r) return;
	std::free(r->coefficients);
	r->coefficients = nullptr;
}

void oracle_free_result_inference(AnofoxFitResultInference *r) {
	if (!r) return;
	std::free(r->std_errors);
	std::free(r->t_values);
	std::free(r->p_values);
	std::free(r->ci_lower);
	std::free(r->ci_upper);
	r->std_errors = r->t_values = r->p_values = r->ci_lower = r->ci_upper = nullptr;
}

double oracle_t_critical(double confidence_level, size_t df) { return t_critical<long double>(confidence_level, df); }

bool oracle_predict_with_interval(const double *coefficients, size_t coefficients_len, double intercept,
                                  const double *x_new, size_t x_len, double residual_std_error,
                                  size_t n_observations, double confidence_level, AnofoxPredictionResult *out) {
	if (!out) return false;
	out->yhat = out->yhat_lower = out->yhat_upper = kNaN;
	if (!coefficients || coefficients_len == 0) return false;
	if (!x_new || x_len != coefficients_len) return false;
	predict_row<long double>(coefficients, coefficients_len, intercept, x_new, residual_std_error, n_observations,
	                         confidence_level, kNaN, out);
	return true;
}

bool oracle_compute_aic(double rss, size_t n, size_t k, double *out_aic, AnofoxError *err) {
	if (err) err->code = ANOFOX_ERROR_SUCCESS;
	if (!out_aic || n == 0 || rss < 0) {
		set_err(err, ANOFOX_ERROR_INVALID_INPUT, "invalid input");
		return false;
	}
	double a, b;
	info_criteria(rss, n, k, a, b);
	*out_aic = a;
	return true;
}

bool oracle_compute_bic(double rss, size_t n, size_t k, double *out_bic, AnofoxError *err) {
	if (err) err->code = ANOFOX_ERROR_SUCCESS;
	if (!out_bic || n == 0 || rss < 0) {
		set_err(err, ANOFOX_ERROR_INVALID_INPUT, "invalid input");
		return false;
	}
	double a, b;
	info_criteria(rss, n, k, a, b);
	*out_bic = b;
	return true;
}

// Distribution helpers exposed for the special-function tests.
double oracle_t_two_sided_p(double t, double df) { return (double)t_two_sided_p<long double>(t, df); }
double oracle_f_upper_p(double f, double d1, double d2) { return (double)f_upper_p<long double>(f, d1, d2); }

// Batch entry points over the same CSR layout the product takes (HOST pointers).
// precision: 0 = long double (checker), 1 = double (timed CPU baseline). threads <= 0: OpenMP default.
int32_t oracle_fit_batch(const int64_t *offsets, int64_t n_groups, const double *y, const double *x, const double *w,
                         int64_t n_rows, int32_t n_features, const AnofoxB200FitOptions *options,
                         const AnofoxB200FitOutputs *out, int32_t precision, int32_t threads) {
	(void)n_rows;
	if (!offsets || !y || !x || !options || !out || n_features <= 0) return ANOFOX_ERROR_INVALID_INPUT;
	if (options->method == ANOFOX_B200_RIDGE && options->alpha < 0.0) return ANOFOX_ERROR_INVALID_ALPHA;
	if (options->method == ANOFOX_B200_WLS && !w) return ANOFOX_ERROR_INVALID_INPUT;
	if (precision == 0)
		fit_batch<long double>(offsets, n_groups, y, x, w, n_features, options, out, nullptr, nullptr, threads);
	else
		fit_batch<double>(offsets, n_groups, y, x, w, n_features, options, out, nullptr, nullptr, threads);
	return 0;
}

int32_t oracle_fit_predict_batch(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                 const double *w, const uint8_t *train, int64_t n_rows, int32_t n_features,
                                 const AnofoxB200FitOptions *options, const AnofoxB200PredictOutputs *pred,
                                 const AnofoxB200FitOutputs *fit, int32_t precision, int32_t threads) {
	(void)n_rows;
	if (!offsets || !y || !x || !options || !pred || n_features <= 0) return ANOFOX_ERROR_INVALID_INPUT;
	if (options->method == ANOFOX_B200_RIDGE && options->alpha < 0.0) return ANOFOX_ERROR_INVALID_ALPHA;
	if (options->method == ANOFOX_B200_WLS && !w) return ANOFOX_ERROR_INVALID_INPUT;
	if (precision == 0)
		fit_batch<long double>(offsets, n_groups, y, x, w, n_features, options, fit, train, pred, threads);
	else
		fit_batch<double>(offsets, n_groups, y, x, w, n_features, options, fit, train, pred, threads);
	return 0;
}

int32_t oracle_window_fit_predict(const int64_t *offsets, int64_t n_groups, const double *y, const double *x,
                                  const double *w, int64_t n_rows, int32_t n_features,
                                  const AnofoxB200FitOptions *options, int64_t preceding, int64_t end_offset,
                                  const AnofoxB200PredictOutputs *pred, int32_t precision, int32_t threads) {
	(void)n_rows;
	if (precision == 1)
		window_fit_predict<double>(offsets, n_groups, y, x, w, n_features, options, preceding, end_offset, pred, threads);
	else
		window_fit_predict<long double>(offsets, n_groups, y, x, w, n_features, options, preceding, end_offset, pred,
		                                threads);
	return 0;
}

int32_t oracle_max_threads(void) {
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}

} // extern "C"
