// specfun.cuh -- Student-t and F distribution functions on the device (f64).
//
// Replaces what the reference obtains from statrs 0.18 (crates/anofox-stats-ffi/src/lib.rs:26,
// :2223-2230 StudentsT::inverse_cdf) and from anofox-regression's p-value code. Written from
// the standard formulas: regularised incomplete beta by a Lentz continued fraction, log-beta
// with the Stirling-correction form (as in R's lbeta) so that large degrees of freedom do not
// cancel, and a safeguarded Newton iteration for the quantile.
#pragma once
#include "common.cuh"

namespace b200 {

// Stirling correction  lgamma(x) - [(x - 1/2) ln x - x + ln sqrt(2 pi)],  x >= 10.
__device__ inline double lgamma_corr(double x) {
	double r = 1.0 / x, r2 = r * r;
	return r * (8.3333333333333333e-2 +
	            r2 * (-2.7777777777777778e-3 +
	                  r2 * (7.9365079365079365e-4 +
	                        r2 * (-5.9523809523809524e-4 + r2 * (8.4175084175084175e-4 + r2 * (-1.9175269175269175e-3))))));
}

__device__ inline double lbeta_dev(double a, double b) {
	double p = fmin(a, b), q = fmax(a, b);
	const double ln_sqrt_2pi = 0.91893853320467274178;
	if (p >= 10.0) {
		double corr = lgamma_corr(p) + lgamma_corr(q) - lgamma_corr(p + q);
		return -0.5 * log(q) + ln_sqrt_2pi + corr + (p - 0.5) * log(p / (p + q)) + q * log1p(-p / (p + q));
	}
	if (q >= 10.0) {
		double corr = lgamma_corr(q) - lgamma_corr(p + q);
		return lgamma(p) + corr + p - p * log(p + q) + (q - 0.5) * log1p(-p / (p + q));
	}
	return lgamma(p) + lgamma(q) - lgamma(p + q);
}

// Continued fraction for I_x(a,b) (modified Lentz); converges in O(sqrt(max(a,b))) steps for
// x < (a+1)/(a+b+2).
__device__ inline double beta_cf(double a, double b, double x) {
	const double tiny = 1e-300;
	double qab = a + b, qap = a + 1.0, qam = a - 1.0;
	double c = 1.0, d = 1.0 - qab * x / qap;
	if (fabs(d) < tiny) d = tiny;
	d = 1.0 / d;
	double h = d;
	for (int m = 1; m <= 5000; ++m) {
		double m2 = 2.0 * m;
		double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
		d = 1.0 + aa * d;
		if (fabs(d) < tiny) d = tiny;
		c = 1.0 + aa / c;
		if (fabs(c) < tiny) c = tiny;
		d = 1.0 / d;
		h *= d * c;
		aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
		d = 1.0 + aa * d;
		if (fabs(d) < tiny) d = tiny;
		c = 1.0 + aa / c;
		if (fabs(c) < tiny) c = tiny;
		d = 1.0 / d;
		double del = d * c;
		h *= del;
		if (fabs(del - 1.0) < 3e-16) break;
	}
	return h;
}

// I_x(a,b) with xc = 1 - x given by the caller (no cancellation near x = 1).
__device__ inline double beta_inc(double a, double b, double x, double xc) {
	if (!(x > 0.0)) return 0.0;
	if (!(xc > 0.0)) return 1.0;
	double lnpre = a * log(x) + b * log(xc) - lbeta_dev(a, b);
	if (x < (a + 1.0) / (a + b + 2.0)) return exp(lnpre) * beta_cf(a, b, x) / a;
	return 1.0 - exp(lnpre) * beta_cf(b, a, xc) / b;
}

// Large degrees of freedom: the continued fraction needs O(sqrt(df)) steps (a 10^8-row group: ~10^4 per
// evaluation), so above kLargeDf the two-sided tail comes from Hill's normalising expansion (CACM Algorithm
// 395, 1970): P(|T_df| > t) = erfc(z / sqrt 2) with z a series in 1 / (48 (df - 1/2)^2). Relative error
// against a 60-digit reference <= 8e-11 for df >= 4097 down to tails of 1e-300 (tests/test_gpu_parity.py).
constexpr double kLargeDf = 4096.0;
__device__ inline double t_two_sided_p_large(double t2, double df) {
	const double a = df - 0.5, b = 48.0 * a * a;
	const double y = a * log1p(t2 / df);
	const double z = (((((-0.4 * y - 3.3) * y - 24.0) * y - 85.5) / (0.8 * y * y + 100.0 + b) + y + 3.0) / b + 1.0) * sqrt(y);
	return erfc(z * 0.70710678118654752440);
}

// 2 * (1 - T_df(|t|)) = I_{df/(df+t^2)}(df/2, 1/2)
__device__ inline double t_two_sided_p(double t, double df) {
	if (isnan(t) || !(df > 0.0)) return nan_f64();
	if (isinf(t)) return 0.0;
	double t2 = t * t;
	if (df > kLargeDf) return t_two_sided_p_large(t2, df);
	return beta_inc(0.5 * df, 0.5, df / (df + t2), t2 / (df + t2));
}

// Upper tail of F(d1, d2) for integer d1 and d2 > kLargeDf: I_x(a, b), a = d2/2, b = d1/2, built up from
// b = 1 (x^a) or b = 1/2 (the t tail above) with I_x(a, b+1) = I_x(a, b) + x^a y^b / (b B(a, b)): at most
// d1/2 positive terms, x^a = exp(-a log1p(d1 f / d2)) applied last (the sum can be 1e50 times x^a).
__device__ inline double f_upper_p_large(double f, double d1, double d2) {
	const double a = 0.5 * d2, b = 0.5 * d1;
	const double y = d1 * f / (d2 + d1 * f);
	const double lnxa = -a * log1p(d1 * f / d2);
	// result = base + exp(lnbase) S, S = 1 + r_1 + r_2 + ... kept apart from the (possibly underflowing) x^a
	double base = 0.0, lnbase = lnxa, bp = 1.0;
	if (fmod(d1, 2.0) != 0.0) {
		base = t_two_sided_p_large(d1 * f, d2); // I_x(a, 1/2)
		if (b < 1.0) return base;
		lnbase = 0.69314718055994530942 + lnxa + 0.5 * log(y) - lbeta_dev(a, 0.5); // T_{1/2}
		bp = 1.5;
	}
	double S = 1.0, r = 1.0, lnscale = 0.0;
	while (bp < b - 0.25) {
		r *= y * (a + bp - 1.0) / bp;
		S += r;
		bp += 1.0;
		if (S > 1e200) {
			S *= 1e-200;
			r *= 1e-200;
			lnscale += 460.51701859880913680; // 200 ln 10
		}
	}
	return fmin(base + exp(lnbase + lnscale + log(S)), 1.0);
}

// Upper tail of F(d1,d2)
__device__ inline double f_upper_p(double f, double d1, double d2) {
	if (isnan(f) || !(d1 > 0.0) || !(d2 > 0.0)) return nan_f64();
	if (f <= 0.0) return 1.0;
	if (isinf(f)) return 0.0;
	if (d2 > kLargeDf && d1 == floor(d1) && d1 <= 4096.0) return f_upper_p_large(f, d1, d2);
	double den = d2 + d1 * f;
	return beta_inc(0.5 * d2, 0.5 * d1, d2 / den, d1 * f / den);
}

// t such that P(|T_df| <= t) = level  (anofox_t_critical, lib.rs:2217-2231).
__device__ inline double t_critical_dev(double level, double df) {
	if (!(df > 0.0) || !(level > 0.0) || !(level < 1.0)) return nan_f64();
	const double alpha = 1.0 - level; // two-sided tail mass
	if (df == 1.0) return tan(1.5707963267948966 * level);
	if (df == 2.0) return level * sqrt(2.0 / ((1.0 - level) * (1.0 + level)));
	// Cornish-Fisher start from the normal quantile
	double z = normcdfinv(1.0 - 0.5 * alpha);
	double z2 = z * z;
	double t = z + z * (z2 + 1.0) / (4.0 * df) + z * (5.0 * z2 * z2 + 16.0 * z2 + 3.0) / (96.0 * df * df) +
	           z * (3.0 * z2 * z2 * z2 + 19.0 * z2 * z2 + 17.0 * z2 - 15.0) / (384.0 * df * df * df);
	if (!(t > 0.0) || isinf(t)) t = z > 0.0 ? z : 1.0;
	const double lnc = -lbeta_dev(0.5, 0.5 * df) - 0.5 * log(df); // log of the pdf normaliser
	double lo = 0.0, hi = -1.0;                                   // bracket: f(lo) > 0 > f(hi)
	for (int it = 0; it < 100; ++it) {
		double f = t_two_sided_p(t, df) - alpha;
		if (f > 0.0)
			lo = t;
		else
			hi = t;
		double pdf = exp(lnc - 0.5 * (df + 1.0) * log1p(t * t / df));
		double step = f / (2.0 * pdf); // Newton: t_new = t + f / (2 pdf)  (f' = -2 pdf)
		double tn = t + step;
		if (!(tn > lo) || (hi > 0.0 && !(tn < hi)) || isnan(tn)) tn = hi > 0.0 ? 0.5 * (lo + hi) : 2.0 * t + 1.0;
		if (fabs(tn - t) <= 4e-16 * tn) {
			t = tn;
			break;
		}
		t = tn;
	}
	return t;
}

// ---------------------------------------------------------------------------------------------
// Fast paths for INTEGER degrees of freedom (every p-value of a fit has df = n - rank).
// Tables built once per confidence level (df_tables_kernel, cached by the host):
//   tcrit[df]  = t_{(1+level)/2, df}
//   gratio[df] = Gamma((df+1)/2) / Gamma(df/2)
//   rcp[m]     = 1 / m
//   ce[i] = prod_{k=1..i} (2k-1)/(2k),  co[i] = prod_{k=1..i} 2k/(2k+1)      (level-independent)
// ce / co are the coefficients of the binomial series (1-x)^(-1/2) = sum ce[j] x^j and of its
// odd-df counterpart. With y = 1 - x = t^2/(df + t^2):
//   df = 2m   : P(|T| > t) = 1 - sqrt(y) sum_{j<m} ce[j] x^j      = sqrt(y) sum_{j>=m} ce[j] x^j
//   df = 2m+1 : P(|T| > t) = (2/pi)(atan2(sqrt x, sqrt y) - sqrt(xy) sum_{j<m} co[j] x^j)
//                          = (2/pi) sqrt(xy) sum_{j>=m} co[j] x^j
// i.e. the classical finite sum (A&S 26.7.3-4, used in the body of the distribution, m terms) and its
// own remainder (all terms positive: full relative accuracy in the far tail, geometric convergence
// with ratio x). Both are evaluated by Horner's rule straight from the coefficient table: one load
// and one multiply-add per term, no division, no loop-carried exit test.
// ---------------------------------------------------------------------------------------------
constexpr int kRcpTable = 2 * kTcritTable + 1024;
constexpr int kSeriesTerms = 500;                             // tail terms at most (x <= 0.85 needs 242)
constexpr int kSeriesTable = kTcritTable / 2 + kSeriesTerms + 32; // ce / co entries (+ Horner round-up slack)

struct DfTables {
	const double *tcrit;
	const double *gratio;
	const double *rcp;
	const double *ce;
	const double *co;
};

// sum_{n=0}^{N8-1} tab[m+n] x^n with N8 = N rounded up to a multiple of 8 (the extra terms are genuine
// terms of the series, so rounding up only adds accuracy; the tables carry the slack).
__device__ __forceinline__ double horner_from(const double *__restrict__ tab, int m, int N, double x) {
	const int N8 = (N + 7) & ~7;
	const double *t = tab + m + N8;
	double acc = 0.0;
	for (int n = N8; n > 0; n -= 8) {
		t -= 8;
		const double c7 = __ldg(t + 7), c6 = __ldg(t + 6), c5 = __ldg(t + 5), c4 = __ldg(t + 4);
		const double c3 = __ldg(t + 3), c2 = __ldg(t + 2), c1 = __ldg(t + 1), c0 = __ldg(t);
		acc = fma(acc, x, c7);
		acc = fma(acc, x, c6);
		acc = fma(acc, x, c5);
		acc = fma(acc, x, c4);
		acc = fma(acc, x, c3);
		acc = fma(acc, x, c2);
		acc = fma(acc, x, c1);
		acc = fma(acc, x, c0);
	}
	return acc;
}

// sum_{n=0}^{N-1} tab[n] x^n, exactly N terms (the finite sums of the body of the distribution).
__device__ __forceinline__ double horner_exact(const double *__restrict__ tab, int N, double x) {
	const double *t = tab + N;
	double acc = 0.0;
	const int r = N & 7;
	for (int i = 0; i < r; ++i) {
		--t;
		acc = fma(acc, x, __ldg(t));
	}
	for (int n = N - r; n > 0; n -= 8) {
		t -= 8;
		const double c7 = __ldg(t + 7), c6 = __ldg(t + 6), c5 = __ldg(t + 5), c4 = __ldg(t + 4);
		const double c3 = __ldg(t + 3), c2 = __ldg(t + 2), c1 = __ldg(t + 1), c0 = __ldg(t);
		acc = fma(acc, x, c7);
		acc = fma(acc, x, c6);
		acc = fma(acc, x, c5);
		acc = fma(acc, x, c4);
		acc = fma(acc, x, c3);
		acc = fma(acc, x, c2);
		acc = fma(acc, x, c1);
		acc = fma(acc, x, c0);
	}
	return acc;
}

// x^m, m >= 0, by repeated squaring: <= 2 log2(m) multiplications, relative error ~ m ulp (m <= 2048).
__device__ __forceinline__ double pow_int(double x, int m) {
	double r = 1.0;
	while (m > 0) {
		if (m & 1) r *= x;
		x *= x;
		m >>= 1;
	}
	return r;
}

// Remainder series: I_x(nu/2, 1/2) = x^(nu/2) sqrt(y) {1 | 2/pi} sum_{n>=0} tab[m+n] x^n. c = sqrt(x).
__device__ inline double betai_half_tail_series(int nu, double x, double y, double c, const DfTables &T) {
	const int m = nu >> 1;
	// terms until x^N < 1e-17: N = 39.2 / -ln x; single precision is plenty for a term count
	int N = (int)(-39.2f / __logf(fmaxf((float)x, 1e-30f))) + 3;
	if (N > kSeriesTerms) N = kSeriesTerms;
	const double xm = pow_int(x, m);
	if (nu & 1) return 0.63661977236758134308 * xm * c * sqrt(y) * horner_from(T.co, m, N, x);
	return xm * sqrt(y) * horner_from(T.ce, m, N, x);
}

// I_x(nu/2, 1/2) = P(|T_nu| > t) with x = nu / (nu + t^2), y = 1 - x, nu integer in [1, kTcritTable].
__device__ inline double betai_half_fast(int nu, double x, double y, const DfTables &T) {
	if (!(x > 0.0)) return 0.0;
	if (!(y > 0.0)) return 1.0;
	const double c = sqrt(x);
	if (x <= 0.6) return betai_half_tail_series(nu, x, y, c, T);
	if (nu <= 400) {
		const double s = sqrt(y);
		const int m = nu >> 1;
		double p;
		if (nu & 1)
			p = 0.63661977236758134308 * (atan2(c, s) - (m > 0 ? s * c * horner_exact(T.co, m, x) : 0.0)); // 2/pi
		else
			p = 1.0 - s * horner_exact(T.ce, m, x);
		if (p > 1e-4) return p; // 1 - A (df/2 rounded terms) keeps >= 11 digits
	}
	if (x <= 0.85) return betai_half_tail_series(nu, x, y, c, T);
	return beta_inc(0.5 * (double)nu, 0.5, x, y);
}

__device__ inline double t_two_sided_p_fast(double t, double df, const DfTables &T) {
	if (isnan(t) || !(df > 0.0)) return nan_f64();
	if (isinf(t)) return 0.0;
	const double t2 = t * t, den = df + t2;
	if (df > kLargeDf) return t_two_sided_p_large(t2, df);
	if (T.ce == nullptr || df != floor(df)) return beta_inc(0.5 * df, 0.5, df / den, t2 / den);
	return betai_half_fast((int)df, df / den, t2 / den, T);
}

// Upper tail of F(d1, d2), integer d1 >= 1 and d2 in [1, kTcritTable]:
// I_x(a, b) built up from b = 1/2 (or 1) with the recurrence I_x(a, b+1) = I_x(a, b) + T_b,
// T_b = x^a y^b Gamma(a+b) / (Gamma(a) Gamma(b+1)) -- all terms positive, no cancellation.
__device__ inline double f_upper_p_fast(double f, double d1, double d2, const DfTables &T) {
	if (isnan(f) || !(d1 > 0.0) || !(d2 > 0.0)) return nan_f64();
	if (f <= 0.0) return 1.0;
	if (isinf(f)) return 0.0;
	const double den = d2 + d1 * f;
	const double x = d2 / den, y = d1 * f / den;
	if (d2 > kLargeDf && d1 == floor(d1) && d1 <= 4096.0) return f_upper_p_large(f, d1, d2);
	if (T.ce == nullptr || d2 > (double)kTcritTable || d2 != floor(d2) || d1 != floor(d1) || d1 > 1024.0)
		return beta_inc(0.5 * d2, 0.5 * d1, x, y);
	const int nu = (int)d2, m = (int)d1;
	const double a = 0.5 * d2;
	double acc, term = 0.0;
	int twob; // 2b of the term about to be added
	if (m & 1) {
		acc = betai_half_fast(nu, x, y, T); // m == 1: the two-sided t test, f = t^2
		if (m > 1) { // x^a sqrt(y) Gamma(a + 1/2) / (Gamma(a) Gamma(3/2)): T_{1/2}
			const double xa = (nu & 1) ? pow_int(x, nu >> 1) * sqrt(x) : pow_int(x, nu >> 1);
			term = xa * sqrt(y) * T.gratio[nu] * 1.12837916709551257390; // 2/sqrt(pi)
		}
		twob = 1;
	} else {
		const double xa = (nu & 1) ? pow_int(x, nu >> 1) * sqrt(x) : pow_int(x, nu >> 1);
		acc = xa;
		term = xa * y * a; // T_1
		twob = 2;
	}
	while (twob < m) {
		acc += term;
		// T_{b+1} = T_b * y * (a+b)/(b+1),  (a+b)/(b+1) = (nu + 2b) / (2b + 2)
		term *= y * (double)(nu + twob) * T.rcp[twob + 2];
		twob += 2;
	}
	return fmin(acc, 1.0);
}

// The unified p-value entry point of the fit kernels: upper tail of F(d1, d2); the two-sided t test with df
// degrees of freedom is pvalue_f(t^2, 1, df). Kept out of line: one copy of the (large) distribution code
// per kernel instead of one per call site.
static __device__ __noinline__ double pvalue_f(double f, double d1, double d2, const DfTables T) {
	return f_upper_p_fast(f, d1, d2, T);
}

// One thread per table entry.
static __global__ void df_tables_kernel(double level, double *__restrict__ tcrit, double *__restrict__ gratio,
                                 double *__restrict__ rcp, double *__restrict__ ce, double *__restrict__ co) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i <= kRcpTable) rcp[i] = i > 0 ? 1.0 / (double)i : 0.0;
	if (i <= kTcritTable) {
		tcrit[i] = i == 0 ? nan_f64() : t_critical_dev(level, (double)i);
		// Gamma((df+1)/2) / Gamma(df/2) = sqrt(pi) / B(df/2, 1/2)
		gratio[i] = i == 0 ? 0.0 : 1.7724538509055160273 * exp(-lbeta_dev(0.5 * (double)i, 0.5));
	}
	if (i < kSeriesTable) {
		// ce[i] = Gamma(i+1/2) / (sqrt(pi) i!) = 1 / (i B(i, 1/2));  co[i] = pi / ((2i+1) B(i+1/2, 1/2))
		const double di = (double)i;
		ce[i] = i == 0 ? 1.0 : exp(-lbeta_dev(di, 0.5)) / di;
		co[i] = 3.14159265358979323846 * exp(-lbeta_dev(di + 0.5, 0.5)) / (2.0 * di + 1.0);
	}
}

} // namespace b200
