// common.cuh -- shared definitions of the B200 group-fit kernels (sm_100a only).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/anofox_b200.h"

namespace b200 {

// -------------------------------------------------------------------------------------------
// Moment record: what the accumulation stage hands to the solve stage, one per group.
// All sums are taken over VALID rows (finite y and x, w > 0 and finite, training flag set)
// of d = z - shift, z = (x_1 .. x_p, y), shift = the group's first valid row (which is also
// the row the reference's constant-column test compares against, ols.rs:76-87).
//   [0] n_valid   [1] sum w   [2] n_state (rows that entered the aggregate state)   [3] spare
//   shift[k]      k = p + 1
//   cdev[p]       max |x_j - shift_j| over valid rows (then 0..3 pad fields)
//   S1[k]         sum w d
//   S2[k(k+1)/2]  sum w d d^T, packed upper triangle, row-major: (i,j), i <= j
// n, sw, S1, S2 add across row shards that share a shift; cdev combines with max.
// -------------------------------------------------------------------------------------------
__host__ __device__ inline int rec_off_shift(int) { return 4; }
__host__ __device__ inline int rec_off_cdev(int p) { return 4 + (p + 1); }
// the sums S1 | S2 start at a multiple of 4 fields (the small-group kernel leaves sum j in lane j % 4)
__host__ __device__ inline int rec_off_s1(int p) { return (4 + (p + 1) + p + 3) & ~3; }
__host__ __device__ inline int rec_off_s2(int p) { return rec_off_s1(p) + (p + 1); }
__host__ __device__ inline int rec_len(int p) { return rec_off_s2(p) + (p + 1) * (p + 2) / 2; }
__host__ __device__ inline int tri_idx(int i, int j, int k) { return i * k - (i * (i - 1)) / 2 + (j - i); } // i <= j

constexpr int kMaxFeatures = 64;
constexpr int kMaxStageGroups = 512; // groups packed into one smem stage at most
constexpr int kTcritTable = 4096;    // t-critical values precomputed for df = 1 .. kTcritTable

struct FitParams {
	int p;
	int method;
	int fit_intercept;
	int compute_inference;
	int lambda_scaling;
	double level;
	double alpha;
	const double *tcrit;  // [kTcritTable + 1] t_{(1+level)/2, df}, index = df   (may be null: evaluate directly)
	const double *gratio; // [kTcritTable + 1] Gamma((df+1)/2) / Gamma(df/2)
	const double *rcp;    // [kRcpTable + 1]  1 / m
	const double *ce;     // [kSeriesTable] binomial-series coefficients for even df (specfun.cuh)
	const double *co;     // [kSeriesTable] ... and for odd df
	int predict_mode;    // 1: df for t-critical follows anofox_predict_with_interval
	int min_state_rows;  // 2 = aggregate rule (ols_aggregate.cpp:263-267); 0 = direct anofox_*_fit call
	// QR fallback queue (qr.cuh): groups whose Cholesky met a small pivot are appended here by the solve
	// kernels and re-fitted from their ROWS by qr_fit_kernel. null = no queue (moment-only entry points).
	int *qr_count;
	int64_t *qr_list;
};

// ---- ill-conditioning probe of the Gram-Cholesky solve ("QR fallback") ---------------
// The solve kernels factor the shifted Gram matrix, whose condition number is cond(X)^2. The ratio d / a0 of a
// column's remaining pivot to its centred diagonal is 1 - R^2 of that column on the earlier ones; the smallest
// ratio bounds the condition number of the equilibrated Gram from below by its reciprocal and, loosely, from
// above by p times it. With unit roundoff 1.1e-16 and an error constant of ~10 a Cholesky solve keeps 1e-9 up
// to cond(Gram) ~ 1e6: below kQrPivotRatio = 1e-5 it can no longer promise that. Such a group -- and any group
// where a column's aliasing decision d >= 1e-14 * orig (R's tol = 1e-7, squared) sits within half a decade of its threshold, where the Gram's own
// rounding could flip it -- is queued for the Householder QR on the rows themselves (cond(X), not cond(X)^2,
// and the reference default is a decomposition of X: types.rs:121-130 SVD, "falls back to QR").
constexpr double kQrPivotRatio = 1e-5;
__device__ __forceinline__ bool chol_pivot_suspicious(double d, double a0, double orig, bool ridge) {
	if (!(d > 0.0)) return false; // exactly dependent (or worse): aliased for certain
	if (!ridge) {
		if (d >= 3e-15 * orig && d < 3e-14 * orig) return true; // decision band around 1e-14 * orig
		if (d < 1e-14 * orig) return false;                     // aliased, and clear of the band
	}
	return d < kQrPivotRatio * a0;
}
__device__ __forceinline__ void queue_qr(const FitParams &P, int64_t g) {
	if (P.qr_count) P.qr_list[atomicAdd(P.qr_count, 1)] = g;
}

__device__ __forceinline__ bool finite_f64(double v) {
	return (__double2hiint(v) & 0x7ff00000) != 0x7ff00000;
}

__device__ __forceinline__ double nan_f64() { return __longlong_as_double(0x7ff8000000000000LL); }

// ---- mbarrier / bulk-copy (TMA) primitives ------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
	asm volatile(
	    "{\n"
	    " .reg .pred p;\n"
	    "WAIT_%=:\n"
	    " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
	    " @p bra DONE_%=;\n"
	    " bra WAIT_%=;\n"
	    "DONE_%=:\n"
	    "}\n" ::"r"(smem_u32(bar)),
	    "r"(parity), "r"(20000u) // suspend-time hint (ns): sleep in hardware instead of spinning
	    : "memory");
}
// Same with a run-time suspend hint; hint 0 = pure spin on test_wait (experiments: wake-up latency of a suspended warp)
__device__ __forceinline__ void mbar_wait_hint(uint64_t *bar, uint32_t parity, uint32_t hint_ns) {
	if (hint_ns == 0u) {
		uint32_t ok;
		do {
			asm volatile(
			    "{\n"
			    " .reg .pred p;\n"
			    " mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
			    " selp.u32 %0, 1, 0, p;\n"
			    "}\n"
			    : "=r"(ok)
			    : "r"(smem_u32(bar)), "r"(parity)
			    : "memory");
		} while (ok == 0u);
		return;
	}
	asm volatile(
	    "{\n"
	    " .reg .pred p;\n"
	    "WAIT_%=:\n"
	    " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
	    " @p bra DONE_%=;\n"
	    " bra WAIT_%=;\n"
	    "DONE_%=:\n"
	    "}\n" ::"r"(smem_u32(bar)),
	    "r"(parity), "r"(hint_ns)
	    : "memory");
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t bar_addr, uint32_t parity) {
	asm volatile(
	    "{\n"
	    " .reg .pred p;\n"
	    "WAIT_%=:\n"
	    " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
	    " @p bra DONE_%=;\n"
	    " bra WAIT_%=;\n"
	    "DONE_%=:\n"
	    "}\n" ::"r"(bar_addr),
	    "r"(parity), "r"(20000u)
	    : "memory");
}
// non-blocking: has the phase with this parity completed?
__device__ __forceinline__ bool mbar_test_addr(uint32_t bar_addr, uint32_t parity) {
	uint32_t ok;
	asm volatile(
	    "{\n"
	    " .reg .pred p;\n"
	    " mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
	    " selp.u32 %0, 1, 0, p;\n"
	    "}\n"
	    : "=r"(ok)
	    : "r"(bar_addr), "r"(parity)
	    : "memory");
	return ok != 0u;
}
__device__ __forceinline__ void st_release_s64(long long *p, long long v) {
	asm volatile("st.release.cta.shared.s64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release_s32(int *p, int v) {
	asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_volatile_s32(const int *p) {
	int v;
	asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
	return v;
}
__device__ __forceinline__ longlong2 ld_volatile_s64x2(const long long *p) {
	longlong2 v;
	asm volatile("ld.volatile.shared.v2.s64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(smem_u32(p)));
	return v;
}
__device__ __forceinline__ void fence_acq_rel_cta() { asm volatile("fence.acq_rel.cta;" ::: "memory"); }
// One contiguous global -> shared bulk copy (SASS: UBLKCP). dst/src 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
	                 smem_u32(dst)),
	             "l"(src), "r"(bytes), "r"(smem_u32(bar))
	             : "memory");
}

// FP64 tensor-core MMA, D(8x8) += A(8x4, row) * B(4x8, col).  SASS: DMMA.8x8x4.
// Fragment layout (PTX ISA): a = A[lane/4][lane%4], b = B[lane%4][lane/4],
// c0,c1 = C[lane/4][2*(lane%4) + {0,1}].
__device__ __forceinline__ void dmma_884(double &c0, double &c1, double a, double b) {
	asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
	             : "+d"(c0), "+d"(c1)
	             : "d"(a), "d"(b));
}

} // namespace b200
