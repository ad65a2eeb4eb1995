// fused_launch.h -- host-side launchers of the fused small-group kernel, one translation unit per feature
// count (fused_p*.cu) so that the fully unrolled instantiations compile in parallel.
#pragma once
#include <cstdlib>
#include "fused_small.cuh"

namespace b200 {

constexpr int kSmallMaxFeatures = 8;
// consumer warps per CTA: bounded by the registers the per-lane moment triangle needs
constexpr int small_warps(int p) { return p <= 5 ? 12 : 8; }
// stash kernel: bytes of one 32-slot table of parked sums; accumulate and solve warps per CTA. Every accumulate
// warp owns two tables; the ring gets the rest of the shared memory.
constexpr int stash_bytes(int p, bool has_w) { return (2 + (has_w ? 1 : 0) + 2 * (p + 1) + (p + 1) * (p + 2) / 2) * 32 * 8; }
#ifndef ANOFOX_STASH_NACC
#define ANOFOX_STASH_NACC 3
#endif
#ifndef ANOFOX_STASH_NSOL
#define ANOFOX_STASH_NSOL 12
#endif
constexpr int stash_acc_warps(int p) { return p <= 5 ? ANOFOX_STASH_NACC : 2; }
constexpr int stash_solve_warps(int p) { return p <= 5 ? ANOFOX_STASH_NSOL : 9; }
constexpr int stash_pool_bytes(int p, bool has_w) { return 2 * stash_acc_warps(p) * stash_bytes(p, has_w); }

#define B200_DECLARE_SMALL(P_)                                                                                       \
	cudaError_t launch_small_p##P_(int mode, bool has_w, bool has_t, const SmallArgs &a, int grid, cudaStream_t st);   \
	cudaError_t launch_solve_small_p##P_(const double *records, int64_t n_groups, const FitParams &P,                \
	                                     const AnofoxB200FitOutputs &o, int *refine_count, int64_t *refine_list,     \
	                                     int defer_p, cudaStream_t st);
B200_DECLARE_SMALL(1)
B200_DECLARE_SMALL(2)
B200_DECLARE_SMALL(3)
B200_DECLARE_SMALL(4)
B200_DECLARE_SMALL(5)
B200_DECLARE_SMALL(6)
B200_DECLARE_SMALL(7)
B200_DECLARE_SMALL(8)
#undef B200_DECLARE_SMALL

template <int P, int NCW, bool HASW, bool HAST, bool FUSED>
cudaError_t launch_small_t(const SmallArgs &a, int grid, cudaStream_t st) {
	auto kern = fit_small_kernel<P, NCW, HASW, HAST, FUSED>;
	const size_t dyn = (size_t)a.acc.n_stages * a.acc.stage_bytes;
	cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
	if (e != cudaSuccess) return e;
	kern<<<grid, 32 * (NCW + 1), dyn, st>>>(a);
	return cudaGetLastError();
}

template <int P, bool HASW, bool HAST> cudaError_t launch_stash_t(const SmallArgs &a, int grid, cudaStream_t st) {
	constexpr int NA = stash_acc_warps(P), NSV = stash_solve_warps(P);
	static_assert(SmallStash<P, HASW>::BYTES == stash_bytes(P, HASW), "stash size");
	auto kern = fit_small_stash_kernel<P, NA, NSV, HASW, HAST>;
	const size_t dyn = (size_t)a.acc.n_stages * a.acc.stage_bytes + (size_t)stash_pool_bytes(P, HASW);
	cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
	if (e != cudaSuccess) return e;
	kern<<<grid, 32 * (1 + NA + NSV), dyn, st>>>(a);
	return cudaGetLastError();
}

#define B200_DEFINE_SMALL(P_)                                                                                        \
	cudaError_t launch_small_p##P_(int mode, bool has_w, bool has_t, const SmallArgs &a, int grid,                   \
	                               cudaStream_t st) {                                                                \
		constexpr int W = small_warps(P_);                                                                           \
		const bool fused = mode == 1;                                                                                \
		if (mode == 2) {                                                                                             \
			switch ((has_w ? 2 : 0) | (has_t ? 1 : 0)) {                                                             \
			case 0: return launch_stash_t<P_, false, false>(a, grid, st);                                            \
			case 1: return launch_stash_t<P_, false, true>(a, grid, st);                                             \
			case 2: return launch_stash_t<P_, true, false>(a, grid, st);                                             \
			default: return launch_stash_t<P_, true, true>(a, grid, st);                                             \
			}                                                                                                        \
		}                                                                                                            \
		const int sel = (fused ? 4 : 0) | (has_w ? 2 : 0) | (has_t ? 1 : 0);                                         \
		switch (sel) {                                                                                               \
		case 0: return launch_small_t<P_, W, false, false, false>(a, grid, st);                                      \
		case 1: return launch_small_t<P_, W, false, true, false>(a, grid, st);                                       \
		case 2: return launch_small_t<P_, W, true, false, false>(a, grid, st);                                       \
		case 3: return launch_small_t<P_, W, true, true, false>(a, grid, st);                                        \
		case 4: return launch_small_t<P_, W, false, false, true>(a, grid, st);                                       \
		case 5: return launch_small_t<P_, W, false, true, true>(a, grid, st);                                        \
		case 6: return launch_small_t<P_, W, true, false, true>(a, grid, st);                                        \
		default: return launch_small_t<P_, W, true, true, true>(a, grid, st);                                        \
		}                                                                                                            \
	}                                                                                                                \
	cudaError_t launch_solve_small_p##P_(const double *records, int64_t n_groups, const FitParams &P,                \
	                                     const AnofoxB200FitOutputs &o, int *refine_count, int64_t *refine_list,     \
	                                     int defer_p, cudaStream_t st) {                                             \
		static const int thr = getenv("ANOFOX_B200_SOLVE_THREADS") ? atoi(getenv("ANOFOX_B200_SOLVE_THREADS")) : 128; \
		const unsigned blocks = (unsigned)((n_groups + thr - 1) / thr);                                              \
		static const int dyn_kb = getenv("ANOFOX_B200_SOLVE_DYN_KB") ? atoi(getenv("ANOFOX_B200_SOLVE_DYN_KB")) : 0; \
		if (dyn_kb > 0)                                                                                              \
			cudaFuncSetAttribute(solve_small_kernel<P_>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_kb * 1024); \
		solve_small_kernel<P_><<<blocks, thr, (size_t)dyn_kb * 1024, st>>>(records, n_groups, P, o, refine_count,    \
		                                                                   refine_list, defer_p);                    \
		return cudaGetLastError();                                                                                   \
	}

} // namespace b200
