// fused_launch.h -- host-side launchers of the fused small-group kernel, one translation unit per feature
// count (fused_p*.cu) so that the fully unrolled instantiations compile in parallel.
#pragma once
#include "fused_small.cuh"

namespace b200 {

constexpr int kSmallMaxFeatures = 8;
// consumer warps per CTA: bounded by the registers the per-lane moment triangle needs
constexpr int small_warps(int p) { return p <= 5 ? 12 : 8; }

#define B200_DECLARE_SMALL(P_)                                                                                       \
	cudaError_t launch_small_p##P_(bool fused, bool has_w, bool has_t, const SmallArgs &a, int grid, cudaStream_t st); \
	cudaError_t launch_solve_small_p##P_(const double *records, int64_t n_groups, const FitParams &P,                \
	                                     const AnofoxB200FitOutputs &o, int *refine_count, int64_t *refine_list,     \
	                                     cudaStream_t st);
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

#define B200_DEFINE_SMALL(P_)                                                                                        \
	cudaError_t launch_small_p##P_(bool fused, bool has_w, bool has_t, const SmallArgs &a, int grid,                 \
	                               cudaStream_t st) {                                                                \
		constexpr int W = small_warps(P_);                                                                           \
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
	                                     cudaStream_t st) {                                                          \
		const unsigned blocks = (unsigned)((n_groups + 127) / 128);                                                  \
		solve_small_kernel<P_><<<blocks, 128, 0, st>>>(records, n_groups, P, o, refine_count, refine_list);          \
		return cudaGetLastError();                                                                                   \
	}

} // namespace b200
