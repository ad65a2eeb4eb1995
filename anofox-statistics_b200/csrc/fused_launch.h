// fused_launch.h -- host-side launchers of the fused small-group kernel, one translation unit per feature
// count (fused_p*.cu) so that the fully unrolled instantiations compile in parallel.
#pragma once
#include "fused_small.cuh"

namespace b200 {

constexpr int kSmallMaxFeatures = 8;
// consumer warps per CTA: bounded by the registers the per-lane moment triangle needs
constexpr int small_warps(int p) { return p <= 5 ? 12 : 8; }

#define B200_DECLARE_SMALL(P_) cudaError_t launch_small_p##P_(bool has_w, bool has_t, const SmallArgs &a, int grid, cudaStream_t st);
B200_DECLARE_SMALL(1)
B200_DECLARE_SMALL(2)
B200_DECLARE_SMALL(3)
B200_DECLARE_SMALL(4)
B200_DECLARE_SMALL(5)
B200_DECLARE_SMALL(6)
B200_DECLARE_SMALL(7)
B200_DECLARE_SMALL(8)
#undef B200_DECLARE_SMALL

template <int P, int NCW, bool HASW, bool HAST> cudaError_t launch_small_t(const SmallArgs &a, int grid, cudaStream_t st) {
	auto kern = fit_small_kernel<P, NCW, HASW, HAST>;
	const size_t dyn = (size_t)a.acc.n_stages * a.acc.stage_bytes;
	cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
	if (e != cudaSuccess) return e;
	kern<<<grid, 32 * (NCW + 1), dyn, st>>>(a);
	return cudaGetLastError();
}

#define B200_DEFINE_SMALL(P_)                                                                                        \
	cudaError_t launch_small_p##P_(bool has_w, bool has_t, const SmallArgs &a, int grid, cudaStream_t st) {          \
		constexpr int W = small_warps(P_);                                                                           \
		if (has_w) {                                                                                                 \
			if (has_t) return launch_small_t<P_, W, true, true>(a, grid, st);                                        \
			return launch_small_t<P_, W, true, false>(a, grid, st);                                                  \
		}                                                                                                            \
		if (has_t) return launch_small_t<P_, W, false, true>(a, grid, st);                                           \
		return launch_small_t<P_, W, false, false>(a, grid, st);                                                     \
	}

} // namespace b200
