// resident_launch.h -- host-side launchers of fit_predict_resident_kernel (resident.cuh), one translation unit per
// feature count (resident_p*.cu) so that the unrolled instantiations compile in parallel.
#pragma once
#include "common.cuh"

namespace b200 {
struct ResidentArgs;
constexpr int kResidentMaxFeatures = 8;
// static shared memory of the kernel (barriers, stage headers, per-warp record / scratch), rounded up: what the
// ring of stages cannot use
constexpr int kResidentStaticBytes = 6 * 1024;
#define B200_DECLARE_RESIDENT(P_) cudaError_t launch_resident_fp_p##P_(bool has_w, const ResidentArgs &a, int grid, cudaStream_t st);
B200_DECLARE_RESIDENT(1)
B200_DECLARE_RESIDENT(2)
B200_DECLARE_RESIDENT(3)
B200_DECLARE_RESIDENT(4)
B200_DECLARE_RESIDENT(5)
B200_DECLARE_RESIDENT(6)
B200_DECLARE_RESIDENT(7)
B200_DECLARE_RESIDENT(8)
#undef B200_DECLARE_RESIDENT

#define B200_DEFINE_RESIDENT(P_)                                                                                     \
	template <bool HASW> static cudaError_t launch_resident_fp_t##P_(const ResidentArgs &a, int grid, cudaStream_t st) { \
		auto kern = fit_predict_resident_kernel<P_, HASW>;                                                           \
		const size_t dyn = (size_t)a.acc.n_stages * a.acc.stage_bytes;                                               \
		cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);            \
		if (e != cudaSuccess) return e;                                                                              \
		kern<<<grid, 32 * (kResidentWarps + 1), dyn, st>>>(a);                                                       \
		return cudaGetLastError();                                                                                   \
	}                                                                                                                \
	cudaError_t launch_resident_fp_p##P_(bool has_w, const ResidentArgs &a, int grid, cudaStream_t st) {             \
		return has_w ? launch_resident_fp_t##P_<true>(a, grid, st) : launch_resident_fp_t##P_<false>(a, grid, st);   \
	}
} // namespace b200
