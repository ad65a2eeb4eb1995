// hc_launch.h -- host-side launcher of hc_small_kernel (hc_small.cuh), compiled in its own translation unit
// (hc_small.cu).
#pragma once
#include "common.cuh"

namespace b200 {
constexpr int64_t kHcSmallMaxRows = 2048; // longer groups go to hc_kernel (hc.cuh) through big_list
struct HcSmallArgs {
	const int64_t *offsets;
	const double *y, *x, *w;
	const uint8_t *train;
	int64_t n_groups;
	int hc_type;       // AnofoxHcType 1 .. 4
	int64_t *big_list; // groups above kHcSmallMaxRows rows (for hc_kernel)
	int *big_count;
};
constexpr int kHcSmallMaxFeatures = 8;
cudaError_t launch_hc_small(int p, const HcSmallArgs &a, const FitParams &P, const AnofoxB200FitOutputs &o, int grid, cudaStream_t st);
} // namespace b200
