#include "hc_small.cuh"

namespace b200 {

cudaError_t launch_hc_small(int p, const HcSmallArgs &a, const FitParams &P, const AnofoxB200FitOutputs &o, int grid, cudaStream_t st) {
	switch (p) {
	case 1: hc_small_kernel<1><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 2: hc_small_kernel<2><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 3: hc_small_kernel<3><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 4: hc_small_kernel<4><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 5: hc_small_kernel<5><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 6: hc_small_kernel<6><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 7: hc_small_kernel<7><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	case 8: hc_small_kernel<8><<<grid, kHcSmallThreads, 0, st>>>(a, P, o); break;
	default: return cudaErrorInvalidValue;
	}
	return cudaGetLastError();
}

} // namespace b200
