#include "window.cuh"
#include "window_launch.h"

namespace b200 {

template <int P> static cudaError_t launch_window_p(bool has_w, const WindowArgs &a, const FitParams &Pm, const AnofoxB200FitOutputs &o,
                                                    int grid, cudaStream_t st) {
	if (has_w)
		window_fit_predict_kernel<P, true><<<grid, 128, 0, st>>>(a, Pm, o);
	else
		window_fit_predict_kernel<P, false><<<grid, 128, 0, st>>>(a, Pm, o);
	return cudaGetLastError();
}

cudaError_t launch_window(int p, bool has_w, const WindowArgs &a, const FitParams &P, const AnofoxB200FitOutputs &o, int grid,
                          cudaStream_t st) {
	switch (p) {
	case 1: return launch_window_p<1>(has_w, a, P, o, grid, st);
	case 2: return launch_window_p<2>(has_w, a, P, o, grid, st);
	case 3: return launch_window_p<3>(has_w, a, P, o, grid, st);
	case 4: return launch_window_p<4>(has_w, a, P, o, grid, st);
	case 5: return launch_window_p<5>(has_w, a, P, o, grid, st);
	case 6: return launch_window_p<6>(has_w, a, P, o, grid, st);
	case 7: return launch_window_p<7>(has_w, a, P, o, grid, st);
	case 8: return launch_window_p<8>(has_w, a, P, o, grid, st);
	}
	return cudaErrorInvalidValue;
}

} // namespace b200
