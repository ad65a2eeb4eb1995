// window_launch.h -- host-side launcher of window_fit_predict_kernel (window.cuh), compiled in its own
// translation unit (window.cu).
#pragma once
#include "common.cuh"

namespace b200 {
struct WindowArgs;
constexpr int kWindowMaxFeatures = 8;
cudaError_t launch_window(int p, bool has_w, const WindowArgs &a, const FitParams &P, const AnofoxB200FitOutputs &o, int grid,
                          cudaStream_t st);
} // namespace b200
