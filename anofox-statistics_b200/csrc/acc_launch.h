// acc_launch.h -- host-side launchers of the accumulation kernels, one translation unit per tile count
// (acc_nt*.cu) so that the template instantiations compile in parallel.
#pragma once
#include "accumulate.cuh"
#include "ring.cuh"
#include "slice.cuh"

namespace b200 {

constexpr int kStreamWarps = 8;

#define B200_DECLARE_NT(NT_)                                                                                         \
	cudaError_t launch_resident_nt##NT_(bool side, bool has_w, bool has_t, const AccArgs &a, int grid,               \
	                                    cudaStream_t st);                                                            \
	cudaError_t launch_stream_nt##NT_(bool side, const StreamArgs &a, int grid, cudaStream_t st);                    \
	cudaError_t launch_big_nt##NT_(bool side, const StreamArgs &a, int grid, cudaStream_t st);                       \
	cudaError_t launch_ring_nt##NT_(bool side, bool has_w, bool has_t, const RingArgs &a, int grid, size_t dyn,      \
	                                cudaStream_t st);                                                                \
	cudaError_t launch_redo_nt##NT_(bool side, const RingArgs &a, int grid, cudaStream_t st);                       \
	cudaError_t launch_slice_nt##NT_(bool side, bool has_w, const SliceArgs &a, int grid, size_t dyn, cudaStream_t st);
B200_DECLARE_NT(1)
B200_DECLARE_NT(2)
B200_DECLARE_NT(3)
B200_DECLARE_NT(4)
B200_DECLARE_NT(5)
B200_DECLARE_NT(6)
B200_DECLARE_NT(7)
B200_DECLARE_NT(8)
B200_DECLARE_NT(9)
#undef B200_DECLARE_NT

// consumer warps per CTA of the resident kernel: register budget of the DMMA accumulators
constexpr int resident_warps(int nt) { return nt <= 2 ? 16 : (nt <= 4 ? 8 : 4); }

template <int NT, bool SIDE, int NCW, bool HASW, bool HAST>
cudaError_t launch_resident_t(const AccArgs &a, int grid, cudaStream_t st) {
	auto kern = acc_resident_kernel<NT, SIDE, NCW, HASW, HAST>;
	const size_t dyn = (size_t)a.n_stages * a.stage_bytes;
	cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
	if (e != cudaSuccess) return e;
	kern<<<grid, 32 * (NCW + 1), dyn, st>>>(a);
	return cudaGetLastError();
}

template <int NT, bool SIDE, bool HASW, bool HAST>
cudaError_t launch_ring_t(const RingArgs &a, int grid, size_t dyn, cudaStream_t st) {
	constexpr int NCW = ring_warps(NT);
	auto kern = acc_ring_kernel<NT, SIDE, NCW, HASW, HAST, ring_sets(NT)>;
	cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
	if (e != cudaSuccess) return e;
	kern<<<grid, 32 * (NCW + 1), dyn, st>>>(a);
	return cudaGetLastError();
}

// wide shapes only (five or more tiles): narrower shards stay on the ring kernel's slice mode
constexpr int kSliceMinTiles = 5;
template <int NT, bool SIDE, bool HASW>
cudaError_t launch_slice_t(const SliceArgs &a, int grid, size_t dyn, cudaStream_t st) {
	if constexpr (NT >= kSliceMinTiles) {
		auto kern = acc_slice_kernel<NT, SIDE, HASW>;
		cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
		if (e != cudaSuccess) return e;
		kern<<<grid, 32 * (kSliceWarps + 1), dyn, st>>>(a);
		return cudaGetLastError();
	} else {
		return cudaErrorInvalidValue;
	}
}

#define B200_DEFINE_NT(NT_)                                                                                          \
	cudaError_t launch_resident_nt##NT_(bool side, bool has_w, bool has_t, const AccArgs &a, int grid,               \
	                                    cudaStream_t st) {                                                           \
		constexpr int W = resident_warps(NT_);                                                                       \
		const int sel = (side ? 4 : 0) | (has_w ? 2 : 0) | (has_t ? 1 : 0);                                          \
		switch (sel) {                                                                                               \
		case 0: return launch_resident_t<NT_, false, W, false, false>(a, grid, st);                                  \
		case 1: return launch_resident_t<NT_, false, W, false, true>(a, grid, st);                                   \
		case 2: return launch_resident_t<NT_, false, W, true, false>(a, grid, st);                                   \
		case 3: return launch_resident_t<NT_, false, W, true, true>(a, grid, st);                                    \
		case 4: return launch_resident_t<NT_, true, W, false, false>(a, grid, st);                                   \
		case 5: return launch_resident_t<NT_, true, W, false, true>(a, grid, st);                                    \
		case 6: return launch_resident_t<NT_, true, W, true, false>(a, grid, st);                                    \
		default: return launch_resident_t<NT_, true, W, true, true>(a, grid, st);                                    \
		}                                                                                                            \
	}                                                                                                                \
	cudaError_t launch_ring_nt##NT_(bool side, bool has_w, bool has_t, const RingArgs &a, int grid, size_t dyn,      \
	                                cudaStream_t st) {                                                               \
		const int sel = (side ? 4 : 0) | (has_w ? 2 : 0) | (has_t ? 1 : 0);                                          \
		switch (sel) {                                                                                               \
		case 0: return launch_ring_t<NT_, false, false, false>(a, grid, dyn, st);                                    \
		case 1: return launch_ring_t<NT_, false, false, true>(a, grid, dyn, st);                                     \
		case 2: return launch_ring_t<NT_, false, true, false>(a, grid, dyn, st);                                     \
		case 3: return launch_ring_t<NT_, false, true, true>(a, grid, dyn, st);                                      \
		case 4: return launch_ring_t<NT_, true, false, false>(a, grid, dyn, st);                                     \
		case 5: return launch_ring_t<NT_, true, false, true>(a, grid, dyn, st);                                      \
		case 6: return launch_ring_t<NT_, true, true, false>(a, grid, dyn, st);                                      \
		default: return launch_ring_t<NT_, true, true, true>(a, grid, dyn, st);                                      \
		}                                                                                                            \
	}                                                                                                                \
	cudaError_t launch_redo_nt##NT_(bool side, const RingArgs &a, int grid, cudaStream_t st) {                       \
		if (side)                                                                                                    \
			redo_kernel<NT_, true><<<grid, 256, 0, st>>>(a);                                                         \
		else                                                                                                         \
			redo_kernel<NT_, false><<<grid, 256, 0, st>>>(a);                                                        \
		return cudaGetLastError();                                                                                   \
	}                                                                                                                \
	cudaError_t launch_slice_nt##NT_(bool side, bool has_w, const SliceArgs &a, int grid, size_t dyn,                \
	                                 cudaStream_t st) {                                                              \
		if (side) return has_w ? launch_slice_t<NT_, true, true>(a, grid, dyn, st)                                   \
		                       : launch_slice_t<NT_, true, false>(a, grid, dyn, st);                                 \
		return has_w ? launch_slice_t<NT_, false, true>(a, grid, dyn, st)                                            \
		             : launch_slice_t<NT_, false, false>(a, grid, dyn, st);                                          \
	}                                                                                                                \
	cudaError_t launch_big_nt##NT_(bool side, const StreamArgs &a, int grid, cudaStream_t st) {                      \
		if (side)                                                                                                    \
			acc_big_kernel<NT_, true, kStreamWarps><<<grid, 32 * kStreamWarps, 0, st>>>(a);                          \
		else                                                                                                         \
			acc_big_kernel<NT_, false, kStreamWarps><<<grid, 32 * kStreamWarps, 0, st>>>(a);                         \
		return cudaGetLastError();                                                                                   \
	}                                                                                                                \
	cudaError_t launch_stream_nt##NT_(bool side, const StreamArgs &a, int grid, cudaStream_t st) {                   \
		if (side)                                                                                                    \
			acc_stream_kernel<NT_, true, kStreamWarps><<<grid, 32 * kStreamWarps, 0, st>>>(a);                       \
		else                                                                                                         \
			acc_stream_kernel<NT_, false, kStreamWarps><<<grid, 32 * kStreamWarps, 0, st>>>(a);                      \
		return cudaGetLastError();                                                                                   \
	}

} // namespace b200
