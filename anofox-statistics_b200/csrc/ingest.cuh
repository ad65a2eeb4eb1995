// ingest.cuh -- unsorted (group_id, y, x[, w]) rows -> group-sorted CSR on the device.
//
// Replaces what *AggUpdate / *AggCombine do in the reference (src/aggregate_functions/ols_aggregate.cpp:120-234):
// there every row is appended to its group's heap vectors through a per-row state pointer (random access across
// groups, millions of small allocations) and thread-local states are concatenated on combine. Here the rows of a
// query stay where the scan put them and a STABLE radix partition by group id produces the layout the fit kernels
// stream: offsets[G+1] plus y / X / w in group order. Stability matters for parity: within a group the rows keep
// their arrival ("state") order, which is the order the reference's constant-column rule refers to
// (|x_i - x_first_valid| < 1e-10, ols.rs:76-87) and the order of the *_fit_predict output lists.
//
// Algorithm: least-significant-digit radix sort of (group id -> row index) pairs, 8 bits per pass,
// ceil(log2 G / 8) passes; each pass = per-tile digit histogram, one exclusive scan over (digit, tile), stable
// scatter (ranks inside a tile from warp match/popc + per-warp digit counts in shared memory); then one gather of
// the rows by the sorted row index. All HBM-bound byte/integer work: coalesced key reads, no tensor cores.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kRadixThreads = 256;
constexpr int kRadixItems = 8;
constexpr int kRadixTile = kRadixThreads * kRadixItems; // keys per CTA
constexpr int kScanBlock = 1024;                       // elements per CTA of the scan kernels

// ---- exclusive scan of a uint32 array (three phases; n up to 2^30 entries) -------------------------------------
__global__ void __launch_bounds__(256) scan_local_kernel(uint32_t *__restrict__ a, int64_t n, uint32_t *__restrict__ sums) {
	__shared__ uint32_t wsum[8];
	const int64_t base = (int64_t)blockIdx.x * kScanBlock + (int64_t)threadIdx.x * 4;
	uint32_t v[4], t = 0;
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		v[i] = base + i < n ? a[base + i] : 0u;
		t += v[i];
	}
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t inc = t;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t u = __shfl_up_sync(0xffffffffu, inc, d);
		if (lane >= d) inc += u;
	}
	if (lane == 31) wsum[warp] = inc;
	__syncthreads();
	uint32_t woff = 0;
	for (int w = 0; w < warp; ++w) woff += wsum[w];
	uint32_t run = woff + inc - t;
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		if (base + i < n) a[base + i] = run;
		run += v[i];
	}
	if (threadIdx.x == 255) sums[blockIdx.x] = woff + inc;
}
// one CTA scans the block sums in place (sequential over chunks of 256; at most ~1 M sums)
__global__ void __launch_bounds__(256) scan_sums_kernel(uint32_t *__restrict__ sums, int64_t m) {
	__shared__ uint32_t wsum[8];
	__shared__ uint32_t carry;
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int64_t b = 0; b < m; b += 256) {
		const int64_t i = b + threadIdx.x;
		const uint32_t v = i < m ? sums[i] : 0u;
		uint32_t inc = v;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const uint32_t u = __shfl_up_sync(0xffffffffu, inc, d);
			if (lane >= d) inc += u;
		}
		if (lane == 31) wsum[warp] = inc;
		__syncthreads();
		uint32_t woff = carry;
		for (int w = 0; w < warp; ++w) woff += wsum[w];
		if (i < m) sums[i] = woff + inc - v;
		__syncthreads();
		if (threadIdx.x == 255) carry = woff + inc;
		__syncthreads();
	}
}
__global__ void __launch_bounds__(256) scan_add_kernel(uint32_t *__restrict__ a, int64_t n, const uint32_t *__restrict__ sums) {
	const int64_t base = (int64_t)blockIdx.x * kScanBlock + (int64_t)threadIdx.x * 4;
	const uint32_t s = sums[blockIdx.x];
#pragma unroll
	for (int i = 0; i < 4; ++i)
		if (base + i < n) a[base + i] += s;
}

// ---- one radix pass ----------------------------------------------------------------------------------------------
// hist[digit * n_tiles + tile] = number of keys of the tile with that digit
__global__ void __launch_bounds__(kRadixThreads) radix_hist_kernel(const uint32_t *__restrict__ keys, int64_t n, int shift,
                                                                  int64_t n_tiles, uint32_t *__restrict__ hist) {
	__shared__ uint32_t h[256];
	h[threadIdx.x] = 0;
	__syncthreads();
	const int64_t base = (int64_t)blockIdx.x * kRadixTile;
#pragma unroll
	for (int i = 0; i < kRadixItems; ++i) {
		const int64_t e = base + (int64_t)i * kRadixThreads + threadIdx.x;
		if (e < n) atomicAdd(&h[(keys[e] >> shift) & 255u], 1u);
	}
	__syncthreads();
	hist[(int64_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}
// stable scatter: element e of the tile goes to scanned_hist[digit][tile] + (number of earlier keys of the tile
// with the same digit)
__global__ void __launch_bounds__(kRadixThreads) radix_scatter_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                                                     int64_t n, int shift, int64_t n_tiles,
                                                                     const uint32_t *__restrict__ hist, uint32_t *__restrict__ keys_out,
                                                                     uint32_t *__restrict__ vals_out, int first_pass) {
	__shared__ uint32_t running[256];     // keys of this digit seen in earlier rounds of the tile
	__shared__ uint32_t wcount[8][256];   // per warp of the current round
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	running[threadIdx.x] = hist[(int64_t)threadIdx.x * n_tiles + blockIdx.x];
	const int64_t base = (int64_t)blockIdx.x * kRadixTile;
	for (int i = 0; i < kRadixItems; ++i) {
#pragma unroll
		for (int w = 0; w < 8; ++w) wcount[w][threadIdx.x] = 0;
		__syncthreads();
		const int64_t e = base + (int64_t)i * kRadixThreads + threadIdx.x;
		const bool in = e < n;
		const uint32_t k = in ? keys[e] : 0xffffffffu;
		const uint32_t d = in ? ((k >> shift) & 255u) : 256u; // 256: a digit no real key has
		const unsigned peers = __match_any_sync(0xffffffffu, d);
		const int rank = __popc(peers & ((1u << lane) - 1u));
		if (in && rank == 0) wcount[warp][d] = (uint32_t)__popc(peers);
		__syncthreads();
		if (in) {
			uint32_t pos = running[d] + (uint32_t)rank;
			for (int w = 0; w < warp; ++w) pos += wcount[w][d];
			keys_out[pos] = k;
			vals_out[pos] = first_pass ? (uint32_t)e : vals[e];
		}
		__syncthreads();
		uint32_t add = 0;
#pragma unroll
		for (int w = 0; w < 8; ++w) add += wcount[w][threadIdx.x];
		running[threadIdx.x] += add;
		__syncthreads();
	}
}

// ---- keys, counts, offsets, gather -------------------------------------------------------------------------------
// group ids (int64, 0 <= id < n_groups) -> 32-bit keys + per-group counts; a row with an id outside the range is
// flagged (bad[0] = 1) and sent to key 0 so that the partition stays well formed
__global__ void ingest_keys_kernel(const int64_t *__restrict__ gid, int64_t n, int64_t n_groups, uint32_t *__restrict__ keys,
                                   uint32_t *__restrict__ counts, int *__restrict__ bad) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	int64_t g = gid[i];
	if (g < 0 || g >= n_groups) {
		*bad = 1;
		g = 0;
	}
	keys[i] = (uint32_t)g;
	atomicAdd(&counts[g], 1u);
}
__global__ void ingest_offsets_kernel(const uint32_t *__restrict__ excl, int64_t n_groups, int64_t n, int64_t *__restrict__ offsets) {
	const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g < n_groups) offsets[g] = (int64_t)excl[g];
	if (g == n_groups) offsets[g] = n;
}
// rows by sorted row index. A CTA takes kGatherRows consecutive OUTPUT rows: their source indices go to shared
// memory once, then thread t walks the CTA's output elements t, t + 256, ... (coalesced stores; the p loads of a
// source row sit side by side in a warp) with 32-bit index arithmetic on the CTA-local element number -- the first
// version spent a 64-bit division per element and was compute-bound (2.9 of the 5.5 ms of a 50 M-row ingest).
constexpr int kGatherRows = 256;
__global__ void __launch_bounds__(256) ingest_gather_kernel(const uint32_t *__restrict__ perm, int64_t n, int p,
                                                            const double *__restrict__ y, const double *__restrict__ x,
                                                            const double *__restrict__ w, double *__restrict__ yo,
                                                            double *__restrict__ xo, double *__restrict__ wo,
                                                            int64_t *__restrict__ row_index) {
	__shared__ uint32_t sperm[kGatherRows];
	const int64_t r0 = (int64_t)blockIdx.x * kGatherRows;
	const int nr = (int)(n - r0 < kGatherRows ? n - r0 : kGatherRows);
	const int t = threadIdx.x;
	uint32_t src = 0;
	if (t < nr) {
		src = perm[r0 + t];
		sperm[t] = src;
		yo[r0 + t] = y[src];
		if (w) wo[r0 + t] = w[src];
		if (row_index) row_index[r0 + t] = (int64_t)src;
	}
	__syncthreads();
	const uint32_t ne = (uint32_t)nr * (uint32_t)p, up = (uint32_t)p;
	double *__restrict__ dst = xo + r0 * p;
	for (uint32_t le = (uint32_t)t; le < ne; le += 256u) {
		const uint32_t row = le / up, col = le - row * up;
		dst[le] = x[(int64_t)sperm[row] * p + col];
	}
}

} // namespace b200
