// slice.cuh -- moment accumulation of ONE row shard with a wide Gram (p >= 31: five or more 8-column tiles),
// the kernel behind anofox_b200_moments_device for BASELINE config 4 (single group, 1 B rows x 64 features,
// row-sharded). It replaces, for those shapes, what the reference does by buffering every row of the group
// in the aggregate state (src/aggregate_functions/ols_aggregate.cpp:120-234) and handing the whole matrix to
// the regressor (crates/anofox-stats-core/src/models/ols.rs:59-87, 155-161).
//
// Why a kernel of its own: at p = 64 a 4-row step is 36 FP64 tensor instructions (DMMA.8x8x4) on 72 accumulator
// doubles per lane -- 144 registers. One warp per scheduler is all that fits, and an in-order warp that also
// loads, subtracts the shift row and keeps the side sums leaves the DMMA pipe idle half of the time (ncu:
// 42 % DMMA pipe utilisation in the ring kernel's slice mode). Here the tile pairs of a step are split between
// TWO warps (roles 0 / 1, alternating pairs): 18 DMMAs and 36 accumulator doubles each, <= 168 registers, so
// two warps share every scheduler and one warp's loads and bookkeeping hide behind the other's tensor work.
//
// Shape: one persistent CTA per slice (slice = equal part of the shard, bounds multiples of 16 rows), 8
// consumer warps + 1 producer warp. Rows flow through a ring of self-contained windows of SR rows: window k
// = slice rows [k SR, (k+1) SR), fetched with one bulk copy per array (cp.async.bulk -> UBLKCP); full / empty
// mbarriers per slot. Consumer warp w has role w & 1 and takes the 4-row steps (w >> 1) + 4 j of every window.
// Nothing is checked per element: non-finite values poison the sums and an undecidable constant column raises
// redo_count, upon which the checked streaming pass (acc_stream_kernel) recomputes the shard.
#pragma once
#include "accumulate.cuh"

namespace b200 {

constexpr int kSliceWarps = 8;   // consumer warps: 4 step lanes x 2 roles
constexpr int kSliceMaxWin = 8;  // ring slots at most
constexpr int kSliceRegs = 168;  // 9 warps: the fullest scheduler holds 3 -> 16384 / 96 = 170 registers
constexpr int kSliceRows = 64;   // rows of a window (p <= 64: 4 windows of 64 rows are 135 KB at most)

struct SliceArgs {
	const int64_t *bounds; // [gridDim.x + 1] row bounds of the slices
	const double *y, *x, *w;
	int p;
	const double *shift; // [p + 1] common shift row
	double *partials;    // [gridDim.x * kSliceWarps][rec_len(p)]
	int *redo_count;
	int win_rows, n_win;                 // SR (multiple of 16), ring slots
	uint32_t x_off, y_off, w_off;        // byte offsets of the regions in dynamic shared memory
	uint32_t x_slot, y_slot;             // bytes per slot (window bytes + 16 for the alignment lead)
};

// One 4-row step of role ROLE (-1: all tile pairs), in two halves so that the next step's loads can be issued
// between them: slice_prep turns the raw values into differences against the shift row (HASW masks rows with
// w <= 0, TAIL masks the lanes beyond the last row), slice_mma feeds the tensor pipe and the side sums.
template <int NT> struct SliceVals {
	double d[NT], a[NT];
	double dy, wdy, weff;
};
template <int NT, bool SIDE, bool HASW, int ROLE, bool TAIL>
__device__ __forceinline__ void slice_prep(WarpAcc<NT, SIDE> &A, SliceVals<NT> &s, const double (&v)[NT], double yv,
                                           double wv, const double (&sh)[NT], double shy, bool in_range) {
	bool ok = true;
	if (HASW) ok = wv > 0.0;
	if (TAIL) ok = ok && in_range;
	constexpr bool MASK = HASW || TAIL;
	s.weff = MASK ? (ok ? wv : 0.0) : wv;
	A.nv += MASK ? (ok ? 1 : 0) : 1;
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		double dv = v[t] - sh[t];
		if (MASK) dv = ok ? dv : 0.0;
		s.d[t] = dv;
		if (HASW && ROLE <= 0) A.nc[t] = A.nc[t] || (fabs(dv) >= 1e-10);
		s.a[t] = HASW ? s.weff * dv : dv;
	}
	s.dy = 0.0;
	s.wdy = 0.0;
	if (SIDE) {
		double dy = yv - shy;
		if (MASK) dy = ok ? dy : 0.0;
		s.dy = dy;
		s.wdy = HASW ? s.weff * dy : dy;
	}
}
template <int NT, bool SIDE, bool HASW, int ROLE>
__device__ __forceinline__ void slice_mma(WarpAcc<NT, SIDE> &A, const SliceVals<NT> &s) {
	int pair = 0;
#pragma unroll
	for (int ta = 0; ta < NT; ++ta)
#pragma unroll
		for (int tb = ta; tb < NT; ++tb) {
			if (ROLE < 0 || (pair & 1) == ROLE) dmma_884(A.c0[pair], A.c1[pair], HASW ? s.a[ta] : s.d[ta], s.d[tb]);
			++pair;
		}
	if (SIDE) {
		if (ROLE <= 0) {
#pragma unroll
			for (int t = 0; t < NT; ++t) A.sxy[t] = fma(s.d[t], s.wdy, A.sxy[t]);
			A.syy = fma(s.wdy, s.dy, A.syy);
			A.sy += s.wdy;
			if (HASW) A.sw += s.weff;
		}
		if (ROLE != 0) {
#pragma unroll
			for (int t = 0; t < NT; ++t) A.s1[t] += HASW ? s.a[t] : s.d[t];
		}
	}
}
// One bulk copy of the stream bytes [b0, b1) of an array whose 16-byte aligned stream starts at src0.
__device__ __forceinline__ uint32_t slice_fetch(unsigned char *dst, const unsigned char *src0, int64_t b0, int64_t b1,
                                                uint64_t *bar) {
	if (b1 <= b0) return 0u;
	bulk_g2s(dst, src0 + b0, (uint32_t)(b1 - b0), bar);
	return (uint32_t)(b1 - b0);
}

struct SliceLoads { // per-lane load plan of a step: tiles 0 .. NT-2 from p0 (+64 t), last tile from pl
	uint32_t p0, pl, py, pw;
};

template <int NT, bool SIDE, bool HASW>
__device__ __forceinline__ void slice_load(double (&v)[NT], double &yv, double &wv, const SliceLoads &L) {
	double head[NT > 1 ? NT - 1 : 1];
	TileLoads<(NT > 1 ? NT - 1 : 1), (NT > 1 ? 0 : 1)>::run(head, L.p0);
#pragma unroll
	for (int t = 0; t < NT - 1; ++t) v[t] = head[t];
	v[NT - 1] = lds_f64(L.pl);
	if (SIDE) yv = lds_f64(L.py);
	if (HASW) wv = lds_f64(L.pw);
}

// The whole life of one consumer warp with a compile-time role: every accumulator the role never touches stays
// a constant zero, so the register allocation covers half of the Gram only.
template <int NT, bool SIDE, bool HASW, int ROLE>
__device__ __forceinline__ void slice_consume(const SliceArgs &a, unsigned char *smem, uint64_t *full_bar,
                                              uint64_t *empty_bar, const double *kconst, int cw, int lane, int64_t n,
                                              uint32_t lx, uint32_t ly, uint32_t lw, double *rec) {
	const int p = a.p, SR = a.win_rows, NS = a.n_win;
	const int rq = cw >> 1, rl = lane & 3, cl = lane >> 2;
	const int nwin = (int)((n + SR - 1) / SR);
	double sh[NT], shy;
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		const int c = 8 * t + cl;
		sh[t] = (c < p) ? a.shift[c] : ((!SIDE && c == p) ? a.shift[p] : 0.0);
	}
	shy = a.shift[p];
	// where the lane's element of the last tile comes from: 0 = x, 1 = y, 2 = constant
	const int cL = 8 * (NT - 1) + cl;
	const int kindL = cL < p ? 0 : ((!SIDE && cL == p) ? 1 : 2);
	const uint32_t constL = smem_u32(kconst) + ((!SIDE && cL == p + 1) ? 0u : 8u); // kconst = {1.0, 0.0}
	const uint32_t step_x = 32u * (uint32_t)p; // bytes from one 4-row step to the next
	// lane offsets inside a slot
	const uint32_t ox0 = lx + (uint32_t)(rl * p + cl) * 8u;
	const uint32_t oxL = lx + (uint32_t)(rl * p + cL) * 8u;
	const uint32_t oy = ly + (uint32_t)rl * 8u, ow = lw + (uint32_t)rl * 8u;
	const uint32_t sx = smem_u32(smem + a.x_off), sy = smem_u32(smem + a.y_off), sw_ = smem_u32(smem + a.w_off);

	WarpAcc<NT, SIDE> A;
	A.clear();
	int slot = 0;
	uint32_t ph = 0;
	for (int k = 0; k < nwin; ++k) {
		mbar_wait(&full_bar[slot], ph);
		const int64_t left = n - (int64_t)k * SR;
		const int rows_in = left < (int64_t)SR ? (int)left : SR;
		const uint32_t bx = sx + (uint32_t)slot * a.x_slot, by = sy + (uint32_t)slot * a.y_slot,
		               bw = sw_ + (uint32_t)slot * a.y_slot;
		auto plan = [&](int step) {
			SliceLoads L;
			L.p0 = bx + ox0 + (uint32_t)step * step_x;
			L.pl = kindL == 0 ? bx + oxL + (uint32_t)step * step_x : (kindL == 1 ? by + oy + (uint32_t)step * 32u : constL);
			L.py = by + oy + (uint32_t)step * 32u;
			L.pw = bw + ow + (uint32_t)step * 32u;
			return L;
		};
		const int nfull = rows_in >> 2;
		double v[NT], yv = 0.0, wv = 1.0;
		if (rows_in == kSliceRows) {
			// full window: this warp's kSliceRows / 16 steps, straight-line. Per step: differences, then the next
			// step's loads into the freed registers, then the DMMAs (which hide the load latency).
			SliceLoads L = plan(rq);
			slice_load<NT, SIDE, HASW>(v, yv, wv, L);
#pragma unroll
			for (int j = 0; j < kSliceRows / 16; ++j) {
				SliceVals<NT> sv;
				slice_prep<NT, SIDE, HASW, ROLE, false>(A, sv, v, yv, wv, sh, shy, true);
				if (j + 1 < kSliceRows / 16) {
					L.p0 += 4u * step_x;
					if (kindL == 0) L.pl += 4u * step_x;
					if (kindL == 1) L.pl += 128u;
					L.py += 128u;
					L.pw += 128u;
					slice_load<NT, SIDE, HASW>(v, yv, wv, L);
				}
				slice_mma<NT, SIDE, HASW, ROLE>(A, sv);
			}
		} else if (rq < nfull) {
			int step = rq;
			slice_load<NT, SIDE, HASW>(v, yv, wv, plan(step));
			for (;;) {
				const int next = step + 4;
				SliceVals<NT> sv;
				slice_prep<NT, SIDE, HASW, ROLE, false>(A, sv, v, yv, wv, sh, shy, true);
				const bool more = next < nfull;
				if (more) slice_load<NT, SIDE, HASW>(v, yv, wv, plan(next));
				slice_mma<NT, SIDE, HASW, ROLE>(A, sv);
				if (!more) break;
				step = next;
			}
		}
		if ((rows_in & 3) && (nfull & 3) == rq) { // the slice's last, partial step
			const bool in_range = rl < (rows_in & 3);
			SliceLoads L = plan(nfull);
			if (!in_range) { // keep the loads inside the slice's rows
				L.p0 -= (uint32_t)rl * (step_x >> 2);
				if (kindL == 0) L.pl -= (uint32_t)rl * (step_x >> 2);
				if (kindL == 1) L.pl -= (uint32_t)rl * 8u;
				L.py -= (uint32_t)rl * 8u;
				L.pw -= (uint32_t)rl * 8u;
			}
			slice_load<NT, SIDE, HASW>(v, yv, wv, L);
			SliceVals<NT> sv;
			slice_prep<NT, SIDE, HASW, ROLE, true>(A, sv, v, yv, wv, sh, shy, in_range);
			slice_mma<NT, SIDE, HASW, ROLE>(A, sv);
		}
		__syncwarp();
		if (lane == 0) mbar_arrive(&empty_bar[slot]);
		if (++slot == NS) {
			slot = 0;
			ph ^= 1u;
		}
	}
	if (SIDE && !HASW) A.sw = (double)A.nv;
	// constant-column flags from the diagonal of the Gram (unweighted: max d^2 <= sum d^2 <= n max d^2), poison check
	int nvt = A.nv + __shfl_xor_sync(0xffffffffu, A.nv, 1);
	nvt += __shfl_xor_sync(0xffffffffu, nvt, 2);
	bool bad = acc_poisoned<NT, SIDE>(A);
	if (!HASW) {
		bool band = false;
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			const int dp = t * NT - (t * (t - 1)) / 2; // pair (t, t)
			if (ROLE < 0 || (dp & 1) == ROLE) {
				const double dv = (cl & 1) ? A.c1[dp] : A.c0[dp];
				const bool mine = rl == (cl >> 1); // this lane holds C[cl][cl] of the diagonal tile
				const bool sure = dv > 1.01e-20 * (double)nvt;
				A.nc[t] = mine && sure;
				band = band || (mine && !sure && !(dv < 0.99e-20) && 8 * t + cl < p);
			}
		}
		bad = bad || __any_sync(0xffffffffu, band);
	}
	if (bad && lane == 0) atomicAdd(a.redo_count, 1);
	if (ROLE > 0) { // the counts belong to role 0's record
		A.nv = 0;
		A.sw = 0.0;
	}
	store_record<NT, SIDE>(A, rec, p, lane, sh, shy, cw == 0 ? (int)n : 0);
}

template <int NT, bool SIDE, bool HASW>
__global__ void __maxnreg__(kSliceRegs) acc_slice_kernel(const SliceArgs a) {
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ uint64_t full_bar[kSliceMaxWin], empty_bar[kSliceMaxWin];
	__shared__ double kconst[2];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int p = a.p, SR = a.win_rows, NS = a.n_win;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NS; ++s) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], kSliceWarps);
		}
		mbar_fence_init();
		kconst[0] = 1.0;
		kconst[1] = 0.0;
	}
	__syncthreads();
	const int64_t R0 = a.bounds[blockIdx.x], R1 = a.bounds[blockIdx.x + 1];
	const int64_t n = R1 - R0;
	const int L = rec_len(p);
	double *rec0 = a.partials + (int64_t)blockIdx.x * kSliceWarps * L;
	if (n <= 0) {
		if (warp < kSliceWarps) store_empty_record(rec0 + (int64_t)warp * L, p, lane, 0);
		return;
	}
	const int64_t esx = (int64_t)p * 8;
	const unsigned char *gx = (const unsigned char *)a.x + R0 * esx;
	const unsigned char *gy = (const unsigned char *)a.y + R0 * 8;
	const unsigned char *gw = HASW ? (const unsigned char *)a.w + R0 * 8 : nullptr;
	const uint32_t lx = (uint32_t)((uintptr_t)gx & 15), ly = (uint32_t)((uintptr_t)gy & 15);
	const uint32_t lw = HASW ? (uint32_t)((uintptr_t)gw & 15) : 0u;
	if (warp == kSliceWarps) {
		// ------------------------------ producer ------------------------------
		if (lane != 0) return;
		const int nwin = (int)((n + SR - 1) / SR);
		auto total = [&](uint32_t lead, int64_t es) { return ((int64_t)lead + n * es + 15) & ~(int64_t)15; };
		const int64_t totx = total(lx, esx), toty = total(ly, 8), totw = HASW ? total(lw, 8) : 0;
		const int64_t WBx = (int64_t)SR * esx, WBy = (int64_t)SR * 8;
		auto lo = [](int64_t v, int64_t cap) { return v < cap ? v : cap; };
		for (int k = 0; k < nwin; ++k) {
			const int s = k % NS;
			if (k >= NS) mbar_wait(&empty_bar[s], (uint32_t)(((k / NS) - 1) & 1));
			uint64_t *bar = &full_bar[s];
			uint32_t tx = 0;
			// the window plus 16 bytes: with a lead of 8 the last row ends 8 bytes into the next window's range
			tx += slice_fetch(smem + a.x_off + (size_t)s * a.x_slot, gx - lx, k * WBx, lo((k + 1) * WBx + 16, totx), bar);
			tx += slice_fetch(smem + a.y_off + (size_t)s * a.y_slot, gy - ly, k * WBy, lo((k + 1) * WBy + 16, toty), bar);
			if (HASW) tx += slice_fetch(smem + a.w_off + (size_t)s * a.y_slot, gw - lw, k * WBy, lo((k + 1) * WBy + 16, totw), bar);
			mbar_arrive_expect_tx(bar, tx);
		}
		return; // every window is waited for by the consumers, so no copy is in flight when the CTA retires
	}
	// ------------------------------ consumers ------------------------------
	double *rec = rec0 + (int64_t)warp * L;
	if (warp & 1)
		slice_consume<NT, SIDE, HASW, 1>(a, smem, full_bar, empty_bar, kconst, warp, lane, n, lx, ly, lw, rec);
	else
		slice_consume<NT, SIDE, HASW, 0>(a, smem, full_bar, empty_bar, kconst, warp, lane, n, lx, ly, lw, rec);
}

} // namespace b200
