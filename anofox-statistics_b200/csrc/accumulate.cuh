// accumulate.cuh -- segmented, group-sorted streaming accumulation of the per-group moment records.
//
// Replaces the reference's raw-row aggregate state (src/aggregate_functions/ols_aggregate.cpp:19-42
// state, :120-186 Update, :189-234 Combine) and the design-matrix build of fit_ols
// (crates/anofox-stats-core/src/models/ols.rs:59-66 finite-row filter, :76-87 constant-column
// scan, :149-152 Mat::from_fn): instead of buffering rows and decomposing them, every row is
// read from HBM exactly once and folded into X'WX, X'Wy, y'Wy, n, sum w and the column sums.
//
// Shape of the kernel (one persistent CTA per SM):
//   * warp 0 = producer. It walks the CTA's windows of the group-sorted row range, packs as many
//     consecutive whole groups as fit into a shared-memory stage and fetches their rows with ONE
//     bulk copy per array (cp.async.bulk -> SASS UBLKCP, the TMA engine; the 16-byte alignment
//     head/tail elements are patched with plain stores). Completion is signalled through an
//     mbarrier transaction count; a ring of stages keeps several copies in flight per SM.
//   * warps 1.. = consumers, one group per warp at a time. A step covers 4 rows: lane l holds the
//     element (row l%4, column 8t + l/4) of each 8-column tile t, which is simultaneously the A
//     fragment (scaled by the weight) and the B fragment of an FP64 tensor-core MMA
//     (mma.m8n8k4.f64 -> SASS DMMA): the reduction over rows happens inside the MMA, so no
//     cross-lane reduction of the Gram matrix is needed. y and the constant-1 column ride in the
//     spare columns of the last tile when p % 8 leaves room, otherwise they are accumulated by
//     a few scalar FMAs per step (SIDE mode).
//   * rows are differenced against the group's first valid row before squaring (exact
//     constant-column statistic, and a shift that removes the cancellation of raw moments).
//   * the fast path assumes finite features and checks the result: a non-finite accumulator
//     re-runs the group with the per-element finite filter of ols.rs:59-66 (CHECKX).
//   * a group too large for a stage is appended to a list and left to the streaming kernel.
#pragma once
#include "common.cuh"

namespace b200 {

struct AccArgs {
	const int64_t *offsets; // [G+1]
	const double *y;        // [N]
	const double *x;        // [N*p]
	const double *w;        // [N] or null
	const uint8_t *train;   // [N] or null
	int64_t n_groups;
	int p;
	const int64_t *tile_g0; // [n_tiles+1] first group starting in each row window
	int64_t n_tiles;
	double *records;        // [G][rec_len(p)]
	int64_t *big_list;      // groups that do not fit a stage
	int *big_count;
	int cap_rows;           // rows per stage
	int stage_bytes;        // bytes per stage (all regions)
	int xcap, ycap, wcap;   // region sizes in bytes (16-byte multiples)
	int n_stages;
};

template <int MAXG> struct StageHeaderT {
	int64_t g_first;
	int count; // groups in the stage; -1 = producer finished
	int next;  // fused kernel: next unclaimed round of the stage (reset by the producer)
	int lead_x, lead_y, lead_w, lead_t; // byte offset of element 0 inside each region
	uint32_t rowstart[MAXG + 1];
};
using StageHeader = StageHeaderT<kMaxStageGroups>;

// Copy `bytes` bytes from global `src` into the shared region `dst` (16-byte aligned) so that byte e lands
// at dst + lead + e, lead = src & 15. ONE bulk copy moves the enclosing 16-byte aligned range
// [src - lead, ceil16(src + bytes)): up to 15 bytes of the neighbouring elements on either side come
// along (never across an allocation granule: device allocations are >= 256-byte aligned and sized), which
// keeps the producer free of dependent patch loads. Returns the bytes the copy will deliver (for the
// mbarrier transaction count).
template <int ELEM>
__device__ __forceinline__ uint32_t stage_fetch(unsigned char *dst, const unsigned char *src, int64_t bytes, int lane,
                                                uint64_t *bar, int *lead_out) {
	const uint32_t lead = (uint32_t)((uintptr_t)src & 15);
	*lead_out = (int)lead;
	if (bytes <= 0) return 0;
	const uint32_t total = (uint32_t)((lead + bytes + 15) & ~(int64_t)15);
	if (lane == 0) bulk_g2s(dst, src - lead, total, bar);
	return total;
}

__device__ __forceinline__ double xor_sum4(double v) { // sum over the 4 lanes that share lane/4
	v += __shfl_xor_sync(0xffffffffu, v, 1);
	v += __shfl_xor_sync(0xffffffffu, v, 2);
	return v;
}
__device__ __forceinline__ double xor_max4(double v) {
	v = fmax(v, __shfl_xor_sync(0xffffffffu, v, 1));
	v = fmax(v, __shfl_xor_sync(0xffffffffu, v, 2));
	return v;
}

// Running state of one warp for one group (or one row shard of a big group).
template <int NT, bool SIDE> struct WarpAcc {
	static constexpr int NTP = NT * (NT + 1) / 2;
	double c0[NTP], c1[NTP]; // DMMA accumulators, tile pairs (ta <= tb) in row-major order
	bool nc[NT];             // some valid row differs from the shift row by >= 1e-10 (ols.rs:76-87)
	double sxy[SIDE ? NT : 1], s1[SIDE ? NT : 1];
	double syy, sy, sw;
	int nv, ns;
	__device__ __forceinline__ void clear() {
#pragma unroll
		for (int i = 0; i < NTP; ++i) c0[i] = c1[i] = 0.0;
#pragma unroll
		for (int t = 0; t < NT; ++t) nc[t] = false;
#pragma unroll
		for (int t = 0; t < (SIDE ? NT : 1); ++t) sxy[t] = s1[t] = 0.0;
		syy = sy = sw = 0.0;
		nv = ns = 0;
	}
};

// Row validity from y / w / train only (the part every lane can evaluate alone).
__device__ __forceinline__ bool row_ok_yw(double yv, double wv, int tr, bool has_w) {
	bool ok = finite_f64(yv) && tr != 0;
	if (has_w) ok = ok && (wv > 0.0) && finite_f64(wv);
	return ok;
}

// Accumulate rows [r_begin, r_end) of a row-major block into `A`, differencing against sh/shy.
// xs/ys/ws/ts point at row 0 of the block; n_clamp = last addressable row.
template <int NT, bool SIDE, bool CHECKX>
__device__ __forceinline__ void acc_rows(WarpAcc<NT, SIDE> &A, const double *__restrict__ xs,
                                         const double *__restrict__ ys, const double *__restrict__ ws,
                                         const uint8_t *__restrict__ ts, int r_begin, int r_end, int p, int lane,
                                         const double (&sh)[NT], double shy, bool nan_unflags = false) {
	const int rl = lane & 3, cl = lane >> 2;
	const bool has_w = ws != nullptr;
	for (int base = r_begin; base < r_end; base += 4) {
		const int r = base + rl;
		const int rc = r < r_end ? r : r_end - 1;
		const double yv = ys[rc];
		const double wv = has_w ? ws[rc] : 1.0;
		const int tr = ts ? (int)ts[rc] : 1;
		double xv[NT];
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			const int c = 8 * t + cl;
			xv[t] = (c < p) ? xs[(int64_t)rc * p + c] : 0.0;
		}
		bool ok = (r < r_end) && row_ok_yw(yv, wv, tr, has_w);
		bool counted = r < r_end && tr != 0;
		if (CHECKX) {
			bool bad = false, nanx = false;
#pragma unroll
			for (int t = 0; t < NT; ++t) {
				bad = bad || !finite_f64(xv[t]);
				nanx = nanx || isnan(xv[t]);
			}
			const unsigned mb = __ballot_sync(0xffffffffu, bad);
			ok = ok && ((mb & (0x11111111u << rl)) == 0u);
			if (nan_unflags) {
				// default training flags derived from y alone (fit_device): a row with a missing feature is not a
				// training row (ols_predict_aggregate.cpp:216-239) -- it does not count towards the state either
				const unsigned mn = __ballot_sync(0xffffffffu, nanx);
				counted = counted && ((mn & (0x11111111u << rl)) == 0u);
			}
		}
		A.ns += counted ? 1 : 0;
		A.nv += ok ? 1 : 0;
		const double weff = ok ? wv : 0.0;
		const double dy = ok ? yv - shy : 0.0;
		double d[NT], a[NT];
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			const int c = 8 * t + cl;
			double dv = ok ? xv[t] - sh[t] : 0.0; // feature column (or zero padding: xv = sh = 0)
			if (!SIDE && t == NT - 1) {
				if (c == p) dv = dy;
				if (c == p + 1) dv = ok ? 1.0 : 0.0;
			}
			d[t] = dv;
			A.nc[t] = A.nc[t] || (fabs(dv) >= 1e-10);
			a[t] = has_w ? weff * dv : dv;
		}
		int pair = 0;
#pragma unroll
		for (int ta = 0; ta < NT; ++ta)
#pragma unroll
			for (int tb = ta; tb < NT; ++tb) {
				dmma_884(A.c0[pair], A.c1[pair], a[ta], d[tb]);
				++pair;
			}
		if (SIDE) {
			const double wdy = weff * dy;
#pragma unroll
			for (int t = 0; t < NT; ++t) {
				A.sxy[t] = fma(d[t], wdy, A.sxy[t]);
				A.s1[t] += a[t];
			}
			A.syy = fma(wdy, dy, A.syy);
			A.sy += wdy;
			A.sw += weff;
		}
	}
}

// First valid row of [0,n) -> index, or -1. Also loads the shift vector.
template <int NT, bool SIDE, bool CHECKX>
__device__ __forceinline__ int find_first_valid(const double *__restrict__ xs, const double *__restrict__ ys,
                                                const double *__restrict__ ws, const uint8_t *__restrict__ ts, int n,
                                                int p, int lane) {
	const int rl = lane & 3, cl = lane >> 2;
	const bool has_w = ws != nullptr;
	for (int base = 0; base < n; base += 4) {
		const int r = base + rl;
		const int rc = r < n ? r : n - 1;
		bool ok = (r < n) && row_ok_yw(ys[rc], has_w ? ws[rc] : 1.0, ts ? (int)ts[rc] : 1, has_w);
		if (CHECKX) {
			bool bad = false;
#pragma unroll
			for (int t = 0; t < NT; ++t) {
				const int c = 8 * t + cl;
				if (c < p) bad = bad || !finite_f64(xs[(int64_t)rc * p + c]);
			}
			const unsigned mb = __ballot_sync(0xffffffffu, bad);
			ok = ok && ((mb & (0x11111111u << rl)) == 0u);
		}
		const unsigned m4 = __ballot_sync(0xffffffffu, ok) & 0xFu;
		if (m4) return base + __ffs(m4) - 1;
	}
	return -1;
}

template <int NT, bool SIDE>
__device__ __forceinline__ void load_shift(const double *__restrict__ xs, const double *__restrict__ ys, int f, int p,
                                           int lane, double (&sh)[NT], double &shy) {
	const int cl = lane >> 2;
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		const int c = 8 * t + cl;
		sh[t] = (c < p) ? xs[(int64_t)f * p + c] : 0.0;
	}
	shy = ys[f];
}

template <int NT, bool SIDE> __device__ __forceinline__ bool acc_poisoned(const WarpAcc<NT, SIDE> &A) {
	bool bad = false;
#pragma unroll
	for (int i = 0; i < WarpAcc<NT, SIDE>::NTP; ++i) bad = bad || !finite_f64(A.c0[i]) || !finite_f64(A.c1[i]);
	if (SIDE) {
#pragma unroll
		for (int t = 0; t < NT; ++t) bad = bad || !finite_f64(A.sxy[t]) || !finite_f64(A.s1[t]);
		bad = bad || !finite_f64(A.syy) || !finite_f64(A.sy);
	}
	return __any_sync(0xffffffffu, bad);
}

// Scatter the warp's accumulators into a record (global or shared memory, element stride 1).
// add = false: plain stores of a complete group; add = true is not used here (partials are
// written whole and reduced by a separate kernel, which keeps the summation order fixed).
template <int NT, bool SIDE>
__device__ __forceinline__ void store_record(WarpAcc<NT, SIDE> &A, double *__restrict__ rec, int p, int lane,
                                             const double (&sh)[NT], double shy, int n_state) {
	const int rl = lane & 3, cl = lane >> 2, k = p + 1;
	double *S1 = rec + rec_off_s1(p), *S2 = rec + rec_off_s2(p);
	int pair = 0;
#pragma unroll
	for (int ta = 0; ta < NT; ++ta)
#pragma unroll
		for (int tb = ta; tb < NT; ++tb) {
			const int i = 8 * ta + cl;
#pragma unroll
			for (int h = 0; h < 2; ++h) {
				const int j = 8 * tb + 2 * rl + h;
				const double v = h ? A.c1[pair] : A.c0[pair];
				if (i <= j) {
					if (SIDE) {
						if (j < p) S2[tri_idx(i, j, k)] = v;
					} else {
						if (j <= p)
							S2[tri_idx(i, j, k)] = v;
						else if (j == p + 1) {
							if (i <= p)
								S1[i] = v;
							else if (i == p + 1)
								rec[1] = v; // sum of weights
						}
					}
				}
			}
			++pair;
		}
	// per-column statistics live replicated over the 4 row lanes
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		const int c = 8 * t + cl;
		const unsigned ncm = __ballot_sync(0xffffffffu, A.nc[t]);
		const double cdm = ((ncm >> (4 * cl)) & 0xFu) ? 1.0 : 0.0;
		if (SIDE) {
			const double vxy = xor_sum4(A.sxy[t]);
			const double v1 = xor_sum4(A.s1[t]);
			if (rl == 0 && c < p) {
				S2[tri_idx(c, p, k)] = vxy;
				S1[c] = v1;
			}
		}
		if (rl == 0 && c < p) {
			rec[rec_off_cdev(p) + c] = cdm;
			rec[rec_off_shift(p) + c] = sh[t];
		}
	}
	const int nv = A.nv + __shfl_xor_sync(0xffffffffu, A.nv, 1);
	const int nvt = nv + __shfl_xor_sync(0xffffffffu, nv, 2);
	const int ns = A.ns + __shfl_xor_sync(0xffffffffu, A.ns, 1);
	const int nst = ns + __shfl_xor_sync(0xffffffffu, ns, 2);
	if (SIDE) {
		const double vyy = xor_sum4(A.syy), vy = xor_sum4(A.sy), vw = xor_sum4(A.sw);
		if (lane == 0) {
			S2[tri_idx(p, p, k)] = vyy;
			S1[p] = vy;
			rec[1] = vw;
		}
	}
	if (lane == 0) {
		rec[0] = (double)nvt;
		rec[2] = (double)(n_state >= 0 ? n_state : nst);
		rec[3] = 0.0;
		rec[rec_off_shift(p) + p] = shy;
	}
}

// A record that says "no valid rows" (n_valid = 0): the solve stage turns it into the right status.
__device__ __forceinline__ void store_empty_record(double *__restrict__ rec, int p, int lane, int n_state) {
	const int L = rec_len(p);
	for (int i = lane; i < L; i += 32) rec[i] = (i == 2) ? (double)n_state : 0.0;
}

// ---------------------------------------------------------------------------------------------
// Fast path of the resident kernel. Every lane owns, per tile, ONE shared-memory address and a
// per-step stride: feature columns walk the row-major X block, the y column walks the y block,
// the constant-1 column and the zero padding point at two constants (stride 0). A 4-row step is
// then: one LDS + one DADD (difference against the shift row) + one DSETP (constant-column flag)
// per tile, and the DMMAs. Nothing else is checked per element: non-finite inputs poison the
// accumulators and the group is redone by the checked path above.
//   HASW: rows with w <= 0 (or NaN) are masked here (they are legitimate input, wls.rs:76-86).
//   HAST: rows whose training flag is 0 are masked (held-out rows of *_fit_predict_agg).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double lds_f64(uint32_t addr) {
	double v;
	asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
	return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
	uint32_t v;
	asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
	return v;
}

struct SmemGroup { // shared-space addresses of one group's rows
	uint32_t x, y, w, t; // byte addresses (w / t unused unless HASW / HAST)
	uint32_t kone, kzero; // addresses of the constants 1.0 and 0.0
};

template <int NT, bool SIDE>
__device__ __forceinline__ void lane_pointers(const SmemGroup &g, int p, int lane, uint32_t (&pa)[NT], uint32_t (&st)[NT]) {
	const int rl = lane & 3, cl = lane >> 2;
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		const int c = 8 * t + cl;
		if (c < p) {
			pa[t] = g.x + (uint32_t)(rl * p + c) * 8u;
			st[t] = 32u * (uint32_t)p;
		} else if (!SIDE && c == p) {
			pa[t] = g.y + (uint32_t)rl * 8u;
			st[t] = 32u;
		} else if (!SIDE && c == p + 1) {
			pa[t] = g.kone;
			st[t] = 0u;
		} else {
			pa[t] = g.kzero;
			st[t] = 0u;
		}
	}
}

// Bank swizzle for rows that are a multiple of 128 bytes (p % 16 == 0): the four rows of a step then start in
// the same bank, and the 8-byte loads of one tile touch only half of the banks four times over. Lanes of odd
// rows therefore load tile t ^ 1 in the slot of tile t (their pointers are swapped once, at set-up) and swap
// the two values back here: every load instruction covers all 32 banks, two wavefronts instead of four.
template <int NT> __device__ __forceinline__ void swizzle_pairs(double (&v)[NT], bool swz) {
#pragma unroll
	for (int t = 0; t + 1 < NT; t += 2) {
		const double a = v[t], b = v[t + 1];
		v[t] = swz ? b : a;
		v[t + 1] = swz ? a : b;
	}
}

// c0 / c1: the DMMA accumulator set this step adds into (A's own, or one of the extra sets the ring kernel
// rotates through to break the DMMA -> DMMA dependency between consecutive steps).
// TRACKNC: keep the exact per-element constant-column flag (one DSETP per tile and step).
template <int NT, bool SIDE, bool HASW, bool HAST, bool TAIL, bool TRACKNC, bool COUNTS = true>
__device__ __forceinline__ void mma_step(WarpAcc<NT, SIDE> &A, double (&c0)[NT * (NT + 1) / 2],
                                         double (&c1)[NT * (NT + 1) / 2], uint32_t (&pa)[NT], const uint32_t (&st)[NT],
                                         uint32_t &py, uint32_t &pw, uint32_t &pt, const double (&sh)[NT], double shy,
                                         bool in_range, bool swz = false) {
	double v[NT];
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		v[t] = lds_f64(pa[t]);
		pa[t] += st[t];
	}
	swizzle_pairs<NT>(v, swz);
	bool ok = true;
	double wv = 1.0, yv = 0.0;
	if (SIDE) {
		yv = lds_f64(py);
		py += 32u;
	}
	if (HASW) {
		wv = lds_f64(pw);
		pw += 32u;
		ok = wv > 0.0;
	}
	if (HAST) {
		const uint32_t tr = lds_u8(pt);
		pt += 4u;
		ok = ok && tr != 0u;
		if (!TAIL || in_range) A.ns += tr != 0u ? 1 : 0;
	}
	if (TAIL) ok = ok && in_range;
	constexpr bool MASK = HASW || HAST || TAIL;
	// COUNTS = false (ring kernel, unweighted): n_valid equals the training-row count and sum w equals n_valid;
	// both are filled in at the end of the group instead of per step
	if (MASK && (COUNTS || HASW || !HAST)) A.nv += ok ? 1 : 0;
	const double weff = MASK ? (ok ? wv : 0.0) : wv;
	double d[NT], a[NT];
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		double dv = v[t] - sh[t];
		// weighted fits: a masked row has weff = 0, so its a-operand (and w dy below) is zero whatever d holds; the
		// B-operand only needs its own mask where nothing else zeroes the product (unweighted masks, tail rows that
		// may read past the data, the constant-column flags)
		if (MASK && (!HASW || TAIL || TRACKNC)) dv = ok ? dv : 0.0;
		d[t] = dv;
		if (TRACKNC) A.nc[t] = A.nc[t] || (fabs(dv) >= 1e-10);
		a[t] = HASW ? weff * dv : dv;
	}
	int pair = 0;
#pragma unroll
	for (int ta = 0; ta < NT; ++ta)
#pragma unroll
		for (int tb = ta; tb < NT; ++tb) {
			dmma_884(c0[pair], c1[pair], a[ta], d[tb]);
			++pair;
		}
	if (SIDE) {
		double dy = yv - shy;
		if (MASK && (!HASW || TAIL)) dy = ok ? dy : 0.0;
		const double wdy = HASW ? weff * dy : dy;
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			A.sxy[t] = fma(d[t], wdy, A.sxy[t]);
			A.s1[t] += a[t];
		}
		A.syy = fma(wdy, dy, A.syy);
		A.sy += wdy;
		if (COUNTS || HASW) A.sw += weff;
	}
}

// Software-pipelined form of mma_step for full (unmasked-tail) steps: the operands of a step are loaded a step
// ahead. The raw values are dead once the differences d = v - shift exist, so the next step's loads reuse their
// registers and sit between the subtractions and the DMMAs of the current step: the shared-memory latency hides
// behind ~NT(NT+1)/2 tensor instructions instead of stalling the warp at the top of every step.
template <int NT, bool SIDE, bool HASW, bool HAST> struct StepIn {
	double v[NT];
	double yv, wv;
	uint32_t tr;
};
// p0: the lane's pointer into tile 0 (tiles 1 .. NT-2 are +64 t bytes, same row stride st0); pl / stl: the
// last tile, whose lanes may read y, the constant 1 or the zero padding instead.
template <int OFF> __device__ __forceinline__ double lds_f64_off(uint32_t addr) {
	double v;
	asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(addr), "n"(OFF));
	return v;
}
template <int NT, int T> struct TileLoads {
	static __device__ __forceinline__ void run(double (&v)[NT], uint32_t p0) {
		v[T] = lds_f64_off<64 * T>(p0);
		TileLoads<NT, T + 1>::run(v, p0);
	}
};
template <int NT> struct TileLoads<NT, NT> {
	static __device__ __forceinline__ void run(double (&)[NT], uint32_t) {}
};
template <int NT, bool SIDE, bool HASW, bool HAST>
__device__ __forceinline__ void step_load(StepIn<NT, SIDE, HASW, HAST> &in, uint32_t &p0, uint32_t &pl, uint32_t st0,
                                          uint32_t stl, uint32_t &py, uint32_t &pw, uint32_t &pt) {
	{
		double head[NT > 1 ? NT - 1 : 1];
		TileLoads<(NT > 1 ? NT - 1 : 1), (NT > 1 ? 0 : 1)>::run(head, p0);
#pragma unroll
		for (int t = 0; t < NT - 1; ++t) in.v[t] = head[t];
		in.v[NT - 1] = lds_f64(pl);
		p0 += st0;
		pl += stl;
	}
	if (SIDE) {
		in.yv = lds_f64(py);
		py += 32u;
	}
	if (HASW) {
		in.wv = lds_f64(pw);
		pw += 32u;
	}
	if (HAST) {
		in.tr = lds_u8(pt);
		pt += 4u;
	}
}
template <int NT, bool SIDE, bool HASW, bool HAST, bool TRACKNC, bool COUNTS>
__device__ __forceinline__ void mma_step_pipe(WarpAcc<NT, SIDE> &A, double (&c0)[NT * (NT + 1) / 2],
                                              double (&c1)[NT * (NT + 1) / 2], StepIn<NT, SIDE, HASW, HAST> &in,
                                              uint32_t &p0, uint32_t &pl, uint32_t st0, uint32_t stl, uint32_t &py,
                                              uint32_t &pw, uint32_t &pt, const double (&sh)[NT], double shy, bool more) {
	bool ok = true;
	double wv = 1.0;
	if (HASW) {
		wv = in.wv;
		ok = wv > 0.0;
	}
	if (HAST) {
		ok = ok && in.tr != 0u;
		A.ns += in.tr != 0u ? 1 : 0;
	}
	constexpr bool MASK = HASW || HAST;
	if (MASK && (COUNTS || HASW || !HAST)) A.nv += ok ? 1 : 0;
	const double weff = MASK ? (ok ? wv : 0.0) : wv;
	double d[NT], a[NT];
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		double dv = in.v[t] - sh[t];
		if (MASK && (!HASW || TRACKNC)) dv = ok ? dv : 0.0; // see mma_step
		d[t] = dv;
		if (TRACKNC) A.nc[t] = A.nc[t] || (fabs(dv) >= 1e-10);
		a[t] = HASW ? weff * dv : dv;
	}
	double dy = 0.0;
	if (SIDE) {
		dy = in.yv - shy;
		if (MASK && !HASW) dy = ok ? dy : 0.0;
	}
	if (more) step_load<NT, SIDE, HASW, HAST>(in, p0, pl, st0, stl, py, pw, pt);
	int pair = 0;
#pragma unroll
	for (int ta = 0; ta < NT; ++ta)
#pragma unroll
		for (int tb = ta; tb < NT; ++tb) {
			dmma_884(c0[pair], c1[pair], a[ta], d[tb]);
			++pair;
		}
	if (SIDE) {
		const double wdy = HASW ? weff * dy : dy;
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			A.sxy[t] = fma(d[t], wdy, A.sxy[t]);
			A.s1[t] += a[t];
		}
		A.syy = fma(wdy, dy, A.syy);
		A.sy += wdy;
		if (COUNTS || HASW) A.sw += weff;
	}
}

template <int NT, bool SIDE, bool HASW, bool HAST, bool TAIL>
__device__ __forceinline__ void fast_step(WarpAcc<NT, SIDE> &A, uint32_t (&pa)[NT], const uint32_t (&st)[NT],
                                          uint32_t &py, uint32_t &pw, uint32_t &pt, const double (&sh)[NT], double shy,
                                          bool in_range) {
	mma_step<NT, SIDE, HASW, HAST, TAIL, true>(A, A.c0, A.c1, pa, st, py, pw, pt, sh, shy, in_range);
}

// Returns false when the result is poisoned (non-finite input somewhere) or no row passed the
// w / train mask with a usable shift; the caller then runs the checked path.
template <int NT, bool SIDE, bool HASW, bool HAST>
__device__ __forceinline__ bool accumulate_group_fast(WarpAcc<NT, SIDE> &A, const SmemGroup &g, int n, int p, int lane,
                                                      double (&sh)[NT], double &shy) {
	const int rl = lane & 3, cl = lane >> 2;
	// first row that passes the mask (row 0 in the plain case)
	int f = 0;
	if (HASW || HAST) {
		f = -1;
		for (int base = 0; base < n && f < 0; base += 32) {
			const int r = base + lane;
			bool ok = r < n;
			if (ok && HASW) ok = lds_f64(g.w + (uint32_t)r * 8u) > 0.0;
			if (ok && HAST) ok = lds_u8(g.t + (uint32_t)r) != 0u;
			const unsigned m = __ballot_sync(0xffffffffu, ok);
			if (m) f = base + __ffs(m) - 1;
		}
		if (f < 0) return false;
	}
	uint32_t pa[NT], st[NT];
	lane_pointers<NT, SIDE>(g, p, lane, pa, st);
#pragma unroll
	for (int t = 0; t < NT; ++t) {
		const int c = 8 * t + cl;
		const bool data = (c < p) || (!SIDE && c == p);
		// row f of this lane's column: undo the lane's own row offset (rl), then add f rows
		sh[t] = data ? lds_f64(pa[t] + (uint32_t)(f - rl) * (st[t] >> 2)) : 0.0;
	}
	shy = lds_f64(g.y + (uint32_t)f * 8u);
	uint32_t py = g.y + (uint32_t)rl * 8u, pw = g.w + (uint32_t)rl * 8u, pt = g.t + (uint32_t)rl;
	A.clear();
	const int nfull = n >> 2;
#pragma unroll 2
	for (int s = 0; s < nfull; ++s) fast_step<NT, SIDE, HASW, HAST, false>(A, pa, st, py, pw, pt, sh, shy, true);
	if (n & 3) {
		const bool in_range = rl < (n & 3);
		if (!in_range) { // keep the loads inside the group's rows
#pragma unroll
			for (int t = 0; t < NT; ++t) pa[t] -= (uint32_t)rl * (st[t] >> 2);
			py -= (uint32_t)rl * 8u;
			pw -= (uint32_t)rl * 8u;
			pt -= (uint32_t)rl;
		}
		fast_step<NT, SIDE, HASW, HAST, true>(A, pa, st, py, pw, pt, sh, shy, in_range);
	}
	// unmasked full steps of the plain case did not count their rows
	if (!(HASW || HAST)) A.nv += nfull;
	return !acc_poisoned<NT, SIDE>(A);
}

// One whole group resident in shared memory -> its record.
template <int NT, bool SIDE, bool HASW, bool HAST>
__device__ __forceinline__ void accumulate_group(const SmemGroup &sg, const double *xs, const double *ys,
                                                 const double *ws, const uint8_t *ts, int n, int p, int lane,
                                                 double *rec) {
	// rows that entered the aggregate state: all of them, or the flagged training rows
	int n_state = n;
	if (HAST) {
		int c = 0;
		for (int r = lane; r < n; r += 32) c += ts[r] ? 1 : 0;
		for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
		n_state = c;
	}
	if (n <= 0) {
		store_empty_record(rec, p, lane, 0);
		return;
	}
	WarpAcc<NT, SIDE> A;
	double sh[NT], shy;
	if (!accumulate_group_fast<NT, SIDE, HASW, HAST>(A, sg, n, p, lane, sh, shy)) {
		// a non-finite value (or no row passing the mask): the checked path of ols.rs:59-66
		const int f = find_first_valid<NT, SIDE, true>(xs, ys, ws, ts, n, p, lane);
		if (f < 0) {
			store_empty_record(rec, p, lane, n_state);
			return;
		}
		load_shift<NT, SIDE>(xs, ys, f, p, lane, sh, shy);
		A.clear();
		acc_rows<NT, SIDE, true>(A, xs, ys, ws, ts, 0, n, p, lane, sh, shy);
	}
	store_record<NT, SIDE>(A, rec, p, lane, sh, shy, n_state);
}

// ---------------------------------------------------------------------------------------------
// Producer warp (shared by the resident DMMA kernel and the fused small-group kernel): walks this
// CTA's row windows, packs consecutive whole groups into a stage (at most max_groups of them, at most
// cap_rows rows) and fetches their rows with one bulk copy per array. Groups larger than a stage go
// to big_list. Ends with a stage whose count is -1.
// ---------------------------------------------------------------------------------------------
template <bool HASW, bool HAST, class Hdr>
__device__ __forceinline__ void producer_loop(const AccArgs &a, unsigned char *smem, uint64_t *full_bar,
                                              uint64_t *empty_bar, Hdr *hdr, int NS, int lane, int max_groups,
                                              int n_end = 1, int nb = 0) {
	// nb > 0 (consumers that take stages by ticket and may be several fills ahead of the producer): the full
	// barrier of fill number k is full_bar[k % nb], nb > stages + consumers, so that a waiter is never a whole
	// barrier phase ahead (a parity wait cannot tell phase n from phase n - 2). The data stage stays k % NS.
	const int p = a.p;
	int s = 0;
	uint32_t ph = 0;
	int fb = 0; // index of the full barrier of the current fill
	for (int64_t t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
		int64_t g = a.tile_g0[t];
		const int64_t gend = a.tile_g0[t + 1];
		if (g >= gend) continue;
		// offsets of the next stage's candidate groups are fetched one stage ahead (pf_*), so that the
		// global-load latency overlaps the copy issue and the wait for a free stage
		int64_t pf_row0 = a.offsets[g];
		int64_t pf_oe = (g + lane < gend) ? a.offsets[g + lane + 1] : 0;
		while (g < gend) {
			mbar_wait(&empty_bar[s], ph ^ 1u);
			const int64_t row0 = pf_row0;
			int cnt = 0;
			int64_t row_end = row0;
			bool closed = false, first = true;
			while (!closed && g + cnt < gend && cnt < max_groups) {
				const int64_t gi = g + cnt + lane;
				const bool in = gi < gend && (cnt + lane) < max_groups;
				int64_t oe = row0;
				if (in) oe = first ? pf_oe : a.offsets[gi + 1];
				first = false;
				const bool fits = in && (oe - row0) <= (int64_t)a.cap_rows;
				const unsigned m = __ballot_sync(0xffffffffu, fits);
				const int nfit = (m == 0xffffffffu) ? 32 : (__ffs(~m) - 1);
				if (lane < nfit) hdr[s].rowstart[cnt + lane + 1] = (uint32_t)(oe - row0);
				if (nfit > 0) row_end = __shfl_sync(0xffffffffu, oe, nfit - 1);
				cnt += nfit;
				if (nfit < 32) closed = true;
			}
			if (cnt == 0) {
				// group g alone exceeds a stage: hand it to the streaming kernel
				if (lane == 0) {
					const int slot = atomicAdd(a.big_count, 1);
					a.big_list[slot] = g;
				}
				g += 1;
				if (g < gend) {
					pf_row0 = a.offsets[g];
					pf_oe = (g + lane < gend) ? a.offsets[g + lane + 1] : 0;
				}
				continue;
			}
			const int64_t g_next = g + cnt;
			pf_row0 = row_end; // == offsets[g_next]
			pf_oe = (g_next + lane < gend) ? a.offsets[g_next + lane + 1] : 0;
			const int64_t rows = row_end - row0;
			unsigned char *base = smem + (size_t)s * a.stage_bytes;
			int lx = 0, ly = 0, lw = 0, lt = 0;
			uint32_t tx = 0;
			tx += stage_fetch<8>(base, (const unsigned char *)(a.x + row0 * p), rows * p * 8, lane, &full_bar[fb], &lx);
			tx += stage_fetch<8>(base + a.xcap, (const unsigned char *)(a.y + row0), rows * 8, lane, &full_bar[fb], &ly);
			if (HASW)
				tx += stage_fetch<8>(base + a.xcap + a.ycap, (const unsigned char *)(a.w + row0), rows * 8, lane,
				                     &full_bar[fb], &lw);
			if (HAST)
				tx += stage_fetch<1>(base + a.xcap + a.ycap + a.wcap, a.train + row0, rows, lane, &full_bar[fb], &lt);
			if (lane == 0) {
				hdr[s].rowstart[0] = 0;
				hdr[s].g_first = g;
				hdr[s].count = cnt;
				hdr[s].next = 0;
				hdr[s].lead_x = lx;
				hdr[s].lead_y = ly;
				hdr[s].lead_w = lw;
				hdr[s].lead_t = lt;
			}
			__syncwarp();
			if (lane == 0) {
				if (tx)
					mbar_arrive_expect_tx(&full_bar[fb], tx);
				else
					mbar_arrive(&full_bar[fb]);
			}
			g = g_next;
			if (++s == NS) {
				s = 0;
				ph ^= 1u;
			}
			fb = nb > 0 ? (fb + 1 == nb ? 0 : fb + 1) : s;
		}
	}
	// end marker(s): one per consumer that takes stages by ticket (each leaves on its own marker), otherwise one
	// that every consumer sees
	for (int e = 0; e < n_end; ++e) {
		mbar_wait(&empty_bar[s], ph ^ 1u);
		if (lane == 0) {
			hdr[s].count = -1;
			mbar_arrive(&full_bar[fb]);
		}
		if (++s == NS) {
			s = 0;
			ph ^= 1u;
		}
		fb = nb > 0 ? (fb + 1 == nb ? 0 : fb + 1) : s;
	}
}

// ---------------------------------------------------------------------------------------------
// The same producer with the group offsets of a whole tile prefetched into shared memory (cp.async, one tile
// ahead) and the tile bounds read two tiles ahead. In producer_loop the offsets of the next stage are a global
// load issued ~0.1 us before they are needed: its latency (~0.7 us) sits on the per-stage critical path of the
// one warp that feeds the SM, next to two dependent loads per tile -- about 1.2 us per 39 KB stage, i.e. the
// 4.7 TB/s at which the small-group kernels used to saturate whatever the consumers did.
// obuf: shared memory, 2 * kOffsetTile int64 entries.
// ---------------------------------------------------------------------------------------------
constexpr int kOffsetTile = 1024; // offsets held per tile; groups of a tile beyond that read global memory
__device__ __forceinline__ void cp_async8(void *dst_smem, const void *src) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool HASW, bool HAST, class Hdr>
__device__ __forceinline__ void producer_loop_pf(const AccArgs &a, unsigned char *smem, uint64_t *full_bar,
                                                 uint64_t *empty_bar, Hdr *hdr, int NS, int lane, int max_groups,
                                                 int64_t *obuf, int n_end = 1, int nb = 0, uint32_t hint = 20000u) {
	const int p = a.p;
	int s = 0;
	uint32_t ph = 0;
	int fb = 0; // index of the full barrier of the current fill (see producer_loop)
	const int64_t stride = gridDim.x;
	// offsets[g0 .. g1] (at most kOffsetTile of them) -> obuf[b]; always commits a group
	auto prefetch = [&](int b, int64_t g0, int64_t g1) {
		int64_t n = g1 - g0 + 1;
		if (n > kOffsetTile) n = kOffsetTile;
		for (int64_t i = lane; i < n; i += 32) cp_async8(obuf + b * kOffsetTile + i, a.offsets + g0 + i);
		cp_async_commit();
	};
	int64_t t = blockIdx.x;
	int64_t g = 0, gend = 0, gn = 0, gendn = 0;
	if (t < a.n_tiles) {
		g = a.tile_g0[t];
		gend = a.tile_g0[t + 1];
		prefetch(0, g, gend);
	}
	if (t + stride < a.n_tiles) {
		gn = a.tile_g0[t + stride];
		gendn = a.tile_g0[t + stride + 1];
	}
	int b = 0;
	for (; t < a.n_tiles; t += stride) {
		// next tile: its offsets start streaming now, the bounds of the tile after it go to registers
		const bool have_next = t + stride < a.n_tiles;
		if (have_next)
			prefetch(b ^ 1, gn, gendn);
		else
			cp_async_commit();
		int64_t gnn = 0, gendnn = 0;
		if (t + 2 * stride < a.n_tiles) {
			gnn = a.tile_g0[t + 2 * stride];
			gendnn = a.tile_g0[t + 2 * stride + 1];
		}
		cp_async_wait<1>(); // this tile's offsets have landed (everything but the group just committed)
		__syncwarp();
		const int64_t gt0 = g;
		const int64_t *ob = obuf + b * kOffsetTile;
		auto off_at = [&](int64_t gi) -> int64_t { // offsets[gi], gt0 <= gi <= gend
			const int64_t i = gi - gt0;
			return i < kOffsetTile ? ob[i] : a.offsets[gi];
		};
		while (g < gend) {
			mbar_wait_hint(&empty_bar[s], ph ^ 1u, hint);
			const int64_t row0 = off_at(g);
			int cnt = 0;
			int64_t row_end = row0;
			bool closed = false;
			while (!closed && g + cnt < gend && cnt < max_groups) {
				const int64_t gi = g + cnt + lane;
				const bool in = gi < gend && (cnt + lane) < max_groups;
				int64_t oe = row0;
				if (in) oe = off_at(gi + 1);
				const bool fits = in && (oe - row0) <= (int64_t)a.cap_rows;
				const unsigned m = __ballot_sync(0xffffffffu, fits);
				const int nfit = (m == 0xffffffffu) ? 32 : (__ffs(~m) - 1);
				if (lane < nfit) hdr[s].rowstart[cnt + lane + 1] = (uint32_t)(oe - row0);
				if (nfit > 0) row_end = __shfl_sync(0xffffffffu, oe, nfit - 1);
				cnt += nfit;
				if (nfit < 32) closed = true;
			}
			if (cnt == 0) {
				// group g alone exceeds a stage: hand it to the streaming kernel
				if (lane == 0) {
					const int slot = atomicAdd(a.big_count, 1);
					a.big_list[slot] = g;
				}
				g += 1;
				continue;
			}
			const int64_t rows = row_end - row0;
			unsigned char *base = smem + (size_t)s * a.stage_bytes;
			int lx = 0, ly = 0, lw = 0, lt = 0;
			uint32_t tx = 0;
			tx += stage_fetch<8>(base, (const unsigned char *)(a.x + row0 * p), rows * p * 8, lane, &full_bar[fb], &lx);
			tx += stage_fetch<8>(base + a.xcap, (const unsigned char *)(a.y + row0), rows * 8, lane, &full_bar[fb], &ly);
			if (HASW)
				tx += stage_fetch<8>(base + a.xcap + a.ycap, (const unsigned char *)(a.w + row0), rows * 8, lane,
				                     &full_bar[fb], &lw);
			if (HAST)
				tx += stage_fetch<1>(base + a.xcap + a.ycap + a.wcap, a.train + row0, rows, lane, &full_bar[fb], &lt);
			if (lane == 0) {
				hdr[s].rowstart[0] = 0;
				hdr[s].g_first = g;
				hdr[s].count = cnt;
				hdr[s].next = 0;
				hdr[s].lead_x = lx;
				hdr[s].lead_y = ly;
				hdr[s].lead_w = lw;
				hdr[s].lead_t = lt;
			}
			__syncwarp();
			if (lane == 0) {
				if (tx)
					mbar_arrive_expect_tx(&full_bar[fb], tx);
				else
					mbar_arrive(&full_bar[fb]);
			}
			g += cnt;
			if (++s == NS) {
				s = 0;
				ph ^= 1u;
			}
			fb = nb > 0 ? (fb + 1 == nb ? 0 : fb + 1) : s;
		}
		__syncwarp(); // every lane is done with obuf[b] before the tile after next overwrites it
		g = gn;
		gend = gendn;
		gn = gnn;
		gendn = gendnn;
		b ^= 1;
	}
	cp_async_wait<0>();
	for (int e = 0; e < n_end; ++e) {
		mbar_wait_hint(&empty_bar[s], ph ^ 1u, hint);
		if (lane == 0) {
			hdr[s].count = -1;
			mbar_arrive(&full_bar[fb]);
		}
		if (++s == NS) {
			s = 0;
			ph ^= 1u;
		}
		fb = nb > 0 ? (fb + 1 == nb ? 0 : fb + 1) : s;
	}
}

// ---------------------------------------------------------------------------------------------
// Resident kernel: groups that fit a stage.
// ---------------------------------------------------------------------------------------------
template <int NT, bool SIDE, int NCW, bool HASW, bool HAST>
__global__ void __launch_bounds__(32 * (NCW + 1), 1) acc_resident_kernel(const AccArgs a) {
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ uint64_t full_bar[8], empty_bar[8];
	__shared__ StageHeader hdr[8];
	__shared__ double kconst[2];
	const int NS = a.n_stages;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NS; ++s) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], NCW);
		}
		mbar_fence_init();
		kconst[0] = 1.0;
		kconst[1] = 0.0;
	}
	__syncthreads();
	const int p = a.p;
	const int L = rec_len(p);
	if (warp == 0) {
		producer_loop<HASW, HAST>(a, smem, full_bar, empty_bar, hdr, NS, lane, kMaxStageGroups);
	} else {
		// ------------------------------ consumers -----------------------------
		const int cw = warp - 1;
		for (int it = 0;; ++it) {
			const int s = it % NS;
			mbar_wait(&full_bar[s], (uint32_t)((it / NS) & 1));
			const int cnt = hdr[s].count;
			if (cnt < 0) break;
			unsigned char *base = smem + (size_t)s * a.stage_bytes;
			const double *sx = (const double *)(base + hdr[s].lead_x);
			const double *sy = (const double *)(base + a.xcap + hdr[s].lead_y);
			const double *swp = HASW ? (const double *)(base + a.xcap + a.ycap + hdr[s].lead_w) : nullptr;
			const uint8_t *st = HAST ? (const uint8_t *)(base + a.xcap + a.ycap + a.wcap + hdr[s].lead_t) : nullptr;
			const int64_t g0 = hdr[s].g_first;
			SmemGroup sg;
			sg.kone = smem_u32(&kconst[0]);
			sg.kzero = smem_u32(&kconst[1]);
			const uint32_t ax = smem_u32(sx), ay = smem_u32(sy), aw = HASW ? smem_u32(swp) : 0u, at = HAST ? smem_u32(st) : 0u;
			// rotate the warp -> group assignment from stage to stage so that a stage holding fewer
			// groups than there are consumer warps does not always idle the same warps
			for (int i = (cw + it * 5) % NCW; i < cnt; i += NCW) {
				const int rb = (int)hdr[s].rowstart[i];
				const int n = (int)hdr[s].rowstart[i + 1] - rb;
				sg.x = ax + (uint32_t)rb * (uint32_t)p * 8u;
				sg.y = ay + (uint32_t)rb * 8u;
				sg.w = aw + (uint32_t)rb * 8u;
				sg.t = at + (uint32_t)rb;
				accumulate_group<NT, SIDE, HASW, HAST>(sg, sx + (int64_t)rb * p, sy + rb, HASW ? swp + rb : nullptr,
				                                       HAST ? st + rb : nullptr, n, p, lane, a.records + (g0 + i) * L);
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(&empty_bar[s]);
		}
	}
}

// ---------------------------------------------------------------------------------------------
// Window bounds: tile_g0[t] = first group whose first row is >= t * window_rows.
// ---------------------------------------------------------------------------------------------
static __global__ void tile_bounds_kernel(const int64_t *__restrict__ offsets, int64_t n_groups, int64_t window_rows,
                                   int64_t n_tiles, int64_t *__restrict__ tile_g0) {
	const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (t > n_tiles) return;
	if (t == n_tiles) {
		tile_g0[t] = n_groups;
		return;
	}
	const int64_t target = t * window_rows;
	int64_t lo = 0, hi = n_groups; // first g in [0, G] with offsets[g] >= target
	while (lo < hi) {
		const int64_t mid = (lo + hi) >> 1;
		if (offsets[mid] >= target)
			hi = mid;
		else
			lo = mid + 1;
	}
	tile_g0[t] = lo;
}

// ---------------------------------------------------------------------------------------------
// Streaming path for groups larger than a stage (and for the row-sharded single group).
//   step 1  first_valid_kernel : one warp per big group finds its first valid row -> shift
//   step 2  acc_stream_kernel  : every CTA takes an equal slice of the group's rows; each warp
//                                accumulates 4-row steps straight from global memory and writes
//                                one partial record
//   step 3  reduce_partials_kernel : fixed-order sum of the partial records
// ---------------------------------------------------------------------------------------------
struct StreamArgs {
	const int64_t *offsets; // group mode: rows of big_list[b] are offsets[g] .. offsets[g+1]
	const int64_t *big_list;
	const int *big_count;
	int64_t fixed_rows;     // shard mode (offsets == nullptr): the block is rows [0, fixed_rows)
	const double *y, *x, *w;
	const uint8_t *train;
	int p;
	double *shift;          // [nbig][p+1]  (shard mode: provided by the caller)
	int *first_ok;          // [nbig] 1 when a valid row exists
	double *partials;       // [nbig][n_part][L]
	int n_part;             // gridDim.x * warps per CTA of acc_stream_kernel
	double *records;        // group mode: [G][L]; shard mode: [1][L]
	int64_t rec_gs, rec_es; // record element (g, i) lives at records[g * rec_gs + i * rec_es]
	// Row-shard mode behind the ring kernel's slices: the checked streaming pass only runs when *run_if is
	// non-zero (a slice met a non-finite value or an undecided constant column); the reduction then reads
	// `partials` / `n_part`, otherwise the ring's partial records `ring_partials` / `ring_n_part`.
	const int *run_if;
	const double *ring_partials;
	int ring_n_part;
	int flags_from_y; // the train flags say "y present" only: rows with a NaN feature do not count as state rows
	// Group mode, decided on the device (no host round trip): big groups of at most huge_rows rows are taken one
	// per CTA by acc_big_kernel; larger ones are appended to huge_list (as indices into big_list) by
	// first_valid_kernel and go through acc_stream_kernel + reduce_partials_kernel in rounds of b_cap entries,
	// round r covering huge_list[b_begin, b_begin + b_cap).
	int64_t *huge_list;
	int *huge_count;
	int64_t huge_rows;
	int b_begin, b_cap;
};

__device__ __forceinline__ bool row_valid_full(const StreamArgs &a, int64_t r) {
	bool ok = row_ok_yw(a.y[r], a.w ? a.w[r] : 1.0, a.train ? (int)a.train[r] : 1, a.w != nullptr);
	for (int c = 0; ok && c < a.p; ++c) ok = finite_f64(a.x[r * a.p + c]);
	return ok;
}

static __global__ void first_valid_kernel(const StreamArgs a) {
	const int nbig = *a.big_count;
	const int lane = threadIdx.x;
	for (int b = blockIdx.x; b < nbig; b += gridDim.x) {
		const int64_t g = a.big_list[b];
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		int64_t f = -1;
		for (int64_t base = r0; base < r1 && f < 0; base += 32) {
			const int64_t r = base + lane;
			const bool ok = r < r1 && row_valid_full(a, r);
			const unsigned m = __ballot_sync(0xffffffffu, ok);
			if (m) f = base + __ffs(m) - 1;
		}
		const int k = a.p + 1;
		for (int c = lane; c < a.p; c += 32) a.shift[(int64_t)b * k + c] = f >= 0 ? a.x[f * a.p + c] : 0.0;
		if (lane == 0) {
			a.shift[(int64_t)b * k + a.p] = f >= 0 ? a.y[f] : 0.0;
			a.first_ok[b] = f >= 0 ? 1 : 0;
			if (r1 - r0 > a.huge_rows) a.huge_list[atomicAdd(a.huge_count, 1)] = b;
		}
	}
}

template <int NT, bool SIDE, int NW> __global__ void __launch_bounds__(32 * NW) acc_stream_kernel(const StreamArgs a) {
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, cl = lane >> 2;
	const int p = a.p, k = p + 1, L = rec_len(p);
	int nbig = 1;
	if (a.offsets) {
		nbig = *a.huge_count - a.b_begin;
		nbig = nbig < a.b_cap ? nbig : a.b_cap;
	}
	if (a.run_if && *a.run_if == 0) return;
	for (int b = 0; b < nbig; ++b) {
		int64_t r0 = 0, r1 = a.fixed_rows;
		int bb = 0; // index of the group in big_list / shift / first_ok
		if (a.offsets) {
			bb = (int)a.huge_list[a.b_begin + b];
			const int64_t g = a.big_list[bb];
			r0 = a.offsets[g];
			r1 = a.offsets[g + 1];
		}
		double *part = a.partials + ((int64_t)b * a.n_part + (int64_t)blockIdx.x * NW + warp) * L;
		// a group without any valid row still runs: every row is masked, the pass counts the state rows
		// equal contiguous slices per warp, multiples of 4 rows
		const int64_t nrows = r1 - r0;
		const int64_t nw_total = (int64_t)gridDim.x * NW;
		int64_t per = (nrows + nw_total - 1) / nw_total;
		per = (per + 3) & ~(int64_t)3;
		const int64_t wid = (int64_t)blockIdx.x * NW + warp;
		int64_t wb = r0 + wid * per, we = wb + per;
		if (wb > r1) wb = r1;
		if (we > r1) we = r1;
		double sh[NT], shy;
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			const int c = 8 * t + cl;
			sh[t] = (c < p) ? a.shift[(int64_t)bb * k + c] : 0.0;
		}
		shy = a.shift[(int64_t)bb * k + p];
		WarpAcc<NT, SIDE> A;
		A.clear();
		// chunks keep the int row index of acc_rows in range
		for (int64_t cb = wb; cb < we; cb += (1 << 20)) {
			const int64_t ce = cb + (1 << 20) < we ? cb + (1 << 20) : we;
			acc_rows<NT, SIDE, true>(A, a.x + cb * p, a.y + cb, a.w ? a.w + cb : nullptr, a.train ? a.train + cb : nullptr,
			                         0, (int)(ce - cb), p, lane, sh, shy, a.flags_from_y != 0);
		}
		store_record<NT, SIDE>(A, part, p, lane, sh, shy, -1);
	}
}

// Sum the partial records of each big group in a fixed order: n, sw, n_state, S1, S2 add; cdev
// takes the max; the shift is common to all partials.
// field i of the sum of n_part partial records
// (no __restrict__ on part: acc_big_kernel reads partials its own CTA has just written)
__device__ __forceinline__ double reduce_field(const double *part, int n_part, int L, int p, int i, const double *shift) {
	double v = 0.0;
	if (i >= rec_off_shift(p) && i < rec_off_cdev(p)) {
		v = shift[i - rec_off_shift(p)];
	} else if (i >= rec_off_cdev(p) && i < rec_off_cdev(p) + p) {
		for (int q = 0; q < n_part; ++q) v = fmax(v, part[(int64_t)q * L + i]);
	} else if ((i >= rec_off_cdev(p) && i < rec_off_s1(p)) || i == 3) {
		v = 0.0; // pad
	} else {
		for (int q = 0; q < n_part; ++q) v += part[(int64_t)q * L + i];
	}
	return v;
}
static __global__ void reduce_partials_kernel(const StreamArgs a) {
	const int b = blockIdx.y;
	int nbig = 1;
	if (a.offsets) {
		nbig = *a.huge_count - a.b_begin;
		nbig = nbig < a.b_cap ? nbig : a.b_cap;
	}
	if (b >= nbig) return;
	const int p = a.p, L = rec_len(p);
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= L) return;
	const int bb = a.offsets ? (int)a.huge_list[a.b_begin + b] : 0;
	const int64_t g = a.offsets ? a.big_list[bb] : 0;
	const bool ring = a.run_if && *a.run_if == 0;
	const int n_part = ring ? a.ring_n_part : a.n_part;
	const double *part = ring ? a.ring_partials : a.partials + (int64_t)b * a.n_part * L;
	a.records[g * a.rec_gs + (int64_t)i * a.rec_es] = reduce_field(part, n_part, L, p, i, a.shift + (int64_t)bb * (p + 1));
}

// Big groups of at most huge_rows rows: one CTA per group (grid-stride over the device-side list), the CTA's
// warps take equal row slices with the checked per-element filter, their partial records meet in a fixed order.
template <int NT, bool SIDE, int NW> __global__ void __launch_bounds__(32 * NW) acc_big_kernel(const StreamArgs a) {
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, cl = lane >> 2;
	const int p = a.p, k = p + 1, L = rec_len(p);
	const int nbig = *a.big_count;
	double *parts = a.partials + (int64_t)blockIdx.x * NW * L;
	for (int b = blockIdx.x; b < nbig; b += gridDim.x) {
		const int64_t g = a.big_list[b];
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		if (r1 - r0 > a.huge_rows) continue; // sliced over the whole grid instead (acc_stream_kernel)
		int64_t per = (r1 - r0 + NW - 1) / NW;
		per = (per + 3) & ~(int64_t)3;
		int64_t wb = r0 + warp * per, we = wb + per;
		if (wb > r1) wb = r1;
		if (we > r1) we = r1;
		double sh[NT], shy;
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			const int c = 8 * t + cl;
			sh[t] = (c < p) ? a.shift[(int64_t)b * k + c] : 0.0;
		}
		shy = a.shift[(int64_t)b * k + p];
		WarpAcc<NT, SIDE> A;
		A.clear();
		for (int64_t cb = wb; cb < we; cb += (1 << 20)) {
			const int64_t ce = cb + (1 << 20) < we ? cb + (1 << 20) : we;
			acc_rows<NT, SIDE, true>(A, a.x + cb * p, a.y + cb, a.w ? a.w + cb : nullptr, a.train ? a.train + cb : nullptr,
			                         0, (int)(ce - cb), p, lane, sh, shy, a.flags_from_y != 0);
		}
		store_record<NT, SIDE>(A, parts + (int64_t)warp * L, p, lane, sh, shy, -1);
		__syncthreads(); // partials visible to the block (global stores of this CTA, read by this CTA)
		for (int i = threadIdx.x; i < L; i += 32 * NW)
			a.records[g * a.rec_gs + (int64_t)i * a.rec_es] = reduce_field(parts, NW, L, p, i, a.shift + (int64_t)b * k);
		__syncthreads();
	}
}

} // namespace b200
