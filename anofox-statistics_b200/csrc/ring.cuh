// ring.cuh -- streaming ring accumulation: the moment records of groups of ANY size, with the rows of a
// CTA's group range flowing once through a shared-memory ring of fixed-size row windows.
//
// Replaces the same reference code as accumulate.cuh (src/aggregate_functions/ols_aggregate.cpp:120-234
// Update/Combine, crates/anofox-stats-core/src/models/ols.rs:59-87 row filter + constant-column scan),
// for the shapes where a group is not small against an SM's shared memory (p > 8, or hundreds of rows per
// group): there the whole-group stages of acc_resident_kernel leave only two or three warps busy per SM.
//
// Shape of the kernel (one persistent CTA per SM, a contiguous range of groups per CTA):
//   * warp 0, lane 0 = producer. Window k of the CTA's row range is rows [k SR, (k+1) SR), whatever the
//     group boundaries are; it is fetched with one bulk copy per array (cp.async.bulk -> UBLKCP) into slot
//     k mod NS of the ring. The first rows of every NS-th window are mirrored behind the ring end (the
//     "apron"), so that a 16-row quad never has to wrap. No offsets are read by the producer.
//   * consumer warps form teams of TW warps (TW = 1, 2, 4, ... chosen from the mean group size). A team owns
//     a group; warp i of the team takes the quads i, i + TW, ... of the group (a quad = NSET 4-row DMMA
//     steps, aligned to the group's first row, so that the summation order does not depend on where the group
//     lies in the batch). Consecutive steps of a quad add into different accumulator sets, which breaks the
//     ~63-cycle DMMA -> DMMA dependency. At the end of the group the team's partial sums meet in shared
//     memory in a fixed order and the team's first warp stores the record.
//   * every consumer warp passes every window in order (wait "full", later arrive on "empty"), holding only
//     the windows its next quad needs: rows stream, nothing has to be resident as a whole.
//   * non-finite values poison the sums, and a column whose sum of squares cannot decide the constant-column
//     rule lands in a narrow band: both are detected at the end of the group, which is then appended to the
//     redo list and recomputed from global memory with the full per-element checks (redo_kernel).
//   * slice mode (shift_common != null): the units are equal row slices of ONE group with a given shift row
//     (the row-sharded single group and groups larger than big_rows); every warp writes its own partial record
//     and reduce_partials_kernel adds them in a fixed order.
#pragma once
#include "accumulate.cuh"

namespace b200 {

// tile counts whose CTAs hold only four consumer warps take the software-pipelined step (accumulate.cuh)
template <int NT> constexpr bool kRingPipelined = NT >= 5;
constexpr int kRingApron = 16;  // rows mirrored behind the ring end (>= rows of a quad - 1)
constexpr int kRingMaxWin = 24; // windows in the ring at most

struct RingArgs {
	const int64_t *offsets; // [n_units + 1] row bounds of the units (groups, or slices of one group)
	const double *y, *x, *w;
	const uint8_t *train;
	int p;
	const int64_t *tile_g0; // [gridDim.x + 1] first unit of every CTA
	double *records;        // [n_units][rec_len(p)]   (slice mode: [n_units * NCW][rec_len(p)])
	const double *shift_common; // slice mode: the shift row [p + 1] shared by all units
	int64_t *big_list;      // units with more than big_rows rows are appended here and skipped
	int *big_count;
	int64_t big_rows;
	int64_t *redo_list;     // units that need the checked path
	int *redo_count;
	int win_rows, n_win;    // SR (multiple of 16), NS
	int team;               // TW: warps per unit, a power of two dividing the consumer warp count
	uint32_t x_off, y_off, w_off, t_off, slot_off; // byte offsets of the regions in dynamic shared memory
	int flags_from_y; // train = "y present" only: redo_kernel drops rows with a NaN feature from the state count
};

// Number of deposited values per lane when a team combines through shared memory.
template <int NT, bool SIDE> __host__ __device__ constexpr int ring_deposit_values() {
	return NT * (NT + 1) + (SIDE ? 2 * NT + 3 : 0) + 3;
}
__host__ __device__ constexpr int ring_sets(int nt) { return nt == 1 ? 4 : (nt <= 4 ? 2 : 1); }
__host__ __device__ constexpr int ring_warps(int nt) { return nt == 1 ? 16 : (nt == 2 ? 12 : (nt <= 4 ? 8 : 4)); }
// registers per thread that still let one CTA of (ncw + 1) warps launch: the register file is split over the
// four SM sub-partitions (16384 each), and the fullest one holds ceil((ncw + 1) / 4) warps
__host__ __device__ constexpr int ring_maxreg(int ncw) {
	return 16384 / ((ncw + 1 + 3) / 4) / 32 / 8 * 8 > 255 ? 255 : 16384 / ((ncw + 1 + 3) / 4) / 32 / 8 * 8;
}

template <int NT, bool SIDE, class F> __device__ __forceinline__ void ring_for_each(WarpAcc<NT, SIDE> &A, F f) {
#pragma unroll
	for (int i = 0; i < WarpAcc<NT, SIDE>::NTP; ++i) {
		f(A.c0[i]);
		f(A.c1[i]);
	}
	if (SIDE) {
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			f(A.sxy[t]);
			f(A.s1[t]);
		}
		f(A.syy);
		f(A.sy);
		f(A.sw);
	}
}

// One bulk copy of the stream bytes [b0, b1) of an array whose stream starts at `src0` (16-byte aligned)
// into `dst`. Returns the bytes it will deliver.
__device__ __forceinline__ uint32_t ring_fetch(unsigned char *dst, const unsigned char *src0, int64_t b0, int64_t b1,
                                               uint64_t *bar) {
	if (b1 <= b0) return 0u;
	bulk_g2s(dst, src0 + b0, (uint32_t)(b1 - b0), bar);
	return (uint32_t)(b1 - b0);
}

template <int NT, bool SIDE, int NCW, bool HASW, bool HAST, int NSET>
__global__ void __maxnreg__(ring_maxreg(NCW)) acc_ring_kernel(const RingArgs a) {
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ uint64_t full_bar[kRingMaxWin];
	__shared__ uint64_t dep_bar[NCW], rel_bar[NCW];
	__shared__ int issued_s; // windows the producer has issued so far
	__shared__ __align__(16) long long pos_s[NCW + (NCW & 1)]; // per consumer warp: it will not read rows below this again
	__shared__ double kconst[2];
	constexpr int NTP = NT * (NT + 1) / 2;
	constexpr int QR = 4 * NSET; // rows of a quad
	constexpr int NV = ring_deposit_values<NT, SIDE>();
	constexpr bool MASK = HASW || HAST;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int p = a.p, SR = a.win_rows, NS = a.n_win, TW = a.team;
	const int ring_rows = SR * NS;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NS; ++s) {
			mbar_init(&full_bar[s], 1);
		}
		for (int t = 0; t < NCW + (NCW & 1); ++t) pos_s[t] = t < NCW ? 0 : ((long long)1 << 62);
		for (int t = 0; t < NCW; ++t) {
			mbar_init(&dep_bar[t], TW > 1 ? TW - 1 : 1);
			mbar_init(&rel_bar[t], 1);
		}
		mbar_fence_init();
		kconst[0] = 1.0;
		kconst[1] = 0.0;
		issued_s = 0;
	}
	__syncthreads();
	const int64_t ga = a.tile_g0[blockIdx.x], gb = a.tile_g0[blockIdx.x + 1];
	if (ga >= gb) return;
	const int64_t R0 = a.offsets[ga], R1 = a.offsets[gb];
	const int64_t trows = R1 - R0;
	// + 15: the lead of the one-byte flag array can push its last bytes into one more window
	const int nwin = (int)((trows + 15) / SR + 1);
	// stream geometry of the four arrays: element (row) size, 16-byte aligned stream start, lead, total bytes
	const int64_t esx = (int64_t)p * 8;
	const unsigned char *gx = (const unsigned char *)a.x + R0 * esx;
	const unsigned char *gy = (const unsigned char *)a.y + R0 * 8;
	const unsigned char *gw = HASW ? (const unsigned char *)a.w + R0 * 8 : nullptr;
	const unsigned char *gt = HAST ? (const unsigned char *)a.train + R0 : nullptr;
	const uint32_t lx = (uint32_t)((uintptr_t)gx & 15), ly = (uint32_t)((uintptr_t)gy & 15);
	const uint32_t lw = HASW ? (uint32_t)((uintptr_t)gw & 15) : 0u, lt = HAST ? (uint32_t)((uintptr_t)gt & 15) : 0u;

	if (warp == 0) {
		// ------------------------------ producer ------------------------------
		if (lane != 0) return;
		auto total = [&](uint32_t lead, int64_t es) { return trows > 0 ? ((int64_t)lead + trows * es + 15) & ~(int64_t)15 : (int64_t)0; };
		const int64_t totx = total(lx, esx), toty = total(ly, 8), totw = HASW ? total(lw, 8) : 0, tott = HAST ? total(lt, 1) : 0;
		const int64_t WBx = (int64_t)SR * esx, WBy = (int64_t)SR * 8, WBt = SR;
		const int64_t ABx = kRingApron * esx + 16, ABy = kRingApron * 8 + 16, ABt = kRingApron + 16;
		auto lo = [](int64_t v, int64_t cap) { return v < cap ? v : cap; };
		int64_t minpos = 0;
		for (int k = 0; k < nwin; ++k) {
			const int s = k % NS;
			if (k >= NS) {
				// the slot's previous copy has landed (also when nobody waited for it: skipped big groups) ...
				mbar_wait(&full_bar[s], (uint32_t)(((k / NS) - 1) & 1));
				// ... and every consumer warp has moved past the window it held
				const int64_t need = (int64_t)(k - NS + 1) * SR;
				while (minpos < need) {
					long long m = (long long)1 << 62;
#pragma unroll
					for (int t = 0; t < NCW + (NCW & 1); t += 2) {
						const longlong2 v = ld_volatile_s64x2(&pos_s[t]);
						m = v.x < m ? v.x : m;
						m = v.y < m ? v.y : m;
					}
					minpos = m;
					if (minpos < need) __nanosleep(32);
				}
				fence_acq_rel_cta();
			}
			uint64_t *bar = &full_bar[s];
			uint32_t tx = 0;
			const int64_t kb = k;
			tx += ring_fetch(smem + a.x_off + (size_t)s * WBx, gx - lx, lo(kb * WBx, totx), lo((kb + 1) * WBx, totx), bar);
			tx += ring_fetch(smem + a.y_off + (size_t)s * WBy, gy - ly, lo(kb * WBy, toty), lo((kb + 1) * WBy, toty), bar);
			if (HASW) tx += ring_fetch(smem + a.w_off + (size_t)s * WBy, gw - lw, lo(kb * WBy, totw), lo((kb + 1) * WBy, totw), bar);
			if (HAST) tx += ring_fetch(smem + a.t_off + (size_t)s * WBt, gt - lt, lo(kb * WBt, tott), lo((kb + 1) * WBt, tott), bar);
			if (s == 0 && k > 0) { // the apron: the first rows of this window once more, behind the ring end
				tx += ring_fetch(smem + a.x_off + (size_t)NS * WBx, gx - lx, lo(kb * WBx, totx), lo(kb * WBx + ABx, totx), bar);
				tx += ring_fetch(smem + a.y_off + (size_t)NS * WBy, gy - ly, lo(kb * WBy, toty), lo(kb * WBy + ABy, toty), bar);
				if (HASW) tx += ring_fetch(smem + a.w_off + (size_t)NS * WBy, gw - lw, lo(kb * WBy, totw), lo(kb * WBy + ABy, totw), bar);
				if (HAST) tx += ring_fetch(smem + a.t_off + (size_t)NS * WBt, gt - lt, lo(kb * WBt, tott), lo(kb * WBt + ABt, tott), bar);
			}
			if (tx)
				mbar_arrive_expect_tx(bar, tx);
			else
				mbar_arrive(bar);
			st_release_s32(&issued_s, k + 1);
		}
		// no copy may be in flight when the CTA retires
		for (int k = nwin > NS ? nwin - NS : 0; k < nwin; ++k) mbar_wait(&full_bar[k % NS], (uint32_t)((k / NS) & 1));
		return;
	}

	// ------------------------------ consumers ------------------------------
	const int cw = warp - 1;
	const bool slice = a.shift_common != nullptr;
	// teams of TW warps (a power of two dividing NCW); slice mode: the whole CTA is one team (TW == NCW)
	const int ltw = 31 - __clz(TW);
	const int team = slice ? 0 : cw >> ltw, ti = slice ? cw : cw & (TW - 1), nteams = slice ? 1 : NCW / TW;
	const int rl = lane & 3, cl = lane >> 2;
	const uint32_t xs = smem_u32(smem + a.x_off) + lx, ys = smem_u32(smem + a.y_off) + ly;
	const uint32_t ws = HASW ? smem_u32(smem + a.w_off) + lw : 0u, ts = HAST ? smem_u32(smem + a.t_off) + lt : 0u;
	const uint32_t kone = smem_u32(&kconst[0]), kzero = smem_u32(&kconst[1]);
	const int L = rec_len(p);
	// Window cursor, kept without divisions: kn = next window to wait for (slot kn_s, parity kn_ph), `edge` =
	// its nominal end row (kn + 1) SR; rows [0, ready) have landed as far as this warp reads them. A warp waits
	// only for the windows it reads; the windows of other teams' groups are stepped over without waiting.
	int kn = 0, kn_s = 0, issued = 0;
	uint32_t kn_ph = 0;
	int64_t edge = SR, ready = -15;
	const uint32_t full0 = smem_u32(&full_bar[0]);
	const int64_t la_rows = ring_rows / 4 > SR ? ring_rows / 4 : SR;
	auto advance = [&]() {
		++kn;
		if (++kn_s == NS) {
			kn_s = 0;
			kn_ph ^= 1u;
		}
		edge += SR;
		// a row's last byte may sit up to 15 bytes (<= 15 rows of the flag array) into the next window
		ready = (kn == nwin) ? ((int64_t)1 << 62) : edge - SR - 15;
	};
	// rows [0, need) of the CTA's range have landed
	auto ensure = [&](int64_t need) {
		while (need > ready) {
			// a warp that stepped over other teams' windows may be more than a lap ahead of the producer: the
			// parity wait is only meaningful once window kn itself has been issued
			while (issued <= kn) {
				issued = ld_volatile_s32(&issued_s);
				if (issued <= kn) __nanosleep(64);
			}
			mbar_wait_addr(full0 + 8u * (uint32_t)kn_s, kn_ph);
			advance();
		}
		// look ahead: take further windows that have already landed, so that the batch below runs over more than
		// the one window that was asked for -- but not over the whole ring: rows are released at the end of a
		// batch, and the producer needs released windows early enough to keep copies in flight
		const int64_t la_end = need + la_rows;
		while (kRingPipelined<NT> && kn < nwin && ready < la_end) {
			if (issued <= kn) {
				issued = ld_volatile_s32(&issued_s);
				if (issued <= kn) break;
			}
			if (!mbar_test_addr(full0 + 8u * (uint32_t)kn_s, kn_ph)) break;
			advance();
		}
	};
	// this warp will not read rows < pos again: step over the windows that lie entirely below and publish the
	// position (the producer refills a slot once every warp's position has passed the window it held)
	auto release = [&](int64_t pos) {
		while (edge <= pos && kn < nwin) advance();
		__syncwarp();
		if (lane == 0) st_release_s64(&pos_s[cw], pos);
	};
	// per-lane address pattern of the tile columns: base + ring_row * rmul, step stride 4 * rmul
	// Tiles 0 .. NT-2 hold feature columns only, so one pointer serves them all (tile t = +64 t bytes, row stride
	// 8 p); only the last tile has lanes that read y, the constant 1 or the zero padding.
	const uint32_t base0 = xs + (uint32_t)(rl * p + cl) * 8u, st0 = 32u * (uint32_t)p;
	uint32_t baseL, stL;
	{
		const int c = 8 * (NT - 1) + cl;
		if (c < p) {
			baseL = xs + (uint32_t)(rl * p + c) * 8u;
			stL = 32u * (uint32_t)p;
		} else if (!SIDE && c == p) {
			baseL = ys + (uint32_t)rl * 8u;
			stL = 32u;
		} else if (!SIDE && c == p + 1) {
			baseL = kone;
			stL = 0u;
		} else {
			baseL = kzero;
			stL = 0u;
		}
	}
	// the per-tile form the (rare) tail and shift code uses
	auto tile_ptrs = [&](uint32_t (&b)[NT], uint32_t (&s)[NT]) {
#pragma unroll
		for (int t = 0; t < NT; ++t) {
			b[t] = t < NT - 1 ? base0 + 64u * (uint32_t)t : baseL;
			s[t] = t < NT - 1 ? st0 : stL;
		}
	};
	// rows of a multiple of 128 bytes: odd rows load the tiles of a pair in swapped order (swizzle_pairs)
	const bool swz = (p % 16 == 0) && (rl & 1);
	auto tile_ptrs_swz = [&](uint32_t (&b)[NT], uint32_t (&s)[NT]) {
		tile_ptrs(b, s);
		if (p % 16 == 0) {
#pragma unroll
			for (int t = 0; t + 1 < NT; t += 2) {
				const uint32_t lo = b[t], hi = b[t + 1];
				b[t] = swz ? hi : lo;
				b[t + 1] = swz ? lo : hi;
			}
		}
	};
	uint32_t n_comb = 0; // team combines so far (parity of dep_bar / rel_bar)
	int64_t g = ga + team;
	int64_t nr0 = 0, nr1 = 0;
	if (g < gb) {
		nr0 = a.offsets[g];
		nr1 = a.offsets[g + 1];
	}
	for (; g < gb; g += nteams) {
		const int64_t r0 = nr0 - R0, n64 = nr1 - nr0;
		if (g + nteams < gb) { // the next group's bounds, a group ahead of their use
			nr0 = a.offsets[g + nteams];
			nr1 = a.offsets[g + nteams + 1];
		}
		double *rec = a.records + (slice ? (g * TW + ti) : g) * L;
		if (n64 <= 0) {
			if (ti == 0 || slice) store_empty_record(rec, p, lane, 0);
			continue;
		}
		if (!slice && n64 > a.big_rows) {
			if (ti == 0 && lane == 0) {
				const int slot = atomicAdd(a.big_count, 1);
				a.big_list[slot] = g;
			}
			continue;
		}
		const int n = (int)n64;
		release(r0);
		const int rp0 = r0 < 0x7fffffff ? (int)((uint32_t)r0 % (uint32_t)ring_rows) : (int)(r0 % ring_rows);
		double sh[NT], shy;
		int start = 0, ns_pre = 0;
		if (slice) {
#pragma unroll
			for (int t = 0; t < NT; ++t) {
				const int c = 8 * t + cl;
				sh[t] = (c < p) ? a.shift_common[c] : ((!SIDE && c == p) ? a.shift_common[p] : 0.0);
			}
			shy = a.shift_common[p];
		} else {
			int f = 0;
			if (MASK) {
				// first row that passes the cheap mask, searched 32 rows at a time
				f = -1;
				for (int sb = 0; sb < n; sb += 32) {
					const int hi = sb + 32 < n ? sb + 32 : n;
					ensure(r0 + hi);
					const int r = sb + lane;
					bool ok = r < n, trn = r < n;
					if (ok) {
						const uint32_t rr = (uint32_t)((rp0 + r) % ring_rows);
						if (HASW) ok = lds_f64(ws + rr * 8u) > 0.0;
						if (HAST) {
							trn = lds_u8(ts + rr) != 0u;
							ok = ok && trn;
						}
					}
					const unsigned m = __ballot_sync(0xffffffffu, ok);
					if (m) {
						f = sb + __ffs(m) - 1;
						start = sb;
						break;
					}
					if (HAST) ns_pre += __popc(__ballot_sync(0xffffffffu, trn));
					release(r0 + hi);
				}
				if (f < 0) { // no row passes the mask: "no valid data" with the state's row count
					if (ti == 0) store_empty_record(rec, p, lane, HAST ? ns_pre : n);
					continue;
				}
			}
			ensure(r0 + f + 1);
			uint32_t rf = (uint32_t)(rp0 + f);
			if (rf >= (uint32_t)ring_rows) rf %= (uint32_t)ring_rows;
			uint32_t base[NT], st[NT];
			tile_ptrs(base, st);
#pragma unroll
			for (int t = 0; t < NT; ++t) {
				const int c = 8 * t + cl;
				const bool data = (c < p) || (!SIDE && c == p);
				// the lane's own column in row rf: undo the lane's row offset rl
				sh[t] = data ? lds_f64(base[t] + (rf - (uint32_t)rl) * (st[t] >> 2)) : 0.0;
			}
			shy = lds_f64(ys + rf * 8u);
		}

		WarpAcc<NT, SIDE> A;
		A.clear();
		double e0[NSET > 1 ? NSET - 1 : 1][NTP], e1[NSET > 1 ? NSET - 1 : 1][NTP];
#pragma unroll
		for (int s = 0; s < (NSET > 1 ? NSET - 1 : 1); ++s)
#pragma unroll
			for (int i = 0; i < NTP; ++i) e0[s][i] = e1[s][i] = 0.0;

		// This warp's quads are start + (ti + j TW) QR, j = 0, 1, ...: full quads run in batches (as many as the
		// landed windows and the ring end allow) with nothing but pointer increments between them.
		int rp = rp0 + start + ti * QR;
		if (rp >= ring_rows) rp %= ring_rows;
		const int rstep = TW * QR;
		int row = start + ti * QR;
		const uint32_t skip_rows = (uint32_t)(rstep - QR);
		while (row + QR <= n) {
			ensure(r0 + row + QR);
			const int64_t av64 = ready - r0;
			const int avail = av64 > (int64_t)n ? n : (int)av64;
			const int span = avail - row - QR; // >= 0
			const int nq = (slice ? span / rstep : span >> (ltw + (NSET == 4 ? 4 : (NSET == 2 ? 3 : 2)))) + 1;
			uint32_t py = ys + (uint32_t)(rp + rl) * 8u, pw = ws + (uint32_t)(rp + rl) * 8u, pt = ts + (uint32_t)(rp + rl);
			// quads of this batch: as many as have landed, up to the ring end
			// (group mode: rstep = TW * QR is a power of two -- a shift, like nq above, not a division per batch)
			const int to_end = slice ? (ring_rows - rp + rstep - 1) / rstep
			                         : (ring_rows - rp + rstep - 1) >> (ltw + (NSET == 4 ? 4 : (NSET == 2 ? 3 : 2)));
			const int nqe = nq < to_end ? nq : to_end;
			int q = 0;
			if constexpr (kRingPipelined<NT>) {
				// few warps per SM (wide Gram): loads a step ahead, one pointer for all full tiles
				uint32_t p0 = base0 + (uint32_t)rp * (st0 >> 2), pl = baseL + (uint32_t)rp * (stL >> 2);
				StepIn<NT, SIDE, HASW, HAST> in;
				step_load<NT, SIDE, HASW, HAST>(in, p0, pl, st0, stL, py, pw, pt);
				do {
#pragma unroll
					for (int s = 0; s < NSET; ++s) {
						if (s == NSET - 1) { // the next step is the first of the next quad
							p0 += skip_rows * (st0 >> 2);
							pl += skip_rows * (stL >> 2);
							py += skip_rows * 8u;
							pw += skip_rows * 8u;
							pt += skip_rows;
						}
						const bool more = s + 1 < NSET || q + 1 < nqe;
						if (s == 0)
							mma_step_pipe<NT, SIDE, HASW, HAST, HASW, false>(A, A.c0, A.c1, in, p0, pl, st0, stL, py, pw, pt, sh, shy, more);
						else
							mma_step_pipe<NT, SIDE, HASW, HAST, HASW, false>(A, e0[s - 1], e1[s - 1], in, p0, pl, st0, stL, py, pw, pt, sh, shy, more);
					}
					rp += rstep;
					++q;
				} while (q < nqe);
			} else {
				// many warps per SM hide the shared-memory latency by themselves: plain steps
				uint32_t pa[NT], st[NT];
				tile_ptrs_swz(pa, st);
#pragma unroll
				for (int t = 0; t < NT; ++t) pa[t] += (uint32_t)rp * (st[t] >> 2);
				do {
#pragma unroll
					for (int s = 0; s < NSET; ++s) {
						if (s == 0)
							mma_step<NT, SIDE, HASW, HAST, false, HASW, false>(A, A.c0, A.c1, pa, st, py, pw, pt, sh, shy, true, swz);
						else
							mma_step<NT, SIDE, HASW, HAST, false, HASW, false>(A, e0[s - 1], e1[s - 1], pa, st, py, pw, pt, sh, shy, true, swz);
					}
#pragma unroll
					for (int t = 0; t < NT; ++t) pa[t] += skip_rows * (st[t] >> 2);
					py += skip_rows * 8u;
					pw += skip_rows * 8u;
					pt += skip_rows;
					rp += rstep;
					++q;
				} while (q < nqe);
			}
			if (!MASK) A.nv += NSET * q;
			row += q * rstep;
			if (rp >= ring_rows) rp -= ring_rows;
			release(r0 + (row < n ? row : n));
		}
		if (row < n) { // the group's last, partial quad
			const int cnt = n - row;
			ensure(r0 + n);
			uint32_t pa[NT], st[NT];
			tile_ptrs(pa, st);
#pragma unroll
			for (int t = 0; t < NT; ++t) pa[t] += (uint32_t)rp * (st[t] >> 2);
			uint32_t py = ys + (uint32_t)(rp + rl) * 8u, pw = ws + (uint32_t)(rp + rl) * 8u, pt = ts + (uint32_t)(rp + rl);
			const int nfs = cnt >> 2;
#pragma unroll
			for (int s = 0; s < NSET - 1; ++s) {
				if (s < nfs) {
					if (s == 0)
						mma_step<NT, SIDE, HASW, HAST, false, HASW, false>(A, A.c0, A.c1, pa, st, py, pw, pt, sh, shy, true);
					else
						mma_step<NT, SIDE, HASW, HAST, false, HASW, false>(A, e0[s - 1], e1[s - 1], pa, st, py, pw, pt, sh, shy, true);
				}
			}
			if (!MASK) A.nv += nfs;
			if (cnt & 3) {
				const bool in_range = rl < (cnt & 3);
				if (!in_range) { // keep the loads inside the group's rows
#pragma unroll
					for (int t = 0; t < NT; ++t) pa[t] -= (uint32_t)rl * (st[t] >> 2);
					py -= (uint32_t)rl * 8u;
					pw -= (uint32_t)rl * 8u;
					pt -= (uint32_t)rl;
				}
#pragma unroll
				for (int s = 0; s < NSET; ++s) {
					if (s == nfs) {
						if (s == 0)
							mma_step<NT, SIDE, HASW, HAST, true, HASW, false>(A, A.c0, A.c1, pa, st, py, pw, pt, sh, shy, in_range);
						else
							mma_step<NT, SIDE, HASW, HAST, true, HASW, false>(A, e0[s - 1], e1[s - 1], pa, st, py, pw, pt, sh, shy, in_range);
					}
				}
			}
			release(r0 + n);
		}
		// the per-step counts the unweighted steps left out: valid rows = training rows, sum w = valid rows
		if (HAST && !HASW) A.nv = A.ns;
		if (SIDE && !HASW) A.sw = (double)A.nv;
		if (ti == 0 && lane == 0) A.ns += ns_pre;
		// fold the accumulator sets in a fixed order
#pragma unroll
		for (int s = 0; s < NSET - 1; ++s)
#pragma unroll
			for (int i = 0; i < NTP; ++i) {
				A.c0[i] += e0[s][i];
				A.c1[i] += e1[s][i];
			}

		if (slice) {
			// every warp keeps its own partial record (reduce_partials_kernel adds them); a poisoned or
			// undecided partial only raises the redo count: the caller falls back to the checked path
			int nvt = A.nv + __shfl_xor_sync(0xffffffffu, A.nv, 1);
			nvt += __shfl_xor_sync(0xffffffffu, nvt, 2);
			bool bad = acc_poisoned<NT, SIDE>(A);
			if (!HASW) {
				bool band = false;
#pragma unroll
				for (int t = 0; t < NT; ++t) {
					const int dp = t * NT - (t * (t - 1)) / 2;
					const double v = (cl & 1) ? A.c1[dp] : A.c0[dp];
					const bool mine = rl == (cl >> 1);
					const bool sure = v > 1.01e-20 * (double)nvt;
					A.nc[t] = mine && sure;
					band = band || (mine && !sure && !(v < 0.99e-20) && 8 * t + cl < p);
				}
				bad = bad || __any_sync(0xffffffffu, band);
			}
			if (bad && lane == 0) atomicAdd(a.redo_count, 1);
			store_record<NT, SIDE>(A, rec, p, lane, sh, shy, HAST ? -1 : (ti == 0 ? n : 0));
			continue;
		}

		if (TW > 1) {
			double *slots = (double *)(smem + a.slot_off);
			if (ti != 0) {
				// member: wait until the lead has read the previous deposit, then leave this one
				double *slot = slots + (size_t)(team * (TW - 1) + (ti - 1)) * NV * 32 + lane;
				if (n_comb > 0) mbar_wait(&rel_bar[team], (n_comb - 1) & 1u);
				int i = 0;
				ring_for_each<NT, SIDE>(A, [&](double &v) {
					slot[i * 32] = v;
					++i;
				});
				unsigned ncb = 0u;
#pragma unroll
				for (int t = 0; t < NT; ++t) ncb |= A.nc[t] ? (1u << t) : 0u;
				slot[i * 32] = (double)A.nv;
				slot[(i + 1) * 32] = (double)A.ns;
				slot[(i + 2) * 32] = (double)ncb;
				__syncwarp();
				if (lane == 0) mbar_arrive(&dep_bar[team]);
				++n_comb;
				continue;
			}
			mbar_wait(&dep_bar[team], n_comb & 1u);
			for (int m = 1; m < TW; ++m) {
				const double *slot = slots + (size_t)(team * (TW - 1) + (m - 1)) * NV * 32 + lane;
				int i = 0;
				ring_for_each<NT, SIDE>(A, [&](double &v) {
					v += slot[i * 32];
					++i;
				});
				A.nv += (int)slot[i * 32];
				A.ns += (int)slot[(i + 1) * 32];
				const unsigned ncb = (unsigned)slot[(i + 2) * 32];
#pragma unroll
				for (int t = 0; t < NT; ++t) A.nc[t] = A.nc[t] || ((ncb >> t) & 1u);
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(&rel_bar[team]);
			++n_comb;
		}
		// lead warp: decide the constant-column flags, check for poison, store
		int nvt = A.nv + __shfl_xor_sync(0xffffffffu, A.nv, 1);
		nvt += __shfl_xor_sync(0xffffffffu, nvt, 2);
		bool bad = acc_poisoned<NT, SIDE>(A);
		if (!HASW) {
			// unweighted: sum d^2 (the diagonal of the Gram) brackets max d^2: max d^2 <= sum d^2 <= n max d^2
			bool band = false;
#pragma unroll
			for (int t = 0; t < NT; ++t) {
				const int dp = t * NT - (t * (t - 1)) / 2; // pair (t, t)
				const double v = (cl & 1) ? A.c1[dp] : A.c0[dp];
				const bool mine = rl == (cl >> 1); // this lane holds C[cl][cl] of the diagonal tile
				const bool sure = v > 1.01e-20 * (double)nvt;
				A.nc[t] = mine && sure;
				band = band || (mine && !sure && !(v < 0.99e-20) && 8 * t + cl < p);
			}
			bad = bad || __any_sync(0xffffffffu, band);
		}
		if (bad) {
			if (lane == 0) {
				const int slot = atomicAdd(a.redo_count, 1);
				a.redo_list[slot] = g;
			}
			continue;
		}
		store_record<NT, SIDE>(A, rec, p, lane, sh, shy, HAST ? -1 : n);
	}
	// pass the remaining windows on, so that the producer can finish for the other warps
	release((int64_t)1 << 62);
}

// Checked path for the groups the ring kernel could not finish: one warp per listed group, rows straight from
// global memory with the per-element filter of ols.rs:59-66 and the exact constant-column flag.
template <int NT, bool SIDE> __global__ void __launch_bounds__(256) redo_kernel(const RingArgs a) {
	const int lane = threadIdx.x & 31;
	const int nw = (gridDim.x * blockDim.x) >> 5;
	const int n_list = *a.redo_count;
	const int p = a.p, L = rec_len(p);
	for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n_list; i += nw) {
		const int64_t g = a.redo_list[i];
		const int64_t r0 = a.offsets[g], r1 = a.offsets[g + 1];
		const int64_t n = r1 - r0;
		const double *xs = a.x + r0 * p, *ys = a.y + r0, *ws = a.w ? a.w + r0 : nullptr;
		const uint8_t *ts = a.train ? a.train + r0 : nullptr;
		double *rec = a.records + g * L;
		// first fully valid row, searched in chunks that keep the int row index in range
		int64_t f = -1;
		for (int64_t cb = 0; cb < n && f < 0; cb += (1 << 20)) {
			const int cn = (int)(n - cb < (1 << 20) ? n - cb : (1 << 20));
			const int fc = find_first_valid<NT, SIDE, true>(xs + cb * p, ys + cb, ws ? ws + cb : nullptr, ts ? ts + cb : nullptr, cn, p, lane);
			if (fc >= 0) f = cb + fc;
		}
		WarpAcc<NT, SIDE> A;
		A.clear();
		double sh[NT], shy = 0.0;
#pragma unroll
		for (int t = 0; t < NT; ++t) sh[t] = 0.0;
		if (f >= 0) load_shift<NT, SIDE>(xs + f * p, ys + f, 0, p, lane, sh, shy);
		// without a valid row every row is masked below (the shift is irrelevant); the pass still counts n_state
		for (int64_t cb = 0; cb < n; cb += (1 << 20)) {
			const int cn = (int)(n - cb < (1 << 20) ? n - cb : (1 << 20));
			acc_rows<NT, SIDE, true>(A, xs + cb * p, ys + cb, ws ? ws + cb : nullptr, ts ? ts + cb : nullptr, 0, cn, p, lane, sh, shy,
			                         a.flags_from_y != 0);
		}
		if (f < 0) {
			int ns = A.ns + __shfl_xor_sync(0xffffffffu, A.ns, 1);
			ns += __shfl_xor_sync(0xffffffffu, ns, 2);
			store_empty_record(rec, p, lane, __shfl_sync(0xffffffffu, ns, 0));
		} else {
			store_record<NT, SIDE>(A, rec, p, lane, sh, shy, -1);
		}
	}
}

} // namespace b200
