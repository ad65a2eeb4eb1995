// resident.cuh -- fit + predict in ONE pass over the rows for groups that fit a shared-memory stage (p <= 8).
//
// Replaces, for the *_fit_predict_agg / *_fit_predict_by path (src/aggregate_functions/ols_predict_aggregate.cpp:
// 322-426: fit the group's training rows, then one anofox_predict_with_interval call per row of the group), the
// three-launch sequence accumulate -> solve -> predict, which reads X from HBM twice. Here a group's rows arrive
// once, by TMA bulk copies into a ring of stages (producer_loop, accumulate.cuh), and stay in shared memory until
// the group is finished:
//   warp 0            producer: whole groups -> stages (one cp.async.bulk per array), big groups -> big_list
//   warps 1 .. 6      three teams of two warps; a group belongs to ONE team from start to end:
//     1. moments      lane q of the team's 64 folds rows q, q+64, ... into the full triangle of sum w d d' in registers
//                     (small_accumulate, DFMA; for p = 8 / 4 the lanes walk the columns in rotated order, see there),
//                     butterfly inside each warp, the second warp hands its sums over through shared memory
//     2. solve        first warp: small_solve<P>, all arrays in registers; the rows are at hand, so a near-exact fit
//                     takes its residuals from them (no refine queue); ill-conditioned groups go to the QR queue
//     3. predict      both warps: yhat and the interval of anofox_predict_with_interval (lib.rs:2292-2347) for every
//                     row of the group, training or not, from shared memory; lane = row, coefficients in registers
//                     in the lane's rotated column order (so that the 8P-byte row stride does not fold the half-warp
//                     onto a few banks)
// A single warp runs a group's ~10 k instructions at a fraction of an instruction per cycle (dependent FP64 chains),
// and only three groups fit the shared memory: the team halves the chain of the two row passes.
// HBM traffic per row: 8(p+1) [+8 weight] + 1 flag byte in, 25 bytes out -- the algorithmic figure of SURVEY 8(d).
// Groups queued for the QR fallback are predicted again after qr_fit_kernel (list mode of predict_kernel).
#pragma once
#include "fused_small.cuh"

namespace b200 {

constexpr int kResidentTeams = 3;          // two consumer warps each
constexpr int kResidentWarps = 2 * kResidentTeams;
constexpr int kResidentMaxStages = 8;

struct ResidentArgs {
	AccArgs acc;
	FitParams params;
	AnofoxB200FitOutputs out;
	AnofoxB200PredictOutputs pred;
	int write_is_training; // the flags came from y alone: is_training = flag && no NaN feature is written here
	int max_stage_groups;
	unsigned long long *dbg; // experiments: per-phase clock totals of warp 1 / 2 of every CTA (null = off)
};

// what the first warp of a team leaves for the prediction pass of both
struct ResidentPredict {
	double coef[8];
	double b0, margin;
	unsigned live;
	int ok;
};

template <int P, bool HASW>
__global__ void __launch_bounds__(32 * (kResidentWarps + 1), 1) fit_predict_resident_kernel(const ResidentArgs args) {
	constexpr int NCW = kResidentWarps, NTEAM = kResidentTeams;
	constexpr bool ROT = (P == 8 || P == 4); // rows of 64 / 32 bytes: bank-conflict-free column order (small_accumulate)
	constexpr int PVW = P + 3;
	static_assert(P >= 1 && P <= 8, "resident kernel: p <= 8");
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ uint64_t full_bar[kResidentMaxStages], empty_bar[kResidentMaxStages];
	__shared__ SmallStageHeader hdr[kResidentMaxStages];
	__shared__ double pv_scratch[NTEAM][PVW];
	__shared__ double team_x[NTEAM][kSmallTeamDoubles];
	__shared__ ResidentPredict rp[NTEAM];
	__shared__ double fit_x[NTEAM][4 + 8 + 1];
	const AccArgs &a = args.acc;
	const FitParams &Pm = args.params;
	const AnofoxB200FitOutputs &o = args.out;
	const int NS = a.n_stages;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NS; ++s) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], NCW);
		}
		mbar_fence_init();
	}
	__syncthreads();
	if (warp == 0) {
		producer_loop<HASW, true>(a, smem, full_bar, empty_bar, hdr, NS, lane, args.max_stage_groups);
		return;
	}
	const int tm = (warp - 1) >> 1, tw = (warp - 1) & 1; // team, warp inside the team
	const int tbar = 1 + tm;                              // the team's named barrier (0 is __syncthreads)
	double *pvw = pv_scratch[tm];
	const bool want_p = Pm.compute_inference && o.std_errors;
	const DfTables Tb{Pm.tcrit, Pm.gratio, Pm.rcp, Pm.ce, Pm.co};
	const double nanv = nan_f64();
	const int rot = ROT ? small_rot<P>(lane) : 0;
	long long t_wait = 0, t_fit = 0, t_prep = 0, t_bar = 0, t_pred = 0, t_groups = 0;
	const bool dbg = args.dbg != nullptr && tm == 0;
	for (int it = 0;; ++it) {
		const int s = it % NS;
		long long c0 = dbg ? clock64() : 0;
		mbar_wait(&full_bar[s], (uint32_t)((it / NS) & 1));
		if (dbg) t_wait += clock64() - c0;
		const int cnt = hdr[s].count;
		if (cnt < 0) break;
		unsigned char *base = smem + (size_t)s * a.stage_bytes;
		const uint32_t ax = smem_u32(base + hdr[s].lead_x);
		const uint32_t ay = smem_u32(base + a.xcap + hdr[s].lead_y);
		const uint32_t aw = HASW ? smem_u32(base + a.xcap + a.ycap + hdr[s].lead_w) : 0u;
		const uint32_t at = smem_u32(base + a.xcap + a.ycap + a.wcap + hdr[s].lead_t);
		const int64_t g0 = hdr[s].g_first;
		for (int i = (tm + it * 5) % NTEAM; i < cnt; i += NTEAM) {
			const int rb = (int)hdr[s].rowstart[i];
			const int n = (int)hdr[s].rowstart[i + 1] - rb;
			const int64_t g = g0 + i;
			SmallGroupView v;
			v.sg.kone = v.sg.kzero = 0u;
			v.sg.x = ax + (uint32_t)rb * (uint32_t)(P * 8);
			v.sg.y = ay + (uint32_t)rb * 8u;
			v.sg.w = aw + (uint32_t)rb * 8u;
			v.sg.t = at + (uint32_t)rb;
			v.n = n;
			v.q = lane;
			v.T = 32;
			v.mask = 0xffffffffu;
			v.pv = pvw;
			v.refine_count = nullptr;
			v.refine_list = nullptr;
			v.rec_out = nullptr;
			v.stash = nullptr;
			v.team_w = tw;
			v.team_bar = tbar;
			v.team_x = team_x[tm];
			v.fit_x = fit_x[tm];
			// ---- 1 + 2. moments (both warps) and solve (first warp) ----
			if (tw == 0 && want_p && lane == 0) pvw[P + 1] = 0.0;
			__syncwarp();
			c0 = dbg ? clock64() : 0;
			small_group<P, HASW, true, 1, ROT, 2>(v, g, Pm, o, nullptr, a.n_groups);
			if (dbg) {
				const long long c1 = clock64();
				t_fit += c1 - c0;
				c0 = c1;
				t_groups += 1;
			}
			if (tw == 0) {
				__syncwarp();
				if (want_p) {
					const double d1 = pvw[P + 1];
					if (lane <= P && d1 > 0.0) {
						const double pval = pvalue_f(pvw[lane], lane < P ? 1.0 : d1, pvw[P + 2], Tb);
						if (lane < P)
							o.p_values[g * P + lane] = pval;
						else
							o.f_pvalue[g] = pval;
					}
					__syncwarp();
				}
				// what the prediction pass needs: small_solve left it in fit_x (no round trip through the outputs)
				const double *fx = fit_x[tm];
				const bool ok = fx[0] == 0.0;
				double b0 = 0.0, margin = 0.0;
				unsigned live = 0u;
				if (ok) {
					const double icpt = fx[1];
					const bool has_i = !isnan(icpt);
					b0 = has_i ? icpt : 0.0;
					const double rse = fx[2];
					const int64_t nobs = (int64_t)fx[3];
					if (!(isnan(rse) || rse <= 0.0 || nobs <= (int64_t)P + 1)) {
						const int64_t kk = (int64_t)P + (has_i ? 1 : 0);
						const int64_t df = nobs > kk ? nobs - kk : 0;
						if (df > 0) {
							const double tc = (Pm.tcrit && df <= kTcritTable) ? Pm.tcrit[df] : t_critical_dev(Pm.level, (double)df);
							if (!isnan(tc)) margin = tc * (rse * sqrt(1.0 + 1.0 / (double)nobs));
						}
					}
					if (lane < P) {
						const double c = fx[4 + lane];
						rp[tm].coef[lane] = isnan(c) ? 0.0 : c; // aliased / constant columns contribute nothing (lib.rs:2292-2300)
						live = isnan(c) ? 0u : (1u << lane);
					}
					live = __reduce_or_sync(0xffffffffu, live);
				}
				if (lane == 0) {
					rp[tm].b0 = b0;
					rp[tm].margin = margin;
					rp[tm].live = live;
					rp[tm].ok = ok ? 1 : 0;
				}
			}
			if (dbg) {
				const long long c1 = clock64();
				t_prep += c1 - c0;
				c0 = c1;
			}
			named_bar_sync(tbar, 64);
			if (dbg) {
				const long long c1 = clock64();
				t_bar += c1 - c0;
				c0 = c1;
			}
			// ---- 3. predict every row of the group from shared memory, both warps ----
			{
				const bool ok = rp[tm].ok != 0;
				const double b0 = rp[tm].b0, margin = rp[tm].margin;
				const unsigned live = rp[tm].live;
				const bool all_live = live == ((1u << P) - 1u);
				double cf[P];      // coefficient of the column in slot c of this lane
				uint32_t co[P];    // its byte offset inside a row
				unsigned lslot = 0u;
#pragma unroll
				for (int c = 0; c < P; ++c) {
					int col = c + rot;
					col = col >= P ? col - P : col;
					cf[c] = rp[tm].coef[col];
					co[c] = (uint32_t)col * 8u;
					lslot |= ((live >> col) & 1u) << c;
				}
				const int64_t row0 = a.offsets[g];
				for (int r = lane + 32 * tw; r < n; r += 64) {
					const uint32_t xr = v.sg.x + (uint32_t)(r * P) * 8u;
					double sacc = b0;
					bool nanx = false;
					if (all_live) {
#pragma unroll
						for (int c = 0; c < P; ++c) {
							const double xv = lds_f64(xr + co[c]);
							nanx = nanx || isnan(xv);
							sacc = fma(cf[c], xv, sacc);
						}
					} else {
#pragma unroll
						for (int c = 0; c < P; ++c) {
							const double xv = lds_f64(xr + co[c]);
							nanx = nanx || isnan(xv);
							if ((lslot >> c) & 1u) sacc = fma(cf[c], xv, sacc);
						}
					}
					double yh = nanv, lo = nanv, hi = nanv;
					if (ok && finite_f64(sacc)) { // ols_predict_aggregate.cpp:409-417
						yh = sacc;
						lo = sacc - margin;
						hi = sacc + margin;
					}
					args.pred.yhat[row0 + r] = yh;
					args.pred.yhat_lower[row0 + r] = lo;
					args.pred.yhat_upper[row0 + r] = hi;
					if (args.write_is_training) args.pred.is_training[row0 + r] = (lds_u8(v.sg.t + (uint32_t)r) != 0u && !nanx) ? 1 : 0;
				}
			}
			// the team's scratch (team_x, rp) is rewritten for the next group only after both warps are here
			named_bar_sync(tbar, 64);
			if (dbg) t_pred += clock64() - c0;
		}
		__syncwarp();
		if (lane == 0) mbar_arrive(&empty_bar[s]);
	}
	if (dbg && lane == 0) {
		unsigned long long *d = args.dbg + 8 * tw;
		atomicAdd(d + 0, (unsigned long long)t_wait);
		atomicAdd(d + 1, (unsigned long long)t_fit);
		atomicAdd(d + 2, (unsigned long long)t_prep);
		atomicAdd(d + 3, (unsigned long long)t_bar);
		atomicAdd(d + 4, (unsigned long long)t_pred);
		atomicAdd(d + 5, (unsigned long long)t_groups);
	}
}

} // namespace b200
