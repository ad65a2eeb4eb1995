#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200 group-fit path (contract: task prompt, section 4).

Workload (BASELINE.json configs[1], "C2"): ols_fit_agg over 1 M groups x 50 rows x 5 features, f64,
full inference, synthetic. One "step" = one pass of the hot path (accumulate + solve/inference)
over the whole batch. `value` = fits/s with inputs resident in HBM; `e2e` = the same through the
host-pointer C-ABI call (pinned host buffers, H2D + D2H inside the timed region).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C1|C2|C3r|C3w|C4|C5]
  torchrun --nproc-per-node N bench.py --gpus N ...      (weak scaling: every rank fits its own C2)
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
# stdout carries exactly one JSON line. Libraries print there too (NCCL's version banner on boxes that export
# NCCL_DEBUG), so file descriptor 1 is pointed at stderr for the life of the process and the JSON line goes to the
# saved descriptor of the real stdout.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(obj):
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())

sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (G, n, p, method, inference, options)
    "C1": (10_000, 100, 3, "ols", False, {}),
    "C2": (1_000_000, 50, 5, "ols", True, {}),
    "C2n": (1_000_000, 50, 5, "ols", False, {}),  # experiments: C2 without inference
    "C2a": (1_000_000, 48, 5, "ols", True, {}),   # experiments: 128-byte aligned groups
    "C3r": (1_000_000, 200, 16, "ridge", False, {"alpha": 1.0, "rank_deficient": True}),
    "C3w": (1_000_000, 200, 16, "wls", False, {"weights": True, "rank_deficient": True}),
    "C5": (100_000, 1_000, 8, "ols", False, {"holdout": 0.2, "predict": True}),
    # one group, rows sharded over the ranks (125 M rows x 64 features per GPU; 1 B rows on 8 GPUs)
    "C4": (1, 125_000_000, 64, "ols", True, {"row_sharded": True}),
}


def algorithmic_bytes(G, N, p, method, inference, predict=False):
    """compulsory traffic, each input element once, results once."""
    per_row = 8 * (p + 1) + (8 if method == "wls" else 0)
    inp = N * per_row + 8 * (G + 1)
    out = G * (8 * p + 48 + ((40 * p + 16) if inference else 0))
    if predict:
        out += N * 25
    return inp, out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_reference_run(wl, host, steps, warmup, threads, budget_s=12.0):
    """Time the CPU oracle (port of the reference path; the Rust crates are unavailable) on a bounded sample."""
    import numpy as np
    import oracle  # test infrastructure, allowed here as the timed CPU baseline only
    G, n, p, method, inference, extra = wl
    opt = dict(method=method, compute_inference=inference, alpha=extra.get("alpha", 0.0))
    # every host core this process may use -- not omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1
    threads = threads or len(os.sched_getaffinity(0)) or oracle.max_threads()
    # pilot to size the sample
    gs = min(G, 2000)
    t0 = time.perf_counter()
    oracle.fit_batch(host["offsets"][:gs + 1], host["y"][:gs * n], host["x"][:gs * n], host["w"][:gs * n] if host["w"] is not None else None,
                     precision=1, threads=threads, **opt)
    rate = gs / max(time.perf_counter() - t0, 1e-6)
    per_step = budget_s / max(steps + warmup, 1)
    gs = int(max(1000, min(G, rate * per_step)))
    sub = (host["offsets"][:gs + 1], host["y"][:gs * n], host["x"][:gs * n],
           host["w"][:gs * n] if host["w"] is not None else None)
    for _ in range(warmup):
        oracle.fit_batch(*sub, precision=1, threads=threads, **opt)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.fit_batch(*sub, precision=1, threads=threads, **opt)
    dt = (time.perf_counter() - t0) / steps
    return gs / dt, dt, gs, threads


def make_c4_shard(rows, p, seed, device="cuda", chunk=2_000_000):
    """x ~ N(0,1) with AR(1) correlation 0.3 across features, y linear + noise."""
    import math
    import torch
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    x = torch.empty((rows, p), dtype=torch.float64, device=device)
    y = torch.empty(rows, dtype=torch.float64, device=device)
    gb = torch.Generator(device=device)
    gb.manual_seed(4)  # the coefficients are the same on every rank: it is ONE group
    beta = torch.rand(p, generator=gb, dtype=torch.float64, device=device) * 2 - 1
    c = math.sqrt(1 - 0.09)
    for r0 in range(0, rows, chunk):
        r1 = min(rows, r0 + chunk)
        z = torch.randn((r1 - r0, p), generator=gen, dtype=torch.float64, device=device)
        for j in range(1, p):
            z[:, j].mul_(c).add_(z[:, j - 1], alpha=0.3)
        x[r0:r1] = z
        y[r0:r1] = 1.5 + z @ beta + torch.randn(r1 - r0, generator=gen, dtype=torch.float64, device=device) * 0.7
        del z
    return x, y


def run_c4(args, rank, world, local):
    """Row-sharded single group: moments per rank + NCCL all-reduce of the moment block + replicated solve."""
    import torch
    import torch.distributed as dist
    from importlib import import_module
    from __graft_entry__ import load_package
    pkg = load_package()
    pkg.device_check()
    D = import_module("anofox_statistics_b200.distributed")
    G, n, p, method, inference, extra = WORKLOADS["C4"]
    rows = max(4096, int(n * args.scale))
    x, y = make_c4_shard(rows, p, seed=40 + rank)
    fit = D.ShardedSingleGroupFit(rows, p, inference=inference)

    def step():
        return fit.fit(y, x, None, method=method, compute_inference=inference)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    pkg.stage_timing(True)
    l0 = pkg.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = pkg.launch_count() - l0
    calls, acc_ms, red_ms = pkg.stage_timing_read()
    pkg.stage_timing(False)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    # the accumulate kernel's launch time on every rank: a slow step with equal kernel times is the exchange, a slow
    # rank shows here
    acc_all = torch.tensor([acc_ms / max(calls, 1)], dtype=torch.float64, device="cuda")
    acc_ranks = [acc_all.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(acc_ranks, acc_all)
    acc_ranks = [float(a.item()) for a in acc_ranks]
    total_rows = world * rows
    value = total_rows / (ms_per_step * 1e-3)
    # every rank must hold the same result (replicated solve on identical reduced blocks)
    coef = out["coefficients"].clone()
    same = True
    if world > 1:
        ref = coef.clone()
        dist.broadcast(ref, 0)
        flag = torch.tensor([1.0 if torch.equal(ref, coef) else 0.0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        same = bool(flag.item() == 1.0)
    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        k = p + 2
        acc_launch_ms = acc_ms / max(calls, 1)
        acc_bytes = rows * 8 * (p + 1)
        achieved = acc_bytes / (acc_launch_ms * 1e-3) / 1e9 if acc_launch_ms > 0 else None
        flops = rows * k * (k + 1)  # symmetric Gram with the y and 1 columns, 2 flop per MAC on the triangle
        tf = flops / (acc_launch_ms * 1e-3) / 1e12 if acc_launch_ms > 0 else None
        # p = 64 is ~8.5 flop/B: the FP64 tensor pipe bounds this kernel (floor 0.55 TFLOP / 37 TFLOP/s = 15 ms per
        # GPU), HBM is the lower floor (65 GB / 6.5 TB/s = 9.9 ms). Peak = DMMA.8x8x4 issue rate measured on this
        # pool by tools/microbench.cu (profiles/r01d_microbench.txt): 37.0 TFLOP/s at 1965 MHz.
        dmma_peak = 37.0
        roofline = {"bound": "tensor", "kernel": "acc_slice_kernel<8,SIDE> (FP64 DMMA Gram of one row shard, "
                                                 "split-role warps, TMA-fed ring)",
                    "achieved": tf, "peak": dmma_peak,
                    "peak_source": "builder-measured FP64 DMMA m8n8k4 issue rate (tools/microbench.cu, profiles/r01d_microbench.txt); "
                                   "MEASURED_PEAKS.json has no FP64 figure; nominal 40 TFLOP/s quoted beside it",
                    "unit": "TFLOP/s", "frac": (tf / dmma_peak) if tf else None,
                    "traffic": None, "traffic_source": load_traffic().get("C4"),  # captured at 1/10 of the rows: not per-launch here
                    "frac_of_nominal_40_tflops": (tf / 40.0) if tf else None,
                    "algorithmic_flops_per_launch": flops, "algorithmic_bytes_per_launch": acc_bytes,
                    "avg_launch_ms": acc_launch_ms, "avg_launch_ms_per_rank": acc_ranks,
                    "hbm_achieved_gbs": achieved, "hbm_peak_gbs": peak, "hbm_peak_source": peak_src,
                    "hbm_frac": (achieved / peak) if achieved else None,
                    "note": "algorithmic flops = rows * k(k+1), k = p + 2 (symmetric Gram with the y and 1 columns); "
                            "the kernel executes 36 8x8x4 tile products per 4 rows = 4608 flop/row (+4 %)",
                    "reduce_and_split_ms": red_ms / max(calls, 1)}
        line = {"metric": "rows_per_sec", "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "single-group OLS, %d rows x %d features per GPU, row-sharded, "
                                       "NCCL all-reduce of the moment block" % (rows, p),
                           "rows_per_gpu": rows, "features": p, "method": method, "inference": inference,
                           "collectives_per_step": "all_gather(p+2) + all_reduce SUM(%d f64) + all_reduce MAX(%d f64)"
                                                   % (pkg.moment_block_len(p), p),
                           "l2_policy": "inputs (%.1f GB/step/GPU) exceed the 126 MB L2" % (acc_bytes / 1e9)},
                "fits_per_sec": 1.0 / (ms_per_step * 1e-3), "identical_on_all_ranks": same,
                "status": int(out["status"][0].item()), "n_observations": int(out["n_observations"][0].item()),
                "clocks": clocks, "e2e": None, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": None}
        return line
    return None


def bind_to_gpu_numa(local):
    """Pin this process to the CPU cores NVML reports as closest to its GPU, so that the pinned host buffers of the
    e2e leg are first-touched on that GPU's NUMA node. Returns a short description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1}
        allowed = os.sched_getaffinity(0)
        use = sorted(cpus & allowed)
        if use:
            os.sched_setaffinity(0, use)
            return "nvml cpu affinity of gpu %d: %d cores (%d..%d)" % (local, len(use), use[0], use[-1])
        return "nvml affinity disjoint from the cpuset (%d allowed cores): unchanged" % len(allowed)
    except Exception as e:  # NVML missing or not permitted: report, do not fail the bench
        return "unchanged (%s)" % type(e).__name__


def load_traffic():
    """profiles/traffic.json: dram bytes per launch of each workload's dominant kernel from an `ncu --set full`
    capture, written by tools/ncu_traffic.py together with the capture's own kernel duration and source file."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


NAMES = {"C1": "ols_fit_agg 10k groups x 100 rows x 3 features",
         "C2": "ols_fit_agg 1M groups x 50 rows x 5 features, full inference",
         "C2n": "ols_fit_agg 1M groups x 50 rows x 5 features, no inference (experiment)",
         "C2a": "ols_fit_agg 1M groups x 48 rows x 5 features (experiment: 128-byte aligned groups)",
         "C3r": "ridge_fit_agg(lambda=1) 1M groups x 200 rows x 16 features incl. rank-deficient groups",
         "C3w": "wls_fit_agg 1M groups x 200 rows x 16 features incl. rank-deficient groups",
         "C5": "ols_fit_predict_by 100k groups x 1k rows x 8 features, 20% held-out"}


def run_window_workload(args, rank, world, local, steps, warmup, G=1_000_000, n=100, p=3):
    """The reference's own headline benchmark (examples/performance_1m_groups/benchmark_ols.sql:6-22):
    ols_fit_predict(y, [x1, x2, x3]) OVER (PARTITION BY g ORDER BY t ROWS BETWEEN UNBOUNDED PRECEDING AND 1 PRECEDING)
    over 1 M partitions x 100 rows = 100 M predictions; published there: 178.6 s (other hardware, DuckDB CPU)."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from importlib import import_module
    from __graft_entry__ import load_package
    pkg = load_package()
    synth = import_module("anofox_statistics_b200.synth")
    G = max(64, int(G * args.scale))
    N = G * n
    d = synth.make_torch(G, n, p, seed=11 + rank)
    out = {k: torch.empty(N, dtype=torch.float64, device="cuda") for k in ("yhat", "yhat_lower", "yhat_upper")}
    ps = pkg.abi.B200PredictOutputs()
    for nm, ctype in pkg.abi.B200PredictOutputs._fields_:
        setattr(ps, nm, C.cast(out[nm].data_ptr(), ctype) if nm in out else None)
    wsb = pkg.window_workspace_bytes(N, p)
    ws = torch.empty(wsb + 256, dtype=torch.uint8, device="cuda")
    wp = (ws.data_ptr() + 255) & ~255
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        pkg.window_fit_predict_device(d["offsets"].data_ptr(), d["y"].data_ptr(), d["x"].data_ptr(), None, N, G, p, ps,
                                      wp, wsb, stream, preceding=-1, end_offset=1, method="ols")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    l0 = pkg.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    n_pred = int(torch.isfinite(out["yhat"]).sum().item())
    res = {"workload": "ols_fit_predict OVER (PARTITION BY g ORDER BY t ROWS BETWEEN UNBOUNDED PRECEDING AND 1 PRECEDING), "
                       "%d partitions x %d rows x %d features" % (G, n, p),
           "ms_per_step": ms, "predictions_per_sec": world * N / (ms * 1e-3), "rows": N, "non_null_predictions": n_pred,
           "steps": steps, "gpu_launches": pkg.launch_count() - l0,
           "algorithmic_bytes": N * (8 * (p + 1) + 24) + 8 * (G + 1),
           "frac_of_hbm_peak": (N * (8 * (p + 1) + 24) + 8 * (G + 1)) / (ms * 1e-3) / 1e9 / measured_peak_gbs()[0],
           "reference_published_s": 178.6,
           "reference_published_note": "examples/performance_1m_groups/README.md (DuckDB CPU, other hardware)"}
    del d, out, ws
    torch.cuda.empty_cache()
    return res if rank == 0 else None


def run_finalize_pattern(args, rank, threads_list=(1, 8, 16), groups_per_call=2048, calls_per_thread=24, n=50, p=5):
    """DuckDB's own call pattern against the host-pointer entry point: Finalize hands over at most 2 048 group states
    per call (ols_aggregate.cpp:249-257) and runs on many threads at once (radix_partitioned_hashtable.cpp:990).
    T host threads each make `calls_per_thread` calls of 2 048 groups x 50 rows x 5 features (C2's shape, ~4.9 MB in,
    0.6 MB out per call) from pinned buffers; ctypes releases the GIL for the duration of a call."""
    import ctypes as C
    import threading
    import numpy as np
    import torch
    from importlib import import_module
    from __graft_entry__ import load_package
    pkg = load_package()
    abi = pkg.abi
    synth = import_module("anofox_statistics_b200.synth")
    G = groups_per_call
    N = G * n
    res = {}
    for T in threads_list:
        bufs = []
        for t in range(T):
            d = synth.make_numpy(G, n, p, seed=100 + t)
            h = {k: torch.from_numpy(np.ascontiguousarray(d[k])).pin_memory() for k in ("offsets", "y", "x")}
            fit = {"coefficients": torch.empty((G, p), dtype=torch.float64).pin_memory()}
            for k in abi.FIT_F64_G + abi.INF_F64_G + abi.EXTRA_F64_G:
                fit[k] = torch.empty(G, dtype=torch.float64).pin_memory()
            for k in abi.INF_F64_GP:
                fit[k] = torch.empty((G, p), dtype=torch.float64).pin_memory()
            fit["n_observations"] = torch.empty(G, dtype=torch.int64).pin_memory()
            fit["status"] = torch.empty(G, dtype=torch.int32).pin_memory()
            fit["has_inference"] = torch.empty(G, dtype=torch.uint8).pin_memory()
            hs = abi.B200FitOutputs()
            for nm, ctype in abi.B200FitOutputs._fields_:
                setattr(hs, nm, None if nm not in fit else C.cast(fit[nm].data_ptr(), ctype))
            bufs.append((h, fit, hs))

        def work(t, reps):
            h, fit, hs = bufs[t]
            torch.cuda.set_device(torch.cuda.current_device())
            for _ in range(reps):
                pkg.fit_batch_host_into(h["offsets"].data_ptr(), h["y"].data_ptr(), h["x"].data_ptr(), None, hs, G, N, p,
                                        method="ols", compute_inference=True)

        for phase, reps in (("warm", 2), ("timed", calls_per_thread)):
            ths = [threading.Thread(target=work, args=(t, reps)) for t in range(T)]
            t0 = time.perf_counter()
            for th in ths:
                th.start()
            for th in ths:
                th.join()
            dt = time.perf_counter() - t0
        ok = all(int((b[1]["status"] == 0).sum()) == G for b in bufs)
        res["threads_%d" % T] = {"fits_per_sec": T * calls_per_thread * G / dt, "ms_per_call": dt / calls_per_thread * 1e3,
                                 "calls": T * calls_per_thread, "all_groups_ok": ok}
        del bufs
    res["shape"] = "%d groups x %d rows x %d features per call, full inference, pinned host buffers" % (G, n, p)
    return res if rank == 0 else None


def run_next_rows(args, rank, d, G, n, p):
    """The rows SURVEY section 8(f) marks "next", timed on the headline's data (C2's shape), a few steps each:
    device ingest of scan-order rows (f-3), HC1 robust standard errors (f-4), column-major array fits (f-2)."""
    import torch
    from __graft_entry__ import load_package
    pkg = load_package()
    N = G * n
    res = {}

    def timed(fn, steps=3, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    peak = measured_peak_gbs()[0]
    # f-4: HC1 on top of the full-inference fit
    fit = pkg.DeviceFit(G, N, p, inference=True)
    base = timed(lambda: fit.fit(d["offsets"], d["y"], d["x"], None, method="ols", compute_inference=True))
    l0 = pkg.launch_count()
    hc = timed(lambda: fit.fit(d["offsets"], d["y"], d["x"], None, method="ols", compute_inference=True, hc_type="hc1"))
    res["hc1_on_c2"] = {"ms_per_step": hc, "ms_without_hc": base, "hc_pass_ms": hc - base,
                        "gpu_launches_per_step": (pkg.launch_count() - l0) // 5,
                        "finite_std_errors": int(torch.isfinite(fit.out["std_errors"]).all(dim=1).sum().item())}
    # f-2: the same rows as p columns per fit (the layout of ols_fit(y LIST, x LIST(LIST)))
    xc = d["x"].view(G, n, p).transpose(1, 2).contiguous().view(-1)
    col = timed(lambda: fit.fit_columns(d["offsets"], d["y"], xc, None, method="ols", compute_inference=True))
    res["column_fits_on_c2"] = {"ms_per_step": col, "fits_per_sec": G / (col * 1e-3), "transpose_and_fit_vs_fit_ms": [col, base]}
    del xc, fit
    torch.cuda.empty_cache()
    # f-3: scan-order rows -> group-sorted CSR
    gen = torch.Generator(device="cuda")
    gen.manual_seed(5)
    gid = torch.randint(0, G, (N,), generator=gen, dtype=torch.int64, device="cuda")
    ing = pkg.DeviceIngest(N, p, G)
    ms = timed(lambda: ing.run(gid, d["y"], d["x"]))
    byts = N * (8 + 2 * 8 * (p + 1) + 8)  # ids in, rows in + out, row index out
    ok = bool((ing.offsets[1:] - ing.offsets[:-1]).sum().item() == N) and ing.bad_ids() == 0
    res["ingest_50m_rows"] = {"ms_per_step": ms, "rows_per_sec": N / (ms * 1e-3), "algorithmic_bytes": byts,
                              "frac_of_hbm_peak": byts / (ms * 1e-3) / 1e9 / peak, "offsets_consistent": ok,
                              "what": "%d scan-order rows x %d features, %d groups: stable radix partition + gather" % (N, p, G)}
    del ing, gid
    torch.cuda.empty_cache()
    return res if rank == 0 else None


def run_group_workload(args, wname, rank, world, local, steps, warmup, headline, data_cache=None):
    """One group-sharded workload on this rank (weak scaling: every rank fits its own copy of the shape).
    Returns the result dict on rank 0 (None elsewhere)."""
    import torch
    import torch.distributed as dist
    from importlib import import_module
    from __graft_entry__ import load_package
    pkg = load_package()
    synth = import_module("anofox_statistics_b200.synth")
    G, n, p, method, inference, extra = WORKLOADS[wname]
    G = max(64, int(G * args.scale))
    N = G * n
    predict = bool(extra.get("predict"))
    key = (G, n, p, bool(extra.get("rank_deficient")), extra.get("holdout", 0.0))
    if data_cache is not None and data_cache.get("key") == key:
        d = data_cache["d"]
    else:
        if data_cache is not None:
            data_cache.clear()
            torch.cuda.empty_cache()
        # C3r and C3w share one data set (the weights are simply unused by ridge)
        d = synth.make_torch(G, n, p, seed=2 + rank, weights=bool(extra.get("weights")) or wname.startswith("C3"),
                             rank_deficient=extra.get("rank_deficient", False), holdout=extra.get("holdout", 0.0))
        if data_cache is not None:
            data_cache.update(key=key, d=d)
    w_in = d["w"] if method == "wls" else None
    fit = pkg.DeviceFit(G, N, p, inference=inference, predict=predict)
    opt = dict(method=method, compute_inference=inference, alpha=extra.get("alpha", 0.0))

    def step():
        if predict:
            fit.fit_predict(d["offsets"], d["y"], d["x"], w_in, None, **opt)
        else:
            fit.fit(d["offsets"], d["y"], d["x"], w_in, **opt)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    pkg.stage_timing(True)
    l0 = pkg.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = pkg.launch_count() - l0
    calls, acc_ms, solve_ms = pkg.stage_timing_read()
    pkg.stage_timing(False)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / steps
    value = world * G / (ms_per_step * 1e-3)
    n_ok = int((fit.out["status"] == 0).sum().item())
    counters = fit.counters()
    # launch-bound shapes (C1: ~8 launches around 0.05 ms of kernels): the same call captured once into a CUDA graph
    # and replayed -- the library only enqueues work on the caller's stream, so the caller may capture it
    graph_ms = None
    if wname == "C1" and world == 1:
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                step()
            for _ in range(3):
                gr.replay()
            torch.cuda.synchronize()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            for _ in range(20):
                gr.replay()
            g1.record()
            torch.cuda.synchronize()
            graph_ms = g0.elapsed_time(g1) / 20
            if int((fit.out["status"] == 0).sum().item()) != n_ok:
                graph_ms = None
            del gr
        except Exception as e:  # noqa: BLE001
            graph_ms = "capture failed: " + str(e)[:160]

    # ------------------------------------------------------------------ e2e through the host-pointer C ABI
    e2e = None
    if headline and not args.no_e2e:
        import ctypes as C
        abi = pkg.abi
        pin = lambda tt: tt.cpu().pin_memory()  # noqa: E731
        h = {k: (pin(d[k]) if d[k] is not None else None) for k in ("offsets", "y", "x")}
        h["w"] = pin(w_in) if w_in is not None else None
        ho = {k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in fit.out.items()}
        hs = abi.B200FitOutputs()
        for nm, ctype in abi.B200FitOutputs._fields_:
            setattr(hs, nm, None if nm not in ho else C.cast(ho[nm].data_ptr(), ctype))
        hp, hps = None, None
        if predict:
            hp = {k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in fit.pred.items()}
            hps = abi.B200PredictOutputs()
            for nm, ctype in abi.B200PredictOutputs._fields_:
                setattr(hps, nm, C.cast(hp[nm].data_ptr(), ctype))
        ptr = lambda tt: None if tt is None else tt.data_ptr()  # noqa: E731
        e2e_steps = max(1, min(steps, 5))

        def e2e_step():
            if predict:
                pkg.fit_predict_batch_host_into(ptr(h["offsets"]), ptr(h["y"]), ptr(h["x"]), ptr(h["w"]), None, hps, hs, G, N, p, **opt)
            else:
                pkg.fit_batch_host_into(ptr(h["offsets"]), ptr(h["y"]), ptr(h["x"]), ptr(h["w"]), hs, G, N, p, **opt)

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        dt = (time.perf_counter() - t0) / e2e_steps
        # the copies alone, from the same pinned buffers on one stream: what the PCIe / host-memory path gives this rank
        dev_in = [torch.empty_like(d[k]) for k in ("y", "x")]
        torch.cuda.synchronize()
        c0 = time.perf_counter()
        for _ in range(2):
            dev_in[0].copy_(h["y"], non_blocking=True)
            dev_in[1].copy_(h["x"], non_blocking=True)
        torch.cuda.synchronize()
        h2d_gbs = 2 * (h["y"].numel() + h["x"].numel()) * 8 / (time.perf_counter() - c0) / 1e9
        del dev_in
        tt = torch.tensor([dt, -h2d_gbs], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt[0].item())
        h2d = sum(v.numel() * v.element_size() for v in h.values() if v is not None)
        d2h = sum(v.numel() * v.element_size() for v in ho.values())
        if hp:
            d2h += sum(v.numel() * v.element_size() for v in hp.values())
        same = bool(torch.equal(ho["status"], fit.out["status"].cpu()))
        e2e = {"value": world * G / dt, "unit": "fits/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": dt * 1e3, "steps": e2e_steps, "matches_device_path": same,
               "h2d_only_gbs_slowest_rank": -float(tt[1].item()),
               "call_gbs_per_rank": (h2d + d2h) / dt / 1e9}
        del h, ho, hp

    if rank != 0:
        return None
    peak, peak_src = measured_peak_gbs()
    inp, outb = algorithmic_bytes(G, N, p, method, inference, predict)
    acc_launch_ms = acc_ms / max(calls, 1)
    acc_bytes = N * (8 * (p + 1) + (8 if method == "wls" else 0)) + 8 * (G + 1)
    path, lanes = pkg.plan(G, N, p, method == "wls", "predict" if predict else False)
    if path == "resident":
        acc_bytes += N + outb  # one flag byte per row in; fit structs and 25 bytes of predictions per row out
    if path == "stash":
        acc_bytes += outb  # the one-launch kernel also writes the result structs
    achieved = acc_bytes / (acc_launch_ms * 1e-3) / 1e9 if acc_launch_ms > 0 else None
    if path == "stash":
        kname = "fit_small_stash_kernel<%d> (rows -> results in one launch: TMA-staged rows, %d lanes per group, " \
                "sums parked in shared memory, thread-per-group solve in the same warps)" % (p, lanes)
    elif path == "resident":
        kname = "fit_predict_resident_kernel<%d> (rows -> fit results and predictions in one launch: TMA-staged whole " \
                "groups, FP64 DMMA moments, register solve, predictions from shared memory)" % p
    elif path == "fused":
        kname = "fit_small_kernel<%d> (rows -> moment records, %d lanes per group, TMA-staged)" % (p, lanes)
    else:
        kname = "acc_ring_kernel (streaming ring of TMA windows, FP64 DMMA moment accumulation)"
    tr = load_traffic().get(wname) if args.scale == 1.0 else None
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None,
                "traffic": (tr["dram_read_bytes"] + tr["dram_write_bytes"]) if tr else None,
                "traffic_source": ({k: tr.get(k) for k in ("file", "kernel", "capture_duration_us", "commit")} if tr else None),
                "algorithmic_bytes_per_launch": acc_bytes, "avg_launch_ms": acc_launch_ms,
                "solve_stage_ms": solve_ms / max(calls, 1), "solve_share_of_step": (solve_ms / max(calls, 1)) / ms_per_step,
                "step_algorithmic_gbs": (inp + outb) / (ms_per_step * 1e-3) / 1e9,
                "step_frac_of_peak": (inp + outb) / (ms_per_step * 1e-3) / 1e9 / peak}
    cpu = None
    if headline and not args.no_cpu and world == 1:  # the CPU baseline leg runs at N = 1 only (task contract)
        gs = min(G, 200_000)
        host = {"offsets": d["offsets"][:gs + 1].cpu().numpy(), "y": d["y"][:gs * n].cpu().numpy(),
                "x": d["x"][:gs * n].cpu().numpy(), "w": w_in[:gs * n].cpu().numpy() if w_in is not None else None}
        v, dt, used, threads = cpu_reference_run((gs, n, p, method, inference, extra), host, 3, 1, 0, budget_s=12.0)
        cpu = {"value": v, "unit": "fits/s", "cores": threads, "kind": "port",
               "sample": f"{used} groups x {n} rows per step, 3 steps (oracle f64 build, OpenMP over groups)"}
    return {"value": value, "ms_per_step": ms_per_step, "rows_per_sec": value * n, "groups_fitted_ok": n_ok,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
            "device_counters": counters, "G": G, "n": n, "p": p, "method": method, "inference": inference, "N": N,
            "steps": steps, "cuda_graph_replay_ms": graph_ms}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS) + ["W1"])
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the groups (smoke runs only)")
    ap.add_argument("--path", default="auto", help="A/B runs: force a kernel path (anofox_b200_force_path names)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the other BASELINE configs after the headline")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "C4" and args.impl == "ours":
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        line = run_c4(args, rank, world, local)
        if rank == 0:
            emit(line)
        if world > 1:
            dist.destroy_process_group()
        return
    if args.workload == "W1" and args.impl == "ours":
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        line = run_window_workload(args, rank, world, local, args.steps, args.warmup)
        if rank == 0:
            emit(line)
        if world > 1:
            dist.destroy_process_group()
        return
    G, n, p, method, inference, extra = WORKLOADS[args.workload]
    G = max(64, int(G * args.scale))
    N = G * n
    config = {"workload": NAMES[args.workload], "groups_per_gpu": G, "rows_per_group": n, "features": p, "method": method,
              "inference": inference, "sharding": "group-range per rank, no collective",
              "l2_policy": "inputs (%.2f GB/step) exceed the 126 MB L2" % (N * 8 * (p + 1) / 1e9)}

    sys.path.insert(0, os.path.join(ROOT, "oracle"))

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return
        from importlib import import_module
        from __graft_entry__ import load_package
        load_package()
        synth = import_module("anofox_statistics_b200.synth")
        gs = min(G, 200_000)
        d = synth.make_numpy(gs, n, p, seed=2, weights=extra.get("weights", False),
                             rank_deficient=extra.get("rank_deficient", False))
        wl_s = (gs, n, p, method, inference, extra)
        v, dt, used, threads = cpu_reference_run(wl_s, d, max(args.steps, 1), args.warmup, 0, budget_s=60.0)
        # the arm times a bounded sample of the workload (throughput-normalised): say so in the config
        config = dict(config, groups_timed_per_step=used, sample_of_workload="first %d of %d groups, same generator" % (used, G))
        line = {"impl": "reference", "metric": "fits_per_sec", "value": v, "unit": "fits/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "rows_per_sec": v * n,
                "cpu_baseline": {"value": v, "unit": "fits/s", "cores": threads, "kind": "port",
                                 "sample": f"{used} groups x {n} rows per step (oracle f64, OpenMP over groups; "
                                           "the reference's Rust crates cannot be built here)"},
                "e2e": {"value": v, "unit": "fits/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return

    # ------------------------------------------------------------------ our arm (GPU)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    affinity = bind_to_gpu_numa(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from __graft_entry__ import load_package
    pkg = load_package()
    pkg.device_check()
    pkg.force_path(args.path)
    cache = {}
    r = run_group_workload(args, args.workload, rank, world, local, args.steps, args.warmup, True, cache)
    secondary = {}
    if not args.no_secondary and args.workload == "C2" and world == 1 and "d" in cache:
        nr = run_next_rows(args, rank, cache["d"], r["G"], r["n"], r["p"])
        if rank == 0:
            secondary["next_rows"] = nr
    if not args.no_secondary and args.workload == "C2" and args.scale == 1.0:
        # the other BASELINE configs, a few steps each, so that the driver's BENCH / SCALE records time them too
        for wn in ("C1", "C3r", "C3w", "C5"):
            s = run_group_workload(args, wn, rank, world, local, 5, 3, False, cache)
            if rank == 0:
                secondary[wn] = {"workload": NAMES[wn], "ms_per_step": s["ms_per_step"], "fits_per_sec": s["value"],
                                 "rows_per_sec": s["rows_per_sec"], "steps": s["steps"],
                                 "accumulate_kernel_ms": s["roofline"]["avg_launch_ms"],
                                 "accumulate_frac_of_peak": s["roofline"]["frac"],
                                 "solve_share_of_step": s["roofline"]["solve_share_of_step"],
                                 "step_frac_of_peak": s["roofline"]["step_frac_of_peak"],
                                 "gpu_launches": s["gpu_launches"], "device_counters": s["device_counters"],
                                 "groups_fitted_ok": s["groups_fitted_ok"], "clocks": s["clocks"]}
                if s.get("cuda_graph_replay_ms") is not None:
                    secondary[wn]["cuda_graph_replay_ms"] = s["cuda_graph_replay_ms"]
        cache.clear()
        torch.cuda.empty_cache()
        wres = run_window_workload(args, rank, world, local, 3, 3)
        if rank == 0:
            secondary["W1"] = wres
        if world == 1:
            secondary["finalize_pattern"] = run_finalize_pattern(args, rank)
            # CPU-A: the reference's own DuckDB callers (oracle/_ref, built by oracle/build_ref.sh) + the CPU oracle behind
            # the seven FFI symbols, C1's query on the host cores; absent when oracle/_ref was not built
            if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "duckdb")) and not args.no_cpu:
                try:
                    out = subprocess.run([sys.executable, os.path.join(ROOT, "harness", "time_c1.py"), "--flavour", "cpu"],
                                         capture_output=True, text=True, timeout=300).stdout
                    a = json.loads(out)
                    secondary["cpu_baseline_sql_c1"] = {
                        "kind": "reference-callers+port", "value": a["fits_per_sec_best"], "unit": "fits/s", "cores": a["threads"],
                        "seconds_no_inference": a["seconds_no_inference"], "seconds_with_inference": a["seconds_with_inference"],
                        "what": "DuckDB (built from the reference tree) + the reference's unmodified ols_aggregate.cpp + CPU oracle "
                                "behind anofox_ols_fit: " + a["query"]}
                except Exception as e:  # noqa: BLE001
                    secondary["cpu_baseline_sql_c1"] = {"unavailable": str(e)[:200]}
        if world > 1:
            # C4: the single huge group, 125 M rows x 64 features per GPU, three NCCL collectives per fit
            a4 = argparse.Namespace(**vars(args))
            a4.steps, a4.warmup = 5, 3
            c4 = run_c4(a4, rank, world, local)
            if rank == 0:
                secondary["C4"] = {k: c4[k] for k in ("value", "unit", "ms_per_step", "identical_on_all_ranks",
                                                      "n_observations", "roofline", "config", "gpu_launches")}
    if rank == 0:
        line = {"metric": "fits_per_sec", "value": r["value"], "unit": "fits/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "rows_per_sec": r["rows_per_sec"], "groups_fitted_ok": r["groups_fitted_ok"], "clocks": r["clocks"],
                "e2e": r["e2e"], "gpu_launches": r["gpu_launches"], "roofline": r["roofline"],
                "cpu_baseline": r["cpu_baseline"], "device_counters": r["device_counters"],
                "host_affinity": affinity, "secondary": secondary or None}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
