"""anofox-statistics_b200 -- B200-native per-group OLS / Ridge / WLS fit aggregates.

Thin Python binding over libanofox_stats_b200.so (C ABI: include/anofox_b200.h). All numerics run
in the library's CUDA kernels; this module only marshals pointers. There is no CPU fallback: if the
shared library is missing or no sm_100 device is present, calls raise.
"""
import ctypes as C
import os

import numpy as np

from . import _abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ANOFOX_B200_LIB") or os.path.join(_HERE, "libanofox_stats_b200.so")  # env: A/B builds
_LIB = None

# every symbol include/anofox_b200.h declares
HEADER_SYMBOLS = [
    "anofox_ols_fit", "anofox_ridge_fit", "anofox_wls_fit", "anofox_free_result_core",
    "anofox_free_result_inference", "anofox_t_critical", "anofox_predict_with_interval", "anofox_compute_aic",
    "anofox_compute_bic", "anofox_b200_workspace_bytes", "anofox_b200_fit_batch_device",
    "anofox_b200_fit_batch_host", "anofox_b200_fit_predict_batch_device", "anofox_b200_fit_predict_batch_host",
    "anofox_b200_moment_block_len", "anofox_b200_first_valid_row_device", "anofox_b200_moments_device", "anofox_b200_solve_moments_device",
    "anofox_b200_solve_moments_refine_device", "anofox_b200_shard_residual_device",
    "anofox_b200_device_check", "anofox_b200_launch_count", "anofox_b200_version",
]


class B200Error(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"anofox_b200 error {code}: {message}")
        self.code = code
        self.message = message


def load_library():
    """dlopen the CUDA library (no GPU needed for this) and declare its signatures."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is missing: run `make -C anofox-statistics_b200/csrc` "
                                "(or __graft_entry__.build()). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for s in HEADER_SYMBOLS:
        getattr(lib, s)  # AttributeError if a declared symbol is not exported
    abi.declare_reference_symbols(lib, "anofox_")
    E = C.POINTER(abi.AnofoxError)
    lib.anofox_b200_workspace_bytes.restype = C.c_size_t
    lib.anofox_b200_workspace_bytes.argtypes = [C.c_int64, C.c_int64, C.c_int32]
    common = [abi.c_int64_p, C.c_int64, abi.c_double_p, abi.c_double_p, abi.c_double_p]
    lib.anofox_b200_fit_batch_device.restype = C.c_int32
    lib.anofox_b200_fit_batch_device.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                                 C.c_int32, C.POINTER(abi.B200FitOptions),
                                                 C.POINTER(abi.B200FitOutputs), C.c_void_p, C.c_size_t, C.c_void_p, E]
    lib.anofox_b200_fit_batch_host.restype = C.c_int32
    lib.anofox_b200_fit_batch_host.argtypes = common + [C.c_int64, C.c_int32, C.POINTER(abi.B200FitOptions),
                                                        C.POINTER(abi.B200FitOutputs), E]
    lib.anofox_b200_fit_predict_batch_device.restype = C.c_int32
    lib.anofox_b200_fit_predict_batch_device.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                                         C.c_void_p, C.c_int64, C.c_int32,
                                                         C.POINTER(abi.B200FitOptions),
                                                         C.POINTER(abi.B200PredictOutputs),
                                                         C.POINTER(abi.B200FitOutputs), C.c_void_p, C.c_size_t,
                                                         C.c_void_p, E]
    lib.anofox_b200_fit_predict_batch_host.restype = C.c_int32
    lib.anofox_b200_fit_predict_batch_host.argtypes = common + [abi.c_uint8_p, C.c_int64, C.c_int32,
                                                                C.POINTER(abi.B200FitOptions),
                                                                C.POINTER(abi.B200PredictOutputs),
                                                                C.POINTER(abi.B200FitOutputs), E]
    lib.anofox_b200_ingest_workspace_bytes.restype = C.c_size_t
    lib.anofox_b200_ingest_workspace_bytes.argtypes = [C.c_int64, C.c_int64]
    lib.anofox_b200_ingest_device.restype = C.c_int32
    lib.anofox_b200_ingest_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_size_t, C.c_void_p, E]
    lib.anofox_b200_ingest_device_static.restype = C.c_int32
    lib.anofox_b200_ingest_device_static.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_int32, C.c_int64] + [C.c_void_p] * 7 + [E]
    lib.anofox_b200_ingest_bad_ids.restype = C.c_int32
    lib.anofox_b200_ingest_bad_ids.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
    lib.anofox_b200_fit_columns_workspace_bytes.restype = C.c_size_t
    lib.anofox_b200_fit_columns_workspace_bytes.argtypes = [C.c_int64, C.c_int64, C.c_int32]
    lib.anofox_b200_fit_columns_batch_device.restype = C.c_int32
    lib.anofox_b200_fit_columns_batch_device.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                                         C.c_int64, C.c_int32, C.POINTER(abi.B200FitOptions),
                                                         C.POINTER(abi.B200FitOutputs), C.c_void_p, C.c_size_t,
                                                         C.c_void_p, E]
    lib.anofox_b200_window_workspace_bytes.restype = C.c_size_t
    lib.anofox_b200_window_workspace_bytes.argtypes = [C.c_int64, C.c_int32]
    lib.anofox_b200_window_fit_predict_host.restype = C.c_int32
    lib.anofox_b200_window_fit_predict_host.argtypes = common + [C.c_int64, C.c_int32, C.POINTER(abi.B200FitOptions),
                                                                 C.c_int64, C.c_int64,
                                                                 C.POINTER(abi.B200PredictOutputs), E]
    lib.anofox_b200_window_fit_predict_device.restype = C.c_int32
    lib.anofox_b200_window_fit_predict_device.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                                          C.c_int64, C.c_int32, C.POINTER(abi.B200FitOptions),
                                                          C.c_int64, C.c_int64, C.POINTER(abi.B200PredictOutputs),
                                                          C.c_void_p, C.c_size_t, C.c_void_p, E]
    lib.anofox_b200_moment_block_len.restype = C.c_size_t
    lib.anofox_b200_moment_block_len.argtypes = [C.c_int32]
    lib.anofox_b200_first_valid_row_device.restype = C.c_int32
    lib.anofox_b200_first_valid_row_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, E]
    lib.anofox_b200_moments_device.restype = C.c_int32
    lib.anofox_b200_moments_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, E]
    lib.anofox_b200_solve_moments_device.restype = C.c_int32
    lib.anofox_b200_solve_moments_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                                     C.POINTER(abi.B200FitOptions), C.POINTER(abi.B200FitOutputs),
                                                     C.c_void_p, E]
    lib.anofox_b200_solve_moments_refine_device.restype = C.c_int32
    lib.anofox_b200_solve_moments_refine_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                                            C.POINTER(abi.B200FitOptions),
                                                            C.POINTER(abi.B200FitOutputs), C.c_void_p, C.c_void_p,
                                                            C.c_void_p, E]
    lib.anofox_b200_shard_residual_device.restype = C.c_int32
    lib.anofox_b200_shard_residual_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                      C.c_void_p, C.c_size_t, C.c_void_p, E]
    lib.anofox_b200_device_check.restype = C.c_int32
    lib.anofox_b200_device_check.argtypes = [E]
    lib.anofox_b200_launch_count.restype = C.c_int64
    lib.anofox_b200_version.restype = C.c_char_p
    lib.anofox_b200_eval_specfun.restype = C.c_int32
    lib.anofox_b200_eval_specfun.argtypes = [C.c_int32, C.c_int32, abi.c_double_p, abi.c_double_p, abi.c_double_p,
                                             abi.c_double_p, E]
    lib.anofox_b200_stage_timing_enable.restype = None
    lib.anofox_b200_stage_timing_enable.argtypes = [C.c_int32]
    lib.anofox_b200_stage_timing_read.restype = C.c_int64
    lib.anofox_b200_stage_timing_read.argtypes = [abi.c_double_p, abi.c_double_p]
    lib.anofox_b200_force_path.restype = None
    lib.anofox_b200_force_path.argtypes = [C.c_int32]
    lib.anofox_b200_debug_counters.restype = C.c_int32
    lib.anofox_b200_debug_counters.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, abi.c_int32_p]
    lib.anofox_b200_plan.restype = C.c_int32
    lib.anofox_b200_plan.argtypes = [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, abi.c_int32_p]
    _LIB = lib
    return lib


def _check(rc, err):
    if rc != 0:
        raise B200Error(rc, err.message.decode(errors="replace"))


def device_check():
    err = abi.AnofoxError()
    _check(load_library().anofox_b200_device_check(C.byref(err)), err)


def launch_count():
    return int(load_library().anofox_b200_launch_count())


# ---------------------------------------------------------------------------------------------
# drop-in symbols (one group, column arrays) -- what ols_aggregate.cpp:270-296 calls
# ---------------------------------------------------------------------------------------------
def single_fit(method, y, x_cols, w=None, **kw):
    return abi.call_single_fit(load_library(), "anofox_", method, y, x_cols, w, **kw)


def t_critical(level, df):
    return load_library().anofox_t_critical(level, df)


def predict_with_interval(coef, intercept, x_new, rse, n_obs, level):
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    x_new = np.ascontiguousarray(x_new, dtype=np.float64)
    out = abi.PredictionResult()
    ok = load_library().anofox_predict_with_interval(abi.np_ptr(coef, C.c_double), len(coef), intercept,
                                                     abi.np_ptr(x_new, C.c_double), len(x_new), rse, n_obs, level,
                                                     C.byref(out))
    return bool(ok), (out.yhat, out.yhat_lower, out.yhat_upper)


def eval_specfun(mode, a, b, c=None):
    """Device special functions: mode 0 two-sided t p-value(t, df), 1 t-critical(level, df), 2 F upper tail."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    c = None if c is None else np.ascontiguousarray(c, dtype=np.float64)
    out = np.empty_like(a)
    err = abi.AnofoxError()
    _check(load_library().anofox_b200_eval_specfun(mode, a.size, abi.np_ptr(a, C.c_double), abi.np_ptr(b, C.c_double),
                                                   abi.np_ptr(c, C.c_double), abi.np_ptr(out, C.c_double),
                                                   C.byref(err)), err)
    return out


# ---------------------------------------------------------------------------------------------
# batch layer, host pointers (numpy)
# ---------------------------------------------------------------------------------------------
def _prep(offsets, y, x, w):
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    w = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
    N = y.shape[0]
    p = x.shape[1] if x.ndim == 2 else x.size // max(N, 1)
    return offsets, y, x, w, N, p


def fit_batch_host(offsets, y, x, w=None, extras=True, **opt):
    """ols/ridge/wls_fit_agg over a group-sorted CSR held in host memory -> dict of numpy arrays."""
    offsets, y, x, w, N, p = _prep(offsets, y, x, w)
    G = len(offsets) - 1
    o = abi.make_fit_options(**opt)
    arrs, s = abi.alloc_fit_outputs_numpy(G, p, bool(o.compute_inference), extras)
    err = abi.AnofoxError()
    rc = load_library().anofox_b200_fit_batch_host(abi.np_ptr(offsets, C.c_int64), G, abi.np_ptr(y, C.c_double),
                                                   abi.np_ptr(x, C.c_double), abi.np_ptr(w, C.c_double), N, p,
                                                   C.byref(o), C.byref(s), C.byref(err))
    _check(rc, err)
    return arrs


def fit_predict_batch_host(offsets, y, x, w=None, train=None, **opt):
    offsets, y, x, w, N, p = _prep(offsets, y, x, w)
    G = len(offsets) - 1
    o = abi.make_fit_options(**opt)
    fit, fs = abi.alloc_fit_outputs_numpy(G, p, False, True)
    pred, ps = abi.alloc_predict_outputs_numpy(N)
    tr = None if train is None else np.ascontiguousarray(train, dtype=np.uint8)
    err = abi.AnofoxError()
    rc = load_library().anofox_b200_fit_predict_batch_host(abi.np_ptr(offsets, C.c_int64), G,
                                                           abi.np_ptr(y, C.c_double), abi.np_ptr(x, C.c_double),
                                                           abi.np_ptr(w, C.c_double), abi.np_ptr(tr, C.c_uint8), N, p,
                                                           C.byref(o), C.byref(ps), C.byref(fs), C.byref(err))
    _check(rc, err)
    return pred, fit


def window_fit_predict_host(offsets, y, x, w=None, preceding=-1, end_offset=1, **opt):
    """{ols,ridge,wls}_fit_predict(...) OVER (PARTITION BY g ORDER BY t ROWS BETWEEN <preceding> AND <end>) for
    every row of every partition (src/window_functions/ols_fit_predict.cpp). preceding < 0 = UNBOUNDED PRECEDING;
    end_offset 0 = CURRENT ROW, 1 = 1 PRECEDING. Returns dict(yhat, yhat_lower, yhat_upper); NaN = NULL."""
    offsets, y, x, w, N, p = _prep(offsets, y, x, w)
    G = len(offsets) - 1
    o = abi.make_fit_options(**opt)
    pred, ps = abi.alloc_predict_outputs_numpy(N)
    err = abi.AnofoxError()
    rc = load_library().anofox_b200_window_fit_predict_host(abi.np_ptr(offsets, C.c_int64), G,
                                                            abi.np_ptr(y, C.c_double), abi.np_ptr(x, C.c_double),
                                                            abi.np_ptr(w, C.c_double), N, p, C.byref(o),
                                                            int(preceding), int(end_offset), C.byref(ps), C.byref(err))
    _check(rc, err)
    pred.pop("is_training", None)
    return pred


class DeviceIngest:
    """Unsorted (group_id, y, x[, w]) rows -> group-sorted CSR on the device (include/anofox_b200.h, csrc/ingest.cuh).
    torch owns the buffers; the partition itself is the library's radix kernels."""

    def __init__(self, N, p, G, weights=False, device="cuda"):
        import torch
        self.torch, self.N, self.p, self.G = torch, N, p, G
        dev = torch.device(device)
        self.offsets = torch.empty(G + 1, dtype=torch.int64, device=dev)
        self.y = torch.empty(N, dtype=torch.float64, device=dev)
        self.x = torch.empty((N, p), dtype=torch.float64, device=dev)
        self.w = torch.empty(N, dtype=torch.float64, device=dev) if weights else None
        self.row_index = torch.empty(N, dtype=torch.int64, device=dev)
        self.ws_bytes = load_library().anofox_b200_ingest_workspace_bytes(N, G)
        self.ws = torch.empty(self.ws_bytes + 256, dtype=torch.uint8, device=dev)
        self.ws_ptr = (self.ws.data_ptr() + 255) & ~255

    def run(self, group_ids, y, x, w=None, stream=None):
        torch = self.torch
        err = abi.AnofoxError()
        st = torch.cuda.current_stream().cuda_stream if stream is None else stream
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
        rc = load_library().anofox_b200_ingest_device(p(group_ids), p(y), p(x), p(w), self.N, self.p, self.G, p(self.offsets),
                                                      p(self.y), p(self.x), p(self.w) if w is not None else None,
                                                      p(self.row_index), C.c_void_p(self.ws_ptr), self.ws_bytes,
                                                      C.c_void_p(st), C.byref(err))
        _check(rc, err)
        return self

    def bad_ids(self):
        return int(load_library().anofox_b200_ingest_bad_ids(C.c_void_p(self.ws_ptr), self.N, self.G))


INGEST_STATIC_MAX_BYTES = 16 * 1024 * 1024  # ANOFOX_B200_INGEST_STATIC_MAX_BYTES (include/anofox_b200.h)


def ingest_device_static(group_ids, y, x, w, N, p, G, offsets_out, y_out, x_out, w_out, row_index_out, bad_ids_out, stream=0):
    """anofox_b200_ingest_device_static: the ingest with the library's built-in scratch array (no allocation).
    Every argument is a raw device address (or None); raises B200Error beyond the array's size."""
    err = abi.AnofoxError()
    v = lambda a: None if a is None else C.c_void_p(a)  # noqa: E731
    rc = load_library().anofox_b200_ingest_device_static(v(group_ids), v(y), v(x), v(w), N, p, G, v(offsets_out), v(y_out),
                                                         v(x_out), v(w_out), v(row_index_out), v(bad_ids_out),
                                                         C.c_void_p(stream), C.byref(err))
    _check(rc, err)


def window_workspace_bytes(N, p):
    return int(load_library().anofox_b200_window_workspace_bytes(N, p))


def window_fit_predict_device(offsets, y, x, w, N, G, p, pred_struct, workspace, workspace_bytes, stream=0,
                              preceding=-1, end_offset=1, **opt):
    """Device-pointer form (raw addresses; torch owns the memory)."""
    o = abi.make_fit_options(**opt)
    err = abi.AnofoxError()
    rc = load_library().anofox_b200_window_fit_predict_device(offsets, G, y, x, w, N, p, C.byref(o), int(preceding),
                                                              int(end_offset), C.byref(pred_struct), workspace,
                                                              workspace_bytes, stream, C.byref(err))
    _check(rc, err)


# ---------------------------------------------------------------------------------------------
# batch layer, device pointers (torch tensors own the memory; torch is plumbing only)
# ---------------------------------------------------------------------------------------------
class DeviceFit:
    """Preallocated device outputs + workspace for repeated fits of one shape (bench hot loop)."""

    def __init__(self, G, N, p, inference, device="cuda", extras=True, predict=False):
        import torch
        self.torch = torch
        self.G, self.N, self.p = G, N, p
        dev = torch.device(device)
        f64 = dict(dtype=torch.float64, device=dev)
        t = {}
        t["coefficients"] = torch.empty((G, p), **f64)
        for k in abi.FIT_F64_G:
            t[k] = torch.empty(G, **f64)
        t["n_observations"] = torch.empty(G, dtype=torch.int64, device=dev)
        t["status"] = torch.empty(G, dtype=torch.int32, device=dev)
        if inference:
            for k in abi.INF_F64_GP:
                t[k] = torch.empty((G, p), **f64)
            for k in abi.INF_F64_G:
                t[k] = torch.empty(G, **f64)
            t["has_inference"] = torch.empty(G, dtype=torch.uint8, device=dev)
        if extras:
            for k in abi.EXTRA_F64_G:
                t[k] = torch.empty(G, **f64)
        self.out = t
        self.out_struct = abi.B200FitOutputs()
        for name, ctype in abi.B200FitOutputs._fields_:
            v = t.get(name)
            setattr(self.out_struct, name, None if v is None else C.cast(v.data_ptr(), ctype))
        self.pred = None
        if predict:
            self.pred = {"yhat": torch.empty(N, **f64), "yhat_lower": torch.empty(N, **f64),
                         "yhat_upper": torch.empty(N, **f64),
                         "is_training": torch.empty(N, dtype=torch.uint8, device=dev)}
            self.pred_struct = abi.B200PredictOutputs()
            for name, ctype in abi.B200PredictOutputs._fields_:
                setattr(self.pred_struct, name, C.cast(self.pred[name].data_ptr(), ctype))
        wb = load_library().anofox_b200_workspace_bytes(G, N, p)
        self.workspace = torch.empty(wb + 256, dtype=torch.uint8, device=dev)
        self.ws_ptr = (self.workspace.data_ptr() + 255) & ~255
        self.ws_bytes = wb

    def _ptr(self, t):
        return None if t is None else C.c_void_p(t.data_ptr())

    def fit(self, offsets, y, x, w=None, stream=None, **opt):
        torch = self.torch
        o = abi.make_fit_options(**opt)
        err = abi.AnofoxError()
        st = torch.cuda.current_stream().cuda_stream if stream is None else stream
        rc = load_library().anofox_b200_fit_batch_device(self._ptr(offsets), self.G, self._ptr(y), self._ptr(x),
                                                         self._ptr(w), self.N, self.p, C.byref(o),
                                                         C.byref(self.out_struct), C.c_void_p(self.ws_ptr),
                                                         self.ws_bytes, C.c_void_p(st), C.byref(err))
        _check(rc, err)
        return self.out

    def fit_columns(self, offsets, y, x_columns, w=None, stream=None, **opt):
        """Batched scalar array fits ({ols,ridge,wls}_fit(y LIST, x LIST(LIST))): x_columns holds, per fit, p columns
        of n_i values back to back (include/anofox_b200.h). G = number of fits."""
        torch = self.torch
        lib = load_library()
        need = lib.anofox_b200_fit_columns_workspace_bytes(self.G, self.N, self.p)
        if getattr(self, "_cws_bytes", 0) < need:
            self._cws = torch.empty(need + 256, dtype=torch.uint8, device=y.device)
            self._cws_ptr = (self._cws.data_ptr() + 255) & ~255
            self._cws_bytes = need
        o = abi.make_fit_options(**opt)
        err = abi.AnofoxError()
        st = torch.cuda.current_stream().cuda_stream if stream is None else stream
        rc = lib.anofox_b200_fit_columns_batch_device(self._ptr(offsets), self.G, self._ptr(y), self._ptr(x_columns),
                                                      self._ptr(w), self.N, self.p, C.byref(o), C.byref(self.out_struct),
                                                      C.c_void_p(self._cws_ptr), self._cws_bytes, C.c_void_p(st),
                                                      C.byref(err))
        _check(rc, err)
        return self.out

    def counters(self):
        """Device-side counters of the last call: big groups, exact-residual refinements, checked redos, huge
        groups, QR-fallback groups (synchronises)."""
        out = (C.c_int32 * 5)()
        rc = load_library().anofox_b200_debug_counters(C.c_void_p(self.ws_ptr), self.G, self.N, self.p, out)
        if rc != 0:
            raise B200Error(rc, "debug_counters")
        return dict(zip(("big", "refined", "redone", "huge", "qr"), [int(v) for v in out]))

    def debug_raw(self):
        """Experiments: the 192-int counter block of the last call as uint64 pairs (synchronises)."""
        import numpy as np
        out = (C.c_int32 * 192)()
        fn = load_library().anofox_b200_debug_raw
        fn.restype = C.c_int32
        rc = fn(C.c_void_p(self.ws_ptr), C.c_int64(self.G), C.c_int64(self.N), C.c_int32(self.p), out)
        if rc != 0:
            raise B200Error(rc, "debug_raw")
        return np.frombuffer(bytes(out), dtype=np.uint64)

    def fit_predict(self, offsets, y, x, w=None, train=None, stream=None, **opt):
        torch = self.torch
        o = abi.make_fit_options(**opt)
        err = abi.AnofoxError()
        st = torch.cuda.current_stream().cuda_stream if stream is None else stream
        rc = load_library().anofox_b200_fit_predict_batch_device(
            self._ptr(offsets), self.G, self._ptr(y), self._ptr(x), self._ptr(w), self._ptr(train), self.N, self.p,
            C.byref(o), C.byref(self.pred_struct), C.byref(self.out_struct), C.c_void_p(self.ws_ptr), self.ws_bytes,
            C.c_void_p(st), C.byref(err))
        _check(rc, err)
        return self.pred, self.out


def stage_timing(enable):
    load_library().anofox_b200_stage_timing_enable(1 if enable else 0)


def stage_timing_read():
    """(calls, accumulate_ms_total, solve_ms_total) since stage_timing(True); device-timed with CUDA events."""
    a, b = C.c_double(), C.c_double()
    n = load_library().anofox_b200_stage_timing_read(C.byref(a), C.byref(b))
    return int(n), a.value, b.value


def fit_batch_host_into(offsets, y, x, w, out_struct, G, N, p, **opt):
    """Host-pointer batch call with caller-owned (e.g. pinned) buffers given as raw addresses."""
    o = abi.make_fit_options(**opt)
    err = abi.AnofoxError()
    lib = load_library()
    cast = lambda a, t: None if a is None else C.cast(a, t)  # noqa: E731
    rc = lib.anofox_b200_fit_batch_host(cast(offsets, abi.c_int64_p), G, cast(y, abi.c_double_p),
                                        cast(x, abi.c_double_p), cast(w, abi.c_double_p), N, p, C.byref(o),
                                        C.byref(out_struct), C.byref(err))
    _check(rc, err)


def fit_predict_batch_host_into(offsets, y, x, w, train, pred_struct, fit_struct, G, N, p, **opt):
    """Host-pointer fit + predict call with caller-owned (e.g. pinned) buffers given as raw addresses."""
    o = abi.make_fit_options(**opt)
    err = abi.AnofoxError()
    lib = load_library()
    cast = lambda a, t: None if a is None else C.cast(a, t)  # noqa: E731
    rc = lib.anofox_b200_fit_predict_batch_host(cast(offsets, abi.c_int64_p), G, cast(y, abi.c_double_p),
                                                cast(x, abi.c_double_p), cast(w, abi.c_double_p),
                                                cast(train, abi.c_uint8_p), N, p, C.byref(o), C.byref(pred_struct),
                                                C.byref(fit_struct), C.byref(err))
    _check(rc, err)


def force_path(path):
    """Test / A-B hook: "auto", "record" (always the DMMA accumulation kernel + solve kernel) or "fused"
    (small groups through the single rows-to-results kernel instead of accumulate + register solve)."""
    load_library().anofox_b200_force_path({"auto": 0, "record": 1, "fused": 2, "stages": 3, "ringslice": 4, "qr": 5, "stash": 6, "split": 7, "resident": 8}[path])


def plan(G, N, p, has_w=False, has_train=False):
    """("stash" | "fused", lanes_per_group) or ("record", 0): the kernel path a batch of this shape takes.
    "stash" = the one-launch rows-to-results kernel, "fused" = accumulate-to-record kernel + register solve kernel."""
    t = C.c_int32(0)
    r = load_library().anofox_b200_plan(G, N, p, int(has_w), 2 if has_train == "predict" else int(bool(has_train)), C.byref(t))
    return {0: "record", 1: "fused", 2: "stash", 3: "resident"}[r], int(t.value)


def moment_block_len(p):
    return int(load_library().anofox_b200_moment_block_len(p))
