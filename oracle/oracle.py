"""ctypes front end of the CPU oracle -- TEST INFRASTRUCTURE (see groupfit_oracle.cpp header).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs. It borrows the POD layouts from the product's _abi.py (type declarations only).
"""
import ctypes as C
import importlib.util
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)

_spec = importlib.util.spec_from_file_location("_anofox_b200_abi",
                                               os.path.join(_ROOT, "anofox-statistics_b200", "_abi.py"))
abi = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(abi)

_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "libgroupfit_oracle.so")
    src = os.path.join(_HERE, "groupfit_oracle.cpp")
    hdr = os.path.join(_ROOT, "include", "anofox_b200.h")
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libgroupfit_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        abi.declare_reference_symbols(_LIB, "oracle_")
        _LIB.oracle_t_two_sided_p.restype = C.c_double
        _LIB.oracle_t_two_sided_p.argtypes = [C.c_double, C.c_double]
        _LIB.oracle_f_upper_p.restype = C.c_double
        _LIB.oracle_f_upper_p.argtypes = [C.c_double, C.c_double, C.c_double]
        _LIB.oracle_fit_batch.restype = C.c_int32
        _LIB.oracle_fit_batch.argtypes = [abi.c_int64_p, C.c_int64, abi.c_double_p, abi.c_double_p, abi.c_double_p,
                                          C.c_int64, C.c_int32, C.POINTER(abi.B200FitOptions),
                                          C.POINTER(abi.B200FitOutputs), C.c_int32, C.c_int32]
        _LIB.oracle_fit_predict_batch.restype = C.c_int32
        _LIB.oracle_fit_predict_batch.argtypes = [abi.c_int64_p, C.c_int64, abi.c_double_p, abi.c_double_p,
                                                  abi.c_double_p, abi.c_uint8_p, C.c_int64, C.c_int32,
                                                  C.POINTER(abi.B200FitOptions), C.POINTER(abi.B200PredictOutputs),
                                                  C.POINTER(abi.B200FitOutputs), C.c_int32, C.c_int32]
        _LIB.oracle_window_fit_predict.restype = C.c_int32
        _LIB.oracle_window_fit_predict.argtypes = [abi.c_int64_p, C.c_int64, abi.c_double_p, abi.c_double_p,
                                                   abi.c_double_p, C.c_int64, C.c_int32,
                                                   C.POINTER(abi.B200FitOptions), C.c_int64, C.c_int64,
                                                   C.POINTER(abi.B200PredictOutputs), C.c_int32, C.c_int32]
        _LIB.oracle_max_threads.restype = C.c_int32
    return _LIB


def single_fit(method, y, x_cols, w=None, **kw):
    return abi.call_single_fit(lib(), "oracle_", method, y, x_cols, w, **kw)


def t_critical(level, df):
    return lib().oracle_t_critical(level, df)


def t_two_sided_p(t, df):
    return lib().oracle_t_two_sided_p(t, df)


def f_upper_p(f, d1, d2):
    return lib().oracle_f_upper_p(f, d1, d2)


def predict_with_interval(coef, intercept, x_new, rse, n_obs, level):
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    x_new = np.ascontiguousarray(x_new, dtype=np.float64)
    out = abi.PredictionResult()
    ok = lib().oracle_predict_with_interval(abi.np_ptr(coef, C.c_double), len(coef), intercept,
                                            abi.np_ptr(x_new, C.c_double), len(x_new), rse, n_obs, level,
                                            C.byref(out))
    return bool(ok), (out.yhat, out.yhat_lower, out.yhat_upper)


def _prep(offsets, y, x, w):
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    w = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
    N = y.shape[0]
    p = x.shape[1] if x.ndim == 2 else x.size // max(N, 1)
    return offsets, y, x, w, N, p


def fit_batch(offsets, y, x, w=None, precision=0, threads=0, extras=True, **opt):
    """Oracle fits over a group-sorted CSR (host numpy). Returns dict of numpy arrays."""
    offsets, y, x, w, N, p = _prep(offsets, y, x, w)
    G = len(offsets) - 1
    o = abi.make_fit_options(**opt)
    arrs, s = abi.alloc_fit_outputs_numpy(G, p, bool(o.compute_inference), extras)
    rc = lib().oracle_fit_batch(abi.np_ptr(offsets, C.c_int64), G, abi.np_ptr(y, C.c_double),
                                abi.np_ptr(x, C.c_double), abi.np_ptr(w, C.c_double), N, p, C.byref(o), C.byref(s),
                                precision, threads)
    arrs["rc"] = rc
    return arrs


def fit_predict_batch(offsets, y, x, w=None, train=None, precision=0, threads=0, **opt):
    offsets, y, x, w, N, p = _prep(offsets, y, x, w)
    G = len(offsets) - 1
    o = abi.make_fit_options(**opt)
    fit, fs = abi.alloc_fit_outputs_numpy(G, p, False, True)
    pred, ps = abi.alloc_predict_outputs_numpy(N)
    tr = None if train is None else np.ascontiguousarray(train, dtype=np.uint8)
    rc = lib().oracle_fit_predict_batch(abi.np_ptr(offsets, C.c_int64), G, abi.np_ptr(y, C.c_double),
                                        abi.np_ptr(x, C.c_double), abi.np_ptr(w, C.c_double),
                                        abi.np_ptr(tr, C.c_uint8), N, p, C.byref(o), C.byref(ps), C.byref(fs),
                                        precision, threads)
    pred["rc"] = rc
    return pred, fit


def window_fit_predict(offsets, y, x, w=None, preceding=-1, end_offset=1, precision=0, threads=0, **opt):
    """One raw-row fit per output row over its frame (src/window_functions/ols_fit_predict.cpp)."""
    offsets, y, x, w, N, p = _prep(offsets, y, x, w)
    G = len(offsets) - 1
    o = abi.make_fit_options(**opt)
    pred, ps = abi.alloc_predict_outputs_numpy(N)
    rc = lib().oracle_window_fit_predict(abi.np_ptr(offsets, C.c_int64), G, abi.np_ptr(y, C.c_double),
                                         abi.np_ptr(x, C.c_double), abi.np_ptr(w, C.c_double), N, p, C.byref(o),
                                         int(preceding), int(end_offset), C.byref(ps), precision, threads)
    pred["rc"] = rc
    pred.pop("is_training", None)
    return pred


def max_threads():
    return lib().oracle_max_threads()
