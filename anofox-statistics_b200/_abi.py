"""ctypes mirror of include/anofox_b200.h (POD layouts only -- no compute, no library loading)."""
import ctypes as C

import numpy as np

c_double_p = C.POINTER(C.c_double)
c_int64_p = C.POINTER(C.c_int64)
c_int32_p = C.POINTER(C.c_int32)
c_uint8_p = C.POINTER(C.c_uint8)

# AnofoxErrorCode
SUCCESS, INVALID_INPUT, SINGULAR_MATRIX, CONVERGENCE_FAILURE, INVALID_ALPHA = 0, 1, 2, 3, 4
INVALID_L1_RATIO, INSUFFICIENT_DATA, ALLOCATION_FAILURE, SERIALIZATION_ERROR = 5, 6, 7, 8
DIMENSION_MISMATCH, NO_VALID_DATA, INTERNAL = 9, 10, 99
STATUS_OK, STATUS_TOO_FEW_ROWS = 0, 1000

METHOD = {"ols": 0, "ridge": 1, "wls": 2}
SOLVER = {"qr": 0, "svd": 1, "cholesky": 2}
LAMBDA_SCALING = {"raw": 0, "glmnet": 1}
HC = {"none": 0, "hc0": 1, "hc1": 2, "hc2": 3, "hc3": 4}


class AnofoxError(C.Structure):
    _fields_ = [("code", C.c_int), ("message", C.c_char * 256)]


class DataArray(C.Structure):
    _fields_ = [("data", c_double_p), ("validity", c_uint8_p), ("len", C.c_size_t)]


class FitResultCore(C.Structure):
    _fields_ = [("coefficients", c_double_p), ("coefficients_len", C.c_size_t), ("intercept", C.c_double),
                ("r_squared", C.c_double), ("adj_r_squared", C.c_double), ("residual_std_error", C.c_double),
                ("n_observations", C.c_size_t), ("n_features", C.c_size_t)]


class FitResultInference(C.Structure):
    _fields_ = [("std_errors", c_double_p), ("t_values", c_double_p), ("p_values", c_double_p),
                ("ci_lower", c_double_p), ("ci_upper", c_double_p), ("len", C.c_size_t),
                ("confidence_level", C.c_double), ("f_statistic", C.c_double), ("f_pvalue", C.c_double)]


class OlsOptions(C.Structure):  # also AnofoxWlsOptions (same layout)
    _fields_ = [("fit_intercept", C.c_bool), ("compute_inference", C.c_bool), ("confidence_level", C.c_double),
                ("solver", C.c_int), ("hc_type", C.c_int)]


class RidgeOptions(C.Structure):
    _fields_ = [("alpha", C.c_double), ("fit_intercept", C.c_bool), ("compute_inference", C.c_bool),
                ("confidence_level", C.c_double), ("solver", C.c_int), ("lambda_scaling", C.c_int)]


class PredictionResult(C.Structure):
    _fields_ = [("yhat", C.c_double), ("yhat_lower", C.c_double), ("yhat_upper", C.c_double)]


class B200FitOptions(C.Structure):
    _fields_ = [("method", C.c_int32), ("fit_intercept", C.c_int32), ("compute_inference", C.c_int32),
                ("solver", C.c_int32), ("lambda_scaling", C.c_int32), ("hc_type", C.c_int32),
                ("confidence_level", C.c_double), ("alpha", C.c_double)]


FIT_F64_GP = ["coefficients"]
FIT_F64_G = ["intercept", "r_squared", "adj_r_squared", "residual_std_error"]
INF_F64_GP = ["std_errors", "t_values", "p_values", "ci_lower", "ci_upper"]
INF_F64_G = ["f_statistic", "f_pvalue"]
EXTRA_F64_G = ["aic", "bic", "rss"]


class B200FitOutputs(C.Structure):
    _fields_ = [("coefficients", c_double_p), ("intercept", c_double_p), ("r_squared", c_double_p),
                ("adj_r_squared", c_double_p), ("residual_std_error", c_double_p), ("n_observations", c_int64_p),
                ("status", c_int32_p), ("std_errors", c_double_p), ("t_values", c_double_p),
                ("p_values", c_double_p), ("ci_lower", c_double_p), ("ci_upper", c_double_p),
                ("f_statistic", c_double_p), ("f_pvalue", c_double_p), ("has_inference", c_uint8_p),
                ("aic", c_double_p), ("bic", c_double_p), ("rss", c_double_p)]


class B200PredictOutputs(C.Structure):
    _fields_ = [("yhat", c_double_p), ("yhat_lower", c_double_p), ("yhat_upper", c_double_p),
                ("is_training", c_uint8_p)]


def make_fit_options(method="ols", fit_intercept=True, compute_inference=False, confidence_level=0.95, alpha=0.0,
                     solver="svd", lambda_scaling="raw", hc_type="none"):
    """Options with the reference's defaults (ols_aggregate.cpp:47-52: intercept on, inference off, 0.95, SVD)."""
    return B200FitOptions(METHOD[method], int(bool(fit_intercept)), int(bool(compute_inference)), SOLVER[solver],
                          LAMBDA_SCALING[lambda_scaling], HC[hc_type], float(confidence_level), float(alpha))


def np_ptr(a, ctype):
    return None if a is None else a.ctypes.data_as(C.POINTER(ctype))


def alloc_fit_outputs_numpy(G, p, inference, extras=True):
    """Host (numpy) output arrays + the ctypes struct pointing at them."""
    arrs = {}
    for k in FIT_F64_GP:
        arrs[k] = np.full((G, p), np.nan)
    for k in FIT_F64_G:
        arrs[k] = np.full(G, np.nan)
    arrs["n_observations"] = np.zeros(G, dtype=np.int64)
    arrs["status"] = np.full(G, -1, dtype=np.int32)
    if inference:
        for k in INF_F64_GP:
            arrs[k] = np.full((G, p), np.nan)
        for k in INF_F64_G:
            arrs[k] = np.full(G, np.nan)
        arrs["has_inference"] = np.zeros(G, dtype=np.uint8)
    if extras:
        for k in EXTRA_F64_G:
            arrs[k] = np.full(G, np.nan)
    s = B200FitOutputs()
    for name, ctype in B200FitOutputs._fields_:
        a = arrs.get(name)
        setattr(s, name, None if a is None else a.ctypes.data_as(ctype))
    return arrs, s


def alloc_predict_outputs_numpy(N):
    arrs = {"yhat": np.full(N, np.nan), "yhat_lower": np.full(N, np.nan), "yhat_upper": np.full(N, np.nan),
            "is_training": np.zeros(N, dtype=np.uint8)}
    s = B200PredictOutputs()
    for name, ctype in B200PredictOutputs._fields_:
        setattr(s, name, arrs[name].ctypes.data_as(ctype))
    return arrs, s


def declare_reference_symbols(lib, prefix):
    """argtypes/restype for the seven drop-in symbols (+ AIC/BIC) exported under `prefix`."""
    E = C.POINTER(AnofoxError)
    f = getattr(lib, prefix + "ols_fit")
    f.restype = C.c_bool
    f.argtypes = [DataArray, C.POINTER(DataArray), C.c_size_t, OlsOptions, C.POINTER(FitResultCore),
                  C.POINTER(FitResultInference), E]
    f = getattr(lib, prefix + "ridge_fit")
    f.restype = C.c_bool
    f.argtypes = [DataArray, C.POINTER(DataArray), C.c_size_t, RidgeOptions, C.POINTER(FitResultCore),
                  C.POINTER(FitResultInference), E]
    f = getattr(lib, prefix + "wls_fit")
    f.restype = C.c_bool
    f.argtypes = [DataArray, C.POINTER(DataArray), C.c_size_t, DataArray, OlsOptions, C.POINTER(FitResultCore),
                  C.POINTER(FitResultInference), E]
    f = getattr(lib, prefix + "free_result_core")
    f.restype = None
    f.argtypes = [C.POINTER(FitResultCore)]
    f = getattr(lib, prefix + "free_result_inference")
    f.restype = None
    f.argtypes = [C.POINTER(FitResultInference)]
    f = getattr(lib, prefix + "t_critical")
    f.restype = C.c_double
    f.argtypes = [C.c_double, C.c_size_t]
    f = getattr(lib, prefix + "predict_with_interval")
    f.restype = C.c_bool
    f.argtypes = [c_double_p, C.c_size_t, C.c_double, c_double_p, C.c_size_t, C.c_double, C.c_size_t, C.c_double,
                  C.POINTER(PredictionResult)]
    for nm in ("compute_aic", "compute_bic"):
        f = getattr(lib, prefix + nm)
        f.restype = C.c_bool
        f.argtypes = [C.c_double, C.c_size_t, C.c_size_t, c_double_p, E]


class SingleFit:
    """Python view of one anofox_*_fit call (core + optional inference), arrays copied out."""

    def __init__(self, ok, err, core=None, inf=None):
        self.ok = ok
        self.error_code = err.code
        self.error_message = err.message.decode(errors="replace")
        if ok:
            n = core.coefficients_len
            self.coefficients = np.array([core.coefficients[i] for i in range(n)])
            self.intercept = core.intercept
            self.r_squared = core.r_squared
            self.adj_r_squared = core.adj_r_squared
            self.residual_std_error = core.residual_std_error
            self.n_observations = core.n_observations
            self.n_features = core.n_features
            if inf is not None:
                m = inf.len
                g = lambda p: np.array([p[i] for i in range(m)])  # noqa: E731
                self.std_errors, self.t_values, self.p_values = g(inf.std_errors), g(inf.t_values), g(inf.p_values)
                self.ci_lower, self.ci_upper = g(inf.ci_lower), g(inf.ci_upper)
                self.inference_len = m
                self.f_statistic, self.f_pvalue = inf.f_statistic, inf.f_pvalue
                self.confidence_level = inf.confidence_level


def call_single_fit(lib, prefix, method, y, x_cols, w=None, fit_intercept=True, compute_inference=False,
                    confidence_level=0.95, alpha=0.0, lambda_scaling="raw", solver="svd", hc_type="none",
                    y_validity=None, x_validity=None):
    """Drive anofox_{ols,ridge,wls}_fit exactly as ols_aggregate.cpp:270-296 does (column arrays)."""
    y = np.ascontiguousarray(y, dtype=np.float64)
    cols = [np.ascontiguousarray(c, dtype=np.float64) for c in x_cols]
    keep = [y] + cols

    def da(a, validity=None):
        v = None
        if validity is not None:
            bits = np.packbits(np.asarray(validity, dtype=np.uint8), bitorder="little")
            keep.append(bits)
            v = bits.ctypes.data_as(c_uint8_p)
        return DataArray(a.ctypes.data_as(c_double_p), v, len(a))

    xs = (DataArray * max(len(cols), 1))()
    for j, c in enumerate(cols):
        xs[j] = da(c, None if x_validity is None else x_validity[j])
    core, inf, err = FitResultCore(), FitResultInference(), AnofoxError()
    inf_p = C.byref(inf) if compute_inference else None
    xs_p = xs if len(cols) else None
    if method == "ols":
        o = OlsOptions(fit_intercept, compute_inference, confidence_level, SOLVER[solver], HC[hc_type])
        ok = getattr(lib, prefix + "ols_fit")(da(y, y_validity), xs_p, len(cols), o, C.byref(core), inf_p,
                                              C.byref(err))
    elif method == "ridge":
        o = RidgeOptions(alpha, fit_intercept, compute_inference, confidence_level, SOLVER[solver],
                         LAMBDA_SCALING[lambda_scaling])
        ok = getattr(lib, prefix + "ridge_fit")(da(y, y_validity), xs_p, len(cols), o, C.byref(core), inf_p,
                                                C.byref(err))
    else:
        wv = np.ascontiguousarray(w, dtype=np.float64)
        keep.append(wv)
        o = OlsOptions(fit_intercept, compute_inference, confidence_level, SOLVER[solver], HC[hc_type])
        ok = getattr(lib, prefix + "wls_fit")(da(y, y_validity), xs_p, len(cols), da(wv), o, C.byref(core), inf_p,
                                              C.byref(err))
    res = SingleFit(bool(ok), err, core if ok else None, inf if (ok and compute_inference) else None)
    if ok:
        getattr(lib, prefix + "free_result_core")(C.byref(core))
        if compute_inference:
            getattr(lib, prefix + "free_result_inference")(C.byref(inf))
    return res
