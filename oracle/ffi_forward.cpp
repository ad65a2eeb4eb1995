// libanofox_ffi.so, CPU flavour -- TEST / BASELINE INFRASTRUCTURE (like everything under oracle/).
// The seven symbols the reference's DuckDB callers bind (src/include/anofox_stats_ffi.h:155,161,166,378,472,655,686),
// forwarded to the CPU oracle. The DuckDB harness (harness/duckdb) links the reference's UNMODIFIED aggregate sources
// against "libanofox_ffi.so"; LD_LIBRARY_PATH decides whether that is this library (CPU baseline "reference callers +
// port") or the one that pulls in libanofox_stats_b200.so (the drop-in proof on a B200).
#include "../include/anofox_b200.h"

extern "C" {
bool oracle_ols_fit(AnofoxDataArray, const AnofoxDataArray *, size_t, AnofoxOlsOptions, AnofoxFitResultCore *, AnofoxFitResultInference *, AnofoxError *);
bool oracle_ridge_fit(AnofoxDataArray, const AnofoxDataArray *, size_t, AnofoxRidgeOptions, AnofoxFitResultCore *, AnofoxFitResultInference *, AnofoxError *);
bool oracle_wls_fit(AnofoxDataArray, const AnofoxDataArray *, size_t, AnofoxDataArray, AnofoxWlsOptions, AnofoxFitResultCore *, AnofoxFitResultInference *, AnofoxError *);
void oracle_free_result_core(AnofoxFitResultCore *);
void oracle_free_result_inference(AnofoxFitResultInference *);
double oracle_t_critical(double, size_t);
bool oracle_predict_with_interval(const double *, size_t, double, const double *, size_t, double, size_t, double, AnofoxPredictionResult *);

bool anofox_ols_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t n, AnofoxOlsOptions o, AnofoxFitResultCore *c,
                    AnofoxFitResultInference *i, AnofoxError *e) { return oracle_ols_fit(y, x, n, o, c, i, e); }
bool anofox_ridge_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t n, AnofoxRidgeOptions o, AnofoxFitResultCore *c,
                      AnofoxFitResultInference *i, AnofoxError *e) { return oracle_ridge_fit(y, x, n, o, c, i, e); }
bool anofox_wls_fit(AnofoxDataArray y, const AnofoxDataArray *x, size_t n, AnofoxDataArray w, AnofoxWlsOptions o,
                    AnofoxFitResultCore *c, AnofoxFitResultInference *i, AnofoxError *e) { return oracle_wls_fit(y, x, n, w, o, c, i, e); }
void anofox_free_result_core(AnofoxFitResultCore *r) { oracle_free_result_core(r); }
void anofox_free_result_inference(AnofoxFitResultInference *r) { oracle_free_result_inference(r); }
double anofox_t_critical(double level, size_t df) { return oracle_t_critical(level, df); }
bool anofox_predict_with_interval(const double *c, size_t n, double b0, const double *x, size_t m, double rse, size_t nobs,
                                  double level, AnofoxPredictionResult *out) {
	return oracle_predict_with_interval(c, n, b0, x, m, rse, nobs, level, out);
}
}
