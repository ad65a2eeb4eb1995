/* Force-included before the reference's UNMODIFIED aggregate sources (tests/test_reference_callers_compile.py):
 * the reference's own header src/include/anofox_stats_ffi.h is switched off through its include guard, so every
 * type and prototype those files see for the hot path comes from include/anofox_b200.h. A drift in a struct
 * layout, an enumerator or one of the seven prototypes then fails this compile.
 * The two enums below belong to model families outside the path (AFT, GLMM); ffi_enum_converters.hpp converts
 * them in the same file as the solver / hc_type / lambda_scaling converters the path uses, so they must exist. */
#ifndef ANOFOX_B200_REF_COMPILE_SHIM_H
#define ANOFOX_B200_REF_COMPILE_SHIM_H
#define ANOFOX_STATS_FFI_H /* the reference header's guard: its body is skipped */
#include "anofox_b200.h"
typedef enum { ANOFOX_AFT_WEIBULL = 0, ANOFOX_AFT_LOGNORMAL = 1, ANOFOX_AFT_LOGLOGISTIC = 2, ANOFOX_AFT_EXPONENTIAL = 3 } AnofoxAftDistribution;
typedef enum {
	ANOFOX_GLMM_GAUSSIAN = 0, ANOFOX_GLMM_POISSON = 1, ANOFOX_GLMM_BINOMIAL = 2, ANOFOX_GLMM_NEGBINOMIAL = 3,
	ANOFOX_GLMM_GAMMA = 4, ANOFOX_GLMM_TWEEDIE = 5
} AnofoxGlmmFamily;
#endif
