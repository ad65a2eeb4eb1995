#!/usr/bin/env python3
"""Regenerate tests/golden/*.json from the reference checkout (run in the build container only).

Inputs (read-only, /root/reference):
  * test/data/{ols,wls,ridge,inference}_tests/{input/*.csv,expected/*.json}  -- R 4.5.1 golden vectors
    (generators: validation/generators/*.R; tolerances: */metadata.json)
  * known answers transcribed by hand from test/sql/ols_validation.test, test/sql/error_handling.test,
    test/sql/fit_predict_agg/*.test and the recorded session validation/r_seesion_output.txt

Outputs (committed, travel to the GPU box -- /root/reference does not exist there):
  * tests/golden/r_fixtures.json   the R vectors, inputs + expected values
  * tests/golden/kat.json          the known-answer tests

Nothing here calls reference *code*: the reference's numerics are Rust + un-vendored crates that
cannot be imported or compiled in this image, so these data files are the
strongest pins that exist for the path.
"""
import csv
import json
import os
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_csv(path):
    with open(path) as f:
        rows = list(csv.reader(f))
    hdr = rows[0]
    cols = {h: [float(r[i]) for r in rows[1:]] for i, h in enumerate(hdr)}
    return hdr, cols


def na(v):
    if isinstance(v, list):
        return [na(e) for e in v]
    return None if v == "NA" else v


def aslist(v):
    return v if isinstance(v, list) else [v]


def r_case(suite, name, method, fit_intercept=True, weight_col=None, **extra):
    hdr, cols = read_csv(f"{REF}/test/data/{suite}/input/{name}.csv")
    exp = json.load(open(f"{REF}/test/data/{suite}/expected/{name}.json"))
    xn = [h for h in hdr if h not in ("y", weight_col)]
    case = {
        "name": f"{suite}/{name}",
        "source": f"test/data/{suite}/input/{name}.csv + expected/{name}.json",
        "method": method,
        "fit_intercept": fit_intercept,
        "y": cols["y"],
        "x": [cols[h] for h in xn],
        "w": cols[weight_col] if weight_col else None,
    }
    case.update(extra)
    return case, exp


def main():
    fixtures = []

    # ---- OLS (test/data/ols_tests, tolerance strict 1e-10 / relaxed 1e-8 per metadata.json) ----
    for name, icpt in [("simple_linear", True), ("multiple_regression", True), ("no_intercept", False),
                       ("rank_deficient", True), ("perfect_collinearity", True)]:
        case, exp = r_case("ols_tests", name, "ols", icpt)
        coefs = aslist(na(exp["coefficients"]))
        case["expected"] = {
            "intercept": coefs[0] if icpt else None,
            "coefficients": coefs[1:] if icpt else coefs,
            "r_squared": exp["r_squared"],
            "adj_r_squared": exp["adj_r_squared"],
            "residual_std_error": exp["sigma"],
            "df_residual": exp["df_residual"],
            "rss": sum(r * r for r in exp["residuals"]),
            "n_observations": len(exp["residuals"]),
        }
        case["tol"] = 1e-10
        fixtures.append(case)

    # ---- WLS (test/data/wls_tests) ----
    for name in ["wls_equal_weights", "wls_inverse_variance"]:
        case, exp = r_case("wls_tests", name, "wls", True, weight_col="weight")
        coefs = aslist(na(exp["coefficients"]))
        case["expected"] = {
            "intercept": coefs[0],
            "coefficients": coefs[1:],
            "r_squared": exp["r_squared"],
            "adj_r_squared": exp["adj_r_squared"],
            "residual_std_error": exp["sigma"],
            "n_observations": len(exp["residuals"]),
        }
        case["tol"] = 1e-10
        fixtures.append(case)

    # ---- inference (test/data/inference_tests; p-values / intervals tolerance 1e-8) ----
    for name in ["simple_inference", "multiple_inference"]:
        case, exp = r_case("inference_tests", name, "ols", True, compute_inference=True, confidence_level=0.95)
        c = exp["coefficients"]
        ci = exp["confidence_intervals"]
        ms = exp["model_stats"]
        case["expected"] = {
            "intercept": c["estimates"][0],
            "coefficients": c["estimates"][1:],
            "std_errors": c["std_errors"][1:],
            "t_values": c["t_values"][1:],
            "p_values": c["p_values"][1:],
            "ci_lower": ci["lower_95"][1:],
            "ci_upper": ci["upper_95"][1:],
            "r_squared": ms["r_squared"],
            "adj_r_squared": ms["adj_r_squared"],
            "residual_std_error": ms["sigma"],
            "df_residual": ms["df_residual"],
            "f_statistic": aslist(ms["fstatistic"])[0],
            "f_pvalue": ms.get("f_pvalue", ms.get("f_p_value")),
            # the intercept row of R's table is kept for the special-function tests
            "intercept_std_error": c["std_errors"][0],
            "intercept_t": c["t_values"][0],
            "intercept_p": c["p_values"][0],
        }
        case["model_stats_raw"] = ms
        case["tol"] = 1e-10
        case["tol_inference"] = 1e-8
        fixtures.append(case)

    # ---- prediction (fit pinned; R's true intervals differ from the reference's simplified ones,
    #      lib.rs:2340-2347 -- only `fit` is a pin) ----
    hdr, tr = read_csv(f"{REF}/test/data/inference_tests/input/prediction_train.csv")
    _, nw = read_csv(f"{REF}/test/data/inference_tests/input/prediction_new.csv")
    exp = json.load(open(f"{REF}/test/data/inference_tests/expected/prediction_intervals.json"))
    fixtures.append({
        "name": "inference_tests/prediction_intervals",
        "source": "test/data/inference_tests/input/prediction_{train,new}.csv + expected/prediction_intervals.json",
        "method": "ols", "fit_intercept": True, "y": tr["y"], "x": [tr["x"]], "w": None,
        "x_new": [nw["x"]],
        "expected": {
            "intercept": exp["model_coefficients"][0],
            "coefficients": exp["model_coefficients"][1:],
            "residual_std_error": exp["model_sigma"],
            "yhat": exp["predictions"]["fit"],
            "r_prediction_lower": exp["predictions"]["prediction_lower"],
            "r_prediction_upper": exp["predictions"]["prediction_upper"],
        },
        "tol": 1e-10,
    })

    # ---- ridge (glmnet, coordinate descent: loose; lambda_scaling='glmnet') ----
    for name in ["ridge_lambda_0.1", "ridge_lambda_1.0"]:
        case, exp = r_case("ridge_tests", name, "ridge", True, lambda_scaling="glmnet", alpha=exp_lambda(name))
        coefs = exp["coefficients"]
        case["expected"] = {"intercept": coefs[0], "coefficients": coefs[1:]}
        case["tol"] = 5e-6  # glmnet stops at thresh=1e-7: coefficients agree to ~5e-7
        fixtures.append(case)

    with open(os.path.join(HERE, "r_fixtures.json"), "w") as f:
        json.dump({"generated_by": "tests/golden/make_golden.py", "cases": fixtures}, f)

    # ---- hand-transcribed known answers -------------------------------------------------------
    # rnorm(10, 0, .5) with set.seed(42) (validation/r_seesion_output.txt:121-126): the same stream
    # generated test/data/inference_tests/input/simple_inference.csv as y = 5 + 2x + e, so the first
    # ten deviates are recovered from that file to ~1e-14.
    _, si = read_csv(f"{REF}/test/data/inference_tests/input/simple_inference.csv")
    e = [si["y"][i] - 5.0 - 2.0 * si["x"][i] for i in range(10)]
    ya = [b + 0.5 * e[i] for i, b in enumerate([3, 5, 7, 9, 11])]
    yb = [b + 0.5 * e[5 + i] for i, b in enumerate([10, 20, 30, 40, 50])]
    xa, xb = [1.0, 2, 3, 4, 5], [2.0, 4, 6, 8, 10]

    kat = [
        # test/sql/ols_validation.test:10-25
        dict(name="ols_validation/1", src="test/sql/ols_validation.test:10-25", method="ols", fit_intercept=True,
             y=[2.1, 4.2, 5.9, 8.1, 10.0], x=[[1.0, 2, 3, 4, 5]],
             expect=dict(intercept=(0.15, 10), coef0=(1.97, 10), r_squared=(0.9988932359, 10),
                         adj_r_squared=(0.9985243145, 10), residual_std_error=(0.11972, 5))),
        # :43-58
        dict(name="ols_validation/2", src="test/sql/ols_validation.test:43-58", method="ols", fit_intercept=True,
             y=[9.3490142459, 23.0585207096, 20.1443065614, 27.5069089569, 23.3797539876, 40.4297589129,
                36.6737638447, 43.2302304187, 44.8091576842, 59.3127680131],
             x=[[1.0, 2, 3, 4, 5, 6, 7, 8, 9, 10], [5.2, 2.8, 7.1, 1.5, 9.3, 3.6, 8.4, 4.2, 6.7, 10.1],
                [2.3, 8.1, 3.7, 6.9, 1.4, 9.6, 4.5, 7.2, 5.8, 10.3]],
             expect=dict(intercept=(-1.2644666912, 10), coef0=(3.5092902696, 10), coef1=(0.4860377434, 10),
                         coef2=(1.9882987012, 10), r_squared=(0.9997870909, 10))),
        # :63-86 (no intercept: slope 2.0, intercept NaN)
        dict(name="ols_validation/3", src="test/sql/ols_validation.test:63-86", method="ols", fit_intercept=False,
             y=[2.0, 4, 6, 8, 10], x=[[1.0, 2, 3, 4, 5]],
             expect=dict(coef0=(2.0, 10), intercept_is_nan=True)),
        # :91-104 (constant feature)
        dict(name="ols_validation/4", src="test/sql/ols_validation.test:91-104", method="ols", fit_intercept=True,
             y=[2.0, 4, 6, 8, 10], x=[[1.0, 2, 3, 4, 5], [5.0, 5, 5, 5, 5]],
             expect=dict(intercept=(0.0, 10), coef0=(2.0, 10), coef1_is_nan=True)),
        # :106-119 (perfect collinearity x2 = 2 x1)
        dict(name="ols_validation/5", src="test/sql/ols_validation.test:106-119", method="ols", fit_intercept=True,
             y=[3.0, 5, 7, 9, 11], x=[[1.0, 2, 3, 4, 5], [2.0, 4, 6, 8, 10]],
             expect=dict(intercept=(1.0, 10), r_squared=(1.0, 10), n_nan_coef=1)),
        # :121-133 exact fit: rse < 1e-4
        dict(name="ols_validation/6", src="test/sql/ols_validation.test:121-133", method="ols", fit_intercept=True,
             y=[1.0, 2, 3, 4, 5], x=[[1.0, 2, 3, 4, 5]], expect=dict(rse_below=1e-4)),
        # :135-148 n = 3
        dict(name="ols_validation/7", src="test/sql/ols_validation.test:135-148", method="ols", fit_intercept=True,
             y=[2.0, 4, 6], x=[[1.0, 2, 3]],
             expect=dict(intercept=(0.0, 10), coef0=(2.0, 10), r_squared=(1.0, 10))),
        # :186-210 1e-10 recovery
        dict(name="ols_validation/10-11", src="test/sql/ols_validation.test:186-210", method="ols",
             fit_intercept=True, y=[3.0 + 2 * i for i in range(20)], x=[[1.0 + i for i in range(20)]],
             expect=dict(coef0_abs=(2.0, 1e-10), intercept_abs=(1.0, 1e-10))),
        # validation/r_seesion_output.txt:593-611 (DuckDB == R lm to 2e-14)
        dict(name="r_session/ols_group_a", src="validation/r_seesion_output.txt:121-126,597-606", method="ols",
             fit_intercept=True, y=ya, x=[xa],
             expect=dict(intercept_abs=(1.331024882267129, 1e-11), coef0_abs=(1.96320902641729, 1e-11),
                         r_squared_abs=(0.988038930915687, 1e-11), adj_r_squared_abs=(0.984051907887583, 1e-11),
                         n_observations=5)),
        dict(name="r_session/ols_group_b", src="validation/r_seesion_output.txt:121-126,597-606", method="ols",
             fit_intercept=True, y=yb, x=[xb],
             expect=dict(intercept_abs=(0.237586423198462, 1e-11), coef0_abs=(5.01484306376291, 1e-11),
                         r_squared_abs=(0.998953896758364, 1e-11), adj_r_squared_abs=(0.998605195677819, 1e-11))),
        # :621-640 no intercept (r2 uncentred)
        dict(name="r_session/ols_noint_group_a", src="validation/r_seesion_output.txt:621-640", method="ols",
             fit_intercept=False, y=ya, x=[xa],
             expect=dict(coef0_abs=(2.32621581249014, 1e-11), r_squared_abs=(0.993069168187997, 1e-11))),
        dict(name="r_session/ols_noint_group_b", src="validation/r_seesion_output.txt:621-640", method="ols",
             fit_intercept=False, y=yb, x=[xb],
             expect=dict(coef0_abs=(5.04724121238088, 1e-11), r_squared_abs=(0.999802919286896, 1e-11))),
        # :662-671 ridge raw lambda = 1 as recorded from the extension (legacy build)
        dict(name="r_session/ridge_group_a", src="validation/r_seesion_output.txt:668-671", method="ridge",
             alpha=1.0, fit_intercept=True, y=ya, x=[xa],
             expect=dict(intercept_abs=(1.866445525835480, 1e-11), coef0_abs=(1.78473547856117, 1e-11),
                         r_squared_abs=(0.979873319916384, 1e-11))),
        dict(name="r_session/ridge_group_b", src="validation/r_seesion_output.txt:668-671", method="ridge",
             alpha=1.0, fit_intercept=True, y=yb, x=[xb],
             expect=dict(intercept_abs=(0.971465895944252, 1e-11), coef0_abs=(4.89252981830527, 1e-11),
                         r_squared_abs=(0.998359635070822, 1e-11))),
    ]
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump({"generated_by": "tests/golden/make_golden.py", "cases": kat}, f)
    print(f"wrote {len(fixtures)} R fixtures and {len(kat)} known-answer tests")


def exp_lambda(name):
    return float(name.split("_")[-1])


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    main()
