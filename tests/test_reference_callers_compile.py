"""The reference's own callers, UNMODIFIED, compiled against this repo's header (SURVEY.md section 7 step 7,
rows a5 / a7 / b of section 8).

Sources taken where they lie under /root/reference (never copied):
  src/aggregate_functions/{ols,ridge,wls}_aggregate.cpp          (bind + options parser call + Finalize -> anofox_*_fit)
  src/aggregate_functions/{ols,ridge,wls}_predict_aggregate.cpp  (fit + anofox_predict_with_interval per row)
  src/include/map_options_parser.cpp                              (the options parser those binds use)
  src/macros/fit_predict_macros.cpp                               (the *_fit_predict_by SQL macros)
Every hot-path type and prototype they see comes from include/anofox_b200.h: the reference's header is switched
off through its include guard (tests/ref_compile/shim.h). The objects' undefined anofox_* symbols must all be
exported by libanofox_stats_b200.so -- a renamed symbol, a changed struct or a missing enumerator fails here.
Runs where /root/reference exists (the build container); skipped on the GPU box.
"""
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
SOURCES = [
    "src/aggregate_functions/ols_aggregate.cpp",
    "src/aggregate_functions/ridge_aggregate.cpp",
    "src/aggregate_functions/wls_aggregate.cpp",
    "src/aggregate_functions/ols_predict_aggregate.cpp",
    "src/aggregate_functions/ridge_predict_aggregate.cpp",
    "src/aggregate_functions/wls_predict_aggregate.cpp",
    "src/include/map_options_parser.cpp",
    "src/macros/fit_predict_macros.cpp",
]
# the FFI symbols the hot path binds; the predict aggregates use the first six
SEVEN = {"anofox_ols_fit", "anofox_ridge_fit", "anofox_wls_fit", "anofox_free_result_core",
         "anofox_free_result_inference", "anofox_predict_with_interval", "anofox_t_critical"}

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "duckdb", "src", "include")),
                                reason="/root/reference (with DuckDB's headers) is only present in the build container")


def _compile(src, outdir):
    obj = os.path.join(outdir, os.path.basename(src).replace(".cpp", ".o"))
    cmd = ["g++", "-std=c++17", "-O0", "-c", "-fPIC", "-w",
           "-include", os.path.join(ROOT, "tests", "ref_compile", "shim.h"),
           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "ref_compile"),
           "-I", os.path.join(REF, "duckdb", "src", "include"), "-I", os.path.join(REF, "src", "include"),
           "-I", os.path.join(REF, "src"),
           os.path.join(REF, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return src, obj, r.returncode, r.stderr[-3000:]


@pytest.fixture(scope="module")
def objects(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("refobj"))
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        res = list(ex.map(lambda s: _compile(s, out), SOURCES))
    return res


def test_reference_callers_compile_against_b200_header(objects):
    bad = [(s, err) for s, _, rc, err in objects if rc != 0]
    assert not bad, "reference sources no longer compile against include/anofox_b200.h:\n" + \
        "\n".join(f"--- {s}\n{e}" for s, e in bad)


def test_reference_callers_link_to_the_b200_library(objects, b200):
    lib = b200.LIB_PATH
    exported = set()
    for ln in subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True).stdout.splitlines():
        parts = ln.split()
        if len(parts) >= 3:
            exported.add(parts[2])
    wanted = set()
    for s, obj, rc, _ in objects:
        assert rc == 0, s
        for ln in subprocess.run(["nm", "-u", obj], capture_output=True, text=True).stdout.splitlines():
            sym = ln.split()[-1]
            if sym.startswith("anofox_"):
                wanted.add(sym)
    assert wanted, "the reference objects reference no anofox_* symbol at all?"
    assert wanted <= SEVEN, f"hot-path callers bind symbols outside the seven of SURVEY 8(b): {sorted(wanted - SEVEN)}"
    missing = wanted - exported
    assert not missing, f"libanofox_stats_b200.so does not export {sorted(missing)}"
    # the fit aggregates call the three fits and both frees; the predict aggregates the per-row predictor
    assert {"anofox_ols_fit", "anofox_ridge_fit", "anofox_wls_fit", "anofox_free_result_core",
            "anofox_predict_with_interval"} <= wanted
