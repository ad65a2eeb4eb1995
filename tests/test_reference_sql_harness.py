"""The reference's own sqllogictest files for the hot path, run through the reference's UNMODIFIED DuckDB callers
(oracle/_ref, built by oracle/build_ref.sh from /root/reference; harness/run_reference_tests.py prepares and runs
them). CPU flavour: the oracle behind the seven FFI symbols. b200 flavour (-m gpu): libanofox_stats_b200.so behind the
same binary -- the drop-in proof. Skipped where oracle/_ref has not been built (it is a 15-minute DuckDB build)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFDIR = os.path.join(ROOT, "oracle", "_ref")
have = all(os.path.exists(os.path.join(REFDIR, f)) for f in ("unittest", "cpu/libanofox_ffi.so", "tests"))
needs_ref = pytest.mark.skipif(not have, reason="oracle/_ref not built (oracle/build_ref.sh, needs /root/reference)")


def _run(flavour, tmp_path):
    out = str(tmp_path / "ref.json")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "harness", "run_reference_tests.py"), "run", "--flavour", flavour,
                        "--out", out], capture_output=True, text=True, timeout=1200)
    res = json.load(open(out))
    files = {k: v for k, v in res.items() if not k.startswith("_")}
    bad = {k: v for k, v in files.items() if v["status"] != "pass"}
    assert not bad, json.dumps(bad, indent=1)[:3000]
    assert len(files) >= 15 and sum(v["assertions"] for v in files.values()) > 400, (len(files), r.stdout[-500:])


@needs_ref
def test_reference_sql_tests_pass_with_the_oracle_behind_the_reference_callers(tmp_path):
    _run("cpu", tmp_path)


@needs_ref
@pytest.mark.gpu
def test_reference_sql_tests_pass_with_the_b200_library_behind_the_reference_callers(tmp_path, b200):
    b200.device_check()
    assert os.path.exists(os.path.join(REFDIR, "b200", "libanofox_ffi.so"))
    _run("b200", tmp_path)
