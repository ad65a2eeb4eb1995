#!/bin/bash
# oracle/_ref: the reference's own C++ callers, UNMODIFIED, built where they lie under /root/reference into a DuckDB
# with a statically linked, trimmed anofox_statistics extension (harness/duckdb) -- SURVEY.md section 7 step 7.
# Outputs (git-ignored, travel to the GPU box): oracle/_ref/{duckdb,unittest}, oracle/_ref/{cpu,b200}/libanofox_ffi.so,
# oracle/_ref/tests/ (the reference's .test files for the path, prepared by harness/run_reference_tests.py).
# The Rust numerics of the reference are NOT built (no toolchain, un-vendored crates): behind the seven FFI symbols sits
# either the CPU oracle (cpu flavour) or libanofox_stats_b200.so (b200 flavour).
# DuckDB's core is a ~15 minute build on 8 cores; scratch goes to oracle/_ref_build (git- and gpurun-ignored).
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
REF=${ANOFOX_REF:-/root/reference}
[ -d "$REF/duckdb" ] || { echo "no reference tree at $REF"; exit 1; }
make -C "$ROOT/oracle"
mkdir -p "$ROOT/oracle/_ref/cpu" "$ROOT/oracle/_ref/b200" "$ROOT/oracle/_ref_build"
g++ -O2 -std=c++17 -fPIC -shared -o "$ROOT/oracle/_ref/cpu/libanofox_ffi.so" "$ROOT/oracle/ffi_forward.cpp" \
    -L"$ROOT/oracle" -lgroupfit_oracle -Wl,-rpath,'$ORIGIN/../..'
if [ -f "$ROOT/anofox-statistics_b200/libanofox_stats_b200.so" ]; then
  echo 'void anofox_ffi_b200_marker(void) {}' > "$ROOT/oracle/_ref_build/marker.c"
  gcc -fPIC -shared -o "$ROOT/oracle/_ref/b200/libanofox_ffi.so" "$ROOT/oracle/_ref_build/marker.c" -Wl,--no-as-needed \
      -L"$ROOT/anofox-statistics_b200" -lanofox_stats_b200 -Wl,-rpath,'$ORIGIN/../../../anofox-statistics_b200'
fi
cd "$ROOT/oracle/_ref_build"
cmake -G Ninja -DCMAKE_BUILD_TYPE=Release -DOVERRIDE_GIT_DESCRIBE=v1.5.5-0-g0000000000 -DBUILD_UNITTESTS=ON -DBUILD_SHELL=ON \
      -DENABLE_EXTENSION_AUTOLOADING=0 -DENABLE_EXTENSION_AUTOINSTALL=0 \
      -DSKIP_EXTENSIONS="jemalloc;parquet;icu;tpch;tpcds;json;autocomplete" \
      -DDUCKDB_EXTENSION_CONFIGS="$ROOT/harness/duckdb/extension_config.cmake" \
      -DANOFOX_REF="$REF" -DANOFOX_FFI_LIB="$ROOT/oracle/_ref/cpu/libanofox_ffi.so" "$REF/duckdb" > cmake.log 2>&1
ninja -j"${JOBS:-8}" unittest shell > ninja.log 2>&1 || { tail -30 ninja.log; exit 1; }
cp duckdb "$ROOT/oracle/_ref/duckdb"
cp test/unittest "$ROOT/oracle/_ref/unittest"
strip "$ROOT/oracle/_ref/duckdb" "$ROOT/oracle/_ref/unittest"
python "$ROOT/harness/run_reference_tests.py" prepare --out "$ROOT/oracle/_ref/prepare.json"
echo "oracle/_ref ready"
