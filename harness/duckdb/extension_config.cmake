# DuckDB extension config of the drop-in harness: a statically linked extension named
# anofox_statistics whose sources are the reference's UNMODIFIED hot-path callers (taken where they lie under
# /root/reference, never copied) plus this directory's trimmed entry point.
duckdb_extension_load(anofox_statistics
    SOURCE_DIR ${CMAKE_CURRENT_LIST_DIR}
    INCLUDE_DIR ${CMAKE_CURRENT_LIST_DIR}/include
)
