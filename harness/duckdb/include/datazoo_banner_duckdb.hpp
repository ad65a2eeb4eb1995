// Stand-in for the reference's datazoo-banner submodule (not vendored): DATAZOO_GUARD wraps a DuckDB callback so that
// exceptions get an issue-link footer; the harness passes the callback through unchanged.
#pragma once
namespace datazoo {
struct BannerInfo {};
} // namespace datazoo
#define DATAZOO_GUARD(banner, fn) fn
