// No-op stand-in for the reference's telemetry header (its posthog-telemetry submodule is not vendored).
#pragma once
#include <string>
namespace duckdb {
struct PostHogTelemetry {
	static PostHogTelemetry &Instance() { static PostHogTelemetry t; return t; }
	template <class... A> void RecordFunctionCall(A &&...) {}
	template <class... A> void CaptureExtensionLoad(A &&...) {}
};
} // namespace duckdb
