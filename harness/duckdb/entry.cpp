// Trimmed entry point of the drop-in harness. The reference's LoadInternal (src/anofox_statistics_extension.cpp:64-216)
// registers 97 function families that need its whole Rust FFI; this one registers the hot path only: the three fit
// aggregates, the three fit-predict aggregates (engine of the *_fit_predict_by macros), the three window forms and the
// three scalar array fits -- all of them the reference's own, unmodified translation units.
#include "anofox_statistics_extension.hpp"
#include "duckdb/main/extension/extension_loader.hpp"
#include "anofox_statistics_banner.hpp"

const datazoo::BannerInfo ANOFOX_STATISTICS_BANNER {};

namespace duckdb {

static void LoadInternal(ExtensionLoader &loader) {
	RegisterOlsFitFunction(loader);
	RegisterRidgeFitFunction(loader);
	RegisterWlsFitFunction(loader);
	RegisterOlsAggregateFunction(loader);
	RegisterRidgeAggregateFunction(loader);
	RegisterWlsAggregateFunction(loader);
	RegisterOlsFitPredictAggregateFunction(loader);
	RegisterRidgeFitPredictAggregateFunction(loader);
	RegisterWlsFitPredictAggregateFunction(loader);
	RegisterOlsFitPredictFunction(loader);
	RegisterRidgeFitPredictFunction(loader);
	RegisterWlsFitPredictFunction(loader);
	// the *_fit_predict_by table macros are SQL text bound at call time: those of other model families stay unusable
	RegisterFitPredictTableMacros(loader);
}

void AnofoxStatisticsExtension::Load(ExtensionLoader &loader) { LoadInternal(loader); }
std::string AnofoxStatisticsExtension::Name() { return "anofox_statistics"; }
std::string AnofoxStatisticsExtension::Version() const { return "b200-harness"; }

} // namespace duckdb

extern "C" {
DUCKDB_CPP_EXTENSION_ENTRY(anofox_statistics, loader) { duckdb::LoadInternal(loader); }
}
