// Trimmed extension class of the drop-in harness: registers only the hot path's functions.
#pragma once
#include "duckdb.hpp"

namespace duckdb {
class ExtensionLoader;
void RegisterOlsAggregateFunction(ExtensionLoader &loader);
void RegisterRidgeAggregateFunction(ExtensionLoader &loader);
void RegisterWlsAggregateFunction(ExtensionLoader &loader);
void RegisterOlsFitPredictAggregateFunction(ExtensionLoader &loader);
void RegisterRidgeFitPredictAggregateFunction(ExtensionLoader &loader);
void RegisterWlsFitPredictAggregateFunction(ExtensionLoader &loader);
void RegisterOlsFitPredictFunction(ExtensionLoader &loader);
void RegisterRidgeFitPredictFunction(ExtensionLoader &loader);
void RegisterWlsFitPredictFunction(ExtensionLoader &loader);
void RegisterOlsFitFunction(ExtensionLoader &loader);
void RegisterRidgeFitFunction(ExtensionLoader &loader);
void RegisterWlsFitFunction(ExtensionLoader &loader);
void RegisterFitPredictTableMacros(ExtensionLoader &loader);

class AnofoxStatisticsExtension : public Extension {
public:
	void Load(ExtensionLoader &loader) override;
	std::string Name() override;
	std::string Version() const override;
};
} // namespace duckdb
