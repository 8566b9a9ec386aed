// Operator API of the reference, as far as the path's operators and their registration need it.
//   exec::Operator            <- velox/exec/Operator.h:300-520 (needsInput / addInput / noMoreInput /
//                                getOutput / isBlocked / isFinished / close, registerOperator)
//   exec::PlanNodeTranslator  <- velox/exec/Operator.h:256-297
//   exec::DriverAdapter / DriverFactory::registerAdapter / replaceOperators
//                             <- velox/exec/Driver.h:478-536, velox/exec/LocalPlanner.cpp:563-625
//   exec::Driver              <- velox/exec/Driver.cpp (single pipeline, one thread; no blocking states)
// The LocalPlanner analogue here knows only the Values source; every other node must be claimed by a
// registered PlanNodeTranslator — this repository has no CPU operators to fall back to.
#pragma once
#include <functional>
#include <future>
#include <optional>

#include "Plan.h"

namespace vb200 {
namespace exec {

struct QueryConfig {
  int32_t preferredOutputBatchRows = 1024; // velox/core/QueryConfig.h kPreferredOutputBatchRows ("preferred_output_batch_rows")
  // GPU knobs (the reference's QueryConfig is a string map; these would be its "gpu."-prefixed keys)
  int32_t device = 0;                      // gpu.device: CUDA device of this driver (one driver = one GPU stream)
  bool stringViewKeys = true;              // gpu.string_view_keys: VARCHAR keys that arrive as StringViews are mapped to numbers on the device
  int64_t coalesceRows = 1 << 20;          // gpu.coalesce_rows: host batches are gathered on the device until this many rows before a kernel runs
  int32_t errorCheckEvery = 16;            // gpu.error_check_every: device launches between two data-error checks (plus one at noMoreInput)
  int32_t abandonPartialAggregationMinRows = 100000; // abandon_partial_aggregation_min_rows (velox/core/QueryConfig.h)
  int32_t abandonPartialAggregationMinPct = 80;      // abandon_partial_aggregation_min_pct
  int64_t maxPartialAggregationMemory = 1ll << 34; // max_partial_aggregation_memory (device bytes of logged partial groups before a flush)
};

struct DriverCtx {
  int driverId = 0;
  int pipelineId = 0;
  QueryConfig queryConfig;
};

enum class BlockingReason { kNotBlocked };
using ContinueFuture = std::future<void>;   // folly::SemiFuture<folly::Unit> in the reference (velox/exec/Driver.h)
namespace common { struct SpillConfig {}; } // the path never spills; kept for the constructor's shape

class Operator {
 public:
  // velox/exec/Operator.h:309-315
  Operator(DriverCtx* ctx, RowTypePtr outputType, int32_t operatorId, std::string planNodeId, std::string operatorType,
           std::optional<common::SpillConfig> spillConfig = std::nullopt)
      : ctx_(ctx), outputType_(std::move(outputType)), operatorId_(operatorId), planNodeId_(std::move(planNodeId)), operatorType_(std::move(operatorType)) { (void)spillConfig; }
  virtual ~Operator() = default;
  virtual void initialize() {}
  virtual bool needsInput() const = 0;
  virtual void addInput(RowVectorPtr input) = 0;
  virtual void noMoreInput() { noMoreInput_ = true; }
  virtual RowVectorPtr getOutput() = 0; // nullptr or a NON-empty vector (velox/exec/Driver.cpp:550-556)
  virtual BlockingReason isBlocked(ContinueFuture* future) = 0; // velox/exec/Operator.h:361 (pure virtual there, too)
  virtual bool isFinished() = 0;
  virtual void close() { input_ = nullptr; }
  virtual bool isFilter() const { return false; }
  virtual bool preservesOrder() const { return false; }
  const RowTypePtr& outputType() const { return outputType_; }
  const std::string& operatorType() const { return operatorType_; }
  const std::string& planNodeId() const { return planNodeId_; }
  int32_t operatorId() const { return operatorId_; }
  void setOperatorId(int32_t id) { operatorId_ = id; }

  // Operator::registerOperator (velox/exec/Operator.h:507): translators are tried in registration order.
  static void registerOperator(std::unique_ptr<class PlanNodeTranslator> translator);
  static std::vector<std::unique_ptr<class PlanNodeTranslator>>& translators();
  static void unregisterAllOperators();

 protected:
  DriverCtx* ctx_;
  RowTypePtr outputType_;
  int32_t operatorId_;
  std::string planNodeId_;
  std::string operatorType_;
  RowVectorPtr input_;
  bool noMoreInput_ = false;
};

class SourceOperator : public Operator {
 public:
  using Operator::Operator;
  bool needsInput() const override { return false; }
  void addInput(RowVectorPtr) override { throw VeloxRuntimeError("SourceOperator does not support addInput()"); }
  BlockingReason isBlocked(ContinueFuture*) override { return BlockingReason::kNotBlocked; }
};

class PlanNodeTranslator {
 public:
  virtual ~PlanNodeTranslator() = default;
  // nullptr = this translator does not handle the node
  virtual std::unique_ptr<Operator> toOperator(DriverCtx* ctx, int32_t id, const core::PlanNodePtr& node) { return nullptr; }
};

} // namespace exec
namespace core {
// velox/core/PlanFragment.h:40-60 (the fields an adapter's inspect() can look at on this path)
struct PlanFragment {
  PlanNodePtr planNode; // top level (root) plan node
};
} // namespace core
namespace exec {

class Driver;
struct DriverFactory {
  std::vector<core::PlanNodePtr> planNodes; // source first
  // Replaces operators [begin, end) of `driver` and re-numbers the rest (LocalPlanner.cpp:591-625).
  std::vector<std::unique_ptr<Operator>> replaceOperators(Driver& driver, int32_t begin, int32_t end,
                                                           std::vector<std::unique_ptr<Operator>> replaceWith) const;
  struct DriverAdapter {
    std::string label;
    std::function<void(const core::PlanFragment&)> inspect; // velox/exec/Driver.h:482-486
    std::function<bool(const DriverFactory&, Driver&)> adapt;
  };
  static void registerAdapter(DriverAdapter adapter);
  static std::vector<DriverAdapter>& adapters();
};
using DriverAdapter = DriverFactory::DriverAdapter;

class Driver {
 public:
  explicit Driver(std::unique_ptr<DriverCtx> ctx) : ctx_(std::move(ctx)) {}
  DriverCtx* driverCtx() { return ctx_.get(); }
  std::vector<std::unique_ptr<Operator>>& operators() { return operators_; }
  // Runs the pipeline to completion and returns what the last operator produced.
  std::vector<RowVectorPtr> run();
 private:
  void push(size_t from, RowVectorPtr batch, std::vector<RowVectorPtr>& results);
  std::unique_ptr<DriverCtx> ctx_;
  std::vector<std::unique_ptr<Operator>> operators_;
};

// Builds the single-pipeline driver of a linear plan (LocalPlanner::plan + DriverFactory::createDriver)
// and gives the registered adapters a chance to rewrite it.
std::unique_ptr<Driver> createDriver(const core::PlanNodePtr& plan, const QueryConfig& config = {});
// AssertQueryBuilder(plan).copyResults analogue.
std::vector<RowVectorPtr> runPlan(const core::PlanNodePtr& plan, const QueryConfig& config = {});

} // namespace exec
} // namespace vb200
