// The B200 operators behind Velox's Operator API and their registration.
//   GpuFilterProject     replaces exec::FilterProject   (velox/exec/FilterProject.cpp:44-191)
//   GpuHashAggregation   replaces exec::HashAggregation (velox/exec/HashAggregation.cpp:97-472), optionally with
//                        the FilterProject below it fused in (vgpu_agg_add_input evaluates filter + projections
//                        + accumulation in one kernel, the shape of experimental/wave/exec/ToWave.cpp:346-413)
// Both talk to the CUDA side only through include/vgpu.h; a non-zero status becomes VeloxUserError
// (VGPU_ERR_USER: data errors) or VeloxRuntimeError.  There is no CPU path: if libvgpu.so or the GPU is
// missing, construction throws.
#pragma once
#include <unordered_map>
#include "../../include/vgpu.h"
#include "Lowering.h"
#include "Operator.h"

namespace vb200 {
namespace gpu {

void check(int status, const char* what);           // throws on non-zero vgpu_status
struct DeviceBatch;                                 // device copies of the columns a program reads

// What a FilterNode / ProjectNode pair contributes when fused into the aggregation above it.
struct FilterProjectExprs {
  core::TypedExprPtr filter;                                   // may be null
  std::vector<std::string> names;                              // projection names (empty = identity over the input)
  std::vector<core::TypedExprPtr> projections;
};

class GpuFilterProject : public exec::Operator {
 public:
  GpuFilterProject(exec::DriverCtx* ctx, int32_t id, std::shared_ptr<const core::FilterNode> filter,
                   std::shared_ptr<const core::ProjectNode> project);
  ~GpuFilterProject() override;
  bool needsInput() const override { return input_ == nullptr && !noMoreInput_; }
  void addInput(RowVectorPtr input) override { input_ = std::move(input); }
  RowVectorPtr getOutput() override;
  exec::BlockingReason isBlocked(exec::ContinueFuture*) override { return exec::BlockingReason::kNotBlocked; } // device work is awaited inside getOutput
  bool isFinished() override { return noMoreInput_ && input_ == nullptr; }
  bool isFilter() const override { return exprs_.filter != nullptr; }
  bool preservesOrder() const override { return true; }
  // "for accelerator conversion" (velox/exec/FilterProject.h:63-72)
  const FilterProjectExprs& exprsAndProjection() const { return exprs_; }
  const RowTypePtr& inputType() const { return inputType_; }
  // VARCHAR =, <>, IN run on this operator's value-id dictionary; such a FilterProject is not fused into an aggregation
  bool comparesStrings() const { return !lowered_.stringColumns.empty() || !lowered_.stringLiterals.empty() || !lowered_.stringPredicates.empty(); }
 private:
  FilterProjectExprs exprs_;
  RowTypePtr inputType_;
  std::vector<int> identity_;        // per output: input child index for identity projections, else -1
  std::vector<int> programOutput_;   // per output: index into the program's outputs, else -1
  LoweredProgram lowered_;
  // value ids of the strings compared by the program: the literals first (id = position in lowered_.stringLiterals),
  // then every distinct string of the compared columns in order of first appearance
  std::unordered_map<std::string, int32_t> stringIds_;
  vgpu_program_t program_ = nullptr;
  void* stream_ = nullptr;
};

class GpuHashAggregation : public exec::Operator {
 public:
  // `fused` (optional): evaluate this filter + projections on the raw input inside the aggregation kernel.
  GpuHashAggregation(exec::DriverCtx* ctx, int32_t id, std::shared_ptr<const core::AggregationNode> node,
                     const FilterProjectExprs* fused = nullptr, RowTypePtr fusedInputType = nullptr);
  ~GpuHashAggregation() override;
  bool needsInput() const override { return !noMoreInput_ && !partialFull_; }
  void addInput(RowVectorPtr input) override;
  void noMoreInput() override;
  RowVectorPtr getOutput() override;
  exec::BlockingReason isBlocked(exec::ContinueFuture*) override { return exec::BlockingReason::kNotBlocked; }
  bool isFinished() override { return finished_; }
  void close() override;
  // OperatorStats-shaped counters (velox/exec/Operator.h:83-182) plus what only a device operator has
  struct Stats {
    int64_t inputVectors = 0, inputRows = 0, inputBytes = 0, outputVectors = 0, outputRows = 0;
    int64_t deviceLaunches = 0;      // vgpu_agg_add_input calls (coalesced host batches per launch = inputVectors / deviceLaunches)
    int64_t h2dBytes = 0, stagedBytes = 0; // bytes copied to HBM; of those, bytes that went through the pinned staging ring
    int64_t flushes = 0;             // partial-aggregation flushes (HashAggregation.cpp:326-348)
    bool abandonedPartialAggregation = false; // the partial aggregation did not reduce its input (HashAggregation.cpp:229-233,355-364): groups leave after every device launch
    int64_t regrows = 0;             // table rebuilt because a VARCHAR key left the laid-out range (array -> wider array / hash)
    double addInputWallMs = 0, getOutputWallMs = 0;
  };
  const Stats& stats() const;
  struct Impl;
 private:
  std::unique_ptr<Impl> impl_;
  bool finished_ = false;
  bool partialFull_ = false; // a partial aggregation over its memory limit: getOutput drains the groups, then input resumes
};

// Registers the PlanNodeTranslator (FilterNode / ProjectNode / AggregationNode -> GPU operators) and the
// DriverAdapter that fuses [GpuFilterProject, GpuHashAggregation(raw input)] into one operator.
void registerGpuOperators();

} // namespace gpu
} // namespace vb200
