// Registration seams:
//   1. Operator::registerOperator(PlanNodeTranslator)  — velox/exec/Operator.h:507; our translator claims
//      FilterNode / ProjectNode / AggregationNode.
//   2. DriverFactory::registerAdapter(DriverAdapter)    — velox/exec/Driver.h:478-536; our adapter replaces
//      [GpuFilterProject, GpuHashAggregation(raw input)] by ONE fused operator through replaceOperators,
//      the way experimental/wave/exec/ToWave.cpp:346-413 rewrites a pipeline.
#include "GpuOperators.h"

namespace vb200 {
namespace gpu {
using namespace core;

namespace {
class GpuTranslator : public exec::PlanNodeTranslator {
 public:
  std::unique_ptr<exec::Operator> toOperator(exec::DriverCtx* ctx, int32_t id, const PlanNodePtr& node) override {
    if (auto project = std::dynamic_pointer_cast<const ProjectNode>(node)) {
      auto filter = std::dynamic_pointer_cast<const FilterNode>(project->sources()[0]);
      return std::make_unique<GpuFilterProject>(ctx, id, filter, project);
    }
    if (auto filter = std::dynamic_pointer_cast<const FilterNode>(node)) return std::make_unique<GpuFilterProject>(ctx, id, filter, nullptr);
    if (auto agg = std::dynamic_pointer_cast<const AggregationNode>(node)) return std::make_unique<GpuHashAggregation>(ctx, id, agg);
    return nullptr; // not ours: the next translator (or the planner's error) takes it
  }
};

bool fuseFilterProjectIntoAggregation(const exec::DriverFactory& factory, exec::Driver& driver) {
  auto& ops = driver.operators();
  for (size_t i = 0; i + 1 < ops.size(); ++i) {
    auto fp = dynamic_cast<GpuFilterProject*>(ops[i].get());
    auto ha = dynamic_cast<GpuHashAggregation*>(ops[i + 1].get());
    if (!fp || !ha || fp->comparesStrings()) continue;
    std::shared_ptr<const AggregationNode> node;
    for (auto& n : factory.planNodes) if (n->id() == ha->planNodeId()) node = std::dynamic_pointer_cast<const AggregationNode>(n);
    if (!node || !AggregationNode::isRawInput(node->step())) continue;
    std::unique_ptr<exec::Operator> fused;
    try {
      fused = std::make_unique<GpuHashAggregation>(driver.driverCtx(), (int32_t)i, node, &fp->exprsAndProjection(), fp->inputType());
    } catch (const VeloxException&) {
      continue; // a shape the fused kernel does not take (e.g. a computed VARCHAR key): keep the two operators
    }
    std::vector<std::unique_ptr<exec::Operator>> with;
    with.push_back(std::move(fused));
    factory.replaceOperators(driver, (int32_t)i, (int32_t)i + 2, std::move(with));
    return true;
  }
  return false;
}
} // namespace

void registerGpuOperators() {
  exec::Operator::registerOperator(std::make_unique<GpuTranslator>());
  exec::DriverFactory::registerAdapter(exec::DriverAdapter{"b200-fuse-filterproject-aggregation", nullptr, fuseFilterProjectIntoAggregation});
}

} // namespace gpu
} // namespace vb200
