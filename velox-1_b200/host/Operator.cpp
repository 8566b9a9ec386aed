#include "Operator.h"

namespace vb200 {
namespace exec {

std::vector<std::unique_ptr<PlanNodeTranslator>>& Operator::translators() {
  static std::vector<std::unique_ptr<PlanNodeTranslator>> t;
  return t;
}
void Operator::registerOperator(std::unique_ptr<PlanNodeTranslator> translator) { translators().push_back(std::move(translator)); }
void Operator::unregisterAllOperators() { translators().clear(); DriverFactory::adapters().clear(); }

std::vector<DriverAdapter>& DriverFactory::adapters() {
  static std::vector<DriverAdapter> a;
  return a;
}
void DriverFactory::registerAdapter(DriverAdapter adapter) { adapters().push_back(std::move(adapter)); }

std::vector<std::unique_ptr<Operator>> DriverFactory::replaceOperators(Driver& driver, int32_t begin, int32_t end,
                                                                        std::vector<std::unique_ptr<Operator>> replaceWith) const {
  auto& ops = driver.operators();
  VB_CHECK(begin >= 0 && begin <= end && end <= (int32_t)ops.size(), "replaceOperators: bad range");
  std::vector<std::unique_ptr<Operator>> replaced;
  for (int32_t i = begin; i < end; ++i) replaced.push_back(std::move(ops[i]));
  ops.erase(ops.begin() + begin, ops.begin() + end);
  ops.insert(ops.begin() + begin, std::make_move_iterator(replaceWith.begin()), std::make_move_iterator(replaceWith.end()));
  for (size_t i = 0; i < ops.size(); ++i) ops[i]->setOperatorId((int32_t)i);
  return replaced;
}

namespace {
// The one built-in operator: replays the vectors of a ValuesNode (velox/exec/Values.cpp).
class Values : public SourceOperator {
 public:
  Values(DriverCtx* ctx, int32_t id, std::shared_ptr<const core::ValuesNode> node)
      : SourceOperator(ctx, node->outputType(), id, node->id(), "Values"), node_(std::move(node)) {}
  RowVectorPtr getOutput() override {
    while (next_ < node_->values().size()) {
      auto v = node_->values()[next_++];
      if (v->size() > 0) return v;
    }
    return nullptr;
  }
  bool isFinished() override { return next_ >= node_->values().size(); }
 private:
  std::shared_ptr<const core::ValuesNode> node_;
  size_t next_ = 0;
};
} // namespace

std::unique_ptr<Driver> createDriver(const core::PlanNodePtr& plan, const QueryConfig& config) {
  DriverFactory factory;
  for (auto n = plan; n; n = n->sources().empty() ? nullptr : n->sources()[0]) {
    VB_CHECK(n->sources().size() <= 1, "only linear plans are supported (node " + n->id() + ")");
    factory.planNodes.insert(factory.planNodes.begin(), n);
  }
  auto ctx = std::make_unique<DriverCtx>();
  ctx->queryConfig = config;
  auto driver = std::make_unique<Driver>(std::move(ctx));
  auto& ops = driver->operators();
  for (size_t n = 0; n < factory.planNodes.size(); ++n) {
    const auto& node = factory.planNodes[n];
    const int32_t id = (int32_t)ops.size();
    // Filter directly below Project becomes ONE FilterProject operator, built from the ProjectNode
    // (velox/exec/LocalPlanner.cpp: the planner looks ahead for this pair).
    if (std::dynamic_pointer_cast<const core::FilterNode>(node) && n + 1 < factory.planNodes.size() &&
        std::dynamic_pointer_cast<const core::ProjectNode>(factory.planNodes[n + 1])) {
      continue;
    }
    if (auto values = std::dynamic_pointer_cast<const core::ValuesNode>(node)) {
      ops.push_back(std::make_unique<Values>(driver->driverCtx(), id, values));
      continue;
    }
    std::unique_ptr<Operator> op;
    for (auto& t : Operator::translators()) { // registration order (velox/exec/Operator.cpp:107-126)
      op = t->toOperator(driver->driverCtx(), id, node);
      if (op) break;
    }
    VB_CHECK(op != nullptr, "No operator registered for plan node " + node->name() + " (id " + node->id() + "): call registerGpuOperators()");
    ops.push_back(std::move(op));
  }
  core::PlanFragment fragment;
  fragment.planNode = plan;
  for (auto& a : DriverFactory::adapters()) {
    if (a.inspect) a.inspect(fragment);
    if (a.adapt && a.adapt(factory, *driver)) break; // first adapter that rewrites wins (LocalPlanner.cpp:582-588)
  }
  for (auto& op : ops) op->initialize();
  return driver;
}

void Driver::push(size_t from, RowVectorPtr batch, std::vector<RowVectorPtr>& results) {
  if (from == operators_.size()) { results.push_back(std::move(batch)); return; }
  VB_CHECK(batch->size() > 0, "operators must not produce empty vectors");
  auto& op = *operators_[from];
  if (op.isFinished()) return;
  ContinueFuture future;
  VB_CHECK(op.isBlocked(&future) == BlockingReason::kNotBlocked, op.operatorType() + " blocked: this single-threaded driver has no one to wait for");
  VB_CHECK(op.needsInput(), op.operatorType() + " does not accept input");
  op.addInput(std::move(batch));
  while (auto out = op.getOutput()) push(from + 1, std::move(out), results);
}

std::vector<RowVectorPtr> Driver::run() {
  std::vector<RowVectorPtr> results;
  VB_CHECK(!operators_.empty(), "empty pipeline");
  auto& source = *operators_[0];
  while (!source.isFinished()) {
    auto batch = source.getOutput();
    if (batch) push(1, std::move(batch), results);
  }
  for (size_t i = 1; i < operators_.size(); ++i) { // end of input travels down the pipeline
    operators_[i]->noMoreInput();
    while (auto out = operators_[i]->getOutput()) push(i + 1, std::move(out), results);
  }
  for (auto& op : operators_) op->close();
  return results;
}

std::vector<RowVectorPtr> runPlan(const core::PlanNodePtr& plan, const QueryConfig& config) {
  return createDriver(plan, config)->run();
}

} // namespace exec
} // namespace vb200
