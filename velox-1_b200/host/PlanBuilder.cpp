// PlanBuilder for the path's plan shapes (velox/exec/tests/utils/PlanBuilder.h): values -> filter ->
// project -> partial/intermediate/final/single aggregation, plus addNode for custom nodes.
#include <cctype>

#include "Plan.h"

namespace vb200 {
using namespace core;

PlanBuilder& PlanBuilder::values(std::vector<RowVectorPtr> vectors) {
  node_ = std::make_shared<ValuesNode>(nextId(), std::move(vectors));
  return *this;
}

PlanBuilder& PlanBuilder::filter(const std::string& expr) {
  VB_CHECK(node_ != nullptr, "filter() needs a source node");
  node_ = std::make_shared<FilterNode>(nextId(), parse::parseExpr(expr, node_->outputType()), node_);
  return *this;
}

PlanBuilder& PlanBuilder::project(const std::vector<std::string>& exprs) {
  VB_CHECK(node_ != nullptr, "project() needs a source node");
  std::vector<std::string> names;
  std::vector<TypedExprPtr> projections;
  for (size_t i = 0; i < exprs.size(); ++i) {
    std::string alias;
    auto e = parse::parseExpr(exprs[i], node_->outputType(), &alias);
    if (alias.empty()) {
      alias = e->kind() == ITypedExpr::Kind::kFieldAccess ? static_cast<const FieldAccessTypedExpr*>(e.get())->name() : "p" + std::to_string(i);
    }
    names.push_back(alias);
    projections.push_back(std::move(e));
  }
  node_ = std::make_shared<ProjectNode>(nextId(), std::move(names), std::move(projections), node_);
  return *this;
}

namespace {
// "fn(arg) [as name]" with arg = column | * | integer literal | nothing
void parseAggregate(const std::string& text, const RowTypePtr& in, size_t ordinal, AggregationNode::Aggregate& out, std::string& name) {
  size_t p = 0;
  auto ws = [&] { while (p < text.size() && isspace((unsigned char)text[p])) ++p; };
  auto word = [&] { ws(); size_t b = p; while (p < text.size() && (isalnum((unsigned char)text[p]) || text[p] == '_')) ++p; return text.substr(b, p - b); };
  out.function = word();
  for (auto& c : out.function) c = (char)tolower(c);
  ws();
  VB_USER_CHECK(p < text.size() && text[p] == '(', "Aggregate call expected: " + text);
  ++p;
  ws();
  out.args.clear();
  out.rawInputType = BIGINT();
  if (text[p] == '*') { ++p; }
  else if (text[p] != ')') {
    const std::string arg = word();
    VB_USER_CHECK(!arg.empty(), "Aggregate argument must be a column or a literal: " + text);
    if (isdigit((unsigned char)arg[0])) {
      out.args.push_back(std::make_shared<ConstantTypedExpr>(BIGINT(), false, std::stoll(arg), 0.0));
    } else {
      const int idx = in->getChildIdx(arg);
      out.args.push_back(std::make_shared<FieldAccessTypedExpr>(in->childAt(idx), arg));
      out.rawInputType = in->childAt(idx);
    }
  }
  ws();
  VB_USER_CHECK(p < text.size() && text[p] == ')', "')' expected in aggregate: " + text);
  ++p;
  name = "a" + std::to_string(ordinal);
  std::string as = word();
  for (auto& c : as) c = (char)tolower(c);
  if (as == "as") name = word();
  if (out.function != "count") VB_USER_CHECK(out.args.size() == 1 && out.args[0]->kind() == ITypedExpr::Kind::kFieldAccess, out.function + " takes one column argument: " + text);
  if (out.function == "sum" || out.function == "avg") VB_USER_CHECK(out.rawInputType->isInteger() || out.rawInputType->isFloating(), "Aggregate function signature is not supported: " + out.function + "(" + out.rawInputType->toString() + ")");
  AggregationNode::finalType(out.function, out.rawInputType); // throws for unknown functions
}
} // namespace

PlanBuilder& PlanBuilder::aggregation(const std::vector<std::string>& keys, const std::vector<std::string>& aggregates,
                                      AggregationNode::Step step, bool ignoreNullKeys, const std::vector<std::string>& masks) {
  VB_CHECK(node_ != nullptr, "aggregation needs a source node");
  VB_CHECK(AggregationNode::isRawInput(step), "use intermediateAggregation()/finalAggregation() after a partial aggregation");
  std::vector<AggregationNode::Aggregate> aggs(aggregates.size());
  std::vector<std::string> names(aggregates.size());
  for (size_t i = 0; i < aggregates.size(); ++i) parseAggregate(aggregates[i], node_->outputType(), i, aggs[i], names[i]);
  VB_USER_CHECK(masks.empty() || masks.size() == aggregates.size(), "one mask (or \"\") per aggregate expected");
  for (size_t i = 0; i < masks.size(); ++i) {
    if (masks[i].empty()) continue;
    const auto& in = node_->outputType();
    const TypePtr t = in->childAt(in->getChildIdx(masks[i]));
    VB_USER_CHECK(t->kind() == TypeKind::BOOLEAN, "aggregation mask " + masks[i] + " must be a BOOLEAN column");
    aggs[i].mask = std::make_shared<FieldAccessTypedExpr>(t, masks[i]);
  }
  node_ = std::make_shared<AggregationNode>(nextId(), step, keys, std::move(names), std::move(aggs), ignoreNullKeys, node_);
  return *this;
}

// final / intermediate over the partial aggregation below: same keys and names, each aggregate reads
// its own intermediate column (PlanBuilder::finalAggregation() of the reference).
PlanBuilder& PlanBuilder::followUp(AggregationNode::Step step) {
  auto partial = std::dynamic_pointer_cast<const AggregationNode>(node_);
  VB_CHECK(partial != nullptr && AggregationNode::isPartialOutput(partial->step()), "final/intermediate aggregation must follow a partial or intermediate aggregation");
  std::vector<AggregationNode::Aggregate> aggs;
  const auto& in = partial->outputType();
  for (size_t i = 0; i < partial->aggregates().size(); ++i) {
    AggregationNode::Aggregate a = partial->aggregates()[i];
    a.mask = nullptr; // masks apply to raw input only
    const std::string& col = partial->aggregateNames()[i];
    a.args = {std::make_shared<FieldAccessTypedExpr>(in->childAt(in->getChildIdx(col)), col)};
    aggs.push_back(std::move(a));
  }
  node_ = std::make_shared<AggregationNode>(nextId(), step, partial->groupingKeys(), partial->aggregateNames(), std::move(aggs), partial->ignoreNullKeys(), node_);
  return *this;
}
PlanBuilder& PlanBuilder::intermediateAggregation() { return followUp(AggregationNode::Step::kIntermediate); }
PlanBuilder& PlanBuilder::finalAggregation() { return followUp(AggregationNode::Step::kFinal); }

PlanBuilder& PlanBuilder::orderBy(const std::vector<std::string>& keys, bool isPartial) {
  std::vector<std::string> names;
  std::vector<core::SortOrder> orders;
  for (auto& k : keys) {
    std::vector<std::string> words;
    std::string w;
    for (char c : k + " ") { if (c == ' ') { if (!w.empty()) words.push_back(w); w.clear(); } else w += (char)std::toupper((unsigned char)c); }
    VB_USER_CHECK(!words.empty(), "orderBy: empty key");
    core::SortOrder o;
    size_t first = k.find_first_not_of(' ');
    names.push_back(k.substr(first, k.find(' ', first) == std::string::npos ? std::string::npos : k.find(' ', first) - first));
    for (size_t i = 1; i < words.size(); ++i) {
      if (words[i] == "DESC") { o.ascending = false; o.nullsFirst = false; }
      else if (words[i] == "ASC") { o.ascending = true; }
      else if (words[i] == "NULLS" && i + 1 < words.size()) { o.nullsFirst = words[i + 1] == "FIRST"; ++i; }
      else throw VeloxUserError("orderBy: cannot parse '" + k + "'");
    }
    orders.push_back(o);
  }
  node_ = std::make_shared<core::OrderByNode>(nextId(), std::move(names), std::move(orders), isPartial, node_);
  return *this;
}

PlanBuilder& PlanBuilder::addNode(std::function<PlanNodePtr(std::string, PlanNodePtr)> make) {
  node_ = make(nextId(), node_);
  return *this;
}

} // namespace vb200
