// Typed expressions, plan nodes and a PlanBuilder for the path's plan shapes.
//   core::ITypedExpr family   <- velox/core/Expressions.h (FieldAccessTypedExpr, ConstantTypedExpr,
//                                CallTypedExpr, CastTypedExpr)
//   core::PlanNode family     <- velox/core/PlanNode.h (ValuesNode, FilterNode, ProjectNode,
//                                AggregationNode with Step {kPartial,kFinal,kIntermediate,kSingle})
//   exec::test::PlanBuilder   <- velox/exec/tests/utils/PlanBuilder.h (values / filter / project /
//                                partialAggregation / finalAggregation / singleAggregation / addNode)
//   parse::parseExpr          <- velox/duckdb/conversion/DuckParser.cpp:53-110,239-290 (operator ->
//                                function-name mapping: plus minus multiply divide mod negate eq neq
//                                lt lte gt gte between and or not is_null, cast / ::TYPE)
#pragma once
#include <functional>
#include <map>
#include <optional>

#include "Vectors.h"

namespace vb200 {
namespace core {

// ---- expressions -----------------------------------------------------------------------------------
class ITypedExpr;
using TypedExprPtr = std::shared_ptr<const ITypedExpr>;

class ITypedExpr {
 public:
  enum class Kind { kFieldAccess, kConstant, kCall, kCast };
  ITypedExpr(Kind k, TypePtr type, std::vector<TypedExprPtr> inputs = {}) : kind_(k), type_(std::move(type)), inputs_(std::move(inputs)) {}
  virtual ~ITypedExpr() = default;
  Kind kind() const { return kind_; }
  const TypePtr& type() const { return type_; }
  const std::vector<TypedExprPtr>& inputs() const { return inputs_; }
  virtual std::string toString() const = 0;
 private:
  Kind kind_;
  TypePtr type_;
  std::vector<TypedExprPtr> inputs_;
};

class FieldAccessTypedExpr : public ITypedExpr {
 public:
  FieldAccessTypedExpr(TypePtr type, std::string name) : ITypedExpr(Kind::kFieldAccess, std::move(type)), name_(std::move(name)) {}
  const std::string& name() const { return name_; }
  std::string toString() const override { return name_; }
 private:
  std::string name_;
};

// Scalar constant; integers / dates / booleans in i, floating point in d, strings in s.
class ConstantTypedExpr : public ITypedExpr {
 public:
  ConstantTypedExpr(TypePtr type, bool isNull, int64_t i, double d, std::string s = "")
      : ITypedExpr(Kind::kConstant, std::move(type)), isNull_(isNull), i_(i), d_(d), s_(std::move(s)) {}
  bool isNull() const { return isNull_; }
  int64_t intValue() const { return i_; }
  double doubleValue() const { return d_; }
  const std::string& stringValue() const { return s_; }
  std::string toString() const override {
    if (isNull_) return "null";
    if (type()->kind() == TypeKind::VARCHAR) return "'" + s_ + "'";
    return type()->isFloating() ? std::to_string(d_) : std::to_string(i_);
  }
 private:
  bool isNull_;
  int64_t i_;
  double d_;
  std::string s_;
};

class CallTypedExpr : public ITypedExpr {
 public:
  CallTypedExpr(TypePtr type, std::vector<TypedExprPtr> inputs, std::string name)
      : ITypedExpr(Kind::kCall, std::move(type), std::move(inputs)), name_(std::move(name)) {}
  const std::string& name() const { return name_; }
  std::string toString() const override {
    std::string s = name_ + "(";
    for (size_t i = 0; i < inputs().size(); ++i) s += (i ? "," : "") + inputs()[i]->toString();
    return s + ")";
  }
 private:
  std::string name_;
};

class CastTypedExpr : public ITypedExpr {
 public:
  CastTypedExpr(TypePtr type, TypedExprPtr input) : ITypedExpr(Kind::kCast, std::move(type), {std::move(input)}) {}
  std::string toString() const override { return "cast(" + inputs()[0]->toString() + " as " + type()->toString() + ")"; }
};

// ---- plan nodes --------------------------------------------------------------------------------------
using PlanNodeId = std::string;
class PlanNode;
using PlanNodePtr = std::shared_ptr<const PlanNode>;

class PlanNode {
 public:
  PlanNode(PlanNodeId id, std::vector<PlanNodePtr> sources) : id_(std::move(id)), sources_(std::move(sources)) {}
  virtual ~PlanNode() = default;
  const PlanNodeId& id() const { return id_; }
  const std::vector<PlanNodePtr>& sources() const { return sources_; }
  virtual const RowTypePtr& outputType() const = 0;
  virtual std::string name() const = 0;
 private:
  PlanNodeId id_;
  std::vector<PlanNodePtr> sources_;
};

class ValuesNode : public PlanNode {
 public:
  ValuesNode(PlanNodeId id, std::vector<RowVectorPtr> values) : PlanNode(std::move(id), {}), values_(std::move(values)) {
    VB_CHECK(!values_.empty(), "ValuesNode needs at least one vector");
    type_ = values_[0]->rowType();
  }
  const std::vector<RowVectorPtr>& values() const { return values_; }
  const RowTypePtr& outputType() const override { return type_; }
  std::string name() const override { return "Values"; }
 private:
  std::vector<RowVectorPtr> values_;
  RowTypePtr type_;
};

class FilterNode : public PlanNode {
 public:
  FilterNode(PlanNodeId id, TypedExprPtr filter, PlanNodePtr source) : PlanNode(std::move(id), {source}), filter_(std::move(filter)) {
    VB_USER_CHECK(filter_->type()->kind() == TypeKind::BOOLEAN, "Filter expression must be of type BOOLEAN");
  }
  const TypedExprPtr& filter() const { return filter_; }
  const RowTypePtr& outputType() const override { return sources()[0]->outputType(); }
  std::string name() const override { return "Filter"; }
 private:
  TypedExprPtr filter_;
};

class ProjectNode : public PlanNode {
 public:
  ProjectNode(PlanNodeId id, std::vector<std::string> names, std::vector<TypedExprPtr> projections, PlanNodePtr source)
      : PlanNode(std::move(id), {source}), names_(std::move(names)), projections_(std::move(projections)) {
    std::vector<TypePtr> types;
    for (auto& p : projections_) types.push_back(p->type());
    type_ = ROW(names_, types);
  }
  const std::vector<std::string>& names() const { return names_; }
  const std::vector<TypedExprPtr>& projections() const { return projections_; }
  const RowTypePtr& outputType() const override { return type_; }
  std::string name() const override { return "Project"; }
 private:
  std::vector<std::string> names_;
  std::vector<TypedExprPtr> projections_;
  RowTypePtr type_;
};

class AggregationNode : public PlanNode {
 public:
  enum class Step { kPartial, kFinal, kIntermediate, kSingle };
  struct Aggregate {
    std::string function;                 // sum count avg min max
    std::vector<TypedExprPtr> args;       // field accesses (or one constant for count(1)); empty = count(*)
    TypePtr rawInputType;                 // type of the ORIGINAL argument (final steps resolve result types from it)
    // Optional BOOLEAN input column: only rows where it is TRUE feed this aggregate (raw-input steps only;
    // core::AggregationNode::aggregateMasks, velox/exec/AggregationMasks.cpp).  nullptr = no mask.
    TypedExprPtr mask;
  };
  AggregationNode(PlanNodeId id, Step step, std::vector<std::string> groupingKeys, std::vector<std::string> aggregateNames,
                  std::vector<Aggregate> aggregates, bool ignoreNullKeys, PlanNodePtr source);
  Step step() const { return step_; }
  const std::vector<std::string>& groupingKeys() const { return groupingKeys_; }
  const std::vector<std::string>& aggregateNames() const { return aggregateNames_; }
  const std::vector<Aggregate>& aggregates() const { return aggregates_; }
  bool ignoreNullKeys() const { return ignoreNullKeys_; }
  const RowTypePtr& outputType() const override { return type_; }
  std::string name() const override { return "Aggregation"; }
  static bool isRawInput(Step s) { return s == Step::kPartial || s == Step::kSingle; }        // velox/exec/Aggregate.h:35-41
  static bool isPartialOutput(Step s) { return s == Step::kPartial || s == Step::kIntermediate; }
  // Result / intermediate types of the Presto aggregates on the path
  // (SumAggregate.cpp:24-108, CountAggregate.cpp:27-149, AverageAggregate.cpp:33-141, MinMaxAggregates.cpp).
  static TypePtr intermediateType(const std::string& fn, const TypePtr& raw);
  static TypePtr finalType(const std::string& fn, const TypePtr& raw);
 private:
  Step step_;
  std::vector<std::string> groupingKeys_, aggregateNames_;
  std::vector<Aggregate> aggregates_;
  bool ignoreNullKeys_;
  RowTypePtr type_;
};

// velox/core/PlanNode.h OrderByNode: sorting keys with ascending / nulls-first flags.  The B200 path ends at the
// aggregation; this node exists so that TpchQueryBuilder's plans (…finalAggregation().orderBy(keys, false),
// velox/exec/tests/utils/TpchQueryBuilder.cpp:232) can be written unchanged — whoever runs the plan supplies the
// operator (Velox's own exec::OrderBy in a deployment; tests/host registers a stand-in through a PlanNodeTranslator).
struct SortOrder { bool ascending = true; bool nullsFirst = true; };
class OrderByNode : public PlanNode {
 public:
  OrderByNode(PlanNodeId id, std::vector<std::string> keys, std::vector<SortOrder> orders, bool isPartial, PlanNodePtr source)
      : PlanNode(std::move(id), {source}), keys_(std::move(keys)), orders_(std::move(orders)), isPartial_(isPartial) {
    for (auto& k : keys_) sources()[0]->outputType()->getChildIdx(k); // throws on unknown columns
  }
  const std::vector<std::string>& sortingKeys() const { return keys_; }
  const std::vector<SortOrder>& sortingOrders() const { return orders_; }
  bool isPartial() const { return isPartial_; }
  const RowTypePtr& outputType() const override { return sources()[0]->outputType(); }
  std::string name() const override { return "OrderBy"; }
 private:
  std::vector<std::string> keys_;
  std::vector<SortOrder> orders_;
  bool isPartial_;
};

} // namespace core

// ---- expression text -> typed expression ----------------------------------------------------------------
namespace parse {
// Parses one SQL expression over the columns of `rowType`.  A trailing "AS name" is returned in *alias.
core::TypedExprPtr parseExpr(const std::string& text, const RowTypePtr& rowType, std::string* alias = nullptr);
// 'YYYY-MM-DD' -> days since 1970-01-01 (velox/type/Date / DATE()->toDays)
int32_t parseDate(const std::string& s);
} // namespace parse

// ---- PlanBuilder ------------------------------------------------------------------------------------------
class PlanBuilder {
 public:
  PlanBuilder& values(std::vector<RowVectorPtr> vectors);
  PlanBuilder& filter(const std::string& expr);
  PlanBuilder& project(const std::vector<std::string>& exprs);
  // masks: per aggregate, the name of a BOOLEAN column or "" (PlanBuilder::partialAggregation(keys, aggregates, masks))
  PlanBuilder& partialAggregation(const std::vector<std::string>& keys, const std::vector<std::string>& aggregates,
                                  const std::vector<std::string>& masks = {}) {
    return aggregation(keys, aggregates, core::AggregationNode::Step::kPartial, false, masks);
  }
  PlanBuilder& singleAggregation(const std::vector<std::string>& keys, const std::vector<std::string>& aggregates,
                                 const std::vector<std::string>& masks = {}) {
    return aggregation(keys, aggregates, core::AggregationNode::Step::kSingle, false, masks);
  }
  PlanBuilder& intermediateAggregation();
  PlanBuilder& finalAggregation();
  PlanBuilder& aggregation(const std::vector<std::string>& keys, const std::vector<std::string>& aggregates,
                           core::AggregationNode::Step step, bool ignoreNullKeys, const std::vector<std::string>& masks = {});
  // keys: "name [ASC|DESC] [NULLS FIRST|NULLS LAST]" (PlanBuilder::orderBy, velox/exec/tests/utils/PlanBuilder.h)
  PlanBuilder& orderBy(const std::vector<std::string>& keys, bool isPartial);
  // custom nodes (PlanBuilder::addNode, velox/exec/tests/utils/PlanBuilder.h:787-792)
  PlanBuilder& addNode(std::function<core::PlanNodePtr(std::string nodeId, core::PlanNodePtr source)> make);
  core::PlanNodePtr planNode() const { return node_; }
 private:
  std::string nextId() { return std::to_string(nextId_++); }
  PlanBuilder& followUp(core::AggregationNode::Step step);
  core::PlanNodePtr node_;
  int nextId_ = 0;
};

} // namespace vb200
