// Typed expressions -> one vgpu program (include/vgpu.h): the ExprSet compile step of the GPU
// operators (velox/exec/FilterProject.cpp:59-103 builds an ExprSet; velox/expression/ExprCompiler.cpp
// :54-58,160-163 shares common sub-expressions, :131-158 flattens AND/OR — both are kept here).
#pragma once
#include <map>
#include <set>

#include "../../include/vgpu.h"
#include "Plan.h"

namespace vb200 {

struct LoweredProgram {
  std::vector<vgpu_op> ops;
  vgpu_operand filter{};                // kind NONE when there is no filter
  std::vector<vgpu_operand> outputs;
  std::vector<std::string> columns;     // input column names, in program column order
  // VARCHAR equality (eq / neq / IN) runs on value ids: literal k of this list has id k, and the operator must bind the
  // columns named in stringColumns through ONE string -> id dictionary that starts with these literals.
  std::vector<std::string> stringLiterals;
  std::set<std::string> stringColumns;
  // Predicates over ONE VARCHAR column and a literal (like, lt, lte, gt, gte): the operator evaluates them on the HOST,
  // once per BASE string of the batch's dictionary (the reference's peeled evaluation, velox/expression/PeeledEncoding.cpp:
  // the result of a deterministic function over a dictionary is a dictionary over its results), and binds the outcome
  // as a BOOLEAN column that keeps the batch's indices and nulls.  The program reads it as the input column
  // `virtualName` = "$sp<k>".
  struct StringPredicate { std::string virtualName, column, fn, literal; };
  std::vector<StringPredicate> stringPredicates;
  const StringPredicate* predicateOf(const std::string& virtualName) const {
    for (auto& p : stringPredicates) if (p.virtualName == virtualName) return &p;
    return nullptr;
  }
  // SQL LIKE with % and _ (no escape character) / bytewise ordering like StringView::compare.
  static bool evalStringPredicate(const std::string& fn, const std::string& value, const std::string& literal);
  vgpu_program_desc desc() const {
    vgpu_program_desc d{};
    d.ops = ops.data(); d.n_ops = (int32_t)ops.size(); d.filter = filter;
    d.outputs = outputs.data(); d.n_outputs = (int32_t)outputs.size(); d.n_columns = (int32_t)columns.size();
    return d;
  }
};

// vgpu_type of a Velox type on the device (DATE = INTEGER; VARCHAR travels as INTEGER value ids).
int deviceType(const TypePtr& t);

// Lowers `filter` (may be null) and `outputs`.  Columns are numbered in order of first use.
LoweredProgram lowerExpressions(const core::TypedExprPtr& filter, const std::vector<core::TypedExprPtr>& outputs);

} // namespace vb200
