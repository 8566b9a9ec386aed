#include "Plan.h"

#include <cctype>
#include <cmath>

namespace vb200 {
namespace core {

TypePtr AggregationNode::intermediateType(const std::string& fn, const TypePtr& raw) {
  if (fn == "count" || fn == "count_if") return BIGINT(); // CountIfAggregate.cpp: BIGINT accumulator, never NULL
  if (fn == "sum") return raw->isFloating() ? DOUBLE() : BIGINT(); // sum(real) keeps a double accumulator
  if (fn == "avg") return ROW({"sum", "count"}, {DOUBLE(), BIGINT()});
  if (fn == "min" || fn == "max") return raw;
  if (fn == "bool_and" || fn == "bool_or") return BOOLEAN(); // functions/prestosql/aggregates/BoolAggregates.cpp
  throw VeloxUserError("Aggregate function not registered: " + fn);
}
TypePtr AggregationNode::finalType(const std::string& fn, const TypePtr& raw) {
  if (fn == "count" || fn == "count_if") return BIGINT(); // CountIfAggregate.cpp: BIGINT accumulator, never NULL
  if (fn == "sum") return raw->kind() == TypeKind::REAL ? REAL() : raw->isFloating() ? DOUBLE() : BIGINT();
  if (fn == "avg") return raw->kind() == TypeKind::REAL ? REAL() : DOUBLE();
  if (fn == "min" || fn == "max") return raw;
  if (fn == "bool_and" || fn == "bool_or") return BOOLEAN();
  throw VeloxUserError("Aggregate function not registered: " + fn);
}

AggregationNode::AggregationNode(PlanNodeId id, Step step, std::vector<std::string> groupingKeys, std::vector<std::string> aggregateNames,
                                 std::vector<Aggregate> aggregates, bool ignoreNullKeys, PlanNodePtr source)
    : PlanNode(std::move(id), {source}), step_(step), groupingKeys_(std::move(groupingKeys)), aggregateNames_(std::move(aggregateNames)),
      aggregates_(std::move(aggregates)), ignoreNullKeys_(ignoreNullKeys) {
  const auto& in = source->outputType();
  std::vector<std::string> names;
  std::vector<TypePtr> types;
  for (auto& k : groupingKeys_) { names.push_back(k); types.push_back(in->childAt(in->getChildIdx(k))); }
  for (size_t i = 0; i < aggregates_.size(); ++i) {
    const auto& a = aggregates_[i];
    names.push_back(aggregateNames_[i]);
    const TypePtr raw = a.rawInputType ? a.rawInputType : BIGINT();
    types.push_back(isPartialOutput(step_) ? intermediateType(a.function, raw) : finalType(a.function, raw));
  }
  type_ = ROW(std::move(names), std::move(types));
}

} // namespace core

// ======================================================================================================
// Expression parser (recursive descent over the DuckParser subset the path needs)
// ======================================================================================================
namespace parse {
using namespace core;

int32_t parseDate(const std::string& s) {
  int y = 0, m = 0, d = 0;
  VB_USER_CHECK(sscanf(s.c_str(), "%d-%d-%d", &y, &m, &d) == 3 && m >= 1 && m <= 12 && d >= 1 && d <= 31, "Unable to parse date value: \"" + s + "\"");
  // days from civil (proleptic Gregorian)
  y -= m <= 2;
  const int era = (y >= 0 ? y : y - 399) / 400;
  const unsigned yoe = (unsigned)(y - era * 400);
  const unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + (int)doe - 719468;
}

namespace {

TypedExprPtr constInt(TypePtr t, int64_t v) { return std::make_shared<ConstantTypedExpr>(std::move(t), false, v, (double)v); }
TypedExprPtr constDouble(TypePtr t, double v) { return std::make_shared<ConstantTypedExpr>(std::move(t), false, (int64_t)v, v); }
TypedExprPtr constNull(TypePtr t) { return std::make_shared<ConstantTypedExpr>(std::move(t), true, 0, 0.0); }
const ConstantTypedExpr* asConst(const TypedExprPtr& e) { return e->kind() == ITypedExpr::Kind::kConstant ? static_cast<const ConstantTypedExpr*>(e.get()) : nullptr; }

TypePtr typeByName(std::string n) {
  for (auto& c : n) c = (char)toupper(c);
  if (n == "BOOLEAN") return BOOLEAN();
  if (n == "TINYINT") return TINYINT();
  if (n == "SMALLINT") return SMALLINT();
  if (n == "INTEGER" || n == "INT") return INTEGER();
  if (n == "BIGINT") return BIGINT();
  if (n == "REAL" || n == "FLOAT") return REAL();
  if (n == "DOUBLE") return DOUBLE();
  if (n == "VARCHAR") return VARCHAR();
  if (n == "DATE") return DATE();
  throw VeloxUserError("Unknown type: " + n);
}

// cast with constant folding (ExprCompiler.cpp:302-335 folds constant sub-trees at compile time)
TypedExprPtr makeCast(const TypedExprPtr& e, const TypePtr& to) {
  if (e->type()->equals(*to)) return e;
  if (auto c = asConst(e)) {
    if (c->isNull()) return constNull(to);
    const auto from = e->type();
    if (from->kind() == TypeKind::VARCHAR && to->isDate()) return constInt(to, parseDate(c->stringValue()));
    if (from->kind() != TypeKind::VARCHAR && to->kind() != TypeKind::VARCHAR) {
      if (to->isFloating()) return constDouble(to, from->isFloating() ? c->doubleValue() : (double)c->intValue());
      if (from->isFloating()) { // Conversions.h:237-268: round half away from zero, range-checked at runtime for columns
        const double r = std::round(c->doubleValue());
        VB_USER_CHECK(std::isfinite(r), "Cannot cast NaN/Infinity to integer");
        return constInt(to, (int64_t)r);
      }
      return constInt(to, c->intValue());
    }
  }
  VB_USER_CHECK(e->type()->kind() != TypeKind::VARCHAR && to->kind() != TypeKind::VARCHAR, "cast between VARCHAR and " + to->toString() + " is only supported for literals");
  return std::make_shared<CastTypedExpr>(to, e);
}

class Parser {
 public:
  Parser(const std::string& text, const RowTypePtr& row) : s_(text), row_(row) {}
  TypedExprPtr parseTop(std::string* alias) {
    auto e = parseOr();
    if (acceptWord("as")) { const std::string a = ident(); if (alias) *alias = a; }
    skipWs();
    VB_USER_CHECK(pos_ == s_.size(), "Unexpected text after expression: '" + s_.substr(pos_) + "' in '" + s_ + "'");
    return e;
  }

 private:
  void skipWs() { while (pos_ < s_.size() && isspace((unsigned char)s_[pos_])) ++pos_; }
  bool accept(const char* tok) {
    skipWs();
    const size_t n = strlen(tok);
    if (s_.compare(pos_, n, tok) == 0) { pos_ += n; return true; }
    return false;
  }
  bool peekWord(const char* w) {
    skipWs();
    const size_t n = strlen(w);
    if (pos_ + n > s_.size()) return false;
    for (size_t i = 0; i < n; ++i) if (tolower((unsigned char)s_[pos_ + i]) != w[i]) return false;
    return pos_ + n == s_.size() || !(isalnum((unsigned char)s_[pos_ + n]) || s_[pos_ + n] == '_');
  }
  bool acceptWord(const char* w) { if (peekWord(w)) { pos_ += strlen(w); return true; } return false; }
  std::string ident() {
    skipWs();
    size_t b = pos_;
    while (pos_ < s_.size() && (isalnum((unsigned char)s_[pos_]) || s_[pos_] == '_')) ++pos_;
    VB_USER_CHECK(pos_ > b, "Identifier expected at '" + s_.substr(b) + "'");
    return s_.substr(b, pos_ - b);
  }

  // Binary call with literal coercion: a literal operand takes the other operand's type.
  TypedExprPtr binary(const std::string& fn, TypedExprPtr a, TypedExprPtr b, bool boolResult) {
    if (!a->type()->equals(*b->type())) {
      if (asConst(b)) b = makeCast(b, a->type());
      else if (asConst(a)) a = makeCast(a, b->type());
    }
    VB_USER_CHECK(a->type()->equals(*b->type()), "Scalar function signature is not supported: " + fn + "(" + a->type()->toString() + ", " + b->type()->toString() + ")");
    const auto ca = asConst(a), cb = asConst(b);
    if (!boolResult && ca && cb && !ca->isNull() && !cb->isNull()) { // constant folding of literal arithmetic
      if (a->type()->isFloating()) {
        const double x = ca->doubleValue(), y = cb->doubleValue();
        return constDouble(a->type(), fn == "plus" ? x + y : fn == "minus" ? x - y : fn == "multiply" ? x * y : x / y);
      }
      const int64_t x = ca->intValue(), y = cb->intValue();
      if (fn != "divide" && fn != "mod") return constInt(a->type(), fn == "plus" ? x + y : fn == "minus" ? x - y : x * y);
    }
    return std::make_shared<CallTypedExpr>(boolResult ? BOOLEAN() : a->type(), std::vector<TypedExprPtr>{a, b}, fn);
  }
  static TypedExprPtr conjunct(const std::string& fn, TypedExprPtr a, TypedExprPtr b) {
    VB_USER_CHECK(a->type()->kind() == TypeKind::BOOLEAN && b->type()->kind() == TypeKind::BOOLEAN, fn + " expects BOOLEAN arguments");
    std::vector<TypedExprPtr> args; // flattened like ExprCompiler.cpp:131-158
    for (auto& x : {a, b}) {
      auto c = x->kind() == ITypedExpr::Kind::kCall ? static_cast<const CallTypedExpr*>(x.get()) : nullptr;
      if (c && c->name() == fn) args.insert(args.end(), c->inputs().begin(), c->inputs().end());
      else args.push_back(x);
    }
    return std::make_shared<CallTypedExpr>(BOOLEAN(), std::move(args), fn);
  }

  // Brings the value branches of if / switch / coalesce to one type: literals (and the untyped NULL) take the type
  // of the first non-literal branch, or of the first typed literal.
  static void unifyBranches(std::vector<TypedExprPtr*> branches, const std::string& what) {
    TypePtr t;
    for (auto* b : branches) if (!asConst(*b)) { t = (*b)->type(); break; }
    if (!t) for (auto* b : branches) { auto c = asConst(*b); if (c && !c->isNull()) { t = (*b)->type(); break; } }
    if (!t) t = (*branches[0])->type();
    for (auto* b : branches) {
      if (!(*b)->type()->equals(*t)) {
        VB_USER_CHECK(asConst(*b) != nullptr, "Scalar function signature is not supported: " + what + " with branches of " + t->toString() + " and " + (*b)->type()->toString());
        *b = asConst(*b)->isNull() ? constNull(t) : makeCast(*b, t);
      }
    }
  }
  // if(c, a[, b]) / switch(c1, v1, c2, v2, ...[, else]) — the names DuckParser.cpp gives CASE expressions
  // (special forms of velox/expression/SwitchExpr.cpp).
  static TypedExprPtr makeSwitch(std::vector<TypedExprPtr> args) {
    VB_USER_CHECK(args.size() >= 2, "if / switch takes at least a condition and a value");
    std::vector<TypedExprPtr*> values;
    for (size_t i = 0; i + 1 < args.size(); i += 2) {
      VB_USER_CHECK(args[i]->type()->kind() == TypeKind::BOOLEAN, "Condition of if / switch must be BOOLEAN");
      values.push_back(&args[i + 1]);
    }
    if (args.size() % 2 == 1) values.push_back(&args.back());
    unifyBranches(values, "switch");
    const TypePtr t = (*values[0])->type();
    const std::string name = args.size() <= 3 ? "if" : "switch";
    return std::make_shared<CallTypedExpr>(t, std::move(args), name);
  }
  static TypedExprPtr makeCoalesce(std::vector<TypedExprPtr> args) {
    VB_USER_CHECK(args.size() >= 2, "coalesce takes at least 2 arguments");
    std::vector<TypedExprPtr*> values;
    for (auto& a : args) values.push_back(&a);
    unifyBranches(values, "coalesce");
    const TypePtr t = args[0]->type();
    return std::make_shared<CallTypedExpr>(t, std::move(args), "coalesce");
  }

  TypedExprPtr parseOr() {
    auto e = parseAnd();
    while (acceptWord("or")) e = conjunct("or", e, parseAnd());
    return e;
  }
  TypedExprPtr parseAnd() {
    auto e = parseNot();
    while (acceptWord("and")) e = conjunct("and", e, parseNot());
    return e;
  }
  TypedExprPtr parseNot() {
    if (acceptWord("not")) {
      auto e = parseNot();
      VB_USER_CHECK(e->type()->kind() == TypeKind::BOOLEAN, "not expects a BOOLEAN argument");
      return std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{e}, "not");
    }
    return parseCmp();
  }
  TypedExprPtr parseCmp() {
    auto a = parseAdd();
    if (acceptWord("between")) {
      auto lo = parseAdd();
      VB_USER_CHECK(acceptWord("and"), "between: AND expected");
      auto hi = parseAdd();
      lo = makeCast(lo, a->type());
      hi = makeCast(hi, a->type());
      return std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{a, lo, hi}, "between");
    }
    {
      // x [NOT] IN (v1, v2, ...): a literal list becomes OR of eq — what the device evaluates
      // (velox/functions/prestosql/InPredicate.cpp builds a set; same three-valued result)
      const size_t save = pos_;
      const bool negated = acceptWord("not");
      if (acceptWord("in")) {
        VB_USER_CHECK(accept("("), "IN: '(' expected");
        TypedExprPtr any;
        do {
          auto eq = binary("eq", a, parseAdd(), true);
          any = any ? conjunct("or", any, eq) : eq;
        } while (accept(","));
        VB_USER_CHECK(accept(")"), "IN: ')' expected");
        return negated ? std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{any}, "not") : any;
      }
      pos_ = save;
    }
    {
      // x [NOT] LIKE 'pattern' (DuckParser.cpp maps the ~~ operator / LIKE to the function "like")
      const size_t save = pos_;
      const bool negated = acceptWord("not");
      if (acceptWord("like")) {
        auto pattern = parseAdd();
        VB_USER_CHECK(a->type()->kind() == TypeKind::VARCHAR && pattern->type()->kind() == TypeKind::VARCHAR, "like expects VARCHAR arguments");
        TypedExprPtr e = std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{a, pattern}, "like");
        return negated ? std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{e}, "not") : e;
      }
      pos_ = save;
    }
    if (acceptWord("is")) {
      const bool neg = acceptWord("not");
      VB_USER_CHECK(acceptWord("null"), "IS [NOT] NULL expected");
      TypedExprPtr e = std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{a}, "is_null");
      return neg ? std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{e}, "not") : e;
    }
    static const std::pair<const char*, const char*> ops[] = {{"<>", "neq"}, {"!=", "neq"}, {"<=", "lte"}, {">=", "gte"}, {"=", "eq"}, {"<", "lt"}, {">", "gt"}};
    for (auto& [tok, fn] : ops) if (accept(tok)) return binary(fn, a, parseAdd(), true);
    return a;
  }
  TypedExprPtr parseAdd() {
    auto e = parseMul();
    for (;;) {
      if (accept("+")) e = binary("plus", e, parseMul(), false);
      else if (accept("-")) e = binary("minus", e, parseMul(), false);
      else return e;
    }
  }
  TypedExprPtr parseMul() {
    auto e = parseUnary();
    for (;;) {
      if (accept("*")) e = binary("multiply", e, parseUnary(), false);
      else if (accept("/")) e = binary("divide", e, parseUnary(), false);
      else if (accept("%")) e = binary("mod", e, parseUnary(), false);
      else return e;
    }
  }
  TypedExprPtr parseUnary() {
    if (accept("-")) {
      auto e = parseUnary();
      if (auto c = asConst(e)) return e->type()->isFloating() ? constDouble(e->type(), -c->doubleValue()) : constInt(e->type(), -c->intValue());
      return std::make_shared<CallTypedExpr>(e->type(), std::vector<TypedExprPtr>{e}, "negate");
    }
    return parsePostfix();
  }
  TypedExprPtr parsePostfix() {
    auto e = parsePrimary();
    while (accept("::")) e = makeCast(e, typeByName(ident()));
    return e;
  }
  TypedExprPtr parsePrimary() {
    skipWs();
    VB_USER_CHECK(pos_ < s_.size(), "Unexpected end of expression: " + s_);
    const char c = s_[pos_];
    if (c == '(') { ++pos_; auto e = parseOr(); VB_USER_CHECK(accept(")"), "')' expected in " + s_); return e; }
    if (c == '\'') {
      const size_t e = s_.find('\'', pos_ + 1);
      VB_USER_CHECK(e != std::string::npos, "Unterminated string literal in " + s_);
      std::string v = s_.substr(pos_ + 1, e - pos_ - 1);
      pos_ = e + 1;
      return std::make_shared<ConstantTypedExpr>(VARCHAR(), false, 0, 0.0, v);
    }
    if (isdigit((unsigned char)c) || c == '.') {
      size_t b = pos_;
      bool fp = false;
      while (pos_ < s_.size() && (isdigit((unsigned char)s_[pos_]) || s_[pos_] == '.' || s_[pos_] == 'e' || s_[pos_] == 'E' ||
                                  ((s_[pos_] == '+' || s_[pos_] == '-') && (s_[pos_ - 1] == 'e' || s_[pos_ - 1] == 'E')))) {
        fp = fp || !isdigit((unsigned char)s_[pos_]);
        ++pos_;
      }
      const std::string num = s_.substr(b, pos_ - b);
      return fp ? constDouble(DOUBLE(), std::stod(num)) : constInt(BIGINT(), std::stoll(num)); // DuckDB literal typing
    }
    if (acceptWord("null")) return constNull(BIGINT());
    if (acceptWord("true")) return constInt(BOOLEAN(), 1);
    if (acceptWord("false")) return constInt(BOOLEAN(), 0);
    if (acceptWord("cast")) {
      VB_USER_CHECK(accept("("), "cast: '(' expected");
      auto e = parseOr();
      VB_USER_CHECK(acceptWord("as"), "cast: AS expected");
      auto t = typeByName(ident());
      VB_USER_CHECK(accept(")"), "cast: ')' expected");
      return makeCast(e, t);
    }
    if (acceptWord("case")) { // CASE WHEN c THEN v ... [ELSE v] END
      std::vector<TypedExprPtr> args;
      while (acceptWord("when")) {
        args.push_back(parseOr());
        VB_USER_CHECK(acceptWord("then"), "case: THEN expected in " + s_);
        args.push_back(parseOr());
      }
      VB_USER_CHECK(!args.empty(), "case: WHEN expected in " + s_);
      if (acceptWord("else")) args.push_back(parseOr());
      VB_USER_CHECK(acceptWord("end"), "case: END expected in " + s_);
      return makeSwitch(std::move(args));
    }
    if (acceptWord("date")) { // DATE 'yyyy-mm-dd'
      auto lit = parsePrimary();
      return makeCast(lit, DATE());
    }
    const std::string name = ident();
    if (accept("(")) { // named scalar function call: the registry names themselves
      std::vector<TypedExprPtr> args;
      if (!accept(")")) { do args.push_back(parseOr()); while (accept(",")); VB_USER_CHECK(accept(")"), "')' expected after arguments of " + name); }
      static const char* arith[] = {"plus", "minus", "multiply", "divide", "mod"};
      static const char* cmp[] = {"eq", "neq", "lt", "lte", "gt", "gte"};
      for (auto f : arith) if (name == f) { VB_USER_CHECK(args.size() == 2, name + " takes 2 arguments"); return binary(name, args[0], args[1], false); }
      for (auto f : cmp) if (name == f) { VB_USER_CHECK(args.size() == 2, name + " takes 2 arguments"); return binary(name, args[0], args[1], true); }
      if (name == "negate" && args.size() == 1) return std::make_shared<CallTypedExpr>(args[0]->type(), args, "negate");
      if ((name == "not" || name == "is_null") && args.size() == 1) return std::make_shared<CallTypedExpr>(BOOLEAN(), args, name);
      if (name == "between" && args.size() == 3) return std::make_shared<CallTypedExpr>(BOOLEAN(), std::vector<TypedExprPtr>{args[0], makeCast(args[1], args[0]->type()), makeCast(args[2], args[0]->type())}, "between");
      if (name == "if" || name == "switch") return makeSwitch(std::move(args));
      if (name == "coalesce") return makeCoalesce(std::move(args));
      throw VeloxUserError("Scalar function doesn't exist: " + name);
    }
    const int idx = row_->getChildIdx(name);
    return std::make_shared<FieldAccessTypedExpr>(row_->childAt(idx), name);
  }

  const std::string& s_;
  RowTypePtr row_;
  size_t pos_ = 0;
};

} // namespace

TypedExprPtr parseExpr(const std::string& text, const RowTypePtr& rowType, std::string* alias) {
  return Parser(text, rowType).parseTop(alias);
}

} // namespace parse
} // namespace vb200
