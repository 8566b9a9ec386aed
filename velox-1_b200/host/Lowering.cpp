#include "Lowering.h"

#include <set>

namespace vb200 {
using namespace core;

int deviceType(const TypePtr& t) {
  switch (t->kind()) {
    case TypeKind::BOOLEAN: return VGPU_BOOLEAN;
    case TypeKind::TINYINT: return VGPU_TINYINT;
    case TypeKind::SMALLINT: return VGPU_SMALLINT;
    case TypeKind::INTEGER: return VGPU_INTEGER;
    case TypeKind::BIGINT: return VGPU_BIGINT;
    case TypeKind::REAL: return VGPU_REAL;
    case TypeKind::DOUBLE: return VGPU_DOUBLE;
    case TypeKind::VARCHAR: return VGPU_INTEGER; // value ids
    default: throw VeloxRuntimeError("type " + t->toString() + " has no device representation");
  }
}

namespace {

struct Value {            // what gen() hands back
  vgpu_operand opnd{};
  bool hasImm = false;
  vgpu_scalar imm{};
};

int opcodeOf(const std::string& fn) {
  static const std::map<std::string, int> m = {
      {"plus", VGPU_OP_ADD}, {"minus", VGPU_OP_SUB}, {"multiply", VGPU_OP_MUL}, {"divide", VGPU_OP_DIV}, {"mod", VGPU_OP_MOD},
      {"eq", VGPU_OP_EQ}, {"neq", VGPU_OP_NE}, {"lt", VGPU_OP_LT}, {"lte", VGPU_OP_LE}, {"gt", VGPU_OP_GT}, {"gte", VGPU_OP_GE}};
  auto it = m.find(fn);
  return it == m.end() ? -1 : it->second;
}

class Lowerer {
 public:
  LoweredProgram out;

  void count(const TypedExprPtr& e) {
    const int n = ++uses_[key(e)];
    if (n == 1 && isCall(e)) for (auto& a : e->inputs()) count(a);
  }

  Value gen(const TypedExprPtr& e, bool wantSlot) {
    if (!isCall(e)) return leaf(e);
    const std::string k = key(e);
    auto w = where_.find(k);
    if (w != where_.end()) { // shared value computed earlier
      Value v;
      v.opnd = opnd(VGPU_OPND_SLOT, deviceType(e->type()), w->second);
      if (--uses_[k] == 0) { pinned_.erase({isBool(e), w->second}); where_.erase(w); }
      return v;
    }
    const bool shared = uses_[k] > 1;
    Value res;
    const std::string fn = nameOf(e);
    res.opnd = (fn == "and" || fn == "or") ? genConjunct(e, fn) : genCall(e, fn);
    // a call that lowered to a plain input column (a string predicate bound by the operator) or a NULL is a leaf
    const bool leafLike = res.opnd.kind == VGPU_OPND_COL || res.opnd.kind == VGPU_OPND_NULL;
    if ((shared || wantSlot) && !leafLike) {
      if (res.opnd.kind != VGPU_OPND_SLOT) {
        const int slot = alloc(isBool(e));
        out.ops.back().dst_slot = (int16_t)slot;
        res.opnd = opnd(VGPU_OPND_SLOT, deviceType(e->type()), slot);
      }
      if (shared) {
        --uses_[k];
        where_[k] = res.opnd.index;
        pinned_.insert({isBool(e), res.opnd.index});
      }
    }
    return res;
  }

  void keep(const vgpu_operand& o) { if (o.kind == VGPU_OPND_SLOT) keep_.insert({o.type == VGPU_BOOLEAN, o.index}); }
  int alloc(bool b) {
    auto& pool = b ? freeB_ : freeV_;
    if (pool.empty()) throw VeloxRuntimeError("expression needs more than 12 scratch slots");
    const int s = pool.back();
    pool.pop_back();
    return s;
  }
  vgpu_operand emit(int opcode, int type, vgpu_operand a, vgpu_operand b, vgpu_operand c, const Value* immSrc, const Value* imm2Src) {
    vgpu_op op{};
    op.opcode = (uint8_t)opcode; op.type = (uint8_t)type; op.dst_slot = -1;
    op.a = a; op.b = b; op.c = c;
    if (immSrc && immSrc->hasImm) op.imm = immSrc->imm;
    if (imm2Src && imm2Src->hasImm) op.imm2 = imm2Src->imm;
    out.ops.push_back(op);
    const bool boolean = opcode >= VGPU_OP_EQ && opcode <= VGPU_OP_IS_NULL;
    return opnd(VGPU_OPND_ACC, boolean ? VGPU_BOOLEAN : type, 0);
  }
  static vgpu_operand opnd(int kind, int type, int index) {
    vgpu_operand o{};
    o.kind = (uint8_t)kind; o.type = (uint8_t)type; o.index = (uint16_t)index;
    return o;
  }

 private:
  static bool isCall(const TypedExprPtr& e) { return e->kind() == ITypedExpr::Kind::kCall || e->kind() == ITypedExpr::Kind::kCast; }
  static bool isBool(const TypedExprPtr& e) { return e->type()->kind() == TypeKind::BOOLEAN; }
  static std::string nameOf(const TypedExprPtr& e) {
    return e->kind() == ITypedExpr::Kind::kCast ? "cast" : static_cast<const CallTypedExpr*>(e.get())->name();
  }
  static std::string key(const TypedExprPtr& e) { return e->toString() + ":" + e->type()->toString(); }

  Value leaf(const TypedExprPtr& e) {
    Value v;
    const int t = deviceType(e->type());
    if (e->kind() == ITypedExpr::Kind::kFieldAccess) {
      const auto& name = static_cast<const FieldAccessTypedExpr*>(e.get())->name();
      int idx = -1;
      for (size_t i = 0; i < out.columns.size(); ++i) if (out.columns[i] == name) idx = (int)i;
      if (idx < 0) { idx = (int)out.columns.size(); out.columns.push_back(name); }
      v.opnd = opnd(VGPU_OPND_COL, t, idx);
      return v;
    }
    auto c = static_cast<const ConstantTypedExpr*>(e.get());
    if (c->isNull()) { v.opnd = opnd(VGPU_OPND_NULL, t, 0); return v; }
    if (e->type()->kind() == TypeKind::VARCHAR) { // the literal's value id (see LoweredProgram::stringLiterals)
      const std::string& sv = c->stringValue();
      size_t id = 0;
      while (id < out.stringLiterals.size() && out.stringLiterals[id] != sv) ++id;
      if (id == out.stringLiterals.size()) out.stringLiterals.push_back(sv);
      v.opnd = opnd(VGPU_OPND_IMM, t, 0);
      v.hasImm = true;
      v.imm.i = (int64_t)id;
      return v;
    }
    v.opnd = opnd(VGPU_OPND_IMM, t, 0);
    v.hasImm = true;
    if (e->type()->isFloating()) v.imm.d = c->doubleValue(); else v.imm.i = c->intValue();
    return v;
  }

  void release(const vgpu_operand& o) {
    if (o.kind != VGPU_OPND_SLOT) return;
    const std::pair<bool, int> k{o.type == VGPU_BOOLEAN, o.index};
    if (pinned_.count(k) || keep_.count(k)) return;
    (k.first ? freeB_ : freeV_).push_back(o.index);
  }

  // Arguments are evaluated so that at most the last non-leaf one stays in the accumulator.
  std::vector<Value> genArgs(const std::vector<TypedExprPtr>& args) {
    int lastCall = -1;
    for (size_t i = 0; i < args.size(); ++i) if (isCall(args[i]) && !where_.count(key(args[i]))) lastCall = (int)i;
    std::vector<Value> v(args.size());
    for (size_t i = 0; i < args.size(); ++i) v[i] = gen(args[i], isCall(args[i]) && (int)i != lastCall);
    return v;
  }

  // like / < / <= / > / >= between a VARCHAR column and a literal: a BOOLEAN virtual column (LoweredProgram::stringPredicates)
  bool stringPredicate(const TypedExprPtr& e, const std::string& fn, vgpu_operand& res) {
    if (fn != "like" && fn != "lt" && fn != "lte" && fn != "gt" && fn != "gte") return false;
    const auto& in = e->inputs();
    if (in.size() != 2 || in[0]->type()->kind() != TypeKind::VARCHAR) return false;
    if (in[0]->kind() != ITypedExpr::Kind::kFieldAccess || in[1]->kind() != ITypedExpr::Kind::kConstant)
      throw VeloxUserError("Scalar function " + fn + " over VARCHAR is supported between a column and a literal only");
    auto lit = static_cast<const ConstantTypedExpr*>(in[1].get());
    if (lit->isNull()) { res = opnd(VGPU_OPND_NULL, VGPU_BOOLEAN, 0); return true; }
    const std::string column = static_cast<const FieldAccessTypedExpr*>(in[0].get())->name();
    std::string name;
    for (auto& p : out.stringPredicates) if (p.column == column && p.fn == fn && p.literal == lit->stringValue()) name = p.virtualName;
    if (name.empty()) {
      name = "$sp" + std::to_string(out.stringPredicates.size());
      out.stringPredicates.push_back({name, column, fn, lit->stringValue()});
    }
    int idx = -1;
    for (size_t i = 0; i < out.columns.size(); ++i) if (out.columns[i] == name) idx = (int)i;
    if (idx < 0) { idx = (int)out.columns.size(); out.columns.push_back(name); }
    res = opnd(VGPU_OPND_COL, VGPU_BOOLEAN, idx);
    return true;
  }

  vgpu_operand genCall(const TypedExprPtr& e, const std::string& fn) {
    { vgpu_operand sp; if (stringPredicate(e, fn, sp)) return sp; }
    auto args = genArgs(e->inputs());
    const vgpu_operand none = opnd(VGPU_OPND_NONE, 0, 0);
    vgpu_operand res;
    const int code = opcodeOf(fn);
    for (auto& in : e->inputs()) {
      if (in->type()->kind() != TypeKind::VARCHAR || fn == "is_null") continue;
      // strings travel as value ids: equality is all that ids preserve
      if (fn != "eq" && fn != "neq") throw VeloxUserError("Scalar function " + fn + " over VARCHAR is not supported on the device (only =, <>, IN, IS NULL)");
      if (in->kind() == ITypedExpr::Kind::kFieldAccess) out.stringColumns.insert(static_cast<const FieldAccessTypedExpr*>(in.get())->name());
      else if (in->kind() != ITypedExpr::Kind::kConstant) throw VeloxUserError("computed VARCHAR values cannot be compared on the device");
    }
    if (code >= 0) {
      VB_CHECK(!(args[0].hasImm && args[1].hasImm), fn + " of two literals must be folded before lowering");
      const int t = deviceType(e->inputs()[0]->type());
      res = emit(code, t, args[0].opnd, args[1].opnd, none, args[0].hasImm ? &args[0] : &args[1], nullptr);
    } else if (fn == "between") {
      VB_CHECK(!args[0].hasImm, "between over a literal must be folded before lowering");
      res = emit(VGPU_OP_BETWEEN, deviceType(e->inputs()[0]->type()), args[0].opnd, args[1].opnd, args[2].opnd, &args[1], &args[2]);
    } else if (fn == "negate") {
      res = emit(VGPU_OP_NEG, deviceType(e->type()), args[0].opnd, none, none, &args[0], nullptr);
    } else if (fn == "cast") {
      res = emit(VGPU_OP_CAST, deviceType(e->type()), args[0].opnd, none, none, &args[0], nullptr);
    } else if (fn == "not") {
      res = emit(VGPU_OP_NOT, VGPU_BOOLEAN, args[0].opnd, none, none, &args[0], nullptr);
    } else if (fn == "is_null") {
      res = emit(VGPU_OP_IS_NULL, VGPU_BOOLEAN, args[0].opnd, none, none, &args[0], nullptr);
    } else if (fn == "if") { // switch arrives here as nested ifs (desugarSwitch)
      VB_CHECK(args.size() == 2 || args.size() == 3, "if takes (condition, then[, else])");
      VB_CHECK(!args[0].hasImm && args[0].opnd.kind != VGPU_OPND_NULL, "if over a constant condition must be folded before lowering");
      if (isBool(e)) throw VeloxUserError("BOOLEAN-valued if / switch is not supported on the device");
      const int t = deviceType(e->type());
      const vgpu_operand other = args.size() > 2 ? args[2].opnd : opnd(VGPU_OPND_NULL, t, 0);
      res = emit(VGPU_OP_SELECT, t, args[0].opnd, args[1].opnd, other, &args[1], args.size() > 2 ? &args[2] : nullptr);
    } else if (fn == "coalesce") {
      // coalesce(a, b, c) = coalesce(a, coalesce(b, c)): the arguments are already evaluated (slots / leaves),
      // the fold runs from the right with the running value in the accumulator
      if (isBool(e)) throw VeloxUserError("BOOLEAN-valued coalesce is not supported on the device");
      const int t = deviceType(e->type());
      vgpu_operand running = args.back().opnd;
      const Value* runningImm = &args.back();
      for (int k = (int)args.size() - 2; k >= 0; --k) {
        VB_CHECK(!args[k].hasImm, "coalesce with a literal before its last argument must be folded before lowering");
        running = emit(VGPU_OP_COALESCE, t, args[k].opnd, running, none, runningImm, nullptr);
        runningImm = nullptr;
      }
      res = running;
    } else {
      throw VeloxUserError("Scalar function is not supported on the device: " + fn);
    }
    for (auto& a : args) release(a.opnd);
    return res;
  }

  // AND / OR over n conjuncts: a running boolean slot combined with one conjunct at a time
  // (three-valued logic and error suppression live in the device op, ConjunctExpr.cpp:96-174).
  vgpu_operand genConjunct(const TypedExprPtr& e, const std::string& fn) {
    const int code = fn == "and" ? VGPU_OP_AND : VGPU_OP_OR;
    const auto& args = e->inputs();
    Value first = gen(args[0], isCall(args[0]));
    vgpu_operand running{};
    bool haveRunning = false;
    const vgpu_operand none = opnd(VGPU_OPND_NONE, 0, 0);
    for (size_t i = 1; i < args.size(); ++i) {
      Value x = gen(args[i], false);
      emit(code, VGPU_BOOLEAN, haveRunning ? running : first.opnd, x.opnd, none, x.hasImm ? &x : (haveRunning ? nullptr : &first), nullptr);
      if (!haveRunning) {
        release(first.opnd);
        running = opnd(VGPU_OPND_SLOT, VGPU_BOOLEAN, alloc(true));
        haveRunning = true;
      }
      if (x.opnd.kind == VGPU_OPND_SLOT && x.opnd.index != running.index) release(x.opnd);
      out.ops.back().dst_slot = (int16_t)running.index;
    }
    return haveRunning ? running : first.opnd;
  }

  std::map<std::string, int> uses_;
  std::map<std::string, int> where_;
  std::set<std::pair<bool, int>> pinned_, keep_;
  std::vector<int> freeV_{11, 10, 9, 8, 7, 6, 5, 4, 3, 2, 1, 0}, freeB_{11, 10, 9, 8, 7, 6, 5, 4, 3, 2, 1, 0};
};

} // namespace

// switch(c1, v1, c2, v2, ..., [else]) -> if(c1, v1, if(c2, v2, ... else)): SwitchExpr.cpp evaluates condition k on
// the rows no earlier condition took, which is what the nesting does.  Applied to the whole tree before lowering.
static TypedExprPtr desugarSwitch(const TypedExprPtr& e) {
  if (e->kind() != ITypedExpr::Kind::kCall && e->kind() != ITypedExpr::Kind::kCast) return e;
  std::vector<TypedExprPtr> in;
  bool changed = false;
  for (auto& x : e->inputs()) { in.push_back(desugarSwitch(x)); changed = changed || in.back() != x; }
  if (e->kind() == ITypedExpr::Kind::kCast) return changed ? std::make_shared<CastTypedExpr>(e->type(), in[0]) : e;
  const auto& name = static_cast<const CallTypedExpr*>(e.get())->name();
  if (name != "switch") return changed ? std::make_shared<CallTypedExpr>(e->type(), std::move(in), name) : e;
  VB_USER_CHECK(in.size() >= 2, "switch takes at least a condition and a value");
  TypedExprPtr tail = in.size() % 2 == 1 ? in.back() : nullptr;
  for (int k = (int)(in.size() / 2) - 1; k >= 0; --k) {
    std::vector<TypedExprPtr> args{in[2 * k], in[2 * k + 1]};
    if (tail) args.push_back(tail);
    tail = std::make_shared<CallTypedExpr>(e->type(), std::move(args), "if");
  }
  return tail;
}

bool LoweredProgram::evalStringPredicate(const std::string& fn, const std::string& v, const std::string& lit) {
  if (fn == "like") { // iterative wildcard match: % = any run of characters, _ = exactly one
    size_t i = 0, j = 0, star = std::string::npos, mark = 0;
    while (i < v.size()) {
      if (j < lit.size() && (lit[j] == '_' || lit[j] == v[i])) { ++i; ++j; }
      else if (j < lit.size() && lit[j] == '%') { star = j++; mark = i; }
      else if (star != std::string::npos) { j = star + 1; i = ++mark; }
      else return false;
    }
    while (j < lit.size() && lit[j] == '%') ++j;
    return j == lit.size();
  }
  const int c = v.compare(lit); // bytewise, like StringView::compare
  if (fn == "lt") return c < 0;
  if (fn == "lte") return c <= 0;
  if (fn == "gt") return c > 0;
  return c >= 0;
}

LoweredProgram lowerExpressions(const TypedExprPtr& filterIn, const std::vector<TypedExprPtr>& outputsIn) {
  Lowerer lw;
  const TypedExprPtr filter = filterIn ? desugarSwitch(filterIn) : filterIn;
  std::vector<TypedExprPtr> outputs;
  for (auto& o : outputsIn) outputs.push_back(desugarSwitch(o));
  if (filter) {
    VB_USER_CHECK(filter->type()->kind() == TypeKind::BOOLEAN, "Filter expression must be of type BOOLEAN");
    lw.count(filter);
  }
  for (auto& o : outputs) lw.count(o);
  lw.out.filter = Lowerer::opnd(VGPU_OPND_NONE, 0, 0);
  const vgpu_operand none = Lowerer::opnd(VGPU_OPND_NONE, 0, 0);
  if (filter) {
    const bool call = filter->kind() == ITypedExpr::Kind::kCall || filter->kind() == ITypedExpr::Kind::kCast;
    Value f = lw.gen(filter, call);
    VB_CHECK(f.opnd.kind != VGPU_OPND_IMM && f.opnd.kind != VGPU_OPND_NULL, "constant filters must be folded by the planner");
    lw.keep(f.opnd);
    lw.out.filter = f.opnd;
  }
  for (auto& o : outputs) {
    const bool call = o->kind() == ITypedExpr::Kind::kCall || o->kind() == ITypedExpr::Kind::kCast;
    Value v = lw.gen(o, call);
    if (v.opnd.kind == VGPU_OPND_IMM || v.opnd.kind == VGPU_OPND_NULL) { // literal output: MOV into a slot for the sink
      VB_CHECK(o->type()->kind() != TypeKind::BOOLEAN, "boolean literal outputs are not supported");
      const int slot = lw.alloc(false);
      lw.emit(VGPU_OP_MOV, deviceType(o->type()), v.opnd, none, none, &v, nullptr);
      lw.out.ops.back().dst_slot = (int16_t)slot;
      v.opnd = Lowerer::opnd(VGPU_OPND_SLOT, deviceType(o->type()), slot);
    }
    lw.keep(v.opnd);
    lw.out.outputs.push_back(v.opnd);
  }
  return std::move(lw.out);
}

} // namespace vb200
