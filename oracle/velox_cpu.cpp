// TEST INFRASTRUCTURE — CPU restatement of the reference's FilterProject -> HashAggregation path.
// It is the timed CPU baseline (bench.py --impl reference / cpu_baseline) and a second checker
// next to oracle/np_oracle.py.  The product never links it.
//
// It keeps the reference's execution shape so its timing means something:
//   * batch at a time, one "driver" thread per contiguous row range (exec/Driver.cpp:429-700),
//   * a tree-walking vectorised interpreter that materialises every intermediate and narrows a
//     SelectivityVector for default-null functions (expression/Expr.cpp:371-441, 1107-1144,
//     1349-1402) and for AND/OR (expression/ConjunctExpr.cpp:96-174, 217-245),
//   * processFilterResults -> selectedIndices, projections evaluated on the passing rows only and
//     handed on by index (exec/FilterProject.cpp:125-191, exec/OperatorUtils.cpp:187-255),
//   * HashAggregation with value-id "array" mode for small key domains and a row-pointer hash
//     table otherwise (exec/HashTable.cpp:426-562, 1376-1461; exec/VectorHasher.cpp:110-332),
//     group rows in a RowContainer-like arena (exec/RowContainer.cpp:115-252), accumulators
//     updated function by function over the batch in input order
//     (functions/lib/aggregates/SimpleNumericAggregate.h:90-142),
//   * partial aggregation per driver, then a single final merge (TpchQueryBuilder.cpp:220-232).
// Deliberate simplifications, none observable in results for the configs it is used on:
// nulls are a byte per row; integer arithmetic errors abort the query at once (the deferred-error
// suppression of ConjunctExpr.cpp:24-93 is covered by np_oracle.py); types are BOOLEAN, INTEGER,
// BIGINT, DOUBLE.
#include "velox_cpu.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace {

enum Kind { kBool = 0, kInt = 3, kBigint = 4, kDouble = 6 };
thread_local std::string g_err;
std::string g_last_error;

struct UserError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

// ---- SelectivityVector (vector/SelectivityVector.h:43-455): bits + [begin,end) + allSelected ----
struct Selectivity {
  std::vector<uint64_t> bits;
  int32_t size = 0, begin = 0, end = 0;
  bool all = true;
  void resizeAll(int32_t n) {
    size = n; begin = 0; end = n; all = true;
    bits.assign((n + 63) / 64, ~0ull);
    if (n & 63) bits.back() = (1ull << (n & 63)) - 1;
  }
  bool isValid(int32_t i) const { return (bits[i >> 6] >> (i & 63)) & 1; }
  void clear(int32_t i) { bits[i >> 6] &= ~(1ull << (i & 63)); }
  void updateBounds() { // :277-287
    begin = size; end = 0;
    int64_t count = 0;
    for (size_t w = 0; w < bits.size(); ++w) {
      if (!bits[w]) continue;
      if (begin == size) begin = (int32_t)(w * 64 + __builtin_ctzll(bits[w]));
      end = (int32_t)(w * 64 + 64 - __builtin_clzll(bits[w]));
      count += __builtin_popcountll(bits[w]);
    }
    if (end == 0) begin = 0;
    all = count == size;
  }
  bool hasSelections() const { return begin < end; }
  template <typename F>
  void forEach(F f) const { // applyToSelected :426-436
    if (all) {
      for (int32_t i = begin; i < end; ++i) f(i);
      return;
    }
    for (int32_t w = begin >> 6; w <= (end - 1) >> 6 && w < (int32_t)bits.size(); ++w) {
      uint64_t word = bits[w];
      while (word) {
        f(w * 64 + __builtin_ctzll(word));
        word &= word - 1;
      }
    }
  }
};

// ---- vectors: flat or constant; bool values are a byte per row ---------------------------------
struct Vec {
  int type = kDouble;
  bool constant = false, const_null = false;
  union { int64_t i; double d; } cval{};
  const void* data = nullptr;       // borrowed (input column) or -> own
  const uint8_t* nulls = nullptr;   // byte per row, 1 = NULL; nullptr = none
  std::vector<char> own;
  std::vector<uint8_t> own_nulls;
  void* alloc(int32_t n, size_t width) {
    own.resize((size_t)n * width);
    data = own.data();
    return own.data();
  }
  uint8_t* allocNulls(int32_t n) {
    own_nulls.assign(n, 0);
    nulls = own_nulls.data();
    return own_nulls.data();
  }
  template <typename T> const T* as() const { return reinterpret_cast<const T*>(data); }
  template <typename T> T at(int32_t i) const {
    if (constant) return std::is_same<T, double>::value ? (T)cval.d : (T)cval.i;
    return as<T>()[i];
  }
  bool isNullAt(int32_t i) const { return constant ? const_null : (nulls && nulls[i]); }
};

size_t widthOf(int type) { return type == kBool ? 1 : type == kInt ? 4 : 8; }

// ---- expression tree ------------------------------------------------------------------------
struct Expr {
  enum Op { kField, kConst, kCall, kAnd, kOr } op = kCall;
  int type = kDouble;
  std::string fn;
  int column = 0;
  bool is_null_const = false;
  union { int64_t i; double d; } value{};
  std::vector<std::unique_ptr<Expr>> args;
};

struct Parser {
  const char* p;
  explicit Parser(const char* s) : p(s) {}
  void ws() { while (*p == ' ' || *p == '\n' || *p == '\t') ++p; }
  static int typeOf(const std::string& t) {
    if (t == "b") return kBool;
    if (t == "i32") return kInt;
    if (t == "i64") return kBigint;
    if (t == "f64") return kDouble;
    throw std::runtime_error("bad type suffix '" + t + "'");
  }
  std::string token() {
    ws();
    const char* s = p;
    while (*p && *p != ' ' && *p != ')' && *p != '(' && *p != '\n') ++p;
    return std::string(s, p);
  }
  std::unique_ptr<Expr> parse() {
    ws();
    auto e = std::make_unique<Expr>();
    if (*p == '(') {
      ++p;
      e->fn = token();
      while (true) {
        ws();
        if (*p == ')') { ++p; break; }
        if (!*p) throw std::runtime_error("unterminated expression");
        e->args.push_back(parse());
      }
      finishCall(*e);
      return e;
    }
    std::string t = token();
    auto colon = t.rfind(':');
    if (colon == std::string::npos) throw std::runtime_error("atom without type: " + t);
    e->type = typeOf(t.substr(colon + 1));
    std::string v = t.substr(0, colon);
    if (v[0] == 'c' && v.size() > 1 && isdigit(v[1])) {
      e->op = Expr::kField;
      e->column = atoi(v.c_str() + 1);
    } else {
      e->op = Expr::kConst;
      if (v == "null") e->is_null_const = true;
      else if (e->type == kDouble) e->value.d = strtod(v.c_str(), nullptr);
      else if (v == "true") e->value.i = 1;
      else if (v == "false") e->value.i = 0;
      else e->value.i = strtoll(v.c_str(), nullptr, 10);
    }
    return e;
  }
  static void finishCall(Expr& e) {
    const std::string& f = e.fn;
    if (f == "and" || f == "or") {
      e.op = f == "and" ? Expr::kAnd : Expr::kOr;
      e.type = kBool;
    } else if (f == "plus" || f == "minus" || f == "multiply" || f == "divide" || f == "negate") {
      e.type = e.args.at(0)->type;
    } else if (f == "if" || f == "switch") {
      if (e.args.size() < 2) throw std::runtime_error("if / switch takes at least (condition, value)");
      e.type = e.args.at(1)->type;
    } else if (f == "coalesce") {
      e.type = e.args.at(0)->type;
    } else if (f == "eq" || f == "neq" || f == "lt" || f == "lte" || f == "gt" || f == "gte" || f == "between" ||
               f == "not" || f == "is_null") {
      e.type = kBool;
    } else {
      throw std::runtime_error("unknown function " + f);
    }
  }
};

// ---- evaluation (expression/Expr.cpp) --------------------------------------------------------
struct Batch {
  const vcpu_column* cols;
  int64_t offset; // first row of the batch inside the columns
  int32_t size;
};

struct Evaluator {
  const Batch& b;
  explicit Evaluator(const Batch& batch) : b(batch) {}

  // flat/flat and flat/constant get tight loops (Comparisons.cpp:105-158 dispatch analogue)
  template <typename T, typename R, typename F>
  static void binaryLoop(const Selectivity& rows, const Vec& x, const Vec& y, R* out, F f) {
    if (!x.constant && y.constant) {
      const T* a = x.as<T>(); const T c = y.at<T>(0);
      rows.forEach([&](int32_t i) { out[i] = f(a[i], c); });
    } else if (x.constant && !y.constant) {
      const T c = x.at<T>(0); const T* a = y.as<T>();
      rows.forEach([&](int32_t i) { out[i] = f(c, a[i]); });
    } else if (!x.constant) {
      const T* a = x.as<T>(); const T* c = y.as<T>();
      rows.forEach([&](int32_t i) { out[i] = f(a[i], c[i]); });
    } else {
      const R r = f(x.at<T>(0), y.at<T>(0));
      rows.forEach([&](int32_t i) { out[i] = r; });
    }
  }

  template <typename T>
  void arithmetic(const std::string& fn, const Selectivity& rows, const Vec& x, const Vec& y, Vec& out) {
    T* o = (T*)out.alloc(b.size, sizeof(T));
    if constexpr (std::is_same<T, double>::value) {
      if (fn == "plus") binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { return a + c; });
      else if (fn == "minus") binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { return a - c; });
      else if (fn == "multiply") binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { return a * c; });
      else binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { return a / c; });
    } else { // common/base/CheckedArithmetic.h:27-68
      if (fn == "plus") binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { T r; if (__builtin_add_overflow(a, c, &r)) throw UserError("integer overflow"); return r; });
      else if (fn == "minus") binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { T r; if (__builtin_sub_overflow(a, c, &r)) throw UserError("integer overflow"); return r; });
      else if (fn == "multiply") binaryLoop<T, T>(rows, x, y, o, [](T a, T c) { T r; if (__builtin_mul_overflow(a, c, &r)) throw UserError("integer overflow"); return r; });
      else binaryLoop<T, T>(rows, x, y, o, [](T a, T c) {
        if (c == 0) throw UserError("division by zero");
        if (a == std::numeric_limits<T>::min() && c == -1) throw UserError("integer overflow");
        return (T)(a / c); });
    }
  }

  template <typename T>
  void comparison(const std::string& fn, const Selectivity& rows, const Vec& x, const Vec& y, uint8_t* o) {
    if (fn == "eq") binaryLoop<T, uint8_t>(rows, x, y, o, [](T a, T c) -> uint8_t { return a == c; });
    else if (fn == "neq") binaryLoop<T, uint8_t>(rows, x, y, o, [](T a, T c) -> uint8_t { return a != c; });
    else if (fn == "lt") binaryLoop<T, uint8_t>(rows, x, y, o, [](T a, T c) -> uint8_t { return a < c; });
    else if (fn == "lte") binaryLoop<T, uint8_t>(rows, x, y, o, [](T a, T c) -> uint8_t { return a <= c; });
    else if (fn == "gt") binaryLoop<T, uint8_t>(rows, x, y, o, [](T a, T c) -> uint8_t { return a > c; });
    else binaryLoop<T, uint8_t>(rows, x, y, o, [](T a, T c) -> uint8_t { return a >= c; });
  }

  template <typename T>
  void between(const Selectivity& rows, const Vec& v, const Vec& lo, const Vec& hi, uint8_t* o) {
    // SimpleFunctionAdapter row loop over Comparisons.h:175-185
    rows.forEach([&](int32_t i) { const T x = v.at<T>(i); o[i] = x >= lo.at<T>(i) && x <= hi.at<T>(i); });
  }

  void eval(const Expr& e, const Selectivity& rows, Vec& out) {
    out.type = e.type;
    switch (e.op) {
      case Expr::kField: { // FieldReference: zero copy
        const vcpu_column& c = b.cols[e.column];
        out.data = (const char*)c.values + (size_t)b.offset * widthOf(c.type);
        out.nulls = c.nulls ? c.nulls + b.offset : nullptr;
        return;
      }
      case Expr::kConst:
        out.constant = true;
        out.const_null = e.is_null_const;
        out.cval.i = e.value.i;
        return;
      case Expr::kAnd:
      case Expr::kOr:
        conjunct(e, rows, out);
        return;
      default: break;
    }
    if (e.fn == "if" || e.fn == "switch") { switchExpr(e, rows, out); return; }
    if (e.fn == "coalesce") { coalesce(e, rows, out); return; }
    // default-null call: evaluate inputs, drop rows with a NULL input (Expr.cpp:371-441)
    std::vector<Vec> in(e.args.size());
    if (e.fn == "is_null") {
      eval(*e.args[0], rows, in[0]);
      uint8_t* o = (uint8_t*)out.alloc(b.size, 1);
      rows.forEach([&](int32_t i) { o[i] = in[0].isNullAt(i); });
      return;
    }
    // evalArgsDefaultNulls: argument k+1 is only evaluated on the rows where arguments 0..k are not NULL
    const Selectivity* active = &rows;
    Selectivity narrowed;
    uint8_t* on = nullptr;
    for (size_t k = 0; k < e.args.size(); ++k) {
      if (!active->hasSelections()) break;
      eval(*e.args[k], *active, in[k]);
      const Vec& v = in[k];
      if (!(v.constant ? v.const_null : v.nulls != nullptr)) continue;
      if (active == &rows) { narrowed = rows; active = &narrowed; }
      if (!on) on = out.allocNulls(b.size);
      bool changed = false;
      Selectivity pass = narrowed;
      pass.forEach([&](int32_t i) { if (v.isNullAt(i)) { narrowed.clear(i); on[i] = 1; changed = true; } });
      if (changed) narrowed.updateBounds();
    }
    if (!active->hasSelections()) { out.alloc(b.size, widthOf(e.type)); return; }
    const int at = e.args[0]->type;
    if (e.type != kBool) {
      if (e.fn == "negate") {
        if (at == kDouble) { double* o = (double*)out.alloc(b.size, 8); active->forEach([&](int32_t i) { o[i] = -in[0].at<double>(i); }); }
        else if (at == kBigint) { int64_t* o = (int64_t*)out.alloc(b.size, 8); active->forEach([&](int32_t i) { int64_t x = in[0].at<int64_t>(i); if (x == INT64_MIN) throw UserError("integer overflow"); o[i] = -x; }); }
        else { int32_t* o = (int32_t*)out.alloc(b.size, 4); active->forEach([&](int32_t i) { int32_t x = in[0].at<int32_t>(i); if (x == INT32_MIN) throw UserError("integer overflow"); o[i] = -x; }); }
        return;
      }
      if (at == kDouble) arithmetic<double>(e.fn, *active, in[0], in[1], out);
      else if (at == kBigint) arithmetic<int64_t>(e.fn, *active, in[0], in[1], out);
      else arithmetic<int32_t>(e.fn, *active, in[0], in[1], out);
      return;
    }
    uint8_t* o = (uint8_t*)out.alloc(b.size, 1);
    if (e.fn == "not") { active->forEach([&](int32_t i) { o[i] = !in[0].at<uint8_t>(i); }); return; }
    if (e.fn == "between") {
      if (at == kDouble) between<double>(*active, in[0], in[1], in[2], o);
      else if (at == kBigint) between<int64_t>(*active, in[0], in[1], in[2], o);
      else between<int32_t>(*active, in[0], in[1], in[2], o);
      return;
    }
    if (at == kDouble) comparison<double>(e.fn, *active, in[0], in[1], o);
    else if (at == kBigint) comparison<int64_t>(e.fn, *active, in[0], in[1], o);
    else if (at == kInt) comparison<int32_t>(e.fn, *active, in[0], in[1], o);
    else comparison<uint8_t>(e.fn, *active, in[0], in[1], o);
  }

  // Copies the non-null values of `v` on `rows` into out's buffer; NULL rows set out's null byte.
  void copyRows(const Selectivity& rows, const Vec& v, int type, char* o, uint8_t*& on, Vec& out) {
    rows.forEach([&](int32_t i) {
      if (v.isNullAt(i)) {
        if (!on) on = out.allocNulls(b.size);
        on[i] = 1;
        return;
      }
      if (type == kInt) ((int32_t*)o)[i] = v.at<int32_t>(i);
      else if (type == kBigint) ((int64_t*)o)[i] = v.at<int64_t>(i);
      else if (type == kDouble) ((double*)o)[i] = v.at<double>(i);
      else ((uint8_t*)o)[i] = v.at<uint8_t>(i);
    });
  }

  // SwitchExpr::evalSpecialForm (SwitchExpr.cpp:64-160): args = (c1, v1, c2, v2, ..., [else]).  Condition k is
  // evaluated on the rows no earlier condition took; its THEN clause only on the rows where it is TRUE; the ELSE
  // clause on what remains (FALSE or NULL everywhere); without ELSE those rows are NULL (:143-156).
  void switchExpr(const Expr& e, const Selectivity& rows, Vec& out) {
    char* o = (char*)out.alloc(b.size, widthOf(e.type));
    uint8_t* on = nullptr;
    Selectivity remaining = rows;
    const size_t cases = e.args.size() / 2;
    for (size_t k = 0; k < cases && remaining.hasSelections(); ++k) {
      Vec cond;
      eval(*e.args[2 * k], remaining, cond);
      Selectivity thenRows = remaining;
      bool any = false;
      remaining.forEach([&](int32_t i) {
        const bool t = !cond.isNullAt(i) && cond.at<uint8_t>(i);
        if (t) any = true; else thenRows.clear(i);
      });
      if (!any) continue;
      thenRows.updateBounds();
      Vec v;
      eval(*e.args[2 * k + 1], thenRows, v);
      copyRows(thenRows, v, e.type, o, on, out);
      thenRows.forEach([&](int32_t i) { remaining.clear(i); });
      remaining.updateBounds();
    }
    if (!remaining.hasSelections()) return;
    if (e.args.size() % 2 == 1) {
      Vec v;
      eval(*e.args.back(), remaining, v);
      copyRows(remaining, v, e.type, o, on, out);
    } else {
      if (!on) on = out.allocNulls(b.size);
      remaining.forEach([&](int32_t i) { on[i] = 1; });
    }
  }

  // CoalesceExpr::evalSpecialForm (CoalesceExpr.cpp:39-81): argument k is evaluated on the rows that
  // are still NULL after arguments 0..k-1.
  void coalesce(const Expr& e, const Selectivity& rows, Vec& out) {
    char* o = (char*)out.alloc(b.size, widthOf(e.type));
    uint8_t* on = nullptr;
    Selectivity active = rows;
    for (auto& arg : e.args) {
      if (!active.hasSelections()) break;
      Vec v;
      eval(*arg, active, v);
      uint8_t* ignored = nullptr;
      bool any = false;
      Selectivity found = active;
      active.forEach([&](int32_t i) { if (v.isNullAt(i)) found.clear(i); else any = true; });
      if (!any) continue;
      found.updateBounds();
      Vec sink; // null marks are not taken from this pass
      copyRows(found, v, e.type, o, ignored, sink);
      found.forEach([&](int32_t i) { active.clear(i); });
      active.updateBounds();
    }
    if (active.hasSelections()) {
      on = out.allocNulls(b.size);
      active.forEach([&](int32_t i) { on[i] = 1; });
    }
  }

  // ConjunctExpr::evalSpecialForm (ConjunctExpr.cpp:96-174): each conjunct is evaluated only on the
  // rows no earlier conjunct decided; FALSE (AND) / TRUE (OR) decides, NULL only if undecided.
  void conjunct(const Expr& e, const Selectivity& rows, Vec& out) {
    const bool is_and = e.op == Expr::kAnd;
    uint8_t* o = (uint8_t*)out.alloc(b.size, 1);
    uint8_t* on = nullptr;
    rows.forEach([&](int32_t i) { o[i] = is_and ? 1 : 0; });
    Selectivity active = rows;
    for (auto& arg : e.args) {
      if (!active.hasSelections()) break;
      Vec v;
      eval(*arg, active, v);
      bool changed = false;
      active.forEach([&](int32_t i) {
        if (v.isNullAt(i)) {
          if (!on) on = out.allocNulls(b.size);
          on[i] = 1;
          return;
        }
        const bool val = v.at<uint8_t>(i);
        if (val != is_and) { // decided
          o[i] = val;
          if (on) on[i] = 0;
          active.clear(i);
          changed = true;
        }
      });
      if (changed) active.updateBounds();
    }
  }
};

// ---- HashAggregation -------------------------------------------------------------------------
// Group row layout (RowContainer.cpp:115-252): [keys 8 B each][key null bytes][acc: value 8 B,
// non-null count 8 B]...
struct AggSpec {
  std::vector<int> key_outputs;
  std::vector<vcpu_acc> accs;
  std::vector<int> out_types;
  bool ignore_null_keys = false;
  int rowBytes() const { return (int)(key_outputs.size() * 8 + ((key_outputs.size() + 7) & ~7u) + accs.size() * 16); }
  int nullOffset() const { return (int)key_outputs.size() * 8; }
  int accOffset(int a) const { return (int)(key_outputs.size() * 8 + ((key_outputs.size() + 7) & ~7u) + a * 16); }
};

struct Arena {
  std::vector<std::unique_ptr<char[]>> chunks;
  size_t used = 0, cap = 0;
  char* alloc(size_t n) {
    if (used + n > cap) {
      cap = std::max<size_t>(1 << 20, n);
      chunks.emplace_back(new char[cap]);
      used = 0;
    }
    char* p = chunks.back().get() + used;
    used += n;
    return p;
  }
};

inline uint64_t mixHash(uint64_t h, uint64_t v) { // bits::hashMix (common/base/BitUtil.h:735)
  const uint64_t kMul = 0x9ddfea08eb382d69ULL;
  uint64_t a = (v ^ h) * kMul;
  a ^= (a >> 47);
  uint64_t b2 = (h ^ a) * kMul;
  b2 ^= (b2 >> 47);
  return b2 * kMul;
}
inline uint64_t hashInt(uint64_t key) { // folly::hasher<int64_t> = twang_mix64
  key = (~key) + (key << 21);
  key = key ^ (key >> 24);
  key = key + (key << 3) + (key << 8);
  key = key ^ (key >> 14);
  key = key + (key << 2) + (key << 4);
  key = key ^ (key >> 28);
  key = key + (key << 31);
  return key;
}

struct GroupingSet {
  const AggSpec& spec;
  Arena arena;
  std::vector<char*> rows;          // every group row, in creation order (RowContainer::listRows)
  // kArray mode: direct index by value ids (HashTable.cpp:514-562)
  bool array_mode = true;
  std::vector<int64_t> mins, maxs, mults;
  std::vector<char*> array_table;
  // kHash mode: open addressing over row pointers, tag = upper hash bits (HashTable.cpp:426-474)
  std::vector<char*> table;
  std::vector<uint64_t> hashes_of_slot;
  uint64_t mask = 0;

  explicit GroupingSet(const AggSpec& s) : spec(s) {}

  char* newRow(const int64_t* keys, const uint8_t* key_nulls) {
    char* r = arena.alloc(spec.rowBytes());
    const int nk = (int)spec.key_outputs.size();
    memcpy(r, keys, nk * 8);
    memset(r + spec.nullOffset(), 0, (nk + 7) & ~7);
    for (int k = 0; k < nk; ++k) r[spec.nullOffset() + k] = key_nulls[k];
    for (size_t a = 0; a < spec.accs.size(); ++a) { // Aggregate::initializeNewGroups
      int64_t init = 0;
      const int t = spec.accs[a].input >= 0 ? spec.out_types[spec.accs[a].input] : kBigint;
      if (spec.accs[a].kind == 3) { if (t == kDouble) { double d = std::numeric_limits<double>::max(); memcpy(&init, &d, 8); } else init = INT64_MAX; }
      if (spec.accs[a].kind == 4) { if (t == kDouble) { double d = std::numeric_limits<double>::lowest(); memcpy(&init, &d, 8); } else init = INT64_MIN; }
      memcpy(r + spec.accOffset((int)a), &init, 8);
      memset(r + spec.accOffset((int)a) + 8, 0, 8);
    }
    rows.push_back(r);
    return r;
  }

  bool sameKeys(const char* r, const int64_t* keys, const uint8_t* nulls) const {
    const int nk = (int)spec.key_outputs.size();
    for (int k = 0; k < nk; ++k) {
      const bool rn = r[spec.nullOffset() + k];
      if (rn != (bool)nulls[k]) return false;
      if (!rn && ((const int64_t*)r)[k] != keys[k]) return false;
    }
    return true;
  }

  void growTable() {
    const uint64_t cap = table.empty() ? 2048 : table.size() * 2; // HashTable.h:569-579
    table.assign(cap, nullptr);
    hashes_of_slot.assign(cap, 0);
    mask = cap - 1;
    const int nk = (int)spec.key_outputs.size();
    for (char* r : rows) {
      uint64_t h = 0;
      for (int k = 0; k < nk; ++k) {
        const uint64_t hk = r[spec.nullOffset() + k] ? 1 : hashInt(((const uint64_t*)r)[k]); // kNullHash = 1
        h = k == 0 ? hk : mixHash(h, hk);
      }
      uint64_t s = h & mask;
      while (table[s]) s = (s + 1) & mask;
      table[s] = r;
      hashes_of_slot[s] = h;
    }
  }

  char* probeHash(const int64_t* keys, const uint8_t* nulls) {
    const int nk = (int)spec.key_outputs.size();
    if (table.empty() || rows.size() * 8 >= table.size() * 7) growTable(); // rehash at 7/8 full
    uint64_t h = 0;
    for (int k = 0; k < nk; ++k) {
      const uint64_t hk = nulls[k] ? 1 : hashInt((uint64_t)keys[k]);
      h = k == 0 ? hk : mixHash(h, hk);
    }
    uint64_t s = h & mask;
    while (true) {
      char* r = table[s];
      if (!r) {
        r = newRow(keys, nulls);
        table[s] = r;
        hashes_of_slot[s] = h;
        return r;
      }
      if (hashes_of_slot[s] == h && sameKeys(r, keys, nulls)) return r;
      s = (s + 1) & mask;
    }
  }

  // Decides / keeps array mode for this batch's key ranges (decideHashMode, HashTable.cpp:1376-1461).
  void analyze(const std::vector<std::vector<int64_t>>& kv, const std::vector<std::vector<uint8_t>>& kn, int n) {
    if (!array_mode) return;
    const int nk = (int)kv.size();
    bool changed = mins.empty();
    if (mins.empty()) { mins.assign(nk, INT64_MAX); maxs.assign(nk, INT64_MIN); }
    for (int k = 0; k < nk; ++k) {
      for (int i = 0; i < n; ++i) {
        if (kn[k][i]) continue;
        if (kv[k][i] < mins[k]) { mins[k] = kv[k][i]; changed = true; }
        if (kv[k][i] > maxs[k]) { maxs[k] = kv[k][i]; changed = true; }
      }
    }
    if (!changed) return;
    __int128 total = 1;
    mults.assign(nk, 0);
    for (int k = 0; k < nk; ++k) {
      mults[k] = (int64_t)total;
      const __int128 range = maxs[k] >= mins[k] ? (__int128)maxs[k] - mins[k] + 2 : 1;
      total *= range;
      if (total > (2 << 20)) { array_mode = false; break; } // kArrayHashMaxSize, HashTable.h:71
    }
    if (!array_mode) { array_table.clear(); growTable(); return; }
    array_table.assign((size_t)total, nullptr);
    for (char* r : rows) {
      int64_t idx = 0;
      for (int k = 0; k < nk; ++k) idx += (r[spec.nullOffset() + k] ? 0 : ((const int64_t*)r)[k] - mins[k] + 1) * mults[k];
      array_table[idx] = r;
    }
  }
};

// Decoded aggregation input: values gathered through selectedIndices (DecodedVector::valueAt).
struct AggInput {
  std::vector<std::vector<int64_t>> key_vals;
  std::vector<std::vector<uint8_t>> key_nulls;
};

void addInput(GroupingSet& gs, const AggSpec& spec, const std::vector<Vec>& outs, const int32_t* sel, int32_t n_sel,
              std::vector<char*>& groups) {
  const int nk = (int)spec.key_outputs.size();
  groups.resize(n_sel);
  if (nk == 0) { // global aggregation: one row (GroupingSet.cpp:513-559)
    if (gs.rows.empty()) gs.newRow(nullptr, nullptr);
    for (int i = 0; i < n_sel; ++i) groups[i] = gs.rows[0];
  } else {
    std::vector<std::vector<int64_t>> kv(nk, std::vector<int64_t>(n_sel));
    std::vector<std::vector<uint8_t>> kn(nk, std::vector<uint8_t>(n_sel));
    for (int k = 0; k < nk; ++k) { // VectorHasher::decode
      const Vec& v = outs[spec.key_outputs[k]];
      for (int i = 0; i < n_sel; ++i) {
        const int32_t r = sel[i];
        kn[k][i] = v.isNullAt(r);
        kv[k][i] = kn[k][i] ? 0 : (v.type == kInt ? (int64_t)v.at<int32_t>(r) : v.type == kBool ? (int64_t)v.at<uint8_t>(r) : v.at<int64_t>(r));
      }
    }
    gs.analyze(kv, kn, n_sel);
    int64_t keys[16]; uint8_t nulls[16];
    for (int i = 0; i < n_sel; ++i) {
      bool any_null = false;
      for (int k = 0; k < nk; ++k) { keys[k] = kv[k][i]; nulls[k] = kn[k][i]; any_null |= nulls[k]; }
      if (any_null && spec.ignore_null_keys) { groups[i] = nullptr; continue; }
      if (gs.array_mode) {
        int64_t idx = 0;
        for (int k = 0; k < nk; ++k) idx += (nulls[k] ? 0 : keys[k] - gs.mins[k] + 1) * gs.mults[k];
        char*& slot = gs.array_table[idx];
        if (!slot) slot = gs.newRow(keys, nulls);
        groups[i] = slot;
      } else {
        groups[i] = gs.probeHash(keys, nulls);
      }
    }
  }
  // one pass per aggregate function over the batch, input order (SimpleNumericAggregate.h:90-142)
  for (size_t a = 0; a < spec.accs.size(); ++a) {
    const vcpu_acc& acc = spec.accs[a];
    const int off = spec.accOffset((int)a);
    if (acc.input < 0) {
      for (int i = 0; i < n_sel; ++i) if (groups[i]) { ++*(int64_t*)(groups[i] + off); ++*(int64_t*)(groups[i] + off + 8); }
      continue;
    }
    const Vec& v = outs[acc.input];
    for (int i = 0; i < n_sel; ++i) {
      char* g = groups[i];
      const int32_t r = sel[i];
      if (!g || v.isNullAt(r)) continue;
      ++*(int64_t*)(g + off + 8);
      switch (acc.kind) {
        case 0: *(double*)(g + off) += v.type == kDouble ? v.at<double>(r) : v.type == kInt ? (double)v.at<int32_t>(r) : (double)v.at<int64_t>(r); break;
        case 1: {
          const int64_t x = v.type == kInt ? (int64_t)v.at<int32_t>(r) : v.at<int64_t>(r);
          if (__builtin_add_overflow(*(int64_t*)(g + off), x, (int64_t*)(g + off))) throw UserError("integer overflow");
          break;
        }
        case 2: ++*(int64_t*)(g + off); break;
        case 3: case 4:
          if (v.type == kDouble) {
            const double x = v.at<double>(r); double& cur = *(double*)(g + off);
            if (acc.kind == 3 ? x < cur : cur < x) cur = x;
          } else {
            const int64_t x = v.type == kInt ? (int64_t)v.at<int32_t>(r) : v.at<int64_t>(r); int64_t& cur = *(int64_t*)(g + off);
            if (acc.kind == 3 ? x < cur : cur < x) cur = x;
          }
          break;
      }
    }
  }
}

// Final aggregation over the partial rows of one driver (addIntermediateResults).
void mergePartial(GroupingSet& final_gs, const AggSpec& spec, const GroupingSet& part) {
  const int nk = (int)spec.key_outputs.size();
  const int n = (int)part.rows.size();
  if (nk > 0) {
    std::vector<std::vector<int64_t>> kv(nk, std::vector<int64_t>(n));
    std::vector<std::vector<uint8_t>> kn(nk, std::vector<uint8_t>(n));
    for (int i = 0; i < n; ++i) for (int k = 0; k < nk; ++k) { kv[k][i] = ((const int64_t*)part.rows[i])[k]; kn[k][i] = part.rows[i][spec.nullOffset() + k]; }
    final_gs.analyze(kv, kn, n);
  }
  for (char* r : part.rows) {
    char* g;
    if (nk == 0) { if (final_gs.rows.empty()) final_gs.newRow(nullptr, nullptr); g = final_gs.rows[0]; }
    else {
      const int64_t* keys = (const int64_t*)r; const uint8_t* nulls = (const uint8_t*)r + spec.nullOffset();
      if (final_gs.array_mode) {
        int64_t idx = 0;
        for (int k = 0; k < nk; ++k) idx += (nulls[k] ? 0 : keys[k] - final_gs.mins[k] + 1) * final_gs.mults[k];
        char*& slot = final_gs.array_table[idx];
        if (!slot) slot = final_gs.newRow(keys, nulls);
        g = slot;
      } else g = final_gs.probeHash(keys, nulls);
    }
    for (size_t a = 0; a < spec.accs.size(); ++a) {
      const int off = spec.accOffset((int)a);
      const int64_t nn = *(const int64_t*)(r + off + 8);
      const int t = spec.accs[a].input >= 0 ? spec.out_types[spec.accs[a].input] : kBigint;
      *(int64_t*)(g + off + 8) += nn;
      if (nn == 0 && spec.accs[a].kind != 2) continue;
      switch (spec.accs[a].kind) {
        case 0: *(double*)(g + off) += *(const double*)(r + off); break;
        case 1: if (__builtin_add_overflow(*(int64_t*)(g + off), *(const int64_t*)(r + off), (int64_t*)(g + off))) throw UserError("integer overflow"); break;
        case 2: *(int64_t*)(g + off) += *(const int64_t*)(r + off); break;
        default:
          if (t == kDouble) { const double x = *(const double*)(r + off); double& c = *(double*)(g + off); if (spec.accs[a].kind == 3 ? x < c : c < x) c = x; }
          else { const int64_t x = *(const int64_t*)(r + off); int64_t& c = *(int64_t*)(g + off); if (spec.accs[a].kind == 3 ? x < c : c < x) c = x; }
      }
    }
  }
}

} // namespace

extern "C" const char* vcpu_error(void) { return g_last_error.c_str(); }

extern "C" int64_t vcpu_run(const char* filter, const char* const* outputs, int32_t n_outputs, const vcpu_column* cols,
                            int32_t n_cols, int64_t n_rows, const int32_t* key_outputs, int32_t n_keys,
                            const vcpu_acc* accs, int32_t n_accs, int32_t ignore_null_keys, int32_t threads,
                            int32_t batch_rows, int64_t* out_rows, int64_t out_cap_rows, double* seconds) {
  try {
    std::unique_ptr<Expr> f;
    if (filter && *filter) f = Parser(filter).parse();
    std::vector<std::unique_ptr<Expr>> outs;
    AggSpec spec;
    for (int i = 0; i < n_outputs; ++i) { outs.push_back(Parser(outputs[i]).parse()); spec.out_types.push_back(outs.back()->type); }
    spec.key_outputs.assign(key_outputs, key_outputs + n_keys);
    spec.accs.assign(accs, accs + n_accs);
    spec.ignore_null_keys = ignore_null_keys != 0;
    if (n_keys > 16) throw std::runtime_error("at most 16 keys");
    threads = std::max(1, threads);
    batch_rows = std::max(1, batch_rows);
    std::vector<std::unique_ptr<GroupingSet>> partials;
    for (int t = 0; t < threads; ++t) partials.emplace_back(new GroupingSet(spec));
    std::vector<std::string> errors(threads);
    const auto t0 = std::chrono::steady_clock::now();
    auto driver = [&](int t) {
      try {
        const int64_t lo = n_rows * t / threads, hi = n_rows * (t + 1) / threads;
        Selectivity all, passing;
        std::vector<int32_t> sel;
        std::vector<char*> groups;
        for (int64_t off = lo; off < hi; off += batch_rows) {
          Batch b{cols, off, (int32_t)std::min<int64_t>(batch_rows, hi - off)};
          Evaluator ev(b);
          all.resizeAll(b.size);
          sel.clear();
          const Selectivity* rows = &all;
          if (f) { // FilterProject::filter + processFilterResults
            Vec fv;
            ev.eval(*f, all, fv);
            passing.resizeAll(b.size);
            for (int32_t i = 0; i < b.size; ++i) {
              if (!fv.isNullAt(i) && fv.at<uint8_t>(i)) sel.push_back(i); else passing.clear(i);
            }
            passing.updateBounds();
            rows = &passing;
            if (sel.empty()) continue;
          } else {
            sel.resize(b.size);
            for (int32_t i = 0; i < b.size; ++i) sel[i] = i;
          }
          std::vector<Vec> ov(outs.size());
          for (size_t k = 0; k < outs.size(); ++k) ev.eval(*outs[k], *rows, ov[k]); // FilterProject::project
          addInput(*partials[t], spec, ov, sel.data(), (int32_t)sel.size(), groups);
        }
      } catch (const std::exception& e) { errors[t] = e.what(); }
    };
    if (threads == 1) driver(0);
    else {
      std::vector<std::thread> pool;
      for (int t = 0; t < threads; ++t) pool.emplace_back(driver, t);
      for (auto& th : pool) th.join();
    }
    for (auto& e : errors) if (!e.empty()) throw UserError(e);
    GroupingSet* result = partials[0].get();
    GroupingSet final_gs(spec);
    if (threads > 1) {
      for (auto& p : partials) mergePartial(final_gs, spec, *p);
      result = &final_gs;
    }
    if (n_keys == 0 && result->rows.empty()) result->newRow(nullptr, nullptr); // global agg: always one row
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const int64_t n = (int64_t)result->rows.size();
    const int words = 2 * n_keys + 2 * n_accs;
    for (int64_t i = 0; i < std::min(n, out_cap_rows); ++i) {
      const char* r = result->rows[i];
      int64_t* o = out_rows + i * words;
      for (int k = 0; k < n_keys; ++k) { o[k] = ((const int64_t*)r)[k]; o[n_keys + k] = r[spec.nullOffset() + k]; }
      for (int a = 0; a < n_accs; ++a) { memcpy(&o[2 * n_keys + 2 * a], r + spec.accOffset(a), 16); }
    }
    return n;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return -1;
  }
}
