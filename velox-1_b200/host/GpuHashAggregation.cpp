// GpuHashAggregation: exec::HashAggregation + GroupingSet on the GPU (velox/exec/HashAggregation.cpp:97-472,
// velox/exec/GroupingSet.cpp:177-307,513-712).
// addInput: the batch's columns are copied into the current slot of the staging ring (GpuDevice.h) on a copy stream —
// VARCHAR keys as int32 value ids (VectorHasher::makeValueIds analogue, velox/exec/VectorHasher.cpp:110-143: one
// dictionary lookup per BASE string on the host, the per-row gather on the device) — and small batches (Velox's
// default scan batch is ~4 k rows) are gathered there until QueryConfig::coalesceRows rows are resident; then ONE
// vgpu_agg_add_input call runs filter + projections + grouping + accumulation over all of them.  No device allocation,
// A VARCHAR key that arrives as StringViews (StringViewVector) is not looked up on the host at all: its 16-byte views
// are copied as they are and the kernels read each as the integer of VectorHasher::stringAsNumber
// (velox/exec/VectorHasher.h:324-329; strings over 7 bytes get host-assigned numbers instead).
// no pageable copy and no stream synchronisation per batch; data errors (sum overflow, checked arithmetic) are
// collected every QueryConfig::errorCheckEvery launches and at noMoreInput.
// After noMoreInput: vgpu_agg_extract, then output batches of preferredOutputBatchRows groups.
// A partial aggregation flushes (HashAggregation.cpp:326-348): when the device state passes
// maxPartialAggregationMemory its groups are emitted downstream and the state is reset.
// A partial aggregation that does not reduce its input is abandoned (:229-233,355-364): its groups leave after every launch.
// Not restated: spilling, sorted / distinct aggregates.
#include <chrono>
#include <cmath>
#include <limits>
#include <unordered_map>

#include "GpuDevice.h"

namespace vb200 {
namespace gpu {
using namespace core;
using Step = AggregationNode::Step;

namespace {
struct KeyInfo {
  std::string name;
  TypePtr type;
  bool varchar = false;
  int64_t lo = 0, hi = 0;
  int bits = 0;
  std::vector<std::string> strings;                 // id -> string
  std::unordered_map<std::string, int32_t> idOf;
  // VARCHAR key carried as numbers instead of ids: stringAsNumber for strings of at most 7 bytes (computed on the
  // device from the StringViews), kLongBase + index into longStrings for longer ones (computed here)
  bool number = false;
  int64_t seenLo = std::numeric_limits<int64_t>::max(), seenHi = std::numeric_limits<int64_t>::min();
  std::vector<std::string> longStrings;
  std::unordered_map<std::string, int64_t> longId;
  // INTEGER / BIGINT key next to other keys: packed as `value - lo + 1` over the range seen so far (VectorHasher's range
  // mode feeding HashTable's kNormalizedKey, velox/exec/VectorHasher.cpp:526-577, HashTable.cpp:1376-1461): the host
  // keeps min / max of the key column per batch (seenLo / seenHi) and the table is rebuilt when a value leaves [lo, hi]
  bool ranged = false;
  std::string rangedColumn;
};
// [lo, hi] widened by its own width on both sides, saturating (so that a slowly growing range does not rebuild every batch)
void padRange(int64_t lo, int64_t hi, int64_t& outLo, int64_t& outHi) {
  const __int128 span = (__int128)hi - lo + 1;
  const __int128 l = (__int128)lo - span, h = (__int128)hi + span;
  outLo = l < (__int128)std::numeric_limits<int64_t>::min() ? std::numeric_limits<int64_t>::min() : (int64_t)l;
  outHi = h > (__int128)std::numeric_limits<int64_t>::max() ? std::numeric_limits<int64_t>::max() : (int64_t)h;
}
constexpr int64_t kLongBase = 1ll << 60;
// What one batch's VARCHAR number key looks like: its views can travel as they are, or numbers per BASE value.
struct NumberBatch {
  bool views = false;
  std::shared_ptr<FlatVector<int64_t>> numbers;
};
const BaseVector* flatBaseOf(const BaseVector* v) {
  while (v->encoding() == VectorEncoding::DICTIONARY) v = static_cast<const DictionaryVector*>(v)->valueVector().get();
  if (v->encoding() == VectorEncoding::CONSTANT) v = static_cast<const ConstantVector*>(v)->valueVector().get();
  VB_CHECK(v->encoding() == VectorEncoding::FLAT, "VARCHAR key over an unsupported encoding");
  return v;
}
int bitsFor(int64_t card) { int b = 1; while (b < 63 && (1ll << b) < card) ++b; return b; }
struct AggInfo {
  std::string fn;
  TypePtr rawType, outType;
  int acc0 = -1, acc1 = -1;
  std::string ordColumn; // min / max over VARCHAR: the accumulator holds order-preserving ids of this input column's strings
};
// min / max over VARCHAR (MinMaxAggregates.cpp:282-534 compare StringViews): every distinct string of the column gets a
// 64-bit id whose order is the strings' byte order, the device runs its integer MIN / MAX over the ids, the output maps
// ids back.  A new string takes the midpoint between its neighbours' ids (2^32 past the ends); when a gap is used up all
// ids are spread out again and the accumulators already on the device are rewritten (remapOrd).
struct OrdDict {
  std::map<std::string, int64_t> idOf;
  std::unordered_map<int64_t, std::string> strOf;
  static constexpr int64_t kStep = 1ll << 32;
};
TypedExprPtr castTo(const TypedExprPtr& e, const TypePtr& t) { return e->type()->equals(*t) ? e : std::make_shared<CastTypedExpr>(t, e); }
int64_t typeLo(TypeKind k) { return k == TypeKind::TINYINT ? -128 : k == TypeKind::SMALLINT ? -32768 : -2147483648ll; }
int64_t typeHi(TypeKind k) { return k == TypeKind::TINYINT ? 127 : k == TypeKind::SMALLINT ? 32767 : 2147483647ll; }
} // namespace

struct GpuHashAggregation::Impl {
  std::shared_ptr<const AggregationNode> node;
  RowTypePtr inputType;       // what addInput receives
  FilterProjectExprs fused;
  bool hasFused = false;
  exec::DriverCtx* ctx = nullptr;

  std::vector<KeyInfo> keys;
  std::vector<AggInfo> aggs;
  std::vector<vgpu_acc_spec> accs;
  std::vector<TypedExprPtr> outputs; // program outputs: keys, then accumulator inputs
  LoweredProgram lowered;
  vgpu_program_t program = nullptr;
  vgpu_agg_t agg = nullptr;
  void* stream = nullptr;
  int mode = VGPU_AGG_GLOBAL;
  bool sawInput = false;
  bool forceHash = false;     // the small-domain table was declined for this shape: group through the HBM table

  std::vector<int64_t> rows; // extracted packed groups
  int64_t nGroups = -1;
  int64_t cursor = 0;
  // ---- staging / coalescing (GpuDevice.h) ----
  std::unique_ptr<Staging> staging;
  struct PendingColumn {
    int type = 0;            // vgpu_type on the device
    size_t width = 0;        // bytes per value
    char* dValues = nullptr; // device array of capacity pendingCap values
    char* hValues = nullptr; // pinned mirror of dValues: small batches are gathered here and copied once per flush
    int64_t hostFrom = -1;   // first gathered row that still lives only in hValues (-1: none)
    bool anyNulls = false;
    std::vector<uint64_t> notNull; // host bitmap of the gathered rows (uploaded at launch when anyNulls)
  };
  std::vector<PendingColumn> pending;
  int64_t pendingRows = 0, pendingCap = 0;
  int64_t launchesSinceCheck = 0;
  int64_t rowsSinceFlush = 0;
  int64_t nextAbandonCheck = 0, launchesAtLastFlush = 0;
  std::map<std::string, OrdDict> ordDicts; // VARCHAR inputs of min / max, by input column name
  GpuHashAggregation::Stats stats;
  int rowWords() const { return 3 + 2 * (int)accs.size(); }
  // ids per key in the device's numbering: [lo, hi] plus one id for NULL unless NULL keys are dropped
  int64_t nullIds() const { return node->ignoreNullKeys() ? 0 : 1; }
  int64_t cardOf(const KeyInfo& k) const { // saturating: a ranged BIGINT key may span most of int64
    const __int128 c = (__int128)k.hi - k.lo + 1 + nullIds();
    return c > (__int128)std::numeric_limits<int64_t>::max() ? std::numeric_limits<int64_t>::max() : (int64_t)c;
  }

  // expression a name refers to on the operator's input (through the fused projections, if any)
  TypedExprPtr exprOf(const std::string& name) const {
    if (hasFused && !fused.names.empty()) {
      for (size_t i = 0; i < fused.names.size(); ++i) if (fused.names[i] == name) return fused.projections[i];
      throw VeloxUserError("Field not found: " + name);
    }
    const size_t dot = name.find('.');
    if (dot != std::string::npos) { // child of a ROW-typed intermediate column (avg)
      const int idx = inputType->getChildIdx(name.substr(0, dot));
      auto row = std::dynamic_pointer_cast<const RowType>(inputType->childAt(idx));
      return std::make_shared<FieldAccessTypedExpr>(row->childAt(row->getChildIdx(name.substr(dot + 1))), name);
    }
    return std::make_shared<FieldAccessTypedExpr>(inputType->childAt(inputType->getChildIdx(name)), name);
  }
  // type of an INPUT column of the operator (what lowered.columns names; "col.field" = a child of a ROW-typed column)
  TypePtr inputColumnType(const std::string& name) const {
    const size_t dot = name.find('.');
    if (dot == std::string::npos) return inputType->childAt(inputType->getChildIdx(name));
    auto row = std::dynamic_pointer_cast<const RowType>(inputType->childAt(inputType->getChildIdx(name.substr(0, dot))));
    VB_CHECK(row != nullptr, "not a ROW column: " + name);
    return row->childAt(row->getChildIdx(name.substr(dot + 1)));
  }
  int addOutput(const TypedExprPtr& e) {
    for (size_t i = keys.size(); i < outputs.size(); ++i) if (outputs[i]->toString() == e->toString() && outputs[i]->type()->equals(*e->type())) return (int)i;
    outputs.push_back(e);
    return (int)outputs.size() - 1;
  }
  int addAcc(int kind, int input) { accs.push_back(vgpu_acc_spec{kind, input}); return (int)accs.size() - 1; }

  void planAccumulators() {
    const bool raw = AggregationNode::isRawInput(node->step());
    for (size_t i = 0; i < node->aggregates().size(); ++i) {
      const auto& a = node->aggregates()[i];
      AggInfo info;
      info.fn = a.function;
      info.rawType = a.rawInputType ? a.rawInputType : BIGINT();
      info.outType = node->outputType()->childAt(keys.size() + i);
      const bool hasColumn = !a.args.empty() && a.args[0]->kind() == ITypedExpr::Kind::kFieldAccess;
      TypedExprPtr x = hasColumn ? exprOf(static_cast<const FieldAccessTypedExpr*>(a.args[0].get())->name()) : nullptr;
      const std::string col = hasColumn ? static_cast<const FieldAccessTypedExpr*>(a.args[0].get())->name() : "";
      bool isCountIf = false;
      if (a.function == "count_if") { // count_if(b) = count(if(b, 1)): TRUE rows only (CountIfAggregate.cpp); final = sum of partials
        isCountIf = true;
        info.fn = "count";
        if (raw) {
          VB_USER_CHECK(x && x->type()->kind() == TypeKind::BOOLEAN, "count_if takes one BOOLEAN argument");
          x = std::make_shared<CallTypedExpr>(BIGINT(), std::vector<TypedExprPtr>{x, std::make_shared<ConstantTypedExpr>(BIGINT(), false, 1, 0.0)}, "if");
        }
      }
      // Aggregation mask: only rows whose mask is TRUE feed the aggregate (AggregationMasks.cpp) == the aggregate over
      // if(mask, x) with no ELSE, since every aggregate here skips NULL inputs; count(*) counts if(mask, 1).
      if (raw && a.mask) {
        TypedExprPtr m = exprOf(static_cast<const FieldAccessTypedExpr*>(a.mask.get())->name());
        TypedExprPtr v = x ? x : TypedExprPtr(std::make_shared<ConstantTypedExpr>(BIGINT(), false, 1, 0.0));
        VB_USER_CHECK(v->type()->kind() != TypeKind::VARCHAR && v->type()->kind() != TypeKind::BOOLEAN, "masked " + a.function + " over " + v->type()->toString() + " is not supported on the GPU");
        x = std::make_shared<CallTypedExpr>(v->type(), std::vector<TypedExprPtr>{m, v}, "if");
      }
      const bool countsColumn = hasColumn || (raw && a.mask);
      if (a.function == "count" || isCountIf) {
        if (raw) info.acc0 = addAcc(VGPU_ACC_COUNT, countsColumn ? addOutput(x) : -1);
        else info.acc0 = addAcc(VGPU_ACC_SUM_I64, addOutput(x)); // final count = sum of partial counts (CountAggregate.cpp:81-90)
      } else if (a.function == "sum") {
        const bool fp = raw ? x->type()->isFloating() : info.rawType->isFloating();
        info.acc0 = addAcc(fp ? VGPU_ACC_SUM_F64 : VGPU_ACC_SUM_I64, addOutput(castTo(x, fp ? DOUBLE() : BIGINT())));
      } else if (a.function == "avg") {
        if (raw) {
          const int in = addOutput(castTo(x, DOUBLE()));
          info.acc0 = addAcc(VGPU_ACC_SUM_F64, in);
          info.acc1 = addAcc(VGPU_ACC_COUNT, in);
        } else { // intermediate ROW(sum DOUBLE, count BIGINT) (AverageAggregateBase.h:36-41)
          info.acc0 = addAcc(VGPU_ACC_SUM_F64, addOutput(exprOf(col + ".sum")));
          info.acc1 = addAcc(VGPU_ACC_SUM_I64, addOutput(exprOf(col + ".count")));
        }
      } else if (a.function == "bool_and" || a.function == "bool_or") {
        // bool_and = min, bool_or = max over FALSE < TRUE; NULL inputs ignored, NULL without input
        // (velox/functions/prestosql/aggregates/BoolAggregates.cpp); the intermediate is the same BOOLEAN
        VB_USER_CHECK(x->type()->kind() == TypeKind::BOOLEAN, a.function + " takes one BOOLEAN argument");
        info.acc0 = addAcc(a.function == "bool_and" ? VGPU_ACC_MIN : VGPU_ACC_MAX, addOutput(x));
      } else if (a.function == "min" || a.function == "max") {
        VB_USER_CHECK(x->type()->kind() != TypeKind::BOOLEAN, a.function + " over " + x->type()->toString() + " is not supported on the GPU");
        if (x->type()->kind() == TypeKind::VARCHAR) {
          VB_USER_CHECK(x->kind() == ITypedExpr::Kind::kFieldAccess && !a.mask, a.function + " over VARCHAR takes a plain, unmasked column on the GPU");
          info.ordColumn = static_cast<const FieldAccessTypedExpr*>(x.get())->name();
          for (size_t k = 0; k < keys.size(); ++k)
            VB_USER_CHECK(!keys[k].varchar || keyColumn(k) != info.ordColumn, a.function + " over a VARCHAR grouping key is not supported on the GPU");
          ordDicts[info.ordColumn];
          info.acc0 = addAcc(a.function == "min" ? VGPU_ACC_MIN : VGPU_ACC_MAX, addOutput(std::make_shared<FieldAccessTypedExpr>(BIGINT(), info.ordColumn)));
        } else
        info.acc0 = addAcc(a.function == "min" ? VGPU_ACC_MIN : VGPU_ACC_MAX, addOutput(x->type()->kind() == TypeKind::REAL ? castTo(x, DOUBLE()) : x));
      } else {
        throw VeloxUserError("Aggregate function not registered: " + a.function);
      }
      aggs.push_back(info);
    }
  }

  // Decides the table mode and key packing once the first batch's VARCHAR dictionaries are known.
  void create() {
    mode = keys.empty() ? VGPU_AGG_GLOBAL : VGPU_AGG_HASH;
    bool small = !keys.empty();
    int64_t product = 1;
    for (auto& k : keys) {
      if (k.number) { // the range seen so far, padded by its own width on both sides (numbers are >= 0)
        const int64_t lo = k.seenLo <= k.seenHi ? k.seenLo : 0, hi = k.seenLo <= k.seenHi ? k.seenHi : 0;
        if (hi >= kLongBase) { k.lo = 0; k.hi = std::numeric_limits<int64_t>::max() / 4; small = false; }
        else { const int64_t span = hi - lo + 1; k.lo = std::max<int64_t>(0, lo - span); k.hi = hi + span; }
      }
      else if (k.varchar) { k.lo = 0; k.hi = std::max<int64_t>(7, 2 * (int64_t)k.strings.size() - 1); }
      else if (k.ranged) {
        if (k.seenLo <= k.seenHi) padRange(k.seenLo, k.seenHi, k.lo, k.hi); else { k.lo = 0; k.hi = 0; }
        if (cardOf(k) > 4096) small = false;
      }
      else if (k.type->kind() == TypeKind::TINYINT) { k.lo = -128; k.hi = 127; }
      else small = false;
      if (small) { product *= cardOf(k); if (product > 4096) small = false; }
    }
    if (small && !forceHash) mode = VGPU_AGG_ARRAY;
    std::vector<vgpu_key_spec> specs(keys.size());
    if (mode == VGPU_AGG_HASH) {
      if (keys.size() == 1 && (!keys[0].varchar || keys[0].number)) {
        keys[0].bits = 64;
        if (!keys[0].varchar) outputs[0] = castTo(outputs[0], BIGINT());
      } else {
        int total = 0;
        for (auto& k : keys) {
          if (k.number) {
            VB_CHECK(k.hi < (1ll << 40), "a VARCHAR key with long strings next to other grouping keys needs multi-word keys (not supported)");
            k.bits = bitsFor(cardOf(k));
          }
          else if (k.varchar) { k.lo = 0; k.hi = (1 << 20) - 2; k.bits = 20; }
          else if (k.ranged) k.bits = cardOf(k) < (1ll << 62) ? bitsFor(cardOf(k)) : 64; // (too wide with its padding: retried without, below)
          else {
            VB_CHECK(k.type->kind() != TypeKind::BIGINT && k.type->isInteger(), "multi-key grouping on " + k.type->toString() + " needs key statistics (not supported yet)");
            k.lo = typeLo(k.type->kind()); k.hi = typeHi(k.type->kind());
            k.bits = k.type->kind() == TypeKind::TINYINT ? 9 : k.type->kind() == TypeKind::SMALLINT ? 17 : 33;
          }
          total += k.bits;
        }
        if (total > 63) { // give up the padding of the ranged keys before giving up
          total = 0;
          for (auto& k : keys) {
            if (k.ranged && k.seenLo <= k.seenHi) {
              k.lo = k.seenLo; k.hi = k.seenHi;
              VB_CHECK(cardOf(k) < (1ll << 62), "the range of grouping key " + k.name + " does not fit a normalized key (multi-word keys are not supported)");
              k.bits = bitsFor(cardOf(k));
            }
            total += k.bits;
          }
        }
        VB_CHECK(total <= 63, "grouping keys do not fit the 63 bits of a normalized key (multi-word keys are not supported)");
      }
    }
    // a number key is read through a BIGINT operand: the column's StringViews (or host-made numbers) are bound under it
    for (size_t i = 0; i < keys.size(); ++i) if (keys[i].number) outputs[i] = std::make_shared<FieldAccessTypedExpr>(BIGINT(), keyColumn(i));
    for (size_t i = 0; i < keys.size(); ++i) specs[i] = vgpu_key_spec{(int32_t)i, mode == VGPU_AGG_HASH ? keys[i].bits : 0, keys[i].lo, keys[i].hi};
    lowered = lowerExpressions(hasFused ? fused.filter : nullptr, outputs);
    // string comparisons need the FilterProject operator's value-id dictionary: the adapter keeps the two operators
    if (!lowered.stringColumns.empty() || !lowered.stringLiterals.empty() || !lowered.stringPredicates.empty()) throw VeloxUserError("VARCHAR comparisons are evaluated by GpuFilterProject, not inside the fused aggregation");
    const auto pdesc = lowered.desc();
    check(vgpu_program_create(&pdesc, &program), "vgpu_program_create");
    vgpu_agg_desc d{};
    d.mode = mode; d.n_keys = (int32_t)keys.size(); d.keys = specs.data();
    d.n_accs = (int32_t)accs.size(); d.accs = accs.data();
    d.ignore_null_keys = node->ignoreNullKeys() ? 1 : 0;
    check(vgpu_agg_create(&d, &agg), "vgpu_agg_create");
  }

  // A VARCHAR key outgrew the range the table was laid out for (ids past the dictionary bound, numbers outside
  // [lo, hi]): the groups are read back, the table is rebuilt for the ranges known now — the small-domain array may
  // become the hash table here, HashTable::decideHashMode's kArray -> kNormalizedKey / kHash step — and the groups
  // are merged back under their re-encoded keys.
  void regrow() {
    checkErrors();
    std::vector<int64_t> old = extractAll();
    const int rw = rowWords();
    const int64_t n = (int64_t)(old.size() / rw);
    std::vector<std::vector<int64_t>> kv;
    std::vector<std::vector<bool>> kn;
    decodeKeys(old.data(), n, kv, kn);
    vgpu_agg_destroy(agg); agg = nullptr;
    vgpu_program_destroy(program); program = nullptr;
    create();
    const int64_t nulls = nullIds();
    for (int64_t g = 0; g < n; ++g) {
      int64_t key = 0, flags = 0, mul = 1;
      int shift = 0;
      for (size_t i = 0; i < keys.size(); ++i) {
        const auto& k = keys[i];
        const bool isNull = kn[i][(size_t)g];
        const int64_t v = kv[i][(size_t)g];
        if (mode == VGPU_AGG_HASH && k.bits == 64) { key = isNull ? 0 : v; flags = isNull ? 1 : 0; continue; }
        VB_CHECK(isNull || (v >= k.lo && v <= k.hi), "group key outside the rebuilt key range");
        const int64_t id = isNull ? 0 : v - k.lo + nulls;
        if (mode == VGPU_AGG_ARRAY) { key += id * mul; mul *= cardOf(k); }
        else { key |= id << shift; shift += k.bits; }
      }
      old[(size_t)g * rw] = key;
      old[(size_t)g * rw + 1] = flags;
    }
    addEmpty();
    if (!old.empty()) {
      DeviceBatch b; b.stream = stream;
      void* d = b.upload(old.data(), old.size() * 8);
      check(vgpu_agg_merge_rows(agg, d, n, stream), "vgpu_agg_merge_rows");
      check(vgpu_stream_sync(stream), "vgpu_stream_sync");
    }
    ++stats.regrows;
  }

  // Packed group rows -> per key: value (an id for VARCHAR id keys, a number for VARCHAR number keys) and NULL flag.
  void decodeKeys(const int64_t* base, int64_t n, std::vector<std::vector<int64_t>>& kv, std::vector<std::vector<bool>>& kn) const {
    const int rw = rowWords();
    kv.assign(keys.size(), std::vector<int64_t>((size_t)n));
    kn.assign(keys.size(), std::vector<bool>((size_t)n, false));
    const int64_t nulls = nullIds();
    for (int64_t g = 0; g < n; ++g) {
      const int64_t key = base[g * rw], flags = base[g * rw + 1];
      if (mode == VGPU_AGG_ARRAY) {
        int64_t id = key;
        for (size_t i = 0; i < keys.size(); ++i) {
          const int64_t card = cardOf(keys[i]), d = id % card;
          id /= card;
          kn[i][(size_t)g] = nulls && d == 0;
          kv[i][(size_t)g] = d - nulls + keys[i].lo;
        }
      } else if (mode == VGPU_AGG_HASH) {
        int shift = 0;
        for (size_t i = 0; i < keys.size(); ++i) {
          if (keys[i].bits == 64) { kn[i][(size_t)g] = flags & 1; kv[i][(size_t)g] = key; continue; }
          const int64_t d = (int64_t)(((uint64_t)key >> shift) & ((1ull << keys[i].bits) - 1));
          shift += keys[i].bits;
          kn[i][(size_t)g] = nulls && d == 0;
          kv[i][(size_t)g] = d - nulls + keys[i].lo;
        }
      }
    }
  }
  // name of the input column a VARCHAR key reads
  std::string keyColumn(size_t i) const { return static_cast<const FieldAccessTypedExpr*>(outputs[i].get())->name(); }
  // device type an input column is bound as by the coalescing path
  int columnDeviceType(const std::string& name) const {
    if (ordDicts.count(name)) return VGPU_BIGINT;
    for (size_t i = 0; i < keys.size(); ++i) if (keys[i].number && keyColumn(i) == name) return VGPU_VARCHAR;
    return deviceType(inputColumnType(name));
  }
  // does a key value outside [lo, hi] break the table's layout?
  bool needsRange(const KeyInfo& k) const { return mode == VGPU_AGG_ARRAY || (mode == VGPU_AGG_HASH && k.bits != 64); }

  void addEmpty() {
    std::vector<vgpu_column> cols(std::max<size_t>(1, lowered.columns.size()));
    for (size_t i = 0; i < lowered.columns.size(); ++i) cols[i].type = columnDeviceType(lowered.columns[i]);
    check(vgpu_agg_add_input(agg, program, cols.data(), (int32_t)lowered.columns.size(), 0, nullptr, 0, stream), "vgpu_agg_add_input");
  }

  std::vector<int64_t> extractAll() {
    int64_t n = 0;
    check(vgpu_agg_num_groups(agg, stream, &n), "vgpu_agg_num_groups");
    std::vector<int64_t> out((size_t)n * rowWords());
    if (n == 0) return out;
    void* d = nullptr;
    check(vgpu_alloc(out.size() * 8, &d), "vgpu_alloc");
    int64_t got = 0;
    int rc = vgpu_agg_extract(agg, 0, n, d, &got, stream);
    if (rc == VGPU_OK) rc = vgpu_d2h_async(out.data(), d, (size_t)got * rowWords() * 8, stream);
    if (rc == VGPU_OK) rc = vgpu_stream_sync(stream);
    vgpu_free(d);
    check(rc, "vgpu_agg_extract");
    out.resize((size_t)got * rowWords());
    return out;
  }

  // ---- order-preserving ids for min / max over VARCHAR -----------------------------------------------------------
  // Id of `s`, inserting it if new.  Returns false when the gap between its neighbours is used up (nothing inserted).
  static bool ordInsert(OrdDict& d, const std::string& s) {
    auto it = d.idOf.lower_bound(s);
    if (it != d.idOf.end() && it->first == s) return true;
    int64_t id = 0;
    if (d.idOf.empty()) id = 0;
    else if (it == d.idOf.begin()) id = it->second - OrdDict::kStep;
    else if (it == d.idOf.end()) id = std::prev(it)->second + OrdDict::kStep;
    else {
      const int64_t lo = std::prev(it)->second, hi = it->second;
      if (hi - lo < 2) return false;
      id = lo + (hi - lo) / 2;
    }
    d.idOf.emplace_hint(it, s, id);
    d.strOf[id] = s;
    return true;
  }
  // Spreads the ids out again (kStep apart, centred on 0) and rewrites the accumulators that hold ids of `column`.
  void ordRenumber(const std::string& column) {
    OrdDict& d = ordDicts[column];
    std::unordered_map<int64_t, int64_t> moved;
    int64_t id = -(int64_t)(d.idOf.size() / 2) * OrdDict::kStep;
    d.strOf.clear();
    for (auto& kv : d.idOf) { moved[kv.second] = id; kv.second = id; d.strOf[id] = kv.first; id += OrdDict::kStep; }
    ++stats.regrows;
    if (!agg || !sawInput) return;
    flushPending();
    checkErrors();
    std::vector<int64_t> rowsNow = extractAll();
    const int rw = rowWords(), na = (int)accs.size();
    const int64_t n = (int64_t)(rowsNow.size() / rw);
    for (auto& a : aggs) {
      if (a.ordColumn != column) continue;
      for (int64_t g = 0; g < n; ++g) {
        if (rowsNow[(size_t)g * rw + 3 + na + a.acc0] == 0) continue; // NULL so far
        auto it = moved.find(rowsNow[(size_t)g * rw + 3 + a.acc0]);
        VB_CHECK(it != moved.end(), "min / max accumulator holds an id that is not in the column's dictionary");
        rowsNow[(size_t)g * rw + 3 + a.acc0] = it->second;
      }
    }
    check(vgpu_agg_reset(agg, stream), "vgpu_agg_reset");
    if (n > 0) {
      DeviceBatch b; b.stream = stream;
      void* dRows = b.upload(rowsNow.data(), rowsNow.size() * 8);
      check(vgpu_agg_merge_rows(agg, dRows, n, stream), "vgpu_agg_merge_rows");
      check(vgpu_stream_sync(stream), "vgpu_stream_sync");
    }
  }
  // Ids of the batch's BASE strings of `column` (a flat BIGINT vector of the base's length; NULL bases get 0).
  std::shared_ptr<FlatVector<int64_t>> ordIdsOf(const std::string& column, const VectorPtr& v) {
    OrdDict& d = ordDicts[column];
    const BaseVector* base = flatBaseOf(v.get());
    const uint64_t* nn = base->mayHaveNulls() ? base->nulls().data() : nullptr;
    for (size_t i = 0; i < base->size(); ++i) {
      if (nn && !bits::isBitSet(nn, i)) continue;
      const std::string s = stringValueAt(base, i);
      if (!ordInsert(d, s)) { ordRenumber(column); VB_CHECK(ordInsert(d, s), "order-preserving id space exhausted"); }
    }
    auto ids = std::make_shared<FlatVector<int64_t>>(BIGINT(), base->size());
    for (size_t i = 0; i < base->size(); ++i) ids->set(i, nn && !bits::isBitSet(nn, i) ? 0 : d.idOf.find(stringValueAt(base, i))->second);
    return ids;
  }

  // min / max of an INTEGER / BIGINT key column's (base) values into the key's seen range
  static void scanRange(KeyInfo& k, const VectorPtr& v) {
    const BaseVector* base = flatBaseOf(v.get());
    const uint64_t* nn = base->mayHaveNulls() ? base->nulls().data() : nullptr;
    int64_t lo = k.seenLo, hi = k.seenHi;
    auto scan = [&](auto* raw) {
      for (size_t i = 0; i < base->size(); ++i) {
        if (nn && !bits::isBitSet(nn, i)) continue;
        const int64_t x = (int64_t)raw[i];
        lo = std::min(lo, x); hi = std::max(hi, x);
      }
    };
    if (base->type()->kind() == TypeKind::BIGINT) scan(static_cast<const FlatVector<int64_t>*>(base)->rawValues());
    else scan(static_cast<const FlatVector<int32_t>*>(base)->rawValues());
    k.seenLo = lo; k.seenHi = hi;
  }

  // VARCHAR key column -> ids of its base strings (new strings get the next id)
  std::vector<int32_t> idsOf(KeyInfo& k, const VectorPtr& v) {
    // the flat vector of strings at the bottom of the wrappers (no per-row work for a dictionary: one lookup per base string)
    const BaseVector* base = flatBaseOf(v.get());
    std::vector<int32_t> ids(base->size());
    for (size_t i = 0; i < ids.size(); ++i) {
      if (base->mayHaveNulls() && !bits::isBitSet(base->nulls().data(), i)) { ids[i] = 0; continue; }
      const std::string value = stringValueAt(base, i);
      auto it = k.idOf.find(value);
      if (it == k.idOf.end()) { it = k.idOf.emplace(value, (int32_t)k.strings.size()).first; k.strings.push_back(value); }
      ids[i] = it->second;
    }
    return ids;
  }

  // VARCHAR number key: looks at the batch's BASE strings once.  Views of strings of at most 7 bytes need nothing but
  // their range (the device computes the numbers); otherwise the numbers are made here, long strings getting
  // kLongBase + their index in the key's own list.
  NumberBatch numbersOf(KeyInfo& k, const VectorPtr& v, bool dryRun = false) {
    const BaseVector* base = flatBaseOf(v.get());
    NumberBatch out;
    int64_t lo = k.seenLo, hi = k.seenHi;
    const uint64_t* nn = base->mayHaveNulls() ? base->nulls().data() : nullptr;
    if (auto sv = dynamic_cast<const StringViewVector*>(base)) {
      const StringView* views = sv->rawValues();
      bool anyLong = false;
      for (size_t i = 0; i < base->size() && !anyLong; ++i) {
        if (nn && !bits::isBitSet(nn, i)) continue;
        const uint32_t size = views[i].size();
        if (size > 7) { anyLong = true; break; }
        uint64_t word;
        memcpy(&word, views[i].data(), 8); // inline bytes; zero past the end
        const int64_t number = size == 0 ? 0 : (int64_t)((word & ((1ull << (8 * size)) - 1)) + (1ull << (8 * size)));
        lo = std::min(lo, number); hi = std::max(hi, number);
      }
      out.views = !anyLong;
    }
    if (!out.views) {
      if (dryRun) return out;
      lo = k.seenLo; hi = k.seenHi;
      out.numbers = std::make_shared<FlatVector<int64_t>>(BIGINT(), base->size());
      for (size_t i = 0; i < base->size(); ++i) {
        if (nn && !bits::isBitSet(nn, i)) { out.numbers->set(i, 0); continue; }
        const std::string value = stringValueAt(base, i);
        int64_t number = vgpu_number_of_string(value.data(), (uint32_t)value.size());
        if (number < 0) {
          auto it = k.longId.find(value);
          if (it == k.longId.end()) { it = k.longId.emplace(value, kLongBase + (int64_t)k.longStrings.size()).first; k.longStrings.push_back(value); }
          number = it->second;
        }
        out.numbers->set(i, number);
        lo = std::min(lo, number); hi = std::max(hi, number);
      }
    }
    if (!dryRun) { k.seenLo = lo; k.seenHi = hi; }
    return out;
  }
  std::string stringOfNumber(const KeyInfo& k, int64_t number) const {
    if (number >= kLongBase) {
      VB_CHECK((size_t)(number - kLongBase) < k.longStrings.size(), "group key number outside the long-string list");
      return k.longStrings[(size_t)(number - kLongBase)];
    }
    char buf[8];
    const int32_t size = vgpu_string_of_number(number, buf);
    VB_CHECK(size >= 0, "group key is not the number of a string");
    return std::string(buf, (size_t)size);
  }

  VectorPtr columnByName(const RowVectorPtr& in, const std::string& name, std::vector<VectorPtr>& keepAlive) {
    const size_t dot = name.find('.');
    if (dot == std::string::npos) return in->childAt(inputType->getChildIdx(name));
    auto parent = in->childAt(inputType->getChildIdx(name.substr(0, dot)));
    auto row = std::dynamic_pointer_cast<RowVector>(parent);
    VB_CHECK(row != nullptr, "intermediate avg column must be a flat ROW vector");
    auto child = row->childAt(row->rowType()->getChildIdx(name.substr(dot + 1)));
    if (!row->mayHaveNulls()) return child;
    // a NULL row hides its fields: fold the row's nulls into a copy of the child
    VB_CHECK(child->encoding() == VectorEncoding::FLAT, "intermediate avg fields must be flat");
    VectorPtr copy;
    if (child->type()->kind() == TypeKind::DOUBLE) copy = std::make_shared<FlatVector<double>>(*static_cast<FlatVector<double>*>(child.get()));
    else copy = std::make_shared<FlatVector<int64_t>>(*static_cast<FlatVector<int64_t>*>(child.get()));
    auto& nl = copy->mutableNulls();
    for (size_t w = 0; w < nl.size(); ++w) nl[w] &= row->nulls()[w];
    keepAlive.push_back(copy);
    return copy;
  }

  // ---- one device launch over `batch` ----------------------------------------------------------------------------
  void launch(DeviceBatch& batch, int64_t nrows) {
    staging->ready();
    int rc = vgpu_agg_add_input(agg, program, batch.cols.data(), (int32_t)batch.cols.size(), nrows, nullptr, 0, stream);
    if (rc == VGPU_ERR_UNSUPPORTED && mode == VGPU_AGG_ARRAY && !sawInput) {
      // HashTable::decideHashMode analogue: fall from the array mode to the hash mode (never to a CPU path)
      vgpu_agg_destroy(agg); agg = nullptr;
      vgpu_program_destroy(program); program = nullptr;
      forceHash = true;
      create();
      rc = vgpu_agg_add_input(agg, program, batch.cols.data(), (int32_t)batch.cols.size(), nrows, nullptr, 0, stream);
    }
    staging->release();
    check(rc, "vgpu_agg_add_input");
    ++stats.deviceLaunches;
    sawInput = true;
    // data errors (sum overflow, checked arithmetic in fused projections) surface like the exception thrown out of
    // HashAggregation::addInput, a few batches late at most
    if (++launchesSinceCheck >= std::max(1, ctx->queryConfig.errorCheckEvery)) checkErrors();
  }
  void checkErrors() {
    launchesSinceCheck = 0;
    check(vgpu_agg_check_errors(agg, stream), "HashAggregation");
  }

  // ---- coalescing ----------------------------------------------------------------------------------------------
  size_t pendingBytes(int64_t cap) const {
    size_t bytes = 0;
    for (auto& name : lowered.columns) {
      const int t = columnDeviceType(name);
      bytes += (size_t)cap * std::max<size_t>(deviceWidth(t), 4) + bits::nwords((size_t)cap) * 8 + 1024;
    }
    return bytes + (1 << 16);
  }
  void openPending(int64_t atLeastRows) {
    pendingCap = std::max<int64_t>(ctx->queryConfig.coalesceRows, atLeastRows);
    // room for the gathered columns plus, per VARCHAR key batch, its raw indices and id table (transient)
    staging->begin(pendingBytes(pendingCap) + (size_t)pendingCap * 4 * keys.size() + (1 << 20));
    pending.assign(lowered.columns.size(), PendingColumn{});
    for (size_t c = 0; c < lowered.columns.size(); ++c) {
      auto& pc = pending[c];
      pc.type = columnDeviceType(lowered.columns[c]);
      pc.width = deviceWidth(pc.type);
      pc.dValues = (char*)staging->reserve((size_t)pendingCap * pc.width);
      pc.hValues = (char*)staging->mirror(pc.dValues);
    }
    pendingRows = 0;
  }
  // rows [pc.hostFrom, pendingRows) of a column wait in the pinned mirror: one DMA moves them
  void pushMirror(PendingColumn& pc) {
    if (pc.hostFrom < 0) return;
    staging->copyMirror(pc.dValues + (size_t)pc.hostFrom * pc.width, (size_t)(pendingRows - pc.hostFrom) * pc.width);
    pc.hostFrom = -1;
  }
  // FLAT numeric columns and VARCHAR keys (flat or one dictionary over flat strings) can be gathered
  bool canCoalesce(const RowVectorPtr& in, const std::vector<VectorPtr>& vecs) const {
    for (size_t c = 0; c < vecs.size(); ++c) {
      const auto& v = vecs[c];
      const auto kind = v->type()->kind();
      if (kind == TypeKind::BOOLEAN) return false; // bit-packed values: appended rarely enough to take the direct path
      if (kind == TypeKind::VARCHAR) {
        if (v->encoding() == VectorEncoding::FLAT) continue;
        if (v->encoding() == VectorEncoding::DICTIONARY && static_cast<const DictionaryVector*>(v.get())->valueVector()->encoding() == VectorEncoding::FLAT) continue;
        return false;
      }
      if (v->encoding() != VectorEncoding::FLAT) return false;
    }
    return in->size() > 0;
  }
  static void appendBits(std::vector<uint64_t>& dst, size_t at, const uint64_t* src, size_t n) {
    dst.resize(bits::nwords(at + n), 0);
    for (size_t i = 0; i < n; ++i) if (!src || bits::isBitSet(src, i)) bits::setBit(dst.data(), at + i, true);
  }
  void appendPending(const RowVectorPtr& in, const std::vector<VectorPtr>& vecs, const std::vector<int>& varcharKeyOf,
                     const std::vector<std::vector<int32_t>>& keyIds) {
    const size_t n = in->size();
    if (pendingCap > 0 && pendingRows + (int64_t)n > pendingCap) flushPending();
    if (pendingCap == 0) openPending((int64_t)n);
    std::shared_ptr<const void> owner = in; // keeps pinned vector memory alive while the DMA is in flight
    // Small batches (Velox's default scan batch is ~4 k rows) cost more in CUDA calls than in bytes: they are gathered
    // on the HOST into the pinned mirror of the device arrays — no CUDA call per batch — and one DMA per column moves
    // them at flush time.  Large batches are copied straight out of their vectors (page-locked) or through the ring.
    const bool small = n < 32768;
    for (size_t c = 0; c < vecs.size(); ++c) {
      auto& pc = pending[c];
      const auto& v = vecs[c];
      char* dst = pc.dValues + (size_t)pendingRows * pc.width;
      char* hdst = pc.hValues + (size_t)pendingRows * pc.width;
      if (!small) pushMirror(pc); // keep the column's rows in order: earlier small batches go first
      else if (pc.hostFrom < 0) pc.hostFrom = pendingRows;
      const uint64_t* rowNulls = nullptr;
      std::vector<uint64_t> merged;
      if (varcharKeyOf[c] >= 0 && keys[(size_t)varcharKeyOf[c]].number) { // flat StringViews, as they are (addBatch checked)
        if (small) memcpy(hdst, rawValuesOf(v.get()), n * pc.width); else staging->copyIn(dst, rawValuesOf(v.get()), n * pc.width, owner);
        if (v->mayHaveNulls()) rowNulls = v->nulls().data();
      } else if (varcharKeyOf[c] >= 0) {
        const auto& ids = keyIds[varcharKeyOf[c]]; // id per BASE string of this batch
        if (v->encoding() == VectorEncoding::FLAT) { // one id per row already
          if (small) memcpy(hdst, ids.data(), n * 4); else staging->copyIn(dst, ids.data(), n * 4);
          if (v->mayHaveNulls()) rowNulls = v->nulls().data();
        } else {
          auto dict = static_cast<const DictionaryVector*>(v.get());
          const BaseVector* base = dict->valueVector().get();
          if (small) {
            int32_t* out = (int32_t*)hdst;
            const int32_t* ix = dict->rawIndices();
            for (size_t i = 0; i < n; ++i) out[i] = ids[(size_t)ix[i]];
          } else {
            void* dIdx = staging->upload(dict->rawIndices(), n * 4, owner);
            void* dTab = staging->upload(ids.data(), ids.size() * 4);
            check(vgpu_translate_i32((int32_t*)dst, (const int32_t*)dIdx, (const int32_t*)dTab, (int32_t)ids.size(), (int64_t)n, staging->copyStream()), "vgpu_translate_i32");
          }
          if (v->mayHaveNulls() || base->mayHaveNulls()) { // NULL when the wrapper or the base string is NULL
            merged.assign(bits::nwords(n), 0);
            for (size_t i = 0; i < n; ++i) {
              const bool ok = (!v->mayHaveNulls() || bits::isBitSet(v->nulls().data(), i)) &&
                              (!base->mayHaveNulls() || bits::isBitSet(base->nulls().data(), (size_t)dict->rawIndices()[i]));
              if (ok) bits::setBit(merged.data(), i, true);
            }
            rowNulls = merged.data();
          }
        }
      } else {
        if (small) memcpy(hdst, rawValuesOf(v.get()), n * pc.width); else staging->copyIn(dst, rawValuesOf(v.get()), n * pc.width, owner);
        if (v->mayHaveNulls()) rowNulls = v->nulls().data();
      }
      if (rowNulls && !pc.anyNulls) { pc.anyNulls = true; pc.notNull.assign(bits::nwords((size_t)pendingRows), ~0ull); }
      if (pc.anyNulls) appendBits(pc.notNull, (size_t)pendingRows, rowNulls, n);
    }
    pendingRows += (int64_t)n;
    if (pendingRows >= pendingCap) flushPending();
  }
  void flushPending() {
    if (pendingCap == 0) return;
    if (pendingRows > 0) {
      DeviceBatch batch;
      batch.stream = stream;
      batch.staging = staging.get();
      for (auto& pc : pending) {
        pushMirror(pc);
        vgpu_column c{};
        c.type = pc.type;
        c.encoding = VGPU_FLAT;
        c.values = pc.dValues;
        if (pc.anyNulls) c.nulls = (const uint64_t*)staging->upload(pc.notNull.data(), bits::nwords((size_t)pendingRows) * 8);
        batch.cols.push_back(c);
      }
      launch(batch, pendingRows);
    } else {
      staging->ready();
      staging->release();
    }
    pendingCap = 0;
    pendingRows = 0;
    pending.clear();
  }

  void addBatch(const RowVectorPtr& in) {
    const auto t0 = std::chrono::steady_clock::now();
    ++stats.inputVectors;
    stats.inputRows += (int64_t)in->size();
    std::vector<std::vector<int32_t>> keyIds(keys.size());
    std::vector<NumberBatch> keyNumbers(keys.size());
    std::vector<VectorPtr> keyVecs(keys.size());
    bool grew = false, direct = false;
    for (size_t i = 0; i < keys.size(); ++i) {
      auto& k = keys[i];
      if (k.ranged) {
        scanRange(k, in->childAt(inputType->getChildIdx(k.rangedColumn)));
        if (agg && k.seenLo <= k.seenHi && (k.seenLo < k.lo || k.seenHi > k.hi)) grew = true;
        continue;
      }
      if (!k.varchar) continue;
      keyVecs[i] = in->childAt(inputType->getChildIdx(keyColumn(i)));
      if (!agg && !sawInput && ctx->queryConfig.stringViewKeys) {
        // first batch: StringViews of short strings make this a number key — on its own, or next to other keys when
        // its range is small enough to pack
        const NumberBatch probe = numbersOf(k, keyVecs[i], /*dryRun=*/true);
        if (probe.views) {
          KeyInfo tmp = k;
          numbersOf(tmp, keyVecs[i]);
          k.number = keys.size() == 1 || tmp.seenLo > tmp.seenHi || tmp.seenHi - tmp.seenLo < 65536;
        }
      }
      if (k.number) {
        keyNumbers[i] = numbersOf(k, keyVecs[i]);
        if (!(keyNumbers[i].views && keyVecs[i]->encoding() == VectorEncoding::FLAT)) direct = true;
        if (agg && needsRange(k) && k.seenLo <= k.seenHi && (k.seenLo < k.lo || k.seenHi > k.hi)) grew = true;
      } else {
        keyIds[i] = idsOf(k, keyVecs[i]);
        if (agg && needsRange(k) && (int64_t)k.strings.size() - 1 > k.hi) grew = true;
      }
    }
    if (!staging) staging.reset(new Staging(stream));
    if (!agg) { create(); }
    else if (grew) {
      flushPending(); // rows gathered so far carry keys of the old layout
      regrow();
    }
    if (!ordDicts.empty()) direct = true; // their ids are made per batch: no gathering of small batches
    std::vector<VectorPtr> keepAlive;
    std::vector<VectorPtr> vecs(lowered.columns.size());
    std::vector<int> varcharKeyOf(lowered.columns.size(), -1);
    for (size_t c = 0; c < lowered.columns.size(); ++c) {
      const auto& name = lowered.columns[c];
      for (size_t i = 0; i < keys.size(); ++i) if (keys[i].varchar && keyColumn(i) == name) varcharKeyOf[c] = (int)i;
      vecs[c] = varcharKeyOf[c] >= 0 ? keyVecs[varcharKeyOf[c]] : columnByName(in, name, keepAlive);
    }
    std::vector<std::shared_ptr<FlatVector<int64_t>>> ordIds(lowered.columns.size());
    for (size_t c = 0; c < lowered.columns.size(); ++c) if (ordDicts.count(lowered.columns[c])) ordIds[c] = ordIdsOf(lowered.columns[c], vecs[c]);
    if (keepAlive.empty() && !direct && canCoalesce(in, vecs)) {
      appendPending(in, vecs, varcharKeyOf, keyIds);
    } else {
      // other encodings (dictionary / constant wrappers over numbers, booleans): this batch alone, through a slot of its own
      flushPending();
      size_t bytes = 1 << 16;
      for (auto& v : vecs) bytes += bindBytes(v);
      staging->begin(bytes);
      DeviceBatch batch;
      batch.stream = stream;
      batch.staging = staging.get();
      for (size_t c = 0; c < vecs.size(); ++c) {
        const int ki = varcharKeyOf[c];
        if (ki >= 0 && keys[(size_t)ki].number) {
          if (keyNumbers[(size_t)ki].views) bindColumn(batch, vecs[c], nullptr, nullptr, /*stringViews=*/true);
          else bindColumn(batch, vecs[c], nullptr, keyNumbers[(size_t)ki].numbers.get());
        }
        else if (ki >= 0) bindColumn(batch, vecs[c], &keyIds[(size_t)ki]);
        else if (ordDicts.count(lowered.columns[c])) bindColumn(batch, vecs[c], nullptr, ordIds[c].get());
        else bindColumn(batch, vecs[c]);
      }
      launch(batch, (int64_t)in->size());
    }
    rowsSinceFlush += (int64_t)in->size();
    stats.h2dBytes = staging->h2dBytes;
    stats.stagedBytes = staging->stagedBytes;
    stats.addInputWallMs += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
};

GpuHashAggregation::GpuHashAggregation(exec::DriverCtx* ctx, int32_t id, std::shared_ptr<const AggregationNode> node,
                                       const FilterProjectExprs* fused, RowTypePtr fusedInputType)
    : Operator(ctx, node->outputType(), id, node->id(), fused ? "GpuFilterProjectAggregation" : "GpuHashAggregation"), impl_(new Impl) {
  auto& m = *impl_;
  m.node = node;
  m.ctx = ctx;
  m.hasFused = fused != nullptr;
  if (fused) { m.fused = *fused; VB_CHECK(AggregationNode::isRawInput(node->step()), "only raw-input aggregations can absorb a FilterProject"); }
  m.inputType = fused ? fusedInputType : node->sources()[0]->outputType();
  check(vgpu_init(ctx->queryConfig.device), "vgpu_init");
  check(vgpu_stream_create(&m.stream), "vgpu_stream_create");
  for (auto& name : node->groupingKeys()) {
    KeyInfo k;
    k.name = name;
    auto e = m.exprOf(name);
    k.type = e->type();
    k.varchar = k.type->kind() == TypeKind::VARCHAR;
    VB_USER_CHECK(!k.varchar || e->kind() == ITypedExpr::Kind::kFieldAccess, "VARCHAR grouping keys must be plain columns");
    VB_USER_CHECK(k.varchar || k.type->isInteger(), "grouping on " + k.type->toString() + " is not supported on the GPU");
    m.keys.push_back(std::move(k));
    m.outputs.push_back(e);
  }
  if (m.keys.size() > 1) {
    for (size_t i = 0; i < m.keys.size(); ++i) {
      auto& k = m.keys[i];
      const auto kind = k.type->kind();
      if (k.varchar || (kind != TypeKind::INTEGER && kind != TypeKind::BIGINT) || m.outputs[i]->kind() != ITypedExpr::Kind::kFieldAccess) continue;
      const std::string column = static_cast<const FieldAccessTypedExpr*>(m.outputs[i].get())->name();
      if (m.inputType->getChildIdxIfExists(column) < 0) continue;
      k.ranged = true;
      k.rangedColumn = column;
    }
  }
  m.planAccumulators();
}

GpuHashAggregation::~GpuHashAggregation() { close(); }
const GpuHashAggregation::Stats& GpuHashAggregation::stats() const { return impl_->stats; }

void GpuHashAggregation::close() {
  auto& m = *impl_;
  m.staging.reset(); // waits for the copies / kernels still reading its slots
  if (m.agg) { vgpu_agg_destroy(m.agg); m.agg = nullptr; }
  if (m.program) { vgpu_program_destroy(m.program); m.program = nullptr; }
  if (m.stream) { vgpu_stream_destroy(m.stream); m.stream = nullptr; }
  Operator::close();
}

void GpuHashAggregation::addInput(RowVectorPtr input) {
  VB_CHECK(!noMoreInput_, "addInput after noMoreInput");
  VB_CHECK(!partialFull_, "addInput while the partial aggregation is flushing");
  auto& m = *impl_;
  m.addBatch(input);
  // Partial aggregation flushes when its state passes max_partial_aggregation_memory (HashAggregation.cpp:326-348): the
  // groups so far go downstream as partial results (the final aggregation merges duplicates) and the state starts over.
  if (m.node->step() == Step::kPartial && !m.keys.empty() && m.agg) {
    // Abandoning a partial aggregation that does not reduce its input (HashAggregation.cpp:229-233,253-258,355-364): once
    // more than abandon_partial_aggregation_min_rows rows came in and the groups are at least
    // abandon_partial_aggregation_min_pct % of them, the groups so far go downstream and from then on every device launch's
    // groups follow at once — the device still folds what one launch holds, which a partial stage may always do.  The
    // group count needs a finalize + read-back, so it is looked at when the row count passes the threshold and then
    // each time it has doubled, not per batch.
    auto& qc = m.ctx->queryConfig;
    bool flushNow = false;
    if (!m.stats.abandonedPartialAggregation && m.rowsSinceFlush > qc.abandonPartialAggregationMinRows && m.rowsSinceFlush >= m.nextAbandonCheck) {
      m.flushPending();
      m.checkErrors();
      int64_t groups = 0;
      check(vgpu_agg_num_groups(m.agg, m.stream, &groups), "vgpu_agg_num_groups");
      if (100 * groups / m.rowsSinceFlush >= qc.abandonPartialAggregationMinPct) { m.stats.abandonedPartialAggregation = true; flushNow = true; }
      else m.nextAbandonCheck = 2 * m.rowsSinceFlush;
    } else if (m.stats.abandonedPartialAggregation && m.stats.deviceLaunches != m.launchesAtLastFlush) {
      flushNow = true;
    }
    int64_t bytes = 0;
    check(vgpu_agg_memory_bytes(m.agg, &bytes), "vgpu_agg_memory_bytes");
    if (flushNow || bytes > m.ctx->queryConfig.maxPartialAggregationMemory) {
      m.flushPending();
      m.checkErrors();
      m.rows = m.extractAll();
      m.nGroups = (int64_t)(m.rows.size() / m.rowWords());
      m.cursor = 0;
      partialFull_ = m.nGroups > 0;
      if (!partialFull_) { check(vgpu_agg_reset(m.agg, m.stream), "vgpu_agg_reset"); m.nGroups = -1; }
      ++m.stats.flushes;
      m.launchesAtLastFlush = m.stats.deviceLaunches;
      if (!m.stats.abandonedPartialAggregation) { m.rowsSinceFlush = 0; m.nextAbandonCheck = 0; } // HashAggregation::resetPartialOutputIfNeed
    }
  }
}

void GpuHashAggregation::noMoreInput() {
  Operator::noMoreInput();
  auto& m = *impl_;
  if (m.agg) { m.flushPending(); if (m.sawInput) m.checkErrors(); }
  if (!m.sawInput) {
    if (!m.keys.empty()) { finished_ = true; return; } // grouped aggregation over no input: no rows (SumTest.cpp:509-516)
    m.create();                                        // global: exactly one row, even on empty input
    m.addEmpty();
  }
  m.rows = m.extractAll();
  m.nGroups = (int64_t)(m.rows.size() / m.rowWords());
  if (m.nGroups == 0) finished_ = true;
}

namespace {
template <typename T>
std::shared_ptr<FlatVector<T>> flatOf(const TypePtr& t, size_t n) { return std::make_shared<FlatVector<T>>(t, n); }
double asDouble(int64_t bitsValue) { double d; memcpy(&d, &bitsValue, 8); return d; }
} // namespace

RowVectorPtr GpuHashAggregation::getOutput() {
  auto& m = *impl_;
  if (partialFull_ && m.cursor >= m.nGroups) { // the flush is drained: start over
    check(vgpu_agg_reset(m.agg, m.stream), "vgpu_agg_reset");
    m.nGroups = -1; m.cursor = 0; m.rows.clear();
    partialFull_ = false;
    return nullptr;
  }
  if ((!noMoreInput_ && !partialFull_) || finished_ || m.nGroups < 0) return nullptr;
  const int64_t n = std::min<int64_t>(m.ctx->queryConfig.preferredOutputBatchRows, m.nGroups - m.cursor);
  const int rw = m.rowWords(), na = (int)m.accs.size();
  const int64_t* base = m.rows.data() + m.cursor * rw;
  std::vector<VectorPtr> children;
  // ---- grouping keys (normalized key -> values; NULL is a key value of its own, GroupingSet.cpp:240-307)
  std::vector<std::vector<int64_t>> kv;
  std::vector<std::vector<bool>> kn;
  m.decodeKeys(base, n, kv, kn);
  for (size_t i = 0; i < m.keys.size(); ++i) {
    const auto& k = m.keys[i];
    VectorPtr v;
    if (k.number) { auto f = flatOf<std::string>(k.type, (size_t)n); for (int64_t g = 0; g < n; ++g) if (!kn[i][g]) f->set(g, m.stringOfNumber(k, kv[i][g])); v = f; }
    else if (k.varchar) { auto f = flatOf<std::string>(k.type, (size_t)n); for (int64_t g = 0; g < n; ++g) if (!kn[i][g]) { VB_CHECK(kv[i][g] >= 0 && (size_t)kv[i][g] < k.strings.size(), "group key id outside the VARCHAR dictionary"); f->set(g, k.strings[(size_t)kv[i][g]]); } v = f; }
    else if (k.type->kind() == TypeKind::BIGINT) { auto f = flatOf<int64_t>(k.type, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, kv[i][g]); v = f; }
    else if (k.type->kind() == TypeKind::INTEGER) { auto f = flatOf<int32_t>(k.type, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (int32_t)kv[i][g]); v = f; }
    else if (k.type->kind() == TypeKind::SMALLINT) { auto f = flatOf<int16_t>(k.type, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (int16_t)kv[i][g]); v = f; }
    else { auto f = flatOf<int8_t>(k.type, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (int8_t)kv[i][g]); v = f; }
    for (int64_t g = 0; g < n; ++g) if (kn[i][g]) v->setNull(g, true);
    children.push_back(v);
  }
  // ---- aggregates: accumulators (partial / intermediate steps) or final values (Aggregate.h:35-41)
  const bool partialOut = AggregationNode::isPartialOutput(m.node->step());
  for (auto& a : m.aggs) {
    auto value = [&](int64_t g, int acc) { return base[g * rw + 3 + acc]; };
    auto nonNull = [&](int64_t g, int acc) { return base[g * rw + 3 + na + acc]; };
    VectorPtr v;
    if (a.fn == "count") {
      auto f = flatOf<int64_t>(BIGINT(), (size_t)n);
      for (int64_t g = 0; g < n; ++g) f->set(g, value(g, a.acc0)); // never NULL (CountAggregate.cpp:37-44)
      v = f;
    } else if (a.fn == "avg") {
      if (partialOut) {
        auto s = flatOf<double>(DOUBLE(), (size_t)n);
        auto c = flatOf<int64_t>(BIGINT(), (size_t)n);
        for (int64_t g = 0; g < n; ++g) { s->set(g, asDouble(value(g, a.acc0))); c->set(g, value(g, a.acc1)); }
        auto row = std::make_shared<RowVector>(std::dynamic_pointer_cast<const RowType>(a.outType), (size_t)n, std::vector<VectorPtr>{s, c});
        for (int64_t g = 0; g < n; ++g) if (value(g, a.acc1) == 0) row->setNull(g, true);
        v = row;
      } else {
        const bool real = a.outType->kind() == TypeKind::REAL;
        auto fd = real ? nullptr : flatOf<double>(a.outType, (size_t)n);
        auto ff = real ? flatOf<float>(a.outType, (size_t)n) : nullptr;
        for (int64_t g = 0; g < n; ++g) {
          const int64_t cnt = value(g, a.acc1);
          const double avg = cnt ? asDouble(value(g, a.acc0)) / (double)cnt : 0.0; // AverageAggregateBase.h:110
          if (real) ff->set(g, (float)avg); else fd->set(g, avg);
        }
        v = real ? VectorPtr(ff) : VectorPtr(fd);
        for (int64_t g = 0; g < n; ++g) if (value(g, a.acc1) == 0) v->setNull(g, true);
      }
    } else { // sum / min / max: NULL until a non-null input arrived (Aggregate.h:305-337)
      const int kind = m.accs[a.acc0].kind;
      const auto outKind = a.outType->kind();
      const bool accIsDouble = kind == VGPU_ACC_SUM_F64 || ((kind == VGPU_ACC_MIN || kind == VGPU_ACC_MAX) && m.outputs[m.accs[a.acc0].input]->type()->isFloating());
      switch (outKind) {
        case TypeKind::VARCHAR: { // min / max over VARCHAR: ids back to strings
          VB_CHECK(!a.ordColumn.empty(), "VARCHAR aggregate result without a dictionary");
          const OrdDict& dict = m.ordDicts.at(a.ordColumn);
          auto f = flatOf<std::string>(a.outType, (size_t)n);
          for (int64_t g = 0; g < n; ++g) {
            if (nonNull(g, a.acc0) == 0) continue;
            auto it = dict.strOf.find(value(g, a.acc0));
            VB_CHECK(it != dict.strOf.end(), "min / max result is not an id of the column's dictionary");
            f->set(g, it->second);
          }
          v = f;
          break;
        }
        case TypeKind::BOOLEAN: { auto f = std::make_shared<BoolVector>((size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, value(g, a.acc0) != 0); v = f; break; }
        case TypeKind::DOUBLE: { auto f = flatOf<double>(a.outType, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, asDouble(value(g, a.acc0))); v = f; break; }
        case TypeKind::REAL: { auto f = flatOf<float>(a.outType, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (float)asDouble(value(g, a.acc0))); v = f; break; }
        case TypeKind::BIGINT: { auto f = flatOf<int64_t>(a.outType, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, value(g, a.acc0)); v = f; break; }
        case TypeKind::INTEGER: { auto f = flatOf<int32_t>(a.outType, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (int32_t)value(g, a.acc0)); v = f; break; }
        case TypeKind::SMALLINT: { auto f = flatOf<int16_t>(a.outType, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (int16_t)value(g, a.acc0)); v = f; break; }
        case TypeKind::TINYINT: { auto f = flatOf<int8_t>(a.outType, (size_t)n); for (int64_t g = 0; g < n; ++g) f->set(g, (int8_t)value(g, a.acc0)); v = f; break; }
        default: throw VeloxRuntimeError("aggregate result type " + a.outType->toString() + " is not supported");
      }
      VB_CHECK(accIsDouble == a.outType->isFloating(), "accumulator / result type mismatch for " + a.fn);
      (void)outKind;
      for (int64_t g = 0; g < n; ++g) if (nonNull(g, a.acc0) == 0) v->setNull(g, true);
    }
    children.push_back(v);
  }
  m.cursor += n;
  ++m.stats.outputVectors;
  m.stats.outputRows += n;
  if (m.cursor >= m.nGroups && !partialFull_) finished_ = true;
  return std::make_shared<RowVector>(outputType_, (size_t)n, std::move(children));
}

} // namespace gpu
} // namespace vb200
