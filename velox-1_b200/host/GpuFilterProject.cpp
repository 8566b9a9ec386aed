#include "GpuDevice.h"

namespace vb200 {
namespace gpu {
using namespace core;

void check(int status, const char* what) {
  if (status == VGPU_OK) return;
  char msg[1024] = {0};
  vgpu_last_error(msg, sizeof(msg));
  const std::string text = std::string(what) + ": " + msg;
  if (status == VGPU_ERR_USER || status == VGPU_ERR_INVALID) throw VeloxUserError(text); // VELOX_USER_FAIL
  throw VeloxRuntimeError(text);                                                           // VELOX_FAIL
}

size_t deviceWidth(int t) {
  switch (t) {
    case VGPU_BOOLEAN: return 0;
    case VGPU_TINYINT: return 1;
    case VGPU_SMALLINT: return 2;
    case VGPU_INTEGER: case VGPU_REAL: return 4;
    case VGPU_VARCHAR: return 16;
    default: return 8;
  }
}

// DecodedVector::decode + combineWrappers: walk the wrappers from the outside in, composing indices
// and OR-ing the wrappers' extra nulls (indexed by the OUTER row).
Decoded decode(const VectorPtr& v) {
  Decoded d;
  d.size = v->size();
  const BaseVector* cur = v.get();
  bool identity = true;
  auto rowOf = [&](size_t i) -> int32_t { return identity ? (int32_t)i : d.indices[i]; };
  for (;;) {
    if (cur->encoding() == VectorEncoding::DICTIONARY) {
      auto dict = static_cast<const DictionaryVector*>(cur);
      if (cur->mayHaveNulls()) {
        if (d.nulls.empty()) d.nulls.assign(bits::nwords(d.size) ? bits::nwords(d.size) : 1, ~0ull);
        for (size_t i = 0; i < d.size; ++i) {
          if (d.constant ? !bits::isBitSet(cur->nulls().data(), 0) : !bits::isBitSet(cur->nulls().data(), rowOf(i))) bits::setBit(d.nulls.data(), i, false);
        }
      }
      std::vector<int32_t> next(d.size);
      for (size_t i = 0; i < d.size; ++i) next[i] = dict->rawIndices()[d.constant ? 0 : rowOf(i)];
      if (d.constant) { const int32_t only = dict->rawIndices()[0]; std::fill(next.begin(), next.end(), only); d.constant = false; }
      d.indices = std::move(next);
      identity = false;
      cur = dict->valueVector().get();
    } else if (cur->encoding() == VectorEncoding::CONSTANT) {
      auto c = static_cast<const ConstantVector*>(cur);
      if (c->isNullConstant()) { d.constantNull = true; d.constant = true; d.base = c->valueVector().get(); d.indices.clear(); return d; }
      if (identity) { d.constant = true; }
      else { std::fill(d.indices.begin(), d.indices.end(), 0); }
      cur = c->valueVector().get();
      if (identity) { d.base = cur; VB_CHECK(cur->encoding() == VectorEncoding::FLAT, "constant over a non-flat vector"); return d; }
    } else {
      VB_CHECK(cur->encoding() == VectorEncoding::FLAT, "unsupported vector encoding on the GPU path");
      d.base = cur;
      return d;
    }
  }
}

const void* rawValuesOf(const BaseVector* flat) {
  switch (flat->type()->kind()) {
    case TypeKind::BOOLEAN: return static_cast<const BoolVector*>(flat)->rawValues();
    case TypeKind::TINYINT: return static_cast<const FlatVector<int8_t>*>(flat)->rawValues();
    case TypeKind::SMALLINT: return static_cast<const FlatVector<int16_t>*>(flat)->rawValues();
    case TypeKind::INTEGER: return static_cast<const FlatVector<int32_t>*>(flat)->rawValues();
    case TypeKind::BIGINT: return static_cast<const FlatVector<int64_t>*>(flat)->rawValues();
    case TypeKind::REAL: return static_cast<const FlatVector<float>*>(flat)->rawValues();
    case TypeKind::DOUBLE: return static_cast<const FlatVector<double>*>(flat)->rawValues();
    case TypeKind::VARCHAR:
      if (auto sv = dynamic_cast<const StringViewVector*>(flat)) return sv->rawValues();
      throw VeloxRuntimeError("no raw device image for a VARCHAR vector that is not a StringView vector");
    default: throw VeloxRuntimeError("no raw device image for " + flat->type()->toString());
  }
}

size_t bindBytes(const VectorPtr& v) {
  // values + indices + two bitmaps, each padded to the staging alignment
  const BaseVector* cur = v.get();
  size_t n = v->size(), base = n;
  while (cur->encoding() == VectorEncoding::DICTIONARY) { cur = static_cast<const DictionaryVector*>(cur)->valueVector().get(); base = cur->size(); }
  if (cur->encoding() == VectorEncoding::CONSTANT) { cur = static_cast<const ConstantVector*>(cur)->valueVector().get(); base = cur->size(); }
  return base * 16 + n * 4 + 2 * (bits::nwords(std::max(n, base)) * 8) + 5 * 256;
}

void bindColumn(DeviceBatch& b, const VectorPtr& v, const std::vector<int32_t>* idsOfBase, const BaseVector* valuesOverride, bool stringViews) {
  Decoded d = decode(v);
  vgpu_column c{};
  VB_CHECK(!stringViews || dynamic_cast<const StringViewVector*>(d.base) != nullptr, "StringView binding of a vector that holds no StringViews");
  const int t = stringViews ? VGPU_VARCHAR : idsOfBase ? VGPU_INTEGER : deviceType(valuesOverride ? valuesOverride->type() : d.base->type());
  c.type = t;
  const size_t w = deviceWidth(t);
  const size_t baseN = d.base->size();
  auto baseBytes = [&](size_t n) { return w ? n * w : bits::nwords(n) * 8; };
  const void* baseRaw = idsOfBase ? (const void*)idsOfBase->data() : rawValuesOf(valuesOverride ? valuesOverride : d.base);
  if (d.constant) {
    c.encoding = VGPU_CONSTANT;
    c.constant_is_null = d.constantNull || (d.base->mayHaveNulls() && !bits::isBitSet(d.base->nulls().data(), 0));
    if (!c.constant_is_null) {
      if (idsOfBase) c.constant.i = (*idsOfBase)[0];
      else if (valuesOverride && valuesOverride->type()->kind() == TypeKind::BIGINT) c.constant.i = static_cast<const FlatVector<int64_t>*>(valuesOverride)->valueAt(0);
      else if (valuesOverride) c.constant.i = static_cast<const BoolVector*>(valuesOverride)->valueAt(0);
      else if (stringViews) { const auto& sv = static_cast<const StringViewVector*>(d.base)->valueAt(0); c.constant.i = vgpu_number_of_string(sv.data(), sv.size()); VB_CHECK(c.constant.i >= 0, "constant VARCHAR key longer than 7 bytes"); }
      else switch (d.base->type()->kind()) {
        case TypeKind::BOOLEAN: c.constant.i = static_cast<const BoolVector*>(d.base)->valueAt(0); break;
        case TypeKind::TINYINT: c.constant.i = static_cast<const FlatVector<int8_t>*>(d.base)->valueAt(0); break;
        case TypeKind::SMALLINT: c.constant.i = static_cast<const FlatVector<int16_t>*>(d.base)->valueAt(0); break;
        case TypeKind::INTEGER: c.constant.i = static_cast<const FlatVector<int32_t>*>(d.base)->valueAt(0); break;
        case TypeKind::BIGINT: c.constant.i = static_cast<const FlatVector<int64_t>*>(d.base)->valueAt(0); break;
        case TypeKind::REAL: c.constant.d = static_cast<const FlatVector<float>*>(d.base)->valueAt(0); break;
        default: c.constant.d = static_cast<const FlatVector<double>*>(d.base)->valueAt(0); break;
      }
    }
  } else if (d.indices.empty()) {
    c.encoding = VGPU_FLAT;
    c.values = b.upload(baseRaw, baseBytes(baseN));
    if (d.base->mayHaveNulls()) c.nulls = (const uint64_t*)b.upload(d.base->nulls().data(), bits::nwords(baseN) * 8);
  } else {
    c.encoding = VGPU_DICTIONARY;
    c.values = b.upload(baseRaw, baseBytes(baseN));
    c.indices = (const int32_t*)b.upload(d.indices.data(), d.indices.size() * 4);
    c.base_length = (int64_t)baseN;
    if (!d.nulls.empty()) c.nulls = (const uint64_t*)b.upload(d.nulls.data(), d.nulls.size() * 8);
    if (d.base->mayHaveNulls()) c.base_nulls = (const uint64_t*)b.upload(d.base->nulls().data(), bits::nwords(baseN) * 8);
  }
  b.cols.push_back(c);
}

// ======================================================================================================
// GpuFilterProject
// ======================================================================================================
GpuFilterProject::GpuFilterProject(exec::DriverCtx* ctx, int32_t id, std::shared_ptr<const FilterNode> filter,
                                   std::shared_ptr<const ProjectNode> project)
    : Operator(ctx, project ? project->outputType() : filter->outputType(), id, project ? project->id() : filter->id(), "GpuFilterProject") {
  VB_CHECK(filter || project, "GpuFilterProject needs a FilterNode or a ProjectNode");
  inputType_ = filter ? filter->sources()[0]->outputType() : project->sources()[0]->outputType();
  if (filter) exprs_.filter = filter->filter();
  if (project) { exprs_.names = project->names(); exprs_.projections = project->projections(); }
  else for (size_t i = 0; i < inputType_->size(); ++i) { // filter only: identity projection of every column
    exprs_.names.push_back(inputType_->nameOf(i));
    exprs_.projections.push_back(std::make_shared<FieldAccessTypedExpr>(inputType_->childAt(i), inputType_->nameOf(i)));
  }
  std::vector<TypedExprPtr> computed;
  for (auto& p : exprs_.projections) {
    if (p->kind() == ITypedExpr::Kind::kFieldAccess) { // identity projections pass columns through (FilterProject.cpp:59-103)
      identity_.push_back(inputType_->getChildIdx(static_cast<const FieldAccessTypedExpr*>(p.get())->name()));
      programOutput_.push_back(-1);
    } else {
      identity_.push_back(-1);
      programOutput_.push_back((int)computed.size());
      computed.push_back(p);
    }
  }
  lowered_ = lowerExpressions(exprs_.filter, computed);
  for (size_t k = 0; k < lowered_.stringLiterals.size(); ++k) stringIds_.emplace(lowered_.stringLiterals[k], (int32_t)k);
  check(vgpu_init(ctx->queryConfig.device), "vgpu_init");
  const auto desc = lowered_.desc();
  check(vgpu_program_create(&desc, &program_), "vgpu_program_create");
  check(vgpu_stream_create(&stream_), "vgpu_stream_create");
}

GpuFilterProject::~GpuFilterProject() {
  if (program_) vgpu_program_destroy(program_);
  if (stream_) vgpu_stream_destroy(stream_);
}

namespace {
VectorPtr makeFlatFromDevice(const TypePtr& type, size_t n, const void* dValues, const uint64_t* dNulls, void* stream) {
  VectorPtr out;
  void* dst = nullptr;
  size_t bytes = 0;
  switch (type->kind()) {
    case TypeKind::BOOLEAN: { auto v = std::make_shared<BoolVector>(n); dst = v->mutableRawValues(); bytes = bits::nwords(n) * 8; out = v; break; }
    case TypeKind::TINYINT: { auto v = std::make_shared<FlatVector<int8_t>>(type, n); dst = v->mutableRawValues(); bytes = n; out = v; break; }
    case TypeKind::SMALLINT: { auto v = std::make_shared<FlatVector<int16_t>>(type, n); dst = v->mutableRawValues(); bytes = n * 2; out = v; break; }
    case TypeKind::INTEGER: { auto v = std::make_shared<FlatVector<int32_t>>(type, n); dst = v->mutableRawValues(); bytes = n * 4; out = v; break; }
    case TypeKind::BIGINT: { auto v = std::make_shared<FlatVector<int64_t>>(type, n); dst = v->mutableRawValues(); bytes = n * 8; out = v; break; }
    case TypeKind::REAL: { auto v = std::make_shared<FlatVector<float>>(type, n); dst = v->mutableRawValues(); bytes = n * 4; out = v; break; }
    case TypeKind::DOUBLE: { auto v = std::make_shared<FlatVector<double>>(type, n); dst = v->mutableRawValues(); bytes = n * 8; out = v; break; }
    default: throw VeloxRuntimeError("projection of type " + type->toString() + " is not supported");
  }
  check(vgpu_d2h_async(dst, dValues, bytes, stream), "vgpu_d2h_async");
  auto& nulls = out->mutableNulls();
  check(vgpu_d2h_async(nulls.data(), dNulls, bits::nwords(n) * 8, stream), "vgpu_d2h_async");
  return out;
}
} // namespace

RowVectorPtr GpuFilterProject::getOutput() {
  if (!input_) return nullptr;
  RowVectorPtr in = std::move(input_);
  input_ = nullptr;
  const size_t n = in->size();
  DeviceBatch batch;
  batch.stream = stream_;
  std::vector<VectorPtr> predicateBases; // keep the BOOLEAN base vectors alive until the copies are issued
  for (auto& name : lowered_.columns) {
    if (const auto* sp = lowered_.predicateOf(name)) {
      // like / < / > ... on a VARCHAR column: evaluated once per BASE string, bound as a BOOLEAN column over the batch's
      // own indices and nulls (a dictionary-encoded column costs no per-row host work)
      const auto& child = in->childAt(inputType_->getChildIdx(sp->column));
      const auto* base = decode(child).base;
      auto match = std::make_shared<BoolVector>(base->size());
      for (size_t i = 0; i < base->size(); ++i) match->set(i, LoweredProgram::evalStringPredicate(sp->fn, stringValueAt(base, i), sp->literal));
      predicateBases.push_back(match);
      bindColumn(batch, child, nullptr, match.get());
      continue;
    }
    const auto& child = in->childAt(inputType_->getChildIdx(name));
    if (child->type()->kind() == TypeKind::VARCHAR && lowered_.stringColumns.count(name)) {
      // compared with =, <>, IN: the base strings become value ids of the operator's dictionary (one lookup per BASE
      // value: a dictionary-encoded column costs nothing per row and keeps its indices / nulls as they are)
      const auto* base = decode(child).base;
      std::vector<int32_t> ids(base->size());
      for (size_t i = 0; i < ids.size(); ++i) {
        const std::string value = stringValueAt(base, i);
        auto it = stringIds_.find(value);
        if (it == stringIds_.end()) it = stringIds_.emplace(value, (int32_t)stringIds_.size()).first;
        ids[i] = it->second;
      }
      bindColumn(batch, child, &ids);
    } else if (child->type()->kind() == TypeKind::VARCHAR) {
      // only the null flags of a VARCHAR column can matter to a device expression (is_null); values travel as ids
      const std::vector<int32_t> zeros(decode(child).base->size(), 0);
      bindColumn(batch, child, &zeros);
    } else {
      bindColumn(batch, child);
    }
  }
  const size_t words = bits::nwords(n);
  auto* selBits = (uint64_t*)batch.scratch(words * 8, true);
  auto* selIdx = (int32_t*)batch.scratch(n * 4, false);
  auto* numSel = (int64_t*)batch.scratch(8, true);
  std::vector<vgpu_out_column> outs(lowered_.outputs.size());
  for (size_t k = 0; k < outs.size(); ++k) {
    const size_t w = deviceWidth(lowered_.outputs[k].type);
    outs[k].values = batch.scratch(w ? n * w : words * 8, w == 0);
    outs[k].nulls = (uint64_t*)batch.scratch(words * 8, true);
  }
  check(vgpu_filter_project(program_, batch.cols.data(), (int32_t)batch.cols.size(), (int64_t)n, nullptr, 0, selBits, selIdx, numSel,
                            outs.empty() ? nullptr : outs.data(), stream_), "vgpu_filter_project");
  check(vgpu_check_errors(stream_), "FilterProject"); // deferred per-row errors of rows that stayed selected
  int64_t numOut = 0;
  check(vgpu_d2h_async(&numOut, numSel, 8, stream_), "vgpu_d2h_async");
  check(vgpu_stream_sync(stream_), "vgpu_stream_sync");
  if (numOut == 0) return nullptr;
  const bool allSelected = (size_t)numOut == n;
  std::vector<int32_t> selected;
  if (!allSelected) {
    selected.resize((size_t)numOut);
    check(vgpu_d2h_async(selected.data(), selIdx, (size_t)numOut * 4, stream_), "vgpu_d2h_async");
  }
  std::vector<VectorPtr> children(exprs_.projections.size());
  for (size_t i = 0; i < children.size(); ++i) {
    if (identity_[i] >= 0) children[i] = in->childAt(identity_[i]);
    else children[i] = makeFlatFromDevice(exprs_.projections[i]->type(), n, outs[programOutput_[i]].values, outs[programOutput_[i]].nulls, stream_);
  }
  check(vgpu_stream_sync(stream_), "vgpu_stream_sync");
  // fillOutput (velox/exec/Operator.cpp:195-228): unless every row passed, columns are wrapped in a
  // dictionary over selectedIndices instead of being compacted.
  if (!allSelected) for (auto& c : children) c = wrapInDictionary(selected, c);
  return std::make_shared<RowVector>(outputType_, (size_t)numOut, std::move(children));
}

} // namespace gpu
} // namespace vb200
