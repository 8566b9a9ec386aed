// See PrestoPage.h for the layout and the reference lines it follows.
#include "PrestoPage.h"

#include <cstring>

#include "GpuDevice.h"

namespace vb200 {
namespace serializer {
namespace presto {
namespace {

constexpr int8_t kCompressedBitMask = 1, kEncryptedBitMask = 2, kCheckSumBitMask = 4; // PrestoSerializer.cpp:28-31
constexpr size_t kHeaderSize = 4 + 1 + 4 + 4 + 8;                                     // :2093

struct Out {
  std::string s;
  void i8(int8_t v) { s.push_back((char)v); }
  void i32(int32_t v) { s.append(reinterpret_cast<const char*>(&v), 4); }
  void i64(int64_t v) { s.append(reinterpret_cast<const char*>(&v), 8); }
  void raw(const void* p, size_t n) { s.append(reinterpret_cast<const char*>(p), n); }
  void name(const char* n) { i32((int32_t)strlen(n)); raw(n, strlen(n)); }
};

const char* encodingName(TypeKind k) { // typeToEncodingName, PrestoSerializer.cpp:102-137
  switch (k) {
    case TypeKind::BOOLEAN: case TypeKind::TINYINT: return "BYTE_ARRAY";
    case TypeKind::SMALLINT: return "SHORT_ARRAY";
    case TypeKind::INTEGER: case TypeKind::REAL: return "INT_ARRAY";
    case TypeKind::BIGINT: case TypeKind::DOUBLE: return "LONG_ARRAY";
    case TypeKind::VARCHAR: return "VARIABLE_WIDTH";
    case TypeKind::ROW: return "ROW";
    default: throw VeloxRuntimeError("PrestoPage: unsupported type");
  }
}
size_t fixedWidth(TypeKind k) {
  switch (k) {
    case TypeKind::BOOLEAN: case TypeKind::TINYINT: return 1;
    case TypeKind::SMALLINT: return 2;
    case TypeKind::INTEGER: case TypeKind::REAL: return 4;
    case TypeKind::BIGINT: case TypeKind::DOUBLE: return 8;
    default: return 0;
  }
}

// hasNulls(1) [| bits, 1 = NULL, most significant bit first] over `isNull`
void writeNulls(Out& o, const std::vector<bool>& isNull) {
  bool any = false;
  for (bool b : isNull) any = any || b;
  if (!any) { o.i8(0); return; }
  o.i8(1);
  std::string bits((isNull.size() + 7) / 8, '\0');
  for (size_t i = 0; i < isNull.size(); ++i) if (isNull[i]) bits[i >> 3] = (char)((uint8_t)bits[i >> 3] | (0x80u >> (i & 7)));
  o.raw(bits.data(), bits.size());
}

// Column `v` restricted to `rows` (row numbers of v), in that order.
void writeColumn(Out& o, const VectorPtr& v, const std::vector<size_t>& rows) {
  const TypeKind kind = v->type()->kind();
  o.name(encodingName(kind));
  if (kind == TypeKind::ROW) {
    auto row = std::dynamic_pointer_cast<RowVector>(v);
    VB_CHECK(row != nullptr, "PrestoPage: ROW column must be a RowVector");
    std::vector<bool> isNull(rows.size());
    std::vector<size_t> present;
    for (size_t i = 0; i < rows.size(); ++i) {
      isNull[i] = row->mayHaveNulls() && !bits::isBitSet(row->nulls().data(), rows[i]);
      if (!isNull[i]) present.push_back(rows[i]);
    }
    o.i32((int32_t)row->children().size());
    for (auto& child : row->children()) writeColumn(o, child, present);
    o.i32((int32_t)rows.size());
    int32_t off = 0;
    o.i32(0);
    for (size_t i = 0; i < rows.size(); ++i) { off += isNull[i] ? 0 : 1; o.i32(off); }
    writeNulls(o, isNull);
    return;
  }
  gpu::Decoded d = gpu::decode(v);
  auto baseRow = [&](size_t r) { return d.constant ? (size_t)0 : d.indices.empty() ? r : (size_t)d.indices[r]; };
  std::vector<bool> isNull(rows.size());
  for (size_t i = 0; i < rows.size(); ++i) {
    const size_t r = rows[i], b = baseRow(r);
    isNull[i] = d.constantNull || (!d.nulls.empty() && !bits::isBitSet(d.nulls.data(), r)) ||
                (d.base->mayHaveNulls() && !bits::isBitSet(d.base->nulls().data(), b));
  }
  o.i32((int32_t)rows.size());
  if (kind == TypeKind::VARCHAR) {
    std::string bytes;
    std::vector<int32_t> ends(rows.size());
    for (size_t i = 0; i < rows.size(); ++i) {
      if (!isNull[i]) bytes += stringValueAt(d.base, baseRow(rows[i]));
      ends[i] = (int32_t)bytes.size();
    }
    o.raw(ends.data(), ends.size() * 4);
    writeNulls(o, isNull);
    o.i32((int32_t)bytes.size());
    o.raw(bytes.data(), bytes.size());
    return;
  }
  writeNulls(o, isNull);
  const size_t w = fixedWidth(kind);
  if (kind == TypeKind::BOOLEAN) {
    auto bv = static_cast<const BoolVector*>(d.base);
    for (size_t i = 0; i < rows.size(); ++i) if (!isNull[i]) o.i8(bv->valueAt(baseRow(rows[i])) ? 1 : 0);
    return;
  }
  const char* raw = static_cast<const char*>(gpu::rawValuesOf(d.base));
  for (size_t i = 0; i < rows.size(); ++i) if (!isNull[i]) o.raw(raw + baseRow(rows[i]) * w, w);
}

std::string pageOf(const RowVectorPtr& rows, const std::vector<size_t>& which, bool withChecksum) {
  Out data;
  data.i32((int32_t)rows->children().size());
  for (auto& child : rows->children()) writeColumn(data, child, which);
  const int8_t codec = withChecksum ? kCheckSumBitMask : 0;
  const int32_t numRows = (int32_t)which.size(), size = (int32_t)data.s.size();
  int64_t checksum = 0;
  if (withChecksum) {
    uint32_t c = crc32(data.s.data(), data.s.size());
    c = crc32(&codec, 1, c);
    c = crc32(&numRows, 4, c);
    c = crc32(&size, 4, c);
    checksum = (int64_t)c;
  }
  Out page;
  page.i32(numRows); page.i8(codec); page.i32(size); page.i32(size); page.i64(checksum);
  page.s += data.s;
  return page.s;
}

} // namespace

uint32_t crc32(const void* data, size_t n, uint32_t seed) {
  static uint32_t table[256];
  static bool init = false;
  if (!init) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      table[i] = c;
    }
    init = true;
  }
  uint32_t c = ~seed;
  const uint8_t* p = static_cast<const uint8_t*>(data);
  for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xffu] ^ (c >> 8);
  return ~c;
}

std::string serializePage(const RowVectorPtr& rows, bool withChecksum) {
  std::vector<size_t> all(rows->size());
  for (size_t i = 0; i < all.size(); ++i) all[i] = i;
  return pageOf(rows, all, withChecksum);
}

std::vector<std::string> partitionAndSerialize(const RowVectorPtr& rows, int keyChannel, int numPartitions, bool withChecksum) {
  VB_CHECK(numPartitions >= 1 && keyChannel >= 0 && (size_t)keyChannel < rows->children().size(), "partitionAndSerialize: bad arguments");
  const auto& key = rows->childAt((size_t)keyChannel);
  VB_CHECK(key->type()->kind() == TypeKind::BIGINT, "partitionAndSerialize: the partition key must be BIGINT");
  gpu::Decoded d = gpu::decode(key);
  std::vector<std::vector<size_t>> which((size_t)numPartitions);
  for (size_t r = 0; r < rows->size(); ++r) {
    const size_t b = d.constant ? 0 : d.indices.empty() ? r : (size_t)d.indices[r];
    const bool isNull = d.constantNull || (!d.nulls.empty() && !bits::isBitSet(d.nulls.data(), r)) ||
                        (d.base->mayHaveNulls() && !bits::isBitSet(d.base->nulls().data(), b));
    int32_t part = 0;
    gpu::check(vgpu_partition_of(isNull ? 0 : static_cast<const FlatVector<int64_t>*>(d.base)->valueAt(b), isNull ? 1 : 0, numPartitions, 1, &part), "vgpu_partition_of");
    which[(size_t)part].push_back(r);
  }
  std::vector<std::string> pages((size_t)numPartitions);
  for (int p = 0; p < numPartitions; ++p) if (!which[(size_t)p].empty()) pages[(size_t)p] = pageOf(rows, which[(size_t)p], withChecksum);
  return pages;
}

// ---- reading ---------------------------------------------------------------------------------------------------------
namespace {
struct In {
  const std::string& s;
  size_t at = 0;
  explicit In(const std::string& str) : s(str) {}
  void need(size_t n) const { VB_CHECK(at + n <= s.size(), "PrestoPage: truncated page"); }
  int8_t i8() { need(1); return (int8_t)s[at++]; }
  int32_t i32() { need(4); int32_t v; memcpy(&v, s.data() + at, 4); at += 4; return v; }
  int64_t i64() { need(8); int64_t v; memcpy(&v, s.data() + at, 8); at += 8; return v; }
  const char* raw(size_t n) { need(n); const char* p = s.data() + at; at += n; return p; }
  std::string name() { const int32_t n = i32(); VB_CHECK(n >= 0 && n < 64, "PrestoPage: bad encoding name"); return std::string(raw((size_t)n), (size_t)n); }
};

std::vector<bool> readNulls(In& in, size_t n) {
  std::vector<bool> isNull(n, false);
  if (in.i8() == 0) return isNull;
  const uint8_t* bits = reinterpret_cast<const uint8_t*>(in.raw((n + 7) / 8));
  for (size_t i = 0; i < n; ++i) isNull[i] = (bits[i >> 3] & (0x80u >> (i & 7))) != 0;
  return isNull;
}

template <typename T>
VectorPtr readFixed(In& in, const TypePtr& type, size_t n, const std::vector<bool>& isNull) {
  auto v = std::make_shared<FlatVector<T>>(type, n);
  for (size_t i = 0; i < n; ++i) {
    if (isNull[i]) { v->setNull(i, true); continue; }
    T x;
    memcpy(&x, in.raw(sizeof(T)), sizeof(T));
    v->set(i, x);
  }
  return v;
}

VectorPtr readColumn(In& in, const TypePtr& type) {
  const TypeKind kind = type->kind();
  const std::string enc = in.name();
  VB_CHECK(enc == encodingName(kind), "PrestoPage: column encoding " + enc + " does not match type " + type->toString());
  if (kind == TypeKind::ROW) {
    auto rowType = std::dynamic_pointer_cast<const RowType>(type);
    const int32_t nChildren = in.i32();
    VB_CHECK(rowType && (size_t)nChildren == rowType->size(), "PrestoPage: ROW children do not match the type");
    std::vector<VectorPtr> dense;
    for (int32_t c = 0; c < nChildren; ++c) dense.push_back(readColumn(in, rowType->childAt((size_t)c)));
    const size_t n = (size_t)in.i32();
    std::vector<int32_t> offsets(n + 1);
    memcpy(offsets.data(), in.raw((n + 1) * 4), (n + 1) * 4);
    const std::vector<bool> isNull = readNulls(in, n);
    // children hold the non-null rows only: spread them back over all positions
    std::vector<int32_t> index(n, 0);
    Nulls extra(bits::nwords(n) ? bits::nwords(n) : 1, ~0ull);
    bool anyNull = false;
    for (size_t i = 0; i < n; ++i) {
      VB_CHECK(offsets[i] >= 0 && offsets[i + 1] - offsets[i] == (isNull[i] ? 0 : 1), "PrestoPage: bad ROW offsets");
      index[i] = isNull[i] ? 0 : offsets[i];
      if (isNull[i]) { bits::setBit(extra.data(), i, false); anyNull = true; }
    }
    std::vector<VectorPtr> children;
    for (auto& d : dense) {
      VB_CHECK(d->size() == (size_t)offsets[n], "PrestoPage: ROW child size");
      if (!anyNull) { children.push_back(d); continue; }
      children.push_back(wrapInDictionary(index, d, extra)); // (rows that are NULL never look at their index)
    }
    auto row = std::make_shared<RowVector>(rowType, n, std::move(children));
    for (size_t i = 0; i < n; ++i) if (isNull[i]) row->setNull(i, true);
    return row;
  }
  const size_t n = (size_t)in.i32();
  if (kind == TypeKind::VARCHAR) {
    std::vector<int32_t> ends(n);
    memcpy(ends.data(), in.raw(n * 4), n * 4);
    const std::vector<bool> isNull = readNulls(in, n);
    const int32_t total = in.i32();
    const char* bytes = in.raw((size_t)total);
    auto v = std::make_shared<FlatVector<std::string>>(type, n);
    int32_t prev = 0;
    for (size_t i = 0; i < n; ++i) {
      VB_CHECK(ends[i] >= prev && ends[i] <= total, "PrestoPage: bad VARIABLE_WIDTH offsets");
      if (isNull[i]) v->setNull(i, true); else v->set(i, std::string(bytes + prev, (size_t)(ends[i] - prev)));
      prev = ends[i];
    }
    return v;
  }
  const std::vector<bool> isNull = readNulls(in, n);
  switch (kind) {
    case TypeKind::BOOLEAN: {
      auto v = std::make_shared<BoolVector>(n);
      for (size_t i = 0; i < n; ++i) { if (isNull[i]) v->setNull(i, true); else v->set(i, in.i8() != 0); }
      return v;
    }
    case TypeKind::TINYINT: return readFixed<int8_t>(in, type, n, isNull);
    case TypeKind::SMALLINT: return readFixed<int16_t>(in, type, n, isNull);
    case TypeKind::INTEGER: return readFixed<int32_t>(in, type, n, isNull);
    case TypeKind::BIGINT: return readFixed<int64_t>(in, type, n, isNull);
    case TypeKind::REAL: return readFixed<float>(in, type, n, isNull);
    case TypeKind::DOUBLE: return readFixed<double>(in, type, n, isNull);
    default: throw VeloxRuntimeError("PrestoPage: unsupported type " + type->toString());
  }
}
} // namespace

RowVectorPtr deserializePage(const std::string& page, const RowTypePtr& type) {
  In in(page);
  VB_CHECK(page.size() >= kHeaderSize, "PrestoPage: shorter than its header");
  const int32_t numRows = in.i32();
  const int8_t codec = in.i8();
  const int32_t uncompressedSize = in.i32(), size = in.i32();
  const int64_t checksum = in.i64();
  VB_CHECK(!(codec & kCompressedBitMask) && !(codec & kEncryptedBitMask), "PrestoPage: compressed / encrypted pages are not supported");
  VB_CHECK(numRows >= 0 && size == uncompressedSize && (size_t)size == page.size() - kHeaderSize, "PrestoPage: sizes in the header do not match the page");
  if (codec & kCheckSumBitMask) {
    uint32_t c = crc32(page.data() + kHeaderSize, (size_t)size);
    c = crc32(&codec, 1, c);
    c = crc32(&numRows, 4, c);
    c = crc32(&uncompressedSize, 4, c);
    VB_CHECK((int64_t)c == checksum, "PrestoPage: checksum mismatch");
  }
  const int32_t numColumns = in.i32();
  VB_CHECK((size_t)numColumns == type->size(), "PrestoPage: column count does not match the type");
  std::vector<VectorPtr> children;
  for (int32_t c = 0; c < numColumns; ++c) {
    children.push_back(readColumn(in, type->childAt((size_t)c)));
    VB_CHECK(children.back()->size() == (size_t)numRows, "PrestoPage: column length does not match the page");
  }
  VB_CHECK(in.at == page.size(), "PrestoPage: trailing bytes");
  return std::make_shared<RowVector>(type, (size_t)numRows, std::move(children));
}

} // namespace presto
} // namespace serializer
} // namespace vb200
