// Host-side mirror of the slice of Velox's type / vector model that crosses the operator boundary
// of the FilterProject -> HashAggregation path.  Same names and meaning as the reference so that
// plans and tests read like the reference's own:
//   TypeKind, Type, RowType            <- velox/type/Type.h:62-84, 480-560, 820-900
//   BaseVector / FlatVector / DictionaryVector / ConstantVector / RowVector
//                                      <- velox/vector/BaseVector.h, FlatVector.h,
//                                         DictionaryVector.h, ConstantVector.h, ComplexVector.h
//   nulls bitmaps                      <- velox/common/base/Nulls.h:26-35 (bit = 1: NOT null)
// This is interface plumbing, not a port: buffers are plain std::vector storage, there is no memory
// pool, no lazy vectors, no string views.  A maintainer binding the real Velox types replaces this
// header with the reference's headers.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace vb200 {

// ---- errors (velox/common/base/Exceptions.h) -----------------------------------------------------
struct VeloxException : std::runtime_error { using std::runtime_error::runtime_error; };
struct VeloxUserError : VeloxException { using VeloxException::VeloxException; };       // bad input / data
struct VeloxRuntimeError : VeloxException { using VeloxException::VeloxException; };    // internal / unsupported
#define VB_USER_CHECK(cond, msg) do { if (!(cond)) throw ::vb200::VeloxUserError(std::string(msg)); } while (0)
#define VB_CHECK(cond, msg) do { if (!(cond)) throw ::vb200::VeloxRuntimeError(std::string(msg)); } while (0)

// ---- types ---------------------------------------------------------------------------------------
enum class TypeKind : int8_t { BOOLEAN = 0, TINYINT = 1, SMALLINT = 2, INTEGER = 3, BIGINT = 4, REAL = 5, DOUBLE = 6, VARCHAR = 7, ROW = 32 };

class Type;
using TypePtr = std::shared_ptr<const Type>;

class Type {
 public:
  Type(TypeKind kind, bool date = false) : kind_(kind), date_(date) {}
  virtual ~Type() = default;
  TypeKind kind() const { return kind_; }
  bool isDate() const { return date_; } // DATE is INTEGER days since epoch (velox/type/Type.h DateType)
  bool isInteger() const { return kind_ >= TypeKind::TINYINT && kind_ <= TypeKind::BIGINT; }
  bool isFloating() const { return kind_ == TypeKind::REAL || kind_ == TypeKind::DOUBLE; }
  virtual std::string toString() const {
    static const char* n[] = {"BOOLEAN", "TINYINT", "SMALLINT", "INTEGER", "BIGINT", "REAL", "DOUBLE", "VARCHAR"};
    return date_ ? "DATE" : n[(int)kind_];
  }
  bool equals(const Type& o) const { return toString() == o.toString(); }
 private:
  TypeKind kind_;
  bool date_;
};

class RowType : public Type {
 public:
  RowType(std::vector<std::string> names, std::vector<TypePtr> types) : Type(TypeKind::ROW), names_(std::move(names)), types_(std::move(types)) {}
  size_t size() const { return names_.size(); }
  const std::string& nameOf(size_t i) const { return names_[i]; }
  const TypePtr& childAt(size_t i) const { return types_[i]; }
  const std::vector<std::string>& names() const { return names_; }
  const std::vector<TypePtr>& children() const { return types_; }
  int getChildIdxIfExists(const std::string& name) const {
    for (size_t i = 0; i < names_.size(); ++i) if (names_[i] == name) return (int)i;
    return -1;
  }
  int getChildIdx(const std::string& name) const {
    const int i = getChildIdxIfExists(name);
    VB_USER_CHECK(i >= 0, "Field not found: " + name);
    return i;
  }
  std::string toString() const override {
    std::string s = "ROW<";
    for (size_t i = 0; i < names_.size(); ++i) s += (i ? "," : "") + names_[i] + ":" + types_[i]->toString();
    return s + ">";
  }
 private:
  std::vector<std::string> names_;
  std::vector<TypePtr> types_;
};
using RowTypePtr = std::shared_ptr<const RowType>;

inline TypePtr BOOLEAN() { static TypePtr t = std::make_shared<Type>(TypeKind::BOOLEAN); return t; }
inline TypePtr TINYINT() { static TypePtr t = std::make_shared<Type>(TypeKind::TINYINT); return t; }
inline TypePtr SMALLINT() { static TypePtr t = std::make_shared<Type>(TypeKind::SMALLINT); return t; }
inline TypePtr INTEGER() { static TypePtr t = std::make_shared<Type>(TypeKind::INTEGER); return t; }
inline TypePtr BIGINT() { static TypePtr t = std::make_shared<Type>(TypeKind::BIGINT); return t; }
inline TypePtr REAL() { static TypePtr t = std::make_shared<Type>(TypeKind::REAL); return t; }
inline TypePtr DOUBLE() { static TypePtr t = std::make_shared<Type>(TypeKind::DOUBLE); return t; }
inline TypePtr VARCHAR() { static TypePtr t = std::make_shared<Type>(TypeKind::VARCHAR); return t; }
inline TypePtr DATE() { static TypePtr t = std::make_shared<Type>(TypeKind::INTEGER, true); return t; }
inline RowTypePtr ROW(std::vector<std::string> names, std::vector<TypePtr> types) {
  return std::make_shared<RowType>(std::move(names), std::move(types));
}

// ---- host memory of flat buffers (velox/common/memory/MemoryPool.h analogue) ---------------------------------
// Vectors allocate their value arrays through a process-wide hook.  gpu::usePinnedHostMemory(true) points it at
// page-locked memory, which lets the GPU operators copy a batch with one asynchronous DMA straight out of the vector
// (what a Velox deployment gets by giving the scan's MemoryPool a pinned allocator); the default is malloc.
namespace memory {
struct HostAllocator { void* (*alloc)(size_t bytes); void (*dealloc)(void* p); };
HostAllocator& hostAllocator();
template <typename T>
struct PoolAllocator {
  using value_type = T;
  HostAllocator backing;
  PoolAllocator() : backing(hostAllocator()) {}
  template <typename U> PoolAllocator(const PoolAllocator<U>& o) : backing(o.backing) {}
  T* allocate(size_t n) {
    void* p = backing.alloc(n * sizeof(T));
    if (!p) throw std::bad_alloc();
    return static_cast<T*>(p);
  }
  void deallocate(T* p, size_t) { backing.dealloc(p); }
  template <typename U> bool operator==(const PoolAllocator<U>& o) const { return backing.alloc == o.backing.alloc; }
  template <typename U> bool operator!=(const PoolAllocator<U>& o) const { return !(*this == o); }
};
} // namespace memory
template <typename T>
using Buffer = std::vector<T, memory::PoolAllocator<T>>;

// ---- null bitmaps ----------------------------------------------------------------------------------
using Nulls = std::vector<uint64_t>; // empty = no nulls; bit i = 1 -> row i is NOT null
namespace bits {
inline size_t nwords(size_t n) { return (n + 63) / 64; }
inline bool isBitSet(const uint64_t* w, size_t i) { return (w[i >> 6] >> (i & 63)) & 1ull; }
inline void setBit(uint64_t* w, size_t i, bool v) { if (v) w[i >> 6] |= 1ull << (i & 63); else w[i >> 6] &= ~(1ull << (i & 63)); }
} // namespace bits

// ---- vectors ---------------------------------------------------------------------------------------
enum class VectorEncoding { FLAT, DICTIONARY, CONSTANT, ROW };

class BaseVector;
using VectorPtr = std::shared_ptr<BaseVector>;

class BaseVector {
 public:
  BaseVector(TypePtr type, VectorEncoding enc, size_t size) : type_(std::move(type)), encoding_(enc), size_(size) {}
  virtual ~BaseVector() = default;
  const TypePtr& type() const { return type_; }
  VectorEncoding encoding() const { return encoding_; }
  size_t size() const { return size_; }
  // nulls of THIS layer (a dictionary's extra nulls; a flat vector's nulls)
  const Nulls& nulls() const { return nulls_; }
  Nulls& mutableNulls() { if (nulls_.empty()) nulls_.assign(bits::nwords(size_) ? bits::nwords(size_) : 1, ~0ull); return nulls_; }
  void setNull(size_t i, bool isNull) { bits::setBit(mutableNulls().data(), i, !isNull); }
  bool mayHaveNulls() const { return !nulls_.empty(); }
 protected:
  TypePtr type_;
  VectorEncoding encoding_;
  size_t size_;
  Nulls nulls_;
};

template <typename T>
class FlatVector : public BaseVector {
 public:
  FlatVector(TypePtr type, size_t size) : BaseVector(std::move(type), VectorEncoding::FLAT, size), values_(size) {}
  FlatVector(TypePtr type, const std::vector<T>& values) : BaseVector(std::move(type), VectorEncoding::FLAT, values.size()), values_(values.begin(), values.end()) {}
  const T* rawValues() const { return values_.data(); }
  T* mutableRawValues() { return values_.data(); }
  const T& valueAt(size_t i) const { return values_[i]; }
  void set(size_t i, T v) { values_[i] = std::move(v); }
 private:
  Buffer<T> values_;
};

// velox/type/StringView.h: 16 bytes — the size, the first 4 bytes, then the rest of a string of at most 12 bytes inline or
// a pointer to the whole string.  The bytes after a short string are zero.
struct StringView {
  static constexpr uint32_t kInlineSize = 12;
  StringView() { memset(this, 0, sizeof(*this)); }
  StringView(const char* data, size_t size) {
    memset(this, 0, sizeof(*this));
    size_ = (uint32_t)size;
    if (size <= kInlineSize) { if (size) memcpy(prefix_, data, size); } // prefix_ and value_.inlined are contiguous
    else { memcpy(prefix_, data, 4); value_.data = data; }
  }
  uint32_t size() const { return size_; }
  bool isInline() const { return size_ <= kInlineSize; }
  const char* data() const { return isInline() ? prefix_ : value_.data; }
  std::string str() const { return std::string(data(), size_); }
 private:
  uint32_t size_;
  char prefix_[4];
  union { char inlined[8]; const char* data; } value_;
};
static_assert(sizeof(StringView) == 16, "StringView layout");

// FlatVector<StringView> of the reference: the views plus the buffers that own the bytes of the long strings.  This is
// the form a scan hands VARCHAR columns over in; FlatVector<std::string> is the test-friendly form of the same thing.
class StringViewVector : public BaseVector {
 public:
  explicit StringViewVector(const std::vector<std::string>& strings) : BaseVector(VARCHAR(), VectorEncoding::FLAT, strings.size()), values_(strings.size()) {
    for (size_t i = 0; i < strings.size(); ++i) set(i, strings[i]);
  }
  StringViewVector(TypePtr type, size_t size) : BaseVector(std::move(type), VectorEncoding::FLAT, size), values_(size) {}
  const StringView* rawValues() const { return values_.data(); }
  const StringView& valueAt(size_t i) const { return values_[i]; }
  void set(size_t i, const std::string& s) {
    if (s.size() <= StringView::kInlineSize) { values_[i] = StringView(s.data(), s.size()); return; }
    stringBuffers_.push_back(std::make_shared<std::string>(s));
    values_[i] = StringView(stringBuffers_.back()->data(), s.size());
  }
 private:
  Buffer<StringView> values_;
  std::vector<std::shared_ptr<std::string>> stringBuffers_;
};

// BOOLEAN flat vectors are bit-packed like the reference's FlatVector<bool>.
class BoolVector : public BaseVector {
 public:
  BoolVector(size_t size) : BaseVector(BOOLEAN(), VectorEncoding::FLAT, size), bits_(bits::nwords(size) ? bits::nwords(size) : 1, 0) {}
  const uint64_t* rawValues() const { return bits_.data(); }
  uint64_t* mutableRawValues() { return bits_.data(); }
  bool valueAt(size_t i) const { return bits::isBitSet(bits_.data(), i); }
  void set(size_t i, bool v) { bits::setBit(bits_.data(), i, v); }
 private:
  std::vector<uint64_t> bits_;
};

// wrapInfo = int32 indices into `base`; optional extra nulls on this layer (per OUTER row).
class DictionaryVector : public BaseVector {
 public:
  DictionaryVector(const std::vector<int32_t>& indices, VectorPtr base, Nulls extraNulls = {})
      : BaseVector(base->type(), VectorEncoding::DICTIONARY, indices.size()), indices_(indices.begin(), indices.end()), base_(std::move(base)) {
    nulls_ = std::move(extraNulls);
  }
  const int32_t* rawIndices() const { return indices_.data(); }
  const VectorPtr& valueVector() const { return base_; }
 private:
  Buffer<int32_t> indices_;
  VectorPtr base_;
};

// One value (or NULL) repeated `size` times; the value is element 0 of a flat vector of size 1.
class ConstantVector : public BaseVector {
 public:
  ConstantVector(VectorPtr oneValue, size_t size, bool isNull = false)
      : BaseVector(oneValue->type(), VectorEncoding::CONSTANT, size), value_(std::move(oneValue)), isNull_(isNull) {}
  const VectorPtr& valueVector() const { return value_; }
  bool isNullConstant() const { return isNull_; }
 private:
  VectorPtr value_;
  bool isNull_;
};

class RowVector : public BaseVector {
 public:
  RowVector(RowTypePtr type, size_t size, std::vector<VectorPtr> children)
      : BaseVector(type, VectorEncoding::ROW, size), rowType_(std::move(type)), children_(std::move(children)) {
    VB_CHECK(children_.size() == rowType_->size(), "RowVector: children do not match the row type");
    for (auto& c : children_) VB_CHECK(c == nullptr || c->size() == size_, "RowVector: child size differs");
  }
  const RowTypePtr& rowType() const { return rowType_; }
  const VectorPtr& childAt(size_t i) const { return children_[i]; }
  const std::vector<VectorPtr>& children() const { return children_; }
 private:
  RowTypePtr rowType_;
  std::vector<VectorPtr> children_;
};
using RowVectorPtr = std::shared_ptr<RowVector>;

// Value i of a FLAT VARCHAR vector of either form.
inline std::string stringValueAt(const BaseVector* flat, size_t i) {
  if (auto sv = dynamic_cast<const StringViewVector*>(flat)) return sv->valueAt(i).str();
  return static_cast<const FlatVector<std::string>*>(flat)->valueAt(i);
}

// ---- small builders (velox/vector/tests/utils/VectorMaker.h analogue) ---------------------------
template <typename T>
inline std::shared_ptr<FlatVector<T>> makeFlatVector(TypePtr type, std::vector<T> values) {
  return std::make_shared<FlatVector<T>>(std::move(type), std::move(values));
}
inline VectorPtr wrapInDictionary(std::vector<int32_t> indices, VectorPtr base, Nulls extraNulls = {}) {
  return std::make_shared<DictionaryVector>(std::move(indices), std::move(base), std::move(extraNulls));
}
inline RowVectorPtr makeRowVector(std::vector<std::string> names, std::vector<VectorPtr> children) {
  std::vector<TypePtr> types;
  for (auto& c : children) types.push_back(c->type());
  const size_t n = children.empty() ? 0 : children[0]->size();
  return std::make_shared<RowVector>(ROW(std::move(names), std::move(types)), n, std::move(children));
}

} // namespace vb200
