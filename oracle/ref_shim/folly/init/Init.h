// TEST INFRASTRUCTURE — stand-in for folly/init/Init.h (folly is not in this image), only so that the reference's own
// stand-alone GPU hash-table benchmark (velox/experimental/gpu/tests/HashTableTest.cu) compiles where it lies.
#pragma once
namespace folly {
inline void init(int*, char***) {}
} // namespace folly
