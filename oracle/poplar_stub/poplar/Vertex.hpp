// Host-side stand-in for <poplar/Vertex.hpp>, used ONLY to compile the reference's
// codelets/splat/codelets.cpp (lines 1-642) with g++ for oracle/_ref (test infrastructure).
// It is not product code and contains no splat arithmetic.  Poplar's Vector is a (pointer,size)
// view that the graph builder binds to tensors (src/splat/ipu_rasteriser.cpp:372-383); here the
// oracle driver binds it to std::vector storage.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cmath>

// popc resolves the codelet's unqualified exp/floor/ceil/sqrt/abs/tan on float arguments to
// the float overloads (the IPU has no double unit); glibc's global names are double-only, which
// would not even compile at codelets.cpp:400 (glm::min(float,double)).  See DESIGN.md "oracle".
using std::exp;
using std::floor;
using std::ceil;
using std::sqrt;
using std::abs;
using std::tan;

namespace poplar {

template <typename T> struct Vector {
  T *ptr = nullptr;
  std::size_t count = 0;
  void bind(T *p, std::size_t n) { ptr = p; count = n; }
  std::size_t size() const { return count; }
  T &operator[](std::size_t i) { return ptr[i]; }
  const T &operator[](std::size_t i) const { return ptr[i]; }
};

template <typename V> struct Input : V {};
template <typename V> struct Output : V {};
template <typename V> struct InOut : V {};

struct Vertex {};
struct MultiVertex {
  static constexpr unsigned numWorkers() { return 6; }
};

} // namespace poplar
