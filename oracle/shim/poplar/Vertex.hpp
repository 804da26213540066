// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
//
// Minimal stand-in for Graphcore's <poplar/Vertex.hpp> so that the reference's
// device code (/root/reference/src/codelets/codelets.cpp) compiles *unmodified*
// with g++ as a CPU program. Only the surface the codelets touch is provided:
//   Vertex / MultiVertex::numWorkers()   (codelets.cpp:36,48,93,235,312,361)
//   Vector<T, Layout> with size()/operator[]   (codelets.cpp:39-41,96-98,238,314,363-364)
//   Input/Output/InOut<T> for vector payloads and for scalars (`*fov`, bare
//   `imageWidth`)   (codelets.cpp:42-45,68,71-73,99-101,176,178,209,315,334)
// The IPU has 6 hardware worker contexts per tile; MultiVertex codelets stride by it.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_fp16.h>

using half = __half;

namespace poplar {

enum class VectorLayout { SPAN, ONE_PTR };

struct Vertex {};

struct MultiVertex {
  static constexpr unsigned numWorkers() { return 6u; }
};

template <class T, VectorLayout L = VectorLayout::SPAN>
struct Vector {
  T* ptr = nullptr;
  std::size_t count = 0;
  Vector() {}
  Vector(T* p, std::size_t n) : ptr(p), count(n) {}
  std::size_t size() const { return count; }
  T& operator[](std::size_t i) { return ptr[i]; }
  const T& operator[](std::size_t i) const { return ptr[i]; }
};

// Vector payloads: the wrapper *is* the vector.
template <class T> struct Input : T { using T::T; Input() {} Input(const T& t) : T(t) {} };
template <class T> struct Output : T { using T::T; Output() {} Output(const T& t) : T(t) {} };
template <class T> struct InOut : T { using T::T; InOut() {} InOut(const T& t) : T(t) {} };

// Scalar payloads: dereference and implicit conversion both appear in the codelets.
#define B200PT_SHIM_SCALAR(TYPE)                    \
  template <> struct Input<TYPE> {                  \
    TYPE value;                                     \
    Input() : value() {}                            \
    Input(TYPE v) : value(v) {}                     \
    const TYPE& operator*() const { return value; } \
    operator TYPE() const { return value; }         \
  };
B200PT_SHIM_SCALAR(unsigned)
B200PT_SHIM_SCALAR(unsigned short)
B200PT_SHIM_SCALAR(float)
B200PT_SHIM_SCALAR(half)
#undef B200PT_SHIM_SCALAR

} // namespace poplar
