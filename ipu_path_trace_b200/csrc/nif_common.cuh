// NIF pieces shared by the SIMT and tcgen05 kernels: the Fourier-feature encoding and the
// output decoding, following the reference's graph construction
//   encode  src/neural_networks/NifModel.cpp:185-218 (host twin :404-432, coefficients :466-472)
//   decode  src/neural_networks/NifModel.cpp:221-245 (eps folded into mean: NifMetaData.cpp:48-53)
#pragma once

#include "kernels.h"

namespace b2 {

// p = (uv - 1) * 2 ; p * 2^i is cast to binary16 BEFORE sin/cos are taken (NifModel.cpp:211-213),
// so for i >= 9 the argument grid is coarser than the period: the trained network depends on
// exactly this rounding. sin/cos are evaluated accurately (Cody-Waite sinf/cosf, not __sinf)
// on the binary16 argument and the result is rounded to binary16.
// Feature order (:216, host twin :427-430): [sin(u c_0..c_E-1), sin(v c_..), cos(u c_..), cos(v c_..)].
__device__ __forceinline__ void nifFeaturePair(float p, float coeff, __half& s, __half& c) {
  const float x = __half2float(__float2half_rn(p * coeff));
  float sv, cv;
  sincosf(x, &sv, &cv);
  s = __float2half_rn(sv);
  c = __float2half_rn(cv);
}

// Writes 4E features through `put(k, value)`.
template <class Put>
__device__ __forceinline__ void nifEncode(float u, float v, int E, Put&& put) {
  const float pu = (u - 1.f) * 2.f;
  const float pv = (v - 1.f) * 2.f;
  float coeff = 1.f;
  for (int i = 0; i < E; ++i) {
    __half su, cu, sv, cv;
    nifFeaturePair(pu, coeff, su, cu);
    nifFeaturePair(pv, coeff, sv, cv);
    put(i, su);
    put(E + i, sv);
    put(2 * E + i, cu);
    put(3 * E + i, cv);
    coeff *= 2.f;
  }
}

// NifModel.cpp:226-242: fp32 x * max + mean, optional exp.
__device__ __forceinline__ float nifDecode(float x, float maxv, float mean, int logToneMap) {
  const float o = x * maxv + mean;
  return logToneMap ? expf(o) : o;
}

// Result handling common to both kernels: either store the network output (B,G,R) or
// apply PostProcessEscapedRays' BGR->RGB swap (codelets.cpp:375-379) and add
// throughput (*) radiance to the work item (AccumulateContributions, :255-298).
__device__ __forceinline__ void nifEmit(const NifIo& io, uint32_t row, float b, float g, float r, float4 thrSlot) {
  if (io.bgrOut) {
    io.bgrOut[3 * (size_t)row] = b;
    io.bgrOut[3 * (size_t)row + 1] = g;
    io.bgrOut[3 * (size_t)row + 2] = r;
  }
  if (io.accum) {
    accumAdd(io.accum, __float_as_uint(thrSlot.w), r * thrSlot.x, g * thrSlot.y, b * thrSlot.z);
  }
}

} // namespace b2
