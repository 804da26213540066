// Shared device/host definitions for libb200pt: binary16 rounding, the Philox4x32-10 per-path
// stream, anti-alias noise, record layouts and the device-side settings block.
//
// The reference draws its randomness from the IPU hardware RNG (poprand, PathTracerApp.cpp:29-45,
// 267-299), which cannot be reproduced; this library uses a counter-based stream keyed on
// (seed, pixel, sample) so an image does not depend on how pixels are sharded over GPUs.
// The draw schedule (2 words of anti-alias noise, then one word per primary sample, every
// value rounded to binary16 like the reference's half buffers) is the contract the parity tests
// check word for word.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

// ------------------------------------------------------------------------------------------
// Records
// ------------------------------------------------------------------------------------------
#pragma pack(push, 1)
struct TraceRecord {            // src/codelets/TraceRecord.hpp:7-19
  uint16_t u, v;
  float r, g, b;
  uint16_t sampleCount, pathLength;
};
#pragma pack(pop)
static_assert(sizeof(TraceRecord) == 20, "TraceRecord must stay 20 bytes");

// Work-item radiance accumulator. The reference folds a pixel's samples in a fixed order (codelets.cpp:235-306), so its
// image is a function of the seed. Here samples of one pixel finish on different SMs in no particular order; added as
// fp32 the sum would depend on that order (run to run, and on how many GPUs share the image). Integer addition commutes:
// every contribution is converted once to signed 64-bit fixed point with 32 fraction bits (2^-32 = 2.3e-10 absolute
// resolution; 2^31 = 2.1e9 of headroom for the sum of a step) and added with a 64-bit integer atomic, and the sum is
// converted back to fp32 once, when the records are packed or the film is accumulated. Same bits on every run and every
// sharding.
struct __align__(32) AccumSlot { long long r, g, b, unused; };
constexpr float kAccumScale = 4294967296.f;            // 2^32
constexpr float kAccumInvScale = 2.3283064365386963e-10f;
__device__ __forceinline__ long long accumToFixed(float x) { return __float2ll_rn(x * kAccumScale); }
__device__ __forceinline__ float accumToFloat(long long v) { return __ll2float_rn(v) * kAccumInvScale; }     // one rounding: the scale is a power of two
__device__ __forceinline__ void accumAdd(AccumSlot* __restrict__ accum, uint32_t slot, float r, float g, float b) {
  unsigned long long* a = reinterpret_cast<unsigned long long*>(accum + slot);
  atomicAdd(a, (unsigned long long)accumToFixed(r));
  atomicAdd(a + 1, (unsigned long long)accumToFixed(g));
  atomicAdd(a + 2, (unsigned long long)accumToFixed(b));
}

// Render settings as the kernels see them. Quantities the reference streams as `half`
// (PathTracerApp.cpp:339-346, IpuPathTraceJob.cpp:133-138) are stored already rounded.
struct DeviceSettings {
  uint32_t width, height;
  float aaScale;            // pixels
  float tanHalfFov;         // tanf(fov/2), fov binary16-rounded radians (light.hpp:199)
  float aspect;             // width / height (light.hpp:198)
  float azimuth;            // radians (codelets.cpp:334)
  float refractiveIndex;
  float stopProb;
  float rouletteWeight;     // 1 / (1 - stopProb) (light.hpp:28)
  uint32_t rouletteDepth;
  uint32_t maxPathLength;
  uint32_t seedLo, seedHi;
  int32_t aaNoiseType;
};

// Primitive table, SoA; built on the host from the built-in scene (5 spheres + 1 disc, codelets.cpp:110-144) or from
// b200pt_set_scene (the reference anticipates the scene as data, codelets.cpp:90-92).
constexpr int kMaxPrims = 16;
struct SceneSoA {
  int32_t count;
  int32_t shape[kMaxPrims];     // 0 sphere, 1 disc
  int32_t bsdf[kMaxPrims];      // 0 diffuse, 1 specular, 2 refractive
  float cx[kMaxPrims], cy[kMaxPrims], cz[kMaxPrims];
  float nx[kMaxPrims], ny[kMaxPrims], nz[kMaxPrims];   // disc normal (unit)
  float radius2[kMaxPrims];
  float planeD[kMaxPrims];      // |centre . n| (light.hpp:101)
  float colr[kMaxPrims], colg[kMaxPrims], colb[kMaxPrims];
  int32_t emissive[kMaxPrims];  // light.hpp Material::emissive
  float emr[kMaxPrims], emg[kMaxPrims], emb[kMaxPrims];
};

// ------------------------------------------------------------------------------------------
// binary16 rounding
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float roundToHalf(float f) { return __half2float(__float2half_rn(f)); }
__device__ __forceinline__ uint16_t halfBits(float f) { return __half_as_ushort(__float2half_rn(f)); }
__device__ __forceinline__ float halfBitsToFloat(uint16_t b) { return __half2float(__ushort_as_half(b)); }

// ------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11)
// ------------------------------------------------------------------------------------------
struct Philox4 { uint32_t v[4]; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int round = 0; round < 10; ++round) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0;
    const uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  Philox4 r; r.v[0] = c0; r.v[1] = c1; r.v[2] = c2; r.v[3] = c3;
  return r;
}

// counter = (pixel id = v*W+u, sample index, block, 0); word 0,1 of block 0 -> AA noise.
struct PathRng {
  uint32_t pixel, sample, k0, k1, block;
  int next;
  uint32_t w0, w1, w2, w3;
  __device__ __forceinline__ PathRng(uint32_t seedLo, uint32_t seedHi, uint32_t pixelId, uint32_t sampleIdx)
      : pixel(pixelId), sample(sampleIdx), k0(seedLo), k1(seedHi), block(0), next(0) { refill(); }
  // Resume a stream inside block 0 without recomputing it: the last two words of the block (the first two fed the
  // anti-alias noise) travel with a live ray between the two phases of the wavefront kernel.
  struct Resume {};
  __device__ __forceinline__ PathRng(Resume, uint32_t seedLo, uint32_t seedHi, uint32_t pixelId, uint32_t sampleIdx, uint32_t word2, uint32_t word3)
      : pixel(pixelId), sample(sampleIdx), k0(seedLo), k1(seedHi), block(0), next(2), w0(0), w1(0), w2(word2), w3(word3) {}
  __device__ __forceinline__ void refill() {
    const Philox4 p = philox4x32_10(pixel, sample, block, 0u, k0, k1);
    w0 = p.v[0]; w1 = p.v[1]; w2 = p.v[2]; w3 = p.v[3];
  }
  __device__ __forceinline__ uint32_t word() {
    if (next == 4) { block += 1; refill(); next = 0; }
    const uint32_t r = next == 0 ? w0 : (next == 1 ? w1 : (next == 2 ? w2 : w3));
    next += 1;
    return r;
  }
  // U[0,1) on the 24-bit grid rounded to binary16 (can return 1.0), like the reference's
  // half-precision primary-sample buffer (PathTracerApp.cpp:293-297).
  __device__ __forceinline__ float uniformHalf() { return roundToHalf((float)(word() >> 8) * 5.9604644775390625e-8f); }
};

// ------------------------------------------------------------------------------------------
// Anti-alias noise. Only IEEE + - * / sqrt in a fixed order (this file is compiled with
// -fmad=false), so the CPU oracle reproduces it bit for bit.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float reproLog(float x) {            // x in (0,1]
  const uint32_t b = __float_as_uint(x);
  int e = (int)(b >> 23) - 127;
  float m = __uint_as_float((b & 0x7fffffu) | 0x3f800000u);     // [1,2)
  if (m > 1.41421356f) { m = m * 0.5f; e += 1; }
  const float s = (m - 1.f) / (m + 1.f);
  const float s2 = s * s;
  float p = 0.11111111f;
  p = p * s2 + 0.14285714f;
  p = p * s2 + 0.2f;
  p = p * s2 + 0.33333333f;
  p = p * s2 + 1.f;
  return (float)e * 0.69314718f + 2.f * s * p;
}

__device__ __forceinline__ void reproSinCosTurn(float t, float& s, float& c) {   // angle 2*pi*t
  const float x = t * 4.f;
  const int k = (int)(x + 0.5f);
  const float a = (x - (float)k) * 1.57079633f;
  const float a2 = a * a;
  float sp = 2.75573192e-6f;
  sp = sp * a2 - 1.98412698e-4f;
  sp = sp * a2 + 8.33333333e-3f;
  sp = sp * a2 - 1.66666667e-1f;
  sp = sp * a2 + 1.f;
  const float sa = a * sp;
  float cp = -2.75573192e-7f;
  cp = cp * a2 + 2.48015873e-5f;
  cp = cp * a2 - 1.38888889e-3f;
  cp = cp * a2 + 4.16666667e-2f;
  cp = cp * a2 - 0.5f;
  const float ca = cp * a2 + 1.f;
  switch (k & 3) {
    case 0: s = sa; c = ca; break;
    case 1: s = ca; c = -sa; break;
    case 2: s = -sa; c = -ca; break;
    default: s = -ca; c = sa; break;
  }
}

// normal: N(0,1) (poprand::normal, PathTracerApp.cpp:34-37); uniform: U(-1,1) (:30-33);
// truncated-normal: N(0,1) conditioned on |x| <= 3 sigma (:38-41, poprand::truncatedNormal with alpha 3) by re-drawing: a pair
// with a component beyond 3 sigma (0.54 % of the pairs) is rejected and replaced by the next candidate pair, taken from
// Philox blocks keyed by the rejected words themselves (counter = (w0, w1, attempt, "trun")), so the path's own stream, and
// with it every later draw of the path, stays where it is.
__device__ __forceinline__ void boxMullerPair(uint32_t w0, uint32_t w1, float& a, float& b) {
  const float u1 = (float)((w0 >> 8) + 1u) * 5.9604644775390625e-8f;   // (0,1]
  const float u2 = (float)(w1 >> 8) * 5.9604644775390625e-8f;          // [0,1)
  const float r = sqrtf(-2.f * reproLog(u1));
  float s, c;
  reproSinCosTurn(u2, s, c);
  a = r * c; b = r * s;
}
__device__ __forceinline__ void aaNoisePair(uint32_t w0, uint32_t w1, int type, float& n0, float& n1) {
  if (type == 1) {
    const float u1 = (float)((w0 >> 8) + 1u) * 5.9604644775390625e-8f;   // (0,1]
    const float u2 = (float)(w1 >> 8) * 5.9604644775390625e-8f;          // [0,1)
    n0 = roundToHalf(2.f * u1 - 1.f);
    n1 = roundToHalf(2.f * u2 - 1.f);
    return;
  }
  float a, b;
  boxMullerPair(w0, w1, a, b);
  if (type == 2) {
    for (uint32_t attempt = 0; fabsf(a) > 3.f || fabsf(b) > 3.f; ++attempt) {
      const Philox4 p = philox4x32_10(w0, w1, attempt, 0x7472756eu, 0u, 0u);
      boxMullerPair(p.v[0], p.v[1], a, b);
      if (fabsf(a) > 3.f || fabsf(b) > 3.f) boxMullerPair(p.v[2], p.v[3], a, b);
    }
  }
  n0 = roundToHalf(a);
  n1 = roundToHalf(b);
}

} // namespace b2
