// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, imported by or executed from
// the product path (ipu_path_trace_b200/, include/, libb200pt.so). Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// Shared definitions for the CPU restatement of the reference hot path:
//   * IEEE binary16 conversions (the reference stores camera rays, AA noise, uniforms and
//     several scalars as `half`: codelets.cpp:39-45,96-100; PathTracerApp.cpp:293,339-346)
//   * Philox4x32-10 counter RNG and the per-path draw schedule (the IPU's hardware RNG is
//     not reproducible; both this oracle and the CUDA kernels use this stream instead)
//   * record layouts that mirror TraceRecord (src/codelets/TraceRecord.hpp:7-19) and
//     light::Contribution (external/light/src/light.hpp:263-277)
//
// PARITY STATUS: the path-tracing restatement is PINNED against the unmodified reference
// codelets compiled from /root/reference (oracle/_ref, see tests/test_oracle_vs_ref.py and
// tests/golden/). The NIF restatement is UNPINNED: the reference has no CPU inference, no
// test vectors and its weights blob is missing from the snapshot (.MISSING_LARGE_BLOBS:1).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

// ---------------------------------------------------------------------------------
// binary16 <-> binary32, round-to-nearest-even, subnormals and inf/nan handled.
// ---------------------------------------------------------------------------------
inline std::uint32_t f32Bits(float f) { std::uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float bitsF32(std::uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

inline std::uint16_t floatToHalfBits(float f) {
  const std::uint32_t x = f32Bits(f);
  const std::uint32_t sign = (x >> 16) & 0x8000u;
  const std::uint32_t absx = x & 0x7fffffffu;
  if (absx >= 0x7f800000u) {                       // inf / nan
    return (std::uint16_t)(sign | 0x7c00u | (absx > 0x7f800000u ? 0x200u : 0u));
  }
  if (absx >= 0x477ff000u) {                       // >= 65520 rounds to inf
    return (std::uint16_t)(sign | 0x7c00u);
  }
  if (absx < 0x33000001u) {                        // <= 2^-25 rounds to zero
    return (std::uint16_t)sign;
  }
  int exp = (int)(absx >> 23) - 127;
  std::uint32_t man = (absx & 0x7fffffu) | 0x800000u;   // 24-bit significand
  int shift;                                        // bits to drop
  std::uint32_t halfExp;
  if (exp < -14) {                                  // subnormal half
    shift = 13 + (-14 - exp);
    halfExp = 0;
  } else {
    shift = 13;
    halfExp = (std::uint32_t)(exp + 15);
  }
  const std::uint32_t dropped = man & ((1u << shift) - 1u);
  const std::uint32_t halfway = 1u << (shift - 1);
  std::uint32_t kept = man >> shift;
  if (dropped > halfway || (dropped == halfway && (kept & 1u))) kept += 1;
  // For normals `kept` carries the implicit bit (0x400); adding the exponent field lets a
  // mantissa carry ripple into the exponent, which is exactly what IEEE rounding needs.
  std::uint32_t h = (halfExp == 0) ? kept : ((halfExp - 1) << 10) + kept;
  return (std::uint16_t)(sign | h);
}

inline float halfBitsToFloat(std::uint16_t h) {
  const std::uint32_t sign = ((std::uint32_t)h & 0x8000u) << 16;
  const std::uint32_t exp = (h >> 10) & 0x1fu;
  const std::uint32_t man = h & 0x3ffu;
  if (exp == 0) {
    if (man == 0) return bitsF32(sign);
    float v = (float)man * 5.9604644775390625e-8f;  // 2^-24
    return sign ? -v : v;
  }
  if (exp == 31) return bitsF32(sign | 0x7f800000u | (man << 13));
  return bitsF32(sign | ((exp + 112u) << 23) | (man << 13));
}

inline float roundToHalf(float f) { return halfBitsToFloat(floatToHalfBits(f)); }

// ---------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11). key = 64-bit seed, counter = 4 x u32.
// ---------------------------------------------------------------------------------
struct Philox4 { std::uint32_t v[4]; };

inline Philox4 philox4x32_10(std::uint32_t c0, std::uint32_t c1, std::uint32_t c2, std::uint32_t c3,
                             std::uint32_t k0, std::uint32_t k1) {
  for (int round = 0; round < 10; ++round) {
    const std::uint64_t p0 = (std::uint64_t)0xD2511F53u * c0;
    const std::uint64_t p1 = (std::uint64_t)0xCD9E8D57u * c2;
    const std::uint32_t n0 = (std::uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const std::uint32_t n1 = (std::uint32_t)p1;
    const std::uint32_t n2 = (std::uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const std::uint32_t n3 = (std::uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return Philox4{{c0, c1, c2, c3}};
}

// Per-path stream: counter = (pixel id = v*W+u, sample index, block, 0). Keyed on the
// pixel (not the worklist slot) so the image does not depend on how pixels are sharded.
// Block 0 words 0,1 feed the anti-alias noise; every later word is one uniform.
struct PathRng {
  std::uint32_t pixel, sample, k0, k1;
  std::uint32_t block;
  int next;               // next unread word in `cur`
  Philox4 cur;
  PathRng(std::uint64_t seed, std::uint32_t pixelId, std::uint32_t sampleIdx)
      : pixel(pixelId), sample(sampleIdx), k0((std::uint32_t)seed), k1((std::uint32_t)(seed >> 32)), block(0), next(0) {
    cur = philox4x32_10(pixel, sample, 0u, 0u, k0, k1);
  }
  std::uint32_t word() {
    if (next == 4) { block += 1; cur = philox4x32_10(pixel, sample, block, 0u, k0, k1); next = 0; }
    return cur.v[next++];
  }
  // U[0,1) on the 24-bit grid, then rounded to binary16 exactly like the reference's
  // half-precision primary-sample buffer (PathTracerApp.cpp:293-297). Can return 1.0.
  float uniformHalf() { return roundToHalf((float)(word() >> 8) * 5.9604644775390625e-8f); }
};

// ---------------------------------------------------------------------------------
// Reproducible elementary functions for the anti-alias noise: only IEEE +,-,*,/ and sqrt
// in a fixed order (no libm), so CPU (no FMA contraction) and CUDA (-fmad=false) agree
// bit for bit. Accuracy ~1e-7 abs, far below the binary16 rounding applied afterwards.
// ---------------------------------------------------------------------------------
inline float reproLog(float x) {           // x in (0, 1]
  std::uint32_t b = f32Bits(x);
  int e = (int)(b >> 23) - 127;
  std::uint32_t mb = (b & 0x7fffffu) | 0x3f800000u;
  float m = bitsF32(mb);                    // [1,2)
  if (m > 1.41421356f) { m = m * 0.5f; e += 1; }   // [0.707,1.414]
  const float s = (m - 1.f) / (m + 1.f);
  const float s2 = s * s;
  float p = 0.11111111f;                    // 1/9
  p = p * s2 + 0.14285714f;                 // 1/7
  p = p * s2 + 0.2f;
  p = p * s2 + 0.33333333f;
  p = p * s2 + 1.f;
  return (float)e * 0.69314718f + 2.f * s * p;
}

inline void reproSinCosTurn(float t, float& s, float& c) {   // angle = 2*pi*t, t in [0,1)
  const float x = t * 4.f;
  const int k = (int)(x + 0.5f);
  const float a = (x - (float)k) * 1.57079633f;    // [-pi/4, pi/4]
  const float a2 = a * a;
  float sp = 2.75573192e-6f;                 // 1/9!
  sp = sp * a2 - 1.98412698e-4f;             // -1/7!
  sp = sp * a2 + 8.33333333e-3f;             // 1/5!
  sp = sp * a2 - 1.66666667e-1f;             // -1/3!
  sp = sp * a2 + 1.f;
  const float sa = a * sp;
  float cp = -2.75573192e-7f;                // -1/10!
  cp = cp * a2 + 2.48015873e-5f;             // 1/8!
  cp = cp * a2 - 1.38888889e-3f;             // -1/6!
  cp = cp * a2 + 4.16666667e-2f;             // 1/4!
  cp = cp * a2 - 0.5f;
  const float ca = cp * a2 + 1.f;
  switch (k & 3) {
    case 0: s = sa; c = ca; break;
    case 1: s = ca; c = -sa; break;
    case 2: s = -sa; c = -ca; break;
    default: s = -ca; c = sa; break;
  }
}

enum AaNoiseType { AA_NORMAL = 0, AA_UNIFORM = 1, AA_TRUNCATED_NORMAL = 2 };

// Two anti-alias noise values (as binary16-rounded floats) from two Philox words.
// normal: N(0,1) (poprand::normal, PathTracerApp.cpp:34-37); uniform: U(-1,1) (:30-33);
// truncated-normal: N(0,1) conditioned on |x| <= 3 sigma (:38-41, poprand::truncatedNormal, alpha 3), by re-drawing like the
// IPU library: a pair with a component beyond 3 sigma is replaced by the next candidate pair from Philox blocks whose
// counter is (w0, w1, attempt, "trun"), key 0 - the path's own stream does not advance. The IPU generator itself is not
// reproducible, so the words are this repository's choice; product and oracle share it.
inline void boxMullerPair(std::uint32_t w0, std::uint32_t w1, float& a, float& b) {
  const float u1 = (float)((w0 >> 8) + 1u) * 5.9604644775390625e-8f;   // (0,1]
  const float u2 = (float)(w1 >> 8) * 5.9604644775390625e-8f;          // [0,1)
  const float r = std::sqrt(-2.f * reproLog(u1));
  float s, c;
  reproSinCosTurn(u2, s, c);
  a = r * c; b = r * s;
}
inline void aaNoisePair(std::uint32_t w0, std::uint32_t w1, int type, float& n0, float& n1) {
  if (type == AA_UNIFORM) {
    const float u1 = (float)((w0 >> 8) + 1u) * 5.9604644775390625e-8f;   // (0,1]
    const float u2 = (float)(w1 >> 8) * 5.9604644775390625e-8f;          // [0,1)
    n0 = roundToHalf(2.f * u1 - 1.f);
    n1 = roundToHalf(2.f * u2 - 1.f);
    return;
  }
  float a, b;
  boxMullerPair(w0, w1, a, b);
  if (type == AA_TRUNCATED_NORMAL) {
    for (std::uint32_t attempt = 0; std::fabs(a) > 3.f || std::fabs(b) > 3.f; ++attempt) {
      const Philox4 p = philox4x32_10(w0, w1, attempt, 0x7472756eu, 0u, 0u);
      boxMullerPair(p.v[0], p.v[1], a, b);
      if (std::fabs(a) > 3.f || std::fabs(b) > 3.f) boxMullerPair(p.v[2], p.v[3], a, b);
    }
  }
  n0 = roundToHalf(a);
  n1 = roundToHalf(b);
}

// ---------------------------------------------------------------------------------
// Record layouts (byte-compatible with the reference structs)
// ---------------------------------------------------------------------------------
#pragma pack(push, 1)
struct TraceRec {                 // src/codelets/TraceRecord.hpp:7-19 (20 bytes)
  std::uint16_t u, v;
  float r, g, b;
  std::uint16_t sampleCount, pathLength;
};
struct ContribRec {               // external/light/src/light.hpp:263-277 (20 bytes)
  float clr[3];
  float weight;
  std::uint8_t type;
  std::uint8_t pad[3];
};
#pragma pack(pop)
static_assert(sizeof(TraceRec) == 20, "TraceRecord layout");
static_assert(sizeof(ContribRec) == 20, "Contribution layout");

enum ContribType : std::uint8_t { C_DIFFUSE = 0, C_EMIT, C_SPECULAR, C_REFRACT, C_ESCAPED, C_DEBUG, C_END, C_SKIP };

// Render settings as the device sees them (values already rounded through binary16
// where the reference streams them as half: PathTracerApp.cpp:339-346, IpuPathTraceJob.cpp:133-138).
struct Settings {
  std::uint32_t width, height;
  float fov;              // radians (binary16-rounded)
  float aaScale;          // pixels (binary16-rounded)
  float azimuth;          // radians
  float refractiveIndex;  // binary16-rounded
  float stopProb;         // binary16-rounded
  std::uint32_t rouletteDepth;
  std::uint32_t maxPathLength;
  std::uint64_t seed;
  std::int32_t aaNoiseType;
};

// NIF description handed over the oracle's C interface.
struct NifLayer {
  std::int32_t in, out;
  std::int32_t relu;       // activation == "relu" (NifModel.cpp:323); "linear" -> none (:73-76)
  std::int32_t hasBias;
  const std::uint16_t* kernel;   // [in,out] row-major binary16 (Keras Dense layout, Hdf5Model.cpp:71-82)
  const std::uint16_t* bias;     // [out] binary16 or null
};
struct Nif {
  std::int32_t numLayers;
  std::int32_t embeddingDim;     // nif_metadata.txt "embedding_dimension"
  float maxv;                    // encode_params.max
  float mean[3];                 // encode_params.mean with eps already folded in (NifMetaData.cpp:48-53)
  std::int32_t logToneMap;
  const NifLayer* layers;
};

} // namespace orc
