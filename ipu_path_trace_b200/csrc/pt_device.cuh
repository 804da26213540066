// Device functions of the per-pixel Monte-Carlo loop. Each one states which reference
// function it stands in for; none is a translation of the Poplar vertex code: the path is
// carried forward as an RGB throughput in registers instead of the reference's 200-byte
// per-path Contribution stack (light.hpp:263-277, codelets.cpp:235-306), and the scene is a
// flat SoA primitive table in shared memory instead of virtual Primitive objects.
//
// Float operations follow the reference's order so that, with -fmad=false, every value
// up to the libm calls (sinf/cosf/acosf/atan2f) is bit-identical to the CPU oracle.
#pragma once

#include "pt_common.cuh"

namespace b2 {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 scale(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float norm2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
// vector.hpp:47 normalises as v * (1/sqrt(|v|^2))
__device__ __forceinline__ V3 unit(V3 a) { return scale(a, 1.f / sqrtf(norm2(a))); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

constexpr float kPi = 3.14159265358979323846264338327950288f;      // light.hpp:16
constexpr float kEps = 1.1920928955078125e-7f;                       // FLT_EPSILON, light.cpp:5
constexpr float kHitEps = 1e-5f;                                     // light.cpp:9

// light.hpp:131-147  Sphere::intersect (geometric form)
__device__ __forceinline__ float hitSphere(V3 centre, float radius2, V3 o, V3 d) {
  const V3 f = sub(centre, o);
  const float tca = dot(f, d);
  if (tca < 0.f) return 0.f;
  const V3 l = sub(centre, add(o, scale(d, tca)));
  const float l2 = norm2(l);
  if (l2 > radius2) return 0.f;
  const float thc = sqrtf(radius2 - l2);
  float t0 = tca - thc, t1 = tca + thc;
  if (t0 > t1) { const float tmp = t0; t0 = t1; t1 = tmp; }
  if (t0 < 0.f) { t0 = t1; if (t0 < 0.f) return 0.f; }
  return t0;
}

// light.hpp:106-120  Disc::intersect
__device__ __forceinline__ float hitDisc(V3 normal, V3 centre, float radius2, float planeD, V3 o, V3 d) {
  const float angle = dot(normal, d);
  if (angle != 0.f) {
    const float t = -(dot(normal, o) + planeD) / angle;
    if (t > kEps) {
      const V3 hp = add(o, scale(d, t));
      if (norm2(sub(hp, centre)) < radius2) return t;
    }
  }
  return 0.f;
}

// light.hpp:173-191  Scene::intersect: closest hit with t > intersectionEpsilon; the origin is
// advanced to the hit point and the normal evaluated there (Sphere::normal :149-151).
__device__ __forceinline__ int closestHit(const SceneSoA& sc, V3& o, V3 d, V3& n) {
  int best = -1;
  float bestT = __int_as_float(0x7f800000);
  const int count = sc.count;
  for (int i = 0; i < count; ++i) {
    const V3 c = v3(sc.cx[i], sc.cy[i], sc.cz[i]);
    const float t = sc.shape[i] == 0 ? hitSphere(c, sc.radius2[i], o, d)
                                     : hitDisc(v3(sc.nx[i], sc.ny[i], sc.nz[i]), c, sc.radius2[i], sc.planeD[i], o, d);
    if (t > kHitEps && t < bestT) { bestT = t; best = i; }
  }
  if (best >= 0) {
    o = add(o, scale(d, bestT));
    n = sc.shape[best] == 0 ? unit(sub(o, v3(sc.cx[best], sc.cy[best], sc.cz[best])))
                            : v3(sc.nx[best], sc.ny[best], sc.nz[best]);
  }
  return best;
}

// light.hpp:230-250  orthonormalSystem (first two axes; the third is the normal)
__device__ __forceinline__ void tangentFrame(V3 n, V3& ax, V3& ay) {
  const float nx2 = n.x * n.x, ny2 = n.y * n.y, nz2 = n.z * n.z;
  if (fabsf(n.x) > fabsf(n.y)) {
    const float inv = 1.f / sqrtf(nx2 + nz2);
    ax = v3(-n.z * inv, 0.f, n.x * inv);
  } else {
    const float inv = 1.f / sqrtf(ny2 + nz2);
    ax = v3(0.f, n.z * inv, -n.y * inv);
  }
  ay = cross(n, ax);
}

// [host-checkable: tests/test_libm_restatement.py compiles this block with g++ and compares it with the host's libm]
// sinf/cosf as the reference's CPU build computes them. glibc (>= 2.28) evaluates both in double: quadrant n =
// round(x * 2/pi) taken from a 2^24-prescaled product, r = x - n*pi/2, then an odd (sin) or even (cos) polynomial of r
// chosen by n, rounded to float once. The result is NOT the correctly rounded sine (it differs from (float)sin((double)x)
// for 1.3 % of the arguments in [0, 2 pi]), and the device's own sinf is 1-2 ulp, so neither reproduces the reference;
// this does, for every float in [0, 2 pi] (checked exhaustively on the host against glibc 2.39). Valid for |x| < 120.
__device__ __forceinline__ float glibcSinCosPoly(double x, double x2, bool flip, int n) {
  // the second table of the routine is the first with the cosine coefficients negated
  const double sg = flip ? -1.0 : 1.0;
  if ((n & 1) == 0) {
    const double x3 = x * x2;
    const double s1 = 0x1.1107605230bc4p-7 + x2 * -0x1.994eb3774cf24p-13;
    const double x7 = x3 * x2;
    const double s = x + x3 * -0x1.555545995a603p-3;
    return (float)(s + x7 * s1);
  }
  const double x4 = x2 * x2;
  const double c2 = sg * -0x1.6c087e89a359dp-10 + x2 * (sg * 0x1.99343027bf8c3p-16);
  const double c1 = sg * 0x1p0 + x2 * (sg * -0x1.ffffffd0c621cp-2);
  const double x6 = x4 * x2;
  const double c = c1 + x4 * (sg * 0x1.55553e1068f19p-5);
  return (float)(c + x6 * c2);
}
__device__ __forceinline__ void glibcSinCosf(float y, float& sinOut, float& cosOut) {
  const double x = (double)y;
  const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ffu;
  if (top < ((0x3f490fdbu >> 20) & 0x7ffu)) {                 // |y| < pi/4 by exponent+top mantissa bits, as glibc tests it
    const double x2 = x * x;
    if (top < ((0x39800000u >> 20) & 0x7ffu)) { sinOut = y; cosOut = 1.f; return; }      // |y| < 2^-12
    sinOut = glibcSinCosPoly(x, x2, false, 0);
    cosOut = glibcSinCosPoly(x, x2, false, 1);
    return;
  }
  const double r = x * 0x1.45F306DC9C883p+23;                  // 2/pi * 2^24
  const int n = (__double2int_rz(r) + 0x800000) >> 24;
  const double xr = x - (double)n * 0x1.921FB54442D18p0;
  const double x2 = xr * xr;
  const int ns = n & 3, nc = (n + 1) & 3;
  const double ss = (ns == 1 || ns == 2) ? -1.0 : 1.0;         // sign table {1, -1, -1, 1}
  const double sc = (nc == 1 || nc == 2) ? -1.0 : 1.0;
  sinOut = glibcSinCosPoly(xr * ss, x2, (ns & 2) != 0, n);
  cosOut = glibcSinCosPoly(xr * sc, x2, (nc & 2) != 0, n ^ 1);
}

// [end host-checkable]

// sdf.hpp:8-31 diffuse() with hemisphere() light.hpp:223-228. Returns the record weight
// cos(theta) * 0.1 * rr (SURVEY 8a quirk iv: not a normalised Lambert estimator; kept).
__device__ __forceinline__ float sampleDiffuse(V3& dir, V3 n, float rr, float u1, float u2) {
  V3 ax, ay;
  tangentFrame(n, ax, ay);
  const float r = sqrtf(1.f - u1 * u1);
  const float phi = 2.f * kPi * u2;
  // The thin clear-coat shell (radii .4 / .4001, codelets.cpp:126-131) and the closed Cornell box amplify a last-bit
  // difference in the sampled direction into a different hit sequence, so sinf/cosf must be the reference's, bit for bit.
  float sp, cp;
  glibcSinCosf(phi, sp, cp);
  const V3 h = v3(cp * r, sp * r, u1);
  dir = v3(dot(v3(ax.x, ay.x, n.x), h), dot(v3(ax.y, ay.y, n.y), h), dot(v3(ax.z, ay.z, n.z), h));
  return dot(dir, n) * .1f * rr;
}

// sdf.hpp:36-40 reflect()
__device__ __forceinline__ void sampleMirror(V3& dir, V3 n) {
  const float c = dot(dir, n);
  dir = unit(sub(dir, scale(n, c * 2.f)));
}

// sdf.hpp:45-67 refract(): Schlick reflectance decides between transmission and reflection.
__device__ __forceinline__ bool sampleGlass(V3& dir, V3 n, float ri, float rnd) {
  float r0 = (1.f - ri) / (1.f + ri);
  r0 = r0 * r0;
  if (dot(n, dir) > 0.f) { n = v3(-n.x, -n.y, -n.z); } else { ri = 1.f / ri; }
  const float c1 = -dot(n, dir);
  const float c2 = 1.f - ri * ri * (1.f - c1 * c1);
  const float sb = 1.f - c1;
  const float sb2 = sb * sb;
  const float rprob = r0 + (1.f - r0) * (sb2 * sb * sb2);
  if (c2 > 0.f && rnd > rprob) {
    dir = unit(add(scale(dir, ri), scale(n, ri * c1 - sqrtf(c2))));
    return true;
  }
  dir = unit(add(dir, scale(n, c1 * 2.f)));
  return false;
}

// light.hpp:194-206 pixelToRay + codelets.cpp:68-75: pixel (+ binary16 product of scale and noise)
// -> camera-plane x,y, rounded to binary16 like the reference's half ray buffer.
__device__ __forceinline__ void cameraRay(uint32_t pu, uint32_t pv, float noise0, float noise1,
                                          const DeviceSettings& s, float& camX, float& camY) {
  const float c = (float)pu + roundToHalf(s.aaScale * noise0);
  const float r = (float)pv + roundToHalf(s.aaScale * noise1);
  const float w = (float)s.width, h = (float)s.height;
  const float x = (c / w) - .5f;
  const float y = (r / h) - .5f;
  camX = roundToHalf(2.f * x * s.aspect * s.tanHalfFov);
  camY = roundToHalf(-2.f * y * s.tanHalfFov);
}

// [host-checkable: tests/test_libm_restatement.py compiles this block with g++ and compares it with the host's libm]
// acosf as the reference's CPU build computes it: glibc's float routine (the fdlibm rational approximation in three
// ranges, all in binary32), which is not the correctly rounded arc cosine - (float)acos((double)x) differs from it
// for a few percent of the arguments. Checked exhaustively on the host against glibc 2.39 for every float in [-1, 1].
__device__ __forceinline__ float glibcAcosf(float x) {
  const float pi = 3.1415925026e+00f, pio2Hi = 1.5707962513e+00f, pio2Lo = 7.5497894159e-08f;
  const float p0 = 1.6666667163e-01f, p1 = -3.2556581497e-01f, p2 = 2.0121252537e-01f, p3 = -4.0055535734e-02f, p4 = 7.9153501429e-04f,
              p5 = 3.4793309169e-05f, q1 = -2.4033949375e+00f, q2 = 2.0209457874e+00f, q3 = -6.8828397989e-01f, q4 = 7.7038154006e-02f;
  const int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
  if (ix == 0x3f800000) return hx > 0 ? 0.f : pi + 2.f * pio2Lo;
  if (ix > 0x3f800000) return __int_as_float(0x7fc00000);
  if (ix < 0x3f000000) {                                   // |x| < 0.5
    if (ix <= 0x32800000) return pio2Hi + pio2Lo;
    const float z = x * x;
    const float p = z * (p0 + z * (p1 + z * (p2 + z * (p3 + z * (p4 + z * p5)))));
    const float q = 1.f + z * (q1 + z * (q2 + z * (q3 + z * q4)));
    const float r = p / q;
    return pio2Hi - (x - (pio2Lo - x * r));
  }
  if (hx < 0) {                                            // x < -0.5
    const float z = (1.f + x) * .5f;
    const float p = z * (p0 + z * (p1 + z * (p2 + z * (p3 + z * (p4 + z * p5)))));
    const float q = 1.f + z * (q1 + z * (q2 + z * (q3 + z * q4)));
    const float s = sqrtf(z);
    const float r = p / q;
    const float w = r * s - pio2Lo;
    return pi - 2.f * (s + w);
  }
  const float z = (1.f - x) * .5f;                         // x > 0.5
  const float s = sqrtf(z);
  const float df = __int_as_float(__float_as_int(s) & 0xfffff000);
  const float c = (z - df * df) / (s + df);
  const float p = z * (p0 + z * (p1 + z * (p2 + z * (p3 + z * (p4 + z * p5)))));
  const float q = 1.f + z * (q1 + z * (q2 + z * (q3 + z * q4)));
  const float r = p / q;
  const float w = r * s + c;
  return 2.f * (df + w);
}

// [end host-checkable]

// codelets.cpp:333-347 PreProcessEscapedRays: direction -> equirectangular lookup (u vertical).
// The reference calls acosf, and `atan2` unqualified on floats (codelets.cpp:334), which on a
// glibc toolchain is the double overload: phi, the azimuth add, the wrap and the final scale run
// in double and v is rounded to float once. The NIF encoding multiplies uv by up to 2^11 and
// rounds to binary16 (NifModel.cpp:211-213), so a last-bit difference in uv can move a feature
// argument by a whole binary16 step; hence acosf is glibc's own algorithm (above) and the azimuth
// stays in double (a 2-ulp double atan2 changes the float it rounds to about once in 10^8 rays).
__device__ __forceinline__ void escapedUv(V3 d, float azimuth, float& u, float& v) {
  // A normalised direction can have |y| = 1 + 1 ulp; acosf then returns NaN (in the reference too), the NIF turns it into
  // a NaN radiance and the pixel never recovers (measured: 5 pixels of 1104x1000 after 6.6e9 paths). Deliberate
  // deviation: clamp, i.e. treat the direction as exactly vertical.
  const float theta = glibcAcosf(fminf(fmaxf(d.y, -1.f), 1.f));
  double phi = atan2((double)d.z, (double)d.x) + (double)azimuth;
  const float twoPi = 2.f * kPi;
  if (phi < 0.0) { phi += (double)twoPi; } else if (phi > (double)twoPi) { phi -= (double)twoPi; }
  u = theta * (1.f / kPi);
  v = (float)(phi * (double)(1.f / twoPi));
}

struct PathResult {
  bool escaped;
  bool emitted;     // the path ended on an emitter (EMIT record, codelets.cpp:192-196): pixel += throughput, no NIF lookup
  int length;       // records the reference would have written (END included)
  V3 dir;           // unit direction of the escaped ray
  V3 throughput;    // escaped: pixel += throughput (*) L_env(dir); emitted: throughput already holds emission (*) path weight
};

// codelets.cpp:159-223 for one camera ray. The reference pushes one Contribution per vertex
// and folds the stack backwards afterwards (:255-298); the fold is linear in the terminal
// colour, so the same product is carried forward here (rounding order differs: < 1e-6 rel).
template <class Draw>
__device__ __forceinline__ PathResult tracePath(float camX, float camY, const DeviceSettings& s,
                                                const SceneSoA& sc, Draw&& draw) {
  PathResult res;
  V3 o = v3(0.f, 0.f, 0.f);
  V3 d = unit(v3(camX, camY, -1.f));          // Ray ctor normalises (light.hpp:35)
  V3 thr = v3(1.f, 1.f, 1.f);
  uint32_t depth = 0;
  int count = 0;
  const int maxLen = (int)s.maxPathLength;
  bool lit = false, emit = false;
  while (count != maxLen) {
    float rr = 1.f;
    if (depth >= s.rouletteDepth) {             // light.hpp:25-29
      const float rnd = draw();
      if (rnd <= s.stopProb) break;
      rr = s.rouletteWeight;
    }
    V3 n;
    const int hit = closestHit(sc, o, d, n);
    if (hit < 0) {                              // ESCAPED record {dir, rr}
      thr = scale(thr, rr);
      count += 1;
      lit = true;
      break;
    }
    if (sc.emissive[hit]) {                     // EMIT record {emission, rr} (codelets.cpp:192-196)
      thr = mul(thr, scale(v3(sc.emr[hit], sc.emg[hit], sc.emb[hit]), rr));
      count += 1;
      lit = true;
      emit = true;
      break;
    }
    const int bsdf = sc.bsdf[hit];
    if (bsdf == 0) {
      const float u1 = draw();
      const float u2 = draw();
      const float w = sampleDiffuse(d, n, rr, u1, u2);
      thr = scale(mul(thr, v3(sc.colr[hit], sc.colg[hit], sc.colb[hit])), w);
    } else if (bsdf == 1) {
      sampleMirror(d, n);
      thr = scale(thr, rr);
    } else {
      const bool through = sampleGlass(d, n, s.refractiveIndex, draw());
      const V3 c = through ? v3(sc.colr[hit], sc.colg[hit], sc.colb[hit]) : v3(1.f, 1.f, 1.f);
      thr = scale(mul(thr, c), 1.15f * rr);
    }
    count += 1;
    depth += 1;
  }
  if (!lit && count == 0) count = 1;            // END overwrites the last record (:219-222)
  res.escaped = lit && !emit;
  res.emitted = emit;
  res.length = count;
  res.dir = d;
  res.throughput = lit ? thr : v3(0.f, 0.f, 0.f);
  return res;
}

// One iteration of the loop above as a step of a per-lane state machine (the wavefront kernel regenerates a lane from
// the live-ray list as soon as its path ends, instead of letting it idle until the longest path of the warp is done).
// Same operations in the same order as tracePath, so the two agree bit for bit.
struct PathState {
  V3 o, d, thr;
  uint32_t depth;
  int count;
};
enum : int { kPathLive = 0, kPathEnded = 1, kPathEscaped = 2, kPathEmitted = 3 };

template <class Draw>
__device__ __forceinline__ int pathBounce(PathState& p, const DeviceSettings& s, const SceneSoA& sc, Draw&& draw) {
  float rr = 1.f;
  if (p.depth >= s.rouletteDepth) {             // light.hpp:25-29
    const float rnd = draw();
    if (rnd <= s.stopProb) {
      if (p.count == 0) p.count = 1;            // END overwrites the last record (:219-222)
      return kPathEnded;
    }
    rr = s.rouletteWeight;
  }
  V3 n;
  const int hit = closestHit(sc, p.o, p.d, n);
  if (hit < 0) {                                // ESCAPED record {dir, rr}
    p.thr = scale(p.thr, rr);
    p.count += 1;
    return kPathEscaped;
  }
  if (sc.emissive[hit]) {                       // EMIT record {emission, rr}
    p.thr = mul(p.thr, scale(v3(sc.emr[hit], sc.emg[hit], sc.emb[hit]), rr));
    p.count += 1;
    return kPathEmitted;
  }
  const int bsdf = sc.bsdf[hit];
  if (bsdf == 0) {
    const float u1 = draw();
    const float u2 = draw();
    const float w = sampleDiffuse(p.d, n, rr, u1, u2);
    p.thr = scale(mul(p.thr, v3(sc.colr[hit], sc.colg[hit], sc.colb[hit])), w);
  } else if (bsdf == 1) {
    sampleMirror(p.d, n);
    p.thr = scale(p.thr, rr);
  } else {
    const bool through = sampleGlass(p.d, n, s.refractiveIndex, draw());
    const V3 c = through ? v3(sc.colr[hit], sc.colg[hit], sc.colb[hit]) : v3(1.f, 1.f, 1.f);
    p.thr = scale(mul(p.thr, c), 1.15f * rr);
  }
  p.count += 1;
  p.depth += 1;
  return p.count == (int)s.maxPathLength ? kPathEnded : kPathLive;
}

} // namespace b2
