// Second built-in scene: the Cornell box of the reference's stand-alone CPU renderer
// (external/light/src/trace.cpp:168-216 scene, trace.hpp:8-107 path loop, jobs.hpp + xoshiro.hpp streams).
// Planes, an emissive ceiling disc and three emissive spheres: paths end on emitters (EMIT) instead of escaping
// to an environment light, so no NIF is involved. Compiled with -fmad=false like trace_kernels.cu: every value is
// bit-identical to the reference built with g++ for the same random stream (tests/test_cornell.py).
//
// Three entry points (include/b200pt.h):
//   b200pt_cornell_trace_paths  one thread per (camera ray, xoshiro state)                       - parity probe
//   b200pt_cornell_render       rng_mode TILE_STREAMS: one thread per tile job, the reference's jumped xoshiro
//                               streams consumed in its pixel order -> the reference's image for a seed, bit for bit
//                               rng_mode PIXEL_STREAMS (default): one stream per (pixel, sample slice), all SMs busy
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200pt.h"
#include "pt_device.cuh"

namespace b2 {

constexpr int kCornellObjects = 13;
constexpr int kCornellMaxDepth = 20;                    // trace.hpp:25

struct CornellScene {                                   // AoS in constant memory: every lane reads the same object
  int32_t shape[kCornellObjects];                       // 0 sphere, 1 disc, 2 plane
  int32_t bsdf[kCornellObjects];                        // 0 diffuse, 1 specular, 2 refractive (light.hpp:39-41)
  int32_t emissive[kCornellObjects];                    // emission.isNonZero() (light.hpp:55)
  float cx[kCornellObjects], cy[kCornellObjects], cz[kCornellObjects];
  float nx[kCornellObjects], ny[kCornellObjects], nz[kCornellObjects];
  float radius2[kCornellObjects], planeD[kCornellObjects];
  float colr[kCornellObjects], colg[kCornellObjects], colb[kCornellObjects];
  float emr[kCornellObjects], emg[kCornellObjects], emb[kCornellObjects];
};
__constant__ CornellScene gCornell;

struct CornellParams {
  uint32_t width, height, tileW, tileH, tilesX, tilesY, spp, slices;
  float sppF, aaScale, refractiveIndex, stopProb, rouletteWeight, eps, hitEps, aspect, tanTheta;
  uint32_t rouletteDepth;
  uint64_t seed;
};

// ---- xoshiro.hpp: xoroshiro128** and its float conversions -------------------------------------------
struct Xo { uint64_t s0, s1; };
__host__ __device__ __forceinline__ uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t z) {          // xoshiro.hpp:15-20
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t xoNext(Xo& s) {                      // :28-38
  const uint64_t a = s.s0;
  uint64_t b = s.s1;
  const uint64_t r = rotl64(a * 5, 7) * 9;
  b ^= a;
  s.s0 = rotl64(a, 24) ^ b ^ (b << 16);
  s.s1 = rotl64(b, 37);
  return r;
}
__device__ __forceinline__ double xoDouble(uint64_t x) { return __longlong_as_double((long long)((0x3FFull << 52) | (x >> 12))) - 1.0; }   // :60-63
__device__ __forceinline__ float xoUniform01(Xo& s) { return (float)xoDouble(xoNext(s)); }                                                  // :71-73
__device__ __forceinline__ float xoUniformNeg1To1(Xo& s) { return (float)(2.0 * xoDouble(xoNext(s)) - 1.0); }                              // :66-68

// ---- light.hpp:76-93 Plane, :95-121 Disc, :123-151 Sphere, :173-191 Scene::intersect ---------------------
// Branch-free: the reference's early returns become selects, so that the 32 lanes of a warp - each on its own path after
// the first bounce - stay on one instruction stream through all 13 tests (with branches ncu showed 9 of 32 lanes active).
// A select discards exactly what the early return skipped, so every surviving t has the reference's bits.
__device__ __forceinline__ float cornellSphereT(V3 centre, float radius2, V3 o, V3 d) {
  const V3 f = sub(centre, o);
  const float tca = dot(f, d);
  const V3 l = sub(centre, add(o, scale(d, tca)));
  const float l2 = norm2(l);
  // misses (discarded below) take sqrtf(1): a negative or zero argument would send the whole warp through sqrtf's slow path
  const float thc = sqrtf(l2 <= radius2 ? radius2 - l2 : 1.f);
  const float t0 = tca - thc, t1 = tca + thc;
  const float lo = t0 > t1 ? t1 : t0, hi = t0 > t1 ? t0 : t1;
  const float t = lo < 0.f ? hi : lo;
  return (!(tca < 0.f) && l2 <= radius2 && !(t < 0.f)) ? t : 0.f;
}

__device__ __forceinline__ int cornellClosestHit(const CornellScene& sc, V3& o, V3 d, V3& n, float eps, float hitEps) {
  int best = -1;
  float bestT = __int_as_float(0x7f800000);
#pragma unroll
  for (int i = 0; i < kCornellObjects; ++i) {            // fully unrolled: shapes and constant-bank offsets are compile-time
    float t;
    if (i < 6) {                                         // objects 0-5 spheres, 6-11 planes, 12 disc (buildCornell)
      t = cornellSphereT(v3(gCornell.cx[i], gCornell.cy[i], gCornell.cz[i]), gCornell.radius2[i], o, d);
    } else {
      const V3 pn = v3(gCornell.nx[i], gCornell.ny[i], gCornell.nz[i]);
      const float angle = dot(pn, d);
      // A ray leaving a wall has numerator exactly 0 for that wall about half of the time (o was rounded onto it): the
      // reference then gets t = -0/angle = +-0, a miss. 0 / x leaves the fast path of the IEEE division, so those lanes
      // divide 1 / angle instead and are marked as misses - same outcome, no slow-path call on every bounce.
      const float num = dot(pn, o) + gCornell.planeD[i];
      const bool onPlane = num == 0.f;
      const float tt = -(onPlane ? 1.f : num) / angle;
      bool ok = !onPlane && angle != 0.f && tt > eps;
      if (i == 12) {
        const V3 hp = add(o, scale(d, tt));
        ok = ok && norm2(sub(hp, v3(gCornell.cx[i], gCornell.cy[i], gCornell.cz[i]))) < gCornell.radius2[i];
      }
      t = ok ? tt : 0.f;
    }
    const bool closer = t > hitEps && t < bestT;
    bestT = closer ? t : bestT;
    best = closer ? i : best;
  }
  if (best >= 0) {
    // per-lane object index: read the shared-memory copy (a constant-bank read with a divergent index is replayed per address)
    o = add(o, scale(d, bestT));
    const V3 sphereN = unit(sub(o, v3(sc.cx[best], sc.cy[best], sc.cz[best])));
    n = best < 6 ? sphereN : v3(sc.nx[best], sc.ny[best], sc.nz[best]);
  }
  return best;
}

__device__ __forceinline__ void cornellStageScene(CornellScene& sc) {
  const uint32_t* src = reinterpret_cast<const uint32_t*>(&gCornell);
  uint32_t* dst = reinterpret_cast<uint32_t*>(&sc);
  for (uint32_t i = threadIdx.x; i < sizeof(CornellScene) / 4; i += blockDim.x) dst[i] = src[i];
  __syncthreads();
}

// trace.hpp:8-107 for one camera ray. The reference records one Contribution per vertex and folds them from the
// emitter back to the camera; the same fold runs here over (weight, object) pairs so that the rounding order matches.
__device__ __forceinline__ V3 cornellTrace(const CornellScene& sc, V3 dirIn, Xo& rng, const CornellParams& p, uint32_t& depthOut) {
  float weights[kCornellMaxDepth];
  int8_t objects[kCornellMaxDepth];        // >= 0: DIFFUSE by that object's colour; -1: SPECULAR / REFRACT
  V3 o = v3(0.f, 0.f, 0.f);
  V3 d = unit(dirIn);                      // Ray ctor normalises (light.hpp:35)
  uint32_t depth = 0;
  int emitter = -1;
  float emitWeight = 0.f;
  while (depth < (uint32_t)kCornellMaxDepth) {
    float rr = 1.f;
    if (depth >= p.rouletteDepth) {        // light.hpp:25-29
      const float rnd = xoUniform01(rng);
      if (rnd <= p.stopProb) break;
      rr = p.rouletteWeight;
    }
    V3 n;
    const int hit = cornellClosestHit(sc, o, d, n, p.eps, p.hitEps);
    if (hit < 0) break;
    if (sc.emissive[hit]) { emitter = hit; emitWeight = rr; break; }
    const int bsdf = sc.bsdf[hit];
    if (bsdf == 0) {
      const float u1 = xoUniform01(rng);
      const float u2 = xoUniform01(rng);
      weights[depth] = sampleDiffuse(d, n, rr, u1, u2);
      objects[depth] = (int8_t)hit;
    } else if (bsdf == 1) {
      sampleMirror(d, n);
      weights[depth] = rr;
      objects[depth] = -1;
    } else {
      sampleGlass(d, n, p.refractiveIndex, xoUniform01(rng));
      weights[depth] = 1.15f * rr;
      objects[depth] = -1;
    }
    depth += 1;
  }
  depthOut = depth;
  V3 total = v3(0.f, 0.f, 0.f);
  if (emitter >= 0) {
    // EMIT: total += emission * weight (total is zero: the sum is exact)
    total = add(total, scale(v3(sc.emr[emitter], sc.emg[emitter], sc.emb[emitter]), emitWeight));
    for (int i = (int)depth - 1; i >= 0; --i) {
      const int ob = objects[i];
      if (ob >= 0) total = scale(mul(total, v3(sc.colr[ob], sc.colg[ob], sc.colb[ob])), weights[i]);
      else total = scale(total, weights[i]);
    }
  }
  return total;
}

// trace.cpp:253-259: pixelToRay (light.hpp:194-206) + anti-alias jitter on the camera plane. The two noise draws are
// arguments of one constructor call in the reference; g++ evaluates them right to left, so the FIRST draw lands in y
// (pinned against the compiled reference by tests/test_cornell.py).
__device__ __forceinline__ V3 cornellCameraRay(uint32_t col, uint32_t row, Xo& rng, const CornellParams& p) {
  const float w = (float)p.width, h = (float)p.height;
  const float x = ((float)col / w) - .5f;
  const float y = ((float)row / h) - .5f;
  V3 cam = v3(2.f * x * p.aspect * p.tanTheta, -2.f * y * p.tanTheta, -1.f);
  const float noiseY = xoUniformNeg1To1(rng);
  const float noiseX = xoUniformNeg1To1(rng);
  cam = add(cam, scale(v3(noiseX, noiseY, 0.f), p.aaScale));
  return cam;
}

__device__ __forceinline__ void cornellStats(uint64_t casts, uint32_t longest, unsigned long long* stats) {
  for (int o = 16; o > 0; o >>= 1) {
    casts += __shfl_xor_sync(0xffffffffu, casts, o);
    longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&stats[0], (unsigned long long)casts); atomicMax(&stats[1], (unsigned long long)longest); }
}

// ---- kernels ---------------------------------------------------------------------------------------------

__global__ void cornell_paths_kernel(size_t n, const float* __restrict__ camXY, uint64_t* __restrict__ state, float* __restrict__ rgb,
                                     uint32_t* __restrict__ depth, CornellParams p) {
  __shared__ CornellScene sc;
  cornellStageScene(sc);
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Xo rng{state[2 * i], state[2 * i + 1]};
  uint32_t dep;
  const V3 c = cornellTrace(sc, v3(camXY[2 * i], camXY[2 * i + 1], -1.f), rng, p, dep);
  rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
  depth[i] = dep;
  state[2 * i] = rng.s0; state[2 * i + 1] = rng.s1;
}

// The reference's schedule (trace.cpp:246-262, jobs.hpp:46-55): a tile job owns one xoshiro stream that its pixels
// consume in row-major order, one sample per pass, spp passes. A job is inherently serial, so parallelism is the
// number of tiles; this mode exists to reproduce the reference's image exactly, not for speed.
__global__ void cornell_tiles_kernel(const uint64_t* __restrict__ jobState, float* __restrict__ film, unsigned long long* stats, CornellParams p) {
  __shared__ CornellScene sc;
  cornellStageScene(sc);
  const uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t casts = 0;
  uint32_t longest = 0;
  if (job < p.tilesX * p.tilesY) {
    const uint32_t r0 = (job / p.tilesX) * p.tileH, c0 = (job % p.tilesX) * p.tileW;
    const uint32_t r1 = min(r0 + p.tileH, p.height), c1 = min(c0 + p.tileW, p.width);
    Xo rng{jobState[2 * job], jobState[2 * job + 1]};
    for (uint32_t s = 0; s < p.spp; ++s)
      for (uint32_t r = r0; r < r1; ++r)
        for (uint32_t c = c0; c < c1; ++c) {
          const V3 cam = cornellCameraRay(c, r, rng, p);
          uint32_t dep;
          const V3 col = cornellTrace(sc, cam, rng, p, dep);
          float* px = film + ((size_t)r * p.width + c) * 3;
          px[0] += col.x / p.sppF; px[1] += col.y / p.sppF; px[2] += col.z / p.sppF;      // p += colour / spp
          casts += dep;
          longest = max(longest, dep);
        }
  }
  cornellStats(casts, longest, stats);
}

// B200 schedule: one stream per (pixel, slice of the samples), seeded by splitmix64 of (seed, pixel, slice); the
// samples of a slice are summed in registers and added to the film with one atomic per channel.
// Paths in the closed box last 0..20 bounces (6.5 on average), so a lane that traced "its" path to the end while the
// rest of the warp waits would leave two thirds of the issue slots empty (measured: 9.4 of 32 lanes active per
// instruction). Instead every lane runs a bounce-level state machine: one loop iteration is one bounce, and a lane whose
// path ended starts its next sample in the same iteration. The draws of a thread keep their order, so the image is the
// one cornellTrace would give sample by sample (tests/test_cornell.py re-derives it through b200pt_cornell_trace_paths).
__global__ void __launch_bounds__(128) cornell_pixels_kernel(float* __restrict__ film, unsigned long long* stats, CornellParams p) {
  __shared__ CornellScene sc;
  cornellStageScene(sc);
  const uint64_t pixels = (uint64_t)p.width * p.height;
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t casts = 0;
  uint32_t longest = 0;
  // consecutive lanes take consecutive pixels of one slice: neighbouring paths stay coherent for a few bounces
  const uint64_t pixel = t % pixels;
  const uint32_t slice = (uint32_t)(t / pixels);
  const uint32_t row = (uint32_t)(pixel / p.width), col = (uint32_t)(pixel % p.width);
  Xo rng;
  rng.s0 = splitmix64(p.seed ^ (0xd1342543de82ef95ull * (t + 1)));
  rng.s1 = splitmix64(rng.s0);
  const uint32_t per = (p.spp + p.slices - 1) / p.slices;
  const uint32_t s0 = slice * per, s1 = min(s0 + per, p.spp);
  uint32_t s = s0;
  bool alive = t < pixels * p.slices && s0 < s1;
  const bool owner = alive;
  V3 sum = v3(0.f, 0.f, 0.f);
  float weights[kCornellMaxDepth];
  int8_t objects[kCornellMaxDepth];
  V3 o = v3(0.f, 0.f, 0.f), d = v3(0.f, 0.f, -1.f);
  uint32_t depth = 0;
  bool fresh = true;
  // The vote keeps the whole warp in one loop: without it the compiler threads the "path ended -> next sample" edge into
  // a private copy of the loop head and the lanes never reconverge (measured 9.3 of 32 lanes per instruction either way).
  while (__any_sync(0xffffffffu, alive)) {
    if (!alive) continue;
    if (fresh) {
      d = unit(cornellCameraRay(col, row, rng, p));
      o = v3(0.f, 0.f, 0.f);
      depth = 0;
      fresh = false;
    }
    // one bounce of trace.hpp:30-70. The roulette draw and the intersection are issued for every live lane and their
    // effects selected afterwards: as branches, ptxas kept the "no roulette yet" and "survived roulette" lanes on two
    // separate copies of the intersection code (every instruction of it issued twice per iteration, 13 lanes each).
    const bool roulette = depth >= p.rouletteDepth;
    Xo drawn = rng;
    const float rnd = xoUniform01(drawn);
    rng.s0 = roulette ? drawn.s0 : rng.s0;
    rng.s1 = roulette ? drawn.s1 : rng.s1;
    const bool stopped = roulette && rnd <= p.stopProb;
    const float rr = roulette ? p.rouletteWeight : 1.f;
    V3 n, hitPoint = o;
    const int hit = cornellClosestHit(sc, hitPoint, d, n, p.eps, p.hitEps);
    const bool emits = !stopped && hit >= 0 && sc.emissive[hit < 0 ? 0 : hit] != 0;
    const int emitter = emits ? hit : -1;
    bool ended = stopped || hit < 0 || emits;
    if (!ended) {
      o = hitPoint;
      const int bsdf = sc.bsdf[hit];
      if (bsdf == 0) {
        const float u1 = xoUniform01(rng);
        const float u2 = xoUniform01(rng);
        weights[depth] = sampleDiffuse(d, n, rr, u1, u2);
        objects[depth] = (int8_t)hit;
      } else if (bsdf == 1) {
        sampleMirror(d, n);
        weights[depth] = rr;
        objects[depth] = -1;
      } else {
        sampleGlass(d, n, p.refractiveIndex, xoUniform01(rng));
        weights[depth] = 1.15f * rr;
        objects[depth] = -1;
      }
      depth += 1;
      ended = depth >= (uint32_t)kCornellMaxDepth;
    }
    if (ended) {
      if (emitter >= 0) {          // fold from the emitter back to the camera (trace.hpp:76-104)
        V3 total = scale(v3(sc.emr[emitter], sc.emg[emitter], sc.emb[emitter]), rr);
        for (int i = (int)depth - 1; i >= 0; --i) {
          const int ob = objects[i];
          if (ob >= 0) total = scale(mul(total, v3(sc.colr[ob], sc.colg[ob], sc.colb[ob])), weights[i]);
          else total = scale(total, weights[i]);
        }
        sum = add(sum, v3(total.x / p.sppF, total.y / p.sppF, total.z / p.sppF));
      }
      casts += depth;
      longest = max(longest, depth);
      s += 1;
      fresh = true;
      alive = s < s1;
    }
  }
  if (owner) {
    float* px = film + pixel * 3;
    if (p.slices == 1) { px[0] = sum.x; px[1] = sum.y; px[2] = sum.z; }
    else { atomicAdd(px, sum.x); atomicAdd(px + 1, sum.y); atomicAdd(px + 2, sum.z); }
  }
  cornellStats(casts, longest, stats);
}

} // namespace b2

// ---- host side ---------------------------------------------------------------------------------------------

namespace {
using namespace b2;

thread_local std::string g_cornellError;
int cornellFail(int code, const std::string& what) { g_cornellError = what; return code; }

// trace.cpp:168-216, with the constructors' arithmetic (Plane/Disc normalise their normal, Disc::d = |c.n|,
// Sphere::radius2 = r*r, colours carry colourGain 15).
void buildCornell(CornellScene& sc) {
  std::memset(&sc, 0, sizeof(sc));
  int n = 0;
  auto material = [&](int i, const float col[3], const float em[3], int bsdf) {
    sc.colr[i] = col[0]; sc.colg[i] = col[1]; sc.colb[i] = col[2];
    sc.emr[i] = em[0]; sc.emg[i] = em[1]; sc.emb[i] = em[2];
    sc.bsdf[i] = bsdf;
    sc.emissive[i] = (em[0] != 0.f || em[1] != 0.f || em[2] != 0.f) ? 1 : 0;
  };
  auto unitNormal = [&](int i, float x, float y, float z) {
    const float inv = 1.f / sqrtf(x * x + y * y + z * z);
    sc.nx[i] = x * inv; sc.ny[i] = y * inv; sc.nz[i] = z * inv;
  };
  const float zero[3] = {0.f, 0.f, 0.f};
  const float gain = 15.f;
  const float sphereColour[3] = {1.f * gain, .89f * gain, .55f * gain};
  const float wall1[3] = {.98f * gain, .76f * gain, .66f * gain};
  const float wall2[3] = {.93f * gain, .43f * gain, .48f * gain};
  const float wall3[3] = {.27f * gain, .31f * gain, .38f * gain};
  const float mirror[3] = {4.f, 8.f, 4.f}, glass[3] = {10.f, 10.f, 1.f};
  const float lightW[3] = {8000, 8000, 8000}, lightR[3] = {1000, 367, 367}, lightG[3] = {467, 1000, 434}, lightB[3] = {500, 600, 1000};
  auto sphere = [&](float x, float y, float z, float r, const float col[3], const float em[3], int bsdf) {
    sc.shape[n] = 0; sc.cx[n] = x; sc.cy[n] = y; sc.cz[n] = z; sc.radius2[n] = r * r;
    material(n, col, em, bsdf); ++n;
  };
  auto plane = [&](float x, float y, float z, float d, const float col[3]) {
    sc.shape[n] = 2; unitNormal(n, x, y, z); sc.planeD[n] = d;
    material(n, col, zero, 0); ++n;
  };
  sphere(-0.75f, -1.45f, -4.4f, 1.05f, mirror, zero, 1);
  sphere(2.0f, -2.05f, -3.7f, 0.5f, glass, zero, 2);
  sphere(-1.75f, -1.95f, -3.1f, 0.6f, sphereColour, zero, 0);
  sphere(-1.12f, -2.3f, -3.5f, 0.2f, zero, lightB, 1);
  sphere(-0.28f, -2.34f, -3.f, 0.2f, zero, lightG, 1);
  sphere(.58f, -2.39f, -2.6f, 0.2f, zero, lightR, 1);
  plane(0.f, 1.f, 0.f, 2.5f, wall1);          // bottom
  plane(0.f, 0.f, 1.f, 5.5f, wall1);          // back
  plane(1.f, 0.f, 0.f, 2.75f, wall2);         // left
  plane(-1.f, -0.f, -0.f, 2.75f, wall3);      // right  (-X of the reference negates every component)
  plane(-0.f, -1.f, -0.f, 3.f, wall1);        // ceiling
  plane(-0.f, -0.f, -1.f, .5f, wall1);        // front
  {                                           // ceiling light: Disc(-Y, (0, 2.9999, -4), .7)
    sc.shape[n] = 1; unitNormal(n, -0.f, -1.f, -0.f);
    sc.cx[n] = 0.f; sc.cy[n] = 2.9999f; sc.cz[n] = -4.f;
    sc.planeD[n] = std::fabs(sc.cx[n] * sc.nx[n] + sc.cy[n] * sc.ny[n] + sc.cz[n] * sc.nz[n]);
    sc.radius2[n] = .7f * .7f;
    material(n, zero, lightW, 0); ++n;
  }
}

int cornellPrepare(const b200pt_cornell_config* cfg, CornellParams& p) {
  if (!cfg) return cornellFail(B200PT_ERR_INVALID, "null config");
  if (cfg->struct_size != sizeof(b200pt_cornell_config)) return cornellFail(B200PT_ERR_INVALID, "b200pt_cornell_config.struct_size mismatch");
  if (!cfg->width || !cfg->height || !cfg->tile_width || !cfg->tile_height || !cfg->samples) return cornellFail(B200PT_ERR_INVALID, "width, height, tile size and samples must be > 0");
  if (cfg->rng_mode != B200PT_CORNELL_PIXEL_STREAMS && cfg->rng_mode != B200PT_CORNELL_TILE_STREAMS) return cornellFail(B200PT_ERR_INVALID, "unknown rng_mode");
  if (!(cfg->stop_prob >= 0.f && cfg->stop_prob < 1.f)) return cornellFail(B200PT_ERR_INVALID, "stop_prob must be in [0, 1)");
  int devCount = 0;
  const cudaError_t e = cudaGetDeviceCount(&devCount);
  if (e != cudaSuccess || devCount == 0) return cornellFail(B200PT_ERR_CUDA, std::string("no CUDA device (this library has no CPU path): ") + cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= devCount) return cornellFail(B200PT_ERR_INVALID, "device ordinal out of range");
  if (cudaSetDevice(cfg->device) != cudaSuccess) return cornellFail(B200PT_ERR_CUDA, "cudaSetDevice failed");
  CornellScene sc;
  buildCornell(sc);
  if (cudaMemcpyToSymbol(gCornell, &sc, sizeof(sc)) != cudaSuccess) return cornellFail(B200PT_ERR_CUDA, std::string("scene upload: ") + cudaGetErrorString(cudaGetLastError()));
  std::memset(&p, 0, sizeof(p));
  p.width = cfg->width; p.height = cfg->height; p.tileW = cfg->tile_width; p.tileH = cfg->tile_height;
  p.tilesX = (cfg->width + cfg->tile_width - 1) / cfg->tile_width;
  p.tilesY = (cfg->height + cfg->tile_height - 1) / cfg->tile_height;
  p.spp = cfg->samples; p.sppF = (float)cfg->samples;
  p.aaScale = cfg->aa_noise_scale; p.refractiveIndex = cfg->refractive_index; p.stopProb = cfg->stop_prob;
  p.rouletteWeight = 1.f / (1.f - cfg->stop_prob);                  // light.hpp:28
  p.eps = cfg->epsilon; p.hitEps = cfg->intersection_epsilon;
  p.aspect = (float)cfg->width / (float)cfg->height;                // light.hpp:198
  p.tanTheta = tanf((3.14159265358979323846264338327950288f / 2.f) / 2.f);      // fov = Pi/2 (trace.cpp:254), light.hpp:199
  p.rouletteDepth = cfg->roulette_depth; p.seed = cfg->seed;
  return B200PT_OK;
}

#define CORNELL_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { rc = cornellFail(B200PT_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); goto done; } } while (0)

} // namespace

extern "C" {

const char* b200pt_cornell_last_error(void) { return g_cornellError.c_str(); }

namespace {
struct DeviceGuard {          // the entry points select cfg->device; the caller's current device is restored on return
  int previous = -1;
  DeviceGuard() { if (cudaGetDevice(&previous) != cudaSuccess) previous = -1; }
  ~DeviceGuard() { if (previous >= 0) cudaSetDevice(previous); }
};
}

int b200pt_cornell_trace_paths(const b200pt_cornell_config* cfg, size_t n, const float* camXY, uint64_t* rngState, float* rgb, uint32_t* depth) {
  DeviceGuard guard;
  CornellParams p;
  int rc = cornellPrepare(cfg, p);
  if (rc != B200PT_OK) return rc;
  if (n == 0) return B200PT_OK;
  if (!camXY || !rngState || !rgb || !depth) return cornellFail(B200PT_ERR_INVALID, "null buffer");
  float *dCam = nullptr, *dRgb = nullptr; uint64_t* dState = nullptr; uint32_t* dDepth = nullptr;
  CORNELL_CUDA(cudaMalloc(&dCam, n * 2 * sizeof(float)));
  CORNELL_CUDA(cudaMalloc(&dRgb, n * 3 * sizeof(float)));
  CORNELL_CUDA(cudaMalloc(&dState, n * 2 * sizeof(uint64_t)));
  CORNELL_CUDA(cudaMalloc(&dDepth, n * sizeof(uint32_t)));
  CORNELL_CUDA(cudaMemcpy(dCam, camXY, n * 2 * sizeof(float), cudaMemcpyHostToDevice));
  CORNELL_CUDA(cudaMemcpy(dState, rngState, n * 2 * sizeof(uint64_t), cudaMemcpyHostToDevice));
  cornell_paths_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, dCam, dState, dRgb, dDepth, p);
  CORNELL_CUDA(cudaGetLastError());
  CORNELL_CUDA(cudaMemcpy(rgb, dRgb, n * 3 * sizeof(float), cudaMemcpyDeviceToHost));
  CORNELL_CUDA(cudaMemcpy(depth, dDepth, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  CORNELL_CUDA(cudaMemcpy(rngState, dState, n * 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
done:
  cudaFree(dCam); cudaFree(dRgb); cudaFree(dState); cudaFree(dDepth);
  return rc;
}

int b200pt_cornell_render(const b200pt_cornell_config* cfg, float* rgbHost, b200pt_cornell_stats* stats) {
  DeviceGuard guard;
  CornellParams p;
  int rc = cornellPrepare(cfg, p);
  if (rc != B200PT_OK) return rc;
  if (!rgbHost) return cornellFail(B200PT_ERR_INVALID, "null image buffer");
  const size_t floats = (size_t)p.width * p.height * 3;
  float* dFilm = nullptr; unsigned long long* dStats = nullptr; uint64_t* dJobs = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned long long hStats[2] = {0, 0};
  float ms = 0.f;
  CORNELL_CUDA(cudaMalloc(&dFilm, floats * sizeof(float)));
  CORNELL_CUDA(cudaMalloc(&dStats, sizeof(hStats)));
  CORNELL_CUDA(cudaMemset(dFilm, 0, floats * sizeof(float)));
  CORNELL_CUDA(cudaMemset(dStats, 0, sizeof(hStats)));
  CORNELL_CUDA(cudaEventCreate(&ev0));
  CORNELL_CUDA(cudaEventCreate(&ev1));
  if (cfg->rng_mode == B200PT_CORNELL_TILE_STREAMS) {
    // createTracingJobs (jobs.hpp:66-90): one stream per tile, 2^64 draws apart (xoshiro.hpp:43-58 jump)
    const size_t jobs = (size_t)p.tilesX * p.tilesY;
    std::vector<uint64_t> states(jobs * 2);
    Xo s;
    s.s0 = splitmix64(cfg->seed);
    s.s1 = splitmix64(s.s0);                                      // xoshiro.hpp:23-26
    for (size_t j = 0; j < jobs; ++j) {
      states[2 * j] = s.s0; states[2 * j + 1] = s.s1;
      static const uint64_t kJump[2] = {0xdf900294d8f554a5ull, 0x170865df4b3201fcull};
      uint64_t a = 0, b = 0;
      for (int w = 0; w < 2; ++w)
        for (int bit = 0; bit < 64; ++bit) {
          if (kJump[w] & (1ull << bit)) { a ^= s.s0; b ^= s.s1; }
          xoNext(s);
        }
      s.s0 = a; s.s1 = b;
    }
    CORNELL_CUDA(cudaMalloc(&dJobs, states.size() * sizeof(uint64_t)));
    CORNELL_CUDA(cudaMemcpy(dJobs, states.data(), states.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
    CORNELL_CUDA(cudaEventRecord(ev0));
    cornell_tiles_kernel<<<(unsigned)((jobs + 31) / 32), 32>>>(dJobs, dFilm, dStats, p);      // one warp per CTA: spread the few jobs over all SMs
  } else {
    const uint64_t pixels = (uint64_t)p.width * p.height;
    const uint64_t want = 148ull * 2048ull * 2ull;                 // two waves of resident threads
    p.slices = (uint32_t)std::min<uint64_t>(p.spp, std::max<uint64_t>(1, (want + pixels - 1) / pixels));
    const uint64_t threads = pixels * p.slices;
    CORNELL_CUDA(cudaEventRecord(ev0));
    cornell_pixels_kernel<<<(unsigned)((threads + 127) / 128), 128>>>(dFilm, dStats, p);
  }
  CORNELL_CUDA(cudaGetLastError());
  CORNELL_CUDA(cudaEventRecord(ev1));
  CORNELL_CUDA(cudaEventSynchronize(ev1));
  CORNELL_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
  CORNELL_CUDA(cudaMemcpy(rgbHost, dFilm, floats * sizeof(float), cudaMemcpyDeviceToHost));
  CORNELL_CUDA(cudaMemcpy(hStats, dStats, sizeof(hStats), cudaMemcpyDeviceToHost));
  if (stats) { stats->total_ray_casts = hStats[0]; stats->max_path_length = hStats[1]; stats->device_ms = ms; }
done:
  cudaFree(dFilm); cudaFree(dStats); cudaFree(dJobs);
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
  return rc;
}

} // extern "C"

// ---- b200pt_cornell_render_noalloc: the tile-stream render without a single allocation -------------------------
//
// For callers that cannot call cudaMalloc while they render (see include/b200pt.h: a plug-in living on a buffer and a memory
// share given to it, a frame loop that must not take cudaMalloc's device-wide synchronisation, a neighbour allocator that
// owns the rest of the GPU). cornell_tiles_kernel needs working memory that is neither input nor output: one xoroshiro128**
// state per tile job and the two statistics counters. b200pt_cornell_render gets it from cudaMalloc; here it is this array,
// which is part of the library's device image, so every device that loads the library has its own copy for as long as the
// process lives. Layout: bytes [0,16) the counters (ray casts, longest path), bytes [256, 256 + 16 * tiles) the states.
//
// In host code the name gCornellScratch is only a handle for the symbol, not an address a kernel could use: the per-device
// address comes from cudaGetSymbolAddress, asked once per device with that device current, and it is that pointer the
// kernel receives.
static __device__ __align__(256) unsigned char gCornellScratch[256 + (size_t)B200PT_CORNELL_NOALLOC_MAX_TILES * 2 * sizeof(uint64_t)];

namespace {

constexpr int kScratchMaxDevices = 64;
constexpr size_t kScratchStatesOffset = 256;
std::mutex g_scratchMutex;                                  // guards the table below, not the array's contents
unsigned char* g_scratchAddress[kScratchMaxDevices] = {};

// Device address of gCornellScratch on `device`, which must be the current device (cornellPrepare has selected it).
int cornellScratchAddress(int device, unsigned char** out) {
  if (device < 0 || device >= kScratchMaxDevices) return cornellFail(B200PT_ERR_INVALID, "device ordinal beyond the scratch table (64 devices)");
  std::lock_guard<std::mutex> lock(g_scratchMutex);
  if (!g_scratchAddress[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, gCornellScratch);
    if (e != cudaSuccess || !p) return cornellFail(B200PT_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch array): ") + cudaGetErrorString(e));
    g_scratchAddress[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratchAddress[device];
  return B200PT_OK;
}

} // namespace

extern "C" int b200pt_cornell_render_noalloc(const b200pt_cornell_config* cfg, void* rgbDevice, b200pt_cornell_stats* stats) {
  DeviceGuard guard;
  CornellParams p;
  int rc = cornellPrepare(cfg, p);                          // validates cfg, selects cfg->device, uploads the scene constants
  if (rc != B200PT_OK) return rc;
  if (cfg->rng_mode != B200PT_CORNELL_TILE_STREAMS)
    return cornellFail(B200PT_ERR_INVALID, "b200pt_cornell_render_noalloc renders rng_mode = B200PT_CORNELL_TILE_STREAMS only");
  const uint64_t jobs = (uint64_t)p.tilesX * p.tilesY;
  if (jobs > (uint64_t)B200PT_CORNELL_NOALLOC_MAX_TILES) {
    char msg[200];
    snprintf(msg, sizeof(msg), "%llu tiles exceed the %d tiles the built-in scratch array holds (B200PT_CORNELL_NOALLOC_MAX_TILES): use larger tiles or b200pt_cornell_render",
             (unsigned long long)jobs, (int)B200PT_CORNELL_NOALLOC_MAX_TILES);
    return cornellFail(B200PT_ERR_INVALID, msg);
  }
  if (!rgbDevice) return cornellFail(B200PT_ERR_INVALID, "null image buffer");
  {
    cudaPointerAttributes attr{};
    const cudaError_t e = cudaPointerGetAttributes(&attr, rgbDevice);
    if (e != cudaSuccess) { cudaGetLastError(); return cornellFail(B200PT_ERR_INVALID, std::string("rgb_device is not a CUDA pointer: ") + cudaGetErrorString(e)); }
    if (attr.type != cudaMemoryTypeDevice && attr.type != cudaMemoryTypeManaged)
      return cornellFail(B200PT_ERR_INVALID, "rgb_device must be device memory (a host pointer was passed; b200pt_cornell_render takes host buffers)");
    if (attr.type == cudaMemoryTypeDevice && attr.device != cfg->device)
      return cornellFail(B200PT_ERR_INVALID, "rgb_device lives on another device than cfg->device");
  }
  unsigned char* scratch = nullptr;
  rc = cornellScratchAddress(cfg->device, &scratch);
  if (rc != B200PT_OK) return rc;
  unsigned long long* dStats = reinterpret_cast<unsigned long long*>(scratch);
  uint64_t* dJobs = reinterpret_cast<uint64_t*>(scratch + kScratchStatesOffset);
  float* dFilm = static_cast<float*>(rgbDevice);
  const size_t floats = (size_t)p.width * p.height * 3;
  unsigned long long hStats[2] = {0, 0};

  // The array still holds the previous call's counters and states. The counters are accumulated with atomics and the film
  // with +=, so both must start at zero; the states of the tiles this call uses are all rewritten below.
  CORNELL_CUDA(cudaMemset(dStats, 0, sizeof(hStats)));
  CORNELL_CUDA(cudaMemset(dFilm, 0, floats * sizeof(float)));
  {
    // createTracingJobs (jobs.hpp:66-90) like b200pt_cornell_render, but generated in stack-sized pieces so that the host
    // heap is not touched either: one stream per tile, 2^64 draws apart (xoshiro.hpp:43-58 jump).
    constexpr size_t kPiece = 512;
    uint64_t piece[kPiece * 2];
    Xo s;
    s.s0 = splitmix64(cfg->seed);
    s.s1 = splitmix64(s.s0);
    for (uint64_t j0 = 0; j0 < jobs; j0 += kPiece) {
      const size_t count = (size_t)std::min<uint64_t>(kPiece, jobs - j0);
      for (size_t j = 0; j < count; ++j) {
        piece[2 * j] = s.s0; piece[2 * j + 1] = s.s1;
        static const uint64_t kJump[2] = {0xdf900294d8f554a5ull, 0x170865df4b3201fcull};
        uint64_t a = 0, b = 0;
        for (int w = 0; w < 2; ++w)
          for (int bit = 0; bit < 64; ++bit) {
            if (kJump[w] & (1ull << bit)) { a ^= s.s0; b ^= s.s1; }
            xoNext(s);
          }
        s.s0 = a; s.s1 = b;
      }
      // pageable source: the copy has left `piece` when cudaMemcpy returns, so the buffer can be refilled
      CORNELL_CUDA(cudaMemcpy(dJobs + 2 * j0, piece, count * 2 * sizeof(uint64_t), cudaMemcpyHostToDevice));
    }
  }
  cornell_tiles_kernel<<<(unsigned)((jobs + 31) / 32), 32>>>(dJobs, dFilm, dStats, p);
  CORNELL_CUDA(cudaGetLastError());
  CORNELL_CUDA(cudaMemcpy(hStats, dStats, sizeof(hStats), cudaMemcpyDeviceToHost));     // waits for the kernel: the image is complete
  if (stats) { stats->total_ray_casts = hStats[0]; stats->max_path_length = hStats[1]; stats->device_ms = 0.0; }
done:
  return rc;
}
