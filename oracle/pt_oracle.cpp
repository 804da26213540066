// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, imported by or executed from
// the product path. See oracle_common.hpp for the usage rule and parity status.
//
// CPU restatement of the reference's per-pixel Monte-Carlo path-tracing loop:
//   camera rays      src/codelets/codelets.cpp:36-80, external/light/src/light.hpp:194-206
//   path loop        src/codelets/codelets.cpp:93-227
//   scene            src/codelets/codelets.cpp:110-144
//   intersection     external/light/src/light.hpp:95-121 (disc), :123-152 (sphere), :165-192 (scene)
//   BSDF sampling    external/light/src/sdf.hpp:8-31 (diffuse), :36-40 (reflect), :45-67 (refract)
//   roulette         external/light/src/light.hpp:25-29
//   escaped-ray uv   src/codelets/codelets.cpp:312-358
//   accumulate       src/codelets/codelets.cpp:235-306, src/codelets/WrappedArray.hpp:45-68
//   stage order      src/PathTracerApp.cpp:432-458
// It is written as a table-driven state machine over a flat primitive table (no virtual
// primitives, no stack wrapper classes) but performs the same float operations in the
// same order, so that it reproduces the reference's records bit for bit on a CPU.
// Pinned against oracle/_ref (the unmodified codelets) by tests/test_oracle_vs_ref.py.

#include "oracle_common.hpp"

#include <cfloat>
#include <limits>
#include <vector>

#include <omp.h>

namespace orc {

struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 scale(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float norm2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
// vector.hpp:47 : v * (1/sqrt(|v|^2))
static inline V3 unit(V3 a) { return scale(a, 1.f / sqrtf(norm2(a))); }
static inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }

static constexpr float kPi = 3.14159265358979323846264338327950288f;   // light.hpp:16
static const float kEps = std::numeric_limits<float>::epsilon();   // light.cpp:5
static const float kHitEps = 1e-5f;                                // light.cpp:9

enum Shape { SPHERE = 0, DISC = 1 };
enum Bsdf { DIFFUSE = 0, SPECULAR = 1, REFRACTIVE = 2 };

struct Prim {
  int shape;
  V3 centre;       // sphere centre / disc centre
  V3 normal;       // disc only (unit)
  float radius2;   // r*r
  float planeD;    // disc only: |centre . n|   (light.hpp:101, note the abs)
  V3 colour;
  int bsdf;
  int emissive;    // light.hpp Material::emissive: emission != 0
  V3 emission;
};

// codelets.cpp:110-144. Colours carry the 2x gain (:127-130). No emitters (:117,142).
static const int kNumPrims = 6;
static Prim makeSphere(float cx, float cy, float cz, float r, V3 colour, int bsdf) {
  Prim p; p.shape = SPHERE; p.centre = v3(cx, cy, cz); p.normal = v3(0.f, 1.f, 0.f);
  p.radius2 = r * r; p.planeD = 0.f; p.colour = colour; p.bsdf = bsdf; p.emissive = 0; p.emission = v3(0.f, 0.f, 0.f); return p;
}
static Prim makeDisc(V3 n, V3 c, float r, V3 colour, int bsdf) {
  Prim p; p.shape = DISC; p.centre = c; p.normal = unit(n);
  p.radius2 = r * r; p.planeD = std::abs(dot(c, p.normal)); p.colour = colour; p.bsdf = bsdf; p.emissive = 0; p.emission = v3(0.f, 0.f, 0.f); return p;
}
// The reference anticipates passing the scene as data (codelets.cpp:90-92 "currently hard coded"); orc_set_scene
// replaces the built-in table (the product's counterpart is b200pt_set_scene). Not thread-safe: set it between calls.
static std::vector<Prim> g_userScene;
static const Prim* builtinScene();
static const Prim* sceneTable() { return g_userScene.empty() ? builtinScene() : g_userScene.data(); }
static int sceneCount() { return g_userScene.empty() ? kNumPrims : (int)g_userScene.size(); }
static const Prim* builtinScene() {
  static Prim prims[kNumPrims];
  static bool built = false;
  if (!built) {
    const float gain = 2.f;
    // The double-typed literals of the reference (codelets.cpp:114-116: -3.23, -0.19931)
    // are narrowed to float by the Vec constructor, which (float)literal reproduces.
    prims[0] = makeSphere(-1.8575f, -0.98714f, -3.6f, 0.6f, scale(v3(1.f, .89f, .55f), gain), DIFFUSE);
    prims[1] = makeSphere(0.74795f, -0.55f, -4.3816f, 1.05f, v3(1.f, 1.f, 1.f), SPECULAR);
    prims[2] = makeSphere(1.9929f, -1.08666f, (float)-3.23, 0.5f, v3(0.75f, 0.75f, 0.75f), REFRACTIVE);
    prims[3] = makeSphere((float)-0.19931, -1.183f, -2.75f, 0.4f, scale(v3(.8f, .06f, .391f), gain), DIFFUSE);
    prims[4] = makeSphere((float)-0.19931, -1.183f, -2.75f, 0.4001f, v3(1.f, 1.f, 1.f), REFRACTIVE);
    prims[5] = makeDisc(v3(0.f, 1.f, 0.f), v3(0.f, -1.6f, -5.22f), 3.5f, scale(v3(.98f, .76f, .66f), gain), DIFFUSE);
    built = true;
  }
  return prims;
}

// light.hpp:131-147
static float hitSphere(const Prim& p, V3 o, V3 d) {
  const V3 f = sub(p.centre, o);
  const float tca = dot(f, d);
  if (tca < 0.f) return 0.f;
  const V3 l = sub(p.centre, add(o, scale(d, tca)));
  const float l2 = norm2(l);
  if (l2 > p.radius2) return 0.f;
  const float thc = sqrtf(p.radius2 - l2);
  float t0 = tca - thc, t1 = tca + thc;
  if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
  if (t0 < 0) { t0 = t1; if (t0 < 0) return 0.f; }
  return t0;
}

// light.hpp:106-120
static float hitDisc(const Prim& p, V3 o, V3 d) {
  const float angle = dot(p.normal, d);
  if (angle != 0.f) {
    const float t = -(dot(p.normal, o) + p.planeD) / angle;
    if (t > kEps) {
      const V3 hp = add(o, scale(d, t));
      if (norm2(sub(hp, p.centre)) < p.radius2) return t;
    }
  }
  return 0.f;
}

// light.hpp:173-191: closest hit with t > intersectionEpsilon; advances origin in place.
static int closestHit(V3& o, V3 d, V3& normalOut) {
  const Prim* prims = sceneTable();
  int best = -1;
  float bestT = std::numeric_limits<float>::infinity();
  const int count = sceneCount();
  for (int i = 0; i < count; ++i) {
    const float t = prims[i].shape == SPHERE ? hitSphere(prims[i], o, d) : hitDisc(prims[i], o, d);
    if (t > kHitEps && t < bestT) { bestT = t; best = i; }
  }
  if (best >= 0) {
    o = add(o, scale(d, bestT));   // ray.origin += ray.direction * t
    normalOut = prims[best].shape == SPHERE ? unit(sub(o, prims[best].centre)) : prims[best].normal;
  }
  return best;
}

// light.hpp:230-250 (first two axes only; third is the normal itself)
static void tangentFrame(V3 n, V3& ax, V3& ay) {
  const float nx2 = n.x * n.x, ny2 = n.y * n.y, nz2 = n.z * n.z;
  if (fabsf(n.x) > fabsf(n.y)) {
    const float inv = 1.f / std::sqrt(nx2 + nz2);
    ax = v3(-n.z * inv, 0.f, n.x * inv);
  } else {
    const float inv = 1.f / std::sqrt(ny2 + nz2);
    ax = v3(0.f, n.z * inv, -n.y * inv);
  }
  ay = cross(n, ax);
}

// sdf.hpp:8-31 with hemisphere() light.hpp:223-228. Returns the record weight.
static float sampleDiffuse(V3& dir, V3 n, float rr, float u1, float u2) {
  V3 ax, ay;
  tangentFrame(n, ax, ay);
  const float r = sqrtf(1.f - u1 * u1);
  const float phi = 2.f * kPi * u2;
  const V3 h = v3(cosf(phi) * r, sinf(phi) * r, u1);
  dir = v3(dot(v3(ax.x, ay.x, n.x), h), dot(v3(ax.y, ay.y, n.y), h), dot(v3(ax.z, ay.z, n.z), h));
  return dot(dir, n) * .1f * rr;
}

// sdf.hpp:36-40
static void sampleMirror(V3& dir, V3 n) {
  const float c = dot(dir, n);
  dir = unit(sub(dir, scale(n, c * 2.f)));
}

// sdf.hpp:45-67; returns true when the ray was transmitted
static bool sampleGlass(V3& dir, V3 n, float ri, float rnd) {
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

static inline void putRecord(ContribRec& r, V3 c, float w, std::uint8_t type) {
  std::memset(&r, 0, sizeof(r));
  r.clr[0] = c.x; r.clr[1] = c.y; r.clr[2] = c.z; r.weight = w; r.type = type;
}

// codelets.cpp:159-223 for one camera ray. `draw` yields the next primary sample (already
// binary16-valued). Writes up to maxLen records, returns how many were written.
template <class Draw>
static int tracePathRecords(float camX, float camY, const Settings& s, Draw&& draw, ContribRec* rec) {
  const Prim* prims = sceneTable();
  V3 o = v3(0.f, 0.f, 0.f);
  V3 d = unit(v3(camX, camY, -1.f));       // Ray ctor normalises (light.hpp:35)
  std::uint32_t depth = 0;
  int count = 0;
  const int maxLen = (int)s.maxPathLength;
  bool lit = false;
  while (count != maxLen) {
    float rr = 1.f;
    if (depth >= s.rouletteDepth) {         // light.hpp:25-29
      const float rnd = draw();
      if (rnd <= s.stopProb) break;
      rr = 1.f / (1.f - s.stopProb);
    }
    V3 n;
    const int hit = closestHit(o, d, n);
    if (hit < 0) {
      putRecord(rec[count++], d, rr, C_ESCAPED);
      lit = true;
      break;
    }
    const Prim& p = prims[hit];
    if (p.emissive) {                       // codelets.cpp:192-196
      putRecord(rec[count++], p.emission, rr, C_EMIT);
      lit = true;
      break;
    }
    if (p.bsdf == DIFFUSE) {
      const float u1 = draw();
      const float u2 = draw();
      const float w = sampleDiffuse(d, n, rr, u1, u2);
      putRecord(rec[count++], p.colour, w, C_DIFFUSE);
    } else if (p.bsdf == SPECULAR) {
      sampleMirror(d, n);
      putRecord(rec[count++], v3(0.f, 0.f, 0.f), rr, C_SPECULAR);
    } else {
      const bool through = sampleGlass(d, n, s.refractiveIndex, draw());
      putRecord(rec[count++], through ? p.colour : v3(1.f, 1.f, 1.f), 1.15f * rr, C_REFRACT);
    }
    depth += 1;
  }
  if (!lit) {
    // codelets.cpp:219-222: the LAST record is overwritten by END. With rouletteDepth==0 a
    // depth-0 stop leaves the stack empty and the reference writes to index -1 (SURVEY §8a
    // quirk ii); the oracle refuses that configuration instead of reproducing the overrun.
    if (count == 0) { count = 1; }
    putRecord(rec[count - 1], v3(0.f, 0.f, 0.f), 0.f, C_END);
  }
  return count;
}

// WrappedArray.hpp:45-68: number of records up to and including the terminator.
static int recordCount(const ContribRec* rec, bool& contributes) {
  int n = 0;
  contributes = false;
  for (;;) {
    const std::uint8_t t = rec[n++].type;
    if (t == C_EMIT || t == C_DEBUG || t == C_ESCAPED) { contributes = true; break; }
    if (t == C_END) break;
  }
  return n;
}

// codelets.cpp:255-298 backward walk. Returns the path's RGB.
static V3 walkBack(const ContribRec* rec, int n) {
  V3 total = v3(0.f, 0.f, 0.f);
  for (int i = n - 1; i >= 0; --i) {
    const ContribRec& c = rec[i];
    const V3 clr = v3(c.clr[0], c.clr[1], c.clr[2]);
    switch (c.type) {
      case C_DIFFUSE:
      case C_REFRACT: total = scale(mul(total, clr), c.weight); break;
      case C_EMIT:
      case C_ESCAPED: total = add(total, scale(clr, c.weight)); break;
      case C_SPECULAR: total = scale(total, c.weight); break;
      case C_DEBUG: return clr;
      default: break;
    }
  }
  return total;
}

// codelets.cpp:333-347
static void escapedUv(V3 d, float azimuth, float& u, float& v) {
  // `atan2` is called unqualified on floats in the reference (codelets.cpp:334); with
  // glibc/libstdc++ that resolves to the double overload, so phi and the wrap run in double
  // and v is rounded to float once at the end. `auto` keeps whatever the toolchain picks.
  // Reference-exact (codelets.cpp:333): |d.y| can be 1 + 1 ulp after normalisation, for which acosf - and hence the
  // pixel, for ever - is NaN in the reference, and so it is here. The CUDA path deliberately clamps d.y to [-1, 1]
  // (pt_device.cuh escapedUv); tests assert that difference (tests/test_gpu_kat.py::test_escaped_uv) instead of
  // bending the oracle to the product.
  const auto theta = acosf(d.y);
  auto phi = atan2(d.z, d.x) + azimuth;
  constexpr auto twoPi = 2.f * kPi;
  constexpr auto invPi = 1.f / kPi;
  constexpr auto inv2Pi = 1.f / twoPi;
  if (phi < 0.f) { phi += twoPi; } else if (phi > twoPi) { phi -= twoPi; }
  u = theta * invPi;
  v = phi * inv2Pi;
}

// light.hpp:194-206 + codelets.cpp:68-75: pixel (+ scaled AA noise) -> camera-plane x,y,
// stored as binary16. `noise*` are binary16-valued; the product with aaScale is rounded
// to binary16 before it is added (half * half, codelets.cpp:71-72).
static void cameraRay(std::uint32_t pu, std::uint32_t pv, float noise0, float noise1, const Settings& s, float& camX, float& camY) {
  const float c = (float)pu + roundToHalf(s.aaScale * noise0);
  const float r = (float)pv + roundToHalf(s.aaScale * noise1);
  const float w = (float)s.width, h = (float)s.height;
  const float aspect = w / h;
  const float tanTheta = tanf(s.fov / 2.f);
  const float x = (c / w) - .5f;
  const float y = (r / h) - .5f;
  camX = roundToHalf(2.f * x * aspect * tanTheta);
  camY = roundToHalf(-2.f * y * tanTheta);
}

} // namespace orc

using namespace orc;

extern "C" {

int orc_sizeof_settings() { return (int)sizeof(Settings); }

// --- stage-level entry points (same buffers as the reference vertices) ---------------

// Camera rays for n work items. noise: 2 binary16 values per item (item i uses 2i, 2i+1).
void orc_camera_rays(int n, const TraceRec* items, const std::uint16_t* noise, const Settings* s, std::uint16_t* rays) {
  for (int i = 0; i < n; ++i) {
    float x, y;
    cameraRay(items[i].u, items[i].v, halfBitsToFloat(noise[2 * i]), halfBitsToFloat(noise[2 * i + 1]), *s, x, y);
    rays[2 * i] = floatToHalfBits(x);
    rays[2 * i + 1] = floatToHalfBits(y);
  }
}

// Path records for n camera rays; path i reads uniforms[i*perPath ...] sequentially.
void orc_trace_records(int n, const std::uint16_t* rays, const std::uint16_t* uniforms, int perPath, const Settings* s, ContribRec* records) {
  for (int i = 0; i < n; ++i) {
    int next = 0;
    const std::uint16_t* mine = uniforms + (std::size_t)i * perPath;
    auto draw = [&]() { float v = halfBitsToFloat(mine[next]); next += 1; if (next == perPath) next = 0; return v; };
    ContribRec* row = records + (std::size_t)i * s->maxPathLength;
    std::memset(row, 0, sizeof(ContribRec) * s->maxPathLength);
    tracePathRecords(halfBitsToFloat(rays[2 * i]), halfBitsToFloat(rays[2 * i + 1]), *s, draw, row);
  }
}

// u,v for the terminal ESCAPED record of each path, (0,0) otherwise (codelets.cpp:348-352)
void orc_escaped_uv(int n, const ContribRec* records, int maxPathLength, float azimuth, float* u, float* v) {
  for (int i = 0; i < n; ++i) {
    const ContribRec* row = records + (std::size_t)i * maxPathLength;
    bool contributes;
    const int cnt = recordCount(row, contributes);
    const ContribRec& last = row[cnt - 1];
    if (last.type == C_ESCAPED) {
      escapedUv(v3(last.clr[0], last.clr[1], last.clr[2]), azimuth, u[i], v[i]);
    } else { u[i] = 0.f; v[i] = 0.f; }
  }
}

// Apply env light (bgr per path, codelets.cpp:375-379) and accumulate into TraceRecords
// (codelets.cpp:249-301). records are modified in place like the reference does.
void orc_light_and_accumulate(int n, ContribRec* records, int maxPathLength, const float* bgr, TraceRec* items) {
  for (int i = 0; i < n; ++i) {
    ContribRec* row = records + (std::size_t)i * maxPathLength;
    bool contributes;
    const int cnt = recordCount(row, contributes);
    ContribRec& last = row[cnt - 1];
    if (last.type == C_ESCAPED) {
      last.clr[0] = bgr[3 * i + 2]; last.clr[1] = bgr[3 * i + 1]; last.clr[2] = bgr[3 * i];
    }
    items[i].pathLength = (std::uint16_t)(items[i].pathLength + cnt);
    if (contributes) {
      const V3 t = walkBack(row, cnt);
      items[i].r += t.x; items[i].g += t.y; items[i].b += t.z;
    }
    items[i].sampleCount = (std::uint16_t)(items[i].sampleCount + 1);
  }
}

// Forward view of a finished path, the form the CUDA wavefront queue carries:
// escaped flag, unit direction, RGB throughput T such that pixel += T (*) L_env, length.
void orc_path_summary(int n, const ContribRec* records, int maxPathLength, int* escaped, float* dir3, float* throughput3, int* length) {
  for (int i = 0; i < n; ++i) {
    const ContribRec* row = records + (std::size_t)i * maxPathLength;
    bool contributes;
    const int cnt = recordCount(row, contributes);
    length[i] = cnt;
    const ContribRec& last = row[cnt - 1];
    escaped[i] = last.type == C_ESCAPED ? 1 : 0;
    dir3[3 * i] = dir3[3 * i + 1] = dir3[3 * i + 2] = 0.f;
    throughput3[3 * i] = throughput3[3 * i + 1] = throughput3[3 * i + 2] = 0.f;
    if (escaped[i]) {
      dir3[3 * i] = last.clr[0]; dir3[3 * i + 1] = last.clr[1]; dir3[3 * i + 2] = last.clr[2];
      std::vector<ContribRec> tmp(row, row + cnt);
      tmp[cnt - 1].clr[0] = tmp[cnt - 1].clr[1] = tmp[cnt - 1].clr[2] = 1.f;   // unit radiance
      const V3 t = walkBack(tmp.data(), cnt);
      throughput3[3 * i] = t.x; throughput3[3 * i + 1] = t.y; throughput3[3 * i + 2] = t.z;
    }
  }
}

// --- Philox-driven pipeline stage: everything before the NIF ----------------------------
// For work items `items[0..n)` and sample index `sample`, generates the camera ray and the
// path with the shared Philox schedule (oracle_common.hpp PathRng) and returns the forward
// summary plus the env-map lookup coordinates. Pad items (u or v out of range,
// LoadBalancer.cpp:66-71) are reported as length 0, not escaped.
void orc_trace_philox(int n, const TraceRec* items, std::uint32_t sample, const Settings* s,
                      int* escaped, float* uv2, float* throughput3, int* length, std::uint16_t* camRays) {
#pragma omp parallel for schedule(dynamic, 1024)
  for (int i = 0; i < n; ++i) {
    escaped[i] = 0; length[i] = 0;
    uv2[2 * i] = uv2[2 * i + 1] = 0.f;
    throughput3[3 * i] = throughput3[3 * i + 1] = throughput3[3 * i + 2] = 0.f;
    if (camRays) { camRays[2 * i] = camRays[2 * i + 1] = 0; }
    if (items[i].u >= s->width || items[i].v >= s->height) continue;
    PathRng rng(s->seed, (std::uint32_t)items[i].v * s->width + items[i].u, sample);
    const std::uint32_t w0 = rng.word(), w1 = rng.word();
    float n0, n1;
    aaNoisePair(w0, w1, s->aaNoiseType, n0, n1);
    float cx, cy;
    cameraRay(items[i].u, items[i].v, n0, n1, *s, cx, cy);
    if (camRays) { camRays[2 * i] = floatToHalfBits(cx); camRays[2 * i + 1] = floatToHalfBits(cy); }
    ContribRec rec[64];
    auto draw = [&]() { return rng.uniformHalf(); };
    const int cnt = tracePathRecords(cx, cy, *s, draw, rec);
    length[i] = cnt;
    if (rec[cnt - 1].type == C_ESCAPED) {
      escaped[i] = 1;
      escapedUv(v3(rec[cnt - 1].clr[0], rec[cnt - 1].clr[1], rec[cnt - 1].clr[2]), s->azimuth, uv2[2 * i], uv2[2 * i + 1]);
      rec[cnt - 1].clr[0] = rec[cnt - 1].clr[1] = rec[cnt - 1].clr[2] = 1.f;
      const V3 t = walkBack(rec, cnt);
      throughput3[3 * i] = t.x; throughput3[3 * i + 1] = t.y; throughput3[3 * i + 2] = t.z;
    } else if (rec[cnt - 1].type == C_EMIT) {
      // the path ended on an emitter: escaped = 2, throughput3 = the radiance the backward walk yields (no NIF lookup)
      escaped[i] = 2;
      const V3 t = walkBack(rec, cnt);
      throughput3[3 * i] = t.x; throughput3[3 * i + 1] = t.y; throughput3[3 * i + 2] = t.z;
    }
  }
}

// Scene as data: n primitives {shape (0 sphere, 1 disc), bsdf (0 diffuse, 1 specular, 2 refractive), centre[3], normal[3]
// (disc; normalised here like light::Disc's constructor), radius, colour[3], emission[3]} as 15 floats each, emissive iff
// any emission component is non-zero (light.hpp Material). n = 0 restores the built-in scene of codelets.cpp:110-144.
void orc_set_scene(int n, const float* prims15) {
  g_userScene.clear();
  for (int i = 0; i < n; ++i) {
    const float* a = prims15 + 15 * (std::size_t)i;
    const V3 c = v3(a[2], a[3], a[4]), nn = v3(a[5], a[6], a[7]), col = v3(a[9], a[10], a[11]), em = v3(a[12], a[13], a[14]);
    Prim p = (int)a[0] == SPHERE ? makeSphere(c.x, c.y, c.z, a[8], col, (int)a[1]) : makeDisc(nn, c, a[8], col, (int)a[1]);
    p.emissive = (em.x != 0.f || em.y != 0.f || em.z != 0.f) ? 1 : 0;
    p.emission = em;
    g_userScene.push_back(p);
  }
}

int orc_num_threads() { return omp_get_max_threads(); }
// bench.py under torch.distributed.run inherits OMP_NUM_THREADS=1; the CPU arm asks for every host thread.
void orc_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_num_procs() { return omp_get_num_procs(); }

// Raw generator access so tests can compare the CUDA Philox / noise code word for word.
void orc_philox(std::uint32_t c0, std::uint32_t c1, std::uint32_t c2, std::uint32_t c3, std::uint32_t k0, std::uint32_t k1, std::uint32_t* out4) {
  const Philox4 r = philox4x32_10(c0, c1, c2, c3, k0, k1);
  for (int i = 0; i < 4; ++i) out4[i] = r.v[i];
}
void orc_aa_noise(int n, const std::uint32_t* words2, int type, std::uint16_t* out2) {
  for (int i = 0; i < n; ++i) {
    float a, b;
    aaNoisePair(words2[2 * i], words2[2 * i + 1], type, a, b);
    out2[2 * i] = floatToHalfBits(a); out2[2 * i + 1] = floatToHalfBits(b);
  }
}
void orc_float_to_half(int n, const float* in, std::uint16_t* out) { for (int i = 0; i < n; ++i) out[i] = floatToHalfBits(in[i]); }
void orc_half_to_float(int n, const std::uint16_t* in, float* out) { for (int i = 0; i < n; ++i) out[i] = halfBitsToFloat(in[i]); }

} // extern "C"
