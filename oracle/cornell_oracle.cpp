// TEST INFRASTRUCTURE ONLY - never linked into or loaded by the product.
//
// CPU restatement of the reference's stand-alone Cornell-box renderer (external/light/src): scene trace.cpp:168-216,
// path loop trace.hpp:8-107, primitives light.hpp:76-151, Scene::intersect :173-191, BSDFs sdf.hpp:8-67, random streams
// xoshiro.hpp + jobs.hpp:66-90, pixel loop trace.cpp:246-262. Pinned bit for bit against the compiled reference
// (oracle/_ref/libref_cornell.so) and the golden vectors in tests/test_cornell.py. Plain scalar code, same operation
// order as the reference; built without FMA contraction (oracle/Makefile).
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

struct V { float x, y, z; };                                              // vector.hpp:17-58 (non-Eigen build)
inline V v(float x, float y, float z) { return V{x, y, z}; }
inline V operator+(V a, V b) { return v(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V operator-(V a, V b) { return v(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V operator*(V a, float s) { return v(a.x * s, a.y * s, a.z * s); }
inline V operator/(V a, float s) { return v(a.x / s, a.y / s, a.z / s); }
inline V cw(V a, V b) { return v(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float sq(V a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
inline V unit(V a) { return a * (1.f / sqrtf(sq(a))); }                  // vector.hpp:47
inline V cross(V a, V b) { return v(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
constexpr float kPi = 3.14159265358979323846264338327950288f;            // light.hpp:16

// ---- xoshiro.hpp ------------------------------------------------------------------------------------------
struct Rng { uint64_t s0, s1; };
inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
inline uint64_t splitmix(uint64_t z) {                                   // :15-20
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
inline uint64_t next(Rng& r) {                                           // :28-38
  const uint64_t a = r.s0;
  uint64_t b = r.s1;
  const uint64_t out = rotl(a * 5, 7) * 9;
  b ^= a;
  r.s0 = rotl(a, 24) ^ b ^ (b << 16);
  r.s1 = rotl(b, 37);
  return out;
}
inline void jump(Rng& r) {                                               // :43-58
  static const uint64_t kJump[2] = {0xdf900294d8f554a5ull, 0x170865df4b3201fcull};
  uint64_t a = 0, b = 0;
  for (int w = 0; w < 2; ++w)
    for (int bit = 0; bit < 64; ++bit) {
      if (kJump[w] & (1ull << bit)) { a ^= r.s0; b ^= r.s1; }
      next(r);
    }
  r.s0 = a; r.s1 = b;
}
inline double unitDouble(uint64_t x) { const uint64_t u = (0x3FFull << 52) | (x >> 12); double d; std::memcpy(&d, &u, 8); return d - 1.0; }   // :60-63
inline float uniform01(Rng& r) { return (float)unitDouble(next(r)); }                     // :71-73
inline float uniformNeg1To1(Rng& r) { return (float)(2.0 * unitDouble(next(r)) - 1.0); }  // :66-68

// ---- scene ------------------------------------------------------------------------------------------------
struct Object {
  int shape;             // 0 sphere, 1 disc, 2 plane
  int bsdf;              // 0 diffuse, 1 specular, 2 refractive
  V c, n;                // sphere/disc centre; plane/disc unit normal
  float r2, d;           // radius^2; plane offset (Disc: |c.n|, light.hpp:101)
  V colour, emission;
  bool emissive;
};
struct Params { float ri, stopProb, eps, hitEps; uint32_t rouletteDepth; };

std::vector<Object> buildScene() {                                      // trace.cpp:168-216
  std::vector<Object> s;
  const V zero = v(0, 0, 0);
  auto material = [&](Object& o, V col, V em, int bsdf) { o.colour = col; o.emission = em; o.bsdf = bsdf; o.emissive = em.x != 0.f || em.y != 0.f || em.z != 0.f; };
  auto sphere = [&](V c, float r, V col, V em, int bsdf) { Object o{}; o.shape = 0; o.c = c; o.n = zero; o.r2 = r * r; o.d = 0.f; material(o, col, em, bsdf); s.push_back(o); };
  auto plane = [&](V n, float d, V col) { Object o{}; o.shape = 2; o.c = zero; o.n = unit(n); o.r2 = 0.f; o.d = d; material(o, col, zero, 0); s.push_back(o); };
  const float gain = 15.f;
  const V wall1 = v(.98f, .76f, .66f) * gain, wall2 = v(.93f, .43f, .48f) * gain, wall3 = v(.27f, .31f, .38f) * gain;
  sphere(v(-0.75f, -1.45f, -4.4f), 1.05f, v(4.f, 8.f, 4.f), zero, 1);
  sphere(v(2.0f, -2.05f, -3.7f), 0.5f, v(10.f, 10.f, 1.f), zero, 2);
  sphere(v(-1.75f, -1.95f, -3.1f), 0.6f, v(1.f, .89f, .55f) * gain, zero, 0);
  sphere(v(-1.12f, -2.3f, -3.5f), 0.2f, zero, v(500, 600, 1000), 1);
  sphere(v(-0.28f, -2.34f, -3.f), 0.2f, zero, v(467, 1000, 434), 1);
  sphere(v(.58f, -2.39f, -2.6f), 0.2f, zero, v(1000, 367, 367), 1);
  const V X = v(1, 0, 0), Y = v(0, 1, 0), Z = v(0, 0, 1);
  auto neg = [](V a) { return v(-a.x, -a.y, -a.z); };
  plane(Y, 2.5f, wall1); plane(Z, 5.5f, wall1); plane(X, 2.75f, wall2);
  plane(neg(X), 2.75f, wall3); plane(neg(Y), 3.f, wall1); plane(neg(Z), .5f, wall1);
  Object disc{};
  disc.shape = 1; disc.n = unit(neg(Y)); disc.c = v(0.f, 2.9999f, -4.f); disc.d = std::fabs(dot(disc.c, disc.n)); disc.r2 = .7f * .7f;
  material(disc, zero, v(8000, 8000, 8000), 0);
  s.push_back(disc);
  return s;
}

float intersect(const Object& o, V org, V dir, float eps) {
  if (o.shape == 0) {                                                    // light.hpp:131-147
    const V f = o.c - org;
    const float tca = dot(f, dir);
    if (tca < 0.f) return 0.f;
    const V l = o.c - (org + dir * tca);
    const float l2 = sq(l);
    if (l2 > o.r2) return 0.f;
    const float thc = sqrtf(o.r2 - l2);
    float t0 = tca - thc, t1 = tca + thc;
    if (t0 > t1) { const float t = t0; t0 = t1; t1 = t; }
    if (t0 < 0) { t0 = t1; if (t0 < 0) return 0.f; }
    return t0;
  }
  const float angle = dot(o.n, dir);                                     // Plane :84-92, Disc :106-120
  if (angle != 0.f) {
    const float t = -(dot(o.n, org) + o.d) / angle;
    if (o.shape == 2) return t > eps ? t : 0.f;
    if (t > eps && sq((org + dir * t) - o.c) < o.r2) return t;
  }
  return 0.f;
}

// trace.hpp:8-107
V tracePath(const std::vector<Object>& scene, const Params& p, V camDir, Rng& rng, uint32_t& depthOut) {
  struct Contribution { V clr; float weight; int type; };               // 0 DIFFUSE, 1 EMIT, 2 SPECULAR, 3 REFRACT
  std::vector<Contribution> stack;
  V org = v(0, 0, 0), dir = unit(camDir);                                // Ray ctor normalises (light.hpp:35)
  uint32_t depth = 0;
  bool hitEmitter = false;
  while (depth < 20) {
    float rr = 1.f;
    if (depth >= p.rouletteDepth) {                                      // light.hpp:25-29
      const float rnd = uniform01(rng);
      if (rnd <= p.stopProb) break;
      rr = 1.f / (1.f - p.stopProb);
    }
    const Object* best = nullptr;                                        // Scene::intersect light.hpp:173-191
    float bestT = INFINITY;
    for (const Object& o : scene) {
      const float t = intersect(o, org, dir, p.eps);
      if (t > p.hitEps && t < bestT) { bestT = t; best = &o; }
    }
    if (!best) break;
    org = org + dir * bestT;
    const V n = best->shape == 0 ? unit(org - best->c) : best->n;
    if (best->emissive) { stack.push_back({best->emission, rr, 1}); hitEmitter = true; break; }
    if (best->bsdf == 0) {                                               // sdf.hpp:8-31, hemisphere light.hpp:223-228, frame :230-250
      const float u1 = uniform01(rng), u2 = uniform01(rng);
      V ax;
      if (std::fabs(n.x) > std::fabs(n.y)) { const float inv = 1.f / std::sqrt(n.x * n.x + n.z * n.z); ax = v(-n.z * inv, 0.f, n.x * inv); }
      else { const float inv = 1.f / std::sqrt(n.y * n.y + n.z * n.z); ax = v(0.f, n.z * inv, -n.y * inv); }
      const V ay = cross(n, ax);
      const float r = sqrtf(1.f - u1 * u1), phi = 2.f * kPi * u2;
      const V h = v(cosf(phi) * r, sinf(phi) * r, u1);
      dir = v(dot(v(ax.x, ay.x, n.x), h), dot(v(ax.y, ay.y, n.y), h), dot(v(ax.z, ay.z, n.z), h));
      stack.push_back({best->colour, dot(dir, n) * .1f * rr, 0});
    } else if (best->bsdf == 1) {                                        // sdf.hpp:36-40
      const float c = dot(dir, n);
      dir = unit(dir - n * (c * 2.f));
      stack.push_back({v(0, 0, 0), rr, 2});
    } else {                                                             // sdf.hpp:45-67
      const float rnd = uniform01(rng);
      float ri = p.ri;
      V nn = n;
      float r0 = (1.f - ri) / (1.f + ri);
      r0 = r0 * r0;
      if (dot(nn, dir) > 0.f) nn = v(-nn.x, -nn.y, -nn.z); else ri = 1.f / ri;
      const float c1 = -dot(nn, dir);
      const float c2 = 1.f - ri * ri * (1.f - c1 * c1);
      const float sb = 1.f - c1, sb2 = sb * sb;
      const float rprob = r0 + (1.f - r0) * (sb2 * sb * sb2);
      if (c2 > 0.f && rnd > rprob) dir = unit((dir * ri) + (nn * (ri * c1 - sqrtf(c2))));
      else dir = unit(dir + nn * (c1 * 2.f));
      stack.push_back({v(0, 0, 0), 1.15f * rr, 3});
    }
    depth += 1;
  }
  depthOut = depth;
  V total = v(0, 0, 0);
  if (hitEmitter)
    for (int i = (int)stack.size() - 1; i >= 0; --i) {                   // :76-104
      const Contribution& c = stack[i];
      if (c.type == 0) total = cw(total, c.clr) * c.weight;
      else if (c.type == 1) total = total + c.clr * c.weight;
      else total = total * c.weight;
    }
  return total;
}

}  // namespace

extern "C" {

void orc_cornell_trace_paths(size_t n, const float* camXY, uint64_t* rngState, float refractiveIndex, uint32_t rouletteDepth, float stopProb,
                             float epsilon, float intersectionEpsilon, float* rgb, uint32_t* depth) {
  const std::vector<Object> scene = buildScene();
  const Params p{refractiveIndex, stopProb, epsilon, intersectionEpsilon, rouletteDepth};
#pragma omp parallel for schedule(static)
  for (long i = 0; i < (long)n; ++i) {
    Rng rng{rngState[2 * i], rngState[2 * i + 1]};
    const V c = tracePath(scene, p, v(camXY[2 * i], camXY[2 * i + 1], -1.f), rng, depth[i]);
    rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    rngState[2 * i] = rng.s0; rngState[2 * i + 1] = rng.s1;
  }
}

// trace.cpp:246-262 with createTracingJobs (jobs.hpp:66-90). The two jitter draws are arguments of one constructor call in
// the reference; g++ evaluates them right to left, so the first draw is the y jitter (pinned by the image comparison).
void orc_cornell_render(uint32_t width, uint32_t height, uint32_t tileWidth, uint32_t tileHeight, uint32_t spp, uint64_t seed, float aaScale,
                        float refractiveIndex, uint32_t rouletteDepth, float stopProb, float epsilon, float intersectionEpsilon, float* rgb,
                        uint64_t* stats) {
  const std::vector<Object> scene = buildScene();
  const Params p{refractiveIndex, stopProb, epsilon, intersectionEpsilon, rouletteDepth};
  struct Job { uint32_t r0, c0, r1, c1; Rng rng; uint64_t casts, longest; };
  std::vector<Job> jobs;
  Rng state;
  state.s0 = splitmix(seed); state.s1 = splitmix(state.s0);             // xoshiro.hpp:23-26
  for (uint32_t r = 0; r < height; r += tileHeight)
    for (uint32_t c = 0; c < width; c += tileWidth) {
      jobs.push_back({r, c, std::min(r + tileHeight, height), std::min(c + tileWidth, width), state, 0, 0});
      jump(state);
    }
  std::memset(rgb, 0, sizeof(float) * 3 * width * height);
  const float w = (float)width, h = (float)height, aspect = w / h, tanTheta = tanf((kPi / 2.f) / 2.f);   // light.hpp:194-206, fov = Pi/2
#pragma omp parallel for schedule(dynamic)
  for (long j = 0; j < (long)jobs.size(); ++j) {
    Job& job = jobs[j];
    for (uint32_t s = 0; s < spp; ++s)
      for (uint32_t row = job.r0; row < job.r1; ++row)
        for (uint32_t col = job.c0; col < job.c1; ++col) {
          const float x = ((float)col / w) - .5f, y = ((float)row / h) - .5f;
          V cam = v(2.f * x * aspect * tanTheta, -2.f * y * tanTheta, -1.f);
          const float ny = uniformNeg1To1(job.rng);
          const float nx = uniformNeg1To1(job.rng);
          cam = cam + v(nx, ny, 0.f) * aaScale;
          uint32_t depth;
          const V colour = tracePath(scene, p, cam, job.rng, depth);
          float* px = rgb + ((size_t)row * width + col) * 3;
          const V add = colour / (float)spp;
          px[0] += add.x; px[1] += add.y; px[2] += add.z;
          job.casts += depth;
          job.longest = std::max<uint64_t>(job.longest, depth);
        }
  }
  stats[0] = stats[1] = 0;
  for (const Job& job : jobs) { stats[0] += job.casts; stats[1] = std::max(stats[1], job.longest); }
}

}  // extern "C"
