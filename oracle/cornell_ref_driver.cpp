// TEST INFRASTRUCTURE ONLY - never linked into or loaded by the product.
//
// Compiles the UNMODIFIED reference CPU path tracer of the Cornell-box scene (external/light/src/trace.hpp
// with light.hpp, sdf.hpp, jobs.hpp, xoshiro.hpp, resolved through -I/root/reference/external/light/src) and
// exposes it to the tests. Only the scene table (data of external/light/src/trace.cpp:168-216, whose main() is
// an OpenCV application and cannot be built here) and the pixel loop of :251-262 are restated.
#include <cfloat>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <functional>
#include <iostream>
#include <vector>

#include <trace.hpp>

namespace light {
// off-IPU these are extern (light.hpp:17-23); run-time options of trace.cpp:32-33,222-223
float epsilon = FLT_EPSILON;
float intersectionEpsilon = 1e-5f;
}

namespace {
using Vec = light::Vector;

struct Cornell {
  // trace.cpp:168-216
  light::Sphere spheres[6] = {light::Sphere(Vec(-0.75f, -1.45f, -4.4f), 1.05f), light::Sphere(Vec(2.0f, -2.05f, -3.7f), 0.5f),
                              light::Sphere(Vec(-1.75f, -1.95f, -3.1f), 0.6f),  light::Sphere(Vec(-1.12f, -2.3f, -3.5f), 0.2f),
                              light::Sphere(Vec(-0.28f, -2.34f, -3.f), 0.2f),   light::Sphere(Vec(.58f, -2.39f, -2.6f), 0.2f)};
  light::Plane planes[6] = {light::Plane(Vec(0, 1, 0), 2.5f),  light::Plane(Vec(0, 0, 1), 5.5f),  light::Plane(Vec(1, 0, 0), 2.75f),
                            light::Plane(Vec(-1, 0, 0), 2.75f), light::Plane(Vec(0, -1, 0), 3.f), light::Plane(Vec(0, 0, -1), .5f)};
  light::Disc discs[1] = {light::Disc(Vec(0, -1, 0), Vec(0.f, 2.9999f, -4.f), .7f)};
  light::Scene<13> scene;
  static std::array<light::Object, 13> objects(Cornell& c) {
    const Vec zero(0.f, 0.f, 0.f);
    const float gain = 15.f;
    const Vec wall1 = Vec(.98f, .76f, .66f) * gain, wall2 = Vec(.93f, .43f, .48f) * gain, wall3 = Vec(.27f, .31f, .38f) * gain;
    using T = light::Material::Type;
    return {light::Object(&c.spheres[0], Vec(4.f, 8.f, 4.f), zero, T::specular),
            light::Object(&c.spheres[1], Vec(10.f, 10.f, 1.f), zero, T::refractive),
            light::Object(&c.spheres[2], Vec(1.f, .89f, .55f) * gain, zero, T::diffuse),
            light::Object(&c.spheres[3], zero, Vec(500, 600, 1000), T::specular),
            light::Object(&c.spheres[4], zero, Vec(467, 1000, 434), T::specular),
            light::Object(&c.spheres[5], zero, Vec(1000, 367, 367), T::specular),
            light::Object(&c.planes[0], wall1, zero, T::diffuse),
            light::Object(&c.planes[1], wall1, zero, T::diffuse),
            light::Object(&c.planes[2], wall2, zero, T::diffuse),
            light::Object(&c.planes[3], wall3, zero, T::diffuse),
            light::Object(&c.planes[4], wall1, zero, T::diffuse),
            light::Object(&c.planes[5], wall1, zero, T::diffuse),
            light::Object(&c.discs[0], zero, Vec(8000, 8000, 8000), T::diffuse)};
  }
  Cornell() : scene(objects(*this)) {}
};
}  // namespace

extern "C" {

// One path per (camera-plane x,y ; xoshiro state): trace.hpp:8-107 for Ray(0, (x, y, -1)).
// depth[i] = ray casts of the path (what the job adds to totalRayCasts); state is advanced in place.
void ref_cornell_trace_paths(size_t n, const float* camXY, uint64_t* rngState, float refractiveIndex, uint32_t rouletteDepth, float stopProb,
                             float epsilon, float intersectionEpsilon, float* rgb, uint32_t* depth) {
  light::epsilon = epsilon;
  light::intersectionEpsilon = intersectionEpsilon;
  Cornell cornell;
  light::RayTracerContext<decltype(cornell.scene)> tracer(cornell.scene);
  tracer.refractiveIndex = refractiveIndex;
  tracer.rouletteDepth = rouletteDepth;
  tracer.stopProb = stopProb;
  TraceTileJob job(0, 0, 1, 1, 1, 1, 1);
  for (size_t i = 0; i < n; ++i) {
    job.rngState = {rngState[2 * i], rngState[2 * i + 1]};
    job.totalRayCasts = 0;
    const light::Ray ray(Vec(0.f, 0.f, 0.f), Vec(camXY[2 * i], camXY[2 * i + 1], -1.f));
    const Vec c = trace(ray, tracer, job);
    rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    depth[i] = (uint32_t)job.totalRayCasts;
    rngState[2 * i] = job.rngState[0]; rngState[2 * i + 1] = job.rngState[1];
  }
}

// The whole render of trace.cpp:246-262: tile jobs with jumped xoshiro streams, spp passes, p += colour / spp.
// rgb is height x width x 3 (R,G,B); stats = {total ray casts, max path length}.
void ref_cornell_render(uint32_t width, uint32_t height, uint32_t tileWidth, uint32_t tileHeight, uint32_t spp, uint64_t seed, float aaScale,
                        float refractiveIndex, uint32_t rouletteDepth, float stopProb, float epsilon, float intersectionEpsilon, float* rgb,
                        uint64_t* stats) {
  light::epsilon = epsilon;
  light::intersectionEpsilon = intersectionEpsilon;
  Cornell cornell;
  light::RayTracerContext<decltype(cornell.scene)> tracer(cornell.scene);
  tracer.refractiveIndex = refractiveIndex;
  tracer.rouletteDepth = rouletteDepth;
  tracer.stopProb = stopProb;
  auto jobs = createTracingJobs(width, height, tileWidth, tileHeight, spp, seed);
  for (uint32_t s = 0; s < spp; ++s) {
#pragma omp parallel for schedule(dynamic)
    for (std::size_t j = 0; j < jobs.size(); ++j) {
      auto& job = jobs[j];
      job.visitPixels([&](std::size_t row, std::size_t col, Vec& p) {
        constexpr float fov = light::Pi / 2.f;
        Vec cam = light::pixelToRay(col, row, width, height, fov);
        Vec aaNoise(xoshiro::uniform_neg1_1(job.rngState), xoshiro::uniform_neg1_1(job.rngState), 0.f);
        cam += aaNoise * aaScale;
        const light::Ray ray(Vec(0.f, 0.f, 0.f), cam);
        auto color = trace(ray, tracer, job);
        p += color / spp;
      });
    }
  }
  uint64_t total = 0, longest = 0;
  for (auto& j : jobs) {
    total += j.totalRayCasts;
    longest = std::max<uint64_t>(longest, j.maxPathLength);
    j.visitPixels([&](std::size_t row, std::size_t col, Vec& p) {
      float* o = rgb + (row * width + col) * 3;
      o[0] = p.x; o[1] = p.y; o[2] = p.z;
    });
  }
  stats[0] = total; stats[1] = longest;
}

}  // extern "C"
