// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
//
// C-ABI driver around the *unmodified* reference device code, compiled from where
// it lies under /root/reference (never copied into this repo):
//   /root/reference/src/codelets/codelets.cpp        (5 vertex classes)
//   /root/reference/external/light/src/{vector,light,sdf}.hpp
// Built by oracle/Makefile into oracle/_ref/libref_codelets.so. It is the
// authority the CPU restatement (oracle/pt_oracle.cpp) is pinned against, the
// generator of tests/golden/*.npz, and (kind="reference") a timed CPU baseline for the
// path-tracing stages. The NIF MLP has no CPU implementation in the reference
// (it exists only as a Poplar graph), so it is not here.
//
// Stage order reproduced by ref_trace_iteration: PathTracerApp.cpp:432-458.

#include <cfloat>
#include <cstring>
#include <vector>

#include <codelets.cpp>  // resolved through -I/root/reference/src/codelets

namespace light {
// off-IPU these are extern (light.hpp:17-23); values from external/light/src/light.cpp:5-9
float epsilon = std::numeric_limits<float>::epsilon();
float intersectionEpsilon = 1e-5;
}

namespace {

inline half halfFromBits(std::uint16_t b) {
  __half_raw r; r.x = b; return half(r);
}
inline std::uint16_t halfBits(half h) {
  __half_raw r = h; return r.x;
}

using ByteVec = poplar::Vector<unsigned char>;
constexpr std::size_t kContribBytes = sizeof(light::Contribution);

} // namespace

extern "C" {

int ref_sizeof_trace_record() { return (int)sizeof(TraceRecord); }
int ref_sizeof_contribution() { return (int)kContribBytes; }

// ---- function-level KATs (light.hpp / sdf.hpp) ---------------------------------

// light.hpp:194-206
void ref_pixel_to_ray(int n, const float* x, const float* y, unsigned w, unsigned h, float fov, float* out3) {
  for (int i = 0; i < n; ++i) {
    auto v = light::pixelToRay(x[i], y[i], w, h, fov);
    out3[3 * i] = v.x; out3[3 * i + 1] = v.y; out3[3 * i + 2] = v.z;
  }
}

// light.hpp:131-147 ; ray direction is normalised by the Ray ctor (light.hpp:35)
void ref_sphere_intersect(int n, const float* origin3, const float* dir3, const float* centre3, const float* radius, float* t) {
  for (int i = 0; i < n; ++i) {
    light::Sphere s(Vec(centre3[3 * i], centre3[3 * i + 1], centre3[3 * i + 2]), radius[i]);
    light::Ray ray(Vec(origin3[3 * i], origin3[3 * i + 1], origin3[3 * i + 2]), Vec(dir3[3 * i], dir3[3 * i + 1], dir3[3 * i + 2]));
    t[i] = s.intersect(ray);
  }
}

// light.hpp:95-121
void ref_disc_intersect(int n, const float* origin3, const float* dir3, const float* normal3, const float* centre3, const float* radius, float* t) {
  for (int i = 0; i < n; ++i) {
    light::Disc d(Vec(normal3[3 * i], normal3[3 * i + 1], normal3[3 * i + 2]),
                  Vec(centre3[3 * i], centre3[3 * i + 1], centre3[3 * i + 2]), radius[i]);
    light::Ray ray(Vec(origin3[3 * i], origin3[3 * i + 1], origin3[3 * i + 2]), Vec(dir3[3 * i], dir3[3 * i + 1], dir3[3 * i + 2]));
    t[i] = d.intersect(ray);
  }
}

// sdf.hpp:8-31 (+ hemisphere light.hpp:223-228, orthonormalSystem :230-250)
void ref_diffuse(int n, const float* normal3, const float* u1, const float* u2, float rrFactor, float* outDir3, float* outWeight) {
  light::Material m(Vec(1.f, 1.f, 1.f), Vec(0.f, 0.f, 0.f), light::Material::Type::diffuse);
  for (int i = 0; i < n; ++i) {
    light::Ray ray;
    Vec nrm(normal3[3 * i], normal3[3 * i + 1], normal3[3 * i + 2]);
    light::Intersection isect(nullptr, &m, 1.f);
    auto c = light::diffuse(ray, nrm, isect, rrFactor, u1[i], u2[i]);
    outDir3[3 * i] = ray.direction.x; outDir3[3 * i + 1] = ray.direction.y; outDir3[3 * i + 2] = ray.direction.z;
    outWeight[i] = c.weight;
  }
}

// sdf.hpp:36-40
void ref_reflect(int n, const float* dir3, const float* normal3, float* outDir3) {
  for (int i = 0; i < n; ++i) {
    light::Ray ray(Vec(0.f, 0.f, 0.f), Vec(dir3[3 * i], dir3[3 * i + 1], dir3[3 * i + 2]));
    light::reflect(ray, Vec(normal3[3 * i], normal3[3 * i + 1], normal3[3 * i + 2]));
    outDir3[3 * i] = ray.direction.x; outDir3[3 * i + 1] = ray.direction.y; outDir3[3 * i + 2] = ray.direction.z;
  }
}

// sdf.hpp:45-67
void ref_refract(int n, const float* dir3, const float* normal3, float ri, const float* rnd, float* outDir3, int* refracted) {
  for (int i = 0; i < n; ++i) {
    light::Ray ray(Vec(0.f, 0.f, 0.f), Vec(dir3[3 * i], dir3[3 * i + 1], dir3[3 * i + 2]));
    refracted[i] = light::refract(ray, Vec(normal3[3 * i], normal3[3 * i + 1], normal3[3 * i + 2]), ri, rnd[i]) ? 1 : 0;
    outDir3[3 * i] = ray.direction.x; outDir3[3 * i + 1] = ray.direction.y; outDir3[3 * i + 2] = ray.direction.z;
  }
}

// light.hpp:25-29
void ref_roulette(int n, const float* rnd, float stopProb, int* stop, float* weight) {
  for (int i = 0; i < n; ++i) {
    auto r = light::rouletteWeight(rnd[i], stopProb);
    stop[i] = r.first ? 1 : 0; weight[i] = r.second;
  }
}

// ---- vertex-level drivers (codelets.cpp) ----------------------------------------

// GenerateCameraRays (codelets.cpp:36-80). `records` is TraceRecord[n]; `noise` holds
// at least 2n halfs padded to a multiple of 6 (PathTracerApp.cpp:270-275);
// `rays` receives 2n halfs. All 6 worker contexts are run.
void ref_generate_camera_rays(int n, const unsigned char* records, const std::uint16_t* noise, int noiseCount,
                              std::uint16_t* rays, unsigned width, unsigned height,
                              std::uint16_t aaScaleBits, std::uint16_t fovBits) {
  std::vector<half> noiseH(noiseCount), raysH(2 * (std::size_t)n);
  for (int i = 0; i < noiseCount; ++i) noiseH[i] = halfFromBits(noise[i]);
  GenerateCameraRays v;
  v.antiAliasNoise = poplar::Input<poplar::Vector<half>>(poplar::Vector<half>(noiseH.data(), noiseH.size()));
  v.rays = poplar::Output<poplar::Vector<half>>(poplar::Vector<half>(raysH.data(), raysH.size()));
  v.traceBuffer = poplar::Input<ByteVec>(ByteVec(const_cast<unsigned char*>(records), (std::size_t)n * sizeof(TraceRecord)));
  v.imageWidth = poplar::Input<unsigned>(width);
  v.imageHeight = poplar::Input<unsigned>(height);
  v.antiAliasScale = poplar::Input<half>(halfFromBits(aaScaleBits));
  v.fov = poplar::Input<half>(halfFromBits(fovBits));
  for (unsigned w = 0; w < GenerateCameraRays::numWorkers(); ++w) v.compute(w);
  for (std::size_t i = 0; i < raysH.size(); ++i) rays[i] = halfBits(raysH[i]);
}

// RayTraceKernel (codelets.cpp:93-227). One vertex instance consumes `uniforms`
// sequentially for all its rays, wrapping at the end (codelets.cpp:147-155), exactly
// like one IPU worker. With perPath != 0 every ray instead gets its own vertex over
// the slice uniforms[i*uniformsPerPath, (i+1)*uniformsPerPath) so that paths are
// independent of one another (what the fixed-input GPU KATs compare against).
// `contribs` receives n rows of maxPathLength*sizeof(Contribution) bytes.
void ref_ray_trace(int n, const std::uint16_t* cameraRays, const std::uint16_t* uniforms, int uniformCount,
                   int perPath, int uniformsPerPath, unsigned char* contribs, int maxPathLength,
                   std::uint16_t refractiveIndexBits, std::uint16_t stopProbBits, unsigned short rouletteDepth) {
  const std::size_t rowBytes = (std::size_t)maxPathLength * kContribBytes;
  std::vector<half> raysH(2 * (std::size_t)n), uniH(uniformCount);
  for (std::size_t i = 0; i < raysH.size(); ++i) raysH[i] = halfFromBits(cameraRays[i]);
  for (int i = 0; i < uniformCount; ++i) uniH[i] = halfFromBits(uniforms[i]);
  std::vector<poplar::Output<ByteVec>> rows(n);
  for (int i = 0; i < n; ++i) rows[i] = poplar::Output<ByteVec>(ByteVec(contribs + i * rowBytes, rowBytes));

  auto run = [&](int first, int count, half* uni, std::size_t uniN) {
    RayTraceKernel k;
    k.cameraRays = poplar::Input<poplar::Vector<half>>(poplar::Vector<half>(raysH.data() + 2 * first, 2 * (std::size_t)count));
    k.uniform_0_1 = poplar::Input<poplar::Vector<half>>(poplar::Vector<half>(uni, uniN));
    k.contributionData = poplar::Vector<poplar::Output<ByteVec>, poplar::VectorLayout::ONE_PTR>(rows.data() + first, (std::size_t)count);
    k.refractiveIndex = poplar::Input<half>(halfFromBits(refractiveIndexBits));
    k.stopProb = poplar::Input<half>(halfFromBits(stopProbBits));
    k.rouletteDepth = poplar::Input<unsigned short>(rouletteDepth);
    k.compute();
  };
  if (perPath) {
    for (int i = 0; i < n; ++i) run(i, 1, uniH.data() + (std::size_t)i * uniformsPerPath, (std::size_t)uniformsPerPath);
  } else {
    run(0, n, uniH.data(), uniH.size());
  }
}

// PreProcessEscapedRays (codelets.cpp:312-358)
void ref_preprocess_escaped(int n, unsigned char* contribs, int maxPathLength, float azimuth, float* u, float* v) {
  const std::size_t rowBytes = (std::size_t)maxPathLength * kContribBytes;
  std::vector<poplar::InOut<ByteVec>> rows(n);
  for (int i = 0; i < n; ++i) rows[i] = poplar::InOut<ByteVec>(ByteVec(contribs + i * rowBytes, rowBytes));
  PreProcessEscapedRays p;
  p.contributionData = poplar::Vector<poplar::InOut<ByteVec>>(rows.data(), rows.size());
  p.azimuthalOffset = poplar::Input<float>(azimuth);
  p.u = poplar::Output<poplar::Vector<float>>(poplar::Vector<float>(u, n));
  p.v = poplar::Output<poplar::Vector<float>>(poplar::Vector<float>(v, n));
  for (unsigned w = 0; w < PreProcessEscapedRays::numWorkers(); ++w) p.compute(w);
}

// PostProcessEscapedRays (codelets.cpp:361-385); bgr is n x 3 floats
void ref_postprocess_escaped(int n, unsigned char* contribs, int maxPathLength, float* bgr) {
  const std::size_t rowBytes = (std::size_t)maxPathLength * kContribBytes;
  std::vector<poplar::InOut<ByteVec>> rows(n);
  std::vector<poplar::Input<poplar::Vector<float>>> bgrRows(n);
  for (int i = 0; i < n; ++i) {
    rows[i] = poplar::InOut<ByteVec>(ByteVec(contribs + i * rowBytes, rowBytes));
    bgrRows[i] = poplar::Input<poplar::Vector<float>>(poplar::Vector<float>(bgr + 3 * (std::size_t)i, 3));
  }
  PostProcessEscapedRays p;
  p.contributionData = poplar::Vector<poplar::InOut<ByteVec>>(rows.data(), rows.size());
  p.bgr = poplar::Vector<poplar::Input<poplar::Vector<float>>>(bgrRows.data(), bgrRows.size());
  for (unsigned w = 0; w < PostProcessEscapedRays::numWorkers(); ++w) p.compute(w);
}

// AccumulateContributions (codelets.cpp:235-306); records is TraceRecord[n] in/out
void ref_accumulate(int n, const unsigned char* contribs, int maxPathLength, unsigned char* records) {
  const std::size_t rowBytes = (std::size_t)maxPathLength * kContribBytes;
  std::vector<poplar::Input<ByteVec>> rows(n);
  for (int i = 0; i < n; ++i) rows[i] = poplar::Input<ByteVec>(ByteVec(const_cast<unsigned char*>(contribs) + i * rowBytes, rowBytes));
  AccumulateContributions a;
  a.contributionData = poplar::Vector<poplar::Input<ByteVec>>(rows.data(), rows.size());
  a.traceBuffer = poplar::InOut<ByteVec>(ByteVec(records, (std::size_t)n * sizeof(TraceRecord)));
  a.compute();
}

} // extern "C"
