// Path-tracing wavefront kernels (compiled with -fmad=false, see pt_device.cuh).
//
// trace_wavefront_kernel is stages S0-S4 of the reference's pathTraceIteration
// (PathTracerApp.cpp:432-458): anti-alias noise, GenerateCameraRays (codelets.cpp:36-80), primary
// samples, RayTraceKernel (:93-227) and PreProcessEscapedRays (:312-358) for `samplesPerLaunch`
// samples of every work item. The reference evaluates the NIF for every path and discards
// the result for those that did not escape (codelets.cpp:348-352); here only escaped paths are
// appended to the NIF queue, compacted with warp ballots and one atomic per block and sample.
#include "kernels.h"
#include "pt_device.cuh"

namespace b2 {

__device__ __forceinline__ void stageScene(SceneSoA& dst, const SceneSoA* __restrict__ scene) {
  // SceneSoA is a few hundred bytes: word-wise copy global -> shared, once per block.
  const uint32_t* src = reinterpret_cast<const uint32_t*>(scene);
  uint32_t* d = reinterpret_cast<uint32_t*>(&dst);
  for (int i = threadIdx.x; i < (int)(sizeof(SceneSoA) / 4); i += blockDim.x) d[i] = __ldg(src + i);
  __syncthreads();
}

constexpr int kTraceThreads = 256;

__global__ void __launch_bounds__(kTraceThreads)
trace_wavefront_kernel(const SceneSoA* __restrict__ scene, const DeviceSettings s, const uint32_t* __restrict__ pix, uint32_t nItems, uint32_t sampleBegin, uint32_t samplesPerLaunch,
                       float2* __restrict__ qUv, float4* __restrict__ qThrSlot, uint32_t* __restrict__ qCount,
                       uint32_t* __restrict__ sampleCount, uint32_t* __restrict__ pathLength,
                       unsigned long long* __restrict__ stats /* [0] paths [1] escaped [2] segments */) {
  __shared__ SceneSoA sc;
  __shared__ uint32_t warpCount[kTraceThreads / 32];
  __shared__ uint32_t blockBase;
  stageScene(sc, scene);

  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t pu = 0xffffu, pv = 0xffffu;
  if (slot < nItems) { const uint32_t p = pix[slot]; pu = p & 0xffffu; pv = p >> 16; }
  const bool valid = pu < s.width && pv < s.height;       // padding items: LoadBalancer.cpp:66-71
  const uint32_t pixelId = pv * s.width + pu;

  uint32_t myLength = 0, myEscaped = 0;
  for (uint32_t k = 0; k < samplesPerLaunch; ++k) {
    bool esc = false;
    float u = 0.f, v = 0.f;
    V3 thr = v3(0.f, 0.f, 0.f);
    if (valid) {
      PathRng rng(s.seedLo, s.seedHi, pixelId, sampleBegin + k);
      const uint32_t w0 = rng.word(), w1 = rng.word();
      float n0, n1;
      aaNoisePair(w0, w1, s.aaNoiseType, n0, n1);
      float cx, cy;
      cameraRay(pu, pv, n0, n1, s, cx, cy);
      const PathResult r = tracePath(cx, cy, s, sc, [&]() { return rng.uniformHalf(); });
      myLength += (uint32_t)r.length;
      if (r.escaped) {
        esc = true;
        escapedUv(r.dir, s.azimuth, u, v);
        thr = r.throughput;
      }
    }
    // compaction: warp ballot -> block prefix -> one atomic per block
    const uint32_t m = __ballot_sync(0xffffffffu, esc);
    if (lane == 0) warpCount[warp] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t total = 0;
#pragma unroll
      for (int w = 0; w < kTraceThreads / 32; ++w) { const uint32_t c = warpCount[w]; warpCount[w] = total; total += c; }
      blockBase = total ? atomicAdd(qCount, total) : 0u;
    }
    __syncthreads();
    if (esc) {
      const uint32_t idx = blockBase + warpCount[warp] + __popc(m & ((1u << lane) - 1u));
      qUv[idx] = make_float2(u, v);
      qThrSlot[idx] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(slot));
      myEscaped += 1;
    }
    __syncthreads();      // warpCount / blockBase are rewritten by the next sample
  }
  if (valid) {
    sampleCount[slot] += samplesPerLaunch;     // codelets.cpp:300-301: every path counts
    pathLength[slot] += myLength;
  }
  // statistics: block reduction, 3 atomics per block
  __shared__ unsigned long long red[3];
  if (threadIdx.x < 3) red[threadIdx.x] = 0ull;
  __syncthreads();
  uint32_t a = valid ? samplesPerLaunch : 0u, b = myEscaped, c = myLength;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if (lane == 0) { atomicAdd(&red[0], (unsigned long long)a); atomicAdd(&red[1], (unsigned long long)b); atomicAdd(&red[2], (unsigned long long)c); }
  __syncthreads();
  if (threadIdx.x < 3 && red[threadIdx.x]) atomicAdd(&stats[threadIdx.x], red[threadIdx.x]);
}

cudaError_t launchTraceWavefront(const SceneSoA* scene, const DeviceSettings& s, const uint32_t* pix, uint32_t nItems, uint32_t sampleBegin, uint32_t samplesPerLaunch,
                                 float2* qUv, float4* qThrSlot, uint32_t* qCount, uint32_t* sampleCount,
                                 uint32_t* pathLength, unsigned long long* stats, cudaStream_t stream) {
  if (nItems == 0) return cudaSuccess;
  const uint32_t blocks = (nItems + kTraceThreads - 1) / kTraceThreads;
  trace_wavefront_kernel<<<blocks, kTraceThreads, 0, stream>>>(scene, s, pix, nItems, sampleBegin, samplesPerLaunch, qUv, qThrSlot,
                                                              qCount, sampleCount, pathLength, stats);
  return cudaGetLastError();
}

// One sample, no compaction, everything written per item: the stage-level parity entry point.
__global__ void __launch_bounds__(kTraceThreads)
trace_items_kernel(const SceneSoA* __restrict__ scene, const DeviceSettings s, const uint32_t* __restrict__ pix, uint32_t nItems, uint32_t sample, int32_t* __restrict__ escaped,
                   float* __restrict__ uv2, float* __restrict__ thr3, int32_t* __restrict__ length, uint16_t* __restrict__ cam2) {
  __shared__ SceneSoA sc;
  stageScene(sc, scene);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nItems) return;
  const uint32_t p = pix[i];
  const uint32_t pu = p & 0xffffu, pv = p >> 16;
  int esc = 0, len = 0;
  float u = 0.f, v = 0.f;
  V3 thr = v3(0.f, 0.f, 0.f);
  uint16_t c0 = 0, c1 = 0;
  if (pu < s.width && pv < s.height) {
    PathRng rng(s.seedLo, s.seedHi, pv * s.width + pu, sample);
    const uint32_t w0 = rng.word(), w1 = rng.word();
    float n0, n1;
    aaNoisePair(w0, w1, s.aaNoiseType, n0, n1);
    float cx, cy;
    cameraRay(pu, pv, n0, n1, s, cx, cy);
    c0 = halfBits(cx); c1 = halfBits(cy);
    const PathResult r = tracePath(cx, cy, s, sc, [&]() { return rng.uniformHalf(); });
    len = r.length;
    if (r.escaped) { esc = 1; escapedUv(r.dir, s.azimuth, u, v); thr = r.throughput; }
  }
  if (escaped) escaped[i] = esc;
  if (length) length[i] = len;
  if (uv2) { uv2[2 * i] = u; uv2[2 * i + 1] = v; }
  if (thr3) { thr3[3 * i] = thr.x; thr3[3 * i + 1] = thr.y; thr3[3 * i + 2] = thr.z; }
  if (cam2) { cam2[2 * i] = c0; cam2[2 * i + 1] = c1; }
}

cudaError_t launchTraceItems(const SceneSoA* scene, const DeviceSettings& s, const uint32_t* pix, uint32_t nItems, uint32_t sample, int32_t* escaped, float* uv2,
                             float* thr3, int32_t* length, uint16_t* cam2, cudaStream_t stream) {
  if (nItems == 0) return cudaSuccess;
  trace_items_kernel<<<(nItems + kTraceThreads - 1) / kTraceThreads, kTraceThreads, 0, stream>>>(scene, s, pix, nItems, sample, escaped,
                                                                                             uv2, thr3, length, cam2);
  return cudaGetLastError();
}

// RayTraceKernel on caller-supplied camera rays and uniform buffers (codelets.cpp:147-155: sequential
// reads with wrap-around).
__global__ void __launch_bounds__(kTraceThreads)
trace_fixed_kernel(const SceneSoA* __restrict__ scene, const DeviceSettings s, uint32_t n, const uint16_t* __restrict__ cam2, const uint16_t* __restrict__ uniforms, uint32_t perPath,
                   int32_t* __restrict__ escaped, float* __restrict__ dir3, float* __restrict__ thr3, int32_t* __restrict__ length) {
  __shared__ SceneSoA sc;
  stageScene(sc, scene);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint16_t* mine = uniforms + (size_t)i * perPath;
  uint32_t next = 0;
  const PathResult r = tracePath(halfBitsToFloat(cam2[2 * i]), halfBitsToFloat(cam2[2 * i + 1]), s, sc, [&]() {
    const float v = halfBitsToFloat(mine[next]);
    next += 1;
    if (next == perPath) next = 0;
    return v;
  });
  escaped[i] = r.escaped ? 1 : 0;
  length[i] = r.length;
  const V3 d = r.escaped ? r.dir : v3(0.f, 0.f, 0.f);
  dir3[3 * i] = d.x; dir3[3 * i + 1] = d.y; dir3[3 * i + 2] = d.z;
  thr3[3 * i] = r.throughput.x; thr3[3 * i + 1] = r.throughput.y; thr3[3 * i + 2] = r.throughput.z;
}

cudaError_t launchTraceFixed(const SceneSoA* scene, const DeviceSettings& s, uint32_t n, const uint16_t* cam2, const uint16_t* uniforms, uint32_t perPath, int32_t* escaped,
                             float* dir3, float* thr3, int32_t* length, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  trace_fixed_kernel<<<(n + kTraceThreads - 1) / kTraceThreads, kTraceThreads, 0, stream>>>(scene, s, n, cam2, uniforms, perPath, escaped,
                                                                                        dir3, thr3, length);
  return cudaGetLastError();
}

// Function-level known-answer kernel: calls the very device functions the wavefront kernel uses.
__global__ void kat_kernel(int op, uint32_t n, const float* __restrict__ in, int inStride, float* __restrict__ out, int outStride) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* a = in + (size_t)i * inStride;
  float* o = out + (size_t)i * outStride;
  switch (op) {
    case 0: {   // pixelToRay light.hpp:194-206 (unrounded, fp32 in and out)
      const float w = a[2], h = a[3];
      const float aspect = w / h;
      const float tanTheta = tanf(a[4] / 2.f);
      const float x = (a[0] / w) - .5f, y = (a[1] / h) - .5f;
      o[0] = 2.f * x * aspect * tanTheta; o[1] = -2.f * y * tanTheta; o[2] = -1.f;
      break;
    }
    case 1: {   // Sphere::intersect; the Ray ctor normalises the direction
      const V3 d = unit(v3(a[3], a[4], a[5]));
      o[0] = hitSphere(v3(a[6], a[7], a[8]), a[9] * a[9], v3(a[0], a[1], a[2]), d);
      break;
    }
    case 2: {   // Disc::intersect; the Disc ctor normalises the normal and takes |c.n|
      const V3 d = unit(v3(a[3], a[4], a[5]));
      const V3 nn = unit(v3(a[6], a[7], a[8]));
      const V3 c = v3(a[9], a[10], a[11]);
      o[0] = hitDisc(nn, c, a[12] * a[12], fabsf(dot(c, nn)), v3(a[0], a[1], a[2]), d);
      break;
    }
    case 3: {   // diffuse
      V3 d = v3(0.f, 0.f, 0.f);
      const float w = sampleDiffuse(d, v3(a[0], a[1], a[2]), a[5], a[3], a[4]);
      o[0] = d.x; o[1] = d.y; o[2] = d.z; o[3] = w;
      break;
    }
    case 4: {   // reflect (Ray ctor normalises the incoming direction)
      V3 d = unit(v3(a[0], a[1], a[2]));
      sampleMirror(d, v3(a[3], a[4], a[5]));
      o[0] = d.x; o[1] = d.y; o[2] = d.z;
      break;
    }
    case 5: {   // refract
      V3 d = unit(v3(a[0], a[1], a[2]));
      const bool t = sampleGlass(d, v3(a[3], a[4], a[5]), a[6], a[7]);
      o[0] = d.x; o[1] = d.y; o[2] = d.z; o[3] = t ? 1.f : 0.f;
      break;
    }
    case 6: {   // rouletteWeight light.hpp:25-29
      const bool stop = a[0] <= a[1];
      o[0] = stop ? 1.f : 0.f;
      o[1] = stop ? 1.f : 1.f / (1.f - a[1]);
      break;
    }
    case 7: {   // escaped-ray uv
      float u, v;
      escapedUv(v3(a[0], a[1], a[2]), a[3], u, v);
      o[0] = u; o[1] = v;
      break;
    }
    case 8: {   // Philox: 6 input words, 4 output words, raw bits in float slots
      const Philox4 r = philox4x32_10(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]),
                                      __float_as_uint(a[4]), __float_as_uint(a[5]));
      for (int k = 0; k < 4; ++k) o[k] = __uint_as_float(r.v[k]);
      break;
    }
    case 9: {   // anti-alias noise pair
      float n0, n1;
      aaNoisePair(__float_as_uint(a[0]), __float_as_uint(a[1]), (int)__float_as_uint(a[2]), n0, n1);
      o[0] = n0; o[1] = n1;
      break;
    }
    default: break;
  }
}

cudaError_t launchKat(int op, uint32_t n, const float* in, int inStride, float* out, int outStride, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  kat_kernel<<<(n + 127) / 128, 128, 0, stream>>>(op, n, in, inStride, out, outStride);
  return cudaGetLastError();
}

} // namespace b2
