// Work-item records <-> device SoA, and the fp32 HDR film that replaces the reference's host-side
// AccumulatedImage (src/AccumulatedImage.cpp). All of these are pure streaming kernels (HBM bound).
#include "kernels.h"

namespace b2 {

// "setup": TraceRecord[] (20-byte AoS as the reference streams it, PathTracerApp.cpp:426) -> SoA.
__global__ void unpack_records_kernel(const TraceRecord* __restrict__ rec, uint32_t n, uint32_t* __restrict__ pix,
                                      AccumSlot* __restrict__ accum, uint32_t* __restrict__ sampleCount, uint32_t* __restrict__ pathLength) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const TraceRecord r = rec[i];
  pix[i] = (uint32_t)r.u | ((uint32_t)r.v << 16);
  accum[i] = AccumSlot{accumToFixed(r.r), accumToFixed(r.g), accumToFixed(r.b), 0};
  sampleCount[i] = r.sampleCount;
  pathLength[i] = r.pathLength;
}

// "read_results": SoA -> TraceRecord[]; the two counters wrap like the reference's uint16 fields.
__global__ void pack_records_kernel(TraceRecord* __restrict__ rec, uint32_t n, const uint32_t* __restrict__ pix,
                                    const AccumSlot* __restrict__ accum, const uint32_t* __restrict__ sampleCount,
                                    const uint32_t* __restrict__ pathLength) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  TraceRecord r;
  const uint32_t p = pix[i];
  const AccumSlot a = accum[i];
  r.u = (uint16_t)(p & 0xffffu); r.v = (uint16_t)(p >> 16);
  r.r = accumToFloat(a.r); r.g = accumToFloat(a.g); r.b = accumToFloat(a.b);
  r.sampleCount = (uint16_t)sampleCount[i];
  r.pathLength = (uint16_t)pathLength[i];
  rec[i] = r;
}

// AccumulatedImage::accumulate (AccumulatedImage.cpp:59-74): hdr(v,u) += (b,g,r)/sampleCount for
// valid items, then LoadBalancer::clearInactiveAccumulators. One work item owns one pixel, so
// no atomics are needed.
__global__ void film_accumulate_kernel(uint32_t n, const uint32_t* __restrict__ pix, AccumSlot* __restrict__ accum,
                                       uint32_t* __restrict__ sampleCount, uint32_t* __restrict__ pathLength,
                                       float* __restrict__ film, uint32_t width, uint32_t height) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t p = pix[i];
  const uint32_t u = p & 0xffffu, v = p >> 16;
  const uint32_t sc = sampleCount[i];
  if (u < width && v < height && sc > 0) {
    const AccumSlot a = accum[i];
    const float inv = 1.f / (float)sc;
    float* px = film + 3 * ((size_t)v * width + u);
    px[0] += accumToFloat(a.b) * inv; px[1] += accumToFloat(a.g) * inv; px[2] += accumToFloat(a.r) * inv;      // stored B,G,R (:70)
  }
  accum[i] = AccumSlot{0, 0, 0, 0};
  sampleCount[i] = 0;
  pathLength[i] = 0;
}

__global__ void film_scale_kernel(const float* __restrict__ film, float* __restrict__ out, size_t n, float scale) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = film[i] * scale;
}

// AccumulatedImage::updateLdrImage (:23-46): (hdr * 2^exposure)^(1/gamma) * 255, saturated.
__global__ void film_tonemap_kernel(const float* __restrict__ film, uint8_t* __restrict__ out, size_t n, float scale,
                                    float exposureScale, float invGamma) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v = film[i] * scale * exposureScale;
  v = powf(fmaxf(v, 0.f), invGamma) * 255.f;
  v = fminf(fmaxf(v, 0.f), 255.f);
  out[i] = (uint8_t)__float2int_rn(v);
}

static inline uint32_t blocksFor(size_t n, int t) { return (uint32_t)((n + t - 1) / t); }

cudaError_t launchUnpackRecords(const TraceRecord* rec, uint32_t n, uint32_t* pix, AccumSlot* accum, uint32_t* sampleCount,
                                uint32_t* pathLength, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  unpack_records_kernel<<<blocksFor(n, 256), 256, 0, stream>>>(rec, n, pix, accum, sampleCount, pathLength);
  return cudaGetLastError();
}
cudaError_t launchPackRecords(TraceRecord* rec, uint32_t n, const uint32_t* pix, const AccumSlot* accum, const uint32_t* sampleCount,
                              const uint32_t* pathLength, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  pack_records_kernel<<<blocksFor(n, 256), 256, 0, stream>>>(rec, n, pix, accum, sampleCount, pathLength);
  return cudaGetLastError();
}
cudaError_t launchFilmAccumulate(uint32_t n, const uint32_t* pix, AccumSlot* accum, uint32_t* sampleCount, uint32_t* pathLength,
                                 float* filmBgr, uint32_t width, uint32_t height, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  film_accumulate_kernel<<<blocksFor(n, 256), 256, 0, stream>>>(n, pix, accum, sampleCount, pathLength, filmBgr, width, height);
  return cudaGetLastError();
}
cudaError_t launchFilmScale(const float* filmBgr, float* out, size_t n, float scale, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  film_scale_kernel<<<blocksFor(n, 256), 256, 0, stream>>>(filmBgr, out, n, scale);
  return cudaGetLastError();
}
cudaError_t launchFilmTonemap(const float* filmBgr, uint8_t* out, size_t n, float scale, float exposure, float gamma, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  film_tonemap_kernel<<<blocksFor(n, 256), 256, 0, stream>>>(filmBgr, out, n, scale, exp2f(exposure), 1.f / gamma);
  return cudaGetLastError();
}

} // namespace b2
