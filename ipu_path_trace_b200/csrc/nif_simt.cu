// NIF inference on the CUDA cores, for any Dense stack (compiled with -fmad=false).
//
// This is the general-topology kernel: it takes whatever layer list the weights file
// holds (skip-concat autodetect NifModel.cpp:303-308, activations :73-76,323) and follows the
// reference's numerics step for step - binary16 features, binary16 parameters, fp32
// accumulation in k order, every Dense output rounded to binary16 (poplin::matMul is
// asked for the kernel's type, :314), bias added in binary16 (:320), ReLU, fp32 decode
// (:226-242). The tcgen05 kernel (nif_tc.cu) is the fast path for the shipped topology;
// this one also serves as its on-device cross-check.
//
// One thread per row; activations live in shared memory transposed ([k][row]) so a warp's
// reads are conflict-free, weights are read through the read-only path as warp-uniform
// 16-byte loads.
#include "nif_common.cuh"

namespace b2 {

constexpr int kSimtRows = 128;
constexpr int kSimtTile = 16;       // outputs computed per pass over k

__global__ void __launch_bounds__(kSimtRows)
nif_simt_kernel(const NifSimtModel m, const NifIo io) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  __half* feat = reinterpret_cast<__half*>(smemRaw);                       // [nFeat][128]
  __half* bufA = feat + (size_t)m.nFeat * kSimtRows;                       // [maxWidth][128]
  __half* bufB = bufA + (size_t)m.maxWidth * kSimtRows;

  const uint32_t count = io.count ? *io.count : io.countHost;
  const uint32_t r = threadIdx.x;
  for (uint32_t base = blockIdx.x * kSimtRows; base < count; base += gridDim.x * kSimtRows) {
    const uint32_t row = base + r;
    const bool live = row < count;
    float2 uv = make_float2(0.f, 0.f);
    if (live) uv = io.uv[row];
    nifEncode(uv.x, uv.y, m.embeddingDim, [&](int k, __half v) { feat[k * kSimtRows + r] = v; });

    const __half* x = feat;      // current input, [width][128]
    int width = m.nFeat;
    __half* outBuf = bufA;
    float result[4] = {0.f, 0.f, 0.f, 0.f};
    for (int l = 0; l < m.numLayers; ++l) {
      const int in = m.in[l], out = m.out[l];
      const float* __restrict__ W = m.w[l];
      const float* __restrict__ B = m.b[l];
      const int nAct = m.concat[l] ? width : in;       // rows of W fed by x; the rest by the features
      for (int j0 = 0; j0 < out; j0 += kSimtTile) {
        float acc[kSimtTile];
#pragma unroll
        for (int j = 0; j < kSimtTile; ++j) acc[j] = 0.f;
        const bool full = (j0 + kSimtTile <= out) && ((out & 3) == 0);
        for (int k = 0; k < in; ++k) {
          const float a = __half2float(k < nAct ? x[k * kSimtRows + r] : feat[(k - nAct) * kSimtRows + r]);
          const float* wrow = W + (size_t)k * out + j0;
          if (full) {
#pragma unroll
            for (int q = 0; q < kSimtTile / 4; ++q) {
              const float4 w4 = __ldg(reinterpret_cast<const float4*>(wrow) + q);
              acc[4 * q] = acc[4 * q] + a * w4.x;
              acc[4 * q + 1] = acc[4 * q + 1] + a * w4.y;
              acc[4 * q + 2] = acc[4 * q + 2] + a * w4.z;
              acc[4 * q + 3] = acc[4 * q + 3] + a * w4.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < kSimtTile; ++j)
              if (j0 + j < out) acc[j] = acc[j] + a * __ldg(wrow + j);
          }
        }
#pragma unroll
        for (int j = 0; j < kSimtTile; ++j) {
          if (j0 + j < out) {
            __half h = __float2half_rn(acc[j]);
            if (m.hasBias[l]) h = __float2half_rn(__half2float(h) + __ldg(B + j0 + j));
            if (m.relu[l] && __half2float(h) < 0.f) h = __float2half_rn(0.f);
            if (l + 1 < m.numLayers) outBuf[(j0 + j) * kSimtRows + r] = h;
            else if (j0 + j < 4) result[j0 + j] = __half2float(h);
          }
        }
      }
      x = outBuf;
      width = out;
      outBuf = (outBuf == bufA) ? bufB : bufA;
    }
    if (live) {
      const float b = nifDecode(result[0], m.maxv, m.mean[0], m.logToneMap);
      const float g = nifDecode(result[1], m.maxv, m.mean[1], m.logToneMap);
      const float rr = nifDecode(result[2], m.maxv, m.mean[2], m.logToneMap);
      float4 ts = make_float4(0.f, 0.f, 0.f, 0.f);
      if (io.thrSlot) ts = io.thrSlot[row];
      nifEmit(io, row, b, g, rr, ts);
    }
  }
}

size_t nifSimtSmemBytes(const NifSimtModel& m) {
  return ((size_t)m.nFeat + 2 * (size_t)m.maxWidth) * kSimtRows * sizeof(__half);
}

cudaError_t launchNifSimt(const NifSimtModel& m, const NifIo& io, cudaStream_t stream) {
  const size_t smem = nifSimtSmemBytes(m);
  static size_t configured = 0;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(nif_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  // Grid: enough blocks to cover the device-side row count's upper bound, capped so that
  // blocks loop (the count itself stays on the device - no host sync inside the sample loop).
  uint32_t blocks = (io.countHost + kSimtRows - 1) / kSimtRows;
  if (blocks == 0) return cudaSuccess;
  if (blocks > 148u * 64u) blocks = 148u * 64u;
  nif_simt_kernel<<<blocks, kSimtRows, smem, stream>>>(m, io);
  return cudaGetLastError();
}

// ---- small helpers shared by the API ------------------------------------------------------
__global__ void nif_features_kernel(uint32_t n, const float* __restrict__ in, int inStride, float* __restrict__ out, int outStride, int E) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float* o = out + (size_t)i * outStride;
  nifEncode(in[(size_t)i * inStride], in[(size_t)i * inStride + 1], E, [&](int k, __half v) { o[k] = __half2float(v); });
}

cudaError_t launchNifFeatures(uint32_t n, const float* uvIn, int inStride, float* out, int outStride, int embeddingDim, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  nif_features_kernel<<<(n + 127) / 128, 128, 0, stream>>>(n, uvIn, inStride, out, outStride, embeddingDim);
  return cudaGetLastError();
}

__global__ void pack_uv_kernel(uint32_t n, const float* __restrict__ u, const float* __restrict__ v, float2* __restrict__ uv) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) uv[i] = make_float2(u[i], v[i]);
}

cudaError_t launchPackUv(uint32_t n, const float* u, const float* v, float2* uv, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  pack_uv_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, u, v, uv);
  return cudaGetLastError();
}

} // namespace b2
