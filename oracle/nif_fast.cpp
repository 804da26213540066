// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, imported by or executed from
// the product path. See oracle_common.hpp for the usage rule.
//
// TIMING ONLY: the fast CPU statement of the NIF that bench.py's `cpu_baseline` and
// `--impl reference` legs time (orc_nif_forward mode 2, orc_render nifMode 2). Parity tests keep
// using the scalar, bit-reproducible restatement in nif_oracle.cpp (modes 0/1); this file is
// cross-checked against it in tests/test_nif_oracle.py and is otherwise never a checker.
//
// Same network definition as nif_oracle.cpp (NifModel.cpp:185-245,299-326): binary16 Fourier
// features, binary16 weights, fp32 accumulation, every Dense output and the bias sum rounded to
// an 11-bit significand (nearest-even on the bit pattern: equals binary16 rounding for normal-range
// values; sub-normal and overflowing binary16 cases are not reproduced), ReLU, fp32 decode. What differs is the schedule:
//   * rays are processed in batches of `kBatch` rows so that each layer is a blocked
//     [rows x in] . [in x out] GEMM instead of a per-ray GEMV ("AVX-friendly blocked GEMM");
//   * the micro-kernel keeps an MR x 32 accumulator block in vector registers and is compiled for
//     x86-64-v4 (AVX-512), x86-64-v3 (AVX2+FMA) and baseline through target_clones, so one library
//     runs at full width on whatever host the GPU box has (FMA contraction is confined to this file);
//   * OpenMP over batches with dynamic scheduling, like the reference's CPU tracer schedules its
//     tile jobs (external/light/src/trace.cpp:260).

#include "oracle_common.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

namespace orc {

typedef float vf16 __attribute__((vector_size(64), aligned(4)));
typedef int vi16 __attribute__((vector_size(64), aligned(4)));

constexpr int kBatch = 384;   // rows per GEMM batch: X (384 x 368 fp32 = 565 KB) + one layer (471 KB) sit in L2
constexpr int kMR = 8;        // rows per micro-kernel call
constexpr int kNR = 32;       // columns per micro-kernel call (two 16-lane vectors)

#define ORC_CLONES __attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))

struct FastLayer {
  int in, out, outPad, relu, hasBias;
  std::vector<float> w;   // [in][outPad], zero padded
  std::vector<float> b;   // [outPad]
};

static std::vector<FastLayer> unpackFast(const Nif& nif) {
  std::vector<FastLayer> ls(nif.numLayers);
  for (int l = 0; l < nif.numLayers; ++l) {
    const NifLayer& src = nif.layers[l];
    FastLayer& d = ls[l];
    d.in = src.in; d.out = src.out; d.relu = src.relu; d.hasBias = src.hasBias;
    d.outPad = (src.out + kNR - 1) / kNR * kNR;
    d.w.assign((std::size_t)d.in * d.outPad, 0.f);
    for (int k = 0; k < d.in; ++k)
      for (int j = 0; j < d.out; ++j) d.w[(std::size_t)k * d.outPad + j] = halfBitsToFloat(src.kernel[(std::size_t)k * src.out + j]);
    d.b.assign(d.outPad, 0.f);
    if (src.hasBias) for (int j = 0; j < d.out; ++j) d.b[j] = halfBitsToFloat(src.bias[j]);
  }
  return ls;
}

// round to an 11-bit significand (binary16's), nearest-even, on the bit pattern (an FMA-contracted
// Veltkamp split would not round at all): add half an ulp of the 13 dropped bits (+ the kept lsb), truncate.
__attribute__((always_inline)) static inline vf16 roundSig11(vf16 x) {
  vi16 b = (vi16)x;
  b += 0xFFF + ((b >> 13) & 1);
  b &= ~0x1FFF;
  return (vf16)b;
}

// Y[r][j0..j0+31] = act( round( round(sum_k X[r][k] W[k][j]) + b[j] ) ) for r in [r0, r0+kMR)
ORC_CLONES
static void gemmBlock(const float* __restrict X, int ldx, const float* __restrict Wp, int ldw, const float* __restrict bias,
                      float* __restrict Y, int ldy, int K, int relu, int hasBias, int roundHalf) {
  vf16 acc[kMR][2];
  for (int r = 0; r < kMR; ++r) { acc[r][0] = vf16{}; acc[r][1] = vf16{}; }
  for (int k = 0; k < K; ++k) {
    const vf16 w0 = *(const vf16*)(Wp + (std::size_t)k * ldw);
    const vf16 w1 = *(const vf16*)(Wp + (std::size_t)k * ldw + 16);
#pragma GCC unroll 8
    for (int r = 0; r < kMR; ++r) {
      const float xv = X[(std::size_t)r * ldx + k];
      acc[r][0] += xv * w0;
      acc[r][1] += xv * w1;
    }
  }
  const vf16 b0 = *(const vf16*)bias, b1 = *(const vf16*)(bias + 16);
  const vf16 zero = vf16{};
  for (int r = 0; r < kMR; ++r) {
    for (int h = 0; h < 2; ++h) {
      vf16 v = acc[r][h];
      if (roundHalf) v = roundSig11(v);
      if (hasBias) { v = v + (h ? b1 : b0); if (roundHalf) v = roundSig11(v); }
      if (relu) v = v > zero ? v : zero;
      *(vf16*)(Y + (std::size_t)r * ldy + 16 * h) = v;
    }
  }
}

struct FastScratch {
  std::vector<float> x, y, feat;
  std::vector<std::uint16_t> featBits;
};

// one batch of `rows` (<= kBatch) rays: features -> layers -> decoded B,G,R
static void forwardBatch(const std::vector<FastLayer>& ls, const Nif& nif, int rows, const float* u, const float* v,
                         FastScratch& s, float* bgr) {
  const int E = nif.embeddingDim, nFeat = 4 * E;
  const int rowsPad = (rows + kMR - 1) / kMR * kMR;
  int maxW = nFeat;
  for (const FastLayer& L : ls) maxW = std::max(maxW, std::max(L.in, L.outPad + nFeat));
  const int ld = (maxW + 15) / 16 * 16;
  s.x.assign((std::size_t)rowsPad * ld, 0.f);
  s.y.assign((std::size_t)rowsPad * ld, 0.f);
  s.feat.assign((std::size_t)rowsPad * nFeat, 0.f);
  // encode (NifModel.cpp:200-216): binary16 argument, binary16 result; order [sin u, sin v, cos u, cos v]
  for (int r = 0; r < rows; ++r) {
    const float pu = (u[r] - 1.f) * 2.f, pv = (v[r] - 1.f) * 2.f;
    float coeff = 1.f;
    float* f = s.feat.data() + (std::size_t)r * nFeat;
    for (int i = 0; i < E; ++i) {
      const float xu = roundToHalf(pu * coeff), xv = roundToHalf(pv * coeff);
      f[i] = roundToHalf(sinf(xu));
      f[E + i] = roundToHalf(sinf(xv));
      f[2 * E + i] = roundToHalf(cosf(xu));
      f[3 * E + i] = roundToHalf(cosf(xv));
      coeff *= 2.f;
    }
    std::copy(f, f + nFeat, s.x.data() + (std::size_t)r * ld);
  }
  int width = nFeat;
  float* X = s.x.data();
  float* Y = s.y.data();
  for (const FastLayer& L : ls) {
    if (width != L.in) {   // NifModel.cpp:303-308: concat the features behind the activations
      for (int r = 0; r < rows; ++r) std::copy(s.feat.data() + (std::size_t)r * nFeat, s.feat.data() + (std::size_t)(r + 1) * nFeat, X + (std::size_t)r * ld + width);
      width += nFeat;
    }
    for (int r0 = 0; r0 < rowsPad; r0 += kMR)
      for (int j0 = 0; j0 < L.outPad; j0 += kNR)
        gemmBlock(X + (std::size_t)r0 * ld, ld, L.w.data() + j0, L.outPad, L.b.data() + j0, Y + (std::size_t)r0 * ld + j0, ld, L.in, L.relu, L.hasBias, 1);
    std::swap(X, Y);
    width = L.out;
  }
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < 3; ++c) {
      const float o = X[(std::size_t)r * ld + c] * nif.maxv + nif.mean[c];   // NifModel.cpp:226-242
      bgr[3 * (std::size_t)r + c] = nif.logToneMap ? expf(o) : o;
    }
}

int nifForwardFast(const Nif* nif, int n, const float* u, const float* v, float* bgr) {
  const std::vector<FastLayer> ls = unpackFast(*nif);
  const int nBatches = (n + kBatch - 1) / kBatch;
#pragma omp parallel
  {
    FastScratch s;
#pragma omp for schedule(dynamic, 1)
    for (int b = 0; b < nBatches; ++b) {
      const int r0 = b * kBatch, rows = std::min(kBatch, n - r0);
      forwardBatch(ls, *nif, rows, u + r0, v + r0, s, bgr + 3 * (std::size_t)r0);
    }
  }
  return 0;
}

} // namespace orc
