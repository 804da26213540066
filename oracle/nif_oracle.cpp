// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, imported by or executed from
// the product path. See oracle_common.hpp for the usage rule.
//
// PARITY UNPINNED: the reference evaluates the NIF only as a Poplar graph (poplibs is a
// closed SDK, absent here), ships no test vectors for it and its weights blob
// (nif_models/.../converted.hdf5) is missing from the snapshot. This file restates the
// published algorithm from the reference's graph-construction code:
//   encode   src/neural_networks/NifModel.cpp:185-218  (host twin :404-432, coefficients :466-472)
//   layers   src/neural_networks/NifModel.cpp:299-326  (skip-concat autodetect :303-308)
//   decode   src/neural_networks/NifModel.cpp:221-245  (host twin :492-509; eps fold NifMetaData.cpp:48-53)
//   output   channel order B,G,R ; swapped to RGB by codelets.cpp:378
//
// Two precisions:
//   mode 0 "ref-faithful": Fourier features rounded to binary16, binary16 weights, fp32
//          accumulation, every Dense output rounded to binary16 (poplin::matMul is asked
//          for the kernel's type, NifModel.cpp:314), bias added in binary16 (:320), ReLU;
//          decode in fp32 (:226).
//   mode 1 "fp64": same network definition (binary16 features and weights are the model's
//          inputs/parameters) but all arithmetic in double with no intermediate rounding -
//          the yardstick for the error budget.

#include "oracle_common.hpp"

#include <vector>

namespace orc {

// NifModel.cpp:200-216: p = (uv - 1) * 2 ; p * 2^i cast to binary16 ; cos and sin evaluated
// on that binary16 value with a binary16 result. Feature order (:216, host twin :427-430):
// [sin(u*c_0..c_E-1), sin(v*c_..), cos(u*c_..), cos(v*c_..)].
static void encodeFeatures(float u, float v, int E, std::uint16_t* out) {
  const float pu = (u - 1.f) * 2.f;
  const float pv = (v - 1.f) * 2.f;
  float coeff = 1.f;
  for (int i = 0; i < E; ++i) {
    const float xu = roundToHalf(pu * coeff);
    const float xv = roundToHalf(pv * coeff);
    out[i] = floatToHalfBits((float)std::sin((double)xu));
    out[E + i] = floatToHalfBits((float)std::sin((double)xv));
    out[2 * E + i] = floatToHalfBits((float)std::cos((double)xu));
    out[3 * E + i] = floatToHalfBits((float)std::cos((double)xv));
    coeff *= 2.f;
  }
}

struct LayerF32 {
  int in, out, relu, hasBias;
  std::vector<float> w;   // [in,out]
  std::vector<float> b;
};

static std::vector<LayerF32> unpack(const Nif& nif) {
  std::vector<LayerF32> ls(nif.numLayers);
  for (int l = 0; l < nif.numLayers; ++l) {
    const NifLayer& src = nif.layers[l];
    LayerF32& d = ls[l];
    d.in = src.in; d.out = src.out; d.relu = src.relu; d.hasBias = src.hasBias;
    d.w.resize((std::size_t)src.in * src.out);
    for (std::size_t i = 0; i < d.w.size(); ++i) d.w[i] = halfBitsToFloat(src.kernel[i]);
    d.b.assign(src.out, 0.f);
    if (src.hasBias) for (int i = 0; i < src.out; ++i) d.b[i] = halfBitsToFloat(src.bias[i]);
  }
  return ls;
}

template <class Acc, bool RoundHalf>
static void forwardOne(const std::vector<LayerF32>& ls, const Nif& nif, const std::uint16_t* feat, int nFeat,
                       std::vector<Acc>& x, std::vector<Acc>& y, float* bgr) {
  x.assign(nFeat, Acc(0));
  for (int i = 0; i < nFeat; ++i) x[i] = (Acc)halfBitsToFloat(feat[i]);
  for (const LayerF32& L : ls) {
    if ((int)x.size() != L.in) {          // NifModel.cpp:303-308: concat activations with the features
      for (int i = 0; i < nFeat; ++i) x.push_back((Acc)halfBitsToFloat(feat[i]));
    }
    y.assign(L.out, Acc(0));
    for (int k = 0; k < L.in; ++k) {
      const Acc xv = x[k];
      if (xv == Acc(0)) continue;
      const float* wrow = L.w.data() + (std::size_t)k * L.out;
      for (int j = 0; j < L.out; ++j) y[j] += xv * (Acc)wrow[j];
    }
    for (int j = 0; j < L.out; ++j) {
      Acc v = y[j];
      if (RoundHalf) v = (Acc)roundToHalf((float)v);
      if (L.hasBias) { v = v + (Acc)L.b[j]; if (RoundHalf) v = (Acc)roundToHalf((float)v); }
      if (L.relu && v < Acc(0)) v = Acc(0);
      y[j] = v;
    }
    x.swap(y);
  }
  // NifModel.cpp:226-242: fp32 x*max + mean, optional exp
  for (int c = 0; c < 3; ++c) {
    if (RoundHalf) {
      float o = (float)x[c] * nif.maxv + nif.mean[c];
      bgr[c] = nif.logToneMap ? expf(o) : o;
    } else {
      double o = (double)x[c] * (double)nif.maxv + (double)nif.mean[c];
      bgr[c] = (float)(nif.logToneMap ? std::exp(o) : o);
    }
  }
}

int nifForwardFast(const Nif* nif, int n, const float* u, const float* v, float* bgr);   // nif_fast.cpp (timing only)

} // namespace orc

using namespace orc;

extern "C" {

int orc_sizeof_nif() { return (int)sizeof(Nif); }
int orc_sizeof_nif_layer() { return (int)sizeof(NifLayer); }

// features: n x 4E binary16
void orc_nif_encode(int n, const float* u, const float* v, int E, std::uint16_t* features) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < n; ++i) encodeFeatures(u[i], v[i], E, features + (std::size_t)i * 4 * E);
}

// bgr: n x 3 fp32 in the network's channel order (B,G,R). mode 0 ref-faithful, 1 fp64,
// 2 = the batched GEMM statement of mode 0 (nif_fast.cpp; the timed CPU baseline, never a checker).
int orc_nif_forward(const Nif* nif, int n, const float* u, const float* v, int mode, float* bgr) {
  if (nif->numLayers < 1 || nif->layers[nif->numLayers - 1].out < 3) return -1;
  if (mode == 2) return nifForwardFast(nif, n, u, v, bgr);
  const int E = nif->embeddingDim;
  const int nFeat = 4 * E;
  const std::vector<LayerF32> ls = unpack(*nif);
#pragma omp parallel
  {
    std::vector<std::uint16_t> feat(nFeat);
    std::vector<float> xf, yf;
    std::vector<double> xd, yd;
#pragma omp for schedule(dynamic, 256)
    for (int i = 0; i < n; ++i) {
      encodeFeatures(u[i], v[i], E, feat.data());
      if (mode == 0) forwardOne<float, true>(ls, *nif, feat.data(), nFeat, xf, yf, bgr + 3 * (std::size_t)i);
      else forwardOne<double, false>(ls, *nif, feat.data(), nFeat, xd, yd, bgr + 3 * (std::size_t)i);
    }
  }
  return 0;
}

// The decode alone (NifModel.cpp:226-242; host twin :492-509): x * max + mean (eps already folded into mean), optional exp.
void orc_nif_decode(int n, const float* x3, float maxv, const float* mean3, int logToneMap, float* out3) {
  for (int i = 0; i < n; ++i)
    for (int c = 0; c < 3; ++c) {
      const float o = x3[3 * i + c] * maxv + mean3[c];
      out3[3 * i + c] = logToneMap ? expf(o) : o;
    }
}

// Σ (2·in·out + out) — the reference's own FLOP count (NifModel.cpp:122-134)
long long orc_nif_flops_per_eval(const Nif* nif) {
  long long f = 0;
  for (int l = 0; l < nif->numLayers; ++l) {
    f += 2LL * nif->layers[l].in * nif->layers[l].out;
    if (nif->layers[l].hasBias) f += nif->layers[l].out;
  }
  return f;
}

} // extern "C"
