// "init_nif_weights" and the NIF inference entry points of the C ABI.
//
// b200pt_upload_nif takes the Dense stack exactly as the reference's loader holds it
// (NifModel::Data, NifModel.cpp:40-85: per layer a Keras [in,out] kernel and optional bias in
// binary16 or binary32, activation renamed to relu/none :73-76) and prepares two device forms:
//   * fp32 copies for the general SIMT kernel (any topology);
//   * for the shipped topology (hidden width H, optional skip-concat of the features, final layer
//     of <= 4 outputs, padded to N = 16) the weight stream of the tcgen05 kernel: every (layer, N-half, K-range)
//     chunk is stored as the exact shared-memory image of a K-major UMMA B operand, in MMA issue
//     order, so the kernel's producer warp only issues linear bulk copies.
#include "ctx.h"

namespace {

struct HostLayer {
  int in = 0, out = 0, relu = 0, hasBias = 0;
  std::vector<uint16_t> w;   // [in,out] binary16 bits
  std::vector<uint16_t> b;
};

// Why the tensor-core path cannot take a model ("" = it can).
std::string tcEligibility(const std::vector<HostLayer>& ls, int nFeat) {
  if (ls.size() < 2) return "needs at least one hidden layer and a final layer";
  if ((int)ls.size() - 1 > kMaxNifLayers) return "too many layers";
  const int H = ls[0].out;
  if (H % 64 != 0 || H < 64 || H > 320) return "hidden width must be a multiple of 64 in [64,320]";
  if (nFeat > kMaxFeatures) return "more than 64 input features";
  if (ls[0].in != nFeat) return "first layer does not take the features";
  for (size_t l = 1; l + 1 < ls.size(); ++l) {
    if (ls[l].out != H) return "hidden layers of different widths";
    if (ls[l].in != H && ls[l].in != H + nFeat) return "hidden layer input is neither H nor H + features";
  }
  const HostLayer& last = ls.back();
  if (last.in != H) return "final layer must read the last hidden layer only";
  if (last.out < 3 || last.out > 4) return "final layer must have 3 or 4 outputs";
  return "";
}

// Builds the tcgen05 kernel's model: the chunk table and the packed weight stream. Every (layer,
// N-range, K-range) chunk is the exact shared-memory image of a K-major UMMA B operand, in MMA issue
// order; for a CTA pair (cg == 2) a chunk is stored as [image of CTA 0's N/2 rows | image of CTA 1's].
bool buildTcModel(const std::vector<HostLayer>& ls, int nFeat, int E, const b200pt_nif_desc& nif, int cg, int ts, int kc, int mc, NifTcModel& tm,
                  std::vector<uint8_t>& stream) {
  std::memset(&tm, 0, sizeof(tm));
  tm.cg = cg; tm.ts = ts; tm.kc = kc; tm.mc = mc;
  const int numSlots = ts ? 2 : 3;
  const int numHidden = (int)ls.size() - 1;
  const int H = ls[0].out, NH = H / 2;
  const int featKSteps = (nFeat + 15) / 16;
  const int halfKSteps = NH / 16;
  tm.numHidden = numHidden; tm.H = H; tm.nFeat = nFeat; tm.embeddingDim = E; tm.featKSteps = featKSteps;
  tm.maxv = nif.max; tm.mean[0] = nif.mean[0]; tm.mean[1] = nif.mean[1]; tm.mean[2] = nif.mean[2];
  tm.logToneMap = nif.log_tone_map ? 1 : 0;
  int nChunks = 0;
  bool overflow = false;
  // Appends the chunks of one (layer, N-range, K-range) block of a job. N = `nCols` output columns
  // starting at `col0` (columns >= L.out are zero padding); kstepBegin is the first K-step inside
  // the A buffer (features or activations), rowOfK maps a K index of that buffer to a row of the
  // Keras kernel (-1 = zero pad).
  auto addBlock = [&](int layer, int half, int col0, int nCols, bool srcFeat, int kstepBegin, int kstepCount, auto rowOfK,
                      bool firstOfJob, bool lastOfJob, int waitAct, bool isOut) {
    for (int ks0 = 0; ks0 < kstepCount; ks0 += kc) {
      if (nChunks == kTcMaxChunks) { overflow = true; return; }
      const int ks = std::min(kc, kstepCount - ks0);
      NifTcChunk& ch = tm.chunks[nChunks++];
      std::memset(&ch, 0, sizeof(ch));
      ch.offset = (uint32_t)stream.size();
      const size_t bytes = (size_t)ks * nCols * 32;
      ch.bytes16 = (uint16_t)(bytes / 16);
      ch.aOff16 = (uint16_t)((kstepBegin + ks0) * 4096 / 16);      // one K-step of A = 128 rows x 32 B
      ch.ksteps = (uint8_t)ks; ch.n8 = (uint8_t)(nCols / 8);
      ch.flags = (uint8_t)(((firstOfJob && ks0 == 0) ? kChFirst : 0) | ((lastOfJob && ks0 + ks == kstepCount) ? kChLast : 0) |
                           (srcFeat ? kChSrcFeat : 0) | (isOut ? kChOut : 0));
      ch.waitAct = (ks0 == 0) ? (uint8_t)waitAct : 0;
      ch.layer = (uint8_t)layer; ch.half = (uint8_t)half;
      ch.slot = (uint8_t)((2 * layer + half) % numSlots); ch.slotUse = (uint8_t)((2 * layer + half) / numSlots);
      stream.resize(stream.size() + bytes, 0);
      uint16_t* dst = reinterpret_cast<uint16_t*>(stream.data() + ch.offset);
      const HostLayer& L = ls[layer];
      // K-major UMMA B operand, no swizzle: per K-step two K-cores of [rows x 8 halfs];
      // LBO (K-core to K-core) = rows * 16 B, SBO (8-row group to group) = 128 B.
      const int rows = nCols / cg;                                  // B rows (output columns) held by one CTA
      for (int rk = 0; rk < cg; ++rk)
        for (int j = 0; j < ks; ++j)
          for (int kc = 0; kc < 2; ++kc)
            for (int nn = 0; nn < rows; ++nn)
              for (int ke = 0; ke < 8; ++ke) {
                const int k = (kstepBegin + ks0 + j) * 16 + kc * 8 + ke;     // K index inside the A buffer
                const int row = rowOfK(k);
                const int col = col0 + rk * rows + nn;
                const uint16_t v = (row < 0 || col >= L.out) ? (uint16_t)0 : L.w[(size_t)row * L.out + col];
                dst[(size_t)rk * ks * rows * 16 + ((size_t)j * 2 + kc) * rows * 8 + (size_t)nn * 8 + ke] = v;
              }
    }
  };
  int lastFeatChunk = -1;
  for (int l = 0; l < numHidden && !overflow; ++l) {
    const HostLayer& L = ls[l];
    const bool usesAct = (l > 0);
    const bool usesFeat = (l == 0) || (L.in == H + nFeat);
    tm.relu[l] = L.relu;
    // concat order is [activations, features] (NifModel.cpp:305): feature k -> kernel row H + k
    const int featRow0 = usesAct ? H : 0;
    auto featRow = [&](int k) { return k < nFeat ? featRow0 + k : -1; };
    auto actRow = [&](int k) { return k; };
    if (usesFeat) {
      for (int half = 0; half < 2; ++half)
        addBlock(l, half, half * NH, NH, true, 0, featKSteps, featRow, true, !usesAct, 0, false);
      lastFeatChunk = nChunks - 1;
    }
    if (usesAct) {
      // K order: columns [0,H/2) for both halves, then [H/2,H) for both halves, so the layer can
      // start as soon as the first half of the previous layer's epilogue has been written.
      for (int part = 0; part < 2; ++part)
        for (int half = 0; half < 2; ++half)
          addBlock(l, half, half * NH, NH, false, part * halfKSteps, halfKSteps, actRow, !usesFeat && part == 0, part == 1,
                   half == 0 ? part + 1 : 0, false);
    }
  }
  if (!overflow) {
    // final Dense layer (<= 4 outputs) as one N = 16 job over the last hidden activations
    auto actRow = [&](int k) { return k; };
    for (int part = 0; part < 2; ++part)
      addBlock(numHidden, 0, 0, 16, false, part * halfKSteps, halfKSteps, actRow, part == 0, part == 1, part + 1, true);
  }
  if (overflow) return false;
  tm.chunks[lastFeatChunk].flags |= kChLastFeatUse;
  tm.numChunks = nChunks;
  for (int j = 0; j < 2 * numHidden; ++j) tm.slotUses[j % numSlots] += 1;
  const HostLayer& last = ls.back();
  tm.outC = last.out; tm.finalRelu = last.relu; tm.finalHasBias = last.hasBias;
  for (int j = 0; j < 4; ++j) tm.bOut[j] = j < last.out ? hostHalfToFloat(last.b[j]) : 0.f;
  return true;
}

} // namespace

extern "C" {

int b200pt_upload_nif(b200pt_ctx* c, const b200pt_nif_desc* nif) {
  if (!c) return B200PT_ERR_INVALID;
  if (!nif || !nif->layers || nif->num_layers < 1 || nif->num_layers > kMaxNifLayers)
    return c->fail(B200PT_ERR_INVALID, "upload_nif: bad descriptor (1..%d layers)", kMaxNifLayers);
  const int E = nif->embedding_dimension;
  if (E < 1 || 4 * E > kMaxFeatures) return c->fail(B200PT_ERR_INVALID, "upload_nif: embedding_dimension must be 1..%d", kMaxFeatures / 4);
  const int nFeat = 4 * E;
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, cudaStreamSynchronize(c->stream));      // weights may be replaced mid-render (PathTracerApp.cpp:548-557)

  // ---- validate + convert to binary16 bits on the host ----
  std::vector<HostLayer> ls(nif->num_layers);
  int width = nFeat, maxWidth = nFeat;
  int64_t flops = 0;
  for (int l = 0; l < nif->num_layers; ++l) {
    const b200pt_nif_layer& src = nif->layers[l];
    HostLayer& d = ls[l];
    if (src.in < 1 || src.out < 1 || !src.kernel) return c->fail(B200PT_ERR_INVALID, "upload_nif: layer %d is empty", l);
    if (src.dtype != 0 && src.dtype != 1) return c->fail(B200PT_ERR_INVALID, "upload_nif: layer %d dtype must be 0 (f16) or 1 (f32)", l);
    // NifModel.cpp:303-308: when shapes disagree the features are concatenated to the input
    if (src.in != width && src.in != width + nFeat)
      return c->fail(B200PT_ERR_INVALID, "upload_nif: layer %d expects %d inputs, previous layer gives %d (+%d features)", l, src.in, width, nFeat);
    d.in = src.in; d.out = src.out; d.relu = src.relu ? 1 : 0; d.hasBias = src.bias ? 1 : 0;
    const size_t nw = (size_t)src.in * src.out;
    d.w.resize(nw);
    if (src.dtype == 0) std::memcpy(d.w.data(), src.kernel, nw * 2);
    else for (size_t i = 0; i < nw; ++i) d.w[i] = hostFloatToHalf(static_cast<const float*>(src.kernel)[i]);
    d.b.assign(src.out, 0);
    if (src.bias) {
      if (src.dtype == 0) std::memcpy(d.b.data(), src.bias, (size_t)src.out * 2);
      else for (int i = 0; i < src.out; ++i) d.b[i] = hostFloatToHalf(static_cast<const float*>(src.bias)[i]);
    }
    flops += 2LL * src.in * src.out + (src.bias ? src.out : 0);     // NifModel.cpp:122-134
    width = src.out;
    maxWidth = std::max(maxWidth, std::max(src.in, src.out));
  }
  if (ls.back().out < 3) return c->fail(B200PT_ERR_INVALID, "upload_nif: the last layer must produce at least 3 channels");

  // ---- SIMT form ----
  NifSimtModel& sm = c->simt;
  std::memset(&sm, 0, sizeof(sm));
  sm.numLayers = nif->num_layers; sm.nFeat = nFeat; sm.embeddingDim = E; sm.maxWidth = maxWidth;
  sm.maxv = nif->max; sm.mean[0] = nif->mean[0]; sm.mean[1] = nif->mean[1]; sm.mean[2] = nif->mean[2];
  sm.logToneMap = nif->log_tone_map ? 1 : 0;
  size_t total = 0;
  std::vector<size_t> wOff(ls.size()), bOff(ls.size());
  for (size_t l = 0; l < ls.size(); ++l) {
    wOff[l] = total; total += (ls[l].w.size() + 63) & ~(size_t)63;
    bOff[l] = total; total += ((size_t)ls[l].out + 63) & ~(size_t)63;
  }
  std::vector<float> params(total, 0.f);
  width = nFeat;
  for (size_t l = 0; l < ls.size(); ++l) {
    for (size_t i = 0; i < ls[l].w.size(); ++i) params[wOff[l] + i] = hostHalfToFloat(ls[l].w[i]);
    for (int i = 0; i < ls[l].out; ++i) params[bOff[l] + i] = hostHalfToFloat(ls[l].b[i]);
    sm.in[l] = ls[l].in; sm.out[l] = ls[l].out; sm.relu[l] = ls[l].relu; sm.hasBias[l] = ls[l].hasBias;
    sm.concat[l] = (ls[l].in != width) ? 1 : 0;
    width = ls[l].out;
  }
  if (nifSimtSmemBytes(sm) > 200 * 1024) return c->fail(B200PT_ERR_INVALID, "upload_nif: layers wider than the SIMT kernel's shared memory allows");
  CU(c, c->simtParams.alloc(total));
  CU(c, cudaMemcpy(c->simtParams.p, params.data(), total * sizeof(float), cudaMemcpyHostToDevice));
  for (size_t l = 0; l < ls.size(); ++l) { sm.w[l] = c->simtParams.p + wOff[l]; sm.b[l] = c->simtParams.p + bOff[l]; }

  // ---- tcgen05 form ----
  c->tcWhyNot = tcEligibility(ls, nFeat);
  c->tcOk = c->tcWhyNot.empty();
  if (c->tcOk) {
    const int numHidden = (int)ls.size() - 1;
    const int H = ls[0].out;
    std::vector<uint16_t> bias((size_t)numHidden * H, 0);
    for (int l = 0; l < numHidden; ++l) for (int i = 0; i < H; ++i) bias[(size_t)l * H + i] = ls[l].b[i];
    CU(c, c->tcBias.alloc(bias.size()));
    CU(c, cudaMemcpy(c->tcBias.p, bias.data(), bias.size() * 2, cudaMemcpyHostToDevice));
    // Feature table. The encoder's argument of octave i is binary16(p * 2^i) (NifModel.cpp:211-213) and rounding to
    // binary16 commutes with the power of two, so all 4E features of a coordinate are a function of the 16 bits of
    // binary16(p): one row of E (sin, cos) pairs per bit pattern, evaluated here with the host libm exactly as the
    // CPU twin does (NifModel.cpp:404-432). 65536 x E x 4 bytes (3 MB for E = 12), L2 resident.
    {
      std::vector<uint32_t> table((size_t)65536 * E);
      for (uint32_t bits = 0; bits < 65536u; ++bits) {
        const float x0 = hostHalfToFloat((uint16_t)bits);
        float coeff = 1.f;
        for (int i = 0; i < E; ++i, coeff *= 2.f) {
          const float x = hostRoundHalf(x0 * coeff);
          const uint16_t sb = hostFloatToHalf(sinf(x)), cb = hostFloatToHalf(cosf(x));
          table[(size_t)bits * E + i] = (uint32_t)sb | ((uint32_t)cb << 16);
        }
      }
      CU(c, c->tcFeatTable.alloc(table.size()));
      CU(c, cudaMemcpy(c->tcFeatTable.p, table.data(), table.size() * 4, cudaMemcpyHostToDevice));
    }
    // Kernel variant. Default (mc10): activations in tensor memory, 10 K-steps per chunk, weight chunks
    // multicast between the two CTAs of a cluster. B200PT_NIF_MODE = ss1 | ss2 | ts5 | ts10 | tp5 | tp10 |
    // mc5 | mc10 selects another one (A/B measurements and the parity tests of each variant).
    int cg = 1, ts = 1, kc = 10, mc = 1;
    if (const char* e = getenv("B200PT_NIF_MODE")) {
      const std::string mode(e);
      if (mode == "ss1") { cg = 1; ts = 0; kc = 5; mc = 0; }
      else if (mode == "ss2") { cg = 2; ts = 0; kc = 5; mc = 0; }
      else if (mode == "ts5") { cg = 1; ts = 1; kc = 5; mc = 0; }
      else if (mode == "ts10") { cg = 1; ts = 1; kc = 10; mc = 0; }
      else if (mode == "tp5") { cg = 2; ts = 1; kc = 5; mc = 0; }
      else if (mode == "tp10") { cg = 2; ts = 1; kc = 10; mc = 0; }
      else if (mode == "mc10") { cg = 1; ts = 1; kc = 10; mc = 1; }
      else if (mode == "mc5") { cg = 1; ts = 1; kc = 5; mc = 1; }
      else return c->fail(B200PT_ERR_INVALID, "B200PT_NIF_MODE must be ss1, ss2, ts5, ts10, tp5, tp10, mc5 or mc10");
    }
    NifTcModel& tm = c->tcHost;
    std::vector<uint8_t> stream;
    if (!buildTcModel(ls, nFeat, E, *nif, cg, ts, kc, mc, tm, stream)) { c->tcOk = false; c->tcWhyNot = "too many weight chunks"; }
    else {
      CU(c, c->tcStream.alloc(stream.size()));
      CU(c, cudaMemcpy(c->tcStream.p, stream.data(), stream.size(), cudaMemcpyHostToDevice));
      tm.wStream = c->tcStream.p; tm.bias = c->tcBias.p; tm.featTable = c->tcFeatTable.p;
      if (nifTcSmemBytes(tm) > 227 * 1024 || nifTcStages(tm) < 2) { c->tcOk = false; c->tcWhyNot = "tcgen05 kernel shared memory exceeds 227 KB"; }
      else {
        CU(c, prepareNifTc(tm));
        c->tcMaxPairs = (cg == 2 || mc) ? nifTcMaxPairs(tm) : 0;
        if ((cg == 2 || mc) && c->tcMaxPairs < 1) { c->tcOk = false; c->tcWhyNot = "no co-resident CTA pairs on this device"; }
      }
    }
  }
  if (c->cfg.nif_impl == B200PT_NIF_TCGEN05 && !c->tcOk)
    return c->fail(B200PT_ERR_INVALID, "upload_nif: tcgen05 path requested but the model does not fit it: %s", c->tcWhyNot.c_str());
  c->nifFlops = flops;
  c->nifLoaded = true;
  return B200PT_OK;
}

int64_t b200pt_nif_flops_per_eval(const b200pt_ctx* c) { return (c && c->nifLoaded) ? c->nifFlops : 0; }

} // extern "C"

// Chooses and launches the NIF kernel for `io`. impl: B200PT_NIF_*.
int launchNif(b200pt_ctx* c, const NifIo& io, int impl, int* used) {
  if (!c->nifLoaded) return c->fail(B200PT_ERR_STATE, "no NIF weights uploaded (init_nif_weights has not run)");
  if (impl == B200PT_NIF_AUTO) impl = c->cfg.nif_impl;
  if (impl == B200PT_NIF_AUTO) impl = c->tcOk ? B200PT_NIF_TCGEN05 : B200PT_NIF_SIMT;
  if (impl == B200PT_NIF_TCGEN05) {
    if (!c->tcOk) return c->fail(B200PT_ERR_INVALID, "tcgen05 NIF path unavailable for this model: %s", c->tcWhyNot.c_str());
    CU(c, launchNifTc(c->tcHost, io, c->numSms, c->tcMaxPairs, c->stream));
  } else if (impl == B200PT_NIF_SIMT) {
    CU(c, launchNifSimt(c->simt, io, c->stream));
  } else {
    return c->fail(B200PT_ERR_INVALID, "unknown NIF implementation %d", impl);
  }
  if (used) *used = impl;
  return B200PT_OK;
}

extern "C" {

int b200pt_nif_infer_device(b200pt_ctx* c, size_t n, const void* uDev, const void* vDev, void* bgrDev, int impl,
                            int repeats, double* deviceMs) {
  if (!c) return B200PT_ERR_INVALID;
  if (n > 0xffffffffu) return c->fail(B200PT_ERR_INVALID, "nif_infer: too many rows");
  CU(c, cudaSetDevice(c->cfg.device));
  if (n == 0) { if (deviceMs) *deviceMs = 0.0; return B200PT_OK; }
  if (repeats < 1) repeats = 1;
  DevBuf<unsigned char>& uvBuf = c->scratch[0];
  CU(c, uvBuf.ensure(n * sizeof(float2)));
  float2* uv = reinterpret_cast<float2*>(uvBuf.p);
  CU(c, launchPackUv((uint32_t)n, static_cast<const float*>(uDev), static_cast<const float*>(vDev), uv, c->stream));
  NifIo io{};
  io.uv = uv; io.thrSlot = nullptr; io.count = nullptr; io.countHost = (uint32_t)n;
  io.bgrOut = static_cast<float*>(bgrDev); io.accum = nullptr;
  CU(c, cudaEventRecord(c->evBegin, c->stream));
  for (int r = 0; r < repeats; ++r) {
    int rc = launchNif(c, io, impl, nullptr);
    if (rc) return rc;
  }
  CU(c, cudaEventRecord(c->evEnd, c->stream));
  if (deviceMs) {
    CU(c, cudaEventSynchronize(c->evEnd));
    float ms = 0.f;
    CU(c, cudaEventElapsedTime(&ms, c->evBegin, c->evEnd));
    *deviceMs = ms;
  }
  return B200PT_OK;
}

int b200pt_nif_infer(b200pt_ctx* c, size_t n, const float* u, const float* v, float* bgr, int impl) {
  if (!c) return B200PT_ERR_INVALID;
  if (n == 0) return B200PT_OK;
  if (!u || !v || !bgr) return c->fail(B200PT_ERR_INVALID, "nif_infer: null buffer");
  CU(c, cudaSetDevice(c->cfg.device));
  DevBuf<unsigned char>&du = c->scratch[1], &dv = c->scratch[2], &dout = c->scratch[3];
  CU(c, du.ensure(n * 4)); CU(c, dv.ensure(n * 4)); CU(c, dout.ensure(n * 12));
  CU(c, cudaMemcpyAsync(du.p, u, n * 4, cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaMemcpyAsync(dv.p, v, n * 4, cudaMemcpyHostToDevice, c->stream));
  int rc = b200pt_nif_infer_device(c, n, du.p, dv.p, dout.p, impl, 1, nullptr);
  if (rc) return rc;
  CU(c, cudaMemcpyAsync(bgr, dout.p, n * 12, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return B200PT_OK;
}

} // extern "C"
