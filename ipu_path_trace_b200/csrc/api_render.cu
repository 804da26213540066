// The render programs of the C ABI: init_render_settings, setup, path_trace, read_results
// (PathTracerApp.cpp:479-483, driven by the loop at :648-776), the device film and the
// stage-level parity entry points.
#include "ctx.h"

namespace {

int requireSettings(b200pt_ctx* c) {
  if (!c->settingsSet) return c->fail(B200PT_ERR_STATE, "render settings not set (init_render_settings has not run)");
  return B200PT_OK;
}

// Folds the events of the last path_trace into the counters (called after a stream sync).
int foldCounters(b200pt_ctx* c) {
  if (!c->pending) return B200PT_OK;
  c->pending = false;
  unsigned long long st[4] = {0, 0, 0, 0};
  CU(c, cudaMemcpy(st, c->stats.p, sizeof(st), cudaMemcpyDeviceToHost));
  c->counters.paths = st[0]; c->counters.escaped = st[1]; c->counters.path_segments = st[2];
  double tr = 0.0, nf = 0.0;
  for (const auto& sp : c->spans) {
    float ms = 0.f;
    CU(c, cudaEventElapsedTime(&ms, c->evPool[sp.a], c->evPool[sp.b]));
    (sp.kind == 0 ? tr : nf) += ms;
  }
  float total = 0.f;
  CU(c, cudaEventElapsedTime(&total, c->evBegin, c->evEnd));
  c->counters.trace_ms = tr; c->counters.nif_ms = nf; c->counters.total_ms = total;
  return B200PT_OK;
}

} // namespace

extern "C" {

int b200pt_set_render_settings(b200pt_ctx* c, uint64_t seed, float aaScalePx, float fovRad, float azimuthRad,
                               uint32_t samplesPerStep) {
  if (!c) return B200PT_ERR_INVALID;
  if (samplesPerStep == 0) return c->fail(B200PT_ERR_INVALID, "samples_per_step must be > 0");
  DeviceSettings& s = c->settings;
  s.width = c->cfg.width; s.height = c->cfg.height;
  s.aaScale = hostRoundHalf(aaScalePx);                          // PathTracerApp.cpp:339-342 (half)
  const float fov = hostRoundHalf(fovRad);                       // :343-346 (half)
  s.tanHalfFov = tanf(fov / 2.f);                                // light.hpp:199
  s.aspect = (float)s.width / (float)s.height;                   // light.hpp:198
  s.azimuth = azimuthRad;
  s.refractiveIndex = hostRoundHalf(c->cfg.refractive_index);    // IpuPathTraceJob.cpp:133-138 (half)
  s.stopProb = hostRoundHalf(c->cfg.stop_prob);
  s.rouletteWeight = 1.f / (1.f - s.stopProb);                   // light.hpp:28
  s.rouletteDepth = c->cfg.roulette_depth;
  s.maxPathLength = c->cfg.max_path_length;
  s.seedLo = (uint32_t)seed; s.seedHi = (uint32_t)(seed >> 32);
  s.aaNoiseType = c->cfg.aa_noise_type;
  c->samplesPerStep = samplesPerStep;
  c->settingsSet = true;
  return B200PT_OK;
}

int b200pt_set_sample_index(b200pt_ctx* c, uint32_t next) {
  if (!c) return B200PT_ERR_INVALID;
  c->nextSample = next;
  return B200PT_OK;
}

// true only if `p` is ordinary pageable memory, which the driver has staged by the time cudaMemcpyAsync returns. Page-locked,
// managed or partly registered ranges (and anything the query cannot classify) are read later by the copy engine.
static bool isPageable(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeUnregistered;
}

int b200pt_host_alloc(void** out, size_t bytes) {
  if (!out) return B200PT_ERR_INVALID;
  *out = nullptr;
  if (bytes == 0) return B200PT_OK;
  return cudaHostAlloc(out, bytes, cudaHostAllocPortable) == cudaSuccess ? B200PT_OK : B200PT_ERR_CUDA;
}

int b200pt_host_free(void* p) {
  if (!p) return B200PT_OK;
  return cudaFreeHost(p) == cudaSuccess ? B200PT_OK : B200PT_ERR_CUDA;
}

int b200pt_setup_next(b200pt_ctx* c, const b200pt_trace_record* worklist, size_t n) {
  if (!c) return B200PT_ERR_INVALID;
  NvtxRange range("setup_next");
  if (n > c->cfg.max_worklist) return c->fail(B200PT_ERR_INVALID, "setup_next: %zu work items exceed max_worklist %u", n, c->cfg.max_worklist);
  if (n > 0 && !worklist) return c->fail(B200PT_ERR_INVALID, "setup_next: null worklist");
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, c->recStagingNext.ensure(c->cfg.max_worklist));
  // The unpack kernel that path_trace queued for the PREVIOUS staged list may not have run yet (the main stream can be
  // busy with the step before it): this upload must not overwrite the buffer it reads.
  if (c->nextConsumedRecorded) CU(c, cudaStreamWaitEvent(c->copyStream, c->evNextConsumed, 0));
  if (n > 0) CU(c, cudaMemcpyAsync(c->recStagingNext.p, worklist, n * sizeof(TraceRecord), cudaMemcpyHostToDevice, c->copyStream));
  CU(c, cudaEventRecord(c->evNextUploaded, c->copyStream));
  c->nextStaged = true;
  c->nextItems = (uint32_t)n;
  return B200PT_OK;
}

int b200pt_setup(b200pt_ctx* c, const b200pt_trace_record* worklist, size_t n) {
  if (!c) return B200PT_ERR_INVALID;
  NvtxRange range("setup");
  if (n > c->cfg.max_worklist) return c->fail(B200PT_ERR_INVALID, "setup: %zu work items exceed max_worklist %u", n, c->cfg.max_worklist);
  if (n > 0 && !worklist) return c->fail(B200PT_ERR_INVALID, "setup: null worklist");
  CU(c, cudaSetDevice(c->cfg.device));
  static_assert(sizeof(b200pt_trace_record) == sizeof(TraceRecord), "record layouts must agree");
  c->nItems = (uint32_t)n;
  c->nextStaged = false;                       // an explicit setup supersedes a staged list
  if (n == 0) return B200PT_OK;
  CU(c, cudaMemcpyAsync(c->recStaging.p, worklist, n * sizeof(TraceRecord), cudaMemcpyHostToDevice, c->stream));
  CU(c, launchUnpackRecords(c->recStaging.p, (uint32_t)n, c->pix.p, c->accum.p, c->sampleCount.p, c->pathLength.p, c->stream));
  // A pageable host buffer has been consumed when cudaMemcpyAsync returns (the driver stages it); anything else is
  // read by the copy engine later, so wait for the copy - the caller owns the buffer again when setup returns.
  if (!isPageable(worklist)) CU(c, cudaStreamSynchronize(c->stream));
  return B200PT_OK;
}

int b200pt_path_trace(b200pt_ctx* c) {
  if (!c) return B200PT_ERR_INVALID;
  int rc = requireSettings(c);
  if (rc) return rc;
  if (!c->nifLoaded) return c->fail(B200PT_ERR_STATE, "path_trace: no NIF weights uploaded");
  NvtxRange range("path_trace");
  CU(c, cudaSetDevice(c->cfg.device));
  if (c->pending) { CU(c, cudaStreamSynchronize(c->stream)); rc = foldCounters(c); if (rc) return rc; }
  if (c->nextStaged) {
    // the list staged by setup_next becomes the active one: unpack it once its upload has landed
    c->nextStaged = false;
    c->nItems = c->nextItems;
    CU(c, cudaStreamWaitEvent(c->stream, c->evNextUploaded, 0));
    if (c->nItems) CU(c, launchUnpackRecords(c->recStagingNext.p, c->nItems, c->pix.p, c->accum.p, c->sampleCount.p, c->pathLength.p, c->stream));
    CU(c, cudaEventRecord(c->evNextConsumed, c->stream));
    c->nextConsumedRecorded = true;
  }
  c->evUsed = 0; c->spans.clear();
  c->counters = b200pt_counters{};
  CU(c, cudaMemsetAsync(c->stats.p, 0, 4 * sizeof(unsigned long long), c->stream));
  CU(c, cudaEventRecord(c->evBegin, c->stream));
  uint64_t launches = 0;
  int used = 0;
  uint32_t done = 0;
  // Software pipeline over wavefronts: trace(w) fills queue w & 1 on the trace stream while the NIF
  // kernel shades queue (w - 1) & 1 on the main stream. The two kernels touch disjoint fields of the
  // work items (trace: sampleCount/pathLength, NIF: accum). Without overlap everything runs on `stream`.
  const bool ov = c->overlap;
  cudaStream_t ts = ov ? c->traceStream : c->stream;
  if (ov) CU(c, cudaStreamWaitEvent(ts, c->evBegin, 0));       // setup / film work of the main stream comes first
  // Wavefront size: the queue capacity (192 tiles per SM), shrunk towards 64 tiles per SM to keep ~32 wavefronts per step - the first
  // trace and the last NIF launch of a step have nothing to overlap with, and their share grows with the wavefront
  // (8 GPUs at C2, 138 k items each: 27 samples per launch 9.32 G samples/s, 18 -> 9.33, 13 -> 9.42, 9 -> 9.50).
  const uint32_t perLaunch = c->samplesPerLaunchMin ? std::min(c->samplesPerLaunch, std::max(c->samplesPerLaunchMin, c->samplesPerStep / 32u)) : c->samplesPerLaunch;
  for (uint32_t w = 0; done < c->samplesPerStep && c->nItems > 0; ++w) {
    const uint32_t k = std::min(perLaunch, c->samplesPerStep - done);
    const uint32_t q = ov ? (w & 1u) : 0u;
    float2* qUv = q ? c->qUvB.p : c->qUv.p;
    float4* qThr = q ? c->qThrSlotB.p : c->qThrSlot.p;
    uint32_t* qCount = c->qCount.p + q;
    cudaEvent_t e0 = c->nextEvent(), e1 = c->nextEvent(), e2 = c->nextEvent(), e3 = c->nextEvent();
    if (!e0 || !e1 || !e2 || !e3) return c->fail(B200PT_ERR_CUDA, "cudaEventCreate failed");
    const size_t i0 = c->evUsed - 4;
    if (ov && w >= 2) CU(c, cudaStreamWaitEvent(ts, c->evQueueFree[q], 0));    // the NIF kernel of wavefront w - 2 has drained this queue
    CU(c, cudaMemsetAsync(qCount, 0, sizeof(uint32_t), ts));
    CU(c, cudaEventRecord(e0, ts));
    CU(c, launchTraceWavefront(c->scene.p, c->settings, c->pix.p, c->nItems, c->nextSample + done, k, qUv, qThr,
                               qCount, c->sampleCount.p, c->pathLength.p, c->stats.p, c->accum.p, c->liveScratch.p, c->escScratch.p, c->traceVariant, ts));
    CU(c, cudaEventRecord(e1, ts));
    if (ov) {
      CU(c, cudaEventRecord(c->evTraced[q], ts));
      CU(c, cudaStreamWaitEvent(c->stream, c->evTraced[q], 0));
    }
    NifIo io{};
    io.uv = qUv; io.thrSlot = qThr; io.count = qCount;
    io.countHost = (uint32_t)std::min<size_t>((size_t)c->nItems * k, 0xffffffffu);
    io.bgrOut = nullptr; io.accum = c->accum.p;
    CU(c, cudaEventRecord(e2, c->stream));
    rc = launchNif(c, io, B200PT_NIF_AUTO, &used);
    if (rc) return rc;
    CU(c, cudaEventRecord(e3, c->stream));
    if (ov) CU(c, cudaEventRecord(c->evQueueFree[q], c->stream));
    c->spans.push_back({i0, i0 + 1, 0});
    c->spans.push_back({i0 + 2, i0 + 3, 1});
    launches += 2;
    done += k;
  }
  CU(c, cudaEventRecord(c->evEnd, c->stream));
  c->nextSample += c->samplesPerStep;
  c->counters.kernel_launches = launches;
  c->counters.nif_impl_used = used;
  c->pending = true;
  return B200PT_OK;
}

int b200pt_sync(b200pt_ctx* c) {
  if (!c) return B200PT_ERR_INVALID;
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, cudaStreamSynchronize(c->stream));
  return foldCounters(c);
}

int b200pt_get_counters(b200pt_ctx* c, b200pt_counters* out) {
  if (!c || !out) return B200PT_ERR_INVALID;
  int rc = b200pt_sync(c);
  if (rc) return rc;
  *out = c->counters;
  return B200PT_OK;
}

int b200pt_read_results(b200pt_ctx* c, b200pt_trace_record* out, size_t n, b200pt_counters* counters) {
  if (!c) return B200PT_ERR_INVALID;
  if (n > c->nItems) return c->fail(B200PT_ERR_INVALID, "read_results: %zu items requested, %u uploaded", n, c->nItems);
  if (n > 0 && !out) return c->fail(B200PT_ERR_INVALID, "read_results: null buffer");
  NvtxRange range("read_results");
  CU(c, cudaSetDevice(c->cfg.device));
  if (n > 0) {
    CU(c, launchPackRecords(c->recStaging.p, (uint32_t)n, c->pix.p, c->accum.p, c->sampleCount.p, c->pathLength.p, c->stream));
    CU(c, cudaMemcpyAsync(out, c->recStaging.p, n * sizeof(TraceRecord), cudaMemcpyDeviceToHost, c->stream));
  }
  int rc = b200pt_sync(c);
  if (rc) return rc;
  if (counters) *counters = c->counters;
  return B200PT_OK;
}

// ---- film ----------------------------------------------------------------------------------

int b200pt_film_accumulate(b200pt_ctx* c) {
  if (!c) return B200PT_ERR_INVALID;
  NvtxRange range("accumulate_framebuffers");
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, launchFilmAccumulate(c->nItems, c->pix.p, c->accum.p, c->sampleCount.p, c->pathLength.p, c->film.p, c->cfg.width,
                             c->cfg.height, c->stream));
  return B200PT_OK;
}

int b200pt_film_clear(b200pt_ctx* c) {
  if (!c) return B200PT_ERR_INVALID;
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, cudaMemsetAsync(c->film.p, 0, c->film.n * sizeof(float), c->stream));
  return B200PT_OK;
}

int b200pt_film_read_device(b200pt_ctx* c, void* bgrDev, float scale) {
  if (!c || !bgrDev) return B200PT_ERR_INVALID;
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, launchFilmScale(c->film.p, static_cast<float*>(bgrDev), c->film.n, scale, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return foldCounters(c);
}

int b200pt_film_read(b200pt_ctx* c, float* bgrHost, float scale) {
  if (!c || !bgrHost) return B200PT_ERR_INVALID;
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, launchFilmScale(c->film.p, c->filmOut.p, c->film.n, scale, c->stream));
  CU(c, cudaMemcpyAsync(bgrHost, c->filmOut.p, c->film.n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return foldCounters(c);
}

int b200pt_film_tonemap(b200pt_ctx* c, uint8_t* bgr8Host, float scale, float exposure, float gamma) {
  if (!c || !bgr8Host) return B200PT_ERR_INVALID;
  if (!(gamma > 0.f)) return c->fail(B200PT_ERR_INVALID, "tonemap: gamma must be > 0");
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, launchFilmTonemap(c->film.p, c->film8.p, c->film.n, scale, exposure, gamma, c->stream));
  CU(c, cudaMemcpyAsync(bgr8Host, c->film8.p, c->film.n, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return foldCounters(c);
}

// ---- stage-level entry points ----------------------------------------------------------------

int b200pt_trace_wavefront(b200pt_ctx* c, uint32_t sample, size_t n, int32_t* escaped, float* uv2, float* thr3,
                           int32_t* length, uint16_t* cam2) {
  if (!c) return B200PT_ERR_INVALID;
  int rc = requireSettings(c);
  if (rc) return rc;
  if (n > c->nItems) return c->fail(B200PT_ERR_INVALID, "trace_wavefront: %zu items requested, %u uploaded", n, c->nItems);
  if (n == 0) return B200PT_OK;
  CU(c, cudaSetDevice(c->cfg.device));
  DevBuf<unsigned char>&dE = c->scratch[0], &dUv = c->scratch[1], &dT = c->scratch[2], &dL = c->scratch[3], &dC = c->scratch[4];
  CU(c, dE.ensure(n * 4)); CU(c, dUv.ensure(n * 8)); CU(c, dT.ensure(n * 12)); CU(c, dL.ensure(n * 4)); CU(c, dC.ensure(n * 4));
  CU(c, launchTraceItems(c->scene.p, c->settings, c->pix.p, (uint32_t)n, sample, reinterpret_cast<int32_t*>(dE.p),
                         reinterpret_cast<float*>(dUv.p), reinterpret_cast<float*>(dT.p), reinterpret_cast<int32_t*>(dL.p),
                         reinterpret_cast<uint16_t*>(dC.p), c->stream));
  if (escaped) CU(c, cudaMemcpyAsync(escaped, dE.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  if (uv2) CU(c, cudaMemcpyAsync(uv2, dUv.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
  if (thr3) CU(c, cudaMemcpyAsync(thr3, dT.p, n * 12, cudaMemcpyDeviceToHost, c->stream));
  if (length) CU(c, cudaMemcpyAsync(length, dL.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  if (cam2) CU(c, cudaMemcpyAsync(cam2, dC.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return B200PT_OK;
}

int b200pt_trace_fixed(b200pt_ctx* c, size_t n, const uint16_t* cam2, const uint16_t* uniforms, uint32_t perPath,
                       int32_t* escaped, float* dir3, float* thr3, int32_t* length) {
  if (!c) return B200PT_ERR_INVALID;
  int rc = requireSettings(c);
  if (rc) return rc;
  if (n == 0) return B200PT_OK;
  if (!cam2 || !uniforms || perPath == 0 || !escaped || !dir3 || !thr3 || !length)
    return c->fail(B200PT_ERR_INVALID, "trace_fixed: null buffer or per_path == 0");
  if (n > 0xffffffffu) return c->fail(B200PT_ERR_INVALID, "trace_fixed: too many rays");
  CU(c, cudaSetDevice(c->cfg.device));
  DevBuf<unsigned char>&dC = c->scratch[0], &dU = c->scratch[1], &dE = c->scratch[2], &dD = c->scratch[3], &dT = c->scratch[4], &dL = c->scratch[5];
  CU(c, dC.ensure(n * 4)); CU(c, dU.ensure(n * perPath * 2)); CU(c, dE.ensure(n * 4)); CU(c, dD.ensure(n * 12));
  CU(c, dT.ensure(n * 12)); CU(c, dL.ensure(n * 4));
  CU(c, cudaMemcpyAsync(dC.p, cam2, n * 4, cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaMemcpyAsync(dU.p, uniforms, n * perPath * 2, cudaMemcpyHostToDevice, c->stream));
  CU(c, launchTraceFixed(c->scene.p, c->settings, (uint32_t)n, reinterpret_cast<uint16_t*>(dC.p), reinterpret_cast<uint16_t*>(dU.p), perPath,
                         reinterpret_cast<int32_t*>(dE.p), reinterpret_cast<float*>(dD.p), reinterpret_cast<float*>(dT.p),
                         reinterpret_cast<int32_t*>(dL.p), c->stream));
  CU(c, cudaMemcpyAsync(escaped, dE.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(dir3, dD.p, n * 12, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(thr3, dT.p, n * 12, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(length, dL.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return B200PT_OK;
}

int b200pt_kat(b200pt_ctx* c, int op, size_t n, const float* in, int inStride, float* out, int outStride) {
  if (!c) return B200PT_ERR_INVALID;
  if (n == 0) return B200PT_OK;
  if (!in || !out || inStride < 1 || outStride < 1 || n > 0xffffffffu) return c->fail(B200PT_ERR_INVALID, "kat: bad arguments");
  static const int needIn[] = {5, 10, 13, 6, 6, 8, 2, 4, 6, 3, 2};
  if (op < 0 || op > B200PT_KAT_NIF_FEATURES) return c->fail(B200PT_ERR_INVALID, "kat: unknown op %d", op);
  int needOut[] = {3, 1, 1, 4, 3, 4, 2, 2, 4, 2, 0};
  if (op == B200PT_KAT_NIF_FEATURES) {
    if (!c->nifLoaded) return c->fail(B200PT_ERR_STATE, "kat: NIF features need an uploaded model (embedding dimension)");
    needOut[op] = c->simt.nFeat;
  }
  if (inStride < needIn[op] || outStride < needOut[op]) return c->fail(B200PT_ERR_INVALID, "kat: strides too small for op %d", op);
  CU(c, cudaSetDevice(c->cfg.device));
  DevBuf<unsigned char>&dIn = c->scratch[0], &dOut = c->scratch[1];
  CU(c, dIn.ensure(n * inStride * 4)); CU(c, dOut.ensure(n * outStride * 4));
  CU(c, cudaMemcpyAsync(dIn.p, in, n * inStride * 4, cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaMemsetAsync(dOut.p, 0, n * outStride * 4, c->stream));
  if (op == B200PT_KAT_NIF_FEATURES)
    CU(c, launchNifFeatures((uint32_t)n, reinterpret_cast<float*>(dIn.p), inStride, reinterpret_cast<float*>(dOut.p), outStride,
                            c->simt.embeddingDim, c->stream));
  else
    CU(c, launchKat(op, (uint32_t)n, reinterpret_cast<float*>(dIn.p), inStride, reinterpret_cast<float*>(dOut.p), outStride, c->stream));
  CU(c, cudaMemcpyAsync(out, dOut.p, n * outStride * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return B200PT_OK;
}

} // extern "C"
