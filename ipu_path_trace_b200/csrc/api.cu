// libb200pt C ABI (include/b200pt.h): context, the five programs of the reference
// (PathTracerApp.cpp:479-483) and the stage-level entry points the parity tests call.
// Host code only; kernels live in trace_kernels.cu, nif_simt.cu, nif_tc.cu, film_kernels.cu.
#include "ctx.h"

namespace {
thread_local std::string g_createError;
}

namespace {

// The hard-coded scene of the reference (codelets.cpp:110-144): five spheres and a disc, colours
// carry the 2x gain (:127-130), no emitters (:117,142). Double-typed literals of the reference
// (-3.23, -0.19931) narrow to float exactly as (float)literal does.
int appendPrimitive(SceneSoA& sc, const b200pt_primitive& p) {
  const int n = sc.count;
  if (n >= kMaxPrims) return -1;
  sc.shape[n] = p.shape; sc.bsdf[n] = p.material;
  sc.cx[n] = p.centre[0]; sc.cy[n] = p.centre[1]; sc.cz[n] = p.centre[2];
  sc.nx[n] = 0.f; sc.ny[n] = 1.f; sc.nz[n] = 0.f;
  sc.radius2[n] = p.radius * p.radius; sc.planeD[n] = 0.f;
  if (p.shape == 1) {   // Disc ctor normalises n, d = |c.n| (light.hpp:98-101)
    const float nx = p.normal[0], ny = p.normal[1], nz = p.normal[2];
    const float inv = 1.f / std::sqrt(nx * nx + ny * ny + nz * nz);
    const float ux = nx * inv, uy = ny * inv, uz = nz * inv;
    sc.nx[n] = ux; sc.ny[n] = uy; sc.nz[n] = uz;
    sc.planeD[n] = std::fabs(p.centre[0] * ux + p.centre[1] * uy + p.centre[2] * uz);
  }
  sc.colr[n] = p.colour[0]; sc.colg[n] = p.colour[1]; sc.colb[n] = p.colour[2];
  sc.emr[n] = p.emission[0]; sc.emg[n] = p.emission[1]; sc.emb[n] = p.emission[2];
  sc.emissive[n] = (p.emission[0] != 0.f || p.emission[1] != 0.f || p.emission[2] != 0.f) ? 1 : 0;
  sc.count = n + 1;
  return 0;
}

void buildScene(SceneSoA& sc) {
  std::memset(&sc, 0, sizeof(sc));
  auto prim = [&](int shape, float cx, float cy, float cz, float r, float cr, float cg, float cb, int bsdf) {
    b200pt_primitive p{};
    p.shape = shape; p.material = bsdf;
    p.centre[0] = cx; p.centre[1] = cy; p.centre[2] = cz;
    p.normal[0] = 0.f; p.normal[1] = 1.f; p.normal[2] = 0.f;
    p.radius = r;
    p.colour[0] = cr; p.colour[1] = cg; p.colour[2] = cb;
    appendPrimitive(sc, p);
  };
  const float gain = 2.f;
  prim(0, -1.8575f, -0.98714f, -3.6f, 0.6f, 1.f * gain, .89f * gain, .55f * gain, 0);
  prim(0, 0.74795f, -0.55f, -4.3816f, 1.05f, 1.f, 1.f, 1.f, 1);
  prim(0, 1.9929f, -1.08666f, (float)-3.23, 0.5f, 0.75f, 0.75f, 0.75f, 2);
  prim(0, (float)-0.19931, -1.183f, -2.75f, 0.4f, .8f * gain, .06f * gain, .391f * gain, 0);
  prim(0, (float)-0.19931, -1.183f, -2.75f, 0.4001f, 1.f, 1.f, 1.f, 2);
  prim(1, 0.f, -1.6f, -5.22f, 3.5f, .98f * gain, .76f * gain, .66f * gain, 0);      // Disc(Y, (0,-1.6,-5.22), 3.5)
}

} // namespace

extern "C" {

const char* b200pt_version(void) { return "b200pt 0.1 (sm_100a)"; }

const char* b200pt_last_error(const b200pt_ctx* ctx) { return ctx ? ctx->error.c_str() : g_createError.c_str(); }

int b200pt_create(const b200pt_config* cfg, b200pt_ctx** out) {
  if (!cfg || !out) { g_createError = "null argument"; return B200PT_ERR_INVALID; }
  *out = nullptr;
  if (cfg->struct_size != sizeof(b200pt_config)) { g_createError = "b200pt_config.struct_size mismatch"; return B200PT_ERR_INVALID; }
  if (cfg->width == 0 || cfg->height == 0 || cfg->width > 65535 || cfg->height > 65535) {
    g_createError = "image size must be 1..65535 (TraceRecord coordinates are uint16)"; return B200PT_ERR_INVALID;
  }
  if (cfg->roulette_depth == 0) {
    g_createError = "roulette_depth 0 makes the reference write before its record stack (codelets.cpp:176-179,219-222); rejected";
    return B200PT_ERR_INVALID;
  }
  if (cfg->max_path_length == 0 || cfg->max_worklist == 0) { g_createError = "max_path_length and max_worklist must be > 0"; return B200PT_ERR_INVALID; }
  if (cfg->aa_noise_type < 0 || cfg->aa_noise_type > 2) { g_createError = "unknown aa_noise_type"; return B200PT_ERR_INVALID; }
  int devCount = 0;
  cudaError_t e = cudaGetDeviceCount(&devCount);
  if (e != cudaSuccess || devCount == 0) {
    g_createError = std::string("no CUDA device (this library has no CPU path): ") + cudaGetErrorString(e);
    return B200PT_ERR_CUDA;
  }
  if (cfg->device < 0 || cfg->device >= devCount) { g_createError = "device ordinal out of range"; return B200PT_ERR_INVALID; }
  b200pt_ctx* c = new b200pt_ctx();
  c->cfg = *cfg;
  auto bail = [&](cudaError_t err, const char* what) {
    g_createError = std::string(what) + ": " + cudaGetErrorString(err);
    b200pt_destroy(c);
    return B200PT_ERR_CUDA;
  };
  if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) return bail(e, "cudaSetDevice");
  cudaDeviceProp prop{};
  if ((e = cudaGetDeviceProperties(&prop, cfg->device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
  if (prop.major != 10) {
    g_createError = "device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor) + "; libb200pt is built for sm_100a only";
    b200pt_destroy(c);
    return B200PT_ERR_INVALID;
  }
  c->numSms = prop.multiProcessorCount;
  {
    // The NIF stream gets the highest priority so that its persistent blocks are placed first and the
    // path-tracing blocks of the next wavefront fill the remaining register/thread slots of each SM.
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    if ((e = cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, hi)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaStreamCreateWithPriority(&c->traceStream, cudaStreamNonBlocking, lo)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    for (int i = 0; i < 2; ++i) {
      if ((e = cudaEventCreateWithFlags(&c->evQueueFree[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
      if ((e = cudaEventCreateWithFlags(&c->evTraced[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    }
    if (const char* ov = getenv("B200PT_OVERLAP")) c->overlap = std::atoi(ov) != 0;
  }
  if ((e = cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
  if ((e = cudaEventCreateWithFlags(&c->evNextUploaded, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if ((e = cudaEventCreateWithFlags(&c->evNextConsumed, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if ((e = cudaEventCreate(&c->evBegin)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if ((e = cudaEventCreate(&c->evEnd)) != cudaSuccess) return bail(e, "cudaEventCreate");

  const size_t cap = cfg->max_worklist;
  uint32_t spl = cfg->samples_per_launch;
  if (spl == 0) if (const char* e = getenv("B200PT_SPL")) spl = (uint32_t)std::max(0, std::atoi(e));    // A/B runs only
  if (spl == 0) {
    // automatic: enough escaped rays per wavefront to give every SM >= 192 tiles of 128 rows (C2 on one GPU: 4 samples
    // per launch; measured 1, 2, 3, 4, 6 -> 1.169, 1.173, 1.179, 1.182, 1.180 G samples/s), bounded so the queue stays
    // below ~1 GiB.
    const size_t want = (size_t)c->numSms * 192 * 128;
    spl = (uint32_t)std::max<size_t>(1, std::min<size_t>(32, (want + cap - 1) / cap));
    while (spl > 1 && cap * spl * 24 > (size_t)1 << 30) --spl;
    // path_trace shrinks the wavefront towards 64 tiles per SM when a step would otherwise have fewer than ~32 of them
    const size_t least = (size_t)c->numSms * 64 * 128;
    c->samplesPerLaunchMin = (uint32_t)std::min<size_t>(spl, std::max<size_t>(1, (least + cap - 1) / cap));
  }
  c->samplesPerLaunch = spl;
  if ((e = c->recStaging.alloc(cap)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = c->pix.alloc(cap)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = c->sampleCount.alloc(cap)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = c->pathLength.alloc(cap)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = c->accum.alloc(cap)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = c->qUv.alloc(cap * spl)) != cudaSuccess) return bail(e, "cudaMalloc queue");
  if ((e = c->qThrSlot.alloc(cap * spl)) != cudaSuccess) return bail(e, "cudaMalloc queue");
  if (c->overlap) {
    if ((e = c->qUvB.alloc(cap * spl)) != cudaSuccess) return bail(e, "cudaMalloc queue");
    if ((e = c->qThrSlotB.alloc(cap * spl)) != cudaSuccess) return bail(e, "cudaMalloc queue");
  }
  if ((e = c->qCount.alloc(4)) != cudaSuccess) return bail(e, "cudaMalloc");
  if (const char* tv = getenv("B200PT_TRACE")) {                   // A/B runs only
    if (!std::strcmp(tv, "v1")) c->traceVariant = 1;
    else if (!std::strcmp(tv, "v2")) c->traceVariant = 2;
    else if (!std::strncmp(tv, "v2p", 3)) c->traceVariant = 10 + std::max(1, std::min(8, std::atoi(tv + 3)));
  }
  if (c->traceVariant != 1) {
    const size_t entries = traceScratchEntries((uint32_t)cap);
    if ((e = c->liveScratch.alloc(entries)) != cudaSuccess) return bail(e, "cudaMalloc live-ray list");
    if ((e = c->escScratch.alloc(entries * 2)) != cudaSuccess) return bail(e, "cudaMalloc escape list");
  }
  if ((e = c->stats.alloc(4)) != cudaSuccess) return bail(e, "cudaMalloc");
  const size_t filmN = (size_t)cfg->width * cfg->height * 3;
  if ((e = c->film.alloc(filmN)) != cudaSuccess) return bail(e, "cudaMalloc film");
  if ((e = c->filmOut.alloc(filmN)) != cudaSuccess) return bail(e, "cudaMalloc film");
  if ((e = c->film8.alloc(filmN)) != cudaSuccess) return bail(e, "cudaMalloc film");
  if ((e = cudaMemsetAsync(c->film.p, 0, filmN * sizeof(float), c->stream)) != cudaSuccess) return bail(e, "cudaMemset");
  if ((e = cudaMemsetAsync(c->stats.p, 0, 4 * sizeof(unsigned long long), c->stream)) != cudaSuccess) return bail(e, "cudaMemset");

  SceneSoA scene;
  buildScene(scene);
  if ((e = c->scene.alloc(1)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = cudaMemcpyAsync(c->scene.p, &scene, sizeof(scene), cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) return bail(e, "scene upload");
  if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return bail(e, "cudaStreamSynchronize");
  *out = c;
  return B200PT_OK;
}

int b200pt_set_scene(b200pt_ctx* c, const b200pt_primitive* prims, uint32_t n) {
  if (!c) return B200PT_ERR_INVALID;
  if (n > (uint32_t)kMaxPrims) return c->fail(B200PT_ERR_INVALID, "set_scene: %u primitives exceed the limit of %d", n, kMaxPrims);
  if (n > 0 && !prims) return c->fail(B200PT_ERR_INVALID, "set_scene: null primitive list");
  SceneSoA scene;
  if (n == 0) {
    buildScene(scene);
  } else {
    std::memset(&scene, 0, sizeof(scene));
    for (uint32_t i = 0; i < n; ++i) {
      const b200pt_primitive& p = prims[i];
      if (p.shape < 0 || p.shape > 1 || p.material < 0 || p.material > 2 || !(p.radius > 0.f))
        return c->fail(B200PT_ERR_INVALID, "set_scene: primitive %u has an unknown shape/material or a non-positive radius", i);
      if (p.shape == 1 && p.normal[0] == 0.f && p.normal[1] == 0.f && p.normal[2] == 0.f)
        return c->fail(B200PT_ERR_INVALID, "set_scene: disc %u has a zero normal", i);
      appendPrimitive(scene, p);
    }
  }
  CU(c, cudaSetDevice(c->cfg.device));
  CU(c, cudaStreamSynchronize(c->stream));          // a path_trace in flight keeps the scene it started with
  if (c->traceStream) CU(c, cudaStreamSynchronize(c->traceStream));
  CU(c, cudaMemcpy(c->scene.p, &scene, sizeof(scene), cudaMemcpyHostToDevice));
  return B200PT_OK;
}

void b200pt_destroy(b200pt_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->cfg.device);
  if (c->traceStream) cudaStreamSynchronize(c->traceStream);
  if (c->copyStream) { cudaStreamSynchronize(c->copyStream); cudaStreamDestroy(c->copyStream); }
  if (c->evNextUploaded) cudaEventDestroy(c->evNextUploaded);
  if (c->evNextConsumed) cudaEventDestroy(c->evNextConsumed);
  c->recStagingNext.release();
  if (c->stream) cudaStreamSynchronize(c->stream);
  for (cudaEvent_t e : c->evPool) cudaEventDestroy(e);
  for (int i = 0; i < 2; ++i) { if (c->evQueueFree[i]) cudaEventDestroy(c->evQueueFree[i]); if (c->evTraced[i]) cudaEventDestroy(c->evTraced[i]); }
  c->qUvB.release(); c->qThrSlotB.release();
  if (c->traceStream) cudaStreamDestroy(c->traceStream);
  if (c->evBegin) cudaEventDestroy(c->evBegin);
  if (c->evEnd) cudaEventDestroy(c->evEnd);
  c->recStaging.release(); c->pix.release(); c->sampleCount.release(); c->pathLength.release(); c->accum.release();
  c->qUv.release(); c->qThrSlot.release(); c->qCount.release(); c->stats.release();
  c->liveScratch.release(); c->escScratch.release();
  c->scene.release(); c->film.release(); c->filmOut.release(); c->film8.release();
  for (auto& s : c->scratch) s.release();
  c->simtParams.release(); c->tcStream.release(); c->tcBias.release(); c->tcFeatTable.release();
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

} // extern "C"
