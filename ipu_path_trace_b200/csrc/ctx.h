// Internal: the context object behind b200pt_ctx and small host helpers.
#pragma once

#include "../../include/b200pt.h"
#include "kernels.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

using namespace b2;

namespace b2host {

// ---- binary16 on the host (IEEE round-to-nearest-even) -----------------------------------
inline float hostHalfToFloat(uint16_t h) { return __half2float(__ushort_as_half(h)); }
inline uint16_t hostFloatToHalf(float f) { return __half_as_ushort(__float2half_rn(f)); }
inline float hostRoundHalf(float f) { return hostHalfToFloat(hostFloatToHalf(f)); }

template <class T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  cudaError_t alloc(size_t count) {
    release();
    n = count;
    if (count == 0) return cudaSuccess;
    return cudaMalloc(reinterpret_cast<void**>(&p), count * sizeof(T));
  }
  cudaError_t ensure(size_t count) { return count <= n ? cudaSuccess : alloc(count); }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

} // namespace b2host
using namespace b2host;

struct b200pt_ctx {
  b200pt_config cfg{};
  std::string error;
  int numSms = 0;
  cudaStream_t stream = nullptr;
  // Second stream: the path-tracing kernel of wavefront w+1 runs beside the NIF kernel of wavefront w
  // (the NIF kernel leaves registers and shared memory for one trace block per SM).
  cudaStream_t traceStream = nullptr;
  bool overlap = true;
  cudaEvent_t evBegin = nullptr, evEnd = nullptr;
  cudaEvent_t evQueueFree[2] = {nullptr, nullptr}, evTraced[2] = {nullptr, nullptr};
  std::vector<cudaEvent_t> evPool;       // per-launch timing events of the last path_trace
  size_t evUsed = 0;
  struct Span { size_t a, b; int kind; };  // kind 0 trace, 1 nif
  std::vector<Span> spans;

  // render settings
  DeviceSettings settings{};
  bool settingsSet = false;
  uint32_t samplesPerStep = 0;
  uint32_t nextSample = 0;
  uint32_t samplesPerLaunch = 0;      // queue capacity in samples of every item
  uint32_t samplesPerLaunchMin = 0;   // automatic choice only (else 0): lower end of the range path_trace picks from

  // worklist (SoA) and accumulators
  uint32_t nItems = 0;
  DevBuf<TraceRecord> recStaging;
  // setup_next: the next step's work list is uploaded on copyStream into recStagingNext while path_trace runs
  DevBuf<TraceRecord> recStagingNext;
  cudaStream_t copyStream = nullptr;
  cudaEvent_t evNextUploaded = nullptr;
  // recorded on `stream` behind the unpack kernel that reads recStagingNext: the next upload into that buffer waits for it
  cudaEvent_t evNextConsumed = nullptr;
  bool nextConsumedRecorded = false;
  bool nextStaged = false;
  uint32_t nextItems = 0;
  DevBuf<uint32_t> pix, sampleCount, pathLength;
  DevBuf<AccumSlot> accum;
  // wavefront queue
  DevBuf<float2> qUv, qUvB;             // two queues: one being filled while the other is shaded
  DevBuf<float4> qThrSlot, qThrSlotB;
  DevBuf<uint32_t> qCount;               // [0], [1]: row counts of the two queues
  // trace kernel v2: per-block live-ray and escape lists (trace_kernels.cu)
  DevBuf<uint4> liveScratch;
  DevBuf<float4> escScratch;
  // B200PT_TRACE selects the path-tracing kernel: "v1" (default) one thread carries a path per sample; "v2" live-ray
  // compaction + regeneration (bit-identical, measured slower on this scene: profiles/r2_trace_v2_ab.md); "v2p<k>" = v2
  // with k phase-2 warps.
  int traceVariant = 1;
  DevBuf<unsigned long long> stats;
  // film
  DevBuf<SceneSoA> scene;
  DevBuf<float> film, filmOut;
  DevBuf<uint8_t> film8;
  // scratch for stage-level entry points
  DevBuf<unsigned char> scratch[8];

  // NIF
  bool nifLoaded = false;
  int64_t nifFlops = 0;
  NifSimtModel simt{};
  DevBuf<float> simtParams;
  bool tcOk = false;
  std::string tcWhyNot;
  NifTcModel tcHost{};
  DevBuf<uint8_t> tcStream;
  int tcMaxPairs = 0;
  DevBuf<__half> tcBias;
  DevBuf<uint32_t> tcFeatTable;

  b200pt_counters counters{};
  bool pending = false;      // a path_trace is enqueued and not yet folded into counters

  int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    error = buf;
    return code;
  }
  cudaEvent_t nextEvent() {
    if (evUsed == evPool.size()) {
      cudaEvent_t e = nullptr;
      if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
      evPool.push_back(e);
    }
    return evPool[evUsed++];
  }
  int cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return B200PT_OK;
    return fail(B200PT_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  }
};

// NVTX ranges named after the reference's PVTI tracepoints (PathTracerApp.cpp:606-607,690-761); header-only, a no-op
// unless a tool (nsys) is attached.
#include <nvtx3/nvToolsExt.h>
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

#define CU(ctx, call) do { int rc_ = (ctx)->cuda((call), #call); if (rc_) return rc_; } while (0)


// api_nif.cu: chooses and launches the NIF kernel for `io` (impl: B200PT_NIF_*).
int launchNif(b200pt_ctx* c, const NifIo& io, int impl, int* used);
