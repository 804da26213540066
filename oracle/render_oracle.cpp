// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, imported by or executed from
// the product path. See oracle_common.hpp for the usage rule and parity status.
//
// One device "path_trace" program on the CPU: `sampleCount` iterations of the reference's
// pathTraceIteration (src/PathTracerApp.cpp:432-458, looped by countedForLoop :466-468):
//   AA noise -> camera rays -> primary samples -> path trace -> escaped-ray uv -> NIF ->
//   apply env light -> accumulate into the TraceRecords.
// Used for converged-image parity (tests) and as the timed CPU baseline (bench.py). The
// path-tracing stages are the pinned restatement (pt_oracle.cpp); the NIF stage is the
// unpinned restatement (nif_oracle.cpp).
#include "oracle_common.hpp"

#include <vector>

using namespace orc;

extern "C" {
void orc_trace_philox(int n, const TraceRec* items, std::uint32_t sample, const Settings* s,
                      int* escaped, float* uv2, float* throughput3, int* length, std::uint16_t* camRays);
int orc_nif_forward(const Nif* nif, int n, const float* u, const float* v, int mode, float* bgr);

// items: in/out TraceRecord[n] (r,g,b,sampleCount,pathLength are accumulated into, like the
// reference's read_results payload). Returns the number of ESCAPED paths; secondsTrace /
// secondsNif (nullable) receive the split wall-clock.
long long orc_render(int n, TraceRec* items, std::uint32_t sampleBegin, std::uint32_t sampleCount,
                     const Settings* s, const Nif* nif, int nifMode, double* secondsTrace, double* secondsNif);
}

#include <omp.h>

long long orc_render(int n, TraceRec* items, std::uint32_t sampleBegin, std::uint32_t sampleCount,
                     const Settings* s, const Nif* nif, int nifMode, double* secondsTrace, double* secondsNif) {
  std::vector<int> escaped(n), length(n);
  std::vector<float> uv(2 * (std::size_t)n), thr(3 * (std::size_t)n);
  std::vector<float> qu, qv, bgr;
  std::vector<int> qidx;
  long long totalEscaped = 0;
  double tTrace = 0.0, tNif = 0.0;
  for (std::uint32_t k = 0; k < sampleCount; ++k) {
    double t0 = omp_get_wtime();
    orc_trace_philox(n, items, sampleBegin + k, s, escaped.data(), uv.data(), thr.data(), length.data(), nullptr);
    // compact the escaped rays (the reference evaluates the NIF for every path, escaped or
    // not, and discards the rest: codelets.cpp:348-352; the result is the same)
    qu.clear(); qv.clear(); qidx.clear();
    for (int i = 0; i < n; ++i) if (escaped[i] == 1) { qidx.push_back(i); qu.push_back(uv[2 * i]); qv.push_back(uv[2 * i + 1]); }
    double t1 = omp_get_wtime();
    bgr.resize(3 * qidx.size());
    if (!qidx.empty()) orc_nif_forward(nif, (int)qidx.size(), qu.data(), qv.data(), nifMode, bgr.data());
    double t2 = omp_get_wtime();
    for (std::size_t q = 0; q < qidx.size(); ++q) {
      const int i = qidx[q];
      // codelets.cpp:378 (BGR -> RGB) then the backward walk collapses to L (*) throughput
      items[i].r += bgr[3 * q + 2] * thr[3 * i];
      items[i].g += bgr[3 * q + 1] * thr[3 * i + 1];
      items[i].b += bgr[3 * q] * thr[3 * i + 2];
    }
    for (int i = 0; i < n; ++i) {
      if (items[i].u >= s->width || items[i].v >= s->height) continue;
      if (escaped[i] == 2) {           // ended on an emitter (codelets.cpp:192-196, :255-298): radiance without a NIF lookup
        items[i].r += thr[3 * i]; items[i].g += thr[3 * i + 1]; items[i].b += thr[3 * i + 2];
      }
      items[i].pathLength = (std::uint16_t)(items[i].pathLength + length[i]);
      items[i].sampleCount = (std::uint16_t)(items[i].sampleCount + 1);
    }
    totalEscaped += (long long)qidx.size();
    double t3 = omp_get_wtime();
    tTrace += (t1 - t0) + (t3 - t2);
    tNif += (t2 - t1);
  }
  if (secondsTrace) *secondsTrace = tTrace;
  if (secondsNif) *secondsNif = tNif;
  return totalEscaped;
}
