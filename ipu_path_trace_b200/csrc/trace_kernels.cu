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

// A path that ended on an emitter (EMIT record, codelets.cpp:192-196) needs no NIF lookup: its radiance goes straight into the
// work item's accumulator, with the same fixed-point atomics the NIF kernel uses for escaped paths (nif_common.cuh nifEmit).
__device__ __forceinline__ void emitRadiance(AccumSlot* __restrict__ accum, uint32_t slot, V3 radiance) {
  accumAdd(accum, slot, radiance.x, radiance.y, radiance.z);
}

// v1 (default): one thread carries a whole path per sample; lanes whose path has ended idle until the longest path of the
// warp is done (ncu: 19.9 of 32 lanes per instruction); escaped rays are compacted with ballot/prefix per sample.
__global__ void __launch_bounds__(kTraceThreads)
trace_wavefront_kernel_v1(const SceneSoA* __restrict__ scene, const DeviceSettings s, const uint32_t* __restrict__ pix, uint32_t nItems, uint32_t sampleBegin, uint32_t samplesPerLaunch,
                       float2* __restrict__ qUv, float4* __restrict__ qThrSlot, uint32_t* __restrict__ qCount,
                       uint32_t* __restrict__ sampleCount, uint32_t* __restrict__ pathLength,
                       unsigned long long* __restrict__ stats /* [0] paths [1] escaped [2] segments */, AccumSlot* __restrict__ accum) {
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
      } else if (r.emitted) {
        emitRadiance(accum, slot, r.throughput);
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

// v2 (B200PT_TRACE=v2): live-ray compaction and path regeneration. Bit-identical to v1 (tests/test_gpu_trace.py,
// test_gpu_baseline_configs.py run both); measured on B200 it executes 21.4 of 32 lanes per instruction (v1: 19.9) but
// 15 % more thread instructions (the first intersection of a live ray is redone, records are staged twice), so a C2
// step's trace work takes 36.0 ms against v1's 30.6 ms: 71 % of the paths end at the first intersection, where v1 already
// runs on full warps, and the deep bounces that compaction repairs are only ~25 % of the work. Kept as the evidenced
// alternative (profiles/r2_trace_v2_ab.md); v1 stays the default.
//
// A block owns kTraceItems x 256 consecutive work items and walks the launch's samples in groups of (up to) four.
//   phase 1  every thread: Philox -> anti-alias noise -> camera ray -> FIRST intersection, for its item and each sample
//            of the group. 71 % of the paths escape right there: their (u,v) go to the NIF queue (one ballot/prefix
//            compaction and ONE global atomic per item round, not per sample). The rest are LIVE rays: they are
//            compacted (warp ballot + prefix, one shared-memory atomic per warp) into the block's live list as 16-byte
//            records {item, sample, camera ray, the two unused words of the Philox block}.
//   phase 2  the warps drain the live list with a per-lane state machine: one bounce per iteration (roulette, intersect,
//            BSDF sample - pathBounce, bit-identical to tracePath), and a lane whose path ends takes the next live ray in
//            the same iteration (ballot + prefix on a shared-memory cursor), so the deep bounces run on full warps
//            instead of on the 29 / 11 / 4 % of lanes still alive. Rays that escape here are staged (direction,
//            throughput, item) in the block's escape list;
//   phase 3  the escape list is turned into (u,v) - the glibc acosf / double atan2 path, converged - and appended to the
//            NIF queue with one more global atomic.
// Per-item path lengths are summed in shared memory (a live ray is finished by whichever lane picked it up).
#ifndef B200PT_TRACE_ITEMS
#define B200PT_TRACE_ITEMS 2
#endif
constexpr int kTraceItems = B200PT_TRACE_ITEMS;       // work items per thread
constexpr int kGroup = 4;                             // samples per group
constexpr int kBlockItems = kTraceThreads * kTraceItems;
constexpr int kBlockPaths = kBlockItems * kGroup;     // capacity of the live / escape lists of a block

__global__ void __launch_bounds__(kTraceThreads)
trace_wavefront_kernel(const SceneSoA* __restrict__ scene, const DeviceSettings s, const uint32_t* __restrict__ pix, uint32_t nItems, uint32_t sampleBegin, uint32_t samplesPerLaunch,
                       float2* __restrict__ qUv, float4* __restrict__ qThrSlot, uint32_t* __restrict__ qCount,
                       uint32_t* __restrict__ sampleCount, uint32_t* __restrict__ pathLength,
                       unsigned long long* __restrict__ stats /* [0] paths [1] escaped [2] segments */, AccumSlot* __restrict__ accum,
                       uint4* __restrict__ liveScratch, float4* __restrict__ escScratch, uint32_t phase2Warps) {
  __shared__ SceneSoA sc;
  __shared__ uint32_t warpCount[kTraceThreads / 32];
  __shared__ uint32_t blockBase, liveCount, liveNext, escCount;
  __shared__ uint32_t lenAcc[kBlockItems];
  __shared__ unsigned long long red[3];
  stageScene(sc, scene);

  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t ltMask = (1u << lane) - 1u;
  const uint32_t itemBase = blockIdx.x * (uint32_t)kBlockItems;
  uint4* live = liveScratch + (size_t)blockIdx.x * kBlockPaths;
  float4* escBuf = escScratch + (size_t)blockIdx.x * kBlockPaths * 2;
  for (uint32_t i = tid; i < (uint32_t)kBlockItems; i += kTraceThreads) lenAcc[i] = 0u;
  if (tid < 3) red[tid] = 0ull;
  uint32_t myEscaped = 0;

  for (uint32_t k0 = 0; k0 < samplesPerLaunch; k0 += kGroup) {
    const uint32_t kb = min((uint32_t)kGroup, samplesPerLaunch - k0);
    __syncthreads();                 // the previous group's lists are drained; lenAcc / red are initialised
    if (tid == 0) { liveCount = 0u; liveNext = 0u; escCount = 0u; }
    __syncthreads();

    // ---------------- phase 1: camera rays and the first intersection ----------------
#pragma unroll 1
    for (uint32_t p = 0; p < (uint32_t)kTraceItems; ++p) {
      const uint32_t local = p * kTraceThreads + tid, slot = itemBase + local;
      uint32_t pu = 0xffffu, pv = 0xffffu;
      if (slot < nItems) { const uint32_t px = pix[slot]; pu = px & 0xffffu; pv = px >> 16; }
      const bool valid = pu < s.width && pv < s.height;       // padding items: LoadBalancer.cpp:66-71
      const uint32_t pixelId = pv * s.width + pu;
      float2 uv[kGroup];
      uint32_t escMask = 0u;
#pragma unroll
      for (uint32_t j = 0; j < (uint32_t)kGroup; ++j) {
        uv[j] = make_float2(0.f, 0.f);
        bool isLive = false;
        uint4 rec = make_uint4(0u, 0u, 0u, 0u);
        if (valid && j < kb) {
          PathRng rng(s.seedLo, s.seedHi, pixelId, sampleBegin + k0 + j);
          const uint32_t w0 = rng.word(), w1 = rng.word();
          float n0, n1;
          aaNoisePair(w0, w1, s.aaNoiseType, n0, n1);
          float cx, cy;
          cameraRay(pu, pv, n0, n1, s, cx, cy);
          V3 o = v3(0.f, 0.f, 0.f), n;
          const V3 d = unit(v3(cx, cy, -1.f));
          // depth 0 is below the roulette depth (b200pt_create rejects roulette_depth 0), so the first intersection
          // draws nothing: a miss is an ESCAPED record with weight 1, i.e. throughput (1,1,1) and length 1
          if (closestHit(sc, o, d, n) < 0) {
            escapedUv(d, s.azimuth, uv[j].x, uv[j].y);
            escMask |= 1u << j;
          } else {
            isLive = true;
            rec = make_uint4(local | (j << 16), (uint32_t)halfBits(cx) | ((uint32_t)halfBits(cy) << 16), rng.w2, rng.w3);
          }
        }
        const uint32_t lm = __ballot_sync(0xffffffffu, isLive);
        if (lm) {
          uint32_t base = 0u;
          if (lane == 0) base = atomicAdd(&liveCount, (uint32_t)__popc(lm));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (isLive) live[base + __popc(lm & ltMask)] = rec;
        }
      }
      // escaped at depth 0: compaction of up to kGroup records per thread, one global atomic per block
      const uint32_t cnt = (uint32_t)__popc(escMask);
      uint32_t incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)lane >= o) incl += t; }
      if (lane == 31) warpCount[warp] = incl;
      __syncthreads();
      if (tid == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < kTraceThreads / 32; ++w) { const uint32_t c = warpCount[w]; warpCount[w] = total; total += c; }
        blockBase = total ? atomicAdd(qCount, total) : 0u;
      }
      __syncthreads();
      if (cnt) {
        uint32_t idx = blockBase + warpCount[warp] + incl - cnt;
#pragma unroll
        for (uint32_t j = 0; j < (uint32_t)kGroup; ++j) {
          if (escMask & (1u << j)) {
            qUv[idx] = uv[j];
            qThrSlot[idx] = make_float4(1.f, 1.f, 1.f, __uint_as_float(slot));
            idx += 1;
          }
        }
        atomicAdd(&lenAcc[local], cnt);          // one record each (codelets.cpp:181-186)
        myEscaped += cnt;
      }
      __syncthreads();                           // warpCount / blockBase are rewritten by the next round
    }

    // ---------------- phase 2: drain the live list, one bounce per iteration ----------------
    const uint32_t nLive = liveCount;            // (the barrier above ordered every append before this read)
    if (warp < phase2Warps && nLive) {
      bool active = false;
      PathState ps;
      ps.o = v3(0.f, 0.f, 0.f); ps.d = v3(0.f, 0.f, -1.f); ps.thr = v3(1.f, 1.f, 1.f); ps.depth = 0; ps.count = 0;
      PathRng rng(PathRng::Resume(), s.seedLo, s.seedHi, 0u, 0u, 0u, 0u);
      uint32_t local = 0u;
      while (true) {
        const uint32_t need = __ballot_sync(0xffffffffu, !active);
        if (need) {
          uint32_t base = 0u;
          if (lane == 0) base = atomicAdd(&liveNext, (uint32_t)__popc(need));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (!active) {
            const uint32_t idx = base + __popc(need & ltMask);
            if (idx < nLive) {
              const uint4 rec = live[idx];
              local = rec.x & 0xffffu;
              const uint32_t px = pix[itemBase + local];
              const uint32_t pixelId = (px >> 16) * s.width + (px & 0xffffu);
              rng = PathRng(PathRng::Resume(), s.seedLo, s.seedHi, pixelId, sampleBegin + k0 + (rec.x >> 16), rec.z, rec.w);
              ps.o = v3(0.f, 0.f, 0.f);
              ps.d = unit(v3(halfBitsToFloat((uint16_t)(rec.y & 0xffffu)), halfBitsToFloat((uint16_t)(rec.y >> 16)), -1.f));
              ps.thr = v3(1.f, 1.f, 1.f);
              ps.depth = 0; ps.count = 0;
              active = true;
            }
          }
        }
        if (!__any_sync(0xffffffffu, active)) break;
        int state = kPathLive;
        if (active) state = pathBounce(ps, s, sc, [&]() { return rng.uniformHalf(); });
        const bool done = active && state != kPathLive;
        const bool esc = active && state == kPathEscaped;
        if (active && state == kPathEmitted) emitRadiance(accum, itemBase + local, ps.thr);
        if (done) { atomicAdd(&lenAcc[local], (uint32_t)ps.count); active = false; }
        const uint32_t em = __ballot_sync(0xffffffffu, esc);
        if (em) {
          uint32_t base = 0u;
          if (lane == 0) base = atomicAdd(&escCount, (uint32_t)__popc(em));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (esc) {
            const uint32_t idx = base + __popc(em & ltMask);
            escBuf[2 * idx] = make_float4(ps.d.x, ps.d.y, ps.d.z, __uint_as_float(itemBase + local));
            escBuf[2 * idx + 1] = make_float4(ps.thr.x, ps.thr.y, ps.thr.z, 0.f);
            myEscaped += 1;
          }
        }
      }
    }
    __syncthreads();

    // ---------------- phase 3: escaped later than depth 0 -> (u,v) -> NIF queue ----------------
    const uint32_t nEsc = escCount;
    if (nEsc) {
      if (tid == 0) blockBase = atomicAdd(qCount, nEsc);
      __syncthreads();
      for (uint32_t i = tid; i < nEsc; i += kTraceThreads) {
        const float4 a = escBuf[2 * i], b = escBuf[2 * i + 1];
        float u, v;
        escapedUv(v3(a.x, a.y, a.z), s.azimuth, u, v);
        qUv[blockBase + i] = make_float2(u, v);
        qThrSlot[blockBase + i] = make_float4(b.x, b.y, b.z, a.w);
      }
    }
  }
  __syncthreads();

  // per-item counters (codelets.cpp:300-301: every path counts) and the launch statistics
  uint32_t a = 0, c = 0;
  for (uint32_t p = 0; p < (uint32_t)kTraceItems; ++p) {
    const uint32_t local = p * kTraceThreads + tid, slot = itemBase + local;
    if (slot < nItems) {
      const uint32_t px = pix[slot];
      if ((px & 0xffffu) < s.width && (px >> 16) < s.height) {
        sampleCount[slot] += samplesPerLaunch;
        pathLength[slot] += lenAcc[local];
        a += samplesPerLaunch; c += lenAcc[local];
      }
    }
  }
  uint32_t b = myEscaped;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if (lane == 0) { atomicAdd(&red[0], (unsigned long long)a); atomicAdd(&red[1], (unsigned long long)b); atomicAdd(&red[2], (unsigned long long)c); }
  __syncthreads();
  if (tid < 3 && red[tid]) atomicAdd(&stats[tid], red[tid]);
}

size_t traceScratchEntries(uint32_t maxItems) {
  const size_t blocks = ((size_t)maxItems + kBlockItems - 1) / kBlockItems;
  return blocks * (size_t)kBlockPaths;
}

cudaError_t launchTraceWavefront(const SceneSoA* scene, const DeviceSettings& s, const uint32_t* pix, uint32_t nItems, uint32_t sampleBegin, uint32_t samplesPerLaunch,
                                 float2* qUv, float4* qThrSlot, uint32_t* qCount, uint32_t* sampleCount,
                                 uint32_t* pathLength, unsigned long long* stats, AccumSlot* accum, uint4* liveScratch, float4* escScratch, int variant,
                                 cudaStream_t stream) {
  if (nItems == 0) return cudaSuccess;
  if (variant == 1 || !liveScratch || !escScratch) {
    const uint32_t blocks = (nItems + kTraceThreads - 1) / kTraceThreads;
    trace_wavefront_kernel_v1<<<blocks, kTraceThreads, 0, stream>>>(scene, s, pix, nItems, sampleBegin, samplesPerLaunch, qUv, qThrSlot,
                                                                   qCount, sampleCount, pathLength, stats, accum);
    return cudaGetLastError();
  }
  const uint32_t blocks = (nItems + kBlockItems - 1) / kBlockItems;
  const uint32_t p2w = variant >= 10 ? (uint32_t)(variant - 10) : (uint32_t)(kTraceThreads / 32);
  trace_wavefront_kernel<<<blocks, kTraceThreads, 0, stream>>>(scene, s, pix, nItems, sampleBegin, samplesPerLaunch, qUv, qThrSlot,
                                                              qCount, sampleCount, pathLength, stats, accum, liveScratch, escScratch, p2w);
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
    else if (r.emitted) { esc = 2; thr = r.throughput; }        // ended on an emitter: thr is the radiance itself
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
  escaped[i] = r.escaped ? 1 : (r.emitted ? 2 : 0);
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
