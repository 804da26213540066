// NIF inference on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a only.
//
// One persistent CTA per SM walks 128-row tiles of the escaped-ray queue. Per tile:
//   encode warps    (u,v) -> 4E binary16 Fourier features, written straight into the UMMA
//                   K-major core-matrix layout in shared memory (NifModel.cpp:185-218)
//   producer thread streams the pre-packed weight chunks global -> shared with cp.async.bulk
//                   (TMA engine, no tensor map needed: the host packs every chunk as the exact
//                   shared-memory image) through an mbarrier ring
//   MMA thread      issues tcgen05.mma (M=128, K=16, fp16 in, fp32 accumulate in TMEM) for every
//                   Dense layer (:299-326): hidden layers as two N=H/2 jobs, the final (<= 4 output)
//                   layer as one N=16 job. Chunk descriptors are precomputed on the host so the
//                   issue loop is a handful of instructions per MMA.
//   epilogue warps  tcgen05.ld the accumulator (software-pipelined), round to binary16, add the
//                   binary16 bias, ReLU, and write the next layer's A operand back into shared
//                   memory in place; for the output job they decode (:221-245) and accumulate
//                   throughput (*) radiance into the work items (codelets.cpp:235-306,361-385).
//
// Pipelining inside a tile. The two half-jobs of a layer go to a rotation of three TMEM
// accumulators of H/2 columns. The K loop of the next layer is ordered [cols 0..H/2) for both
// halves, then [H/2..H) for both halves], so it starts as soon as the epilogue of the previous
// layer's FIRST half has written its columns; the epilogue of the second half overlaps with
// those MMAs. The activation buffer is rewritten in place: tcgen05.commit ordering guarantees
// that every MMA that read a column range has completed before the epilogue that overwrites it
// is released.
#include "nif_common.cuh"

namespace b2 {

constexpr int kTcRows = 128;
constexpr int kTcStages = 4;
constexpr int kEncodeWarps = 4;
constexpr int kEpiWarps = 8;
constexpr int kTcThreads = 32 * (2 + kEncodeWarps + kEpiWarps);
constexpr int kEpiThreads = 32 * kEpiWarps;
constexpr int kFirstEncodeWarp = 2;
constexpr int kFirstEpiWarp = 2 + kEncodeWarps;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kOutTmemCol = 480;      // the N=16 output accumulator sits after the three H/2 <= 160 column slots

// ---- PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarArrive(uint32_t bar) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbarArriveExpectTx(uint32_t bar, uint32_t bytes) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug traps (CUDA error at the API) instead of hanging the GPU.
// try_wait itself blocks for a hardware-defined time slice, so the spin count stays small.
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  uint32_t spins = 0;
  while (true) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (++spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void bulkCopyG2S(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fenceBarrierInit() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tcFenceBefore() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcFenceAfter() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcCommit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; descriptors passed as (lo, hi) 32-bit halves.
__device__ __forceinline__ void tcMma(uint32_t dTmem, uint32_t aLo, uint32_t bLo, uint32_t descHi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
      ::"r"(dTmem), "r"(aLo), "r"(bLo), "r"(descHi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tcLoad16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ bool electOne() {
  uint32_t pred;
  asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tcWaitLoad() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor): core
// matrices are 8 rows x 16 bytes stored contiguously; LBO is the byte distance between the two
// core matrices of one K=16 step, SBO (= 128 B here) the distance between consecutive 8-row
// groups. lo = start address >> 4 | (LBO >> 4) << 16; hi = SBO >> 4 | version 1 at bit 46.
constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
// Instruction descriptor (cute::UMMA::InstrDescriptor): fp16 x fp16 -> fp32, both K-major, M = 128.
__device__ __forceinline__ uint32_t umInstr(uint32_t n8) { return (1u << 4) | (n8 << 17) | ((uint32_t)(kTcRows >> 4) << 24); }

// Optional in-kernel timers (build with -DB200PT_PROFILE; scratch/nif_prof.py reads them).
#ifdef B200PT_PROFILE
__device__ unsigned long long g_tcProf[148 * 16];
#define PROF_T() clock64()
#define PROF_ADD(var, t0) (var) += clock64() - (t0)
#else
#define PROF_T() 0ll
#define PROF_ADD(var, t0) ((void)(t0))
#endif

struct TcSmem {
  // byte offsets from the dynamic shared memory base
  uint32_t act, feat[2], ring, bias, chunks, bars, tmemSlot, total;
};
__host__ __device__ inline TcSmem tcLayout(int H, int featKSteps, int numHidden, int numChunks) {
  TcSmem s;
  uint32_t o = 0;
  s.act = o; o += (uint32_t)H * kTcRows * 2;
  s.feat[0] = o; o += (uint32_t)featKSteps * 4096;
  s.feat[1] = o; o += (uint32_t)featKSteps * 4096;
  s.ring = o; o += (uint32_t)kTcStages * kTcChunkKSteps * (uint32_t)(H / 2) * 32;
  s.bias = o; o += (uint32_t)numHidden * H * 2;
  s.chunks = o; o += (uint32_t)numChunks * (uint32_t)sizeof(NifTcChunk);
  o = (o + 15u) & ~15u;
  s.bars = o; o += 32 * 8;
  s.tmemSlot = o; o += 16;
  s.total = o;
  return s;
}

// barrier indices inside the bars block
enum { BAR_W_FULL = 0, BAR_W_EMPTY = 4, BAR_FEAT_FULL = 8, BAR_FEAT_EMPTY = 10,
       BAR_ACC_FULL = 12, BAR_ACC_EMPTY = 15, BAR_ACT_READY = 18, BAR_OUT_FULL = 20, BAR_OUT_EMPTY = 21, BAR_COUNT = 22 };

__global__ void __launch_bounds__(kTcThreads, 1)
nif_tc_kernel(const __grid_constant__ NifTcModel m, const NifIo io) {
  // The model (incl. the chunk table) is a kernel parameter: the MMA and producer warps read it
  // through the constant bank with warp-uniform indices, so descriptors, barrier addresses and loop
  // state stay in uniform registers and tcgen05.mma issues back to back. (A chunk table in shared
  // memory, or a single-lane issue loop, makes ptxas wrap every MMA in an R2UR/ELECT waterfall:
  // ~180 cycles per MMA instead of the 80-cycle tensor floor at N = 160; see profiles/.)
  extern __shared__ __align__(1024) unsigned char smem[];
  const int H = m.H, NH = H / 2, numHidden = m.numHidden, numChunks = m.numChunks;
  const TcSmem L = tcLayout(H, m.featKSteps, numHidden, numChunks);
  const uint32_t sbase = smemAddr(smem);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t barBase = sbase + L.bars;
  auto bar = [&](int i) { return barBase + 8u * (uint32_t)i; };

  const uint32_t count = io.count ? *io.count : io.countHost;
  const uint32_t numTiles = (count + kTcRows - 1) / kTcRows;

  // ---- one-time setup ----
  {
    // bias: global -> shared
    const uint32_t* gb = reinterpret_cast<const uint32_t*>(m.bias);
    uint32_t* sb = reinterpret_cast<uint32_t*>(smem + L.bias);
    for (int i = threadIdx.x; i < numHidden * H / 2; i += kTcThreads) sb[i] = gb[i];
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < kTcStages; ++s) { mbarInit(bar(BAR_W_FULL + s), 1); mbarInit(bar(BAR_W_EMPTY + s), 1); }
    for (int b = 0; b < 2; ++b) {
      mbarInit(bar(BAR_FEAT_FULL + b), 32 * kEncodeWarps);
      mbarInit(bar(BAR_FEAT_EMPTY + b), 1);
      mbarInit(bar(BAR_ACT_READY + b), kEpiThreads);
    }
    for (int b = 0; b < 3; ++b) { mbarInit(bar(BAR_ACC_FULL + b), 1); mbarInit(bar(BAR_ACC_EMPTY + b), kEpiThreads); }
    mbarInit(bar(BAR_OUT_FULL), 1);
    mbarInit(bar(BAR_OUT_EMPTY), kEpiThreads / 2);
    fenceBarrierInit();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + L.tmemSlot), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcFenceBefore();
  __syncthreads();
  tcFenceAfter();
  const uint32_t tmemBase = *reinterpret_cast<volatile uint32_t*>(smem + L.tmemSlot);
  const uint32_t stageBytes = kTcChunkKSteps * (uint32_t)NH * 32u;

  if (warp == 0) {
    // ================= weight producer (converged warp, one elected lane issues) =================
    const bool leader = electOne();
    uint32_t stage = 0, phase = 0;
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x) {
      for (int c = 0; c < numChunks; ++c) {
        const uint32_t bytes = (uint32_t)m.chunks[c].bytes16 * 16u;
        mbarWait(bar(BAR_W_EMPTY + stage), phase ^ 1u);
        if (leader) {
          mbarArriveExpectTx(bar(BAR_W_FULL + stage), bytes);
          bulkCopyG2S(sbase + L.ring + stage * stageBytes, m.wStream + m.chunks[c].offset, bytes, bar(BAR_W_FULL + stage));
        }
        if (++stage == kTcStages) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer (converged warp, one elected lane issues) =================
    const bool leader = electOne();
    const uint32_t actLo = ((sbase + L.act) >> 4);
    const uint32_t ringLo = ((sbase + L.ring) >> 4);
    uint32_t stage = 0, phase = 0;
    uint32_t actUse0 = 0, actUse1 = 0;     // completed waits on act_ready[part]
    uint32_t it = 0;
    long long pFeat = 0, pAcc = 0, pAct = 0, pW = 0, pOutE = 0; const long long pStart = PROF_T();
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x, ++it) {
      const uint32_t fb = it & 1u;
      const uint32_t featLo = ((sbase + L.feat[fb]) >> 4);
      const uint32_t jobBase = it * 2u * (uint32_t)numHidden;      // hidden jobs are numbered (tile, layer, half)
      { const long long t0 = PROF_T(); mbarWait(bar(BAR_FEAT_FULL + fb), (it >> 1) & 1u); PROF_ADD(pFeat, t0); }
      for (int c = 0; c < numChunks; ++c) {
        const NifTcChunk& ch = m.chunks[c];
        const uint32_t ksteps = ch.ksteps, n8 = ch.n8, flags = ch.flags, waitAct = ch.waitAct;
        const bool isOut = (flags & kChOut) != 0;
        // the two half-jobs of a layer accumulate concurrently (K order above): slot from the job number
        const uint32_t job = jobBase + 2u * ch.layer + ch.half;
        const uint32_t buf = job % 3u, accPhase = (job / 3u) & 1u;
        if (flags & kChFirst) {                                    // claim the accumulator
          const long long t0 = PROF_T();
          if (isOut) { mbarWait(bar(BAR_OUT_EMPTY), (it & 1u) ^ 1u); PROF_ADD(pOutE, t0); }
          else { mbarWait(bar(BAR_ACC_EMPTY + buf), accPhase ^ 1u); PROF_ADD(pAcc, t0); }
        }
        { const long long t0 = PROF_T();
          if (waitAct == 1) { mbarWait(bar(BAR_ACT_READY + 0), actUse0 & 1u); actUse0 += 1; }
          else if (waitAct == 2) { mbarWait(bar(BAR_ACT_READY + 1), actUse1 & 1u); actUse1 += 1; }
          PROF_ADD(pAct, t0); }
        { const long long t0 = PROF_T(); mbarWait(bar(BAR_W_FULL + stage), phase); PROF_ADD(pW, t0); }
        tcFenceAfter();
        const uint32_t dT = tmemBase + (isOut ? kOutTmemCol : buf * (uint32_t)NH);
        uint32_t aLo = (((flags & kChSrcFeat) ? featLo : actLo) + ch.aOff16) | ((2048u >> 4) << 16);
        uint32_t bLo = (ringLo + stage * (stageBytes >> 4)) | ((n8 * 8u) << 16);     // LBO = N * 16 bytes
        const uint32_t idesc = umInstr(n8);
        const uint32_t bStep = n8 * 16u;                                              // N * 32 bytes per K-step
        uint32_t acc = (flags & kChFirst) ? 0u : 1u;
#pragma unroll
        for (uint32_t j = 0; j < (uint32_t)kTcChunkKSteps; ++j) {
          if (j < ksteps) {
            if (leader) tcMma(dT, aLo, bLo, kDescHi, idesc, acc);
            aLo += 256u; bLo += bStep; acc = 1u;
          }
        }
        if (leader) {
          tcCommit(bar(BAR_W_EMPTY + stage));
          if (flags & kChLast) tcCommit(bar(isOut ? BAR_OUT_FULL : BAR_ACC_FULL + buf));
          if (flags & kChLastFeatUse) tcCommit(bar(BAR_FEAT_EMPTY + fb));
        }
        if (++stage == kTcStages) { stage = 0; phase ^= 1u; }
      }
    }
#ifdef B200PT_PROFILE
    if (lane == 0) { unsigned long long* p = g_tcProf + blockIdx.x * 16; p[0] = clock64() - pStart; p[1] = pFeat; p[2] = pAcc; p[3] = pAct; p[4] = pW; p[5] = pOutE; p[6] = it; }
#endif
    __syncwarp();
  } else if (warp < kFirstEpiWarp) {
    // ================= feature encode =================
    const uint32_t r = (warp - kFirstEncodeWarp) * 32u + lane;      // row inside the tile
    const int E = m.embeddingDim, featCores = m.featKSteps * 2;
    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x, ++it) {
      const uint32_t fb = it & 1u;
      const uint32_t row = tile * kTcRows + r;
      float2 uv = make_float2(0.f, 0.f);
      if (row < count) uv = io.uv[row];
      __half f[kMaxFeatures];
#pragma unroll
      for (int k = 0; k < kMaxFeatures; ++k) f[k] = __float2half_rn(0.f);
      nifEncode(uv.x, uv.y, E, [&](int k, __half v) { f[k] = v; });
      // wait until the MMAs of the tile that last used this buffer are done with it
      mbarWait(bar(BAR_FEAT_EMPTY + fb), ((it >> 1) & 1u) ^ 1u);
      unsigned char* fbuf = smem + L.feat[fb];
      for (int kc = 0; kc < featCores; ++kc) {
        uint4 pk;
        const __half2* h2 = reinterpret_cast<const __half2*>(&f[kc * 8]);
        pk.x = *reinterpret_cast<const uint32_t*>(&h2[0]);
        pk.y = *reinterpret_cast<const uint32_t*>(&h2[1]);
        pk.z = *reinterpret_cast<const uint32_t*>(&h2[2]);
        pk.w = *reinterpret_cast<const uint32_t*>(&h2[3]);
        *reinterpret_cast<uint4*>(fbuf + kc * 2048 + r * 16) = pk;
      }
      fenceProxyAsync();
      mbarArrive(bar(BAR_FEAT_FULL + fb));
    }
  } else {
    // ================= epilogue =================
    const uint32_t ew = warp - kFirstEpiWarp;
    const uint32_t quad = warp & 3u;                      // TMEM lane quadrant this warp may read
    const uint32_t colHalf = ew >> 2;                     // which half of a job's NH columns
    const uint32_t r = quad * 32u + lane;                 // tile row == TMEM lane
    const uint32_t colsMine = (uint32_t)NH / 2u;          // multiple of 16, <= 80
    const uint32_t groups = colsMine / 16u;               // <= 5
    const uint32_t laneBase = tmemBase + ((quad * 32u) << 16);
    const __half2 zero2 = __float2half2_rn(0.f);
    uint32_t buf = 0, accPhase = 0, it = 0;
    long long eWait = 0, eWork = 0, eOutWait = 0, eOutWork = 0; const long long eStart = PROF_T();
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x, ++it) {
      for (int l = 0; l < numHidden; ++l) {
        const bool relu = m.relu[l] != 0;
        const unsigned char* biasRow = smem + L.bias + (uint32_t)l * (uint32_t)H * 2u;
        for (uint32_t h = 0; h < 2; ++h) {
          { const long long t0 = PROF_T(); mbarWait(bar(BAR_ACC_FULL + buf), accPhase); PROF_ADD(eWait, t0); }
          const long long tw = PROF_T();
          tcFenceAfter();
          const uint32_t tBase = laneBase + buf * (uint32_t)NH + colHalf * colsMine;
          const uint32_t colBase = h * (uint32_t)NH + colHalf * colsMine;      // output-neuron index of my first column
          // software pipeline: the load of group g+1 is in flight while group g is converted and stored
          uint32_t v[2][16];
          tcLoad16(tBase, v[0]);
#pragma unroll
          for (uint32_t g = 0; g < 5; ++g) {
            if (g < groups) {
              tcWaitLoad();
              if (g + 1 < groups) tcLoad16(tBase + (g + 1) * 16u, v[(g + 1) & 1]);
              const uint32_t c0 = colBase + g * 16u;
              const uint4 b0 = *reinterpret_cast<const uint4*>(biasRow + c0 * 2u);
              const uint4 b1 = *reinterpret_cast<const uint4*>(biasRow + c0 * 2u + 16u);
              const uint32_t bw[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
              uint32_t packed[8];
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                __half2 hv = __floats2half2_rn(__uint_as_float(v[g & 1][2 * q]), __uint_as_float(v[g & 1][2 * q + 1]));
                hv = __hadd2(hv, *reinterpret_cast<const __half2*>(&bw[q]));
                if (relu) hv = __hmax2(hv, zero2);
                packed[q] = *reinterpret_cast<uint32_t*>(&hv);
              }
              unsigned char* dst = smem + L.act + (c0 / 8u) * 2048u + r * 16u;
              *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
              *reinterpret_cast<uint4*>(dst + 2048) = make_uint4(packed[4], packed[5], packed[6], packed[7]);
            }
          }
          tcFenceBefore();
          mbarArrive(bar(BAR_ACC_EMPTY + buf));
          fenceProxyAsync();
          mbarArrive(bar(BAR_ACT_READY + h));
          PROF_ADD(eWork, tw);
          if (++buf == 3u) { buf = 0; accPhase ^= 1u; }
        }
      }
      // output job: N = 16 accumulator, channels 0..3 are the network outputs (B,G,R[,unused])
      if (colHalf == 0) {
        { const long long t0 = PROF_T(); mbarWait(bar(BAR_OUT_FULL), it & 1u); PROF_ADD(eOutWait, t0); }
        const long long tw = PROF_T();
        tcFenceAfter();
        uint32_t v[16];
        tcLoad16(laneBase + kOutTmemCol, v);
        tcWaitLoad();
        tcFenceBefore();
        mbarArrive(bar(BAR_OUT_EMPTY));
        const uint32_t row = tile * kTcRows + r;
        if (row < count) {
          float out[3];
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            __half hv = __float2half_rn(__uint_as_float(v[c]));           // Dense output in the kernel's type (binary16)
            if (m.finalHasBias) hv = __float2half_rn(__half2float(hv) + m.bOut[c]);
            if (m.finalRelu && __half2float(hv) < 0.f) hv = __float2half_rn(0.f);
            out[c] = nifDecode(__half2float(hv), m.maxv, m.mean[c], m.logToneMap);
          }
          float4 ts = make_float4(0.f, 0.f, 0.f, 0.f);
          if (io.thrSlot) ts = io.thrSlot[row];
          nifEmit(io, row, out[0], out[1], out[2], ts);
        }
        PROF_ADD(eOutWork, tw);
      }
    }
#ifdef B200PT_PROFILE
    if (ew == 0 && lane == 0) { unsigned long long* p = g_tcProf + blockIdx.x * 16 + 8; p[0] = clock64() - eStart; p[1] = eWait; p[2] = eWork; p[3] = eOutWait; p[4] = eOutWork; }
#endif
  }

  // ---- teardown ----
  tcFenceBefore();
  __syncthreads();
  tcFenceAfter();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(kTmemCols) : "memory");
  }
}

size_t nifTcSmemBytes(const NifTcModel& m) {
  return (size_t)tcLayout(m.H, m.featKSteps, m.numHidden, m.numChunks).total + 1024;
}

cudaError_t prepareNifTc(size_t smemBytes) {
  return cudaFuncSetAttribute(nif_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
}

#ifdef B200PT_PROFILE
extern "C" int b200pt_debug_tc_profile(unsigned long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, g_tcProf, sizeof(unsigned long long) * n);
}
#endif

cudaError_t launchNifTc(const NifTcModel* modelDev, const NifTcModel& modelHost, const NifIo& io, int numSms, cudaStream_t stream) {
  if (io.countHost == 0) return cudaSuccess;
  const size_t smem = nifTcSmemBytes(modelHost);
  uint32_t tilesMax = (io.countHost + kTcRows - 1) / kTcRows;
  uint32_t grid = tilesMax < (uint32_t)numSms ? tilesMax : (uint32_t)numSms;
  (void)modelDev;
  nif_tc_kernel<<<grid, kTcThreads, smem, stream>>>(modelHost, io);
  return cudaGetLastError();
}

} // namespace b2
