// NIF inference on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a only.
//
// One persistent CTA per SM walks 128-row tiles of the escaped-ray queue. Per tile:
//   encode warps    (u,v) -> 4E binary16 Fourier features, written straight into the UMMA
//                   K-major core-matrix layout in shared memory (NifModel.cpp:185-218)
//   producer warp   streams the pre-packed weight chunks global -> shared with
//                   cp.async.bulk (TMA engine, no tensor map needed: the host packs every
//                   chunk as the exact shared-memory image) through an mbarrier ring
//   MMA warp        one elected thread issues tcgen05.mma (M=128, N=H/2, K=16, fp16 in,
//                   fp32 accumulate in TMEM) for every hidden Dense layer (:299-326)
//   epilogue warps  tcgen05.ld the accumulator, round to binary16, add the binary16 bias,
//                   ReLU, and write the next layer's A operand back into shared memory in
//                   place; on the last hidden layer they also evaluate the final (3-output)
//                   Dense layer on the CUDA cores, decode (:221-245) and accumulate
//                   throughput (*) radiance into the work items (codelets.cpp:235-306,361-385).
//
// Pipelining inside a tile. A layer's N=H outputs are computed as two halves (N=H/2 each, a
// legal UMMA N) into a rotation of three TMEM accumulators of H/2 columns. The K loop of the
// next layer is ordered [cols 0..H/2) for both halves, then [H/2..H) for both halves], so it
// can start as soon as the epilogue of the previous layer's FIRST half has written its
// columns; the epilogue of the second half overlaps with those MMAs. The activation buffer is
// rewritten in place: tcgen05.commit ordering guarantees that every MMA that read a column
// range has completed before the epilogue that overwrites it is released.
#include "nif_common.cuh"

#include <cstdio>

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
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  const long long t0 = clock64();
  while (true) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (clock64() - t0 > 4000000000ll) {
      printf("b200pt nif_tc: mbarrier timeout block %d thread %d bar %u parity %u\n", blockIdx.x, threadIdx.x, bar, parity);
      __trap();
    }
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
__device__ __forceinline__ void tcMma(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(dTmem), "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tcLoad16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tcWaitLoad() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ bool electOne() {
  uint32_t pred;
  asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void namedBarrier(uint32_t id, uint32_t threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor): core
// matrices are 8 rows x 16 bytes stored contiguously; LBO is the byte distance between the two
// core matrices of one K=16 step, SBO the distance between consecutive 8-row groups.
__device__ __forceinline__ uint64_t umDesc(uint32_t addr, uint32_t lboBytes, uint32_t sboBytes) {
  return (uint64_t)((addr & 0x3ffffu) >> 4) | ((uint64_t)(lboBytes >> 4) << 16) | ((uint64_t)(sboBytes >> 4) << 32) |
         (1ull << 46);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): fp16 x fp16 -> fp32, both K-major.
__device__ __forceinline__ uint32_t umInstr(uint32_t M, uint32_t N) {
  return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

struct TcSmem {
  // byte offsets from the 1024-aligned base
  uint32_t act, feat[2], ring, bias, wOut, meta[2], partial, chunks, bars, tmemSlot, total;
};
__host__ __device__ inline TcSmem tcLayout(int H, int featKSteps, int numHidden, int numChunks) {
  TcSmem s;
  uint32_t o = 0;
  s.act = o; o += (uint32_t)H * kTcRows * 2;
  s.feat[0] = o; o += (uint32_t)featKSteps * 4096;
  s.feat[1] = o; o += (uint32_t)featKSteps * 4096;
  s.ring = o; o += (uint32_t)kTcStages * kTcChunkKSteps * (uint32_t)(H / 2) * 32;
  s.bias = o; o += (uint32_t)numHidden * H * 2;
  s.wOut = o; o += (uint32_t)H * 16;
  s.meta[0] = o; o += kTcRows * 16;
  s.meta[1] = o; o += kTcRows * 16;
  s.partial = o; o += kTcRows * 16;
  s.chunks = o; o += (uint32_t)numChunks * (uint32_t)sizeof(NifTcChunk);
  o = (o + 15u) & ~15u;
  s.bars = o; o += 32 * 8;
  s.tmemSlot = o; o += 16;
  s.total = o;
  return s;
}

// barrier indices inside the bars block
enum { BAR_W_FULL = 0, BAR_W_EMPTY = 4, BAR_FEAT_FULL = 8, BAR_FEAT_EMPTY = 10, BAR_META_EMPTY = 12,
       BAR_ACC_FULL = 14, BAR_ACC_EMPTY = 17, BAR_ACT_READY = 20, BAR_COUNT = 22 };

__global__ void __launch_bounds__(kTcThreads, 1)
nif_tc_kernel(const NifTcModel* __restrict__ mp, const NifIo io) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const NifTcModel& m = *mp;
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
    // bias, final-layer weights and the chunk table: global -> shared
    const uint32_t* gb = reinterpret_cast<const uint32_t*>(m.bias);
    uint32_t* sb = reinterpret_cast<uint32_t*>(smem + L.bias);
    for (int i = threadIdx.x; i < numHidden * H / 2; i += kTcThreads) sb[i] = gb[i];
    float4* sw = reinterpret_cast<float4*>(smem + L.wOut);
    for (int i = threadIdx.x; i < H; i += kTcThreads) sw[i] = m.wOut[i];
    const uint32_t* gc = reinterpret_cast<const uint32_t*>(m.chunks);
    uint32_t* scn = reinterpret_cast<uint32_t*>(smem + L.chunks);
    for (int i = threadIdx.x; i < numChunks * (int)(sizeof(NifTcChunk) / 4); i += kTcThreads) scn[i] = gc[i];
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < kTcStages; ++s) { mbarInit(bar(BAR_W_FULL + s), 1); mbarInit(bar(BAR_W_EMPTY + s), 1); }
    for (int b = 0; b < 2; ++b) {
      mbarInit(bar(BAR_FEAT_FULL + b), 32 * kEncodeWarps);
      mbarInit(bar(BAR_FEAT_EMPTY + b), 1);
      mbarInit(bar(BAR_META_EMPTY + b), kEpiThreads);
      mbarInit(bar(BAR_ACT_READY + b), kEpiThreads);
    }
    for (int b = 0; b < 3; ++b) { mbarInit(bar(BAR_ACC_FULL + b), 1); mbarInit(bar(BAR_ACC_EMPTY + b), kEpiThreads); }
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
  const NifTcChunk* chunks = reinterpret_cast<const NifTcChunk*>(smem + L.chunks);

  if (warp == 0) {
    // ================= weight producer =================
    if (electOne()) {
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x) {
        for (int c = 0; c < numChunks; ++c) {
          const NifTcChunk ch = chunks[c];
          const uint32_t bytes = (uint32_t)ch.bytes16 * 16u;
          mbarWait(bar(BAR_W_EMPTY + stage), phase ^ 1u);
          mbarArriveExpectTx(bar(BAR_W_FULL + stage), bytes);
          bulkCopyG2S(sbase + L.ring + stage * (kTcChunkKSteps * (uint32_t)NH * 32u), m.wStream + ch.offset, bytes, bar(BAR_W_FULL + stage));
          if (++stage == kTcStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = umInstr(kTcRows, (uint32_t)NH);
    uint32_t stage = 0, phase = 0;
    uint32_t actUse[2] = {0, 0};    // completed waits on act_ready[part]
    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x, ++it) {
      const uint32_t fb = it & 1u;
      const uint32_t gTile = it * 2u * (uint32_t)numHidden;     // (layer, half) counter -> accumulator rotation
      mbarWait(bar(BAR_FEAT_FULL + fb), (it >> 1) & 1u);
      for (int c = 0; c < numChunks; ++c) {
        const NifTcChunk ch = chunks[c];
        const uint32_t gi = gTile + 2u * ch.layer + ch.half;
        const uint32_t buf = gi % 3u;
        if (ch.first) mbarWait(bar(BAR_ACC_EMPTY + buf), ((gi / 3u) & 1u) ^ 1u);     // claim the accumulator
        if (ch.waitAct) {
          const uint32_t part = ch.waitAct - 1u;
          mbarWait(bar(BAR_ACT_READY + part), actUse[part] & 1u);
          actUse[part] += 1;
        }
        mbarWait(bar(BAR_W_FULL + stage), phase);
        tcFenceAfter();
        if (electOne()) {
          const uint32_t dT = tmemBase + buf * (uint32_t)NH;
          const uint32_t aBase = sbase + (ch.srcFeat ? L.feat[fb] : L.act) + (uint32_t)ch.aKstep * 4096u;
          const uint32_t bBase = sbase + L.ring + stage * (kTcChunkKSteps * (uint32_t)NH * 32u);
          for (uint32_t j = 0; j < ch.ksteps; ++j) {
            const uint64_t aD = umDesc(aBase + j * 4096u, 2048u, 128u);
            const uint64_t bD = umDesc(bBase + j * ((uint32_t)NH * 32u), (uint32_t)NH * 16u, 128u);
            tcMma(dT, aD, bD, idesc, (ch.first && j == 0) ? 0u : 1u);
          }
          tcCommit(bar(BAR_W_EMPTY + stage));
          if (ch.lastOfHalf) tcCommit(bar(BAR_ACC_FULL + buf));
          if (ch.lastFeatUse) tcCommit(bar(BAR_FEAT_EMPTY + fb));
        }
        __syncwarp();
        if (++stage == kTcStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp < kFirstEpiWarp) {
    // ================= feature encode =================
    const uint32_t r = (warp - kFirstEncodeWarp) * 32u + lane;      // row inside the tile
    const int E = m.embeddingDim, nFeat = m.nFeat, featCores = m.featKSteps * 2;
    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x, ++it) {
      const uint32_t fb = it & 1u;
      const uint32_t row = tile * kTcRows + r;
      float2 uv = make_float2(0.f, 0.f);
      float4 ts = make_float4(0.f, 0.f, 0.f, __uint_as_float(0xffffffffu));
      if (row < count) {
        uv = io.uv[row];
        if (io.thrSlot) ts = io.thrSlot[row];
      }
      __half f[kMaxFeatures];
#pragma unroll
      for (int k = 0; k < kMaxFeatures; ++k) f[k] = __float2half_rn(0.f);
      nifEncode(uv.x, uv.y, E, [&](int k, __half v) { f[k] = v; });
      (void)nFeat;
      // wait until the previous user of this buffer pair is done with it
      mbarWait(bar(BAR_FEAT_EMPTY + fb), ((it >> 1) & 1u) ^ 1u);
      mbarWait(bar(BAR_META_EMPTY + fb), ((it >> 1) & 1u) ^ 1u);
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
      *reinterpret_cast<float4*>(smem + L.meta[fb] + r * 16) = ts;
      fenceProxyAsync();
      mbarArrive(bar(BAR_FEAT_FULL + fb));
    }
  } else {
    // ================= epilogue =================
    const uint32_t ew = warp - kFirstEpiWarp;
    const uint32_t quad = warp & 3u;                      // TMEM lane quadrant this warp may read
    const uint32_t colHalf = ew >> 2;                     // which half of the NH columns
    const uint32_t r = quad * 32u + lane;                 // tile row == TMEM lane
    const uint32_t colsMine = (uint32_t)NH / 2u;          // multiple of 16
    const float4* wOut = reinterpret_cast<const float4*>(smem + L.wOut);
    uint32_t g = 0, it = 0;
    for (uint32_t tile = blockIdx.x; tile < numTiles; tile += gridDim.x, ++it) {
      const uint32_t fb = it & 1u;
      float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
      for (int l = 0; l < numHidden; ++l) {
        const bool lastHidden = (l == numHidden - 1);
        const int relu = m.relu[l];
        const __half2* bias2 = reinterpret_cast<const __half2*>(smem + L.bias + (uint32_t)l * H * 2u);
        for (uint32_t h = 0; h < 2; ++h, ++g) {
          const uint32_t buf = g % 3u;
          mbarWait(bar(BAR_ACC_FULL + buf), (g / 3u) & 1u);
          tcFenceAfter();
          const uint32_t tBase = tmemBase + ((quad * 32u) << 16) + buf * (uint32_t)NH + colHalf * colsMine;
          const uint32_t colBase = h * (uint32_t)NH + colHalf * colsMine;      // output-neuron index of my first column
          for (uint32_t c0 = 0; c0 < colsMine; c0 += 16) {
            uint32_t v[16];
            tcLoad16(tBase + c0, v);
            tcWaitLoad();
            uint32_t packed[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              __half2 hv = __floats2half2_rn(__uint_as_float(v[2 * q]), __uint_as_float(v[2 * q + 1]));
              hv = __hadd2(hv, bias2[(colBase + c0) / 2 + q]);
              if (relu) hv = __hmax2(hv, __float2half2_rn(0.f));
              packed[q] = *reinterpret_cast<uint32_t*>(&hv);
              if (lastHidden) {
                const float2 a = __half22float2(hv);
                const float4 w0 = wOut[colBase + c0 + 2 * q], w1 = wOut[colBase + c0 + 2 * q + 1];
                o0 = fmaf(a.x, w0.x, o0); o1 = fmaf(a.x, w0.y, o1); o2 = fmaf(a.x, w0.z, o2); o3 = fmaf(a.x, w0.w, o3);
                o0 = fmaf(a.y, w1.x, o0); o1 = fmaf(a.y, w1.y, o1); o2 = fmaf(a.y, w1.z, o2); o3 = fmaf(a.y, w1.w, o3);
              }
            }
            if (!lastHidden) {
              unsigned char* dst = smem + L.act + ((colBase + c0) / 8u) * 2048u + r * 16u;
              *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
              *reinterpret_cast<uint4*>(dst + 2048) = make_uint4(packed[4], packed[5], packed[6], packed[7]);
            }
          }
          tcFenceBefore();
          mbarArrive(bar(BAR_ACC_EMPTY + buf));
          if (!lastHidden) {
            fenceProxyAsync();
            mbarArrive(bar(BAR_ACT_READY + h));
          }
        }
      }
      // final Dense layer: combine the two column halves of every row, decode, emit
      float4* partial = reinterpret_cast<float4*>(smem + L.partial);
      if (colHalf == 1) partial[r] = make_float4(o0, o1, o2, o3);
      namedBarrier(1, kEpiThreads);
      if (colHalf == 0) {
        // the meta block of this tile was published together with the features
        mbarWait(bar(BAR_FEAT_FULL + fb), (it >> 1) & 1u);
        const float4 p = partial[r];
        float out[4] = {o0 + p.x, o1 + p.y, o2 + p.z, o3 + p.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          __half hv = __float2half_rn(out[c]);
          if (m.finalHasBias) hv = __float2half_rn(__half2float(hv) + m.bOut[c]);
          if (m.finalRelu && __half2float(hv) < 0.f) hv = __float2half_rn(0.f);
          out[c] = __half2float(hv);
        }
        const uint32_t row = tile * kTcRows + r;
        if (row < count) {
          const float4 ts = *reinterpret_cast<const float4*>(smem + L.meta[fb] + r * 16);
          nifEmit(io, row, nifDecode(out[0], m.maxv, m.mean[0], m.logToneMap), nifDecode(out[1], m.maxv, m.mean[1], m.logToneMap),
                  nifDecode(out[2], m.maxv, m.mean[2], m.logToneMap), ts);
        }
      }
      namedBarrier(1, kEpiThreads);       // partial[] may be rewritten by the next tile
      mbarArrive(bar(BAR_META_EMPTY + fb));
    }
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

cudaError_t launchNifTc(const NifTcModel* modelDev, const NifTcModel& modelHost, const NifIo& io, int numSms, cudaStream_t stream) {
  if (io.countHost == 0) return cudaSuccess;
  const size_t smem = nifTcSmemBytes(modelHost);
  uint32_t tilesMax = (io.countHost + kTcRows - 1) / kTcRows;
  uint32_t grid = tilesMax < (uint32_t)numSms ? tilesMax : (uint32_t)numSms;
  nif_tc_kernel<<<grid, kTcThreads, smem, stream>>>(modelDev, io);
  return cudaGetLastError();
}

} // namespace b2
