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
#include <cstdlib>

namespace b2 {

constexpr int kTcRows = 128;
constexpr int kEncodeWarps = 4;
// Epilogue warps: 8 (two per TMEM lane quadrant, 80 columns of a half-job each) or 16 (four per quadrant, 40 columns
// each). The hand-over of a half-job's activations to the next layer's first MMAs has to fit into the ~800 cycles of the
// layer's last weight chunk; with 8 warps it takes ~800-1000 cycles and the tensor pipe waits ~570 cycles per layer.
#ifndef B200PT_EPI_WARPS
#define B200PT_EPI_WARPS 8
#endif
constexpr int kEpiWarps = B200PT_EPI_WARPS;
static_assert(kEpiWarps == 8 || kEpiWarps == 16, "two or four epilogue warps per TMEM lane quadrant");
constexpr int kEpiColParts = kEpiWarps / 4;      // column parts of a half-job
constexpr int kEpiGroup = kEpiWarps == 8 ? 16 : 8;   // accumulator columns per tcgen05.ld
#ifndef B200PT_EPI_FRONT
#define B200PT_EPI_FRONT (B200PT_EPI_WARPS == 16)
#endif
constexpr bool kEpiFrontLoad = B200PT_EPI_FRONT != 0;   // request every accumulator column of a job before converting
constexpr int kOutWarps = 4;                     // the warps of column part 0 also run the output job
constexpr int kTcThreads = 32 * (3 + kEncodeWarps + kEpiWarps);
constexpr int kSecondIssuerWarp = 2 + kEncodeWarps + kEpiWarps;   // single-CTA MMA variants issue from two warps in turn
constexpr int kFirstEncodeWarp = 2;
constexpr int kFirstEpiWarp = 2 + kEncodeWarps;
constexpr uint32_t kTmemCols = 512;
// weight-ring depth: the CTA pair holds half a chunk per CTA, so the same bytes give twice the stages


// ---- PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t clusterCtaRank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void clusterSync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank`
__device__ __forceinline__ uint32_t mapToCta(uint32_t addr, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarArrive(uint32_t bar) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(bar) : "memory");
}
// arrive on a barrier of another (or this) CTA of the cluster, given its shared::cluster address
__device__ __forceinline__ void mbarArriveCluster(uint32_t clusterAddr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(clusterAddr) : "memory");
}
__device__ __forceinline__ void mbarArriveExpectTx(uint32_t bar, uint32_t bytes) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug traps (CUDA error at the API) instead of hanging the GPU.
// try_wait itself blocks for a hardware-defined time slice, so the spin count stays small.
template <int CG>
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
// Non-blocking poll: 1 if the phase with this parity has completed. The result is a plain register, so
// the poll's latency overlaps whatever is issued before its first use.
__device__ __forceinline__ uint32_t mbarTest(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile("{ .reg .pred p; mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
               : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done;
}
__device__ __forceinline__ void bulkCopyG2S(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// Same, delivered to the same CTA-relative offsets (data and barrier) of every CTA in ctaMask.
__device__ __forceinline__ void bulkCopyG2SMulticast(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint16_t ctaMask) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "h"(ctaMask) : "memory");
}
__device__ __forceinline__ void fenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fenceBarrierInit() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tcFenceBefore() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcFenceAfter() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// MMA-completion arrive. CG == 2: the arrive is multicast to the barrier at this offset in both CTAs.
// MC: single-CTA MMAs whose completion is announced to both CTAs of a weight-sharing cluster.
__device__ __forceinline__ void tcCommitBoth(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}
template <int CG>
__device__ __forceinline__ void tcCommit(uint32_t bar) {
  if (CG == 2)
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
  else
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; descriptors passed as (lo, hi) 32-bit halves.
// CG == 2: one instruction drives the tensor cores of both SMs of the pair (M = 256): each CTA
// supplies its own 128 rows of A and half of the N columns of B from the same smem offsets.
template <int CG>
__device__ __forceinline__ void tcMma(uint32_t dTmem, uint32_t aLo, uint32_t bLo, uint32_t descHi, uint32_t idesc, uint32_t accumulate) {
  if (CG == 2)
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(dTmem), "r"(aLo), "r"(bLo), "r"(descHi), "r"(idesc), "r"(accumulate) : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(dTmem), "r"(aLo), "r"(bLo), "r"(descHi), "r"(idesc), "r"(accumulate) : "memory");
}
// Same with the A operand in tensor memory (128 lanes x 8 columns of packed binary16 pairs per K-step).
template <int CG>
__device__ __forceinline__ void tcMmaTs(uint32_t dTmem, uint32_t aTmem, uint32_t bLo, uint32_t descHi, uint32_t idesc, uint32_t accumulate) {
  if (CG == 2)
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], db, %4, p;\n\t}"
        ::"r"(dTmem), "r"(aTmem), "r"(bLo), "r"(descHi), "r"(idesc), "r"(accumulate) : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n\t}"
        ::"r"(dTmem), "r"(aTmem), "r"(bLo), "r"(descHi), "r"(idesc), "r"(accumulate) : "memory");
}
// registers -> tensor memory: 8 consecutive 32-bit columns of this thread's lane
__device__ __forceinline__ void tcStore8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tcWaitStore() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tcLoad16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tcLoad8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tcStore4(uint32_t taddr, const uint32_t (&v)[4]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
}
__device__ __forceinline__ void tcLoadN(uint32_t taddr, uint32_t (&v)[16]) { tcLoad16(taddr, v); }
__device__ __forceinline__ void tcLoadN(uint32_t taddr, uint32_t (&v)[8]) { tcLoad8(taddr, v); }
__device__ __forceinline__ void tcStoreN(uint32_t taddr, const uint32_t (&v)[8]) { tcStore8(taddr, v); }
__device__ __forceinline__ void tcStoreN(uint32_t taddr, const uint32_t (&v)[4]) { tcStore4(taddr, v); }
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
// Instruction descriptor (cute::UMMA::InstrDescriptor): fp16 x fp16 -> fp32, both K-major, M = 128 per CTA.
template <int CG>
__device__ __forceinline__ uint32_t umInstr(uint32_t n8) { return (1u << 4) | (n8 << 17) | ((uint32_t)((kTcRows * CG) >> 4) << 24); }

// Optional in-kernel timers (build with -DB200PT_PROFILE; scratch/nif_prof.py reads them).
#ifdef B200PT_PROFILE
__device__ unsigned long long g_tcProf[148 * 16];
// Timeline of ONE tile (CTA 0, tile iteration kTraceIt): [chunk * 4 + {0 loop top, 1 barriers passed, 2 turn taken, 3 issued}] for the MMA
// warps, [128 + job * 2 + {0 accumulator full seen, 1 hand-over arrives done}] for epilogue warp 0; scratch/nif_timeline.py prints it.
__device__ unsigned long long g_tcTrace[256];
constexpr unsigned kTraceIt = 8;
#define PROF_MARK(cond, idx) do { if ((cond) && blockIdx.x == 0) g_tcTrace[(idx)] = clock64(); } while (0)
#define PROF_T() clock64()
#define PROF_ADD(var, t0) (var) += clock64() - (t0)
#else
#define PROF_T() 0ll
#define PROF_ADD(var, t0) ((void)(var), (void)(t0))
#define PROF_MARK(cond, idx) ((void)0)
#endif

struct TcSmem {
  // byte offsets from the dynamic shared memory base
  uint32_t act, feat[2], ring, bias, bars, tmemSlot, total;
  uint32_t stages, stageBytes;
};
// cg = CTAs per MMA group (1, or 2 for a CTA pair): the ring holds 1/cg of every chunk per CTA.
// ts = activations live in tensor memory (A operand from TMEM), so no activation buffer in smem and
// the weight ring takes everything that is left. kc = K-steps per chunk = per ring stage.
__host__ __device__ inline TcSmem tcLayout(int H, int featKSteps, int numHidden, int cg, int ts, int kc) {
  TcSmem s;
  uint32_t o = 0;
  s.act = o; if (!ts) o += (uint32_t)H * kTcRows * 2;
  // With activations in tensor memory one feature buffer is enough: its last reader (the skip-concat layer)
  // is done half a tile before the next tile's first layer needs the new features, and the 12 KB buy the
  // weight ring a fourth stage.
  s.feat[0] = o; o += (uint32_t)featKSteps * 4096;
  s.feat[1] = ts ? s.feat[0] : o; if (!ts) o += (uint32_t)featKSteps * 4096;
  s.bias = o; o += (uint32_t)numHidden * H * 2;
  o = (o + 15u) & ~15u;
  s.bars = o; o += 48 * 8;
  s.tmemSlot = o; o += 16;
  o = (o + 127u) & ~127u;
  s.ring = o;
  s.stageBytes = (uint32_t)kc * (uint32_t)(H / 2 / cg) * 32u;
  const uint32_t budget = 232448u - 1024u;                       // 227 KB per CTA minus the alignment slack
  uint32_t st = (budget - o) / s.stageBytes;
  if (!ts) st = cg == 2 ? 8u : 4u;      // the non-TS variants keep their measured configuration
  s.stages = st > 8u ? 8u : st;
  o += s.stages * s.stageBytes;
  s.total = o;
  return s;
}

// barrier indices inside the bars block
enum { BAR_W_FULL = 0, BAR_W_EMPTY = 8, BAR_PEER_W_FULL = 16, BAR_FEAT_FULL = 24, BAR_FEAT_EMPTY = 26,
       BAR_ACC_FULL = 28, BAR_ACC_EMPTY = 31, BAR_ACT_READY = 34, BAR_OUT_FULL = 36, BAR_OUT_EMPTY = 37, BAR_COUNT = 38,
       SLOT_TURN = 40 /* not a barrier: index of the chunk whose MMAs may be issued next */ };

// CG = 1: one CTA per SM, M = 128 MMAs. CG = 2: a cluster of two CTAs (one TPC) runs M = 256 MMAs
// issued by the rank-0 CTA; each CTA encodes, stores and post-processes its own 128-row tile and
// streams only ITS half of every weight chunk, so L2 -> SM weight traffic per row halves.
// Barriers that gate the MMA issuer (feature/activation ready, accumulator drained, peer weights
// landed) live in the rank-0 CTA and collect one arrival per warp from both CTAs; barriers the MMA
// signals (accumulator full, weight stage free, feature buffer free) are multicast commits.
// TS: hidden activations are kept in tensor memory as packed binary16 (the A operand of the next
// layer comes from TMEM, the epilogue writes it with tcgen05.st): no activation traffic in shared
// memory, and the 80 KB it occupied go to the weight ring. TMEM then holds activations (H/2 columns),
// two accumulator slots (one per half-job) and the output accumulator.
// KC: K-steps per weight chunk (= MMAs per ring stage).
// MC (with CG = 1): two CTAs of a cluster run independent M = 128 pipelines on their own tiles but share
// the weight stream: each CTA fetches half of every chunk and multicasts it into the ring of both, so
// L2 -> SM weight traffic per row halves without the cross-CTA MMA. A ring stage is free once the MMAs
// of BOTH CTAs have read it (multicast commit, barrier count 2).
template <int CG, bool TS, int KC, bool MC>
__global__ void __launch_bounds__(kTcThreads, 1)
nif_tc_kernel(const __grid_constant__ NifTcModel m, const NifIo io) {
  // The model (incl. the chunk table) is a kernel parameter: the MMA and producer warps read it
  // through the constant bank with warp-uniform indices, so descriptors, barrier addresses and loop
  // state stay in uniform registers and tcgen05.mma issues back to back. (A chunk table in shared
  // memory, or a single-lane issue loop, makes ptxas wrap every MMA in an R2UR/ELECT waterfall:
  // ~180 cycles per MMA instead of the 80-cycle tensor floor at N = 160; see profiles/.)
  extern __shared__ __align__(1024) unsigned char smem[];
  constexpr uint32_t kSlots = TS ? 2u : 3u;             // TMEM accumulator slots of H/2 columns
  const int H = m.H, NH = H / 2, numHidden = m.numHidden, numChunks = m.numChunks;
  const TcSmem L = tcLayout(H, m.featKSteps, numHidden, CG, TS ? 1 : 0, KC);
  const int kStages = (int)L.stages;
  const uint32_t accCol0 = TS ? (uint32_t)NH : 0u;      // TS: columns [0, H/2) hold the activations
  const uint32_t outCol = 3u * (uint32_t)NH;            // output accumulator (16 columns)
  const uint32_t sbase = smemAddr(smem);
  // The warp index is broadcast from lane 0 so that ptxas knows the role branches below are warp-uniform:
  // only then does it keep the MMA warp's descriptors, barrier addresses and loop state on the uniform
  // datapath (UTCHMMA needs uniform registers) - 9-12 instructions per tcgen05.mma instead of 18-22 with
  // per-MMA R2UR / VOTEU marshalling.
  const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31u;
  static_assert(!(MC && CG == 2), "weight multicast is the single-CTA-MMA alternative to the CTA pair");
  constexpr uint32_t CL = (CG == 2 || MC) ? 2u : 1u;     // CTAs per cluster = tiles per work unit
  const uint32_t rank = (CL == 2) ? clusterCtaRank() : 0u;
  const uint32_t barBase = sbase + L.bars;
  auto bar = [&](int i) { return barBase + 8u * (uint32_t)i; };
  // arrival on a barrier that gates the MMA issuer: lives in the rank-0 CTA
  auto arriveIssuer = [&](int i) {
    if (CG == 2) mbarArriveCluster(mapToCta(bar(i), 0u)); else mbarArrive(bar(i));
  };

  const uint32_t count = io.count ? *io.count : io.countHost;
  const uint32_t numTiles = (count + kTcRows - 1) / kTcRows;
  // work unit = CG consecutive tiles (one per CTA of the group); both CTAs of a pair run the same
  // number of units, a tile index past the end is all padding rows
  const uint32_t numUnits = (numTiles + CL - 1) / CL;
  const uint32_t unit0 = blockIdx.x / CL, unitStride = gridDim.x / CL;

  // ---- one-time setup ----
  {
    // bias: global -> shared
    const uint32_t* gb = reinterpret_cast<const uint32_t*>(m.bias);
    uint32_t* sb = reinterpret_cast<uint32_t*>(smem + L.bias);
    for (int i = threadIdx.x; i < numHidden * H / 2; i += kTcThreads) sb[i] = gb[i];
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbarInit(bar(BAR_W_FULL + s), 1); mbarInit(bar(BAR_W_EMPTY + s), MC ? 2 : 1); mbarInit(bar(BAR_PEER_W_FULL + s), 1); }
    for (int b = 0; b < 2; ++b) {
      mbarInit(bar(BAR_FEAT_FULL + b), CG * kEncodeWarps);
      mbarInit(bar(BAR_FEAT_EMPTY + b), 1);
      mbarInit(bar(BAR_ACT_READY + b), CG * kEpiWarps);
    }
    for (int b = 0; b < (int)kSlots; ++b) { mbarInit(bar(BAR_ACC_FULL + b), 1); mbarInit(bar(BAR_ACC_EMPTY + b), CG * kEpiWarps); }
    *reinterpret_cast<volatile uint32_t*>(smem + L.bars + 8 * SLOT_TURN) = 0u;
    mbarInit(bar(BAR_OUT_FULL), 1);
    mbarInit(bar(BAR_OUT_EMPTY), CG * kOutWarps);
    fenceBarrierInit();
  }
  if (warp == 1) {
    if (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + L.tmemSlot), "r"(kTmemCols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + L.tmemSlot), "r"(kTmemCols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tcFenceBefore();
  __syncthreads();
  if (CL == 2) clusterSync();          // the peer's barriers are initialised before anyone signals them
  tcFenceAfter();
  const uint32_t tmemBase = *reinterpret_cast<volatile uint32_t*>(smem + L.tmemSlot);
  const uint32_t stageBytes = L.stageBytes;

  if (warp == 0) {
    // ================= weight producer (converged warp, one elected lane issues) =================
    const bool leader = electOne();
    uint32_t stage = 0, phase = 0;
    long long qEmpty = 0; const long long qStart = PROF_T(); (void)qStart;
    for (uint32_t unit = unit0; unit < numUnits; unit += unitStride) {
      for (int c = 0; c < numChunks; ++c) {
        const uint32_t bytes = (uint32_t)m.chunks[c].bytes16 * (16u / CG);      // this CTA's share of the chunk
        { const long long t0 = PROF_T(); mbarWait<CG>(bar(BAR_W_EMPTY + stage), phase ^ 1u); PROF_ADD(qEmpty, t0); }
        if (leader) {
          if (MC) {
            // my half of the chunk goes to both rings; the peer sends the other half
            const uint32_t half = bytes / 2u;
            mbarArriveExpectTx(bar(BAR_W_FULL + stage), bytes);
            bulkCopyG2SMulticast(sbase + L.ring + stage * stageBytes + rank * half, m.wStream + m.chunks[c].offset + rank * half, half,
                                 bar(BAR_W_FULL + stage), (uint16_t)3);
          } else {
            mbarArriveExpectTx(bar(BAR_W_FULL + stage), bytes);
            bulkCopyG2S(sbase + L.ring + stage * stageBytes, m.wStream + m.chunks[c].offset + rank * bytes, bytes, bar(BAR_W_FULL + stage));
          }
        }
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
#ifdef B200PT_PROFILE
    if (lane == 0) { unsigned long long* p = g_tcProf + blockIdx.x * 16 + 13; p[0] = qEmpty; p[1] = clock64() - qStart; }
#endif
    __syncwarp();
  } else if (warp == 1 && CG == 2 && rank != 0) {
    // ================= weight-arrival relay (peer CTA of a pair) =================
    // cp.async.bulk can only signal a barrier of its destination CTA, so this warp forwards
    // "my half of the stage has landed" to the issuer's barrier in the rank-0 CTA.
    uint32_t stage = 0, phase = 0;
    for (uint32_t unit = unit0; unit < numUnits; unit += unitStride) {
      for (int c = 0; c < numChunks; ++c) {
        mbarWait<CG>(bar(BAR_W_FULL + stage), phase);
        if (lane == 0) mbarArriveCluster(mapToCta(bar(BAR_PEER_W_FULL + stage), 0u));
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1 || (CG == 1 && warp == (uint32_t)kSecondIssuerWarp)) {
    // ================= MMA issuers (converged warps, one elected lane issues) =================
    // tcgen05.mma issue blocks while the (shallow) tensor queue is full, so whatever the issuing warp does
    // between the last MMA of a chunk and the first of the next - decode the chunk, poll its barriers, build
    // descriptors - leaves the tensor pipe idle. With one CTA per MMA group two warps take the chunks in turn:
    // while one issues, the other has already waited for and prepared the next chunk and only waits for the
    // "turn" word; the hand-over costs one shared-memory round trip.
    constexpr bool kDual = (CG == 1);
    const uint32_t me = (warp == 1) ? 0u : 1u;
    const uint32_t turnAddr = sbase + L.bars + 8u * SLOT_TURN;
    const bool leader = electOne();
    // Descriptor start addresses are 14-bit fields of (address & 0x3FFFF) >> 4: in a cluster launch the
    // shared-window address of an odd-ranked CTA carries the rank above bit 18 and must be masked off.
    const uint32_t actLo = (((sbase + L.act) & 0x3FFFFu) >> 4);
    const uint32_t ringLo = (((sbase + L.ring) & 0x3FFFFu) >> 4);
    uint32_t stage = 0, phase = 0;
    uint32_t actUse0 = 0, actUse1 = 0;     // completed waits on act_ready[part]
    uint32_t wLanded = 0;                  // the current chunk's weights were seen landed by the early poll
    uint32_t it = 0;
    long long pFeat = 0, pAcc = 0, pAct = 0, pW = 0, pOutE = 0, pPeer = 0, pIssue = 0; const long long pStart = PROF_T(); (void)pStart;
    for (uint32_t unit = unit0; unit < numUnits; unit += unitStride, ++it) {
      const uint32_t fb = TS ? 0u : (it & 1u), fpar = TS ? (it & 1u) : ((it >> 1) & 1u);
      const uint32_t featLo = (((sbase + L.feat[fb]) & 0x3FFFFu) >> 4);
      { const long long t0 = PROF_T(); mbarWait<CG>(bar(BAR_FEAT_FULL + fb), fpar); PROF_ADD(pFeat, t0); }
      for (int c = 0; c < numChunks; ++c) {
        const NifTcChunk& ch = m.chunks[c];
        const uint32_t ksteps = ch.ksteps, n8 = ch.n8, flags = ch.flags, waitAct = ch.waitAct;
        const uint32_t g = it * (uint32_t)numChunks + (uint32_t)c;        // running chunk index = turn number
        if (kDual && (g & 1u) != me) {
          // the other warp's chunk: only keep the counters in step
          if (waitAct == 1) actUse0 += 1; else if (waitAct == 2) actUse1 += 1;
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
          continue;
        }
        PROF_MARK(it == kTraceIt && lane == 0, c * 4 + 0);
        const bool isOut = (flags & kChOut) != 0;
        // the two half-jobs of a layer accumulate concurrently (K order above): the accumulator slot
        // and how often the slot was used before in this tile come precomputed with the chunk
        const uint32_t buf = ch.slot, accPhase = (it * m.slotUses[ch.slot] + ch.slotUse) & 1u;
        if (flags & kChFirst) {                                    // claim the accumulator
          const long long t0 = PROF_T();
          if (isOut) { mbarWait<CG>(bar(BAR_OUT_EMPTY), (it & 1u) ^ 1u); PROF_ADD(pOutE, t0); }
          else { mbarWait<CG>(bar(BAR_ACC_EMPTY + buf), accPhase ^ 1u); PROF_ADD(pAcc, t0); }
        }
        { const long long t0 = PROF_T();
          if (waitAct == 1) { mbarWait<CG>(bar(BAR_ACT_READY + 0), actUse0 & 1u); actUse0 += 1; }
          else if (waitAct == 2) { mbarWait<CG>(bar(BAR_ACT_READY + 1), actUse1 & 1u); actUse1 += 1; }
          PROF_ADD(pAct, t0); }
        { const long long t0 = PROF_T();
          if (kDual || !wLanded) mbarWait<CG>(bar(BAR_W_FULL + stage), phase);       // single issuer: usually polled ahead, see below
          PROF_ADD(pW, t0);
          const long long t1 = PROF_T();
          if (CG == 2) mbarWait<CG>(bar(BAR_PEER_W_FULL + stage), phase);
          PROF_ADD(pPeer, t1); }
        tcFenceAfter();
        PROF_MARK(it == kTraceIt && lane == 0, c * 4 + 1);
        const uint32_t dT = tmemBase + (isOut ? outCol : accCol0 + buf * (uint32_t)NH);
        const uint32_t bRows = n8 * (8u / CG);                                        // rows of B this CTA holds
        uint32_t aLo = (((flags & kChSrcFeat) ? featLo : actLo) + ch.aOff16) | ((2048u >> 4) << 16);
        uint32_t aTm = tmemBase + (uint32_t)(ch.aOff16 >> 8) * 8u;                     // TS: K-step k -> columns [8k, 8k+8)
        const bool aFromTmem = TS && !(flags & kChSrcFeat);
        uint32_t bLo = (ringLo + stage * (stageBytes >> 4)) | (bRows << 16);          // LBO = rows * 16 bytes
        const uint32_t idesc = umInstr<CG>(n8);
        const uint32_t bStep = bRows * 2u;                                            // rows * 32 bytes per K-step
        uint32_t acc = (flags & kChFirst) ? 0u : 1u;
        // The next chunk's weights are polled (non-blocking) between two MMAs of this chunk: the producer is
        // normally a full ring ahead, so the blocking wait - a dependent poll in the gap between chunks, where
        // the shallow tensor queue runs dry - is skipped almost always.
        const uint32_t nStage = (stage + 1 == (uint32_t)kStages) ? 0u : stage + 1u;
        const uint32_t nPhase = (stage + 1 == (uint32_t)kStages) ? (phase ^ 1u) : phase;
        if (kDual) {
          // everything is ready: wait until the other warp has issued the previous chunk
          const long long t0 = PROF_T();
          uint32_t cur, spins = 0;
          while (true) {
            asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(cur) : "r"(turnAddr) : "memory");
            if (cur == g) break;
            if (++spins > (1u << 28)) __trap();        // a protocol bug must fault, not hang the GPU
          }
          tcFenceAfter();
          PROF_ADD(pPeer, t0);
        }
        const long long tIssue = PROF_T();
        PROF_MARK(it == kTraceIt && lane == 0, c * 4 + 2);
        // Two separate loops: a predicated-off twin UTCHMMA still occupies a tensor-queue slot.
        if (aFromTmem) {
#pragma unroll
          for (uint32_t j = 0; j < (uint32_t)KC; ++j) {
            if (j == 1 && !kDual) wLanded = mbarTest(bar(BAR_W_FULL + nStage), nPhase);
            if (j < ksteps) {
              if (leader) tcMmaTs<CG>(dT, aTm, bLo, kDescHi, idesc, acc);
              aTm += 8u; bLo += bStep; acc = 1u;
            }
          }
        } else {
#pragma unroll
          for (uint32_t j = 0; j < (uint32_t)KC; ++j) {
            if (j == 1 && !kDual) wLanded = mbarTest(bar(BAR_W_FULL + nStage), nPhase);
            if (j < ksteps) {
              if (leader) tcMma<CG>(dT, aLo, bLo, kDescHi, idesc, acc);
              aLo += 256u; bLo += bStep; acc = 1u;
            }
          }
        }
        if (leader) {
          if (MC) tcCommitBoth(bar(BAR_W_EMPTY + stage)); else tcCommit<CG>(bar(BAR_W_EMPTY + stage));
          if (flags & kChLast) tcCommit<CG>(bar(isOut ? BAR_OUT_FULL : BAR_ACC_FULL + buf));
          if (flags & kChLastFeatUse) tcCommit<CG>(bar(BAR_FEAT_EMPTY + fb));
        }
        if (kDual) {
          tcFenceBefore();
          if (leader) asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(turnAddr), "r"(g + 1u) : "memory");
        }
        PROF_ADD(pIssue, tIssue);
        PROF_MARK(it == kTraceIt && lane == 0, c * 4 + 3);
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
#ifdef B200PT_PROFILE
    if (lane == 0 && me == 0) { unsigned long long* p = g_tcProf + blockIdx.x * 16; p[0] = clock64() - pStart; p[1] = pFeat; p[2] = pAcc; p[3] = pAct; p[4] = pW; p[5] = pOutE; p[6] = it; p[7] = pPeer; p[15] = pIssue; }
#endif
    __syncwarp();
  } else if (warp < kFirstEpiWarp) {
    // ================= feature encode =================
    const uint32_t r = (warp - kFirstEncodeWarp) * 32u + lane;      // row inside the tile
    const int E = m.embeddingDim, featCores = m.featKSteps * 2;
    uint32_t it = 0;
    for (uint32_t unit = unit0; unit < numUnits; unit += unitStride, ++it) {
      const uint32_t fb = TS ? 0u : (it & 1u), fpar = TS ? (it & 1u) : ((it >> 1) & 1u);
      const uint32_t row = (unit * CL + rank) * kTcRows + r;
      float2 uv = make_float2(0.f, 0.f);
      if (row < count) uv = io.uv[row];
      __half f[kMaxFeatures];
#pragma unroll
      for (int k = 0; k < kMaxFeatures; ++k) f[k] = __float2half_rn(0.f);
      // All octaves of one coordinate come from one table row (see b200pt_upload_nif). Values of p whose
      // binary16 form is subnormal or not finite take the arithmetic path (rounding no longer commutes there).
      auto encodeCoord = [&](float t, int base) {
        const float p = (t - 1.f) * 2.f;
        const float ap = fabsf(p);
        if (ap >= 6.103515625e-05f && ap < 65520.f) {
          const uint32_t* row = m.featTable + (size_t)__half_as_ushort(__float2half_rn(p)) * (uint32_t)E;
          for (int i = 0; i < E; ++i) {
            const uint32_t sc = __ldg(row + i);
            f[base + i] = __ushort_as_half((unsigned short)(sc & 0xffffu));
            f[2 * E + base + i] = __ushort_as_half((unsigned short)(sc >> 16));
          }
        } else {
          float coeff = 1.f;
          for (int i = 0; i < E; ++i, coeff *= 2.f) nifFeaturePair(p, coeff, f[base + i], f[2 * E + base + i]);
        }
      };
      encodeCoord(uv.x, 0);
      encodeCoord(uv.y, E);
      // wait until the MMAs of the unit that last used this buffer are done with it
      mbarWait<CG>(bar(BAR_FEAT_EMPTY + fb), fpar ^ 1u);
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
      __syncwarp();
      if (lane == 0) arriveIssuer(BAR_FEAT_FULL + fb);
    }
  } else if (warp < (uint32_t)kSecondIssuerWarp) {
    // ================= epilogue =================
    const uint32_t ew = warp - kFirstEpiWarp;
    const uint32_t quad = warp & 3u;                      // TMEM lane quadrant this warp may read
    const uint32_t colHalf = (warp - (uint32_t)kFirstEpiWarp) >> 2;   // which part of a job's NH columns
    const uint32_t r = quad * 32u + lane;                 // tile row == TMEM lane
    const uint32_t colsMine = (uint32_t)NH / (uint32_t)kEpiColParts;   // multiple of kEpiGroup
    const uint32_t groups = colsMine / (uint32_t)kEpiGroup;            // <= 5
    const uint32_t laneBase = tmemBase + ((quad * 32u) << 16);
    const __half2 one2 = __float2half2_rn(1.f);
    uint32_t it = 0;
    long long eWait = 0, eWork = 0, eOutWait = 0, eOutWork = 0; const long long eStart = PROF_T(); (void)eStart;
    for (uint32_t unit = unit0; unit < numUnits; unit += unitStride, ++it) {
      for (int l = 0; l < numHidden; ++l) {
        const bool relu = m.relu[l] != 0;
        const unsigned char* biasRow = smem + L.bias + (uint32_t)l * (uint32_t)H * 2u;
        for (uint32_t h = 0; h < 2; ++h) {
          const uint32_t jobInTile = 2u * (uint32_t)l + h;
          const uint32_t buf = jobInTile % kSlots, accPhase = (it * m.slotUses[buf] + jobInTile / kSlots) & 1u;
          { const long long t0 = PROF_T(); mbarWait<CG>(bar(BAR_ACC_FULL + buf), accPhase); PROF_ADD(eWait, t0); }
          PROF_MARK(it == kTraceIt && ew == 0 && lane == 0, 128 + jobInTile * 2);
          const long long tw = PROF_T();
          tcFenceAfter();
          const uint32_t tBase = laneBase + accCol0 + buf * (uint32_t)NH + colHalf * colsMine;
          const uint32_t colBase = h * (uint32_t)NH + colHalf * colsMine;      // output-neuron index of my first column
          constexpr int GW = kEpiGroup, GP = kEpiGroup / 2;     // accumulator columns / packed binary16 pairs per group
          // one group: round to binary16, + bias, ReLU, write the next layer's A operand
          auto emit = [&](uint32_t g, const uint32_t (&acc)[GW]) {
            const uint32_t c0 = colBase + g * (uint32_t)GW;
            uint32_t bw[GP];
            {
              const uint4 b0 = *reinterpret_cast<const uint4*>(biasRow + c0 * 2u);
              bw[0] = b0.x; bw[1] = b0.y; bw[2] = b0.z; bw[3] = b0.w;
              if (GP == 8) {
                const uint4 b1 = *reinterpret_cast<const uint4*>(biasRow + c0 * 2u + 16u);
                bw[GP - 4] = b1.x; bw[GP - 3] = b1.y; bw[GP - 2] = b1.z; bw[GP - 1] = b1.w;
              }
            }
            uint32_t packed[GP];
#pragma unroll
            for (int q = 0; q < GP; ++q) {
              __half2 hv = __floats2half2_rn(__uint_as_float(acc[2 * q]), __uint_as_float(acc[2 * q + 1]));
              // x * 1 + bias is one rounding, like the add; the relu variant clamps in the same instruction
              const __half2 bq = *reinterpret_cast<const __half2*>(&bw[q]);
              hv = relu ? __hfma2_relu(hv, one2, bq) : __hadd2(hv, bq);
              packed[q] = *reinterpret_cast<uint32_t*>(&hv);
            }
            if (TS) {
              tcStoreN(laneBase + c0 / 2u, packed);            // neuron c -> TMEM column c / 2 (binary16 pairs)
            } else {
              unsigned char* dst = smem + L.act + (c0 / 8u) * 2048u + r * 16u;
              *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
              if (GP == 8) *reinterpret_cast<uint4*>(dst + 2048) = make_uint4(packed[GP - 4], packed[GP - 3], packed[GP - 2], packed[GP - 1]);
            }
          };
          if (kEpiFrontLoad) {
            // all of this warp's accumulator columns are requested at once: one TMEM round trip per job instead of
            // one per group (the chain of five dependent round trips is what made a job ~720 cycles long)
            uint32_t v[5][GW];
#pragma unroll
            for (uint32_t g = 0; g < 5; ++g) if (g < groups) tcLoadN(tBase + g * (uint32_t)GW, v[g]);
            tcWaitLoad();
#pragma unroll
            for (uint32_t g = 0; g < 5; ++g) if (g < groups) emit(g, v[g]);
          } else {
            // software pipeline: the load of group g+1 is in flight while group g is converted and stored
            uint32_t v[2][GW];
            tcLoadN(tBase, v[0]);
#pragma unroll
            for (uint32_t g = 0; g < 5; ++g) {
              if (g < groups) {
                tcWaitLoad();
                if (g + 1 < groups) tcLoadN(tBase + (g + 1) * (uint32_t)GW, v[(g + 1) & 1]);
                emit(g, v[g & 1]);
              }
            }
          }
          if (TS) tcWaitStore();      // my activation columns are in tensor memory
          else fenceProxyAsync();     // my activation writes are visible to the tensor core (async proxy)
          tcFenceBefore();            // my TMEM accesses are ordered before the arrives below
          __syncwarp();
          if (lane == 0) { arriveIssuer(BAR_ACC_EMPTY + buf); arriveIssuer(BAR_ACT_READY + h); }
          PROF_MARK(it == kTraceIt && ew == 0 && lane == 0, 128 + jobInTile * 2 + 1);
          PROF_ADD(eWork, tw);
        }
      }
      // output job: N = 16 accumulator, channels 0..3 are the network outputs (B,G,R[,unused])
      if (colHalf == 0) {
        { const long long t0 = PROF_T(); mbarWait<CG>(bar(BAR_OUT_FULL), it & 1u); PROF_ADD(eOutWait, t0); }
        const long long tw = PROF_T();
        tcFenceAfter();
        uint32_t v[16];
        tcLoad16(laneBase + outCol, v);
        tcWaitLoad();
        tcFenceBefore();
        __syncwarp();
        if (lane == 0) arriveIssuer(BAR_OUT_EMPTY);
        const uint32_t row = (unit * CL + rank) * kTcRows + r;
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
  if (CL == 2) clusterSync();        // the peer may still be signalling barriers here / reading its TMEM half
  tcFenceAfter();
  if (warp == 1) {
    if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(kTmemCols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(kTmemCols) : "memory");
  }
}

size_t nifTcSmemBytes(const NifTcModel& m) {
  return (size_t)tcLayout(m.H, m.featKSteps, m.numHidden, m.cg, m.ts, m.kc).total + 1024;
}
int nifTcStages(const NifTcModel& m) { return (int)tcLayout(m.H, m.featKSteps, m.numHidden, m.cg, m.ts, m.kc).stages; }

namespace {
using KernelFn = void (*)(const NifTcModel, const NifIo);
KernelFn pickKernel(const NifTcModel& m) {
  if (m.mc) return m.kc == 10 ? nif_tc_kernel<1, true, 10, true> : nif_tc_kernel<1, true, 5, true>;
  if (m.ts && m.cg == 2) return m.kc == 10 ? nif_tc_kernel<2, true, 10, false> : nif_tc_kernel<2, true, 5, false>;
  if (m.ts) return m.kc == 10 ? nif_tc_kernel<1, true, 10, false> : nif_tc_kernel<1, true, 5, false>;
  return m.cg == 2 ? nif_tc_kernel<2, false, 5, false> : nif_tc_kernel<1, false, 5, false>;
}
} // namespace

cudaError_t prepareNifTc(const NifTcModel& m) {
  return cudaFuncSetAttribute(pickKernel(m), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)nifTcSmemBytes(m));
}

#ifdef B200PT_PROFILE
extern "C" int b200pt_debug_tc_profile(unsigned long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, g_tcProf, sizeof(unsigned long long) * n);
}
extern "C" int b200pt_debug_tc_trace(unsigned long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, g_tcTrace, sizeof(unsigned long long) * n);
}
#endif

// How many CTA pairs can be co-resident (cluster launch of 2 CTAs, 1 CTA per SM).
int nifTcMaxPairs(const NifTcModel& m) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2, 1, 1);
  cfg.blockDim = dim3(kTcThreads, 1, 1);
  cfg.dynamicSmemBytes = nifTcSmemBytes(m);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, pickKernel(m), &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

cudaError_t launchNifTc(const NifTcModel& model, const NifIo& io, int numSms, int maxPairs, cudaStream_t stream) {
  if (io.countHost == 0) return cudaSuccess;
  const size_t smem = nifTcSmemBytes(model);
  const uint32_t tilesMax = (io.countHost + kTcRows - 1) / kTcRows;
  uint32_t grid = tilesMax < (uint32_t)numSms ? tilesMax : (uint32_t)numSms;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kTcThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  cfg.attrs = attr; cfg.numAttrs = 0;
  if (model.cg == 2 || model.mc) {
    const uint32_t pairsMax = (tilesMax + 1) / 2;
    grid = 2 * (pairsMax < (uint32_t)maxPairs ? pairsMax : (uint32_t)maxPairs);
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.numAttrs = 1;
  }
  cfg.gridDim = dim3(grid, 1, 1);
  return cudaLaunchKernelEx(&cfg, pickKernel(model), model, io);
}

} // namespace b2
