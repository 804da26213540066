// Microbenchmark: cycles per tcgen05.mma (M=128, K=16, fp16->fp32) for several operand placements.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_bench mma_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory"); }
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0; uint32_t spins = 0;
  while (true) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (++spins > (1u << 24)) __trap();
  }
}
__device__ __forceinline__ void tcCommit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mmaSS(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mmaTS(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((addr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}
// mode: 0 SS no-swizzle, 1 SS swizzle-128B, 2 TS (A in TMEM) + B no-swizzle, 3 TS + B swizzle-128B
// traffic: 0 none, 1 other warps stream STS.128 into a spare smem region, 2 bulk copies global->smem
__global__ void __launch_bounds__(256, 1) bench(int mode, int N, int ksteps, int reps, int traffic, const uint8_t* gsrc, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) unsigned long long bars[4];
  __shared__ uint32_t tmemSlot;
  const uint32_t sbase = smemAddr(smem);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbarInit(smemAddr(&bars[i]), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&tmemSlot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmemSlot;
  const uint32_t aBase = sbase, bBase = sbase + 96 * 1024, spare = sbase + 160 * 1024;
  __shared__ volatile int stopFlag;
  if (threadIdx.x == 0) stopFlag = 0;
  __syncthreads();
  if (warp == 0) {
    // whole warp converged: descriptors stay in uniform registers, only the MMA itself is predicated
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    uint32_t leader; asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(leader));
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      const uint32_t d = tmem + 160 + (r & 1) * 160;
      if (mode == 0) {
        for (int j = 0; j < ksteps; ++j) { if (leader) mmaSS(d, desc(aBase + j * 4096, 2048, 128, 0), desc(bBase + (j & 7) * N * 32, N * 16, 128, 0), idesc, j > 0); }
      } else if (mode == 1) {
        for (int j = 0; j < ksteps; ++j) { if (leader) mmaSS(d, desc(aBase + (j >> 2) * 16384 + (j & 3) * 32, 16, 1024, 2), desc(bBase + ((j >> 2) & 1) * N * 128 + (j & 3) * 32, 16, 1024, 2), idesc, j > 0); }
      } else if (mode == 2) {
        for (int j = 0; j < ksteps; ++j) { if (leader) mmaTS(d, tmem + j * 8, desc(bBase + (j & 7) * N * 32, N * 16, 128, 0), idesc, j > 0); }
      } else {
        for (int j = 0; j < ksteps; ++j) { if (leader) mmaTS(d, tmem + j * 8, desc(bBase + ((j >> 2) & 1) * N * 128 + (j & 3) * 32, 16, 1024, 2), idesc, j > 0); }
      }
    }
    if (leader) tcCommit(smemAddr(&bars[0]));
    mbarWait(smemAddr(&bars[0]), 0);
    const long long t1 = clock64();
    if (lane == 0) { out[blockIdx.x] = (unsigned long long)(t1 - t0); stopFlag = 1; }
  } else if (warp >= 4 && traffic == 1) {
    uint4 v = make_uint4(1, 2, 3, 4);
    uint32_t off = (threadIdx.x - 128) * 16;
    while (!stopFlag) {
#pragma unroll
      for (int k = 0; k < 8; ++k) { asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(spare + ((off + k * 2048) & 0x7fff)), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
    }
  } else if (warp >= 4 && traffic >= 3) {
    // epilogue-like TMEM traffic: every warp keeps loading (traffic 3) or loading+storing (4) its lane quadrant
    const uint32_t lb = tmem + (((warp & 3u) * 32u) << 16);
    uint32_t v[16];
    uint32_t col = 0;
    while (!stopFlag) {
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(lb + 160 + col) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (traffic >= 4) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
          ::"r"(lb + (traffic == 5 ? 80 + (col & 63) : 480 + (col & 15))), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      }
      col = (col + 16) % 144;
    }
  } else if (warp == 4 && lane == 0 && traffic == 2) {
    uint32_t ph = 0;
    while (!stopFlag) {
      const uint32_t bar = smemAddr(&bars[1]);
      asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(bar), "r"(25600) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(spare), "l"(gsrc + (blockIdx.x & 7) * 25600), "r"(25600), "r"(bar) : "memory");
      mbarWait(bar, ph); ph ^= 1;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}
int main() {
  unsigned long long* out; cudaMalloc(&out, 148 * 8);
  uint8_t* gsrc; cudaMalloc(&gsrc, 1 << 20); cudaMemset(gsrc, 0, 1 << 20);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const char* names[] = {"SS noswz", "SS swz128", "TS Bnoswz", "TS Bswz128"};
  for (int traffic : {4, 5})
    for (int mode = 0; mode < 4; ++mode)
      for (int N : {160}) {
        const int ksteps = 10, reps = 400;
        bench<<<148, 256, 200 * 1024>>>(mode, N, ksteps, reps, traffic, gsrc, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d N %d: %s\n", mode, N, cudaGetErrorString(e)); return 1; }
        unsigned long long h[148]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
        printf("traffic %d  %-10s N=%3d : %.1f cycles/MMA (floor %d)\n", traffic, names[mode], N, avg / (ksteps * reps), N / 2);
      }
  return 0;
}
