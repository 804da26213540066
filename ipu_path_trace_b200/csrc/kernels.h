// Host-visible declarations of the CUDA kernels' launchers and of the device-resident
// model descriptions. Internal to libb200pt (the public surface is include/b200pt.h).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "pt_common.cuh"

namespace b2 {

// ---- path tracing (trace_kernels.cu) -----------------------------------------------------
cudaError_t launchTraceWavefront(const SceneSoA* scene, const DeviceSettings& s, const uint32_t* pix, uint32_t nItems, uint32_t sampleBegin, uint32_t samplesPerLaunch,
                                 float2* qUv, float4* qThrSlot, uint32_t* qCount, uint32_t* sampleCount,
                                 uint32_t* pathLength, unsigned long long* stats, AccumSlot* accum, uint4* liveScratch, float4* escScratch, int variant,
                                 cudaStream_t stream);
// entries of the live-ray list (uint4 each) and of the escape list (2 x float4 each) for up to maxItems work items
size_t traceScratchEntries(uint32_t maxItems);
cudaError_t launchTraceItems(const SceneSoA* scene, const DeviceSettings& s, const uint32_t* pix, uint32_t nItems, uint32_t sample, int32_t* escaped, float* uv2,
                             float* thr3, int32_t* length, uint16_t* cam2, cudaStream_t stream);
cudaError_t launchTraceFixed(const SceneSoA* scene, const DeviceSettings& s, uint32_t n, const uint16_t* cam2, const uint16_t* uniforms, uint32_t perPath, int32_t* escaped,
                             float* dir3, float* thr3, int32_t* length, cudaStream_t stream);
cudaError_t launchKat(int op, uint32_t n, const float* in, int inStride, float* out, int outStride, cudaStream_t stream);

// ---- NIF ---------------------------------------------------------------------------------
constexpr int kMaxNifLayers = 16;
constexpr int kMaxFeatures = 64;       // 4 * embedding_dimension

// Where a NIF kernel reads its rows and what it does with the result.
struct NifIo {
  const float2* uv;          // (u,v) per row
  const float4* thrSlot;     // path throughput rgb + work-item slot (bits); null with bgrOut
  const uint32_t* count;     // device-resident row count (the wavefront queue length) or null
  uint32_t countHost;        // used when count == null
  float* bgrOut;             // n x 3 network outputs B,G,R (inference entry points) or null
  AccumSlot* accum;          // work-item accumulators (fixed point, pt_common.cuh) or null
};

// Generic description used by the SIMT kernel: fp32 copies of the binary16 parameters.
struct NifSimtModel {
  int numLayers, nFeat, embeddingDim, maxWidth;
  int in[kMaxNifLayers], out[kMaxNifLayers], relu[kMaxNifLayers], hasBias[kMaxNifLayers], concat[kMaxNifLayers];
  const float* w[kMaxNifLayers];     // [in,out] row-major
  const float* b[kMaxNifLayers];
  float maxv, mean[3];
  int logToneMap;
};
cudaError_t launchNifSimt(const NifSimtModel& m, const NifIo& io, cudaStream_t stream);
size_t nifSimtSmemBytes(const NifSimtModel& m);

// Description used by the tcgen05 kernel: every Dense layer on the tensor cores. A hidden layer's
// N = H outputs are computed as two half "jobs" (N = H/2 each, own TMEM accumulator); the final
// layer (<= 4 outputs) is one job padded to N = 16.
constexpr int kTcMaxChunks = 160;
enum : uint8_t { kChFirst = 1, kChLast = 2, kChSrcFeat = 4, kChLastFeatUse = 8, kChOut = 16 };
struct NifTcChunk {        // 16 bytes, read as one uint4 by the MMA thread
  uint32_t offset;         // byte offset into the packed weight stream
  uint16_t bytes16;        // bytes / 16
  uint16_t aOff16;         // byte offset / 16 of the first K-step inside the A buffer
  uint8_t ksteps;          // MMAs in this chunk
  uint8_t n8;              // N / 8 of those MMAs
  uint8_t flags;           // kCh*: first / last chunk of its job, A operand = feature buffer,
                           // last reader of the feature buffer in the tile, output-layer job
  uint8_t waitAct;         // 0 none, 1 wait for activation columns [0,H/2), 2 for [H/2,H): set on the
                           // first chunk of a layer that reads the range
  uint8_t layer, half;
  uint8_t slot;            // TMEM accumulator slot of the chunk's job: (2 * layer + half) % 3
  uint8_t slotUse;         // how many earlier jobs of the tile used that slot: (2 * layer + half) / 3
};
static_assert(sizeof(NifTcChunk) == 16, "chunk table entries are loaded as uint4");
struct NifTcModel {
  int numHidden, H, nFeat, embeddingDim, featKSteps, numChunks;
  int relu[kMaxNifLayers];
  int outC, finalRelu;
  int slotUses[3];             // jobs per tile on each accumulator slot
  int cg, ts, kc, mc;          // kernel variant: CTAs per MMA group, activations in TMEM, K-steps per chunk,
                               // weight chunks multicast between the two CTAs of a cluster
  const uint8_t* wStream;      // packed chunks (device)
  const __half* bias;          // [numHidden][H] (device; zeros where a layer has no bias)
  const uint32_t* featTable;   // [65536][E] (sin, cos) binary16 pairs of all E octaves for every binary16 value of p = (uv-1)*2
  float bOut[4];
  int finalHasBias;
  float maxv, mean[3];
  int logToneMap;
  NifTcChunk chunks[kTcMaxChunks];
};
cudaError_t launchNifTc(const NifTcModel& model, const NifIo& io, int numSms, int maxPairs, cudaStream_t stream);
size_t nifTcSmemBytes(const NifTcModel& m);
int nifTcStages(const NifTcModel& m);
cudaError_t prepareNifTc(const NifTcModel& m);
int nifTcMaxPairs(const NifTcModel& m);

// n rows of 4E binary16 features (as fp32) for the KAT entry point.
cudaError_t launchNifFeatures(uint32_t n, const float* uvIn, int inStride, float* out, int outStride, int embeddingDim, cudaStream_t stream);
// (u[], v[]) -> float2[]
cudaError_t launchPackUv(uint32_t n, const float* u, const float* v, float2* uv, cudaStream_t stream);

// ---- film / records (film_kernels.cu) ------------------------------------------------------
cudaError_t launchUnpackRecords(const TraceRecord* rec, uint32_t n, uint32_t* pix, AccumSlot* accum, uint32_t* sampleCount,
                                uint32_t* pathLength, cudaStream_t stream);
cudaError_t launchPackRecords(TraceRecord* rec, uint32_t n, const uint32_t* pix, const AccumSlot* accum, const uint32_t* sampleCount,
                              const uint32_t* pathLength, cudaStream_t stream);
cudaError_t launchFilmAccumulate(uint32_t n, const uint32_t* pix, AccumSlot* accum, uint32_t* sampleCount, uint32_t* pathLength,
                                 float* filmBgr, uint32_t width, uint32_t height, cudaStream_t stream);
cudaError_t launchFilmScale(const float* filmBgr, float* out, size_t n, float scale, cudaStream_t stream);
cudaError_t launchFilmTonemap(const float* filmBgr, uint8_t* out, size_t n, float scale, float exposure, float gamma, cudaStream_t stream);

} // namespace b2
