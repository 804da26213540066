/* b200pt.h - C ABI of libb200pt.so, the B200 (sm_100a) path-tracing hot path.
 *
 * This is the drop-in boundary for ONE path of markp-gc/ipu_path_trace: the per-pixel
 * Monte-Carlo loop of src/codelets plus the neural-image-field (NIF) environment-light
 * MLP of src/neural_networks. The reference has no C ABI; its seam is the five string-named
 * Poplar programs registered at src/PathTracerApp.cpp:479-483 and the StreamableTensor
 * host<->device handles of src/ipu_utils.hpp:288-373. One entry point below replaces each.
 * Citations are relative to the reference tree.
 *
 * Conventions: every function returns 0 on success and a negative b200pt_status otherwise;
 * b200pt_last_error() gives the message. No exception crosses this boundary. A context is
 * not thread-safe: one caller thread per context. The library owns all device memory; host
 * pointers are never retained after a call returns. There is no CPU fallback anywhere in
 * the library: without a CUDA device b200pt_create fails.
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200pt_ctx b200pt_ctx;

typedef enum b200pt_status {
  B200PT_OK = 0,
  B200PT_ERR_INVALID = -1,   /* bad argument / unsupported configuration */
  B200PT_ERR_CUDA = -2,      /* a CUDA runtime call or kernel failed */
  B200PT_ERR_STATE = -3,     /* call order violated (e.g. path_trace before upload_nif) */
  B200PT_ERR_NOMEM = -4
} b200pt_status;

/* Byte-compatible with the reference work item, src/codelets/TraceRecord.hpp:7-19 (20 B).
 * Items whose u >= width or v >= height are padding (LoadBalancer.cpp:66-71) and are skipped. */
#pragma pack(push, 1)
typedef struct b200pt_trace_record {
  uint16_t u, v;
  float r, g, b;
  uint16_t sampleCount, pathLength;
} b200pt_trace_record;
#pragma pack(pop)

enum { B200PT_AA_NORMAL = 0, B200PT_AA_UNIFORM = 1, B200PT_AA_TRUNCATED_NORMAL = 2 };
enum { B200PT_NIF_AUTO = 0, B200PT_NIF_TCGEN05 = 1, B200PT_NIF_SIMT = 2 };

/* Graph-build-time constants of the reference (IpuPathTraceJob.cpp:92-95,133-138;
 * RuntimeConfig ipu_utils.hpp:156-165 filled at PathTracerApp.cpp:74-93). */
typedef struct b200pt_config {
  uint32_t struct_size;        /* = sizeof(b200pt_config) */
  int32_t device;              /* CUDA device ordinal this context renders on */
  uint32_t width, height;      /* full image size, -w / -h */
  uint32_t max_path_length;    /* --max-path-length (10) */
  uint32_t roulette_depth;     /* --roulette-depth (3); 0 is rejected (SURVEY 8a quirk ii) */
  float stop_prob;             /* --stop-prob (0.3); rounded through binary16 like the half vertex field */
  float refractive_index;      /* -n (1.5); rounded through binary16 */
  int32_t aa_noise_type;       /* --aa-noise-type */
  uint32_t max_worklist;       /* capacity in work items (pixels incl. padding) of this context */
  uint32_t samples_per_launch; /* samples of every pixel traced per wavefront; 0 = automatic */
  uint32_t nif_impl;           /* B200PT_NIF_* */
} b200pt_config;

/* One Dense layer as Hdf5Model hands it over (src/keras/Hdf5Model.cpp:71-82,96-133):
 * kernel is [in,out] row-major; dtype 0 = binary16, 1 = binary32. */
typedef struct b200pt_nif_layer {
  int32_t in, out;
  int32_t relu;          /* activation == "relu" (NifModel.cpp:323); "linear" -> 0 (:73-76) */
  int32_t dtype;
  const void* kernel;
  const void* bias;      /* [out] or NULL */
} b200pt_nif_layer;

/* NifModel::Data + NifMetaData (NifModel.cpp:40-85, NifMetaData.cpp:11-71). */
typedef struct b200pt_nif_desc {
  int32_t num_layers;
  int32_t embedding_dimension;
  float max;             /* encode_params.max */
  float mean[3];         /* encode_params.mean with eps already folded in (NifMetaData.cpp:48-53) */
  int32_t log_tone_map;
  const b200pt_nif_layer* layers;
} b200pt_nif_desc;

/* Replaces the three cycle counters of read_results (PathTracerApp.cpp:441,449,461,470-475)
 * with device-measured times plus the path statistics the roofline needs. Totals are for the
 * most recent b200pt_path_trace call. */
typedef struct b200pt_counters {
  uint64_t paths;          /* camera paths traced (valid pixels x samples) */
  uint64_t escaped;        /* paths whose terminal record is ESCAPED = NIF evaluations */
  uint64_t path_segments;  /* sum of path lengths ("rays", PathTracerApp.cpp:763-776) */
  double trace_ms;         /* device time in the path-tracing kernels */
  double nif_ms;           /* device time in the NIF kernels */
  double total_ms;         /* device time of the whole program */
  uint64_t kernel_launches;
  int32_t nif_impl_used;   /* B200PT_NIF_TCGEN05 or B200PT_NIF_SIMT */
} b200pt_counters;

/* ---- lifetime -------------------------------------------------------------------------- */
int b200pt_create(const b200pt_config* cfg, b200pt_ctx** out);
void b200pt_destroy(b200pt_ctx* ctx);
/* ctx may be NULL: returns the message of the last failed b200pt_create on this thread. */
const char* b200pt_last_error(const b200pt_ctx* ctx);
const char* b200pt_version(void);

/* ---- the five programs (PathTracerApp.cpp:479-483) ----------------------------------------- */

/* "init_nif_weights" (NifModel.cpp:357-401; re-runnable mid-render, PathTracerApp.cpp:548-557). */
int b200pt_upload_nif(b200pt_ctx* ctx, const b200pt_nif_desc* nif);

/* The scene as data. The reference hard-codes its scene inside RayTraceKernel::compute and says so ("The scene is currently
 * hard coded here", codelets.cpp:110; the struct comment at :90-92 anticipates passing it in). A context starts with that
 * scene (5 spheres + 1 disc, no emitter); b200pt_set_scene replaces it: up to 16 spheres / discs with the reference's three
 * materials, optionally emissive (light::Material, EMIT records: codelets.cpp:192-196). n = 0 restores the built-in scene.
 * Takes effect at the next path_trace. */
typedef struct b200pt_primitive {
  int32_t shape;        /* 0 sphere (light::Sphere), 1 disc (light::Disc) */
  int32_t material;     /* 0 diffuse, 1 specular, 2 refractive (light::Material::Type) */
  float centre[3];
  float normal[3];      /* disc only; normalised like light::Disc's constructor */
  float radius;
  float colour[3];
  float emission[3];    /* emissive iff any component is non-zero */
} b200pt_primitive;
int b200pt_set_scene(b200pt_ctx* ctx, const b200pt_primitive* prims, uint32_t n);

/* "init_render_settings" (PathTracerApp.cpp:333-355,587-595): seed, anti-alias scale in pixels
 * and field of view in radians (both streamed as half by the reference and rounded likewise
 * here), HDRI azimuth in radians, on-device sample count of one path_trace program. */
int b200pt_set_render_settings(b200pt_ctx* ctx, uint64_t seed, float aa_scale_px, float fov_rad,
                               float azimuth_rad, uint32_t samples_per_step);

/* "setup" (PathTracerApp.cpp:426,692): upload the active worklist. r,g,b,sampleCount,pathLength
 * of the uploaded records seed the device accumulators (the reference host zeroes them). */
int b200pt_setup(b200pt_ctx* ctx, const b200pt_trace_record* worklist, size_t n);

/* The reference double-buffers its work list so that the host prepares list k+1 while the device renders list k
 * (PathTracerApp.cpp:703-709). b200pt_setup_next is that second buffer: it starts the upload of the NEXT step's work
 * list on a copy stream and returns at once, also while a path_trace is in flight; the list becomes the active one
 * at the next b200pt_path_trace (which therefore must come after the b200pt_read_results of the running step).
 * `worklist` must stay valid and unmodified until that path_trace call has returned; page-locked memory
 * (b200pt_host_alloc) makes the copy truly asynchronous. A later b200pt_setup discards a staged list. */
int b200pt_setup_next(b200pt_ctx* ctx, const b200pt_trace_record* worklist, size_t n);

/* Page-locked host memory for work lists and results, the counterpart of the host buffers the reference connects to
 * its streams (PathTracerApp.cpp:587-603). Pageable pointers are accepted everywhere; with these the 20-byte-record
 * transfers of setup / read_results run at the full host-link rate and do not block the calling thread. */
int b200pt_host_alloc(void** out, size_t bytes);
int b200pt_host_free(void* p);

/* "path_trace" (PathTracerApp.cpp:432-468,693): samples_per_step iterations of
 * camera ray -> path -> escaped-ray uv -> NIF -> accumulate, with no host involvement.
 * Asynchronous: returns once the work is enqueued; b200pt_sync waits for it.
 * Sample indices continue from the previous call (see b200pt_set_sample_index). */
int b200pt_path_trace(b200pt_ctx* ctx);
int b200pt_sync(b200pt_ctx* ctx);

/* Accumulation is order-independent: between setup and read_results / film_accumulate a work item's r,g,b live on the device as
 * signed 64-bit fixed point with 32 fraction bits and are added with integer atomics, so the same seed gives the same bits on
 * every run and for every split of the image over GPUs (the reference folds a pixel's samples in a fixed order,
 * codelets.cpp:235-306). Consequences for callers: uploaded r,g,b seeds are rounded to multiples of 2^-32, and the sum a work
 * item collects within one step must stay below 2^31 in magnitude. */
/* "read_results" (PathTracerApp.cpp:470-475,598-603): the worklist with accumulated r,g,b,
 * sampleCount and pathLength (the two counts wrap at 65536 exactly like the reference's
 * uint16 fields). counters may be NULL. Implies b200pt_sync. */
int b200pt_read_results(b200pt_ctx* ctx, b200pt_trace_record* out, size_t n, b200pt_counters* counters);

/* ---- native film: replaces AccumulatedImage (src/AccumulatedImage.cpp:59-74,23-56) ----------- */

/* hdr(v,u) += (b,g,r)/sampleCount for every valid work item, then clears the work-item
 * accumulators (LoadBalancer::clearInactiveAccumulators). Device side, no transfer. */
int b200pt_film_accumulate(b200pt_ctx* ctx);
int b200pt_film_clear(b200pt_ctx* ctx);
/* Copies scale * film (height x width x 3 fp32, channel order B,G,R) to a host buffer. */
int b200pt_film_read(b200pt_ctx* ctx, float* bgr_host, float scale);
/* Same into a device buffer on this context's device (e.g. an NCCL gather staging tensor). */
int b200pt_film_read_device(b200pt_ctx* ctx, void* bgr_device, float scale);
/* (hdr * scale * 2^exposure)^(1/gamma) * 255 saturated to uint8, B,G,R (AccumulatedImage.cpp:31-44). */
int b200pt_film_tonemap(b200pt_ctx* ctx, uint8_t* bgr8_host, float scale, float exposure, float gamma);

/* ---- control -------------------------------------------------------------------------- */
int b200pt_set_sample_index(b200pt_ctx* ctx, uint32_t next_sample);
int b200pt_get_counters(b200pt_ctx* ctx, b200pt_counters* counters);
/* FLOPs of one NIF evaluation by the reference's own formula (NifModel.cpp:122-134). */
int64_t b200pt_nif_flops_per_eval(const b200pt_ctx* ctx);

/* ---- stage-level entry points (kernel parity tests and the NIF microbenchmark) -------------- */

/* NIF inference for n (u,v) pairs -> n x 3 fp32 in the network's order B,G,R
 * (encode NifModel.cpp:185-218, layers :299-326, decode :221-245). Host pointers. */
int b200pt_nif_infer(b200pt_ctx* ctx, size_t n, const float* u, const float* v, float* bgr, int impl);
/* Same on device buffers, asynchronous on the context stream; device_ms (nullable) is filled
 * by b200pt_sync-time event timing of `repeats` back-to-back launches. */
int b200pt_nif_infer_device(b200pt_ctx* ctx, size_t n, const void* u_dev, const void* v_dev, void* bgr_dev,
                            int impl, int repeats, double* device_ms);

/* One wavefront of sample `sample` over the uploaded worklist WITHOUT shading: per work item
 * escaped flag, (u,v) of the env-map lookup, RGB path throughput, path length and the binary16
 * camera ray (GenerateCameraRays codelets.cpp:36-80, RayTraceKernel :93-227,
 * PreProcessEscapedRays :312-358). Host output arrays of n items; any may be NULL.
 * escaped: 0 the path ended without light (END), 1 it escaped (uv2 valid, thr3 = throughput to multiply the environment
 * radiance with), 2 it ended on an emitter of a b200pt_set_scene scene (thr3 = the radiance itself, no lookup). */
int b200pt_trace_wavefront(b200pt_ctx* ctx, uint32_t sample, size_t n, int32_t* escaped, float* uv2,
                           float* throughput3, int32_t* length, uint16_t* cam_rays2);

/* RayTraceKernel on caller-supplied inputs: n camera rays (2 binary16 each) and per_path
 * binary16 uniforms per ray, read sequentially with wrap (codelets.cpp:147-155). Outputs as above
 * plus the unit direction of the escaped ray. */
int b200pt_trace_fixed(b200pt_ctx* ctx, size_t n, const uint16_t* cam_rays2, const uint16_t* uniforms,
                       uint32_t per_path, int32_t* escaped, float* dir3, float* throughput3, int32_t* length);

/* Function-level known-answer entry points; each runs the device function the path-tracing
 * kernel itself calls. op selects the function, in/out are packed fp32 rows:
 *   B200PT_KAT_PIXEL_TO_RAY  in [x,y,w,h,fov]            out [dx,dy,dz]     light.hpp:194-206
 *   B200PT_KAT_SPHERE        in [o3,d3,c3,r]             out [t]            light.hpp:131-147
 *   B200PT_KAT_DISC          in [o3,d3,n3,c3,r]          out [t]            light.hpp:106-120
 *   B200PT_KAT_DIFFUSE       in [n3,u1,u2,rr]            out [d3,weight]    sdf.hpp:8-31
 *   B200PT_KAT_REFLECT       in [d3,n3]                  out [d3]           sdf.hpp:36-40
 *   B200PT_KAT_REFRACT       in [d3,n3,ri,rnd]           out [d3,refracted] sdf.hpp:45-67
 *   B200PT_KAT_ROULETTE      in [rnd,stopProb]           out [stop,weight]  light.hpp:25-29
 *   B200PT_KAT_ESCAPED_UV    in [d3,azimuth]             out [u,v]          codelets.cpp:333-347
 *   B200PT_KAT_PHILOX        in 6 words as raw bits      out 4 words        (RNG shared with the oracle)
 *   B200PT_KAT_AA_NOISE      in [w0,w1,type] raw bits    out [n0,n1]        PathTracerApp.cpp:29-45
 *   B200PT_KAT_NIF_FEATURES  in [u,v]                    out 4E features    NifModel.cpp:185-218
 */
enum {
  B200PT_KAT_PIXEL_TO_RAY = 0, B200PT_KAT_SPHERE, B200PT_KAT_DISC, B200PT_KAT_DIFFUSE, B200PT_KAT_REFLECT,
  B200PT_KAT_REFRACT, B200PT_KAT_ROULETTE, B200PT_KAT_ESCAPED_UV, B200PT_KAT_PHILOX, B200PT_KAT_AA_NOISE,
  B200PT_KAT_NIF_FEATURES
};
int b200pt_kat(b200pt_ctx* ctx, int op, size_t n, const float* in, int in_stride, float* out, int out_stride);

/* ---- second built-in scene: the Cornell box of external/light -------------------
 * Replaces the render loop of the reference's stand-alone CPU program external/light/src/trace.cpp
 * (scene :168-216, options :19-35, pixel loop :246-262; path loop trace.hpp:8-107). No context and no
 * NIF: paths end on emitters. Field defaults in comments are the reference's option defaults. */
enum { B200PT_CORNELL_PIXEL_STREAMS = 0,   /* one random stream per (pixel, sample slice): the GPU schedule */
       B200PT_CORNELL_TILE_STREAMS = 1 };  /* the reference's schedule: one jumped xoroshiro128** stream per tile job
                                              (jobs.hpp:66-90) consumed serially -> its image for a seed, bit for bit */
typedef struct b200pt_cornell_config {
  uint32_t struct_size;
  int32_t device;
  uint32_t width, height;            /* 256 x 256 */
  uint32_t tile_width, tile_height;  /* 16 x 16 (TILE_STREAMS only) */
  uint32_t samples;                  /* 32 */
  uint32_t roulette_depth;           /* 3 */
  uint64_t seed;                     /* 1 */
  float aa_noise_scale;              /* 1/700, camera-plane units */
  float refractive_index;            /* 1.5 */
  float stop_prob;                   /* 0.2 */
  float epsilon;                     /* FLT_EPSILON (light.cpp:5) */
  float intersection_epsilon;        /* 1e-5 (light.cpp:9) */
  int32_t rng_mode;                  /* B200PT_CORNELL_* */
} b200pt_cornell_config;
typedef struct b200pt_cornell_stats {
  uint64_t total_ray_casts;          /* sum of path depths (trace.cpp:301-306) */
  uint64_t max_path_length;
  double device_ms;                  /* kernel time (CUDA events) */
} b200pt_cornell_stats;
/* rgb_host: height x width x 3 fp32, R,G,B, the mean over `samples` (Image of jobs.hpp:6; p += colour / spp). */
int b200pt_cornell_render(const b200pt_cornell_config* cfg, float* rgb_host, b200pt_cornell_stats* stats);
/* Parity probe: path i starts at the origin towards (cam_xy[2i], cam_xy[2i+1], -1) with the xoroshiro128** state
 * rng_state[2i..2i+1] (advanced in place); rgb[3i..] = trace() (trace.hpp:8-107), depth[i] = its ray casts. */
int b200pt_cornell_trace_paths(const b200pt_cornell_config* cfg, size_t n, const float* cam_xy, uint64_t* rng_state,
                               float* rgb, uint32_t* depth);
const char* b200pt_cornell_last_error(void);

/* No counterpart in the reference: its renderer is a CPU program and owns its heap. This entry point exists for callers that are
 * not in a position to ask the CUDA driver for memory while they render: a plug-in that was handed one output buffer by its
 * host application and a fixed share of the GPU's memory besides, a frame loop that cannot afford cudaMalloc's device-wide
 * synchronisation, a process running beside an allocator that has already claimed the whole device. It renders what
 * b200pt_cornell_render renders with rng_mode = B200PT_CORNELL_TILE_STREAMS (any other rng_mode is rejected), same bits, but
 *   - the image goes to rgb_device, a DEVICE buffer of height x width x 3 fp32 that the caller owns and that lives on
 *     cfg->device; it is cleared and then written, and is complete when the call returns (the call is synchronous);
 *   - the kernel's working memory (the per-tile generator states and the two statistics counters) lives in an array that is
 *     part of the library image: 256 B + 16 B per tile, sized for B200PT_CORNELL_NOALLOC_MAX_TILES tiles (about 256 KiB).
 *     Every GPU on which this process uses the library keeps that array until the process exits;
 *   - ceil(width / tile_width) * ceil(height / tile_height) above that limit fails with B200PT_ERR_INVALID and a message that
 *     names both numbers; nothing is rendered;
 *   - neither cudaMalloc / cudaFree nor any other allocating call is made, on the device or on the host heap. That is also why
 *     stats->device_ms comes back as 0: timing needs events, and creating events allocates.
 * The first call on each device looks up the array's device address (and, like any first CUDA call, brings up the context and
 * loads the kernels). Make that first call before you begin capturing a stream. The array is shared by all calls on a device
 * and still holds what the previous call left in it; the function rewrites the parts it reads, so two calls in a row are
 * independent, but two calls that run on the same GPU at the same time (two host threads, same cfg->device) overwrite each
 * other's generator states and counters. Serialise them yourself. */
enum { B200PT_CORNELL_NOALLOC_MAX_TILES = 16384 };
int b200pt_cornell_render_noalloc(const b200pt_cornell_config* cfg, void* rgb_device, b200pt_cornell_stats* stats);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
