/*
 * spt_cuda.h -- C ABI of the B200-native (sm_100a) path-tracing hot path that replaces, as a
 * drop-in, the CPU tracers of ilia-glushchenko/SimplePathTracer.
 *
 * Plain C, plain pointers and sizes, no C++/torch types.  Citations are into the reference
 * tree (/root/reference/...).  What each entry point stands in for:
 *
 *   spt_cuda_render        void RenderSegment(RenderSegmentData)      include/SingleThreadPathTracer.hpp:114
 *                          void RenderSegmentTask(RenderSegmentData)  include/TaskBasedPathTracer.hpp:54
 *                          (selected by SptParams.pass_cap: 0 / 10), including io::WritePixel into the
 *                          caller's g_data (include/IOHelpers.hpp:17-22, include/Globals.hpp:19)
 *   spt_cuda_accumulate    the same per-pixel sample loop, leaving linear (sum r, sum g, sum b, count)
 *                          in a DEVICE float4 accumulator so that sample ranges / GPUs can be combined
 *   spt_cuda_resolve       the normalise + io::WritePixel tail of both tracers
 *                          (SingleThreadPathTracer.hpp:133-134, TaskBasedPathTracer.hpp:196-205)
 *   spt_cuda_progressive_* the per-tile / per-pass scheduling around them, include/Renderer.hpp:257-302, on a
 *                          context-owned accumulator (progressive refinement, checkpoint / resume)
 *   spt_cuda_first_hit     cd::FindClosestIntersectionSphere + contact helpers on jitter-free pinhole rays
 *   spt_cuda_closest_hit   cd::FindClosestIntersectionSphere  include/Collision.hpp:87-109
 *   spt_cuda_set_scene     the SoA scene globals  include/Globals.hpp:31-37
 *   spt_cuda_create_multi  the worker pool of RenderImageParallelMain  include/Renderer.hpp:257-302 (RenderJob :242-255):
 *                          ONE call spreads one render over all workers -- here over several GPUs of one node
 *
 * Inputs the reference keeps in globals (scene, camera, image size, sample/bounce counts) are passed
 * explicitly; the host mirror include/CudaPathTracer.hpp gathers them from the globals.
 *
 * There is NO CPU fallback: every function fails with a non-zero status when no CUDA device is usable.
 * All functions return 0 on success, otherwise an SptStatus / CUDA error code; text via spt_cuda_last_error().
 */
#ifndef SPT_CUDA_H
#define SPT_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifndef SPT_API
#if defined(__GNUC__)
#define SPT_API __attribute__((visibility("default"))) /* the library is built with -fvisibility=hidden */
#else
#define SPT_API
#endif
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Material values: include/Definitions.hpp:7-13 (enum class Material : uint8_t) */
enum { SPT_SKYBOX = 0, SPT_REFLECTIVE = 1, SPT_REFRACTIVE = 2, SPT_DIFFUSE = 3 };

enum SptStatus {
    SPT_OK = 0,
    SPT_ERR_INVALID = -1,   /* bad argument (null pointer, empty segment outside the image, w != 0, ...) */
    SPT_ERR_NO_DEVICE = -2, /* no usable CUDA device: there is no CPU path */
    SPT_ERR_NO_SCENE = -3,  /* spt_cuda_set_scene has not been called */
    SPT_ERR_CUDA = -4       /* a CUDA runtime call failed; see spt_cuda_last_error() */
};

/* Scene = the reference's SoA globals, host pointers laid out as the vectors' .data()
 * (Vec4 = 4 packed floats with w = 0, include/Math.hpp:9-14). */
typedef struct SptScene {
    uint32_t n;               /* g_sphereNumber */
    const float *centers4;    /* g_spheres   n*4 */
    const float *radii;       /* g_radii     n   */
    const uint8_t *materials; /* g_materials n   */
    const float *colors4;     /* g_colors    n*4, 0..255 radiance scale */
    const float *fuzz;        /* g_diffuses  n   */
} SptScene;

typedef struct SptCamera {
    float view[16]; /* viewMatrix.array, row-major (Globals.hpp:21; built as Renderer.hpp:321). Row 3 must be 0. */
    float eye[4];   /* eyePos (Globals.hpp:23), w = 0 */
    float sky[4];   /* initColor (Globals.hpp:29) */
} SptCamera;

/* == RenderSegmentData (include/Definitions.hpp:15-21): half-open pixel rectangle */
typedef struct SptSegment {
    uint32_t yBegin, yEnd, xBegin, xEnd;
} SptSegment;

enum { SPT_VARIANT_MEGAKERNEL = 0, SPT_VARIANT_WAVEFRONT = 1 };
enum {
    SPT_FLAG_STRICT_LOOP = 1u, /* closest-hit evaluates every sphere with the reference's exact op sequence
                                  (no conservative pre-filter); results are identical either way */
    SPT_FLAG_COUNTERS = 2u,    /* maintain SptCounters (cheap: per-thread registers, one atomic per warp) */
    SPT_FLAG_CROSS_CHECK = 4u, /* spt_cuda_closest_hit only: run pre-filter AND strict loop, report 0xFFFFFFFF on disagreement */
    SPT_FLAG_STREAMED = 8u,    /* megakernel: stream the sphere list through a TMA chunk ring in shared memory even when it
                                  would fit whole (the default for lists above SPT_RESIDENT_LIMIT_BYTES) */
    SPT_FLAG_SPLIT = 32u,      /* megakernel (the DEFAULT, resident and streamed lists): trace the camera segment of every sample
                                  in a separate, fully convergent kernel with the camera form of the pre-filter (one float4 record
                                  per sample -- direction + first-hit index -- staged in HBM in chunks bounded by the staging
                                  limit), then run the bounce loop on secondary rays only.  Same bits as the single kernel. */
    SPT_FLAG_SINGLE_KERNEL = 64u, /* megakernel: do not split; camera and secondary segments in one kernel */
    SPT_FLAG_RESIDENT = 16u,   /* megakernel: keep the whole list resident in shared memory whenever it fits at all, even at
                                  the price of fewer CTAs per SM */
    SPT_FLAG_SERIAL_CHUNKS = 128u, /* split rendering, measurement: every kernel of every chunk on the main stream, one after the
                                  other (no camera kernel under the tail of the previous bounce kernel), so that
                                  SptTiming.primary_ms / secondary_ms are the kernels' own durations */
    SPT_FLAG_GRID = 512u,      /* OPT-IN, outside the reference's algorithm (the acceleration structure for long lists): the
                                  bounce kernel (and, for lists on the chunk ring, the camera kernel) finds its closest hit by walking a uniform grid over the spheres (built at the first use
                                  after spt_cuda_set_scene) and running the reference's exact sequence on the spheres of the cells the ray
                                  crosses, instead of testing every sphere.  Same result, bit for bit -- the grid only skips spheres that
                                  provably cannot pass, ties and the origin-inside rule are the exact sequence's own -- but NOT the
                                  linear loop the roofline is defined on: SptCounters.tests then still counts segments x spheres, the
                                  work the reference would do.  Split rendering only; ignored by the other kernels.  spt_cuda_closest_hit
                                  takes it too (with SPT_FLAG_CROSS_CHECK: grid against strict loop, 0xFFFFFFFF on disagreement). */
    SPT_FLAG_OWN_PIXELS = 256u /* split rendering, comparison: every lane of the bounce kernel traces the samples of its own pixel
                                  only.  The default shares the samples of a warp's 8x4-pixel tile among its 32 lanes (a lane
                                  whose pixel is through takes samples of the pixel with the most left) and adds each pixel's
                                  samples in order afterwards: same bits, fewer idle lanes */
};
/* geometry + pre-filter bytes (36 B per sphere) up to which the whole list is staged once per CTA; four CTAs per SM
 * stay resident below it.  Longer lists stream (spt_streamed.cuh). */
#define SPT_RESIDENT_LIMIT_BYTES (48u * 1024u)

typedef struct SptParams {
    uint32_t width, height;            /* g_width, g_height (Globals.hpp:14-15) */
    uint32_t depth;                    /* g_bounces (Globals.hpp:12) */
    uint32_t pass_cap;                 /* 0: RenderSegment semantics (specular recursion unbounded);
                                          10: RenderSegmentTask's literal pass count (TaskBasedPathTracer.hpp:81):
                                          paths alive after pass_cap specular events are dropped, not counted */
    uint32_t sample_begin, sample_end; /* sample indices of every pixel; [0, g_samples) for a whole render */
    uint64_t seed;                     /* counter-based RNG seed (replaces Random.hpp:88-89's clock seed) */
    uint32_t variant;                  /* SPT_VARIANT_* */
    uint32_t flags;                    /* SPT_FLAG_* */
    uint32_t band_count, band_index;   /* band_count <= 1: every row of the segment.  Otherwise the segment is cut into
                                          bands of SPT_BAND_ROWS rows (band b = rows [yBegin + 8 b, yBegin + 8 b + 8)) and
                                          the call renders only the bands with b % band_count == band_index: the
                                          interleaved pixel partition for several GPUs (disjoint pixels, so summing the
                                          per-GPU accumulators reproduces the single-GPU bits).  The reference's analogue
                                          is RenderImageParallelMain's static tile grid (Renderer.hpp:264-274).
                                          Must be 0 / 1 on a multi-device context (the library partitions itself). */
    uint32_t split_chunk;              /* split rendering: samples of every pixel traced per pair of kernel launches.
                                          0 = chosen by the library from the staging limit (spt_cuda_set_staging_limit). */
    uint32_t strip_rows;               /* split rendering: 8-row tile rows of the segment covered per pair of kernel launches.
                                          0 = chosen by the library (all of them when the staging limit allows). */
} SptParams;
#define SPT_BAND_ROWS 8u

typedef struct SptCounters {
    uint64_t segments; /* closest-hit calls (one per ray segment) */
    uint64_t tests;    /* ray-sphere boolean tests the algorithm defines = segments * n */
    uint64_t samples;  /* camera samples traced */
    uint64_t dropped;  /* samples dropped by pass_cap */
    uint64_t exact;    /* sphere evaluations that went through the exact (reference-order) sequence */
    uint64_t warp_iterations; /* split rendering: closest-hit trips of the bounce kernel, counted per WARP; the share of lanes
                                 that held a ray in them is (segments - samples) / (32 * warp_iterations) */
} SptCounters;

typedef struct SptTiming {
    float trace_ms;     /* device time of the tracing kernel(s) of the last call, CUDA events on the context stream */
    float resolve_ms;   /* device time of the resolve kernel of the last call */
    uint32_t launches;  /* kernels launched by the last call */
    uint32_t reserved;
    float primary_ms;   /* split rendering: device time of the camera kernels of the last call, summed over its chunks ... */
    float secondary_ms; /* ... and of the bounce kernels.  Each kernel is bracketed by events on its own stream; with several
                           chunks the camera kernel of chunk k+1 overlaps the tail of the bounce kernel of chunk k, so the two
                           sums can exceed trace_ms (SPT_FLAG_SERIAL_CHUNKS makes them additive).  0 for the other kernels. */
} SptTiming;

typedef struct SptDeviceInfo {
    char name[64];
    int sm_count;
    int cc_major, cc_minor;
    int clock_khz;
    size_t smem_per_block_optin;
    size_t global_mem;
} SptDeviceInfo;

typedef struct SptContext SptContext;

/* lifetime.  `device` is a CUDA ordinal.  The context owns a stream, the device copy of the scene and scratch. */
SPT_API int spt_cuda_create(SptContext **out, int device);
/* One context over SEVERAL GPUs of this node (one host process; replaces the worker pool of RenderImageParallelMain,
 * Renderer.hpp:257-302).  devices[0] is the primary device.  The scene is replicated; spt_cuda_render and the
 * progressive entry points split every call over the devices -- sample ranges [k*S/n, (k+1)*S/n) of every pixel, or
 * interleaved 8-row bands when a call has fewer samples than devices (spt_cuda_set_partition) -- each device tracing into
 * its own float4 accumulator on its own stream; the accumulators are then combined over NVLink by a fused
 * reduce-and-resolve kernel (every device sums and encodes one slice of the rows, reading its peers' accumulators
 * through peer-mapped pointers in device order) and each device copies its slice of the image to the host buffer.
 * Band partitions leave the single-GPU bits (disjoint pixels); sample ranges re-associate n partial sums per pixel.
 * The device-pointer entry points (accumulate / resolve / set_stream) and the probes address devices[0] only.
 * Peer access between all pairs is required (SPT_ERR_CUDA otherwise).  The same ordinal may be listed twice (two
 * contexts on one GPU: how the path is tested on a single-GPU box). */
#define SPT_MAX_DEVICES 16
SPT_API int spt_cuda_create_multi(SptContext **out, const int *devices, int n_devices);
SPT_API int spt_cuda_device_count(SptContext *ctx); /* devices behind this context (1 for spt_cuda_create) */
enum { SPT_PARTITION_AUTO = 0, SPT_PARTITION_SAMPLES = 1, SPT_PARTITION_BANDS = 2 };
SPT_API int spt_cuda_set_partition(SptContext *ctx, int partition); /* multi-device contexts; AUTO = samples when a call has >= n of them */
SPT_API void spt_cuda_destroy(SptContext *ctx);
/* Split rendering stages one 16-byte record per camera sample in device memory.  `bytes` bounds the staging per device (two
 * buffers that alternate); 0 restores the default, SPT_DEFAULT_STAGING_BYTES.  A call that does not fit is cut into chunks of
 * samples (whole segment per pair of launches; strips of tile rows only when not even 16 samples of the segment fit); camera
 * kernels, and the bounce kernels of neighbouring strips, overlap on auxiliary streams.  Longer chunks keep more lanes busy
 * (config 3: 8 GiB = 32 samples per launch, 64 GiB = 256: +5 %), so an application that owns the GPU may raise the limit.
 * The library never sizes this from the free device memory. */
#define SPT_DEFAULT_STAGING_BYTES (8ull << 30)
SPT_API int spt_cuda_set_staging_limit(SptContext *ctx, uint64_t bytes);
/* run on a caller stream (cudaStream_t passed as void*; NULL is the legacy default stream) instead of
 * the context's own non-blocking stream; use_own != 0 restores the context's stream */
SPT_API int spt_cuda_set_stream(SptContext *ctx, void *cuda_stream, int use_own);
SPT_API int spt_cuda_device_info(SptContext *ctx, SptDeviceInfo *out);

/* upload the scene (copied; no host pointer is retained) */
SPT_API int spt_cuda_set_scene(SptContext *ctx, const SptScene *scene);

/* Whole hot path with HOST buffers: trace samples [sample_begin,sample_end) of every pixel of `seg`,
 * normalise, gamma-encode and write the segment's bytes into rgb8 (width*height*3, the caller's g_data;
 * pixel (x,y) lands at the reference's flipped slot, SingleThreadPathTracer.hpp:120).  accum4 (optional,
 * width*height*4 floats, row y*width+x NOT flipped) receives the linear sums and counts. Blocking. */
SPT_API int spt_cuda_render(SptContext *ctx, const SptCamera *cam, const SptParams *params, SptSegment seg,
                    uint8_t *rgb8, float *accum4);

/* Device-buffer pieces (asynchronous on the context stream):
 * d_accum4: width*height float4, caller-zeroed; each call ADDS its samples. */
SPT_API int spt_cuda_accumulate(SptContext *ctx, const SptCamera *cam, const SptParams *params, SptSegment seg,
                        float *d_accum4);
SPT_API int spt_cuda_resolve(SptContext *ctx, const float *d_accum4, uint32_t width, uint32_t height, SptSegment seg,
                     uint8_t *d_rgb8);
SPT_API int spt_cuda_synchronize(SptContext *ctx);

/* Progressive rendering on an accumulator OWNED BY THE CONTEXT (host programs that have no CUDA runtime of their own):
 * the per-tile / per-pass call pattern of RenderImageParallelMain (include/Renderer.hpp:257-302) and the live
 * preview that re-reads g_data every frame (Renderer.hpp:157-164,355).  The per-pixel running sums continue across
 * calls, so P passes of spp/P samples leave exactly the bits of one render of spp samples.
 *   begin  : (re)allocate and zero a width*height float4 accumulator
 *   add    : trace samples [sample_begin,sample_end) of `seg` into it (asynchronous)
 *   image  : normalise + gamma + pack the segment and copy it into rgb8 (the caller's g_data; may be NULL) and the
 *            linear sums/counts into accum4 (width*height*4 floats, may be NULL).  Blocking.
 *   load   : overwrite the accumulator from host memory (resume from a checkpoint written from `image`) */
SPT_API int spt_cuda_progressive_begin(SptContext *ctx, uint32_t width, uint32_t height);
SPT_API int spt_cuda_progressive_add(SptContext *ctx, const SptCamera *cam, const SptParams *params, SptSegment seg);
SPT_API int spt_cuda_progressive_image(SptContext *ctx, SptSegment seg, uint8_t *rgb8, float *accum4);
SPT_API int spt_cuda_progressive_load(SptContext *ctx, const float *accum4);

/* Parity probes (host buffers, blocking).  Outputs are indexed (x-xBegin) + (y-yBegin)*(xEnd-xBegin). */
SPT_API int spt_cuda_first_hit(SptContext *ctx, const SptCamera *cam, uint32_t width, uint32_t height, SptSegment seg,
                       uint32_t flags, uint32_t *idx, float *t, float *normal3);
/* exact_evals (optional, count entries): spheres per ray that went through the exact sequence */
SPT_API int spt_cuda_closest_hit(SptContext *ctx, uint32_t count, const float *dir3, const float *org3, uint32_t flags,
                         uint32_t *idx, uint32_t *exact_evals);

/* spt_cuda_closest_hit WITHOUT any device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, same kernel, same results as spt_cuda_closest_hit; the device staging of rays and answers is a
 * file-scope `static __device__` array of the library (SPT_STATIC_SCRATCH_RAYS * 32 bytes = 2 MiB, held by every GPU that
 * loads the library until the process exits) instead of the context's cudaMalloc'ed probe buffer.  The call allocates and
 * frees nothing (no cudaMalloc / cudaFree, no allocating library routine); spt_cuda_set_scene must have been called before.
 *   - count > SPT_STATIC_SCRATCH_RAYS is rejected with SPT_ERR_INVALID and a message naming the limit.
 *   - The array is ONE per device and process, shared by every context and every call, and keeps the last call's contents:
 *     each call first overwrites its own answer region (0xFF) and every word it returns is rewritten by its kernel, so nothing
 *     of an earlier call can be returned; but two calls that overlap on the same GPU (two contexts, or two host threads using
 *     two contexts) RACE on it.  Calls through one context are serialised by the context's lock.  Blocking. */
#define SPT_STATIC_SCRATCH_RAYS 65536u
SPT_API int spt_cuda_closest_hit_static(SptContext *ctx, uint32_t count, const float *dir3, const float *org3, uint32_t flags,
                                uint32_t *idx, uint32_t *exact_evals);

SPT_API int spt_cuda_get_counters(SptContext *ctx, SptCounters *out); /* accumulated since the last reset */
SPT_API int spt_cuda_reset_counters(SptContext *ctx);
SPT_API int spt_cuda_get_timing(SptContext *ctx, SptTiming *out);

/* FP32 CUDA-core peak of this device, measured with dependent-FMA chains on all SMs (the roofline
 * denominator BASELINE.md section 4 asks for).  scalar: fma.rn.f32; packed: fma.rn.f32x2. */
SPT_API int spt_cuda_measure_fp32_peak(SptContext *ctx, float *tflops_scalar, float *tflops_packed);

/* Device self-check of the shared-reciprocal division the kernels use in Normalize and in the camera ray (csrc/spt_device.cuh):
 * `sets` pseudo-random operand sets (rounded up to whole warps), each compared bit for bit with __fdiv_rn.  Blocking. */
SPT_API int spt_cuda_self_check_division(SptContext *ctx, uint64_t seed, uint64_t sets, uint64_t *tested, uint64_t *mismatches);

SPT_API const char *spt_cuda_last_error(void);
SPT_API const char *spt_cuda_version(void);

/* Layout revision of the structs above.  A caller compiled against another revision of this header must not call in
   (revision 2 appended band_count / band_index to SptParams, revision 3 split_chunk / strip_rows, SptTiming.primary_ms / secondary_ms,
   SptCounters.warp_iterations): compare with
   SPT_ABI_VERSION once after loading. */
#define SPT_ABI_VERSION 3u
SPT_API uint32_t spt_cuda_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SPT_CUDA_H */
