// CudaPathTracer.hpp -- the drop-in: RenderSegmentCuda / RenderSegmentCudaWavefront next to the reference's
//
//     void RenderSegment(RenderSegmentData)       /root/reference/include/SingleThreadPathTracer.hpp:114
//     void RenderSegmentTask(RenderSegmentData)   /root/reference/include/TaskBasedPathTracer.hpp:54
//
// Same signature, same implicit inputs (the scene / camera / config globals of Globals.hpp) and the same
// output (8-bit gamma-encoded RGB written in place into g_data at the reference's flipped pixel slots), so it
// slots into RenderJob (Renderer.hpp:242-255) or RenderImage (Renderer.hpp:304-308) by changing one call.
//
// Include it AFTER either
//   * the reference's own headers (Definitions.hpp, Globals.hpp, ...)  -> see INTEGRATION.md, or
//   * this repo's mirror "spt_host/Scene.hpp"                          -> headless builds.
// Both provide: RenderSegmentData, Material, math::Vec4 (.xyzw), math::Mat4 (.array), g_spheres, g_radii,
// g_materials, g_colors, g_diffuses, g_sphereNumber, viewMatrix, eyePos, initColor, g_width, g_height,
// g_samples, g_bounces, g_data.
//
// Link with libspt_cuda.so (include/spt_cuda.h).  There is no CPU fallback: any failure throws.
#pragma once

#include "spt_cuda.h"

#include <cstdint>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace spt_cuda
{
struct State
{
    SptContext *ctx = nullptr;
    std::mutex lock;
    bool sceneUploaded = false;
    std::vector<int> devices{0}; // CUDA ordinals; several = one render spread over several GPUs (spt_cuda_create_multi)
    uint64_t stagingLimit = 0;   // 0 = the library's default (SPT_DEFAULT_STAGING_BYTES per device)
    uint64_t seed = 1;   // counter-RNG seed (the reference seeds from the clock, Random.hpp:88-89)
    uint32_t flags = 0;  // SPT_FLAG_*
};

inline State &GetState()
{
    static State s;
    return s;
}

inline void Check(int rc, char const *what)
{
    if (rc != SPT_OK)
        throw std::runtime_error(std::string("CudaPathTracer: ") + what + ": " + spt_cuda_last_error());
}

// Call after the scene globals change (InitSpheres / GenerateSpheres); the first render uploads lazily.
inline void InvalidateScene()
{
    State &s = GetState();
    std::lock_guard<std::mutex> g(s.lock);
    s.sceneUploaded = false;
}

// Device selection; call before the first render (or after Shutdown()).  SetDevices({0, 1, ..., 7}) is the counterpart of the
// reference's worker pool (g_maxThreads workers in RenderImageParallelMain, Renderer.hpp:257-302): ONE RenderSegmentCuda call
// then spreads its samples (or 8-row bands) over all listed GPUs and the library combines their accumulators over NVLink.
inline void SetDevice(int device) { GetState().devices.assign(1, device); }
inline void SetDevices(std::vector<int> const &devices)
{
    if (devices.empty())
        throw std::runtime_error("CudaPathTracer: SetDevices needs at least one device");
    GetState().devices = devices;
}
inline void SetSeed(uint64_t seed) { GetState().seed = seed; }
// SPT_FLAG_* for every render that follows (e.g. SPT_FLAG_GRID: the opt-in acceleration structure for long sphere lists)
inline void SetFlags(uint32_t flags) { GetState().flags = flags; }
// first-hit staging budget of split rendering per device, in bytes (spt_cuda_set_staging_limit); 0 = library default
inline void SetStagingLimit(uint64_t bytes) { GetState().stagingLimit = bytes; }

inline SptContext *Acquire()
{
    State &s = GetState();
    std::lock_guard<std::mutex> g(s.lock);
    if (!s.ctx)
    {
        if (spt_cuda_abi_version() != SPT_ABI_VERSION) // this TU and the library disagree about the struct layouts
            throw std::runtime_error("CudaPathTracer: libspt_cuda.so was built from another revision of spt_cuda.h");
        if (s.devices.size() > 1)
            Check(spt_cuda_create_multi(&s.ctx, s.devices.data(), int(s.devices.size())), "spt_cuda_create_multi");
        else
            Check(spt_cuda_create(&s.ctx, s.devices[0]), "spt_cuda_create");
        if (s.stagingLimit)
            Check(spt_cuda_set_staging_limit(s.ctx, s.stagingLimit), "spt_cuda_set_staging_limit");
    }
    if (!s.sceneUploaded)
    {
        std::vector<uint8_t> mats(g_sphereNumber);
        for (uint32_t i = 0; i < g_sphereNumber; ++i)
            mats[i] = static_cast<uint8_t>(g_materials[i]);
        SptScene sc{};
        sc.n = g_sphereNumber;
        sc.centers4 = g_spheres[0].xyzw;
        sc.radii = g_radii.data();
        sc.materials = mats.data();
        sc.colors4 = g_colors[0].xyzw;
        sc.fuzz = g_diffuses.data();
        Check(spt_cuda_set_scene(s.ctx, &sc), "spt_cuda_set_scene");
        s.sceneUploaded = true;
    }
    return s.ctx;
}

inline void Shutdown()
{
    State &s = GetState();
    std::lock_guard<std::mutex> g(s.lock);
    if (s.ctx)
        spt_cuda_destroy(s.ctx);
    s.ctx = nullptr;
    s.sceneUploaded = false;
}

inline void FillCameraAndParams(SptCamera &cam, SptParams &p, uint32_t passCap, uint32_t variant)
{
    static_assert(sizeof(math::Vec4) == 16, "Vec4 must be 4 packed floats (Math.hpp:9-14)");
    for (int k = 0; k < 16; ++k)
        cam.view[k] = viewMatrix.array[k];
    // the tracers only ever multiply the matrix with a w = 0 vector and read lanes 0..2 of the result, so row 3
    // and column 3 never reach a pixel; the reference's identity initialiser has a 1 there, its camera basis a 0
    cam.view[3] = cam.view[7] = cam.view[11] = cam.view[12] = cam.view[13] = cam.view[14] = cam.view[15] = 0.f;
    for (int k = 0; k < 3; ++k)
    {
        cam.eye[k] = eyePos.xyzw[k];
        cam.sky[k] = initColor.xyzw[k];
    }
    p.width = g_width;
    p.height = g_height;
    p.depth = g_bounces;
    p.pass_cap = passCap;
    p.sample_begin = 0;
    p.sample_end = g_samples;
    p.seed = GetState().seed;
    p.variant = variant;
    p.flags = GetState().flags;
}

inline void Render(RenderSegmentData segment, uint32_t passCap, uint32_t variant)
{
    SptContext *ctx = Acquire();
    SptCamera cam{};
    SptParams p{};
    FillCameraAndParams(cam, p, passCap, variant);
    SptSegment const seg{segment.yBegin, segment.yEnd, segment.xBegin, segment.xEnd};
    Check(spt_cuda_render(ctx, &cam, &p, seg, g_data, nullptr), "spt_cuda_render");
}

// Progressive rendering of the whole image on the context-owned accumulator: the reference's live preview
// (UpdateTexture re-reads g_data every frame while RenderJob threads fill it, Renderer.hpp:157-164,282-290,355) and its
// progress lines (Renderer.hpp:285), as passes over the SAMPLES instead of tiles.  After pass k, g_data holds the image
// of the first samplesDone samples and onPass(k, passes, samplesDone) is called.  Passes leave the bits of one
// g_samples render.  firstSample > 0 resumes (accum4 = the checkpointed sums, width*height*4 floats).
// accumEveryPass: accumOut4 is refreshed before EVERY onPass call (checkpoint after each pass), not only the last.
template <typename OnPass>
inline void RenderImageProgressiveCuda(uint32_t passes, uint32_t passCap, uint32_t variant, OnPass onPass, uint32_t firstSample = 0,
                                       float const *resumeAccum4 = nullptr, float *accumOut4 = nullptr, bool accumEveryPass = false)
{
    SptContext *ctx = Acquire();
    SptCamera cam{};
    SptParams p{};
    FillCameraAndParams(cam, p, passCap, variant);
    SptSegment const all{0, g_height, 0, g_width};
    Check(spt_cuda_progressive_begin(ctx, g_width, g_height), "spt_cuda_progressive_begin");
    if (resumeAccum4)
        Check(spt_cuda_progressive_load(ctx, resumeAccum4), "spt_cuda_progressive_load");
    if (passes == 0)
        passes = 1;
    uint32_t const todo = g_samples > firstSample ? g_samples - firstSample : 0;
    for (uint32_t k = 0; k < passes; ++k)
    {
        p.sample_begin = firstSample + uint32_t(uint64_t(todo) * k / passes);
        p.sample_end = firstSample + uint32_t(uint64_t(todo) * (k + 1) / passes);
        Check(spt_cuda_progressive_add(ctx, &cam, &p, all), "spt_cuda_progressive_add");
        bool const last = k + 1 == passes;
        Check(spt_cuda_progressive_image(ctx, all, g_data, (last || accumEveryPass) ? accumOut4 : nullptr), "spt_cuda_progressive_image");
        onPass(k + 1, passes, p.sample_end);
    }
}
} // namespace spt_cuda

// RenderSegment on the GPU: recursive-tracer semantics (specular recursion unbounded), megakernel.
inline void RenderSegmentCuda(RenderSegmentData segment) { spt_cuda::Render(segment, 0u, SPT_VARIANT_MEGAKERNEL); }

// RenderSegmentTask on the GPU: queue-tracer semantics (10 passes, drop + per-pixel count,
// TaskBasedPathTracer.hpp:81,196-198), wavefront variant.
inline void RenderSegmentCudaWavefront(RenderSegmentData segment) { spt_cuda::Render(segment, 10u, SPT_VARIANT_WAVEFRONT); }

// RenderSegmentTask semantics on the megakernel (what bench.py times)
inline void RenderSegmentTaskCuda(RenderSegmentData segment) { spt_cuda::Render(segment, 10u, SPT_VARIANT_MEGAKERNEL); }
