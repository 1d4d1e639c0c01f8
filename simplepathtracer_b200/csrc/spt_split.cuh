// spt_split.cuh -- split rendering (SPT_FLAG_SPLIT): the camera segment of every sample in its own kernel.
//
// Why: 44 % of all ray segments on the random-spheres scene are camera segments, and they all start at the eye.  For
// those, everything of the pre-filter's discriminant that does not involve the direction is a per-sphere constant of
// the launch (filter_pair_cam: 4 FFMA2 per record instead of 7 FFMA2 + 1 FADD2), but only a warp whose 32 lanes
// are ALL on camera rays can use the cheaper loop -- which the megakernel's flattened state machine never offers.
//
//   primary_hits_kernel   one lane = one pixel; for every sample of the chunk: jittered camera ray (the same counter-RNG
//                         draws the megakernel makes), closest hit with the camera form, first-hit index stored as
//                         u16 in HBM (2 B per sample: 17 GB per 4K x 1024-spp render, written and read once, in chunks
//                         of <= 2 GiB -- 5 ms of HBM time on a 1.6 s render).  Fully convergent, no shading.
//   trace_secondary       the bounce loop on SECONDARY rays only.  A lane that needs a new sample regenerates its
//                         camera ray, reads the stored index and either adds the sky (miss) and goes on to the next
//                         sample, or has the first hit shaded by a warp-wide shade_step of the fresh lanes and joins
//                         the closest-hit loop with its first secondary ray in the SAME iteration.
//
// Measured on config 3 (B200): primary 21.2 ms + secondary 84.3 ms per 64 spp against 109 ms for the megakernel
// (+5.4 %: 4 861 -> 5 123 Msamples/s).  The secondary kernel pays a second shade_step per iteration (fresh lanes, ~50 %
// of the warp) and the divergent sample-start loop; persistent lanes over a pixel queue were tried here as well and lose
// (94.8 ms: scattered pixels diverge in shading, 22.3 instead of 26.0 active threads per instruction).
#pragma once

#include "spt_kernels.cuh"
#include "spt_streamed.cuh"

namespace spt
{

// The camera segment of every sample of the launch, all lanes in step: jittered primary ray (the same counter-RNG
// draws trace_megakernel makes), closest hit with the camera form of the pre-filter (4 FFMA2 per record), index
// stored for trace_megakernel<.., SPLIT>.  No shading, no divergence outside the exact-candidate loop.
template <int STAGE>
__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) primary_hits_kernel(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ __align__(8) uint64_t mbar;
    uint32_t const geomBytes = p.n * uint32_t(sizeof(float4));
    uint32_t const blobBytes = filter_blob_bytes(p.filterCam.nSlots);
    const float4 *geom = p.geom;
    FilterScene fs = p.filterCam;
    if (STAGE == 2)
    {
        stage_scene_tma(smemRaw, p.geom, geomBytes, smemRaw + geomBytes, p.filterCam.pairs, blobBytes, &mbar);
        geom = reinterpret_cast<const float4 *>(smemRaw);
        rebase_filter(fs, smemRaw + geomBytes);
    }
    uint32_t x, y;
    bool inside;
    lane_pixel(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);
    uint64_t const pixKey = pixel_key(p.seed, pixel);
    V3 const eye = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
    uint32_t const nSamples = p.sampleEnd - p.sampleBegin;
    uint16_t *out = p.firstHits + (size_t(y - p.yBegin) * (p.xEnd - p.xBegin) + (x - p.xBegin)) * nSamples;
    uint32_t exactEvals = 0;
    if (!__any_sync(0xffffffffu, inside))
        return;
    for (uint32_t s = 0; s < nSamples; s += 2) // warp-uniform trip count; two samples of the pixel per trip
    {
        bool const second = s + 1 < nSamples;
        V3 dir[2];
#pragma unroll
        for (int k = 0; k < 2; ++k)
        {
            Rng rng;
            rng.key = stream_key(pixKey, p.sampleBegin + s + (second ? uint32_t(k) : 0u));
            rng.event = 0;
            rng.open_event();
            float xiU, xiV;
            rng.uniform2(-1.f, 1.f, xiU, xiV);
            dir[k] = primary_direction(p.cam, widthF, heightF, x, y, xiU, xiV);
        }
        uint32_t idxA, idxB;
        closest_hit_camera2(fs, geom, p.n, eye, dir[0], dir[1], idxA, idxB, exactEvals, inside, inside && second);
        if (inside)
        {
            out[s] = idxA < p.n ? uint16_t(idxA) : uint16_t(kMissU16);
            if (second)
                out[s + 1] = idxB < p.n ? uint16_t(idxB) : uint16_t(kMissU16);
        }
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, inside ? nSamples : 0u);
        warp_counter_add(p.counters + 1, inside ? nSamples : 0u);
        warp_counter_add(p.counters + 3, exactEvals);
    }
}


template <int STAGE>
__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) trace_secondary(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ ShellSlots shellSlots[kBlockThreads / 32];
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];
    uint32_t const geomBytes = p.n * uint32_t(sizeof(float4));
    uint32_t const blobBytes = filter_blob_bytes(p.filter.nSlots);
    const float4 *geom = p.geom;
    FilterScene fs = p.filter;
    if (STAGE == 2)
    {
        stage_scene_tma(smemRaw, p.geom, geomBytes, smemRaw + geomBytes, p.filter.pairs, blobBytes, &mbar);
        geom = reinterpret_cast<const float4 *>(smemRaw);
        rebase_filter(fs, smemRaw + geomBytes);
    }
    uint32_t x, y;
    bool inside;
    lane_pixel(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);

    Lane L;
    float4 const acc0 = inside ? p.accum[pixel] : make_float4(0.f, 0.f, 0.f, 0.f); // the running sum continues
    L.sum = v3(acc0.x, acc0.y, acc0.z);
    L.count = acc0.w;
    L.mode = MODE_NEW;
    L.sample = p.sampleBegin;
    L.dir = v3(0.f, 0.f, 1.f);
    L.org = v3(0.f, 0.f, 0.f);
    L.chain = v3(0.f, 0.f, 0.f);
    L.rng.state = 0;
    L.rng.key = 0;
    L.rng.event = 0;
    L.rng.word = 0;
    L.bounce = 0;
    L.specular = 0;
    uint32_t segments = 0, dropped = 0, exactEvals = 0;
    uint64_t const pixKey = pixel_key(p.seed, pixel);
    size_t const hitBase = (size_t(y - p.yBegin) * (p.xEnd - p.xBegin) + (x - p.xBegin)) * (p.sampleEnd - p.sampleBegin);

    bool live = inside;
    for (;;)
    {
        // ---- lanes that need a new sample: consume stored first hits until one yields a surface to shade ----
        bool fresh = false;
        uint32_t freshIdx = p.n;
        while (live && L.mode == MODE_NEW)
        {
            if (L.sample >= p.sampleEnd)
            {
                live = false;
                break;
            }
            uint32_t const h = p.firstHits[hitBase + (L.sample - p.sampleBegin)];
            L.rng.key = stream_key(pixKey, L.sample);
            L.rng.event = 0;
            L.rng.open_event();
            float xiU, xiV;
            L.rng.uniform2(-1.f, 1.f, xiU, xiV);
            L.dir = primary_direction(p.cam, widthF, heightF, x, y, xiU, xiV);
            L.org = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
            L.specular = 0;
            ++L.sample;
            if (h == kMissU16)
            {
                // SampleColorSkybox on a miss: terminal, no scatter, no random numbers
                L.sum = vadd(L.sum, sample_sky(p.cam, L.dir));
                L.count = fadd(L.count, 1.f);
                continue;
            }
            L.mode = MODE_TRACE;
            fresh = true;
            freshIdx = h;
        }
        if (!__any_sync(0xffffffffu, live))
            break;
        if (__any_sync(0xffffffffu, fresh))
            shade_step(p, geom, L, freshIdx, dropped, fresh, slots);
        bool const needTrace = live && L.mode != MODE_NEW; // a fresh hit may have ended its path (pass cap, SKYBOX sphere)
        uint32_t const idx = closest_hit_filtered(fs, geom, p.n, L.dir, L.org, exactEvals, needTrace);
        segments += needTrace ? 1u : 0u;
        shade_step(p, geom, L, idx, dropped, needTrace, slots);
    }
    if (inside)
        p.accum[pixel] = make_float4(L.sum.x, L.sum.y, L.sum.z, L.count); // one lane owns one pixel: plain store
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, segments);
        warp_counter_add(p.counters + 2, dropped);
        warp_counter_add(p.counters + 3, exactEvals);
    }
}

// ---- the same two kernels for lists that stream through the TMA chunk ring (spt_streamed.cuh) ---------------------
__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) primary_hits_streamed(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t fullBar[kStages];
    __shared__ __align__(8) uint64_t emptyBar[kStages];
    StreamRing R;
    stream_ring_init(R, ring, fullBar, emptyBar);
    const float4 *geom = p.geom;
    FilterScene const fs = p.filterCam;
    uint32_t x, y;
    bool inside;
    lane_pixel(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);
    uint64_t const pixKey = pixel_key(p.seed, pixel);
    V3 const eye = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
    uint32_t const nSamples = p.sampleEnd - p.sampleBegin;
    uint16_t *out = p.firstHits + (size_t(y - p.yBegin) * (p.xEnd - p.xBegin) + (x - p.xBegin)) * nSamples;
    uint32_t exactEvals = 0;
    bool const warpLive = __any_sync(0xffffffffu, inside);
    for (uint32_t s = 0; s < nSamples; ++s) // CTA-uniform trip count
    {
        __syncthreads(); // no warp reads the ring any more
        Rng rng;
        rng.key = stream_key(pixKey, p.sampleBegin + s);
        rng.event = 0;
        rng.open_event();
        float xiU, xiV;
        rng.uniform2(-1.f, 1.f, xiU, xiV);
        V3 const dir = primary_direction(p.cam, widthF, heightF, x, y, xiU, xiV);
        uint32_t const idx = streamed_closest_hit<true>(fs, geom, p.n, dir, eye, R, exactEvals, inside, warpLive);
        if (inside)
            out[s] = idx < p.n ? uint16_t(idx) : uint16_t(kMissU16);
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, inside ? nSamples : 0u);
        warp_counter_add(p.counters + 1, inside ? nSamples : 0u);
        warp_counter_add(p.counters + 3, exactEvals);
    }
}

__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) trace_secondary_streamed(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t fullBar[kStages];
    __shared__ __align__(8) uint64_t emptyBar[kStages];
    __shared__ ShellSlots shellSlots[kBlockThreads / 32];
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];
    StreamRing R;
    stream_ring_init(R, ring, fullBar, emptyBar);
    const float4 *geom = p.geom;
    FilterScene const fs = p.filter;
    uint32_t x, y;
    bool inside;
    lane_pixel(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);

    Lane L;
    float4 const acc0 = inside ? p.accum[pixel] : make_float4(0.f, 0.f, 0.f, 0.f); // the running sum continues
    L.sum = v3(acc0.x, acc0.y, acc0.z);
    L.count = acc0.w;
    L.mode = MODE_NEW;
    L.sample = p.sampleBegin;
    L.dir = v3(0.f, 0.f, 1.f);
    L.org = v3(0.f, 0.f, 0.f);
    L.chain = v3(0.f, 0.f, 0.f);
    L.rng.state = 0;
    L.rng.key = 0;
    L.rng.event = 0;
    L.rng.word = 0;
    L.bounce = 0;
    L.specular = 0;
    uint32_t segments = 0, dropped = 0, exactEvals = 0;
    uint64_t const pixKey = pixel_key(p.seed, pixel);
    size_t const hitBase = (size_t(y - p.yBegin) * (p.xEnd - p.xBegin) + (x - p.xBegin)) * (p.sampleEnd - p.sampleBegin);

    bool live = inside;
    for (;;)
    {
        bool fresh = false;
        uint32_t freshIdx = p.n;
        while (live && L.mode == MODE_NEW)
        {
            if (L.sample >= p.sampleEnd)
            {
                live = false;
                break;
            }
            uint32_t const h = p.firstHits[hitBase + (L.sample - p.sampleBegin)];
            L.rng.key = stream_key(pixKey, L.sample);
            L.rng.event = 0;
            L.rng.open_event();
            float xiU, xiV;
            L.rng.uniform2(-1.f, 1.f, xiU, xiV);
            L.dir = primary_direction(p.cam, widthF, heightF, x, y, xiU, xiV);
            L.org = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
            L.specular = 0;
            ++L.sample;
            if (h == kMissU16)
            {
                L.sum = vadd(L.sum, sample_sky(p.cam, L.dir));
                L.count = fadd(L.count, 1.f);
                continue;
            }
            L.mode = MODE_TRACE;
            fresh = true;
            freshIdx = h;
        }
        // also the barrier after which no warp reads the ring any more: the producer may start over
        if (!__syncthreads_or(live ? 1 : 0))
            break;
        if (__any_sync(0xffffffffu, fresh))
            shade_step(p, geom, L, freshIdx, dropped, fresh, slots);
        bool const needTrace = live && L.mode != MODE_NEW;
        bool const warpLive = __any_sync(0xffffffffu, needTrace);
        uint32_t const idx = streamed_closest_hit<false>(fs, geom, p.n, L.dir, L.org, R, exactEvals, needTrace, warpLive);
        segments += needTrace ? 1u : 0u;
        shade_step(p, geom, L, idx, dropped, needTrace, slots);
    }
    if (inside)
        p.accum[pixel] = make_float4(L.sum.x, L.sum.y, L.sum.z, L.count);
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, segments);
        warp_counter_add(p.counters + 2, dropped);
        warp_counter_add(p.counters + 3, exactEvals);
    }
}

} // namespace spt
