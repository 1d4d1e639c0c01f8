// spt_streamed.cuh -- the megakernel for sphere lists that do not fit in shared memory next to four resident
// CTAs (BASELINE config 5: 16 384 spheres = 320 KB of pre-filter records + slot map).
//
// Same lane state machine, same device functions and therefore the same bits as trace_megakernel; what changes
// is how the closest-hit loop is fed.  The pre-filter blob is cut into chunks of kChunkSlots slots (10 KB:
// 8 KB of records + 2 KB of slot->index map) that stream through a kStages-deep ring in shared memory:
//
//   producer   one elected thread; per chunk: wait empty[buf] (all 8 warps released the previous tenant),
//              mbarrier.arrive.expect_tx on full[buf], two cp.async.bulk (TMA) copies global -> shared
//   consumers  all 8 warps; per chunk: wait full[buf], run the pre-filter span over it with warp-uniform
//              (broadcast) LDS.128, exact sequence on the candidates (geometry through L1/L2: ~2 per ray),
//              __syncwarp, one mbarrier.arrive on empty[buf] per warp
//
// so the CTA walks the list in loose lockstep (warps may drift by up to kStages - 1 chunks) and every record
// is fetched from L2 once per CTA instead of once per warp.  One __syncthreads_or per ray segment decides
// whether any lane of the CTA still has work; the chunk ring is drained at every segment, so no copy is
// ever in flight when the CTA exits.
#pragma once

#include "spt_kernels.cuh"

namespace spt
{

constexpr uint32_t kChunkSlots = 512;                                           // slots per chunk (16 mask words)
constexpr uint32_t kStages = 4;                                                 // ring depth
constexpr uint32_t kChunkRecordBytes = kChunkSlots / 2u * uint32_t(sizeof(FilterPair)); // 8192
constexpr uint32_t kChunkBytes = kChunkRecordBytes + kChunkSlots * 4u;          // 10240
constexpr uint32_t kStreamSmemBytes = kStages * kChunkBytes;                    // 40960

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t const addr = smem_addr(bar);
    uint32_t done = 0;
    while (!done)
    {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(addr), "r"(parity)
                     : "memory");
    }
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

// producer side, one thread: chunk `c` of the blob as the `seq`-th tenant of the ring
__device__ __forceinline__ void stream_produce(const FilterScene &fs, unsigned char *ring, uint64_t *fullBar, uint64_t *emptyBar,
                                               uint32_t seq, uint32_t c)
{
    uint32_t const buf = seq % kStages, use = seq / kStages;
    if (use > 0u)
        mbar_wait(emptyBar + buf, (use - 1u) & 1u); // every warp has released the previous tenant
    uint32_t const slots = min(kChunkSlots, fs.nSlots - c * kChunkSlots);
    uint32_t const recBytes = slots / 2u * uint32_t(sizeof(FilterPair)), idxBytes = slots * 4u;
    unsigned char *dst = ring + buf * kChunkBytes;
    uint32_t const bar = smem_addr(fullBar + buf);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(recBytes + idxBytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
                 "l"(fs.pairs + size_t(c) * (kChunkSlots / 2u)), "r"(recBytes), "r"(bar)
                 : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst + kChunkRecordBytes)),
                 "l"(fs.slotIdx + size_t(c) * kChunkSlots), "r"(idxBytes), "r"(bar)
                 : "memory");
}

// The ring as the kernels carry it around
struct StreamRing
{
    unsigned char *ring;
    uint64_t *fullBar, *emptyBar;
    uint32_t seq; // chunks this CTA has streamed so far: the ring position
};

__device__ __forceinline__ void stream_ring_init(StreamRing &R, unsigned char *ring, uint64_t *fullBar, uint64_t *emptyBar)
{
    R.ring = ring;
    R.fullBar = fullBar;
    R.emptyBar = emptyBar;
    R.seq = 0;
    if (threadIdx.x == 0)
    {
        for (uint32_t s = 0; s < kStages; ++s)
        {
            mbar_init(fullBar + s, 1u);
            mbar_init(emptyBar + s, kBlockThreads / 32u);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
}

// FindClosestIntersectionSphere (Collision.hpp:87-109) over the streamed list.  Called by EVERY thread of the CTA, after
// a CTA-wide barrier that follows the previous call (no warp reads the ring any more: the producer starts over at
// chunk 0).  live = the lane has a ray; warpLive = some lane of the warp has (a warp without any only keeps the
// barriers moving).  CAM: camera form of the records (fs = the eye-specialised blob).
template <bool CAM>
__device__ __forceinline__ uint32_t streamed_closest_hit(const FilterScene &fs, const float4 *geom, uint32_t n, V3 dir, V3 org,
                                                         StreamRing &R, uint32_t &exactEvals, bool live, bool warpLive)
{
    constexpr uint32_t kAhead = kStages - 1u;
    uint32_t const nChunks = (fs.nSlots + kChunkSlots - 1u) / kChunkSlots;
    if (threadIdx.x == 0)
        for (uint32_t c = 0; c < kAhead && c < nChunks; ++c)
            stream_produce(fs, R.ring, R.fullBar, R.emptyBar, R.seq + c, c);
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    float const od = vdot(org, dir);
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    FilterRay const fr = CAM ? make_filter_ray_cam(dir, od, fs.epsTCam) : make_filter_ray(dir, org, od, fs.amax, fs.nearOO);
    for (uint32_t c = 0; c < nChunks; ++c)
    {
        if (threadIdx.x == 0 && c + kAhead < nChunks)
            stream_produce(fs, R.ring, R.fullBar, R.emptyBar, R.seq + c + kAhead, c + kAhead);
        uint32_t const buf = (R.seq + c) % kStages, use = (R.seq + c) / kStages;
        mbar_wait(R.fullBar + buf, use & 1u);
        if (warpLive)
        {
            const unsigned char *chunk = R.ring + buf * kChunkBytes;
            uint32_t const span = min(kChunkSlots, fs.nSlots - c * kChunkSlots);
            filter_span<CAM>(reinterpret_cast<const FilterPair *>(chunk), reinterpret_cast<const uint32_t *>(chunk + kChunkRecordBytes),
                             span, fr, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
        }
        __syncwarp();
        if ((threadIdx.x & 31u) == 0u)
            mbar_arrive(R.emptyBar + buf);
    }
    R.seq += nChunks;
    return minIndex;
}

// Two camera rays of one lane (the same pixel's next two samples) through each streamed chunk: the camera form is
// bound by the shared-memory wavefronts of the record loads, which two rays share (spt_split.cuh).
__device__ __forceinline__ void streamed_closest_hit_cam2(const FilterScene &fs, const float4 *geom, uint32_t n, V3 org, V3 dirA, V3 dirB,
                                                          uint32_t &idxA, uint32_t &idxB, StreamRing &R, uint32_t &exactEvals, bool liveA,
                                                          bool liveB, bool warpLive)
{
    constexpr uint32_t kAhead = kStages - 1u;
    uint32_t const nChunks = (fs.nSlots + kChunkSlots - 1u) / kChunkSlots;
    if (threadIdx.x == 0)
        for (uint32_t c = 0; c < kAhead && c < nChunks; ++c)
            stream_produce(fs, R.ring, R.fullBar, R.emptyBar, R.seq + c, c);
    idxA = idxB = n;
    float distA = FLT_MAX, distB = FLT_MAX;
    float const odA = vdot(org, dirA), odB = vdot(org, dirB);
    exact_big_spheres(fs, geom, dirA, org, odA, idxA, distA, exactEvals, liveA);
    exact_big_spheres(fs, geom, dirB, org, odB, idxB, distB, exactEvals, liveB);
    FilterRay const fa = make_filter_ray_cam(dirA, odA, fs.epsTCam), fb = make_filter_ray_cam(dirB, odB, fs.epsTCam);
    for (uint32_t c = 0; c < nChunks; ++c)
    {
        if (threadIdx.x == 0 && c + kAhead < nChunks)
            stream_produce(fs, R.ring, R.fullBar, R.emptyBar, R.seq + c + kAhead, c + kAhead);
        uint32_t const buf = (R.seq + c) % kStages, use = (R.seq + c) / kStages;
        mbar_wait(R.fullBar + buf, use & 1u);
        if (warpLive)
        {
            const unsigned char *chunk = R.ring + buf * kChunkBytes;
            uint32_t const span = min(kChunkSlots, fs.nSlots - c * kChunkSlots);
            filter_span_cam2(reinterpret_cast<const FilterPair *>(chunk), reinterpret_cast<const uint32_t *>(chunk + kChunkRecordBytes), span,
                             fa, fb, geom, n, org, dirA, odA, dirB, odB, idxA, distA, idxB, distB, exactEvals, liveA, liveB);
        }
        __syncwarp();
        if ((threadIdx.x & 31u) == 0u)
            mbar_arrive(R.emptyBar + buf);
    }
    R.seq += nChunks;
}

// cone_collect (spt_filter.cuh) over the streamed list: ONE pass of the whole blob through the ring per launch, every warp
// appending the survivors of each chunk to its lanes' lists.  Same calling rules as streamed_closest_hit.
template <int THREADS>
__device__ __forceinline__ void streamed_cone_collect(const FilterScene &fs, uint32_t n, V3 dir, V3 org, StreamRing &R, bool live, bool warpLive,
                                                      uint16_t *list, uint32_t cap, uint32_t &count)
{
    constexpr uint32_t kAhead = kStages - 1u;
    uint32_t const nChunks = (fs.nSlots + kChunkSlots - 1u) / kChunkSlots;
    if (threadIdx.x == 0)
        for (uint32_t c = 0; c < kAhead && c < nChunks; ++c)
            stream_produce(fs, R.ring, R.fullBar, R.emptyBar, R.seq + c, c);
    FilterRay const fr = make_filter_ray_cam(dir, vdot(org, dir), fs.epsTCam);
    for (uint32_t c = 0; c < nChunks; ++c)
    {
        if (threadIdx.x == 0 && c + kAhead < nChunks)
            stream_produce(fs, R.ring, R.fullBar, R.emptyBar, R.seq + c + kAhead, c + kAhead);
        uint32_t const buf = (R.seq + c) % kStages, use = (R.seq + c) / kStages;
        mbar_wait(R.fullBar + buf, use & 1u);
        if (warpLive)
        {
            const unsigned char *chunk = R.ring + buf * kChunkBytes;
            uint32_t const span = min(kChunkSlots, fs.nSlots - c * kChunkSlots);
            cone_collect<THREADS>(reinterpret_cast<const FilterPair *>(chunk), reinterpret_cast<const uint32_t *>(chunk + kChunkRecordBytes), span,
                                  fr, n, live, list, cap, count);
        }
        __syncwarp();
        if ((threadIdx.x & 31u) == 0u)
            mbar_arrive(R.emptyBar + buf);
    }
    R.seq += nChunks;
}

__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) trace_streamed(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t fullBar[kStages];
    __shared__ __align__(8) uint64_t emptyBar[kStages];
    __shared__ ShellSlots shellSlots[kBlockThreads / 32];
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];

    StreamRing R;
    stream_ring_init(R, ring, fullBar, emptyBar);

    const float4 *geom = p.geom; // exact geometry through L1/L2: only candidates and shading touch it
    FilterScene const fs = p.filter;
    uint32_t x, y;
    bool inside;
    lane_pixel(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);

    // the per-pixel running sum CONTINUES from the accumulator (zero for a fresh render): sample ranges traced one
    // after another give the bits of a single launch over the whole range (progressive refinement, resume)
    Lane L;
    float4 const acc0 = inside ? p.accum[pixel] : make_float4(0.f, 0.f, 0.f, 0.f);
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
    bool live = inside;
    for (;;)
    {
        if (live && L.mode == MODE_NEW)
        {
            if (L.sample >= p.sampleEnd)
            {
                live = false;
            }
            else
            {
                // next camera sample (SingleThreadPathTracer.hpp:123-128)
                L.rng.key = stream_key(pixKey, L.sample);
                L.rng.event = 0;
                L.rng.open_event();
                float xiU, xiV;
                L.rng.uniform2(-1.f, 1.f, xiU, xiV);
                L.dir = primary_direction(p.cam, widthF, heightF, x, y, xiU, xiV);
                L.org = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
                L.specular = 0;
                L.mode = MODE_TRACE;
                ++L.sample;
            }
        }
        // also the barrier after which no warp reads the ring any more: the producer may start over
        if (!__syncthreads_or(live ? 1 : 0))
            break;
        bool const warpLive = __any_sync(0xffffffffu, live);

        uint32_t const minIndex = streamed_closest_hit<false>(fs, geom, p.n, L.dir, L.org, R, exactEvals, live, warpLive);
        segments += live ? 1u : 0u;
        shade_step(p, geom, L, minIndex, dropped, live, slots);
    }
    if (inside)
    {
        float4 const a = make_float4(L.sum.x, L.sum.y, L.sum.z, L.count);
        p.accum[pixel] = a;
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, segments);
        warp_counter_add(p.counters + 1, inside ? (p.sampleEnd - p.sampleBegin) : 0u);
        warp_counter_add(p.counters + 2, dropped);
        warp_counter_add(p.counters + 3, exactEvals);
    }
}

} // namespace spt
