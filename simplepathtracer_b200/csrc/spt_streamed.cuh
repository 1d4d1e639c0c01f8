// spt_streamed.cuh -- the megakernel for sphere lists that do not fit in shared memory next to four resident
// CTAs (BASELINE config 5: 16 384 spheres = 320 KB of pre-filter records + slot map).
//
// Same lane state machine, same device functions and therefore the same bits as trace_megakernel; what changes
// is how the closest-hit loop is fed.  The pre-filter blob is cut into chunks of kChunkSlots slots (10 KB:
// 8 KB of records + 2 KB of slot->index map) that stream through a kStages-deep ring in shared memory:
//
//   consumers  all 8 warps; per chunk: wait full[buf] (the bytes have landed), run the pre-filter span over it with
//              warp-uniform (broadcast) LDS.128, exact sequence on the candidates (geometry through L1/L2: ~2 per ray),
//              __syncwarp, one count on released[buf] per warp
//   producer   whoever RELEASES a buffer last: the warp whose count completes released[buf] knows that nobody reads the
//              buffer any more and refills it at once with the chunk kStages further on (mbarrier.arrive.expect_tx on
//              full[buf], two cp.async.bulk (TMA) copies global -> shared).  Nobody ever waits for a buffer to become
//              free.  (Round 1 had one elected thread producing kStages - 1 chunks ahead behind an "empty" mbarrier: ncu on
//              config 5 showed that thread spinning on it -- its warp, a consumer like the others, then ran late for the
//              whole CTA -- and 8 % of all executed instructions in try_wait loops.)
//
// so the CTA walks the list in loose lockstep (warps may drift by up to kStages - 1 chunks) and every record
// is fetched from L2 once per CTA instead of once per warp.  One __syncthreads_or per ray segment decides
// whether any lane of the CTA still has work; the ring is drained at every segment (the last kStages chunks are
// not refilled), so no copy is ever in flight when the CTA exits and thread 0 may start the next segment's first
// kStages chunks without asking.
#pragma once

#include "spt_kernels.cuh"

namespace spt
{

constexpr uint32_t kChunkSlots = 512;                                           // slots per chunk (16 mask words)
#ifndef SPT_RING_STAGES
#define SPT_RING_STAGES 3 // 30 KB of ring + the lane stash: four eight-warp CTAs per SM (four stages: three)
#endif
constexpr uint32_t kStages = SPT_RING_STAGES;                                   // ring depth
constexpr uint32_t kChunkRecordBytes = kChunkSlots / 2u * uint32_t(sizeof(FilterPair)); // 8192
constexpr uint32_t kChunkBytes = kChunkRecordBytes + kChunkSlots * 4u;          // 10240
constexpr uint32_t kStreamSmemBytes = kStages * kChunkBytes;
constexpr uint32_t kRingWarps = kBlockThreads / 32u;

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
// (try_wait suspends the thread until the phase completes or a system time limit passes; an explicit suspendTimeHint makes
// ptxas emit a NANOSLEEP loop that executes MORE instructions while waiting -- session 12 -- so there is none)
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

// The ring as the kernels carry it around
struct StreamRing
{
    unsigned char *ring;
    uint64_t *fullBar;
    uint32_t *released; // per buffer: warps that are through with its current tenant
    uint32_t seq;       // chunks this CTA has streamed so far: the ring position
};

// one thread: chunk `c` of the blob as the `seq`-th tenant of the ring (its buffer is known to be free)
__device__ __forceinline__ void stream_produce(const FilterScene &fs, const StreamRing &R, uint32_t seq, uint32_t c)
{
    uint32_t const buf = seq % kStages;
    uint32_t const slots = min(kChunkSlots, fs.nSlots - c * kChunkSlots);
    uint32_t const recBytes = slots / 2u * uint32_t(sizeof(FilterPair)), idxBytes = slots * 4u;
    unsigned char *dst = R.ring + buf * kChunkBytes;
    uint32_t const bar = smem_addr(R.fullBar + buf);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(recBytes + idxBytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
                 "l"(fs.pairs + size_t(c) * (kChunkSlots / 2u)), "r"(recBytes), "r"(bar)
                 : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst + kChunkRecordBytes)),
                 "l"(fs.slotIdx + size_t(c) * kChunkSlots), "r"(idxBytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void stream_ring_init(StreamRing &R, unsigned char *ring, uint64_t *fullBar, uint32_t *released)
{
    R.ring = ring;
    R.fullBar = fullBar;
    R.released = released;
    R.seq = 0;
    if (threadIdx.x == 0)
    {
        for (uint32_t s = 0; s < kStages; ++s)
        {
            mbar_init(fullBar + s, 1u);
            released[s] = 0u;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
}

// One pass of the whole blob through the ring: body(records, slot map, slots of the chunk) runs in every warp with a live
// lane, chunk after chunk.  Called by EVERY thread of the CTA, after a CTA-wide barrier that follows the previous pass (the
// ring is drained: thread 0 starts the first kStages chunks at once) and that also counted the warps with a live lane
// (stream_live_warps).  A warp without one takes no part: it neither waits for chunks nor counts on them -- it goes straight
// to the CTA's next barrier, which costs no issue slots (ncu, config 5: a tenth of the bounce kernel's issue slots went into
// try_wait loops of warps that had nothing to trace).
template <typename Body>
__device__ __forceinline__ void stream_pass(const FilterScene &fs, StreamRing &R, bool warpLive, uint32_t liveWarps, Body body)
{
    uint32_t const nChunks = (fs.nSlots + kChunkSlots - 1u) / kChunkSlots;
    if (threadIdx.x == 0)
        for (uint32_t c = 0; c < kStages && c < nChunks; ++c)
            stream_produce(fs, R, R.seq + c, c);
    if (warpLive)
        for (uint32_t c = 0; c < nChunks; ++c)
        {
            uint32_t const buf = (R.seq + c) % kStages, use = (R.seq + c) / kStages;
            mbar_wait(R.fullBar + buf, use & 1u);
            const unsigned char *chunk = R.ring + buf * kChunkBytes;
            body(reinterpret_cast<const FilterPair *>(chunk), reinterpret_cast<const uint32_t *>(chunk + kChunkRecordBytes),
                 min(kChunkSlots, fs.nSlots - c * kChunkSlots));
            __syncwarp(); // the warp's reads of the buffer are done
            if ((threadIdx.x & 31u) == 0u && atomicAdd(R.released + buf, 1u) == liveWarps - 1u)
            {
                // the last warp to let go of the buffer refills it (the others are past it; the next round of counts on this
                // buffer starts only after its new tenant has landed and been read)
                R.released[buf] = 0u;
                if (c + kStages < nChunks)
                    stream_produce(fs, R, R.seq + c + kStages, c + kStages);
            }
        }
    R.seq += nChunks;
}
// the CTA-wide barrier in front of a pass: how many warps have a live lane (0: nobody, the caller's loop ends)
__device__ __forceinline__ uint32_t stream_live_warps(bool warpLive)
{
    return uint32_t(__syncthreads_count(((threadIdx.x & 31u) == 0u && warpLive) ? 1 : 0));
}

// FindClosestIntersectionSphere (Collision.hpp:87-109) over the streamed list.  live = the lane has a ray; warpLive = some
// lane of the warp has (a warp without any only keeps the counts moving).  CAM: camera form of the records (fs = the
// eye-specialised blob).
template <bool CAM>
__device__ __forceinline__ uint32_t streamed_closest_hit(const FilterScene &fs, const float4 *geom, uint32_t n, V3 dir, V3 org,
                                                         StreamRing &R, uint32_t &exactEvals, bool live, bool warpLive, uint32_t liveWarps)
{
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    float const od = vdot(org, dir);
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    FilterRay const fr = CAM ? make_filter_ray_cam(dir, od, fs.epsTCam) : make_filter_ray(dir, org, od, fs.amax, fs.nearOO);
    stream_pass(fs, R, warpLive, liveWarps, [&](const FilterPair *rec, const uint32_t *slotIdx, uint32_t span) {
        filter_span<CAM>(rec, slotIdx, span_shape(span), fr, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    });
    return minIndex;
}

// Two camera rays of one lane (the same pixel's next two samples) through each streamed chunk: the camera form is
// bound by the shared-memory wavefronts of the record loads, which two rays share (spt_split.cuh).
__device__ __forceinline__ void streamed_closest_hit_cam2(const FilterScene &fs, const float4 *geom, uint32_t n, V3 org, V3 dirA, V3 dirB,
                                                          uint32_t &idxA, uint32_t &idxB, StreamRing &R, uint32_t &exactEvals, bool liveA,
                                                          bool liveB, bool warpLive, uint32_t liveWarps)
{
    idxA = idxB = n;
    float distA = FLT_MAX, distB = FLT_MAX;
    float const odA = vdot(org, dirA), odB = vdot(org, dirB);
    exact_big_spheres(fs, geom, dirA, org, odA, idxA, distA, exactEvals, liveA);
    exact_big_spheres(fs, geom, dirB, org, odB, idxB, distB, exactEvals, liveB);
    FilterRay const fa = make_filter_ray_cam(dirA, odA, fs.epsTCam), fb = make_filter_ray_cam(dirB, odB, fs.epsTCam);
    stream_pass(fs, R, warpLive, liveWarps, [&](const FilterPair *rec, const uint32_t *slotIdx, uint32_t span) {
        filter_span_cam2(rec, slotIdx, span_shape(span), fa, fb, geom, n, org, dirA, odA, dirB, odB, idxA, distA, idxB, distB, exactEvals, liveA, liveB);
    });
}

// cone_collect (spt_filter.cuh) over the streamed list: ONE pass of the whole blob through the ring per launch, every warp
// appending the survivors of each chunk to its lanes' lists.
template <int THREADS>
__device__ __forceinline__ void streamed_cone_collect(const FilterScene &fs, uint32_t n, V3 dir, V3 org, StreamRing &R, bool live, bool warpLive,
                                                      uint32_t liveWarps, uint16_t *list, uint32_t cap, uint32_t &count)
{
    FilterRay const fr = make_filter_ray_cam(dir, vdot(org, dir), fs.epsTCam);
    stream_pass(fs, R, warpLive, liveWarps, [&](const FilterPair *rec, const uint32_t *slotIdx, uint32_t span) {
        cone_collect<THREADS>(rec, slotIdx, span, fr, n, live, list, cap, count);
    });
}

__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) trace_streamed(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ __align__(8) uint64_t fullBar[kStages];
    __shared__ uint32_t released[kStages];
    __shared__ ShellSlots shellSlots[kBlockThreads / 32];
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];

    StreamRing R;
    stream_ring_init(R, ring, fullBar, released);

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
        // also the barrier after which no warp reads the ring any more: the next pass may start
        bool const warpLive = __any_sync(0xffffffffu, live);
        uint32_t const liveWarps = stream_live_warps(warpLive);
        if (!liveWarps)
            break;

        uint32_t const minIndex = streamed_closest_hit<false>(fs, geom, p.n, L.dir, L.org, R, exactEvals, live, warpLive, liveWarps);
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
