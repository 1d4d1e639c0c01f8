// spt_wavefront.cuh -- the WAVEFRONT variant of the tracing kernel (SPT_VARIANT_WAVEFRONT).
//
// Mirrors the structure of the reference's queue tracer (RenderSegmentTask, TaskBasedPathTracer.hpp:54-206):
// per pass, every live path is intersected, then pushed to one of four per-material queues
// (Tasks{diffuse, reflective, refractive, skybox}, Definitions.hpp:33-39; TraceAndSampleColor(task, tasks),
// TaskBasedPathTracer.hpp:9-30) and each queue is processed on its own.  On the GPU:
//
//   * persistent CTAs: the grid is a fixed number of CTAs per SM; each pulls 32x8-pixel tiles from a global
//     atomic counter until the segment is exhausted;
//   * path state lives in shared memory as SoA, one slot per pixel of the tile (one sample in flight per
//     pixel, so per-pixel sums are added in sample order exactly as in the megakernel and the oracle);
//   * lists too long for shared memory stream through the megakernel family's TMA chunk ring (STAGE 3);
//   * every pass builds the queues with warp ballot + popc prefix + one shared atomic per warp, and the
//     stages run over the COMPACTED queues: thread k takes entry k, so all lanes of a warp intersect, or run
//     the same material handler, together.
//
// It computes exactly the megakernel's per-sample radiance (same device functions, same counter-based RNG),
// so the two variants' accumulators are bit-identical; what differs is where lanes idle, which is the point:
// the divergence cost of the megakernel can be measured against this kernel instead of guessed.
#pragma once

#include "spt_kernels.cuh"
#include "spt_streamed.cuh"

namespace spt
{

constexpr int kWfSlots = kBlockThreads; // one slot per pixel of the 32x8 tile

struct WfState // SoA over slots, in shared memory
{
    float ox[kWfSlots], oy[kWfSlots], oz[kWfSlots];
    float dx[kWfSlots], dy[kWfSlots], dz[kWfSlots];
    float cr[kWfSlots], cg[kWfSlots], cb[kWfSlots];             // SampleColorDiffuse's sampleColor
    float sr[kWfSlots], sg[kWfSlots], sb[kWfSlots], sn[kWfSlots]; // pixel sum and count
    unsigned long long key[kWfSlots];
    uint32_t event[kWfSlots], mode[kWfSlots], bounce[kWfSlots], specular[kWfSlots], sample[kWfSlots], hit[kWfSlots];
    // queues of slot indices (the reference's Tasks)
    uint16_t rayQ[kWfSlots], diffuseQ[kWfSlots], reflectiveQ[kWfSlots], refractiveQ[kWfSlots], skyboxQ[kWfSlots];
    uint32_t nRay, nDiffuse, nReflective, nRefractive, nSkybox;
    uint32_t tile, liveSlots;
};

// append `slot` to queue q when `want`: ballot + popc prefix inside the warp, one shared atomic per warp
__device__ __forceinline__ void wf_push(bool want, uint32_t slot, uint16_t *q, uint32_t *count)
{
    uint32_t const lane = threadIdx.x & 31u;
    uint32_t const votes = __ballot_sync(0xffffffffu, want);
    if (!votes)
        return;
    uint32_t base = 0;
    if (lane == 0)
        base = atomicAdd(count, uint32_t(__popc(votes)));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (want)
        q[base + __popc(votes & ((1u << lane) - 1u))] = uint16_t(slot);
}

template <int STAGE>
__global__ void __launch_bounds__(kBlockThreads, 2) trace_wavefront(const __grid_constant__ TraceParams p, uint32_t *tileCounter)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ ShellSlots shellSlots[kBlockThreads / 32];
    __shared__ WfState S;
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];

    uint32_t const geomBytes = p.n * uint32_t(sizeof(float4));
    uint32_t const pairBytes = filter_blob_bytes(p.filter.nSlots);
    const float4 *geom = p.geom;
    FilterScene fs = p.filter;
    if (STAGE == 2)
    {
        stage_scene_tma(smemRaw, p.geom, geomBytes, smemRaw + geomBytes, p.filter.pairs, pairBytes, &mbar);
        geom = reinterpret_cast<const float4 *>(smemRaw);
        rebase_filter(fs, smemRaw + geomBytes);
    }
    else if (STAGE == 1)
    {
        stage_scene_tma(smemRaw, p.filter.pairs, pairBytes, nullptr, nullptr, 0, &mbar);
        rebase_filter(fs, smemRaw);
    }
    // STAGE 3: lists beyond shared memory stream through the TMA chunk ring (spt_streamed.cuh), as in the megakernel family.  The
    // intersect stage is a CTA affair anyway (thread k traces ray-queue entry k between two barriers), and the warps that hold
    // rays are known without counting: the first ceil(nRay / 32).
    __shared__ __align__(8) uint64_t fullBar[kStages];
    __shared__ uint32_t released[kStages];
    StreamRing R{};
    if (STAGE == 3)
        stream_ring_init(R, smemRaw, fullBar, released);

    uint32_t const tid = threadIdx.x;
    uint32_t const segW = p.xEnd - p.xBegin, segH = p.yEnd - p.yBegin;
    uint32_t const nBands = (segH + kTileH - 1) / kTileH; // tile rows of the segment; this launch takes every bandCount-th
    uint32_t const tilesX = (segW + kTileW - 1) / kTileW,
                   tilesY = nBands > p.bandIndex ? (nBands - p.bandIndex + p.bandCount - 1) / p.bandCount : 0u;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);
    uint32_t segments = 0, dropped = 0, exactEvals = 0, samplesDone = 0;

    for (;;) // persistent: next tile
    {
        __syncthreads();
        if (tid == 0)
            S.tile = atomicAdd(tileCounter, 1u);
        __syncthreads();
        uint32_t const tile = S.tile;
        if (tile >= tilesX * tilesY)
            break;
        // slot <-> pixel, same 8x4-per-warp layout as the megakernel
        uint32_t const lane = tid & 31u, warp = tid >> 5;
        uint32_t const px = p.xBegin + (tile % tilesX) * kTileW + (lane & (kWarpW - 1)) + (warp % kBlockWarpsX) * kWarpW;
        uint32_t const py = p.yBegin + ((tile / tilesX) * p.bandCount + p.bandIndex) * kTileH + (lane / kWarpW) + (warp / kBlockWarpsX) * kWarpH;
        bool const inside = px < p.xEnd && py < p.yEnd;
        uint32_t const pixel = py * p.width + px;
        uint64_t const pixKey = pixel_key(p.seed, pixel);
        {
            // the running sum continues from the accumulator (see trace_megakernel)
            float4 const acc0 = inside ? p.accum[pixel] : make_float4(0.f, 0.f, 0.f, 0.f);
            S.sr[tid] = acc0.x; S.sg[tid] = acc0.y; S.sb[tid] = acc0.z; S.sn[tid] = acc0.w;
        }
        S.mode[tid] = MODE_NEW;
        S.sample[tid] = inside ? p.sampleBegin : p.sampleEnd;
        S.cr[tid] = S.cg[tid] = S.cb[tid] = 0.f;
        S.bounce[tid] = 0;
        S.specular[tid] = 0;

        for (;;) // passes over this tile
        {
            // ---- generate: slots in MODE_NEW start their next camera sample (own slot: no queue needed) ----
            if (tid == 0)
                S.nRay = S.nDiffuse = S.nReflective = S.nRefractive = S.nSkybox = 0;
            bool live = S.mode[tid] != MODE_NEW;
            if (!live && S.sample[tid] < p.sampleEnd)
            {
                Rng rng;
                rng.key = stream_key(pixKey, S.sample[tid]);
                rng.event = 0;
                rng.open_event();
                float xiU, xiV;
                rng.uniform2(-1.f, 1.f, xiU, xiV);
                V3 const d = primary_direction(p.cam, widthF, heightF, px, py, xiU, xiV);
                S.dx[tid] = d.x; S.dy[tid] = d.y; S.dz[tid] = d.z;
                S.ox[tid] = p.cam.eye[0]; S.oy[tid] = p.cam.eye[1]; S.oz[tid] = p.cam.eye[2];
                S.key[tid] = rng.key;
                S.event[tid] = rng.event;
                S.specular[tid] = 0;
                S.mode[tid] = MODE_TRACE;
                S.sample[tid] += 1;
                ++samplesDone;
                live = true;
            }
            __syncthreads();
            wf_push(live, tid, S.rayQ, &S.nRay);
            __syncthreads();
            uint32_t const nRay = S.nRay;
            if (nRay == 0)
                break; // tile finished

            // ---- intersect: thread k traces ray-queue entry k (all lanes of every warp but the last busy) ----
            {
                bool const mine = tid < nRay;
                uint32_t const s = mine ? S.rayQ[tid] : 0u;
                V3 const d = mine ? v3(S.dx[s], S.dy[s], S.dz[s]) : v3(0.f, 0.f, 1.f);
                V3 const o = mine ? v3(S.ox[s], S.oy[s], S.oz[s]) : v3(0.f, 0.f, 0.f);
                uint32_t const idx = STAGE == 3 ? streamed_closest_hit<false>(fs, geom, p.n, d, o, R, exactEvals, mine, (tid & ~31u) < nRay, (nRay + 31u) / 32u)
                                                : closest_hit_filtered(fs, geom, p.n, d, o, exactEvals, mine);
                segments += mine ? 1u : 0u;
                // ---- classify into the four material queues (TraceAndSampleColor(task, tasks)) ----
                uint32_t m = 0;
                bool const inChain = mine && S.mode[s] == MODE_DIFFUSE;
                if (mine)
                {
                    S.hit[s] = idx;
                    if (idx < p.n && !inChain)
                        m = uint32_t(__ldg(p.mat + idx));
                }
                wf_push(mine && (inChain || m == 3u), s, S.diffuseQ, &S.nDiffuse);
                wf_push(mine && !inChain && m == 1u, s, S.reflectiveQ, &S.nReflective);
                wf_push(mine && !inChain && m == 2u, s, S.refractiveQ, &S.nRefractive);
                wf_push(mine && !inChain && m == 0u, s, S.skyboxQ, &S.nSkybox);
            }
            __syncthreads();

            // ---- skybox queue (TaskBasedPathTracer.hpp:182-189) ----
            if (tid < S.nSkybox)
            {
                uint32_t const s = S.skyboxQ[tid];
                V3 const c = sample_sky(p.cam, v3(S.dx[s], S.dy[s], S.dz[s]));
                S.sr[s] = fadd(S.sr[s], c.x); S.sg[s] = fadd(S.sg[s], c.y); S.sb[s] = fadd(S.sb[s], c.z);
                S.sn[s] = fadd(S.sn[s], 1.f);
                S.mode[s] = MODE_NEW;
            }

            // ---- diffuse queue: SampleColorDiffuse's head and every step of its chain (:83-106) ----
            {
                bool const mine = tid < S.nDiffuse;
                uint32_t const s = mine ? S.diffuseQ[tid] : 0u;
                uint32_t const idx = mine ? S.hit[s] : p.n;
                bool const inChain = mine && S.mode[s] == MODE_DIFFUSE;
                bool goes = mine && !inChain; // head always scatters
                if (inChain)
                {
                    uint32_t const b = S.bounce[s] - 1u; // while (--bounceCount && hit)
                    S.bounce[s] = b;
                    goes = b != 0u && idx < p.n;
                    if (!goes)
                    {
                        S.sr[s] = fadd(S.sr[s], S.cr[s]); S.sg[s] = fadd(S.sg[s], S.cg[s]); S.sb[s] = fadd(S.sb[s], S.cb[s]);
                        S.sn[s] = fadd(S.sn[s], 1.f);
                        S.mode[s] = MODE_NEW;
                    }
                }
                float4 sph = make_float4(0.f, 0.f, 0.f, 1.f);
                V3 pnt = v3(0.f, 0.f, 0.f), nrm = v3(0.f, 0.f, 0.f);
                Rng rng;
                rng.state = 0;
                if (goes)
                {
                    sph = geom[idx];
                    pnt = contact_point(sph, v3(S.ox[s], S.oy[s], S.oz[s]), v3(S.dx[s], S.dy[s], S.dz[s]), false);
                    nrm = contact_normal(pnt, sph);
                    rng.key = S.key[s];
                    rng.event = S.event[s];
                    rng.open_event();
                    S.event[s] = rng.event;
                }
                V3 const sh = coop_shell(goes, rng.state, slots);
                if (goes)
                {
                    V3 v;
                    if (inChain)
                    {
                        S.cr[s] = fmul(S.cr[s], 0.5f); S.cg[s] = fmul(S.cg[s], 0.5f); S.cb[s] = fmul(S.cb[s], 0.5f);
                        v = vadd(vadd(pnt, nrm), sh);
                    }
                    else
                    {
                        float4 const col = __ldg(p.shade + idx);
                        S.cr[s] = fmul(col.x, 0.5f); S.cg[s] = fmul(col.y, 0.5f); S.cb[s] = fmul(col.z, 0.5f);
                        v = vadd(nrm, sh);
                        S.bounce[s] = p.depth;
                        S.mode[s] = MODE_DIFFUSE;
                    }
                    V3 const d = vnormalize(v);
                    S.dx[s] = d.x; S.dy[s] = d.y; S.dz[s] = d.z;
                    S.ox[s] = pnt.x; S.oy[s] = pnt.y; S.oz[s] = pnt.z;
                }
            }

            // ---- reflective queue (:108-122) ----
            {
                bool const mine = tid < S.nReflective;
                uint32_t const s = mine ? S.reflectiveQ[tid] : 0u;
                float4 sph = make_float4(0.f, 0.f, 0.f, 1.f);
                V3 pnt = v3(0.f, 0.f, 0.f), nrm = v3(0.f, 0.f, 0.f), din = v3(0.f, 0.f, 1.f);
                Rng rng;
                rng.state = 0;
                uint32_t idx = 0;
                if (mine)
                {
                    idx = S.hit[s];
                    sph = geom[idx];
                    din = v3(S.dx[s], S.dy[s], S.dz[s]);
                    pnt = contact_point(sph, v3(S.ox[s], S.oy[s], S.oz[s]), din, false);
                    nrm = contact_normal(pnt, sph);
                    rng.key = S.key[s];
                    rng.event = S.event[s];
                    rng.open_event();
                    S.event[s] = rng.event;
                }
                V3 const sh = coop_shell(mine, rng.state, slots);
                if (mine)
                {
                    V3 const d = vnormalize(vadd(vreflect(din, nrm), vscale(sh, __ldg(p.shade + idx).w)));
                    S.dx[s] = d.x; S.dy[s] = d.y; S.dz[s] = d.z;
                    S.ox[s] = pnt.x; S.oy[s] = pnt.y; S.oz[s] = pnt.z;
                }
            }

            // ---- refractive queue (:124-180) ----
            if (tid < S.nRefractive)
            {
                uint32_t const s = S.refractiveQ[tid];
                uint32_t const idx = S.hit[s];
                float4 const sph = geom[idx];
                V3 dir = v3(S.dx[s], S.dy[s], S.dz[s]);
                V3 org = contact_point(sph, v3(S.ox[s], S.oy[s], S.oz[s]), dir, false);
                V3 const nrm = contact_normal(org, sph);
                Rng rng;
                rng.key = S.key[s];
                rng.event = S.event[s];
                rng.open_event();
                S.event[s] = rng.event;
                refract_through(sph, nrm, dir, org, rng);
                S.dx[s] = dir.x; S.dy[s] = dir.y; S.dz[s] = dir.z;
                S.ox[s] = org.x; S.oy[s] = org.y; S.oz[s] = org.z;
            }
            __syncthreads();

            // ---- specular bookkeeping: the pass loop's cap (TaskBasedPathTracer.hpp:81) ----
            {
                uint32_t const nSpec = S.nReflective + S.nRefractive;
                if (tid < nSpec)
                {
                    uint32_t const s = tid < S.nReflective ? S.reflectiveQ[tid] : S.refractiveQ[tid - S.nReflective];
                    uint32_t const sp = S.specular[s] + 1u;
                    S.specular[s] = sp;
                    if (p.passCap && sp >= p.passCap)
                    {
                        ++dropped;
                        S.mode[s] = MODE_NEW;
                    }
                    else if (sp >= kSpecularSafetyCap)
                    {
                        S.sn[s] = fadd(S.sn[s], 1.f);
                        S.mode[s] = MODE_NEW;
                    }
                }
            }
            __syncthreads();
        }

        if (inside)
        {
            p.accum[pixel] = make_float4(S.sr[tid], S.sg[tid], S.sb[tid], S.sn[tid]);
        }
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, segments);
        warp_counter_add(p.counters + 1, samplesDone);
        warp_counter_add(p.counters + 2, dropped);
        warp_counter_add(p.counters + 3, exactEvals);
    }
}

} // namespace spt
