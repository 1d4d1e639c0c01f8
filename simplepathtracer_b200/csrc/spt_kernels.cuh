// spt_kernels.cuh -- kernels of the path-tracing hot path (sm_100a only).
//
//   trace_megakernel   camera rays + closest-hit + scatter + bounce loop + per-pixel accumulation
//                      (RenderSegment / RenderSegmentTask semantics, SingleThreadPathTracer.hpp:94-137,
//                      TaskBasedPathTracer.hpp:54-206) as ONE kernel; one lane = one pixel, the sample
//                      loop and the bounce loop are flattened into a single state machine so that
//                      every iteration of the warp is one closest-hit over the sphere list.
//   first_hit_kernel   deterministic pinhole rays -> index / t / normal (parity probe)
//   closest_hit_kernel caller-supplied rays -> index (parity probe)
//   resolve_kernel     (sum, count) -> normalise, gamma 2, round, flip, pack RGB8 (IOHelpers.hpp:17-22)
#pragma once

#include "spt_device.cuh"
#include "spt_filter.cuh"

namespace spt
{

// Block = 8 warps; a warp covers an 8x4 pixel tile (coherent rays), the block a 32x8 tile.
constexpr int kWarpW = 8, kWarpH = 4;
#ifndef SPT_BLOCK_WARPS_X
#define SPT_BLOCK_WARPS_X 4
#endif
#ifndef SPT_BLOCK_WARPS_Y
#define SPT_BLOCK_WARPS_Y 2
#endif
constexpr int kBlockWarpsX = SPT_BLOCK_WARPS_X, kBlockWarpsY = SPT_BLOCK_WARPS_Y;
constexpr int kTileW = kWarpW * kBlockWarpsX; // 32
constexpr int kTileH = kWarpH * kBlockWarpsY; // 8
constexpr int kBlockThreads = kTileW * kTileH; // 256
// CTA shape in warps (each warp an 8x4 pixel tile).  The resident split kernels (spt_split.cuh) take it as a template
// parameter; every other kernel uses the default above.
template <int WX, int WY>
struct BlockShape
{
    static constexpr int kWarpsX = WX, kWarpsY = WY;
    static constexpr int kTileW = kWarpW * WX, kTileH = kWarpH * WY;
    static constexpr int kThreads = kTileW * kTileH;
};
using DefaultShape = BlockShape<kBlockWarpsX, kBlockWarpsY>;
#ifndef SPT_MIN_BLOCKS
#define SPT_MIN_BLOCKS 4 // resident blocks per SM the register allocation aims at
#endif

struct TraceParams
{
    CameraConst cam;
    uint32_t width, height;
    uint32_t yBegin, yEnd, xBegin, xEnd;
    uint32_t sampleBegin, sampleEnd;
    uint32_t depth, passCap;
    uint32_t bandCount, bandIndex; // row r of the grid's tiles is band (r * bandCount + bandIndex) of the segment (SptParams)
    uint64_t seed;
    uint32_t n;
    const float4 *geom;  // (cx, cy, cz, r)
    const float4 *shade; // (colour r, g, b, fuzz)
    const uint8_t *mat;
    FilterScene filter;  // pre-filter records (global memory copies)
    float4 *accum;       // width*height (sum r, sum g, sum b, count), row y*width + x (not flipped)
    unsigned long long *counters; // [0] segments [1] samples [2] dropped [3] exact evaluations (may be null)
    // split rendering (spt_split.cuh): the camera segment of every sample of the launch as one float4 record
    // (direction x, y, z, first-hit sphere index as raw bits; kMissIndex = miss),
    // [(y - yBegin) * segW + (x - xBegin)][sample - sampleBegin]
    float4 *firstHits;
    FilterScene filterCam; // pre-filter records specialised for the eye (camera form)
    uint32_t tileRowOffset; // split rendering launches a strip of tile rows at a time: blockIdx.y counts from this row of the call
};

constexpr uint32_t kMissIndex = 0xFFFFFFFFu;

constexpr uint32_t kSpecularSafetyCap = 65536u; // the reference would overflow its stack long before

// ---- TMA: one bulk copy of the whole sphere list into shared memory ---------------------------
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

// dstA/srcA/bytesA and dstB/srcB/bytesB: up to two regions (bytes multiple of 16; 0 = skip)
__device__ __forceinline__ void stage_scene_tma(void *dstA, const void *srcA, uint32_t bytesA, void *dstB, const void *srcB,
                                                uint32_t bytesB, uint64_t *mbar)
{
    uint32_t const bytes = bytesA + bytesB;
    uint32_t const bar = smem_addr(mbar);
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        if (bytesA)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smem_addr(dstA)),
                         "l"(srcA), "r"(bytesA), "r"(bar)
                         : "memory");
        if (bytesB)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smem_addr(dstB)),
                         "l"(srcB), "r"(bytesB), "r"(bar)
                         : "memory");
    }
    // every thread waits for phase 0 of the barrier (the bytes have landed)
    uint32_t done = 0;
    while (!done)
    {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar)
                     : "memory");
    }
}

// ---- shading (a6-a9 of SURVEY 8a) ---------------------------------------------------------------
// SampleColorSkybox (SingleThreadPathTracer.hpp:11-19): (initColor * (d.y + 1)) * 0.5
__device__ __forceinline__ V3 sample_sky(const CameraConst &cam, V3 dir)
{
    float const k = fadd(dir.y, 1.f);
    return vscale(vscale(v3(cam.sky[0], cam.sky[1], cam.sky[2]), k), 0.5f);
}

// SampleColorRefractive after its contact point / normal (SingleThreadPathTracer.hpp:54-91;
// TaskBasedPathTracer.hpp:135-174): Schlick, refract in, jump to the far side of the SAME sphere (quirk Q9),
// Schlick again, refract out or reflect.  `org` is the near contact point on entry.
__device__ __forceinline__ void refract_through(float4 sphere, V3 n, V3 &dir, V3 &org, Rng &rng)
{
    float const nAir = 1.0f, nGlass = 1.5f;
    float c = vdot(vneg(n), dir);
    float r = fdiv(nAir, nGlass);
    float const q1 = fdiv(fsub(nAir, nGlass), fadd(nAir, nGlass));
    float rSq = fmul(q1, q1); // pow(q, 2.f)
    float schlick = fadd(rSq, fmul(fsub(1.f, rSq), pow5(fsub(1.f, c))));

    if (rng.uniform1(0.f, 1.f) < schlick)
    {
        dir = vreflect(dir, n);
    }
    else if (fmul(r, fsqrt(fsub(1.f, fmul(c, c)))) < 1.f)
    {
        float const k = fsub(fmul(r, c), fsqrt(fsub(1.f, fmul(fmul(r, r), fsub(1.f, fmul(c, c))))));
        dir = vnormalize(vadd(vscale(dir, r), vscale(n, k)));
        org = contact_point(sphere, org, dir, true);
        n = vneg(contact_normal(org, sphere));

        c = vdot(vneg(n), dir);
        r = fdiv(nGlass, nAir);
        float const q2 = fdiv(fsub(nGlass, nAir), fadd(nGlass, nAir));
        rSq = fmul(q2, q2);
        schlick = fadd(rSq, fmul(fsub(1.f, rSq), pow5(fsub(1.f, c))));

        if (rng.uniform1(0.f, 1.f) < schlick)
        {
            dir = vreflect(dir, n);
        }
        else if (fmul(r, fsqrt(fsub(1.f, fmul(c, c)))) < 1.f)
        {
            float const k2 = fsub(fmul(r, c), fsqrt(fsub(1.f, fmul(fmul(r, r), fsub(1.f, fmul(c, c))))));
            dir = vnormalize(vadd(vscale(dir, r), vscale(n, k2)));
        }
        else
        {
            dir = vreflect(dir, n);
        }
    }
    else
    {
        dir = vreflect(dir, n);
    }
}

enum : uint32_t
{
    MODE_NEW = 0,    // lane needs its next camera sample
    MODE_TRACE = 1,  // primary / specular segment: material dispatch follows (TraceAndSampleColor)
    MODE_DIFFUSE = 2 // inside SampleColorDiffuse's terminal chain
};

// Lane state of the flattened (sample x bounce) loop
struct Lane
{
    V3 org, dir;
    V3 chain;   // SampleColorDiffuse's sampleColor
    V3 sum;     // pixelColor (SingleThreadPathTracer.hpp:121) / colors[ci] (TaskBasedPathTracer.hpp:58)
    float count; // samples[ci] (TaskBasedPathTracer.hpp:59)
    Rng rng;
    uint32_t mode, bounce, specular, sample;
};

__device__ __forceinline__ void sts128(float4 *ptr, float4 v)
{
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_addr(ptr)), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds128(const float4 *ptr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(smem_addr(ptr)) : "memory");
    return v;
}

// Where the per-pixel running sum (sum r, g, b, sample count) lives while a kernel runs.  Either way the additions are
// the same float operations in the same order.
struct AccRegs // in the lane's registers (L.sum, L.count)
{
    Lane &L;
    __device__ __forceinline__ void add(V3 v) const
    {
        L.sum = vadd(L.sum, v);
        L.count = fadd(L.count, 1.f);
    }
    __device__ __forceinline__ void count_only() const { L.count = fadd(L.count, 1.f); }
    __device__ __forceinline__ void drop() const {} // a dropped path leaves no trace in the pixel (TaskBasedPathTracer.hpp:196-198)
};
struct AccShared // in a per-lane float4 slot of shared memory: touched once per finished sample, never held in registers
{
    float4 *slot;
    __device__ __forceinline__ void add(V3 v) const
    {
        float4 a = lds128(slot);
        a.x = fadd(a.x, v.x);
        a.y = fadd(a.y, v.y);
        a.z = fadd(a.z, v.z);
        a.w = fadd(a.w, 1.f);
        sts128(slot, a);
    }
    __device__ __forceinline__ void count_only() const
    {
        float4 a = lds128(slot);
        a.w = fadd(a.w, 1.f);
        sts128(slot, a);
    }
    __device__ __forceinline__ void drop() const {}
};

// Warp-cooperative rejection sampling of the shell vector (Random.hpp:115-141).
//
// Each lane that needs a vector owns a counter-addressable stream (state0): attempt t is a pure function of
// (state0, t) (shell_attempt).  Run lane-by-lane the do/while costs the warp max-over-lanes trips (about 6 for
// an acceptance of 0.476) with most lanes idle.  Here the `m` lanes still pending are given per = 2^k <= 32/m
// helper lanes each; in one round helper a of pending lane r evaluates attempt base_r + a, a ballot collects
// the verdicts, and every pending lane takes the FIRST accepted attempt (lowest t) -- exactly the vector the
// sequential loop would have produced -- or moves its base on by `per`.  Two rounds almost always finish.
// Must be called by all 32 lanes of the warp (need = false for lanes that want nothing).
struct ShellSlots
{
    unsigned long long st[32];
    uint32_t base[32];
    uint32_t trips; // bounce kernel: closest-hit trips of this warp (SptCounters.warp_iterations), kept HERE, not in a register
};

#ifndef SPT_SHELL_LOCAL
#define SPT_SHELL_LOCAL 2 // attempts every lane makes on its OWN stream before the cooperative rounds start (0: none)
#endif
__device__ __forceinline__ V3 coop_shell(bool need, unsigned long long state0, ShellSlots &slot)
{
    uint32_t const lane = threadIdx.x & 31u;
    V3 sh = v3(0.f, 0.f, 0.f);
    uint32_t base = 0;
#if SPT_SHELL_LOCAL > 0
    // Round 0 needs no coordination: with most of the warp pending there is at most one helper per lane anyway, so every lane
    // evaluates the first SPT_SHELL_LOCAL attempts of its own stream (independent chains: instruction-level parallelism
    // instead of ballots, shared-memory slots and shuffles) and keeps the FIRST accepted one.  72 % of the lanes are done
    // after two attempts; the cooperative rounds below then serve the few that are left with 4 or more helpers each.
    {
        V3 cand[SPT_SHELL_LOCAL];
        bool ok[SPT_SHELL_LOCAL];
#pragma unroll
        for (int t = 0; t < SPT_SHELL_LOCAL; ++t)
            ok[t] = shell_attempt(state0, uint32_t(t), cand[t]);
        bool got = false;
#pragma unroll
        for (int t = 0; t < SPT_SHELL_LOCAL; ++t)
            if (!got && ok[t])
            {
                sh = cand[t];
                got = true;
            }
        base = SPT_SHELL_LOCAL;
        need = need && !got;
    }
#endif
    uint32_t pending = __ballot_sync(0xffffffffu, need);
    while (pending) // warp-uniform
    {
        uint32_t const m = __popc(pending);
        uint32_t const lg = m <= 1u ? 5u : uint32_t(__clz(int(m - 1u))) - 27u; // per = largest 2^k with m * 2^k <= 32
        uint32_t const per = 1u << lg;
        uint32_t const r = __popc(pending & ((1u << lane) - 1u)); // my rank among the pending lanes
        if (need)
        {
            slot.st[r] = state0;
            slot.base[r] = base;
        }
        __syncwarp();
        uint32_t const g = lane >> lg, a = lane & (per - 1u);
        bool const valid = g < m;
        unsigned long long const s0 = slot.st[valid ? g : 0u];
        uint32_t const b0 = slot.base[valid ? g : 0u];
        __syncwarp();
        V3 cand;
        bool const ok = shell_attempt(s0, b0 + a, cand) && valid;
        uint32_t const verdict = __ballot_sync(0xffffffffu, ok);
        uint32_t src = lane;
        bool got = false;
        if (need)
        {
            uint32_t const field = per == 32u ? verdict : ((verdict >> (r << lg)) & ((1u << per) - 1u));
            if (field)
            {
                src = (r << lg) + uint32_t(__ffs(int(field)) - 1);
                got = true;
            }
            else
                base += per;
        }
        float const sx = __shfl_sync(0xffffffffu, cand.x, src);
        float const sy = __shfl_sync(0xffffffffu, cand.y, src);
        float const sz = __shfl_sync(0xffffffffu, cand.z, src);
        if (got)
        {
            sh = v3(sx, sy, sz);
            need = false;
        }
        pending = __ballot_sync(0xffffffffu, need);
    }
    return sh;
}

// After the closest-hit of the current segment returned `idx`: shade, and either finish the sample or set up
// the next segment.  `geom` may live in shared or global memory.
//
// The reference has four handlers (sky / diffuse / reflective / refractive) plus the diffuse chain; lanes of a
// warp are in different ones at the same time.  To keep the warp converged the work they share is hoisted
// into single sites that every lane needing it executes together: ONE contact point + normal
// (Collision.hpp:49-71), ONE rejection-sampled shell vector (Random.hpp:115-127; the loop whose trip count
// varies most) and ONE final Normalize.  Only the few instructions that differ sit in per-material branches.
// The arithmetic per lane is exactly the reference's.
// What shade_step's first half decides about a lane: which handler (if any) shades a surface next.
struct SurfaceKind
{
    bool chainGoes, isDiffuse, isMirror, isGlass;
    __device__ __forceinline__ bool any() const { return chainGoes || isDiffuse || isMirror || isGlass; }
};
__device__ __forceinline__ SurfaceKind surface_none() { return SurfaceKind{false, false, false, false}; }
// the handler TraceAndSampleColor's switch picks for material m (SingleThreadPathTracer.hpp:98-108)
__device__ __forceinline__ SurfaceKind surface_of_material(uint32_t m) { return SurfaceKind{false, m == 3u, m == 1u, m == 2u}; }

// First half of shade_step: material dispatch / chain control, and the TERMINAL cases (end of SampleColorDiffuse's chain,
// SampleColorSkybox on a miss or a Material::SKYBOX sphere), which add the sample to the pixel and leave the lane in MODE_NEW.
template <typename Acc>
__device__ __forceinline__ SurfaceKind shade_classify(const TraceParams &p, Lane &L, uint32_t idx, bool live, Acc const acc)
{
    bool const hit = idx < p.n;
    bool const inChain = L.mode == MODE_DIFFUSE;
    uint32_t const m = (live && hit && !inChain) ? uint32_t(__ldg(p.mat + idx)) : 0u;
    // while (--bounceCount && sphereIndex < g_sphereNumber) (SingleThreadPathTracer.hpp:28): the decrement
    // happens whether or not the sphere test is reached
    SurfaceKind k = surface_none();
    if (live && inChain)
    {
        --L.bounce;
        k.chainGoes = L.bounce != 0u && hit;
    }
    k.isDiffuse = live && !inChain && m == 3u;
    k.isMirror = live && !inChain && m == 1u;
    k.isGlass = live && !inChain && m == 2u;
    if (live && !k.any())
    {
        // terminal: end of SampleColorDiffuse's chain, or SampleColorSkybox (miss / Material::SKYBOX sphere)
        acc.add(inChain ? L.chain : sample_sky(p.cam, L.dir));
        L.mode = MODE_NEW;
    }
    return k;
}

// Second half of shade_step: the surface handlers of the lanes `k` names, set up the next segment.
template <typename GeomPtr, typename Acc>
__device__ __forceinline__ void shade_surface(const TraceParams &p, GeomPtr geom, Lane &L, uint32_t idx, SurfaceKind const k,
                                              uint32_t &dropped, ShellSlots &slots, Acc const acc)
{
    bool const chainGoes = k.chainGoes, isDiffuse = k.isDiffuse, isMirror = k.isMirror, isGlass = k.isGlass;
    bool const surface = k.any();

    // ---- shared site 1: contact point and normal (every handler starts with these two calls) ----
    float4 s = make_float4(0.f, 0.f, 0.f, 1.f);
    V3 pnt = v3(0.f, 0.f, 0.f), nrm = v3(0.f, 0.f, 0.f);
    if (surface)
    {
        s = geom[idx];
        pnt = contact_point(s, L.org, L.dir, false);
        nrm = contact_normal(pnt, s);
        L.rng.open_event();
    }

    if (isGlass)
    {
        L.org = pnt;
        refract_through(s, nrm, L.dir, L.org, L.rng);
    }

    // ---- shared site 2: the shell vector (drawn even by a mirror with fuzz 0), sampled by the whole warp ----
    bool const needShell = surface && !isGlass;
    V3 const sh = coop_shell(needShell, L.rng.state, slots);

    if (needShell)
    {
        V3 v;
        if (chainGoes)
        {
            // sampleColor * 0.5; Normalize(origin + normal + shell): the origin really is added (quirk Q6)
            L.chain = vscale(L.chain, 0.5f);
            v = vadd(vadd(pnt, nrm), sh);
        }
        else if (isDiffuse)
        {
            // SampleColorDiffuse's head (SingleThreadPathTracer.hpp:23-26)
            float4 const col = __ldg(p.shade + idx);
            L.chain = vscale(v3(col.x, col.y, col.z), 0.5f);
            v = vadd(nrm, sh);
            L.bounce = p.depth;
            L.mode = MODE_DIFFUSE;
        }
        else
        {
            // SampleColorReflective (SingleThreadPathTracer.hpp:39-46): Reflect + shell * g_diffuses[idx]
            v = vadd(vreflect(L.dir, nrm), vscale(sh, __ldg(p.shade + idx).w));
        }
        // ---- shared site 3: the final Normalize ----
        L.dir = vnormalize(v);
        L.org = pnt;
    }

    if (isMirror || isGlass)
    {
        ++L.specular;
        if (p.passCap && L.specular >= p.passCap)
        {
            ++dropped; // the task tracer's pass loop ends (TaskBasedPathTracer.hpp:81): dropped, not counted
            acc.drop();
            L.mode = MODE_NEW;
        }
        else if (L.specular >= kSpecularSafetyCap)
        {
            acc.count_only();
            L.mode = MODE_NEW;
        }
    }
}

template <typename GeomPtr, typename Acc>
__device__ __forceinline__ void shade_step(const TraceParams &p, GeomPtr geom, Lane &L, uint32_t idx, uint32_t &dropped,
                                           bool live, ShellSlots &slots, Acc const acc)
{
    SurfaceKind const k = shade_classify(p, L, idx, live, acc);
    shade_surface(p, geom, L, idx, k, dropped, slots, acc);
}

template <typename GeomPtr>
__device__ __forceinline__ void shade_step(const TraceParams &p, GeomPtr geom, Lane &L, uint32_t idx, uint32_t &dropped,
                                           bool live, ShellSlots &slots)
{
    shade_step(p, geom, L, idx, dropped, live, slots, AccRegs{L});
}

__device__ __forceinline__ void warp_counter_add(unsigned long long *ctr, unsigned long long v)
{
    for (int o = 16; o > 0; o >>= 1)
        v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v)
        atomicAdd(ctr, v);
}

// pixel of this lane inside the segment, in tile row `tileRow` of the launch: warps tile 8x4, blocks 32x8
template <typename S = DefaultShape>
__device__ __forceinline__ void lane_pixel_at(const TraceParams &p, uint32_t tileRow, uint32_t &x, uint32_t &y, bool &inside)
{
    uint32_t const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t const lx = (lane & (kWarpW - 1)) + (warp % S::kWarpsX) * kWarpW;
    uint32_t const ly = (lane / kWarpW) + (warp / S::kWarpsX) * kWarpH;
    x = p.xBegin + blockIdx.x * S::kTileW + lx;
    y = p.yBegin + (tileRow * p.bandCount + p.bandIndex) * S::kTileH + ly;
    inside = x < p.xEnd && y < p.yEnd;
}
template <typename S = DefaultShape>
__device__ __forceinline__ void lane_pixel(const TraceParams &p, uint32_t &x, uint32_t &y, bool &inside)
{
    lane_pixel_at<S>(p, blockIdx.y + p.tileRowOffset, x, y, inside);
}
// row of this lane among the rows the LAUNCH covers (== y - yBegin unless the call renders interleaved bands)
template <typename S = DefaultShape>
__device__ __forceinline__ uint32_t lane_row_local_at(uint32_t tileRow)
{
    uint32_t const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    return tileRow * S::kTileH + (lane / kWarpW) + (warp / S::kWarpsX) * kWarpH;
}
template <typename S = DefaultShape>
__device__ __forceinline__ uint32_t lane_row_local()
{
    return lane_row_local_at<S>(blockIdx.y);
}
static_assert(kTileH == 8 && BlockShape<2, 2>::kTileH == 8, "SPT_BAND_ROWS is one tile row of every tracing kernel");

// STAGE: what is staged into shared memory by TMA bulk copies at kernel start (every lane of a warp then
// reads the same address: a broadcast).  2 = exact geometry + pre-filter records, 1 = pre-filter records only
// (strict: geometry only), 0 = nothing (lists larger than shared memory are read through L1/L2 with
// warp-uniform addresses).  STRICT: no pre-filter, every sphere through the exact sequence.
template <int STAGE, bool STRICT>
__global__ void __launch_bounds__(kBlockThreads, SPT_MIN_BLOCKS) trace_megakernel(const __grid_constant__ TraceParams p)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ __align__(8) uint64_t mbar;
    uint32_t const geomBytes = p.n * uint32_t(sizeof(float4));
    uint32_t const pairBytes = STRICT ? 0u : filter_blob_bytes(p.filter.nSlots); // records + slot->index map
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
        if (STRICT)
        {
            stage_scene_tma(smemRaw, p.geom, geomBytes, nullptr, nullptr, 0, &mbar);
            geom = reinterpret_cast<const float4 *>(smemRaw);
        }
        else
        {
            stage_scene_tma(smemRaw, p.filter.pairs, pairBytes, nullptr, nullptr, 0, &mbar);
            rebase_filter(fs, smemRaw);
        }
    }

    uint32_t x, y;
    bool inside;
    lane_pixel(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);

    __shared__ ShellSlots shellSlots[kBlockThreads / 32];
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];

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

    // All 32 lanes stay in the loop until the whole warp is finished (the shell sampler is a warp
    // collective); a lane whose pixel is complete is carried along with live = false.
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
        if (!__any_sync(0xffffffffu, live))
            break;
        uint32_t idx;
        if (STRICT)
            idx = live ? closest_hit_strict(geom, p.n, L.dir, L.org) : p.n;
        else
            idx = closest_hit_filtered(fs, geom, p.n, L.dir, L.org, exactEvals, live);
        segments += live ? 1u : 0u;
        shade_step(p, geom, L, idx, dropped, live, slots);
    }
    if (inside)
    {
        // one lane owns one pixel for the whole launch: plain load at the start, plain store here, no atomics
        float4 const a = make_float4(L.sum.x, L.sum.y, L.sum.z, L.count);
        p.accum[pixel] = a;
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, segments);
        warp_counter_add(p.counters + 1, inside ? (p.sampleEnd - p.sampleBegin) : 0u);
        warp_counter_add(p.counters + 2, dropped);
        warp_counter_add(p.counters + 3, STRICT ? segments * (unsigned long long)(p.n) : exactEvals);
    }
}

// ---- parity probes ------------------------------------------------------------------------------
struct FirstHitParams
{
    CameraConst cam;
    uint32_t width, height;
    uint32_t yBegin, yEnd, xBegin, xEnd;
    uint32_t n;
    const float4 *geom;
    FilterScene filter;
    uint32_t strict;
    uint32_t *idx;
    float *t;
    float *normal3;
};

// jitter-free pinhole rays exactly as SingleThreadPathTracer.hpp:125-128 with xi = 0
__global__ void first_hit_kernel(const __grid_constant__ FirstHitParams p)
{
    uint32_t const segW = p.xEnd - p.xBegin, segH = p.yEnd - p.yBegin;
    uint32_t const k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= segW * segH)
        return;
    uint32_t const x = p.xBegin + k % segW, y = p.yBegin + k / segW;
    V3 const o = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
    V3 const d = primary_direction(p.cam, __uint2float_rn(p.width), __uint2float_rn(p.height), x, y, 0.f, 0.f);
    uint32_t exactEvals = 0;
    uint32_t const i = p.strict ? closest_hit_strict(p.geom, p.n, d, o) : closest_hit_filtered(p.filter, p.geom, p.n, d, o, exactEvals);
    p.idx[k] = i;
    float t = 0.f;
    V3 nrm = v3(0.f, 0.f, 0.f);
    if (i < p.n)
    {
        float4 const s = p.geom[i];
        t = ray_sphere_t(vsub(v3(s.x, s.y, s.z), o), s.w, d, false);
        nrm = contact_normal(contact_point(s, o, d, false), s);
    }
    p.t[k] = t;
    p.normal3[3 * k + 0] = nrm.x;
    p.normal3[3 * k + 1] = nrm.y;
    p.normal3[3 * k + 2] = nrm.z;
}

// mode 0: pre-filter + exact candidates (the product path); 1: strict loop; 2: both, and idx gets 0xFFFFFFFF
// where they disagree (must never happen).  evals (optional) receives the number of exact evaluations.
__global__ void closest_hit_kernel(const float4 *geom, uint32_t n, FilterScene filter, uint32_t mode, uint32_t count,
                                   const float *dir3, const float *org3, uint32_t *idx, uint32_t *evals)
{
    uint32_t const k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count)
        return;
    V3 const d = v3(dir3[3 * k], dir3[3 * k + 1], dir3[3 * k + 2]);
    V3 const o = v3(org3[3 * k], org3[3 * k + 1], org3[3 * k + 2]);
    uint32_t exactEvals = 0;
    uint32_t r;
    if (mode == 1)
        r = closest_hit_strict(geom, n, d, o);
    else
    {
        r = closest_hit_filtered(filter, geom, n, d, o, exactEvals);
        if (mode == 2 && r != closest_hit_strict(geom, n, d, o))
            r = 0xFFFFFFFFu;
    }
    idx[k] = r;
    if (evals)
        evals[k] = exactEvals;
}

// ---- self-check of the shared-reciprocal division (spt_device.cuh) ------------------------------------------------
// Every thread draws operand pairs from a counter RNG -- random mantissas; exponents mostly inside the fast path's domain,
// a share of them at and beyond its edges, plus zeros / denormals / infinities -- and compares, bit for bit,
//   fdiv_shared(a, b, rcp_refined(b))  with  __fdiv_rn(a, b)            (the camera ray's divisions by the image size)
//   vnormalize(v)                      with  the three __fdiv_rn by sqrt_rn(|v|^2)  (Math.hpp:140-154)
// out[0] += operand sets tested, out[1] += mismatches (must stay 0).
__device__ __forceinline__ float selfcheck_operand(uint64_t z, uint32_t kind)
{
    uint32_t const mant = uint32_t(z) & 0x7FFFFFu, sign = uint32_t(z >> 63) << 31;
    uint32_t e;
    uint32_t const r = uint32_t(z >> 32) & 0xFFFFu;
    if (kind < 12u)
        e = 127u - 60u + r % 121u; // 2^-60 .. 2^60: the fast path's domain
    else if (kind < 14u)
        e = 127u - 64u + r % 9u + (r & 0x100u ? 120u : 0u); // around either edge of the domain
    else if (kind == 14u)
        e = r % 255u; // anything finite, denormals and zero included
    else
        return (r & 1u) ? 0.f : __uint_as_float(sign | 0x7F800000u | ((r & 2u) ? mant : 0u)); // zero, infinity, NaN
    return __uint_as_float(sign | (e << 23) | ((kind == 14u && e == 0u && (r & 0x200u)) ? 0u : mant));
}
__device__ __forceinline__ bool same_bits(float x, float y)
{
    return __float_as_uint(x) == __float_as_uint(y) || (x != x && y != y);
}
__global__ void __launch_bounds__(256) division_selfcheck_kernel(uint64_t seed, uint32_t perThread, unsigned long long *out)
{
    uint64_t const tid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    uint64_t state = mix64(seed + SPT_GOLDEN * (tid + 1ull));
    unsigned long long bad = 0;
    for (uint32_t i = 0; i < perThread; ++i)
    {
        uint64_t const z0 = mix64(state += SPT_GOLDEN), z1 = mix64(state += SPT_GOLDEN), z2 = mix64(state += SPT_GOLDEN);
        uint32_t const kind = uint32_t(z0 >> 48) & 15u;
        float const a = selfcheck_operand(z0, kind), b = selfcheck_operand(z1, (kind + 5u) & 15u);
        // a divisor the callers would hand over: a positive, safe one (image sizes; a vector's length is checked by vnormalize itself)
        float const bs = fabsf(b);
        if (div_operand_safe(bs))
            bad += same_bits(fdiv_shared(a, bs, rcp_refined(bs)), __fdiv_rn(a, bs)) ? 0u : 1u;
        V3 const v = v3(a, b, selfcheck_operand(z2, (kind + 11u) & 15u));
        float const l = fsqrt(vlensq(v));
        V3 const n = vnormalize(v);
        bad += (same_bits(n.x, __fdiv_rn(v.x, l)) && same_bits(n.y, __fdiv_rn(v.y, l)) && same_bits(n.z, __fdiv_rn(v.z, l))) ? 0u : 1u;
    }
    for (int o = 16; o > 0; o >>= 1)
        bad += __shfl_down_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0)
    {
        atomicAdd(out + 0, 32ull * perThread);
        if (bad)
            atomicAdd(out + 1, bad);
    }
}

// ---- resolve ------------------------------------------------------------------------------------
// colors[i] *= 1.f / samples[i]; io::WritePixel at g_size - ((g_width - x)*3 + y*g_width*3)
// (TaskBasedPathTracer.hpp:196-205, SingleThreadPathTracer.hpp:120,133-134, IOHelpers.hpp:17-22)
__global__ void resolve_kernel(const float4 *accum, uint32_t width, uint32_t height, uint32_t yBegin, uint32_t yEnd,
                               uint32_t xBegin, uint32_t xEnd, uint8_t *rgb8)
{
    uint32_t const x = xBegin + blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t const y = yBegin + blockIdx.y;
    if (x >= xEnd || y >= yEnd)
        return;
    float4 const a = accum[size_t(y) * width + x];
    float const inv = fdiv(1.f, a.w);
    uint32_t const size = width * height * 3u;
    uint32_t const index = size - ((width - x) * 3u + y * width * 3u);
    rgb8[index + 0] = encode_channel(fmul(a.x, inv));
    rgb8[index + 1] = encode_channel(fmul(a.y, inv));
    rgb8[index + 2] = encode_channel(fmul(a.z, inv));
}

// ---- multi-device epilogue: reduce + resolve, fused ---------------------------------------------------
// Several GPUs each hold a float4 accumulator of the same image (sample ranges of every pixel, or disjoint row bands with
// zeros elsewhere).  This kernel runs on EVERY device over its own slice of the rows: it sums the n accumulators in device
// order through peer-mapped pointers (loads over NVLink), optionally keeps the total (sumOut, which may alias this device's
// own accumulator: a thread reads and writes only its own pixel), and encodes the pixel exactly as resolve_kernel does.
// With disjoint bands the sum adds zeros to one value: the single-GPU bits.
constexpr int kMaxDevices = 16;
struct ReduceParams
{
    const float4 *acc[kMaxDevices];
    uint32_t n;
};
__global__ void reduce_resolve_kernel(const __grid_constant__ ReduceParams r, uint32_t width, uint32_t height, uint32_t yBegin, uint32_t yEnd,
                                      uint32_t xBegin, uint32_t xEnd, float4 *sumOut, uint8_t *rgb8)
{
    uint32_t const x = xBegin + blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t const y = yBegin + blockIdx.y;
    if (x >= xEnd || y >= yEnd)
        return;
    size_t const i = size_t(y) * width + x;
    float4 a = r.acc[0][i];
    for (uint32_t k = 1; k < r.n; ++k)
    {
        float4 const b = r.acc[k][i];
        a.x = fadd(a.x, b.x);
        a.y = fadd(a.y, b.y);
        a.z = fadd(a.z, b.z);
        a.w = fadd(a.w, b.w);
    }
    if (sumOut)
        sumOut[i] = a;
    if (rgb8)
    {
        float const inv = fdiv(1.f, a.w);
        uint32_t const size = width * height * 3u;
        uint32_t const index = size - ((width - x) * 3u + y * width * 3u);
        rgb8[index + 0] = encode_channel(fmul(a.x, inv));
        rgb8[index + 1] = encode_channel(fmul(a.y, inv));
        rgb8[index + 2] = encode_channel(fmul(a.z, inv));
    }
}

// ---- FP32 peak microbenchmark ---------------------------------------------------------------------
// 8 independent dependent-FMA chains per thread; scalar fma.rn.f32 and packed fma.rn.f32x2.
__global__ void __launch_bounds__(256) fp32_peak_scalar(float *out, uint32_t iters, float b, float c)
{
    float a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
        a[k] = float(threadIdx.x + k);
    for (uint32_t i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
            a[k] = __fmaf_rn(a[k], b, c);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        s += a[k];
    if (s == 123.456f)
        out[0] = s;
}

__global__ void __launch_bounds__(256) fp32_peak_packed(float *out, uint32_t iters, float b, float c)
{
    unsigned long long a[8];
    unsigned long long bb, cc;
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
    asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
#pragma unroll
    for (int k = 0; k < 8; ++k)
    {
        float const v = float(threadIdx.x + k);
        asm("mov.b64 %0, {%1, %1};" : "=l"(a[k]) : "f"(v));
    }
    for (uint32_t i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[k]) : "l"(bb), "l"(cc));
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k)
    {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[k]));
        s += lo + hi;
    }
    if (s == 123.456f)
        out[0] = s;
}

} // namespace spt
