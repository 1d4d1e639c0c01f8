// spt_tensor_filter.cuh -- EXPERIMENTAL, compiled only with -DSPT_TENSOR_FILTER=1 (default 0: not part of the library).
//
// The bounce kernel of split rendering with the pre-filter's discriminant on the 5th-generation tensor cores.  Per iteration
// each RAY TILE (128 rays = four warps, one thread per tensor-memory lane) multiplies its rays' rows with the sphere matrix --
// the first min(nSlots, 128) filter slots -- as ONE M = 128, K = 40 TF32 GEMM (five tcgen05.mma) into tensor memory and reads
// the result back as sign bits; slots past 128 stay with the FFMA2 loop (filter_span).  A CTA holds kTensorTiles ray tiles behind
// ONE staged copy of the scene and of the sphere matrix; tiles run independently (named barriers, one MMA barrier each).
// Layout, algebra and slack: spt_tensor_host.hpp.  The verdict is the discriminant's sign only, as in filter_pair; candidates go
// through exact_sphere<> as always, so the image cannot change -- only a MISSED true hit could change it, which is what the
// bit-exact parity tests and tools/tensor_filter_check.cpp are there to exclude.
//
// Status: the mechanism is the stand-alone prototype tools/umma_filter_microbench.cu, which ran on B200
// (profiles/r1_microbench_umma_filter.txt: no true hit missed; 4.1 - 4.9 T (ray, sphere) pairs/s).  THIS kernel compiles and its
// host half is checked on the CPU, but it has not run on a GPU yet (round 1 had no GPU minutes left): DESIGN.md section 9, item 1.
// Budget per SM with two tiles per CTA and two CTAs: 16 warps, 4 x 128 = 512 tensor-memory columns, ~180 KB shared memory.
// Measured (profiles/r1_ab_tensor_filter.txt): bit-identical, but 9 % slower than the FFMA2 kernel on config 3 -- 16 warps per SM
// do not hide the latencies 32 did.  Hence the time-shared variant below (-DSPT_TENSOR_TILES=4 -DSPT_TENSOR_ACC=2: 32 warps per SM
// at 64 registers, the tiles of a CTA taking turns on two accumulators), which compiles (4 bytes of spill) and has NOT run yet.
// Also NOT run yet: primary_hits_tc, the camera kernel in the same mould (camera form: K = 32, six products of the direction +
// the constant).  It should suit the tensor core better than the bounce kernel does: convergent by construction, no shading
// between closest hits, and ~500 issue cycles per 32 rays instead of ~840 for the FFMA2 camera loop alone.
#pragma once

#if SPT_TENSOR_FILTER

#include "spt_split.cuh"
#include "spt_tensor_host.hpp"

namespace spt
{

#ifndef SPT_TENSOR_TILES
#define SPT_TENSOR_TILES 2
#endif
#ifndef SPT_TENSOR_ACC
#define SPT_TENSOR_ACC SPT_TENSOR_TILES
#endif
constexpr int kTensorTiles = SPT_TENSOR_TILES;
// Accumulators (128 tensor-memory columns + one ray matrix each) per CTA.  Equal to the tile count: every tile owns one.  Fewer:
// the tiles TIME-SHARE them -- an accumulator is busy only from the row stores to the last tcgen05.ld of an iteration, a fraction
// of it -- which is how more than 16 warps per SM fit behind 512 columns (-DSPT_TENSOR_TILES=4 -DSPT_TENSOR_ACC=2: two CTAs of
// sixteen warps at 64 registers).
#ifndef SPT_TENSOR_MERGED_BARRIER
#define SPT_TENSOR_MERGED_BARRIER 0 // 1: the loop's "any lane of the tile alive?" reduction rides on a barrier the closest hit has anyway
#endif
constexpr bool kTensorMergedBarrier = SPT_TENSOR_MERGED_BARRIER != 0;
constexpr int kTensorAcc = SPT_TENSOR_ACC;
constexpr bool kTensorTimeShared = kTensorAcc < kTensorTiles;
static_assert(kTensorTiles == 1 || kTensorTiles == 2 || kTensorTiles == 4, "one, two or four ray tiles per CTA");
static_assert((kTensorAcc == 1 || kTensorAcc == 2 || kTensorAcc == 4) && kTensorAcc <= kTensorTiles, "accumulators per CTA");
using TensorShape = BlockShape<2 * kTensorTiles, 2>; // 16x8, 32x8 or 64x8 pixels
static_assert(TensorShape::kThreads == int(kTensorRays) * kTensorTiles, "one thread per tensor-memory lane");

// UMMA shared-memory descriptor, K-major, SWIZZLE_NONE: start >> 4 in [0,14), LBO >> 4 in [16,30), SBO >> 4 in [32,46), version 1
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo)
{
    return uint64_t((addr >> 4) & 0x3FFFu) | (uint64_t((lbo >> 4) & 0x3FFFu) << 16) | (uint64_t((sbo >> 4) & 0x3FFFu) << 32) | (uint64_t(1) << 46);
}
// instruction descriptor: D = F32 (bit 4), A = B = TF32 (bits 7, 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__device__ __forceinline__ uint32_t umma_idesc_tf32(uint32_t m, uint32_t n)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmemD, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmemD), "l"(ad), "l"(bd), "r"(idesc), "r"(accumulate)
                 : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tensor_mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
}
// the four warps of one ray tile (barrier 0 is __syncthreads)
__device__ __forceinline__ void tile_sync(uint32_t tile)
{
    asm volatile("bar.sync %0, %1;" ::"r"(tile + 1u), "r"(kTensorRays) : "memory");
}
__device__ __forceinline__ bool tile_sync_or(uint32_t tile, bool pred)
{
    uint32_t r;
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %3, 0;\n\tbar.red.or.pred p, %1, %2, q;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(r)
                 : "r"(tile + 1u), "r"(kTensorRays), "r"(pred ? 1u : 0u)
                 : "memory");
    return r != 0u;
}
// 32 consecutive accumulator columns of this thread's TMEM lane -> bit (31 - j) set <=> column j >= 0 (a candidate)
__device__ __forceinline__ uint32_t tmem_signs32(uint32_t taddr)
{
    uint32_t v[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t mask = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j)
        mask = __funnelshift_l(v[j], mask, 1); // sign bit = fail
    return ~mask;
}

// accumulator bookkeeping of a CTA whose tiles time-share them (kTensorTimeShared)
struct TensorPool
{
    uint32_t freeMask;            // bit a set <=> accumulator a is free
    uint32_t holder[4];           // per tile: the accumulator its current iteration holds (written by the tile's first thread)
    uint32_t readers[4];          // per accumulator: warps of the holding tile that have finished reading it
};
// first thread of a tile: take a free accumulator (spins; holders never wait on anything but their own tile, so it frees up)
__device__ __forceinline__ uint32_t pool_acquire(TensorPool *pool)
{
    for (;;)
    {
        uint32_t const m = *reinterpret_cast<volatile uint32_t *>(&pool->freeMask);
        if (m)
        {
            uint32_t const a = uint32_t(__ffs(int(m))) - 1u;
            if (atomicCAS(&pool->freeMask, m, m & ~(1u << a)) == m)
                return a;
        }
        else
            __nanosleep(100);
    }
}
// every warp of the holding tile, once its tcgen05.ld have completed: the last of the four hands the accumulator back
__device__ __forceinline__ void pool_release(TensorPool *pool, uint32_t a)
{
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncwarp();
    if ((threadIdx.x & 31u) == 0u)
    {
        __threadfence_block();
        if (atomicAdd(&pool->readers[a], 1u) == kTensorRays / 32u - 1u)
        {
            atomicExch(&pool->readers[a], 0u);
            __threadfence_block();
            atomicOr(&pool->freeMask, 1u << a);
        }
    }
}

// what a ray tile needs to run the tensor filter
struct TensorFilter
{
    float *sA;       // the tile's rows (kTensorABytes), 128-byte aligned; time-shared: ray matrix of accumulator 0
    TensorPool *pool; // time-shared accumulators only
    uint32_t bAddr;  // shared-memory address of the sphere matrix
    uint32_t bLbo;
    uint32_t rows;   // of the sphere matrix = N: multiple of 16, <= kTensorCols
    uint32_t slots;  // filter slots the matrix covers (the first ones)
    uint32_t tmem;   // base of the tile's accumulator columns
    uint32_t bar;    // shared-memory address of the tile's MMA-complete barrier
    uint32_t phase;  // parity the next wait uses
    uint32_t tile;
};

// This ray's row of the A matrix: ten factors, q rounded so that candidates can only be ADDED, each split head / tail.
__device__ __forceinline__ void tensor_store_ray(float *sA, uint32_t lane, V3 dir, V3 org, float od)
{
    float const ooLo = __fmaf_rd(org.z, org.z, __fmaf_rd(org.y, org.y, __fmul_rd(org.x, org.x))); // <= |O|^2
    float const ooUp = __fmaf_ru(org.z, org.z, __fmaf_ru(org.y, org.y, __fmul_ru(org.x, org.x))); // >= |O|^2
    float const q = __fadd_ru(__fsub_ru(__fmul_ru(od, od), ooLo), __fmul_ru(ooUp, float(kTensorSlackU) * 1.0001f));
    float const od2 = fmul(2.f, od);
    float const f[10] = {fmul(dir.x, dir.x), fmul(dir.x, dir.y), fmul(dir.x, dir.z), fmul(dir.y, dir.y), fmul(dir.y, dir.z),
                         fmul(dir.z, dir.z), fsub(fmul(2.f, org.x), fmul(od2, dir.x)), fsub(fmul(2.f, org.y), fmul(od2, dir.y)),
                         fsub(fmul(2.f, org.z), fmul(od2, dir.z)), q};
    float *row = sA + (lane >> 3) * 32 + (lane & 7) * 4; // (row / 8) * 128 B + (row % 8) * 16 B
#pragma unroll
    for (int c = 0; c < 10; ++c)
    {
        float const h = __uint_as_float(__float_as_uint(f[c]) & 0xFFFFE000u), l = fsub(f[c], h); // exact; the MMA truncates l to TF32
        sts128(reinterpret_cast<float4 *>(row + c * (kTensorALbo / 4)), c < 9 ? make_float4(h, l, h, l) : make_float4(1.f, 1.f, h, l));
    }
}

// Camera form (spt_tensor_host.hpp, build_tensor_matrix_cam): the six products of the direction, head / tail.  Chunks 6 and 7 of
// a row are constants, written once per launch by tensor_init_rows_cam.
__device__ __forceinline__ void tensor_store_ray_cam(float *sA, uint32_t lane, V3 dir)
{
    float const f[6] = {fmul(dir.x, dir.x), fmul(dir.x, dir.y), fmul(dir.x, dir.z), fmul(dir.y, dir.y), fmul(dir.y, dir.z), fmul(dir.z, dir.z)};
    float *row = sA + (lane >> 3) * 32 + (lane & 7) * 4;
#pragma unroll
    for (int c = 0; c < 6; ++c)
    {
        float const h = __uint_as_float(__float_as_uint(f[c]) & 0xFFFFE000u), l = fsub(f[c], h);
        sts128(reinterpret_cast<float4 *>(row + c * (kTensorALbo / 4)), make_float4(h, l, h, l));
    }
}
__device__ __forceinline__ void tensor_init_rows_cam(float *sA, uint32_t lane)
{
    float *row = sA + (lane >> 3) * 32 + (lane & 7) * 4;
    sts128(reinterpret_cast<float4 *>(row + 6 * (kTensorALbo / 4)), make_float4(1.f, 1.f, 0.f, 0.f));
    sts128(reinterpret_cast<float4 *>(row + 7 * (kTensorALbo / 4)), make_float4(0.f, 0.f, 0.f, 0.f));
}

// FindClosestIntersectionSphere with the discriminant of the first T.slots slots from the tensor core.  CAM: the camera form
// (org = the eye the matrix and `fs` were specialised for).  Must be called by ALL
// threads of the ray tile (a tile-wide barrier inside); `live` = false marks a lane that only keeps the others company.
// kTensorMergedBarrier: `alive` (the lane has a path in flight or samples left) is OR-reduced over the tile on the first barrier;
// when nobody is, anyAlive comes back false, nothing was traced and the caller leaves its loop (tile-uniformly).
template <bool CAM = false, typename GeomPtr>
__device__ __forceinline__ uint32_t closest_hit_tensor(const FilterScene &fs, GeomPtr geom, uint32_t n, V3 dir, V3 org, TensorFilter &T,
                                                       uint32_t &exactEvals, bool live, bool alive, bool &anyAlive)
{
    anyAlive = true;
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    float const od = vdot(org, dir);
    uint32_t const lane = threadIdx.x & (kTensorRays - 1u);
    float *sA = T.sA;
    uint32_t tmem = T.tmem, held = 0u;
    if (kTensorTimeShared)
    {
        if (lane == 0)
            T.pool->holder[T.tile] = pool_acquire(T.pool);
        __syncwarp(); // bar.sync is warp-aligned: the spinning lane has to be back
        if (kTensorMergedBarrier)
            anyAlive = tile_sync_or(T.tile, alive);
        else
            tile_sync(T.tile);
        held = *reinterpret_cast<volatile uint32_t *>(&T.pool->holder[T.tile]);
        if (!anyAlive)
        {
            if (lane == 0)
                atomicOr(&T.pool->freeMask, 1u << held); // taken but never used
            return n;
        }
        sA += held * ((CAM ? kTensorABytesCam : kTensorABytes) / 4u);
        tmem += held * kTensorCols;
    }
    if (CAM)
        tensor_store_ray_cam(sA, lane, dir);
    else
        tensor_store_ray(sA, lane, dir, org, od);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); // this thread's earlier tcgen05.ld are done with the accumulator
    if (kTensorMergedBarrier && !kTensorTimeShared)
    {
        anyAlive = tile_sync_or(T.tile, alive);
        if (!anyAlive)
            return n;
    }
    else
        tile_sync(T.tile);
    if (lane == 0)
    {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t const aAddr = smem_addr(sA), idesc = umma_idesc_tf32(kTensorRays, T.rows);
#pragma unroll
        for (uint32_t ks = 0; ks < (CAM ? kTensorKCam : kTensorK) / 8u; ++ks) // one UMMA K step = 8 TF32 = two 16-byte chunks
            umma_tf32(tmem, umma_desc(aAddr + ks * 2u * kTensorALbo, kTensorALbo, 128u),
                      umma_desc(T.bAddr + ks * 2u * T.bLbo, T.bLbo, 128u), idesc, ks > 0u ? 1u : 0u);
        umma_commit(T.bar);
    }
    // the ground and the slots past the matrix while the tensor core works
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    if (fs.nSlots > T.slots) // T.slots == kTensorCols here: a whole number of words
    {
        FilterRay const fr = CAM ? make_filter_ray_cam(dir, od, fs.epsTCam) : make_filter_ray(dir, org, od, fs.amax);
        filter_span<CAM>(fs.pairs + T.slots / 2u, fs.slotIdx + T.slots, fs.nSlots - T.slots, fr, geom, n, dir, org, od, minIndex,
                         minDistanceSq, exactEvals, live);
    }
    tensor_mbar_wait(T.bar, T.phase);
    T.phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- this warp's 32 lanes: sign bits per 32-slot word (at most four: one resolution round), then the exact sequence ----
    uint32_t const lanes = tmem + ((((threadIdx.x >> 5) & 3u) * 32u) << 16);
    uint32_t const nWords = (T.slots + 31u) / 32u;
    uint32_t c0 = 0u, c1 = 0u, c2 = 0u, c3 = 0u;
#pragma unroll 1
    for (uint32_t w = 0; w < nWords; ++w)
    {
        uint32_t cw = tmem_signs32(lanes + w * 32u);
        // columns past T.slots -- padding rows, or columns the MMA never wrote -- must not index slotIdx
        uint32_t const valid = T.slots - w * 32u;
        if (valid < 32u)
            cw &= ~(0xFFFFFFFFu >> valid);
        c3 = c2;
        c2 = c1;
        c1 = c0;
        c0 = live ? cw : 0u;
    }
    if (kTensorTimeShared)
        pool_release(T.pool, held); // the masks are in registers: the accumulator is free while the candidates are resolved
    resolve_candidates(c0, c1, c2, c3, nWords, fs.slotIdx, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals);
    return minIndex;
}

// trace_secondary<2, S> with closest_hit_tensor.  Dynamic shared memory: [geometry | filter blob | pad to 128 | sphere matrix |
// kTensorAcc ray matrices].  Each ray tile's loop is tile-uniform (barrier inside the closest hit).
__global__ void __launch_bounds__(TensorShape::kThreads, kTensorTimeShared ? 512 / (kTensorCols * kTensorAcc) : 2) trace_secondary_tc(const __grid_constant__ TraceParams p)
{
    using S = TensorShape;
    extern __shared__ __align__(128) unsigned char smemTensor[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ __align__(8) uint64_t mmaBar[kTensorTiles];
    __shared__ uint32_t tmemBase;
    __shared__ TensorPool pool;
    __shared__ ShellSlots shellSlots[S::kThreads / 32];
    ShellSlots &slots = shellSlots[threadIdx.x >> 5];
    __shared__ LaneStashT<S::kThreads> stash;

    uint32_t const geomBytes = p.n * uint32_t(sizeof(float4));
    uint32_t const blobBytes = filter_blob_bytes(p.filter.nSlots);
    stage_scene_tma(smemTensor, p.geom, geomBytes, smemTensor + geomBytes, p.filter.pairs, blobBytes, &mbar);
    const float4 *geom = reinterpret_cast<const float4 *>(smemTensor);
    FilterScene fs = p.filter;
    rebase_filter(fs, smemTensor + geomBytes);
    uint32_t const bOff = (geomBytes + blobBytes + 127u) & ~127u;
    uint32_t const bBytes = tensor_b_bytes(p.tensorRows);
    for (uint32_t i = threadIdx.x; i < bBytes / 16u; i += S::kThreads) // the sphere matrix, once per CTA
        reinterpret_cast<float4 *>(smemTensor + bOff)[i] = __ldg(reinterpret_cast<const float4 *>(p.tensorB) + i);
    if (threadIdx.x == 0)
    {
        for (int t = 0; t < kTensorTiles; ++t)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&mmaBar[t])));
        pool.freeMask = (1u << kTensorAcc) - 1u;
        for (int a = 0; a < 4; ++a)
            pool.readers[a] = pool.holder[a] = 0u;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if ((threadIdx.x >> 5) == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&tmemBase)),
                     "n"(kTensorCols * kTensorAcc));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the sphere matrix was written with generic stores
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    TensorFilter T;
    T.tile = threadIdx.x / kTensorRays;
    T.sA = reinterpret_cast<float *>(smemTensor + bOff + bBytes + (kTensorTimeShared ? 0u : T.tile * kTensorABytes));
    T.pool = &pool;
    T.bAddr = smem_addr(smemTensor + bOff);
    T.bLbo = tensor_b_lbo(p.tensorRows);
    T.rows = p.tensorRows;
    T.slots = tensor_slots(p.filter.nSlots);
    T.tmem = tmemBase + (kTensorTimeShared ? 0u : T.tile * kTensorCols);
    T.bar = smem_addr(&mmaBar[T.tile]);
    T.phase = 0u;

    uint32_t x, y;
    bool inside;
    lane_pixel<S>(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    Lane L;
    float4 const acc0 = inside ? p.accum[pixel] : make_float4(0.f, 0.f, 0.f, 0.f); // the running sum continues
    AccShared const acc{&stash.a[threadIdx.x]};
    sts128(acc.slot, acc0);
    L.sum = v3(0.f, 0.f, 0.f); // unused
    L.count = 0.f;
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
    stash_lane_constants(stash, pixel_key(p.seed, pixel),
                         (uint64_t(lane_row_local<S>()) * (p.xEnd - p.xBegin) + (x - p.xBegin)) * (p.sampleEnd - p.sampleBegin));

    bool live = inside;
    uint32_t idx = p.n; // one surface-shading site per iteration, as in trace_secondary
    bool traced = false;
    for (;;)
    {
        SurfaceKind kind = shade_classify(p, L, idx, traced, acc);
        while (live && L.mode == MODE_NEW)
        {
            if (L.sample >= p.sampleEnd)
            {
                live = false;
                break;
            }
            uint64_t pixKey, hitBase;
            lane_constants(stash, pixKey, hitBase);
            float4 const rec = ld_first_hit(p.firstHits + hitBase + (L.sample - p.sampleBegin));
            uint32_t const h = __float_as_uint(rec.w);
            L.dir = v3(rec.x, rec.y, rec.z);
            ++L.sample;
            SurfaceKind const fk = h == kMissIndex ? surface_none() : surface_of_material(uint32_t(__ldg(p.mat + h)));
            if (!fk.any())
            {
                acc.add(sample_sky(p.cam, L.dir));
                continue;
            }
            L.rng.key = stream_key(pixKey, L.sample - 1u);
            L.rng.event = 1;
            L.org = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
            L.specular = 0;
            L.mode = MODE_TRACE;
            kind = fk;
            idx = h;
        }
        if (!kTensorMergedBarrier && !tile_sync_or(T.tile, live)) // tile-uniform: the closest hit below has a tile-wide barrier
            break;
        shade_surface(p, geom, L, idx, kind, dropped, slots, acc); // warp collective (coop_shell); every warp of the tile is here
        bool const needTrace = live && L.mode != MODE_NEW;
        stash_lane(stash, L);
        bool anyAlive;
        idx = closest_hit_tensor(fs, geom, p.n, L.dir, L.org, T, exactEvals, needTrace, live, anyAlive);
        if (!anyAlive) // kTensorMergedBarrier only: every lane of the tile has finished its pixel
            break;
        unstash_lane(stash, L);
        segments += needTrace ? 1u : 0u;
        traced = needTrace;
    }
    if (inside)
    {
        lane_pixel<S>(p, x, y, inside);
        p.accum[y * p.width + x] = lds128(acc.slot);
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, segments);
        warp_counter_add(p.counters + 2, dropped);
        warp_counter_add(p.counters + 3, exactEvals);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads(); // both tiles are done with tensor memory
    if ((threadIdx.x >> 5) == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "n"(kTensorCols * kTensorAcc));
}

// primary_hits_kernel<2, S> with the camera form on the tensor core: one camera sample per lane and trip (the FFMA2 kernel takes
// two, to halve its shared-memory traffic -- no such traffic here), every tile of the CTA in step with itself only.  The loop
// is convergent by construction (same trip count for every lane), which is what the tile barriers want.  Dynamic shared
// memory: [geometry | camera filter blob | pad to 128 | camera sphere matrix | kTensorAcc ray matrices].
__global__ void __launch_bounds__(TensorShape::kThreads, kTensorTimeShared ? 512 / (kTensorCols * kTensorAcc) : 2) primary_hits_tc(const __grid_constant__ TraceParams p)
{
    using S = TensorShape;
    extern __shared__ __align__(128) unsigned char smemTensor[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ __align__(8) uint64_t mmaBar[kTensorTiles];
    __shared__ uint32_t tmemBase;
    __shared__ TensorPool pool;

    uint32_t const geomBytes = p.n * uint32_t(sizeof(float4));
    uint32_t const blobBytes = filter_blob_bytes(p.filterCam.nSlots);
    stage_scene_tma(smemTensor, p.geom, geomBytes, smemTensor + geomBytes, p.filterCam.pairs, blobBytes, &mbar);
    const float4 *geom = reinterpret_cast<const float4 *>(smemTensor);
    FilterScene fs = p.filterCam;
    rebase_filter(fs, smemTensor + geomBytes);
    uint32_t const bOff = (geomBytes + blobBytes + 127u) & ~127u;
    uint32_t const bBytes = tensor_b_bytes_cam(p.tensorRows);
    for (uint32_t i = threadIdx.x; i < bBytes / 16u; i += S::kThreads)
        reinterpret_cast<float4 *>(smemTensor + bOff)[i] = __ldg(reinterpret_cast<const float4 *>(p.tensorBCam) + i);
    float *aBase = reinterpret_cast<float *>(smemTensor + bOff + bBytes);
    for (uint32_t i = threadIdx.x; i < kTensorRays * uint32_t(kTensorAcc); i += S::kThreads) // constant chunks of every ray matrix
        tensor_init_rows_cam(aBase + (i / kTensorRays) * (kTensorABytesCam / 4u), i % kTensorRays);
    if (threadIdx.x == 0)
    {
        for (int t = 0; t < kTensorTiles; ++t)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&mmaBar[t])));
        pool.freeMask = (1u << kTensorAcc) - 1u;
        for (int a = 0; a < 4; ++a)
            pool.readers[a] = pool.holder[a] = 0u;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if ((threadIdx.x >> 5) == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&tmemBase)),
                     "n"(kTensorCols * kTensorAcc));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    TensorFilter T;
    T.tile = threadIdx.x / kTensorRays;
    T.sA = aBase + (kTensorTimeShared ? 0u : T.tile * (kTensorABytesCam / 4u));
    T.pool = &pool;
    T.bAddr = smem_addr(smemTensor + bOff);
    T.bLbo = tensor_b_lbo(p.tensorRows);
    T.rows = p.tensorRows;
    T.slots = tensor_slots(p.filterCam.nSlots);
    T.tmem = tmemBase + (kTensorTimeShared ? 0u : T.tile * kTensorCols);
    T.bar = smem_addr(&mmaBar[T.tile]);
    T.phase = 0u;

    uint32_t x, y;
    bool inside;
    lane_pixel<S>(p, x, y, inside);
    uint32_t const pixel = y * p.width + x;
    float const widthF = __uint2float_rn(p.width), heightF = __uint2float_rn(p.height);
    uint64_t const pixKey = pixel_key(p.seed, pixel);
    V3 const eye = v3(p.cam.eye[0], p.cam.eye[1], p.cam.eye[2]);
    uint32_t const nSamples = p.sampleEnd - p.sampleBegin;
    float4 *out = p.firstHits + (size_t(lane_row_local<S>()) * (p.xEnd - p.xBegin) + (x - p.xBegin)) * nSamples;
    uint32_t exactEvals = 0;
    if (tile_sync_or(T.tile, inside)) // a tile wholly outside the image has nothing to do (tile-uniform)
    {
        for (uint32_t s = 0; s < nSamples; ++s) // the same trip count for every lane of the tile
        {
            Rng rng;
            rng.key = stream_key(pixKey, p.sampleBegin + s);
            rng.event = 0;
            rng.open_event();
            float xiU, xiV;
            rng.uniform2(-1.f, 1.f, xiU, xiV);
            V3 const dir = primary_direction(p.cam, widthF, heightF, x, y, xiU, xiV);
            bool anyAlive;
            uint32_t const idx = closest_hit_tensor<true>(fs, geom, p.n, dir, eye, T, exactEvals, inside, true, anyAlive);
            if (inside)
                out[s] = first_hit_record(dir, idx, p.n);
        }
    }
    if (p.counters)
    {
        __syncwarp();
        warp_counter_add(p.counters + 0, inside ? nSamples : 0u);
        warp_counter_add(p.counters + 1, inside ? nSamples : 0u);
        warp_counter_add(p.counters + 3, exactEvals);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if ((threadIdx.x >> 5) == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "n"(kTensorCols * kTensorAcc));
}

} // namespace spt

#endif // SPT_TENSOR_FILTER
