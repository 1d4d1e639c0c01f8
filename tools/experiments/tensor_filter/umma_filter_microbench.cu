// umma_filter_microbench.cu -- the closest-hit pre-filter's discriminant as ONE tcgen05 (UMMA) GEMM per CTA and
// iteration: groundwork for DESIGN.md section 9, item 1.  Ran on B200 at the very end of round 1
// (profiles/r1_microbench_umma_filter.txt): self-check clean (no true hit missed, 0.107 candidates for 0.103 true hits per
// ray, worst error 19 u(A^2+B^2)); 1 973 cycles per iteration with one CTA per SM (the whole chain is serial inside a CTA),
// 4.07 .. 4.91 T (ray, sphere) pairs/s with two -- above the 3.72 T pairs/s that 100 % of the issue slots would give the FFMA2
// loop.  -DUMMA_CHUNKED=1 (five 32-slot chunks through two 32-column accumulators, 64 TMEM columns per CTA) is correct too but
// slower: ~650 cycles per mma -> commit -> wait -> ld round trip, 4.36 T pairs/s with four CTAs per SM.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_filter_microbench tools/umma_filter_microbench.cu
//   ./umma_filter_microbench [ctas_per_sm=2] [iters=200]      (256 TMEM columns per CTA: two CTAs fit an SM, more serialise on tcgen05.alloc)
//
// What it does.  A CTA of 128 threads holds 128 rays (M = 128), the sphere list is N = 160 filter slots (152 + padding),
// and the conservative discriminant of spt_filter.cuh,
//     v = (C.D - O.D)^2 - (|C|^2 - r^2) + 2 C.O - |O|^2 + S_i + S_ray - 1e-3,
// is written as a K = 40 inner product whose factors separate into a per-ray and a per-sphere part (the square expanded):
//     v = sum_{i<=j} (c_ij C_i C_j)(D_i D_j)  +  sum_i C_i W_i  +  k  +  q,
//     W = 2 O - 2 (O.D) D,   k = -(|C|^2 - r^2) + S_i - 1e-3,   q = (O.D)^2 - |O|^2 + S_ray,   c_ij = 1 (i == j) or 2.
// Every factor is split hi/lo into TF32 pieces and all four cross products are kept (tools/tf32_filter_numerics.py: worst
// error ~ 30 u (A^2 + B^2), inside the shipped slack), so a quadratic term takes 4 K-slots: 6 x 4 + 3 x 4 + 4 = 40.
// Each group of four K-slots is one 16-byte chunk of the UMMA K-major, no-swizzle canonical layout
//     element (row, k) at  base + (k / 4) * LBO + (row / 8) * SBO + (row % 8) * 16 + (k % 4) * 4   (SBO = 128 B),
// so a thread writes its ray as ten conflict-free 16-byte stores, the sphere matrix is laid out once by the host, and one
// elected thread issues five tcgen05.mma.kind::tf32 (M 128, N 160, K 8) per iteration into 160 TMEM columns.  Every warp
// then reads its 32 lanes back (tcgen05.ld 32x32b.x32, five times) and only shifts sign bits into candidate masks.
//
// Prints: the self-check against binary64 (no true hit may be missed; extra candidates per ray; worst |v - v_exact|), and
// cycles per iteration and per (ray, sphere) pair next to the ~10 issue slots per pair of the FFMA2 loop.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <vector>

constexpr int kRays = 128;   // M: rays per CTA (one per thread, one TMEM lane each)
constexpr int kSlots = 160;  // N: filter slots (multiple of 16 for M = 128)
constexpr int kK = 40;       // inner dimension
#ifndef UMMA_CHUNKED
#define UMMA_CHUNKED 0 // 1: the list in five 32-slot chunks through two 32-column accumulators (64 TMEM columns per CTA, so more
#endif                 //    CTAs fit an SM); the MMA of chunk c + 2 is issued as soon as every warp has read chunk c
constexpr int kTmemCols = UMMA_CHUNKED ? 64 : 256; // power of two >= the accumulator columns
constexpr uint32_t kABytes = kRays * kK * 4, kBBytes = kSlots * kK * 4;
constexpr uint32_t kALbo = (kRays / 8) * 128, kBLbo = (kSlots / 8) * 128, kSbo = 128;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return uint32_t(__cvta_generic_to_shared(p)); }

// UMMA shared-memory descriptor, K-major, SWIZZLE_NONE (cute::UMMA::SmemDescriptor): start >> 4 in [0,14), LBO >> 4 in
// [16,30), SBO >> 4 in [32,46), version 1 in [46,48), layout type 0 in [61,64)
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo)
{
    return uint64_t((addr >> 4) & 0x3FFFu) | (uint64_t((lbo >> 4) & 0x3FFFu) << 16) | (uint64_t((sbo >> 4) & 0x3FFFu) << 32) | (uint64_t(1) << 46);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = TF32, both K-major, dense
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kSlots >> 3) << 17) | (uint32_t(kRays >> 4) << 24);
constexpr uint32_t kIdescChunk = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(32 >> 3) << 17) | (uint32_t(kRays >> 4) << 24); // N = 32

__device__ __forceinline__ void umma_tf32(uint32_t tmemD, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmemD), "l"(ad), "l"(bd), "r"(idesc), "r"(accumulate)
                 : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    // arrives on the barrier when every MMA issued so far is complete (implies tcgen05.fence::before_thread_sync)
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane -> sign bits (bit 31 - j set <=> column j >= 0)
__device__ __forceinline__ uint32_t tmem_signs32(uint32_t taddr, float *dump)
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
    if (dump)
        for (int j = 0; j < 32; ++j)
            dump[j] = __uint_as_float(v[j]);
    return ~mask;
}

__device__ __forceinline__ void split_tf32(float x, float &hi, float &lo)
{
    hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
    lo = x - hi; // exact; the tensor core drops its low 13 bits itself
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
}

struct Ray
{
    float ox, oy, oz, dx, dy, dz;
};

// masks: [ray][5] words, bit (31 - j) of word w set <=> slot 32 w + j is a candidate.  vdump (optional): [ray][kSlots]
__global__ void __launch_bounds__(kRays) umma_filter(const float *__restrict__ bmat, const Ray *__restrict__ rays, uint32_t raysTotal,
                                                     uint32_t iters, uint32_t *__restrict__ masks, float *__restrict__ vdump,
                                                     unsigned long long *__restrict__ cycles)
{
    extern __shared__ __align__(128) unsigned char smem[];
    float *sB = reinterpret_cast<float *>(smem);
    float *sA = reinterpret_cast<float *>(smem + kBBytes);
    __shared__ __align__(8) uint64_t mmaBar[2];
    __shared__ uint32_t tmemBase;
    uint32_t const tid = threadIdx.x, warp = tid >> 5;

    // ---- one-time: sphere matrix into shared memory (plain loads here; the product would TMA it), barrier, TMEM ----
    for (uint32_t i = tid; i < kBBytes / 16; i += kRays)
        reinterpret_cast<float4 *>(sB)[i] = reinterpret_cast<const float4 *>(bmat)[i];
    if (tid == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mmaBar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mmaBar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmemBase)), "n"(kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t const tmem = tmemBase;

    uint32_t const rayIdx = (blockIdx.x * kRays + tid) % raysTotal;
    Ray const ray = rays[rayIdx];
    uint32_t m0 = 0, m1 = 0, m2 = 0, m3 = 0, m4 = 0;
    uint32_t phase = 0; // bit b = parity the next wait on mmaBar[b] uses
    long long const t0 = clock64();
    for (uint32_t it = 0; it < iters; ++it)
    {
        // ---- per-ray factors (the product recomputes these per segment; here the ray is perturbed so nothing is hoisted) ----
        float const ox = ray.ox + float(it) * 0.f, oy = ray.oy, oz = ray.oz, dx = ray.dx, dy = ray.dy, dz = ray.dz;
        float const od = __fadd_rn(__fadd_rn(__fmul_rn(ox, dx), __fmul_rn(oy, dy)), __fmul_rn(oz, dz));
        float const oo = __fmaf_rd(oz, oz, __fmaf_rd(oy, oy, __fmul_rd(ox, ox)));                 // <= |O|^2
        float const ooUp = __fmaf_ru(oz, oz, __fmaf_ru(oy, oy, __fmul_ru(ox, ox)));
        float const sRay = __fmul_ru(ooUp, 160.f / 16777216.f);                                     // slack: 160 u |O|^2
        float const q = __fadd_ru(__fsub_ru(__fmul_ru(od, od), oo), sRay);
        float const f[10] = {dx * dx, dx * dy, dx * dz, dy * dy, dy * dz, dz * dz,
                             2.f * ox - 2.f * od * dx, 2.f * oy - 2.f * od * dy, 2.f * oz - 2.f * od * dz, q};
        // chunk c (16 bytes = 4 K-slots) of row tid:  quadratic / linear: (h, l, h, l);  constants: (1, 1, q_h, q_l)
        float *row = sA + (tid >> 3) * 32 + (tid & 7) * 4; // (row / 8) * 128 B + (row % 8) * 16 B
#pragma unroll
        for (int c = 0; c < 9; ++c)
        {
            float h, l;
            split_tf32(f[c], h, l);
            *reinterpret_cast<float4 *>(row + c * (kALbo / 4)) = make_float4(h, l, h, l);
        }
        {
            float h, l;
            split_tf32(f[9], h, l);
            *reinterpret_cast<float4 *>(row + 9 * (kALbo / 4)) = make_float4(1.f, 1.f, h, l);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy stores -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); // earlier tcgen05.ld of this thread are done with TMEM
        __syncthreads();
        uint32_t const aBase = smem_u32(sA), bBase = smem_u32(sB);
        uint32_t const lanes = tmem + ((warp * 32u) << 16); // this warp's 32 TMEM lanes = its 32 rays
        bool const dumping = vdump && it == 0 && blockIdx.x * kRays + tid < raysTotal;
        uint32_t m[5];
#if UMMA_CHUNKED
        auto issue_chunk = [&](uint32_t c) { // elected thread: chunk c (slots 32 c .. 32 c + 31) into accumulator c & 1
#pragma unroll
            for (int ks = 0; ks < kK / 8; ++ks)
                umma_tf32(tmem + (c & 1u) * 32u, umma_desc(aBase + uint32_t(ks) * 2u * kALbo, kALbo, kSbo),
                          umma_desc(bBase + uint32_t(ks) * 2u * kBLbo + c * 4u * kSbo, kBLbo, kSbo), kIdescChunk, ks > 0 ? 1u : 0u);
            umma_commit(smem_u32(&mmaBar[c & 1u]));
        };
        if (tid == 0)
        {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            issue_chunk(0);
            issue_chunk(1);
        }
#pragma unroll
        for (uint32_t c = 0; c < kSlots / 32; ++c)
        {
            uint32_t const b = c & 1u;
            mbar_wait(smem_u32(&mmaBar[b]), (phase >> b) & 1u);
            phase ^= 1u << b;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            m[c] = tmem_signs32(lanes + b * 32u, dumping ? vdump + size_t(rayIdx) * kSlots + c * 32 : nullptr);
            if (c + 2 < kSlots / 32)
            {
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads(); // every warp has read accumulator b
                if (tid == 0)
                {
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    issue_chunk(c + 2);
                }
            }
        }
#else
        if (tid == 0)
        {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ks = 0; ks < kK / 8; ++ks) // one UMMA K step = 8 TF32 = two 16-byte chunks
                umma_tf32(tmem, umma_desc(aBase + uint32_t(ks) * 2u * kALbo, kALbo, kSbo),
                          umma_desc(bBase + uint32_t(ks) * 2u * kBLbo, kBLbo, kSbo), kIdesc, ks > 0 ? 1u : 0u);
            umma_commit(smem_u32(&mmaBar[0]));
        }
        mbar_wait(smem_u32(&mmaBar[0]), phase & 1u);
        phase ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // epilogue: this warp's 32 lanes x 160 columns; only the signs are wanted
#pragma unroll
        for (int w = 0; w < kSlots / 32; ++w)
            m[w] = tmem_signs32(lanes + uint32_t(w) * 32u, dumping ? vdump + size_t(rayIdx) * kSlots + w * 32 : nullptr);
#endif
        m0 |= m[0]; m1 |= m[1]; m2 |= m[2]; m3 |= m[3]; m4 |= m[4];
    }
    long long const t1 = clock64();
    if (blockIdx.x * kRays + tid < raysTotal)
    {
        uint32_t *m = masks + size_t(rayIdx) * 5;
        m[0] = m0; m[1] = m1; m[2] = m2; m[3] = m3; m[4] = m4;
    }
    if (tid == 0)
        atomicMax(cycles, (unsigned long long)(t1 - t0));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols));
}

static float tf32_trunc(float v)
{
    uint32_t u;
    memcpy(&u, &v, 4);
    u &= 0xFFFFE000u;
    memcpy(&v, &u, 4);
    return v;
}
static float round_up(double d)
{
    float f = float(d);
    if (double(f) < d)
        f = std::nextafterf(f, INFINITY);
    return f;
}

int main(int argc, char **argv)
{
    int const ctasPerSm = argc > 1 ? atoi(argv[1]) : 2;
    uint32_t const iters = argc > 2 ? uint32_t(atoi(argv[2])) : 200;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess)
    {
        fprintf(stderr, "no CUDA device\n");
        return 1;
    }
    int const ctas = prop.multiProcessorCount * ctasPerSm;
    uint32_t const nRays = uint32_t(ctas) * kRays;
    srand(11);
    auto uni = [](double lo, double hi) { return lo + (hi - lo) * (rand() / double(RAND_MAX)); };

    // ---- scene: 152 spheres like config 3's small ones, slots padded to 160 with never-pass records ----
    int const nSpheres = 152;
    std::vector<double> C(kSlots * 3, 0.0), R(kSlots, -1.0);
    for (int s = 0; s < nSpheres; ++s)
    {
        R[s] = uni(0.3, 0.5);
        C[3 * s] = uni(-22, 22);
        C[3 * s + 1] = R[s];
        C[3 * s + 2] = uni(0, 19);
    }
    // B matrix in the canonical layout: element (slot n, k) at (k / 4) * LBO + (n / 8) * 128 + (n % 8) * 16 + (k % 4) * 4 bytes
    std::vector<float> B(kSlots * kK, 0.f);
    auto bput = [&](int n, int k, float v) { B[size_t(k / 4) * (kBLbo / 4) + size_t(n / 8) * 32 + size_t(n % 8) * 4 + (k % 4)] = v; };
    int const pi[6] = {0, 0, 0, 1, 1, 2}, pj[6] = {0, 1, 2, 1, 2, 2};
    for (int n = 0; n < kSlots; ++n)
    {
        float fac[10];
        if (R[n] < 0) // padding: v = -1e30
        {
            for (int c = 0; c < 9; ++c)
                fac[c] = 0.f;
            fac[9] = -1e30f;
        }
        else
        {
            double const *c = &C[3 * n];
            double const a2 = c[0] * c[0] + c[1] * c[1] + c[2] * c[2];
            for (int p = 0; p < 6; ++p)
                fac[p] = float(c[pi[p]] * c[pj[p]] * (pi[p] == pj[p] ? 1.0 : 2.0));
            for (int i = 0; i < 3; ++i)
                fac[6 + i] = float(c[i]);
            fac[9] = round_up(-(a2 - R[n] * R[n]) + 160.0 / 16777216.0 * (a2 + R[n] * R[n]) - double(1e-3f));
        }
        for (int c = 0; c < 9; ++c) // sphere side of (h, l, h, l): (h, h, l, l)
        {
            float const h = tf32_trunc(fac[c]), l = fac[c] - h;
            bput(n, 4 * c + 0, h);
            bput(n, 4 * c + 1, h);
            bput(n, 4 * c + 2, l);
            bput(n, 4 * c + 3, l);
        }
        float const h = tf32_trunc(fac[9]), l = fac[9] - h;
        bput(n, 36, h);
        bput(n, 37, l);
        bput(n, 38, 1.f);
        bput(n, 39, 1.f);
    }
    // ---- rays: origins on the ground / on spheres, directions into the upper half space ----
    std::vector<Ray> rays(nRays);
    for (auto &r : rays)
    {
        double d[3] = {uni(-0.6, 0.6), uni(0.05, 1.0), uni(-0.6, 0.6)};
        double const l = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        r = Ray{float(uni(-25, 25)), float(uni(0, 0.4)), float(uni(-4, 25)), float(d[0] / l), float(d[1] / l), float(d[2] / l)};
    }

    float *dB, *dV;
    Ray *dRays;
    uint32_t *dMasks;
    unsigned long long *dCycles;
    cudaMalloc(&dB, B.size() * 4);
    cudaMalloc(&dRays, rays.size() * sizeof(Ray));
    cudaMalloc(&dMasks, size_t(nRays) * 5 * 4);
    cudaMalloc(&dV, size_t(nRays) * kSlots * 4);
    cudaMalloc(&dCycles, 8);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dRays, rays.data(), rays.size() * sizeof(Ray), cudaMemcpyHostToDevice);
    cudaMemset(dMasks, 0, size_t(nRays) * 5 * 4);
    cudaMemset(dCycles, 0, 8);
    size_t const smemBytes = kABytes + kBBytes;
    cudaFuncSetAttribute(umma_filter, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smemBytes));

    // ---- self-check: one iteration, v dumped ----
    umma_filter<<<ctas, kRays, smemBytes>>>(dB, dRays, nRays, 1, dMasks, dV, dCycles);
    cudaError_t ce = cudaDeviceSynchronize();
    if (ce != cudaSuccess)
    {
        fprintf(stderr, "kernel failed: %s\n", cudaGetErrorString(ce));
        return 1;
    }
    std::vector<uint32_t> masks(size_t(nRays) * 5);
    std::vector<float> V(size_t(nRays) * kSlots);
    cudaMemcpy(masks.data(), dMasks, masks.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(V.data(), dV, V.size() * 4, cudaMemcpyDeviceToHost);
    size_t missed = 0, trueHits = 0, cands = 0;
    double worst = 0.0, worstU = 0.0;
    for (uint32_t r = 0; r < nRays; ++r)
    {
        double const o[3] = {rays[r].ox, rays[r].oy, rays[r].oz}, d[3] = {rays[r].dx, rays[r].dy, rays[r].dz};
        double const od = o[0] * d[0] + o[1] * d[1] + o[2] * d[2], oo = o[0] * o[0] + o[1] * o[1] + o[2] * o[2];
        for (int n = 0; n < nSpheres; ++n)
        {
            double const *c = &C[3 * n];
            double const a2 = c[0] * c[0] + c[1] * c[1] + c[2] * c[2];
            double const tc = c[0] * d[0] + c[1] * d[1] + c[2] * d[2] - od;
            double const disc = tc * tc - (a2 - R[n] * R[n]) + 2 * (c[0] * o[0] + c[1] * o[1] + c[2] * o[2]) - oo;
            bool const flagged = (masks[size_t(r) * 5 + n / 32] >> (31 - n % 32)) & 1u;
            bool const hit = disc > 1e-3;
            trueHits += hit;
            cands += flagged;
            missed += hit && !flagged;
            // the device value carries the slack and the threshold: compare like with like
            double const slack = 160.0 / 16777216.0 * (a2 + R[n] * R[n] + oo);
            double const e = std::fabs(double(V[size_t(r) * kSlots + n]) - (disc + slack - 1e-3));
            worst = std::max(worst, e);
            worstU = std::max(worstU, e / ((a2 + oo) / 16777216.0));
        }
    }
    printf("%s: %d CTAs x %d rays x %d spheres\n", prop.name, ctas, kRays, nSpheres);
    printf("self-check: true hits %.3f / ray, candidates %.3f / ray, MISSED %zu (must be 0), worst |v - v_exact| %.3g = %.1f u(A^2+B^2)\n",
           double(trueHits) / nRays, double(cands) / nRays, missed, worst, worstU);

    // ---- timing ----
    cudaMemset(dCycles, 0, 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    umma_filter<<<ctas, kRays, smemBytes>>>(dB, dRays, nRays, iters, dMasks, nullptr, dCycles);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    unsigned long long cyc = 0;
    cudaMemcpy(&cyc, dCycles, 8, cudaMemcpyDeviceToHost);
    double const pairs = double(ctas) * kRays * double(nSpheres) * iters;
    printf("%d CTAs/SM, %u iterations: %.3f ms, %.0f cycles per iteration (slowest CTA), %.2f T (ray, sphere) pairs/s\n", ctasPerSm, iters, ms,
           double(cyc) / iters, pairs / (ms * 1e-3) / 1e12);
    printf("issue budget today: ~10 slots per pair per lane = %.2f T pairs/s at 100 %% of the issue slots of %d SMs\n",
           prop.multiProcessorCount * 4 * 32 * (prop.clockRate * 1e3) / 10.0 / 1e12, prop.multiProcessorCount);
    return missed ? 2 : 0;
}
