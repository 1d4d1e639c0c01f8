// filter_feed_microbench.cu -- how should the pre-filter's sphere records reach the FMA pipe?
// Times the pre-filter inner loop (spt_filter.cuh: 7 FFMA2 + 1 FADD2 + 2 FMNMX + 2 SHF per record) fed three ways:
//   smem   : records in shared memory, warp-uniform LDS.128 (what trace_megakernel does)
//   cmem   : records in __constant__ memory, LDCU.64 into uniform registers, FFMA2 with UR operands
//   param  : records inside the __grid_constant__ kernel parameter (constant bank 0), same instructions as cmem
// Prints FMA-pipe cycles per record per SM sub-partition (the pipe bound is 16).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o filter_feed_microbench filter_feed_microbench.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>

struct __align__(16) FilterPair { float cx[2], cy[2], cz[2], nk[2]; };
constexpr int kMaxPairs = 1024;
__constant__ FilterPair cRecs[kMaxPairs];
__constant__ float4 cSph[kMaxPairs * 2];
template <int CAP> struct ParamRecs { FilterPair r[CAP]; };

__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{ unsigned long long d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b)
{ unsigned long long d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ unsigned long long pack2(float lo, float hi)
{ unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }

struct Ray { unsigned long long dx, dy, dz, ox, oy, oz, nod, noo; };
// nz = -0.0f at run time.  x + nz == x for every x, but ptxas cannot know that, so with nz the two halves of a ray
// constant are distinct registers and FFMA2 takes the ray as a full F32x2 operand, leaving its one uniform /
// scalar-broadcast operand slot to the sphere record (UR).  With the literal 0 the halves are provably equal and
// ptxas uses the scalar-broadcast form for the RAY instead (fewer registers; sphere record must be a vector register).
__device__ __forceinline__ Ray make_ray(float s, float nz)
{
    Ray r;
    float const a = 0.3f + s, b = 0.5f - s, c = 0.8f + s, d = 2.f * s, e = 2.f + s, f = -6.f - s, g = 1.f + s, h = -10.f + s;
    r.dx = pack2(a, a + nz); r.dy = pack2(b, b + nz); r.dz = pack2(c, c + nz);
    r.ox = pack2(d, d + nz); r.oy = pack2(e, e + nz); r.oz = pack2(f, f + nz);
    r.nod = pack2(g, g + nz); r.noo = pack2(h, h + nz);
    return r;
}
__device__ __forceinline__ void filter_pair(const FilterPair *rec, const Ray &r, uint32_t &mask)
{
    ulonglong2 const a = *reinterpret_cast<const ulonglong2 *>(rec->cx);
    ulonglong2 const b = *reinterpret_cast<const ulonglong2 *>(rec->cz);
    unsigned long long const tc = fma2(b.x, r.dz, fma2(a.y, r.dy, fma2(a.x, r.dx, r.nod)));
    unsigned long long const na = fma2(b.x, r.oz, fma2(a.y, r.oy, fma2(a.x, r.ox, b.y)));
    unsigned long long const v = add2(fma2(tc, tc, na), r.noo);
    float t0, t1, v0, v1;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(t0), "=f"(t1) : "l"(tc));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(v0), "=f"(v1) : "l"(v));
    mask = __funnelshift_l(__float_as_uint(fminf(v0, t0)), mask, 1);
    mask = __funnelshift_l(__float_as_uint(fminf(v1, t1)), mask, 1);
}
template <typename P> __device__ __forceinline__ uint32_t sweep(P recs, uint32_t nWords, const Ray &r)
{
    uint32_t acc = 0;
#pragma unroll 1
    for (uint32_t w = 0; w < nWords; ++w)
    {
        uint32_t mask = 0xFFFFFFFFu;
#pragma unroll
        for (uint32_t p = 0; p < 16; ++p)
            filter_pair(&recs[w * 16 + p], r, mask);
        acc += __popc(~mask);
    }
    return acc;
}

// ---- scalar formulation: 7 FFMA + 1 FADD + FMNMX + SHF per sphere, sphere = one float4 -------------------------------
struct RayS { float dx, dy, dz, ox, oy, oz, nod, noo; };
__device__ __forceinline__ RayS make_ray_s(float s)
{
    RayS r; r.dx = 0.3f + s; r.dy = 0.5f - s; r.dz = 0.8f + s; r.ox = 2.f * s; r.oy = 2.f + s; r.oz = -6.f - s; r.nod = 1.f + s; r.noo = -10.f + s;
    return r;
}
__device__ __forceinline__ void filter_one(float4 const c, const RayS &r, uint32_t &mask)
{
    float const tc = __fmaf_rn(c.z, r.dz, __fmaf_rn(c.y, r.dy, __fmaf_rn(c.x, r.dx, r.nod)));
    float const na = __fmaf_rn(c.z, r.oz, __fmaf_rn(c.y, r.oy, __fmaf_rn(c.x, r.ox, c.w)));
    float const v = __fadd_rn(__fmaf_rn(tc, tc, na), r.noo);
    mask = __funnelshift_l(__float_as_uint(fminf(v, tc)), mask, 1);
}
template <typename P> __device__ __forceinline__ uint32_t sweep_s(P sph, uint32_t nWords, const RayS &r)
{
    uint32_t acc = 0;
#pragma unroll 1
    for (uint32_t w = 0; w < nWords; ++w)
    {
        uint32_t mask = 0xFFFFFFFFu;
#pragma unroll
        for (uint32_t p = 0; p < 32; ++p)
            filter_one(sph[w * 32 + p], r, mask);
        acc += __popc(~mask);
    }
    return acc;
}
__global__ void __launch_bounds__(256, 4) k_smem_s(const FilterPair *g, int nWords, int reps, uint32_t *out, float nz)
{
    extern __shared__ __align__(16) unsigned char raw[];
    float4 *s = reinterpret_cast<float4 *>(raw);
    for (int i = threadIdx.x; i < nWords * 32; i += blockDim.x) s[i] = reinterpret_cast<const float4 *>(g)[i];
    __syncthreads();
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep_s(s, nWords, make_ray_s(1e-3f * float(threadIdx.x + k)));
    if (acc == 0xdeadbeef) out[0] = acc;
}
__global__ void __launch_bounds__(256, 4) k_cmem_s(int nWords, int reps, uint32_t *out, float nz)
{
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep_s(cSph, nWords, make_ray_s(1e-3f * float(threadIdx.x + k)));
    if (acc == 0xdeadbeef) out[0] = acc;
}
// no load instructions at all: 5 words fully unrolled, every record a c[0x3][imm] operand of the FFMA
__global__ void __launch_bounds__(256, 4) k_imm_s(int nWords, int reps, uint32_t *out, float nz)
{
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k)
    {
        RayS const r = make_ray_s(1e-3f * float(threadIdx.x + k));
#pragma unroll
        for (uint32_t w = 0; w < 5; ++w)
        {
            uint32_t mask = 0xFFFFFFFFu;
#pragma unroll
            for (uint32_t p = 0; p < 32; ++p)
                filter_one(cSph[w * 32 + p], r, mask);
            acc += __popc(~mask);
        }
    }
    if (acc == 0xdeadbeef) out[0] = acc;
}

__global__ void __launch_bounds__(256, 4) k_smem(const FilterPair *g, int nWords, int reps, uint32_t *out, float nz)
{
    extern __shared__ __align__(16) unsigned char raw[];
    FilterPair *s = reinterpret_cast<FilterPair *>(raw);
    for (int i = threadIdx.x; i < nWords * 16; i += blockDim.x) s[i] = g[i];
    __syncthreads();
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep(s, nWords, make_ray(1e-3f * float(threadIdx.x + k), 0.f));
    if (acc == 0xdeadbeef) out[0] = acc;
}
__global__ void __launch_bounds__(256, 4) k_cmem(int nWords, int reps, uint32_t *out, float nz)
{
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep(cRecs, nWords, make_ray(1e-3f * float(threadIdx.x + k), 0.f));
    if (acc == 0xdeadbeef) out[0] = acc;
}
__global__ void __launch_bounds__(256) k_cmem_u(int nWords, int reps, uint32_t *out, float nz) // no occupancy target: ptxas picks LDCU
{
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep(cRecs, nWords, make_ray(1e-3f * float(threadIdx.x + k), nz));
    if (acc == 0xdeadbeef) out[0] = acc;
}
template <int CAP>
__global__ void __launch_bounds__(256) k_param_u(const __grid_constant__ ParamRecs<CAP> pr, int nWords, int reps, uint32_t *out, float nz)
{
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep(pr.r, nWords, make_ray(1e-3f * float(threadIdx.x + k), nz));
    if (acc == 0xdeadbeef) out[0] = acc;
}
template <int CAP>
__global__ void __launch_bounds__(256, 4) k_param(const __grid_constant__ ParamRecs<CAP> pr, int nWords, int reps, uint32_t *out, float nz)
{
    uint32_t acc = 0;
    for (int k = 0; k < reps; ++k) acc += sweep(pr.r, nWords, make_ray(1e-3f * float(threadIdx.x + k), 0.f));
    if (acc == 0xdeadbeef) out[0] = acc;
}

int main()
{
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    int const sms = prop.multiProcessorCount;
    double const ghz = 1.965; // B200 max SM clock; the ratio between variants is what matters
    std::vector<FilterPair> h(kMaxPairs);
    for (int i = 0; i < kMaxPairs; ++i)
        for (int k = 0; k < 2; ++k) { h[i].cx[k] = float(i % 37) - 18.f; h[i].cy[k] = 0.4f; h[i].cz[k] = float(i % 23); h[i].nk[k] = -float(i % 91); }
    FilterPair *d; cudaMalloc(&d, sizeof(FilterPair) * kMaxPairs); cudaMemcpy(d, h.data(), sizeof(FilterPair) * kMaxPairs, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(cRecs, h.data(), sizeof(FilterPair) * kMaxPairs);
    cudaMemcpyToSymbol(cSph, h.data(), sizeof(FilterPair) * kMaxPairs);
    uint32_t *out; cudaMalloc(&out, 4);
    cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    static ParamRecs<96> p96; static ParamRecs<896> p896;
    for (int i = 0; i < 96; ++i) p96.r[i] = h[i];
    for (int i = 0; i < 896; ++i) p896.r[i] = h[i];
    int const blocks = sms * 4;
    float const nz = -0.0f;
    for (int nWords : {5, 6, 16, 56})
    {
        int const reps = 4000 * 5 / nWords;
        for (int variant = 0; variant < 8; ++variant)
        {
            float best = 1e30f;
            for (int it = 0; it < 3; ++it)
            {
                cudaEventRecord(e0);
                if (variant == 0) k_smem<<<blocks, 256, nWords * 16 * sizeof(FilterPair)>>>(d, nWords, reps, out, nz);
                else if (variant == 1) k_cmem<<<blocks, 256>>>(nWords, reps, out, nz);
                else if (variant == 2 && nWords <= 6) k_param<96><<<blocks, 256>>>(p96, nWords, reps, out, nz);
                else if (variant == 2) k_param<896><<<blocks, 256>>>(p896, nWords, reps, out, nz);
                else if (variant == 5) k_smem_s<<<blocks, 256, nWords * 16 * sizeof(FilterPair)>>>(d, nWords, reps, out, nz);
                else if (variant == 6) k_cmem_s<<<blocks, 256>>>(nWords, reps, out, nz);
                else if (variant == 7) { if (nWords == 5) k_imm_s<<<blocks, 256>>>(nWords, reps, out, nz); }
                else if (variant == 3) k_cmem_u<<<blocks, 256>>>(nWords, reps, out, nz);
                else if (nWords <= 6) k_param_u<96><<<blocks, 256>>>(p96, nWords, reps, out, nz);
                else k_param_u<896><<<blocks, 256>>>(p896, nWords, reps, out, nz);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
            }
            cudaError_t err = cudaGetLastError();
            // per SMSP: 8 warps x reps x nWords x 16 records
            double const recs = 8.0 * reps * nWords * 16;
            printf("%-12s nWords %2d  %8.3f ms  %6.2f cycles/record/SMSP (bound 16)  %s\n", variant == 0 ? "smem" : variant == 1 ? "cmem(LDC)" : variant == 2 ? "param(LDC)" : variant == 3 ? "cmem(LDCU)" : variant == 4 ? "param(LDCU)" : variant == 5 ? "smem scalar" : variant == 6 ? "cmem scalar" : "imm scalar",
                   nWords, best, best * 1e-3 * ghz * 1e9 / recs, err == cudaSuccess ? "" : cudaGetErrorString(err));
        }
    }
    return 0;
}
