// issue_slots_microbench.cu -- does an FFMA2 (fma.rn.f32x2) cost one issue slot or two on sm_100a?
// Per loop iteration: 8 independent FFMA2 (or 16 FFMA) plus M independent integer ALU ops (SHF funnel shifts).
// If FFMA2 took one slot and two FMA-pipe cycles, M <= 8 ALU ops would hide in its shadow (16 cycles/iteration);
// if it blocks the issue port for both cycles, time grows as 16 + M.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o issue_slots_microbench issue_slots_microbench.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int M, bool PACKED>
__global__ void __launch_bounds__(256) k(float b, float c, uint32_t iters, uint32_t *out)
{
    unsigned long long a[8], bb, cc;
    float f[16];
    uint32_t x[8];
    asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b), "f"(b + 1e-7f));
    asm("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(c), "f"(c + 1e-7f));
#pragma unroll
    for (int i = 0; i < 8; ++i) { float v = float(threadIdx.x + i); asm("mov.b64 %0, {%1, %2};" : "=l"(a[i]) : "f"(v), "f"(v + 1.f)); x[i] = threadIdx.x * 7u + i; }
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = float(threadIdx.x + i);
    uint32_t y = threadIdx.x | 1u;
    for (uint32_t it = 0; it < iters; ++it)
    {
        if (PACKED)
        {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(bb), "l"(cc));
        }
        else
        {
#pragma unroll
            for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(b), "f"(c));
        }
#pragma unroll
        for (int j = 0; j < M; ++j) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(x[j & 7]) : "r"(y));
    }
    float s = 0.f; uint32_t xs = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[i])); s += lo + hi; xs ^= x[i]; }
#pragma unroll
    for (int i = 0; i < 16; ++i) s += f[i];
    if (s == 123.456f || xs == 0x12345u) out[0] = xs;
}

template <int M, bool PACKED> void run(int sms, uint32_t *out)
{
    uint32_t const iters = 1u << 16;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 3; ++r)
    {
        cudaEventRecord(e0);
        k<M, PACKED><<<sms * 8, 256>>>(1.0000001f, 1e-9f, iters, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
    }
    // 16 warps per SMSP, each `iters` iterations
    printf("%s M=%2d  %7.3f ms  %6.2f cycles/iteration/SMSP\n", PACKED ? "8 FFMA2 " : "16 FFMA ", M, best, best * 1e-3 * 1.965e9 / (16.0 * iters));
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    uint32_t *out; cudaMalloc(&out, 4);
    run<0, true>(p.multiProcessorCount, out); run<2, true>(p.multiProcessorCount, out); run<4, true>(p.multiProcessorCount, out);
    run<6, true>(p.multiProcessorCount, out); run<8, true>(p.multiProcessorCount, out); run<16, true>(p.multiProcessorCount, out);
    run<0, false>(p.multiProcessorCount, out); run<4, false>(p.multiProcessorCount, out); run<8, false>(p.multiProcessorCount, out);
    run<16, false>(p.multiProcessorCount, out);
    return 0;
}
