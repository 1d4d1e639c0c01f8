// mma_tf32_microbench.cu -- what the legacy warp-level tensor path (mma.sync m16n8k8 tf32, SASS HMMA) delivers on sm_100a,
// alone and next to CUDA-core work, and how it rounds.  Sized for the question "could the two dot products of the
// closest-hit pre-filter (C.D and C.O, K = 3) run as small GEMMs on the tensor pipe while the issue port does the rest?".
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_tf32_microbench tools/mma_tf32_microbench.cu && ./mma_tf32_microbench
//
// Prints: cycles per mma per sub-partition for 1..8 independent accumulator chains and 4 / 8 warps per sub-partition;
// the same with FFMA work interleaved (does the mma hold the issue port?); and the worst error of a K = 16 hi/lo-split
// dot product against binary64, relative to sum |a_i b_i|.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int CHAINS, int FFMA_PER_MMA>
__global__ void __launch_bounds__(256) mma_rate(float *out, uint32_t iters, float x)
{
    float d[CHAINS][4];
    uint32_t a[4], b[2];
    for (int k = 0; k < 4; ++k)
        a[k] = __float_as_uint(x + float(threadIdx.x + k));
    b[0] = __float_as_uint(x * 0.5f);
    b[1] = __float_as_uint(x * 0.25f);
    for (int c = 0; c < CHAINS; ++c)
        for (int k = 0; k < 4; ++k)
            d[c][k] = float(c + k);
    float f[8];
    for (int k = 0; k < 8; ++k)
        f[k] = float(threadIdx.x + k);
    for (uint32_t i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c)
        {
            mma_tf32(d[c], a, b);
#pragma unroll
            for (int k = 0; k < FFMA_PER_MMA; ++k)
                f[k & 7] = __fmaf_rn(f[k & 7], x, 1e-9f);
        }
    }
    float s = 0.f;
    for (int c = 0; c < CHAINS; ++c)
        for (int k = 0; k < 4; ++k)
            s += d[c][k];
    for (int k = 0; k < 8; ++k)
        s += f[k];
    if (s == 123.456f)
        out[0] = s;
}

// accuracy: D[16x8] = A[16x16] * B[16x8] in two k-steps; A rows and B columns hold hi/lo-split operands prepared on the host
__global__ void mma_accuracy(const float *A, const float *B, float *D, int tiles)
{
    int const lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    int const tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (tile >= tiles)
        return;
    const float *a = A + size_t(tile) * 16 * 16; // row-major [16][16]
    const float *b = B + size_t(tile) * 16 * 8;  // [k][n] row-major [16][8]
    float d[4] = {0.f, 0.f, 0.f, 0.f};
    for (int ks = 0; ks < 2; ++ks)
    {
        uint32_t af[4], bf[2];
        af[0] = __float_as_uint(a[g * 16 + ks * 8 + t]);
        af[1] = __float_as_uint(a[(g + 8) * 16 + ks * 8 + t]);
        af[2] = __float_as_uint(a[g * 16 + ks * 8 + t + 4]);
        af[3] = __float_as_uint(a[(g + 8) * 16 + ks * 8 + t + 4]);
        bf[0] = __float_as_uint(b[(ks * 8 + t) * 8 + g]);
        bf[1] = __float_as_uint(b[(ks * 8 + t + 4) * 8 + g]);
        mma_tf32(d, af, bf);
    }
    float *o = D + size_t(tile) * 16 * 8;
    o[g * 8 + 2 * t] = d[0];
    o[g * 8 + 2 * t + 1] = d[1];
    o[(g + 8) * 8 + 2 * t] = d[2];
    o[(g + 8) * 8 + 2 * t + 1] = d[3];
}

static float tf32_trunc(float v)
{
    uint32_t u;
    memcpy(&u, &v, 4);
    u &= 0xFFFFE000u;
    memcpy(&v, &u, 4);
    return v;
}

template <int CHAINS, int FF>
static void run_rate(int warpsPerSm, float *dOut, int sms, double clockGhz)
{
    int const threads = 256, blocksPerSm = warpsPerSm * 32 / threads;
    uint32_t const iters = 1u << 14;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep)
    {
        cudaEventRecord(e0);
        mma_rate<CHAINS, FF><<<sms * blocksPerSm, threads>>>(dOut, iters, 1.0000001f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best)
            best = ms;
    }
    double const mmaPerSubpart = double(iters) * CHAINS * (warpsPerSm / 4.0);
    double const cycles = best * 1e-3 * clockGhz * 1e9;
    printf("chains %d  ffma/mma %2d  warps/SM %2d : %7.2f cycles per mma per sub-partition  (%.1f TF32 TFLOP/s, + %.1f FP32 TFLOP/s)\n", CHAINS, FF,
           warpsPerSm, cycles / mmaPerSubpart, double(iters) * CHAINS * warpsPerSm * sms * 2048.0 / (best * 1e-3) / 1e12,
           double(iters) * CHAINS * FF * warpsPerSm * sms * 64.0 / (best * 1e-3) / 1e12);
}

int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double const ghz = khz * 1e-6;
    printf("%s, %d SMs, %.3f GHz\n", prop.name, prop.multiProcessorCount, ghz);
    float *dOut;
    cudaMalloc(&dOut, 256);
    int const sms = prop.multiProcessorCount;
    run_rate<1, 0>(32, dOut, sms, ghz);
    run_rate<2, 0>(32, dOut, sms, ghz);
    run_rate<4, 0>(32, dOut, sms, ghz);
    run_rate<8, 0>(32, dOut, sms, ghz);
    run_rate<4, 0>(16, dOut, sms, ghz);
    run_rate<4, 2>(32, dOut, sms, ghz);
    run_rate<4, 4>(32, dOut, sms, ghz);
    run_rate<4, 8>(32, dOut, sms, ghz);
    run_rate<4, 16>(32, dOut, sms, ghz);

    // ---- accuracy of a hi/lo-split K = 3 dot product + constant: terms ah*bh, ah*bl, al*bh, al*bl per component, seed hi/lo
    int const tiles = 4096;
    std::vector<float> A(size_t(tiles) * 256), B(size_t(tiles) * 128), D(size_t(tiles) * 128);
    std::vector<double> ref(size_t(tiles) * 128), mag(size_t(tiles) * 128);
    srand(7);
    auto rnd = [](double s) { return float((rand() / double(RAND_MAX) * 2.0 - 1.0) * s); };
    for (int tIdx = 0; tIdx < tiles; ++tIdx)
    {
        float av[16][3], as[16], bv[8][3];
        for (int r = 0; r < 16; ++r)
        {
            for (int c = 0; c < 3; ++c)
                av[r][c] = rnd(40.0);
            as[r] = rnd(900.0);
        }
        for (int n = 0; n < 8; ++n)
            for (int c = 0; c < 3; ++c)
                bv[n][c] = rnd(1.0);
        for (int r = 0; r < 16; ++r)
        {
            float *row = &A[(size_t(tIdx) * 16 + r) * 16];
            for (int c = 0; c < 3; ++c)
            {
                float const h = tf32_trunc(av[r][c]), l = tf32_trunc(av[r][c] - h);
                row[c] = h;      // x bh
                row[3 + c] = h;  // x bl
                row[6 + c] = l;  // x bh
                row[9 + c] = l;  // x bl
            }
            float const sh = tf32_trunc(as[r]), sl = tf32_trunc(as[r] - sh);
            row[12] = sh;
            row[13] = sl;
            row[14] = row[15] = 0.f;
        }
        for (int n = 0; n < 8; ++n)
        {
            float *colBase = &B[size_t(tIdx) * 128];
            for (int c = 0; c < 3; ++c)
            {
                float const h = tf32_trunc(bv[n][c]), l = tf32_trunc(bv[n][c] - h);
                colBase[(c)*8 + n] = h;
                colBase[(3 + c) * 8 + n] = l;
                colBase[(6 + c) * 8 + n] = h;
                colBase[(9 + c) * 8 + n] = l;
            }
            colBase[12 * 8 + n] = 1.f;
            colBase[13 * 8 + n] = 1.f;
            colBase[14 * 8 + n] = colBase[15 * 8 + n] = 0.f;
        }
        for (int r = 0; r < 16; ++r)
            for (int n = 0; n < 8; ++n)
            {
                double s = as[r], m = std::fabs(double(as[r]));
                for (int c = 0; c < 3; ++c)
                {
                    s += double(av[r][c]) * double(bv[n][c]);
                    m += std::fabs(double(av[r][c]) * double(bv[n][c]));
                }
                ref[(size_t(tIdx) * 16 + r) * 8 + n] = s;
                mag[(size_t(tIdx) * 16 + r) * 8 + n] = m;
            }
    }
    float *dA, *dB, *dD;
    cudaMalloc(&dA, A.size() * 4);
    cudaMalloc(&dB, B.size() * 4);
    cudaMalloc(&dD, D.size() * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    mma_accuracy<<<tiles / 8, 256>>>(dA, dB, dD, tiles);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double worst = 0.0, worstAbs = 0.0;
    for (size_t i = 0; i < D.size(); ++i)
    {
        double const e = std::fabs(double(D[i]) - ref[i]);
        worst = std::max(worst, e / mag[i]);
        worstAbs = std::max(worstAbs, e);
    }
    printf("hi/lo split (4 terms per component) K=3 dot + constant: worst |err| / sum|terms| = %.3g = 2^%.2f, worst abs %.3g (%s)\n", worst,
           std::log2(worst), worstAbs, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
