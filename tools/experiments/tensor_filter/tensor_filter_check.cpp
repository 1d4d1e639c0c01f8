// tensor_filter_check.cpp -- CPU check of the host half and the arithmetic of the EXPERIMENTAL tensor-core pre-filter
// (simplepathtracer_b200/csrc/spt_tensor_filter.cuh, spt_tensor_host.hpp).  No GPU, no oracle.
//
//   g++ -O2 -std=c++17 -frounding-math -ffp-contract=off -o /tmp/tfc tools/tensor_filter_check.cpp && /tmp/tfc scene.bin [rays]
//
// With `cam ex ey ez` appended: the camera form (primary_hits_tc): all rays from the eye, build_tensor_matrix_cam, K = 32.
// scene.bin: uint32 n, then n x (cx, cy, cz, r) float32 (tests/test_tensor_filter_host.py writes it from the golden scene).
//
// What it does, for `rays` secondary-like rays (origins on sphere surfaces and on the ground, directions normalised in fp32):
//   * builds the sphere matrix with the product's own builder (build_tensor_matrix) in the UMMA canonical layout;
//   * writes every ray's row with the SAME fp32 operations and the SAME store addressing as tensor_store_ray on the device
//     (directed roundings through fesetround), into a canonical 128-row tile;
//   * multiplies rows through the canonical-layout index function (so the device's store formula and the layout formula must
//     agree), operands truncated to TF32 as the tensor core does, under two accumulator models that bracket the hardware:
//     fp32 truncating after every addition in K order (pessimistic) and binary64 (optimistic);
//   * runs the reference's separately rounded sequence (exact_sphere: Collision.hpp:9-17) on every pair, and FAILS (exit 1) if a
//     pair that sequence accepts (tCenter > 1e-3 and disc > 1e-3) has a negative tensor discriminant under either model;
//   * FAILS if the worst deviation from (real discriminant + slack - 1e-3) exceeds 300 u (A^2 + B^2 + r^2) -- what the 384 u slack
//     leaves after the exact sequence's own error -- or if the candidate rate is out of proportion;
//   * reports candidates per ray next to the shipped FFMA2 filter's slack (96 u) evaluated in binary64, for the cost side.
#include "../simplepathtracer_b200/csrc/spt_tensor_host.hpp"

#include <cfenv>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>

using namespace spt;

namespace
{
// volatile keeps the compiler from folding across fesetround
float mul_r(float a, float b, int mode)
{
    volatile float x = a, y = b;
    std::fesetround(mode);
    volatile float r = x * y;
    std::fesetround(FE_TONEAREST);
    return r;
}
float add_r(float a, float b, int mode)
{
    volatile float x = a, y = b;
    std::fesetround(mode);
    volatile float r = x + y;
    std::fesetround(FE_TONEAREST);
    return r;
}
float fma_r(float a, float b, float c, int mode)
{
    volatile float x = a, y = b, z = c;
    std::fesetround(mode);
    volatile float r = std::fmaf(x, y, z);
    std::fesetround(FE_TONEAREST);
    return r;
}
float fmul(float a, float b)
{
    volatile float r = a * b;
    return r;
}
float fadd(float a, float b)
{
    volatile float r = a + b;
    return r;
}
float fsub(float a, float b)
{
    volatile float r = a - b;
    return r;
}
float dot3(const float *a, const float *b) { return fadd(fadd(fmul(a[0], b[0]), fmul(a[1], b[1])), fmul(a[2], b[2])); }

// tensor_store_ray, operation for operation
void store_ray(float *sA, uint32_t lane, const float *dir, const float *org, float od)
{
    float const ooLo = fma_r(org[2], org[2], fma_r(org[1], org[1], mul_r(org[0], org[0], FE_DOWNWARD), FE_DOWNWARD), FE_DOWNWARD);
    float const ooUp = fma_r(org[2], org[2], fma_r(org[1], org[1], mul_r(org[0], org[0], FE_UPWARD), FE_UPWARD), FE_UPWARD);
    float const q = add_r(add_r(mul_r(od, od, FE_UPWARD), -ooLo, FE_UPWARD), mul_r(ooUp, float(kTensorSlackU) * 1.0001f, FE_UPWARD), FE_UPWARD);
    float const od2 = fmul(2.f, od);
    float const f[10] = {fmul(dir[0], dir[0]), fmul(dir[0], dir[1]), fmul(dir[0], dir[2]), fmul(dir[1], dir[1]), fmul(dir[1], dir[2]),
                         fmul(dir[2], dir[2]), fsub(fmul(2.f, org[0]), fmul(od2, dir[0])), fsub(fmul(2.f, org[1]), fmul(od2, dir[1])),
                         fsub(fmul(2.f, org[2]), fmul(od2, dir[2])), q};
    float *row = sA + (lane >> 3) * 32 + (lane & 7) * 4;
    for (int c = 0; c < 10; ++c)
    {
        float const h = tf32_trunc(f[c]), l = fsub(f[c], h);
        float *d = row + c * (kTensorALbo / 4);
        if (c < 9)
            d[0] = h, d[1] = l, d[2] = h, d[3] = l;
        else
            d[0] = 1.f, d[1] = 1.f, d[2] = h, d[3] = l;
    }
}

// tensor_store_ray_cam + tensor_init_rows_cam, operation for operation
void store_ray_cam(float *sA, uint32_t lane, const float *dir)
{
    float const f[6] = {fmul(dir[0], dir[0]), fmul(dir[0], dir[1]), fmul(dir[0], dir[2]), fmul(dir[1], dir[1]), fmul(dir[1], dir[2]), fmul(dir[2], dir[2])};
    float *row = sA + (lane >> 3) * 32 + (lane & 7) * 4;
    for (int c = 0; c < 6; ++c)
    {
        float const h = tf32_trunc(f[c]), l = fsub(f[c], h);
        float *d = row + c * (kTensorALbo / 4);
        d[0] = h, d[1] = l, d[2] = h, d[3] = l;
    }
    float *d6 = row + 6 * (kTensorALbo / 4), *d7 = row + 7 * (kTensorALbo / 4);
    d6[0] = d6[1] = 1.f, d6[2] = d6[3] = 0.f;
    d7[0] = d7[1] = d7[2] = d7[3] = 0.f;
}

float trunc_to_float(double x) // binary64 -> fp32 toward zero
{
    float f = float(x);
    if (std::fabs(double(f)) > std::fabs(x))
        f = std::nextafterf(f, 0.f);
    return f;
}
} // namespace

int main(int argc, char **argv)
{
    if (argc < 2)
    {
        std::fprintf(stderr, "usage: %s scene.bin [rays [cam ex ey ez]]\n", argv[0]);
        return 2;
    }
    FILE *fp = std::fopen(argv[1], "rb");
    uint32_t n = 0;
    if (!fp || std::fread(&n, 4, 1, fp) != 1)
        return 2;
    std::vector<float> geom(size_t(n) * 4);
    if (std::fread(geom.data(), 16, n, fp) != n)
        return 2;
    std::fclose(fp);
    uint32_t const nRays = argc > 2 ? uint32_t(std::atoi(argv[2])) : 4096u;
    // camera form: every ray starts at the eye, matrix from build_tensor_matrix_cam, K = 32
    bool const cam = argc > 6 && std::string(argv[3]) == "cam";
    double const eyeD[3] = {cam ? std::atof(argv[4]) : 0.0, cam ? std::atof(argv[5]) : 0.0, cam ? std::atof(argv[6]) : 0.0};
    float const eyeF[3] = {float(eyeD[0]), float(eyeD[1]), float(eyeD[2])};
    double const eyeX[3] = {double(eyeF[0]), double(eyeF[1]), double(eyeF[2])}; // the library sees the fp32 eye
    uint32_t const K = cam ? kTensorKCam : kTensorK;

    // filter slots: the rule of build_filter (spt_cuda.cu)
    std::vector<uint32_t> slotIdx;
    for (uint32_t i = 0; i < n; ++i)
    {
        double const cx = geom[4 * i], cy = geom[4 * i + 1], cz = geom[4 * i + 2], r = geom[4 * i + 3];
        if (std::fabs(r) <= 64.0 && std::sqrt(cx * cx + cy * cy + cz * cz) <= 4096.0)
            slotIdx.push_back(i);
    }
    uint32_t const nSlots = uint32_t((slotIdx.size() + 7) / 8 * 8);
    std::vector<double> hostGeom(size_t(nSlots) * 4, -1.0);
    for (uint32_t k = 0; k < slotIdx.size(); ++k)
    {
        for (int e = 0; e < 3; ++e)
            hostGeom[4 * k + e] = geom[4 * slotIdx[k] + e];
        hostGeom[4 * k + 3] = std::fabs(double(geom[4 * slotIdx[k] + 3]));
    }
    std::vector<float> B;
    if (cam)
        build_tensor_matrix_cam(hostGeom, nSlots, eyeX, B);
    else
        build_tensor_matrix(hostGeom, nSlots, B);
    uint32_t const rows = tensor_rows(nSlots), used = tensor_slots(nSlots);
    if (B.size() != size_t(rows) * K || rows % 16u || rows > kTensorCols || used > rows)
    {
        std::printf("FAIL: matrix shape\n");
        return 1;
    }
    // every sphere entry must already be a TF32 value, and the constant column must not round below its binary64 value
    for (uint32_t s = 0; s < rows; ++s)
        for (uint32_t k = 0; k < K; ++k)
            if (tf32_trunc(B[tensor_elem(rows, s, k)]) != B[tensor_elem(rows, s, k)])
            {
                std::printf("FAIL: B(%u,%u) is not a TF32 value\n", s, k);
                return 1;
            }
    for (uint32_t s = 0; s < used && !cam; ++s)
    {
        if (hostGeom[4 * s + 3] < 0.0)
            continue;
        double const cx = hostGeom[4 * s], cy = hostGeom[4 * s + 1], cz = hostGeom[4 * s + 2], r = hostGeom[4 * s + 3];
        double const a2 = cx * cx + cy * cy + cz * cz, k = -(a2 - r * r) + kTensorSlackU * (a2 + r * r) - double(1e-3f);
        double const got = double(B[tensor_elem(rows, s, 36)]) + double(B[tensor_elem(rows, s, 37)]);
        if (got < k || B[tensor_elem(rows, s, 38)] != 1.f || B[tensor_elem(rows, s, 39)] != 1.f)
        {
            std::printf("FAIL: constant column of slot %u\n", s);
            return 1;
        }
    }

    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<double> uni(0.0, 1.0);
    std::normal_distribution<double> gauss(0.0, 1.0);
    std::vector<float> A(size_t(kTensorRays) * kTensorK);
    unsigned long long pairs = 0, hits = 0, missedRz = 0, missedF64 = 0, candRz = 0, candF64 = 0, candShipped = 0;
    double worstRz = 0.0, worstF64 = 0.0; // |v_model - v_real| in units of u (A^2 + B^2 + r^2)
    for (uint32_t r0 = 0; r0 < nRays; r0 += kTensorRays)
    {
        float dirs[kTensorRays][3], orgs[kTensorRays][3], ods[kTensorRays];
        for (uint32_t lane = 0; lane < kTensorRays; ++lane)
        {
            // origin: on a filterable sphere's surface, or on the ground plane out to +-25; direction: normal + shell vector
            double nrm[3] = {gauss(rng), gauss(rng), gauss(rng)};
            double const nl = std::sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
            for (double &c : nrm)
                c /= nl;
            double o[3];
            if (uni(rng) < 0.5 || slotIdx.empty())
            {
                o[0] = -25.0 + 50.0 * uni(rng), o[1] = 0.0, o[2] = -25.0 + 50.0 * uni(rng);
                nrm[0] = 0.0, nrm[1] = 1.0, nrm[2] = 0.0;
            }
            else
            {
                uint32_t const k = uint32_t(uni(rng) * double(slotIdx.size())) % uint32_t(slotIdx.size());
                for (int e = 0; e < 3; ++e)
                    o[e] = hostGeom[4 * k + e] + nrm[e] * hostGeom[4 * k + 3];
            }
            if (lane % 16u == 0u) // some far-away origins: the slack has to scale with |O|^2
                for (double &c : o)
                    c *= 40.0;
            double d[3] = {nrm[0] + uni(rng) - 0.5, nrm[1] + uni(rng) - 0.5, nrm[2] + uni(rng) - 0.5};
            if (cam) // a pinhole at the eye looking down +z, 90 degrees wide as the reference's (SingleThreadPathTracer.hpp:125-128)
            {
                for (int e = 0; e < 3; ++e)
                    o[e] = eyeX[e];
                d[0] = -1.0 + 2.0 * uni(rng), d[1] = -1.0 + 2.0 * uni(rng), d[2] = 1.0;
            }
            if (lane % 5u == 0u) // and some rays aimed at a sphere centre (grazing and head-on hits both occur)
            {
                uint32_t const k = uint32_t(uni(rng) * double(slotIdx.size())) % uint32_t(slotIdx.size());
                for (int e = 0; e < 3; ++e)
                    d[e] = hostGeom[4 * k + e] - o[e] + (uni(rng) - 0.5) * 2.0 * hostGeom[4 * k + 3];
            }
            float df[3] = {float(d[0]), float(d[1]), float(d[2])};
            volatile float len = std::sqrt(dot3(df, df));
            for (int e = 0; e < 3; ++e)
            {
                volatile float q = df[e] / len;
                dirs[lane][e] = q;
                orgs[lane][e] = float(o[e]);
            }
            ods[lane] = dot3(orgs[lane], dirs[lane]);
            if (cam)
                store_ray_cam(A.data(), lane, dirs[lane]);
            else
                store_ray(A.data(), lane, dirs[lane], orgs[lane], ods[lane]);
        }
        for (uint32_t lane = 0; lane < kTensorRays; ++lane)
        {
            const float *D = dirs[lane], *O = orgs[lane];
            double const b2 = double(O[0]) * O[0] + double(O[1]) * O[1] + double(O[2]) * O[2];
            for (uint32_t s = 0; s < used; ++s)
            {
                if (hostGeom[4 * s + 3] < 0.0)
                    continue;
                // the tensor core: operands truncated to TF32, exact products
                float accRz = 0.f;
                double accF64 = 0.0;
                for (uint32_t k = 0; k < K; ++k)
                {
                    double const prod = double(tf32_trunc(A[tensor_elem(kTensorRays, lane, k)])) * double(B[tensor_elem(rows, s, k)]);
                    accRz = trunc_to_float(double(accRz) + prod);
                    accF64 += prod;
                }
                // the reference's sequence (exact_sphere)
                uint32_t const i = slotIdx[s];
                float const C[3] = {geom[4 * i], geom[4 * i + 1], geom[4 * i + 2]}, rad = geom[4 * i + 3];
                float const L[3] = {fsub(C[0], O[0]), fsub(C[1], O[1]), fsub(C[2], O[2])};
                float const tCenter = dot3(L, D);
                float const disc = fsub(fmul(rad, rad), fsub(dot3(L, L), fmul(tCenter, tCenter)));
                bool const hit = tCenter > 1e-3f && disc > 1e-3f;
                // the real discriminant with the slack the filter is entitled to, for the error statistics
                double const cd = double(C[0]) * D[0] + double(C[1]) * D[1] + double(C[2]) * D[2];
                double const odx = double(O[0]) * D[0] + double(O[1]) * D[1] + double(O[2]) * D[2];
                double const co = double(C[0]) * O[0] + double(C[1]) * O[1] + double(C[2]) * O[2];
                double const a2 = double(C[0]) * C[0] + double(C[1]) * C[1] + double(C[2]) * C[2], r2 = double(rad) * rad;
                double const real = (cd - odx) * (cd - odx) - (a2 - r2) + 2.0 * co - b2;
                double const scale = (a2 + b2 + r2) / 16777216.0;
                double const target = real + kTensorSlackU * (a2 + b2 + r2) - double(1e-3f);
                worstRz = std::max(worstRz, std::fabs(double(accRz) - target) / scale);
                worstF64 = std::max(worstF64, std::fabs(accF64 - target) / scale);
                ++pairs;
                hits += hit;
                missedRz += hit && std::signbit(accRz);
                missedF64 += hit && accF64 < 0.0;
                candRz += !std::signbit(accRz);
                candF64 += !(accF64 < 0.0);
                candShipped += real + (96.0 / 16777216.0) * (a2 + b2 + r2) - double(1e-3f) >= 0.0;
            }
        }
    }
    double const rays = double((nRays + kTensorRays - 1) / kTensorRays * kTensorRays);
    if (cam)
        std::printf("camera form, eye (%g, %g, %g), K = %u\n", eyeX[0], eyeX[1], eyeX[2], K);
    std::printf("%u spheres, %u slots (%u on the tensor core, %u matrix rows), %.0f rays, %llu pairs\n", n, nSlots, used, rows, rays, pairs);
    std::printf("accepted by the reference's sequence: %.3f per ray\n", double(hits) / rays);
    std::printf("candidates per ray: tensor form %.3f (fp32 truncating accumulate) / %.3f (binary64 accumulate); shipped slack 96 u, exact arithmetic: %.3f\n",
                double(candRz) / rays, double(candF64) / rays, double(candShipped) / rays);
    std::printf("worst |v - (real + slack - 1e-3)|: %.1f u(A^2+B^2+r^2) truncating, %.1f u binary64 (slack: %.0f u)\n", worstRz, worstF64,
                kTensorSlackU * 16777216.0);
    std::printf("MISSED: %llu truncating, %llu binary64\n", missedRz, missedF64);
    // soundness needs  target - v <= slack - (the exact sequence's own error, <= 33.4 u (A^2 + B^2) + 2 u r^2: spt_filter.cuh);
    // usefulness needs the filter to stay a filter
    bool const bounded = worstRz <= 300.0 && worstF64 <= 300.0;
    bool const selective = double(candRz) <= 2.0 * double(candShipped) + rays;
    if (!bounded)
        std::printf("error above 300 u: the slack no longer covers it\n");
    if (!selective)
        std::printf("the tensor form passes far more candidates than the shipped filter\n");
    if (missedRz || missedF64 || hits == 0 || !bounded || !selective)
    {
        std::printf("FAIL\n");
        return 1;
    }
    std::printf("OK\n");
    return 0;
}
