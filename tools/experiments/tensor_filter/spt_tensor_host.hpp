// spt_tensor_host.hpp -- host half of the EXPERIMENTAL tensor-core pre-filter (spt_tensor_filter.cuh; -DSPT_TENSOR_FILTER=1 builds
// only): constants, the UMMA canonical layout and the sphere matrix.  Plain C++ (no CUDA) so that tools/tensor_filter_check.cpp can
// run the same builder on the CPU against an emulation of the ray rows and of the TF32 accumulate.
//
//   v = (C.D - O.D)^2 - (|C|^2 - r^2) + 2 C.O - |O|^2 + S_i + S_ray - 1e-3
//     = sum_{i<=j} (c_ij C_i C_j)(D_i D_j) + sum_i C_i W_i + k + q,     W = 2 O - 2 (O.D) D,  q = (O.D)^2 - |O|^2 + S_ray.
// Ten factors per side, each split into a TF32 head and tail with all four cross products kept: K = 40.  Chunk c (four K slots =
// 16 bytes) of a ray's row is (h, l, h, l) of its factor c, of a sphere's row (h, h, l, l); the last chunk carries the constants:
// ray (1, 1, q_h, q_l), sphere (k_h, k_l, 1, 1).  Rows sit in the K-major no-swizzle canonical form of tcgen05.mma:
//     element (row, k) at byte  (k / 4) * LBO + (row / 8) * 128 + (row % 8) * 16 + (k % 4) * 4,   LBO = rows / 8 * 128.
// Slack: S_i = kTensorSlackU (|C|^2 + r^2), S_ray = kTensorSlackU |O|^2 (budget in DESIGN.md section 9).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#ifdef __CUDACC__
#define SPT_TENSOR_HD __host__ __device__
#else
#define SPT_TENSOR_HD
#endif

namespace spt
{

constexpr uint32_t kTensorRays = 128;      // M: one ray tile = four warps, one thread per TMEM lane
constexpr uint32_t kTensorCols = 128;      // N <= this: accumulator columns per ray tile (four tiles fill an SM's 512)
constexpr uint32_t kTensorK = 40;
constexpr double kTensorSlackU = 384.0 / 16777216.0;
constexpr uint32_t kTensorKCam = 32;       // camera form: six products + the constant = 28 K slots, padded to a multiple of 8
constexpr uint32_t kTensorABytes = kTensorRays * kTensorK * 4u;
constexpr uint32_t kTensorABytesCam = kTensorRays * kTensorKCam * 4u;
constexpr uint32_t kTensorALbo = (kTensorRays / 8u) * 128u;

// slots the tensor core takes (the first ones; the rest, if any, stay with the FFMA2 loop) and the rows of its matrix
SPT_TENSOR_HD inline uint32_t tensor_slots(uint32_t nSlots) { return nSlots < kTensorCols ? nSlots : kTensorCols; }
SPT_TENSOR_HD inline uint32_t tensor_rows(uint32_t nSlots) { return (tensor_slots(nSlots) + 15u) / 16u * 16u; }
SPT_TENSOR_HD inline uint32_t tensor_b_bytes(uint32_t rows) { return rows * kTensorK * 4u; }
SPT_TENSOR_HD inline uint32_t tensor_b_bytes_cam(uint32_t rows) { return rows * kTensorKCam * 4u; }
SPT_TENSOR_HD inline uint32_t tensor_b_lbo(uint32_t rows) { return (rows / 8u) * 128u; }
// float index of element (row, k) in a canonical matrix of `rows` rows
SPT_TENSOR_HD inline uint32_t tensor_elem(uint32_t rows, uint32_t row, uint32_t k)
{
    return ((k / 4u) * tensor_b_lbo(rows) + (row / 8u) * 128u + (row % 8u) * 16u + (k % 4u) * 4u) / 4u;
}

inline float tf32_trunc(float f)
{
    uint32_t b;
    std::memcpy(&b, &f, 4);
    b &= 0xFFFFE000u;
    std::memcpy(&f, &b, 4);
    return f;
}
// g ~ h + l with h, l TF32 values; up: h + l >= g (the tail is rounded toward +inf), else the tail is truncated
inline void tf32_split(double g, bool up, float &h, float &l)
{
    h = tf32_trunc(float(g));
    double const rest = g - double(h);
    l = tf32_trunc(float(rest));
    if (up)
    {
        while (double(l) < rest) // at most a couple of TF32 steps
        {
            float const mag = l != 0.f ? std::fabs(l) : std::fabs(float(rest));
            float const step = std::ldexp(1.f, std::ilogb(mag) - 10);
            l = tf32_trunc(l + step);
        }
    }
}

// hostGeom: cx, cy, cz, r per slot in binary64, r < 0 for padding slots (spt_cuda.cu, SptContext::hostGeom)
inline void build_tensor_matrix(const std::vector<double> &hostGeom, uint32_t nSlots, std::vector<float> &B)
{
    uint32_t const used = tensor_slots(nSlots), rows = tensor_rows(nSlots);
    B.assign(size_t(rows) * kTensorK, 0.f);
    for (uint32_t s = 0; s < rows; ++s)
    {
        double g[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, -1e30}; // never a candidate
        if (s < used && hostGeom[4 * s + 3] >= 0.0)
        {
            double const cx = hostGeom[4 * s], cy = hostGeom[4 * s + 1], cz = hostGeom[4 * s + 2], r = hostGeom[4 * s + 3];
            double const a2 = cx * cx + cy * cy + cz * cz;
            g[0] = cx * cx, g[1] = 2 * cx * cy, g[2] = 2 * cx * cz, g[3] = cy * cy, g[4] = 2 * cy * cz, g[5] = cz * cz;
            g[6] = cx, g[7] = cy, g[8] = cz;
            g[9] = -(a2 - r * r) + kTensorSlackU * (a2 + r * r) - double(1e-3f);
        }
        for (uint32_t c = 0; c < 10; ++c)
        {
            float h, l;
            tf32_split(g[c], c == 9, h, l);
            float const q[4] = {h, c < 9 ? h : l, c < 9 ? l : 1.f, c < 9 ? l : 1.f};
            for (uint32_t e = 0; e < 4; ++e)
                B[tensor_elem(rows, s, 4 * c + e)] = q[e];
        }
    }
}

// Camera form (all rays start at the eye E): with C' = C - E in binary64,
//   v = (C'.D)^2 - (|C'|^2 - r^2) + S - 1e-3 = sum_{i<=j} (c_ij C'_i C'_j)(D_i D_j) + k',   S = kTensorSlackU (|C|^2 + r^2 + |E|^2):
// six factors + the constant.  Sphere row: chunk c < 6 = (h, h, l, l) of factor c, chunk 6 = (k'_h, k'_l, 0, 0), chunk 7 = 0;
// ray row: chunk c < 6 = (h, l, h, l) of D_i D_j, chunk 6 = (1, 1, 0, 0), chunk 7 = 0 (chunks 6 and 7 never change).
inline void build_tensor_matrix_cam(const std::vector<double> &hostGeom, uint32_t nSlots, const double eye[3], std::vector<float> &B)
{
    uint32_t const used = tensor_slots(nSlots), rows = tensor_rows(nSlots);
    B.assign(size_t(rows) * kTensorKCam, 0.f);
    double const e2 = eye[0] * eye[0] + eye[1] * eye[1] + eye[2] * eye[2];
    for (uint32_t s = 0; s < rows; ++s)
    {
        double g[7] = {0, 0, 0, 0, 0, 0, -1e30}; // never a candidate
        if (s < used && hostGeom[4 * s + 3] >= 0.0)
        {
            double const cx = hostGeom[4 * s], cy = hostGeom[4 * s + 1], cz = hostGeom[4 * s + 2], r = hostGeom[4 * s + 3];
            double const x = cx - eye[0], y = cy - eye[1], z = cz - eye[2];
            double const a2 = cx * cx + cy * cy + cz * cz;
            g[0] = x * x, g[1] = 2 * x * y, g[2] = 2 * x * z, g[3] = y * y, g[4] = 2 * y * z, g[5] = z * z;
            g[6] = -(x * x + y * y + z * z - r * r) + kTensorSlackU * (a2 + r * r + e2) - double(1e-3f);
        }
        for (uint32_t c = 0; c < 7; ++c)
        {
            float h, l;
            tf32_split(g[c], c == 6, h, l);
            float const q[4] = {h, c < 6 ? h : l, c < 6 ? l : 0.f, c < 6 ? l : 0.f};
            for (uint32_t e = 0; e < 4; ++e)
                B[tensor_elem(rows, s, 4 * c + e)] = q[e];
        }
    }
}

} // namespace spt
