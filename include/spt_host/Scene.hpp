// spt_host/Scene.hpp -- host-side mirror of the reference's scene/camera/config state, for programs that use
// the CUDA path WITHOUT the reference tree (headless main, benchmarks).  Same names and meaning as
// /root/reference/include/{Definitions,Globals,Math}.hpp so that include/CudaPathTracer.hpp reads either:
//
//   Material, RenderSegmentData                    include/Definitions.hpp:7-21
//   g_spheres, g_radii, g_materials, g_colors,
//   g_diffuses, g_sphereNumber                     include/Globals.hpp:31-37
//   viewMatrix, lookAt, eyePos, upDir, initColor   include/Globals.hpp:21-29
//   g_width, g_height, g_samples, g_bounces,
//   g_stride, g_size, g_data                       include/Globals.hpp:11-19
//
// Differences, all on purpose: the render constants are run-time variables (the reference bakes ONE
// configuration in at compile time), Vec4/Mat4 are plain float arrays (no SSE needed on the host: the
// arithmetic that matters runs on the GPU), and there are no GL/GLFW globals (no graphics stack here).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <vector>

namespace math
{
struct Vec4
{
    float xyzw[4];
};
struct Mat4
{
    float array[16];
};

inline Vec4 MakeVec4(float x, float y, float z) { return Vec4{{x, y, z, 0.f}}; }
inline Vec4 Sub(Vec4 a, Vec4 b) { return Vec4{{a.xyzw[0] - b.xyzw[0], a.xyzw[1] - b.xyzw[1], a.xyzw[2] - b.xyzw[2], a.xyzw[3] - b.xyzw[3]}}; }
// the reference's summation order (x*x + y*y) + (z*z + w*w), Math.hpp:122-133
inline float LengthSquared(Vec4 v)
{
    float const a = v.xyzw[0] * v.xyzw[0] + v.xyzw[1] * v.xyzw[1];
    float const b = v.xyzw[2] * v.xyzw[2] + v.xyzw[3] * v.xyzw[3];
    return a + b;
}
inline float Length(Vec4 v) { return std::sqrt(LengthSquared(v)); }
inline Vec4 Normalize(Vec4 v)
{
    float const l = Length(v);
    return Vec4{{v.xyzw[0] / l, v.xyzw[1] / l, v.xyzw[2] / l, v.xyzw[3] / l}};
}
// The reference's Cross has a wrong z lane (a.x*b.x - a.y*b.x, Math.hpp:113-120).  It is kept, because the
// view matrix the tracers see is built with it (Renderer.hpp:321); harmless for the default camera.
inline Vec4 Cross(Vec4 a, Vec4 b)
{
    return MakeVec4(a.xyzw[1] * b.xyzw[2] - a.xyzw[2] * b.xyzw[1], a.xyzw[2] * b.xyzw[0] - a.xyzw[0] * b.xyzw[2],
                    a.xyzw[0] * b.xyzw[0] - a.xyzw[1] * b.xyzw[0]);
}
// rows right / up / view, row 3 = 0 (Math.hpp:198-209)
inline Mat4 CreateCameraBasisMatrix(Vec4 eye, Vec4 look, Vec4 up)
{
    Vec4 const viewDir = Normalize(Sub(look, eye));
    Vec4 const right = Normalize(Cross(up, viewDir));
    Vec4 const upDir = Cross(viewDir, right);
    Mat4 m{};
    for (int k = 0; k < 4; ++k)
    {
        m.array[0 + k] = right.xyzw[k];
        m.array[4 + k] = upDir.xyzw[k];
        m.array[8 + k] = viewDir.xyzw[k];
        m.array[12 + k] = 0.f;
    }
    return m;
}
inline Mat4 Transpose(Mat4 m) // Math.hpp:211-231
{
    Mat4 t{};
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c)
            t.array[4 * r + c] = m.array[4 * c + r];
    return t;
}
} // namespace math

enum class Material : uint8_t // values as include/Definitions.hpp:7-13
{
    SKYBOX,
    REFLECTIVE,
    REFRACTIVE,
    DIFFUSE,
};

struct RenderSegmentData // include/Definitions.hpp:15-21, half-open
{
    uint32_t yBegin = 0;
    uint32_t yEnd = 0;
    uint32_t xBegin = 0;
    uint32_t xEnd = 0;
};

// ---- render configuration (run-time here) ----
inline uint32_t g_maxThreads = 4;
inline uint32_t g_bounces = 10;
inline uint32_t g_samples = 100;
inline uint32_t g_width = 1440;
inline uint32_t g_height = 1440;
constexpr uint8_t g_stride = 3;
inline uint32_t g_size = 0;
inline uint8_t *g_data = nullptr;

// ---- camera (include/Globals.hpp:21-29) ----
inline math::Mat4 viewMatrix = {{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}};
inline math::Vec4 lookAt = {{0, 1, 0, 0}};
inline math::Vec4 eyePos = {{0, 1, -3, 0}};
inline math::Vec4 upDir = {{0, 1, 0, 0}};
inline math::Vec4 initColor = {{137, 207, 240, 0}};

// ---- SoA scene (include/Globals.hpp:31-37) ----
inline std::vector<math::Vec4> g_colors;
inline std::vector<math::Vec4> g_spheres;
inline std::vector<float> g_radii;
inline std::vector<Material> g_materials;
inline std::vector<float> g_diffuses;
inline uint32_t g_sphereNumber = 10;

// (re)allocate g_data for a configuration; the reference mallocs it once at static-init time (Globals.hpp:19)
inline void SetRenderConfig(uint32_t width, uint32_t height, uint32_t samples, uint32_t bounces)
{
    g_width = width;
    g_height = height;
    g_samples = samples;
    g_bounces = bounces;
    g_size = width * height * g_stride;
    std::free(g_data);
    g_data = static_cast<uint8_t *>(std::calloc(g_size, 1));
}

// viewMatrix exactly as MainLoop builds it (Renderer.hpp:321)
inline void UpdateViewMatrix() { viewMatrix = math::Transpose(math::CreateCameraBasisMatrix(eyePos, lookAt, upDir)); }
