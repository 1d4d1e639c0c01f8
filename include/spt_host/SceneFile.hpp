// spt_host/SceneFile.hpp -- the scene as a file.  The reference generates its scene from a clock-seeded
// RNG at start-up (SceneGenerators.hpp) and never stores it; a dump of the SoA globals (Globals.hpp:31-37) makes a
// render reproducible and lets the large synthetic scenes (BASELINE configs 4 / 5) travel.
//
// Layout, little-endian, every array 16-byte aligned in the file:
//   0   char[8]  "SPTSCN1\0"
//   8   u32      n                      g_sphereNumber
//   12  u32      flags (0)
//   16  f32[16]  camera: eye xyz0, lookAt xyz0, up xyz0, sky colour rgb0   (Globals.hpp:22-29)
//   80  f32[4n]  g_spheres   (x, y, z, 0)
//   ..  f32[n]   g_radii     padded to a multiple of 4 floats
//   ..  u8[n]    g_materials padded to a multiple of 16 bytes   (Definitions.hpp:7-13 values)
//   ..  f32[4n]  g_colors    (r, g, b, 0), 0..255
//   ..  f32[n]   g_diffuses  padded to a multiple of 4 floats
#pragma once

#include "Scene.hpp"

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace io
{
namespace detail
{
inline size_t Pad16(size_t bytes) { return (bytes + 15u) / 16u * 16u; }
inline bool WritePadded(FILE *f, void const *p, size_t bytes)
{
    static char const zeros[16] = {0};
    return (bytes == 0 || std::fwrite(p, 1, bytes, f) == bytes) && std::fwrite(zeros, 1, Pad16(bytes) - bytes, f) == Pad16(bytes) - bytes;
}
inline bool ReadPadded(FILE *f, void *p, size_t bytes)
{
    char skip[16];
    return (bytes == 0 || std::fread(p, 1, bytes, f) == bytes) && std::fread(skip, 1, Pad16(bytes) - bytes, f) == Pad16(bytes) - bytes;
}
} // namespace detail

// writes the scene / camera globals
inline bool SaveScene(char const *path)
{
    FILE *f = std::fopen(path, "wb");
    if (!f)
        return false;
    uint32_t const n = g_sphereNumber, flags = 0;
    bool ok = std::fwrite("SPTSCN1", 1, 8, f) == 8 && std::fwrite(&n, 4, 1, f) == 1 && std::fwrite(&flags, 4, 1, f) == 1;
    float cam[16];
    std::memcpy(cam + 0, eyePos.xyzw, 16);
    std::memcpy(cam + 4, lookAt.xyzw, 16);
    std::memcpy(cam + 8, upDir.xyzw, 16);
    std::memcpy(cam + 12, initColor.xyzw, 16);
    ok = ok && std::fwrite(cam, 4, 16, f) == 16;
    ok = ok && detail::WritePadded(f, g_spheres.data(), size_t(n) * 16u);
    ok = ok && detail::WritePadded(f, g_radii.data(), size_t(n) * 4u);
    ok = ok && detail::WritePadded(f, g_materials.data(), size_t(n));
    ok = ok && detail::WritePadded(f, g_colors.data(), size_t(n) * 16u);
    ok = ok && detail::WritePadded(f, g_diffuses.data(), size_t(n) * 4u);
    return (std::fclose(f) == 0) && ok;
}

// fills the scene / camera globals (and rebuilds viewMatrix as Renderer.hpp:321 does); false on a malformed file
inline bool LoadScene(char const *path)
{
    FILE *f = std::fopen(path, "rb");
    if (!f)
        return false;
    char magic[8];
    uint32_t n = 0, flags = 0;
    float cam[16];
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "SPTSCN1", 8) == 0 && std::fread(&n, 4, 1, f) == 1 &&
              std::fread(&flags, 4, 1, f) == 1 && std::fread(cam, 4, 16, f) == 16 && n > 0 && n < (1u << 26);
    if (ok)
    {
        std::vector<math::Vec4> centers(n), colors(n);
        std::vector<float> radii(n), fuzz(n);
        std::vector<Material> mats(n);
        ok = detail::ReadPadded(f, centers.data(), size_t(n) * 16u) && detail::ReadPadded(f, radii.data(), size_t(n) * 4u) &&
             detail::ReadPadded(f, mats.data(), size_t(n)) && detail::ReadPadded(f, colors.data(), size_t(n) * 16u) &&
             detail::ReadPadded(f, fuzz.data(), size_t(n) * 4u);
        for (uint32_t i = 0; ok && i < n; ++i)
            ok = static_cast<uint8_t>(mats[i]) <= static_cast<uint8_t>(Material::DIFFUSE);
        if (ok)
        {
            g_spheres.swap(centers);
            g_colors.swap(colors);
            g_radii.swap(radii);
            g_diffuses.swap(fuzz);
            g_materials.swap(mats);
            g_sphereNumber = n;
            std::memcpy(eyePos.xyzw, cam + 0, 16);
            std::memcpy(lookAt.xyzw, cam + 4, 16);
            std::memcpy(upDir.xyzw, cam + 8, 16);
            std::memcpy(initColor.xyzw, cam + 12, 16);
            UpdateViewMatrix();
        }
    }
    std::fclose(f);
    return ok;
}
} // namespace io
