// spt_host/Image.hpp -- headless image output: io::SaveImage() without GLFW / ImGui / stb.
//
// The reference writes g_data with stbi_write_bmp (include/IOHelpers.hpp:24-27,
// third_party/stb/stb_image_write.h:473-499) to "output<spp>s<depth>b.bmp".  The writer below produces the
// same bytes for 3-channel data (checked against a file written by stb itself,
// tests/golden/tiny_7x5.bmp): 14-byte file header + 40-byte BITMAPINFOHEADER, 24 bpp, rows bottom-up,
// BGR order, each row zero-padded to a multiple of 4 bytes.
#pragma once

#include "Scene.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

namespace io
{
inline bool WriteBmp(char const *path, uint32_t width, uint32_t height, uint8_t const *rgb)
{
    FILE *f = std::fopen(path, "wb");
    if (!f)
        return false;
    uint32_t const pad = (4u - (width * 3u) % 4u) % 4u;
    uint32_t const fileSize = 14u + 40u + (width * 3u + pad) * height;
    auto put16 = [f](uint32_t v) { std::fputc(int(v & 0xFF), f); std::fputc(int((v >> 8) & 0xFF), f); };
    auto put32 = [f](uint32_t v) { for (int k = 0; k < 4; ++k) std::fputc(int((v >> (8 * k)) & 0xFF), f); };
    std::fputc('B', f);
    std::fputc('M', f);
    put32(fileSize);
    put16(0);
    put16(0);
    put32(14u + 40u);
    put32(40u);
    put32(width);
    put32(height);
    put16(1);
    put16(24);
    for (int k = 0; k < 6; ++k)
        put32(0);
    for (uint32_t row = height; row-- > 0;)
    {
        uint8_t const *p = rgb + size_t(row) * width * 3u;
        for (uint32_t x = 0; x < width; ++x, p += 3)
        {
            std::fputc(p[2], f);
            std::fputc(p[1], f);
            std::fputc(p[0], f);
        }
        for (uint32_t k = 0; k < pad; ++k)
            std::fputc(0, f);
    }
    return std::fclose(f) == 0;
}

// Linear-float output next to the 8-bit BMP: Portable Float Map, "PF\n<w> <h>\n-1.0\n" followed by
// little-endian float RGB rows, BOTTOM row first -- which is the tracers' own y direction (y = 0 is the bottom
// row, SingleThreadPathTracer.hpp:120), so accumulator rows are written in order.  Values are the mean radiance
// sum / count on the reference's 0..255 scale divided by 255 (linear, no gamma).  accum4 = width*height (sum r,
// sum g, sum b, count) as spt_cuda_render / spt_cuda_progressive_image return it.
inline bool WritePfm(char const *path, uint32_t width, uint32_t height, float const *accum4)
{
    FILE *f = std::fopen(path, "wb");
    if (!f)
        return false;
    std::fprintf(f, "PF\n%u %u\n-1.0\n", width, height);
    std::vector<float> row(size_t(width) * 3u);
    for (uint32_t y = 0; y < height; ++y)
    {
        for (uint32_t x = 0; x < width; ++x)
        {
            float const *a = accum4 + (size_t(y) * width + x) * 4u;
            float const inv = 1.f / a[3];
            for (int k = 0; k < 3; ++k)
                row[size_t(x) * 3u + k] = a[k] * inv / 255.f;
        }
        std::fwrite(row.data(), sizeof(float), row.size(), f);
    }
    return std::fclose(f) == 0;
}

// The other linear-float format SURVEY 8f-3 names: Radiance RGBE (.hdr), what the vendored stbi_write_hdr
// (third_party/stb/stb_image_write.h:183) would serialise.  Header "#?RADIANCE", "FORMAT=32-bit_rle_rgbe", blank line,
// resolution "-Y <h> +X <w>"; then rows TOP first, four bytes per pixel, flat (no run-length coding: every reader takes flat
// scanlines).  A pixel is (r, g, b) * 2^-e rounded down to 8 bits each, e = exponent of the largest channel, stored as
// e + 128; channels below 1e-32 become (0, 0, 0, 0).  Same values as WritePfm (mean radiance / 255, linear).
inline bool WriteHdr(char const *path, uint32_t width, uint32_t height, float const *accum4)
{
    FILE *f = std::fopen(path, "wb");
    if (!f)
        return false;
    std::fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y %u +X %u\n", height, width);
    std::vector<uint8_t> row(size_t(width) * 4u);
    for (uint32_t y = height; y-- > 0;) // accumulator row y = 0 is the BOTTOM of the picture
    {
        for (uint32_t x = 0; x < width; ++x)
        {
            float const *a = accum4 + (size_t(y) * width + x) * 4u;
            float const inv = 1.f / a[3];
            float c[3];
            for (int k = 0; k < 3; ++k)
                c[k] = a[k] * inv / 255.f;
            float const top = std::max(c[0], std::max(c[1], c[2]));
            uint8_t *o = row.data() + size_t(x) * 4u;
            if (!(top >= 1e-32f))
                o[0] = o[1] = o[2] = o[3] = 0;
            else
            {
                int e = 0;
                float const scale = std::frexp(top, &e) * 256.f / top; // top * scale in [128, 256)
                for (int k = 0; k < 3; ++k)
                    o[k] = uint8_t(c[k] > 0.f ? c[k] * scale : 0.f);
                o[3] = uint8_t(e + 128);
            }
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    return std::fclose(f) == 0;
}

// Accumulator checkpoint (resume a render later, or on another GPU): 48-byte header
//   "SPTACC2\0", u32 width, u32 height, u32 samplesDone, u32 depth, u64 seed, u64 identity, u64 reserved
// then width*height float4 (sum r, sum g, sum b, count), little-endian, row y*width + x.  `identity` is the caller's
// fingerprint of everything else the sums depend on (scene arrays, camera, tracer semantics: RenderIdentity() below);
// a loader that passes its own fingerprint refuses a file written for another render.
struct AccumHeader
{
    char magic[8];
    uint32_t width, height, samplesDone, depth;
    uint64_t seed;
    uint64_t identity;
    uint64_t reserved;
};
static_assert(sizeof(AccumHeader) == 48, "checkpoint header layout");

// FNV-1a over the bytes that decide a render besides size / depth / seed: the SoA scene globals, the camera, the sky, and
// the tracer's semantics (pass cap).  Two renders with the same fingerprint accumulate the same per-sample radiance.
inline uint64_t RenderIdentity(uint32_t passCap)
{
    uint64_t h = 1469598103934665603ull;
    auto mix = [&h](void const *p, size_t n) {
        unsigned char const *b = static_cast<unsigned char const *>(p);
        for (size_t i = 0; i < n; ++i)
            h = (h ^ b[i]) * 1099511628211ull;
    };
    uint32_t const n = g_sphereNumber;
    mix(&n, sizeof(n));
    mix(g_spheres.data(), sizeof(g_spheres[0]) * n);
    mix(g_radii.data(), sizeof(float) * n);
    mix(g_materials.data(), sizeof(g_materials[0]) * n);
    mix(g_colors.data(), sizeof(g_colors[0]) * n);
    mix(g_diffuses.data(), sizeof(float) * n);
    mix(viewMatrix.array, sizeof(viewMatrix.array));
    mix(eyePos.xyzw, sizeof(eyePos.xyzw));
    mix(initColor.xyzw, sizeof(initColor.xyzw));
    mix(&passCap, sizeof(passCap));
    return h ? h : 1;
}

inline bool SaveAccum(char const *path, uint32_t width, uint32_t height, uint32_t samplesDone, uint32_t depth, uint64_t seed,
                      float const *accum4, uint64_t identity = 0)
{
    std::string const tmp = std::string(path) + ".tmp"; // written whole, then renamed: an interrupted run leaves the old file
    FILE *f = std::fopen(tmp.c_str(), "wb");
    if (!f)
        return false;
    AccumHeader h{{'S', 'P', 'T', 'A', 'C', 'C', '2', 0}, width, height, samplesDone, depth, seed, identity, 0};
    bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1;
    ok = ok && std::fwrite(accum4, sizeof(float) * 4u, size_t(width) * height, f) == size_t(width) * height;
    ok = (std::fclose(f) == 0) && ok;
    return ok && std::rename(tmp.c_str(), path) == 0;
}

// returns false when the file is missing, malformed, of another image size, or -- identity != 0 -- written for another render
inline bool LoadAccum(char const *path, uint32_t width, uint32_t height, AccumHeader &h, std::vector<float> &accum4, uint64_t identity = 0)
{
    FILE *f = std::fopen(path, "rb");
    if (!f)
        return false;
    bool ok = std::fread(&h, sizeof(h), 1, f) == 1 && std::string(h.magic, 7) == "SPTACC2" && h.width == width && h.height == height;
    ok = ok && (identity == 0 || h.identity == identity);
    if (ok)
    {
        accum4.resize(size_t(width) * height * 4u);
        ok = std::fread(accum4.data(), sizeof(float) * 4u, size_t(width) * height, f) == size_t(width) * height;
    }
    std::fclose(f);
    return ok;
}

// io::SaveImage (include/IOHelpers.hpp:24-27): same file name rule
inline std::string DefaultImageName() { return "output" + std::to_string(g_samples) + "s" + std::to_string(g_bounces) + "b.bmp"; }

inline bool SaveImage(char const *path = nullptr)
{
    std::string const name = path ? std::string(path) : DefaultImageName();
    return WriteBmp(name.c_str(), g_width, g_height, g_data);
}
} // namespace io
