// spt_host/Image.hpp -- headless image output: io::SaveImage() without GLFW / ImGui / stb.
//
// The reference writes g_data with stbi_write_bmp (include/IOHelpers.hpp:24-27,
// third_party/stb/stb_image_write.h:473-499) to "output<spp>s<depth>b.bmp".  The writer below produces the
// same bytes for 3-channel data (checked against a file written by stb itself,
// tests/golden/tiny_7x5.bmp): 14-byte file header + 40-byte BITMAPINFOHEADER, 24 bpp, rows bottom-up,
// BGR order, each row zero-padded to a multiple of 4 bytes.
#pragma once

#include "Scene.hpp"

#include <cstdio>
#include <string>

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

// io::SaveImage (include/IOHelpers.hpp:24-27): same file name rule
inline std::string DefaultImageName() { return "output" + std::to_string(g_samples) + "s" + std::to_string(g_bounces) + "b.bmp"; }

inline bool SaveImage(char const *path = nullptr)
{
    std::string const name = path ? std::string(path) : DefaultImageName();
    return WriteBmp(name.c_str(), g_width, g_height, g_data);
}
} // namespace io
