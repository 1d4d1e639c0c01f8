// spt_host_c.cpp -- C exports (include/spt_host.h) over the header-only host mirror.
#include "../../include/spt_host.h"
#include "../../include/spt_host/Image.hpp"
#include "../../include/spt_host/SceneFile.hpp"
#include "../../include/spt_host/SceneGenerators.hpp"

#include <cstring>
#include <mutex>

namespace
{
std::mutex g_lock; // the mirror keeps the scene in globals, as the reference does

uint32_t Export(uint32_t cap, float *centers4, float *radii, uint8_t *materials, float *colors4, float *fuzz)
{
    uint32_t const n = g_sphereNumber;
    if (cap >= n && centers4 && radii && materials && colors4 && fuzz)
    {
        for (uint32_t i = 0; i < n; ++i)
        {
            std::memcpy(centers4 + 4 * i, g_spheres[i].xyzw, 16);
            std::memcpy(colors4 + 4 * i, g_colors[i].xyzw, 16);
            radii[i] = g_radii[i];
            materials[i] = static_cast<uint8_t>(g_materials[i]);
            fuzz[i] = g_diffuses[i];
        }
    }
    return n;
}
} // namespace

extern "C"
{
uint32_t spt_host_init_spheres(uint32_t seed, uint32_t cap, float *c, float *r, uint8_t *m, float *k, float *f)
{
    std::lock_guard<std::mutex> g(g_lock);
    InitSpheres(seed);
    return Export(cap, c, r, m, k, f);
}
uint32_t spt_host_generate_spheres(uint32_t seed, uint32_t cap, float *c, float *r, uint8_t *m, float *k, float *f)
{
    std::lock_guard<std::mutex> g(g_lock);
    GenerateSpheres(seed);
    return Export(cap, c, r, m, k, f);
}
uint32_t spt_host_generate_metal_glass(uint32_t seed, uint32_t count, uint32_t cap, float *c, float *r, uint8_t *m, float *k,
                                       float *f)
{
    std::lock_guard<std::mutex> g(g_lock);
    GenerateMetalGlassSpheres(seed, count);
    return Export(cap, c, r, m, k, f);
}
uint32_t spt_host_generate_lattice(uint32_t seed, uint32_t side, uint32_t cap, float *c, float *r, uint8_t *m, float *k, float *f)
{
    std::lock_guard<std::mutex> g(g_lock);
    GenerateLatticeSpheres(seed, side);
    return Export(cap, c, r, m, k, f);
}
void spt_host_camera_basis(const float eye[3], const float look[3], const float up[3], float view16[16])
{
    math::Mat4 const v = math::Transpose(math::CreateCameraBasisMatrix(math::MakeVec4(eye[0], eye[1], eye[2]),
                                                                        math::MakeVec4(look[0], look[1], look[2]),
                                                                        math::MakeVec4(up[0], up[1], up[2])));
    std::memcpy(view16, v.array, sizeof(v.array));
}
int spt_host_save_bmp(const char *path, uint32_t width, uint32_t height, const uint8_t *rgb8)
{
    return io::WriteBmp(path, width, height, rgb8) ? 1 : 0;
}
int spt_host_save_scene(const char *path, uint32_t n, const float *c, const float *r, const uint8_t *m, const float *k, const float *f,
                        const float camera16[16])
{
    std::lock_guard<std::mutex> g(g_lock);
    spt_host::ClearScene();
    for (uint32_t i = 0; i < n; ++i)
        spt_host::PushSphere(math::MakeVec4(c[4 * i], c[4 * i + 1], c[4 * i + 2]), r[i], math::MakeVec4(k[4 * i], k[4 * i + 1], k[4 * i + 2]),
                             static_cast<Material>(m[i]), f[i]);
    g_sphereNumber = n;
    std::memcpy(eyePos.xyzw, camera16 + 0, 16);
    std::memcpy(lookAt.xyzw, camera16 + 4, 16);
    std::memcpy(upDir.xyzw, camera16 + 8, 16);
    std::memcpy(initColor.xyzw, camera16 + 12, 16);
    return io::SaveScene(path) ? 1 : 0;
}
uint32_t spt_host_load_scene(const char *path, uint32_t cap, float *c, float *r, uint8_t *m, float *k, float *f, float camera16[16])
{
    std::lock_guard<std::mutex> g(g_lock);
    if (!io::LoadScene(path))
        return 0;
    if (camera16)
    {
        std::memcpy(camera16 + 0, eyePos.xyzw, 16);
        std::memcpy(camera16 + 4, lookAt.xyzw, 16);
        std::memcpy(camera16 + 8, upDir.xyzw, 16);
        std::memcpy(camera16 + 12, initColor.xyzw, 16);
    }
    return Export(cap, c, r, m, k, f);
}
int spt_host_save_pfm(const char *path, uint32_t width, uint32_t height, const float *accum4)
{
    return io::WritePfm(path, width, height, accum4) ? 1 : 0;
}
int spt_host_save_hdr(const char *path, uint32_t width, uint32_t height, const float *accum4)
{
    return io::WriteHdr(path, width, height, accum4) ? 1 : 0;
}
int spt_host_save_accum(const char *path, uint32_t width, uint32_t height, uint32_t samples_done, uint32_t depth, uint64_t seed,
                        const float *accum4)
{
    return io::SaveAccum(path, width, height, samples_done, depth, seed, accum4) ? 1 : 0;
}
int spt_host_load_accum(const char *path, uint32_t width, uint32_t height, uint32_t *samples_done, uint32_t *depth, uint64_t *seed,
                        float *accum4)
{
    io::AccumHeader h{};
    std::vector<float> data;
    if (!io::LoadAccum(path, width, height, h, data))
        return 0;
    if (samples_done) *samples_done = h.samplesDone;
    if (depth) *depth = h.depth;
    if (seed) *seed = h.seed;
    if (accum4) std::memcpy(accum4, data.data(), data.size() * sizeof(float));
    return 1;
}
}
