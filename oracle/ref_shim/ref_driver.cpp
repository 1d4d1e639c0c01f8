/*
 * oracle/ref_shim/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Compiles the UNMODIFIED reference headers, where they lie under
 * /root/reference/include, into oracle/_ref/libspt_ref.so and exposes the
 * reference's own hot-path functions through a small C interface:
 *
 *   cd::FindClosestIntersectionSphere      (include/Collision.hpp:87-109)
 *   cd::CalculateRaySphere* helpers        (include/Collision.hpp:19-71)
 *   RenderSegment                          (include/SingleThreadPathTracer.hpp:114-137)
 *   RenderSegmentTask                      (include/TaskBasedPathTracer.hpp:54-206)
 *   InitSpheres / GenerateSpheres          (include/SceneGenerators.hpp:68 / :6)
 *   CreateCameraBasisMatrix / Transpose    (include/Math.hpp:198-231, used as Renderer.hpp:321)
 *   io::WritePixel / stbi_write_bmp        (include/IOHelpers.hpp:17-27)
 *
 * No reference source is copied or edited.  The MSVC-only constructs are
 * satisfied by shim headers (intrin.h, Gl.hpp in this directory) and by the
 * token-level rebindings below, none of which touches an arithmetic operation:
 *
 *   random           -> spt_ref_random   (namespace name collides with glibc ::random(), Random.hpp:8)
 *   uint8_t          -> ::std::uint8_t   (makes `template <typename uint8_t D>` well-formed, Random.hpp:95,109)
 *   steady_clock     -> spt_seed_clock   (Random.hpp:88-89 seeds from the clock; we hand it a chosen seed)
 *   g_width/g_height/g_samples/g_bounces/g_size/g_data
 *                    -> run-time values  (Globals.hpp:12-19 are compile-time constants of ONE config)
 *   WritePixel       -> tap that also records the linear float colour (IOHelpers.hpp:17-22 only keeps 8 bits)
 *   (SPT_REF_WIDE only) uint8_t -> ::std::uint32_t while Collision.hpp is read, lifting the
 *                    255-sphere ceiling of Collision.hpp:87-92 for the >255-sphere synthetic scenes.
 *
 * Renderer.hpp (GLFW/GL viewer + tile scheduler) is NOT included: no graphics stack here.
 */
#include "ref_prelude.hpp"

/* ------------------------------------------------------------------------ */
extern "C"
{

int sptref_wide_index(void)
{
#ifdef SPT_REF_WIDE
    return 1;
#else
    return 0;
#endif
}

void sptref_set_config(uint32_t width, uint32_t height, uint32_t samples, uint32_t bounces)
{
    g_sptCfg.width = width;
    g_sptCfg.height = height;
    g_sptCfg.samples = samples;
    g_sptCfg.bounces = bounces;
    g_sptCfg.size = width * height * 3u;
}

/* camera exactly as Renderer.hpp:321 */
void sptref_set_camera(float const eye[3], float const look[3])
{
    eyePos = {eye[0], eye[1], eye[2]};
    lookAt = {look[0], look[1], look[2]};
    viewMatrix = Transpose(math::CreateCameraBasisMatrix(eyePos, lookAt, upDir));
}

void sptref_get_camera(float view16[16], float eye4[4])
{
    memcpy(view16, viewMatrix.array, sizeof(float) * 16);
    memcpy(eye4, eyePos.xyzw, sizeof(float) * 4);
}

void sptref_get_sky(float sky4[4]) { memcpy(sky4, initColor.xyzw, sizeof(float) * 4); }

/* scenes: run the generator on a fresh thread so that its thread_local RNG is
 * constructed from `seed` (Random.hpp:88-89) */
static void RunSeeded(uint32_t seed, void (*fn)())
{
    g_sptNextSeed.store(seed);
    std::thread t(fn);
    t.join();
}

void sptref_init_spheres(uint32_t seed)
{
    g_sphereNumber = 10; /* the initialiser of Globals.hpp:37, as in a fresh process */
    g_attenuations.clear();
    g_diffuses.clear();
    RunSeeded(seed, [] { InitSpheres(); });
}

void sptref_generate_spheres(uint32_t seed)
{
    g_attenuations.clear();
    g_diffuses.clear();
    RunSeeded(seed, [] { GenerateSpheres(); });
}

uint32_t sptref_scene_size(void) { return g_sphereNumber; }

void sptref_get_scene(float *centers4, float *radii, unsigned char *materials, float *colors4, float *fuzz)
{
    for (uint32_t i = 0; i < g_sphereNumber; ++i)
    {
        memcpy(centers4 + 4 * i, g_spheres[i].xyzw, 16);
        memcpy(colors4 + 4 * i, g_colors[i].xyzw, 16);
        radii[i] = g_radii[i];
        materials[i] = static_cast<unsigned char>(g_materials[i]);
        fuzz[i] = g_diffuses[i];
    }
}

int sptref_set_scene(uint32_t n, float const *centers4, float const *radii, unsigned char const *materials,
                     float const *colors4, float const *fuzz)
{
#ifndef SPT_REF_WIDE
    if (n > 255)
        return -1; /* Collision.hpp:87-92: uint8_t index never terminates for n >= 256 */
#endif
    g_spheres.resize(n);
    g_colors.resize(n);
    g_radii.resize(n);
    g_materials.resize(n);
    g_diffuses.resize(n);
    g_attenuations.assign(n, {1, 1, 1});
    for (uint32_t i = 0; i < n; ++i)
    {
        g_spheres[i] = {centers4[4 * i + 0], centers4[4 * i + 1], centers4[4 * i + 2], centers4[4 * i + 3]};
        g_colors[i] = {colors4[4 * i + 0], colors4[4 * i + 1], colors4[4 * i + 2], colors4[4 * i + 3]};
        g_radii[i] = radii[i];
        g_materials[i] = static_cast<Material>(materials[i]);
        g_diffuses[i] = fuzz[i];
    }
    g_sphereNumber = n;
    return 0;
}

/* jitter-free pinhole rays built exactly as SingleThreadPathTracer.hpp:125-128 with xi = 0,
 * then the reference's own closest-hit and contact helpers.  Outputs per pixel (x + y*W):
 * idx (g_sphereNumber = miss), t_min, normal xyz, direction xyz. */
void sptref_first_hit(uint32_t xBegin, uint32_t xEnd, uint32_t yBegin, uint32_t yEnd, uint32_t *idx, float *t,
                      float *normal3, float *dir3)
{
    for (uint32_t y = yBegin; y < yEnd; ++y)
    {
        for (uint32_t x = xBegin; x < xEnd; ++x)
        {
            float const u = static_cast<float>(y + 0.f) / g_width;
            float const v = static_cast<float>(x + 0.f) / g_height;
            math::Vec4 const d = math::Normalize(viewMatrix * math::Vec4{-1.f + 2.f * v, -1.f + 2.f * u, 1.f});
            math::Vec4 const o{eyePos};
            uint32_t const i = cd::FindClosestIntersectionSphere(d, o);
            size_t const p = size_t(x - xBegin) + size_t(y - yBegin) * (xEnd - xBegin);
            idx[p] = i;
            if (dir3)
            {
                dir3[3 * p + 0] = d.xyzw[0];
                dir3[3 * p + 1] = d.xyzw[1];
                dir3[3 * p + 2] = d.xyzw[2];
            }
            if (i < g_sphereNumber)
            {
                t[p] = cd::CalculateRaySphereMinIntersectionFactor(g_spheres[i] - o, g_radii[i], d);
                math::Vec4 const c = cd::CalculateRaySphereClosestContactPoint(g_spheres[i], g_radii[i], o, d);
                math::Vec4 const n = cd::CalculateRaySphereContactNormal(c, g_spheres[i]);
                normal3[3 * p + 0] = n.xyzw[0];
                normal3[3 * p + 1] = n.xyzw[1];
                normal3[3 * p + 2] = n.xyzw[2];
            }
            else
            {
                t[p] = 0.f;
                normal3[3 * p + 0] = normal3[3 * p + 1] = normal3[3 * p + 2] = 0.f;
            }
        }
    }
}

/* closest hit for caller-supplied rays (direction xyz, origin xyz) */
void sptref_closest_hit(uint32_t count, float const *dir3, float const *org3, uint32_t *idx)
{
    for (uint32_t k = 0; k < count; ++k)
    {
        math::Vec4 const d{dir3[3 * k], dir3[3 * k + 1], dir3[3 * k + 2]};
        math::Vec4 const o{org3[3 * k], org3[3 * k + 1], org3[3 * k + 2]};
        idx[k] = cd::FindClosestIntersectionSphere(d, o);
    }
}

/* One render of the half-open segment on ONE fresh thread whose RNG is seeded with `seed`:
 * fully deterministic.  tracer 0 = RenderSegment (recursive), 1 = RenderSegmentTask (queues).
 * mean4 (W*H*4, may be null) receives the linear colour handed to io::WritePixel at the
 * reference's own (vertically flipped) pixel slot; rgb8 (W*H*3) is g_data. */
void sptref_render(int tracer, uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, uint32_t seed,
                   float *mean4, unsigned char *rgb8)
{
    g_sptCfg.data = rgb8;
    g_sptCfg.tap = mean4;
    g_sptNextSeed.store(seed);
    RenderSegmentData const seg{yBegin, yEnd, xBegin, xEnd};
    std::thread t([seg, tracer] {
        if (tracer == 0)
            RenderSegment(seg);
        else
            RenderSegmentTask(seg);
    });
    t.join();
    g_sptCfg.data = nullptr;
    g_sptCfg.tap = nullptr;
}

/* Whole image on `threads` workers pulling square `tile` x `tile` segments from an atomic
 * counter (the task tracer mis-strides non-square segments, TaskBasedPathTracer.hpp:103,186).
 * Worker k's RNG seed is seed + (order of first RNG use).  Returns wall seconds of the render. */
double sptref_render_mt(int tracer, uint32_t tile, uint32_t threads, uint32_t seed, float *mean4,
                        unsigned char *rgb8)
{
    g_sptCfg.data = rgb8;
    g_sptCfg.tap = mean4;
    g_sptNextSeed.store(seed);
    uint32_t const tx = (g_width + tile - 1) / tile, ty = (g_height + tile - 1) / tile;
    std::atomic<uint32_t> next{0};
    auto const t0 = std::chrono::system_clock::now();
    std::vector<std::thread> pool;
    for (uint32_t k = 0; k < threads; ++k)
    {
        pool.emplace_back([&] {
            for (;;)
            {
                uint32_t const j = next.fetch_add(1);
                if (j >= tx * ty)
                    break;
                uint32_t const x0 = (j % tx) * tile, y0 = (j / tx) * tile;
                RenderSegmentData const seg{y0, std::min(y0 + tile, g_height), x0, std::min(x0 + tile, g_width)};
                if (tracer == 0)
                    RenderSegment(seg);
                else
                    RenderSegmentTask(seg);
            }
        });
    }
    for (auto &t : pool)
        t.join();
    auto const t1 = std::chrono::system_clock::now();
    g_sptCfg.data = nullptr;
    g_sptCfg.tap = nullptr;
    return std::chrono::duration<double>(t1 - t0).count();
}

/* the reference's RNG front end, for distribution checks (Random.hpp:86-93, :115-127) */
void sptref_rng_uniform(uint32_t seed, float lo, float hi, uint32_t count, float *out)
{
    g_sptNextSeed.store(seed);
    std::thread t([=] {
        for (uint32_t i = 0; i < count; ++i)
            out[i] = spt_ref_random::GenerateUniformRealDist(lo, hi);
    });
    t.join();
}

void sptref_rng_shell(uint32_t seed, uint32_t count, float *out3)
{
    g_sptNextSeed.store(seed);
    std::thread t([=] {
        for (uint32_t i = 0; i < count; ++i)
        {
            math::Vec4 const v = spt_ref_random::GenerateUniformDistInsideSphereVector();
            out3[3 * i + 0] = v.xyzw[0];
            out3[3 * i + 1] = v.xyzw[1];
            out3[3 * i + 2] = v.xyzw[2];
        }
    });
    t.join();
}

/* io::WritePixel on a caller buffer (IOHelpers.hpp:17-22) */
void sptref_write_pixel(float r, float g, float b, unsigned char out[3])
{
    g_sptCfg.data = out;
    io::WritePixel(0, math::Vec4{r, g, b});
    g_sptCfg.data = nullptr;
}

/* the BMP serialiser the reference calls (IOHelpers.hpp:26), at a caller-chosen path */
int sptref_save_bmp(char const *path, uint32_t w, uint32_t h, unsigned char const *rgb8)
{
    return stbi_write_bmp(path, int(w), int(h), 3, rgb8);
}

} /* extern "C" */
