/*
 * oracle/ref_shim/ref_prelude.hpp -- TEST INFRASTRUCTURE ONLY.
 *
 * Brings the UNMODIFIED reference headers (found on the include path under /root/reference/include) into a
 * g++ translation unit.  Shared by ref_driver.cpp (the oracle library) and ref_integration.cpp (the drop-in
 * demonstration).  See ref_driver.cpp for what each token rebinding is for; none touches arithmetic.
 */
#pragma once
#include <intrin.h> /* the shim; pulls every system header the reference needs */

#include <assert.h>
#include <limits.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- chosen-seed clock (Random.hpp:88-89) ------------------------------ */
static std::atomic<long long> g_sptNextSeed{1};
namespace std
{
namespace chrono
{
struct spt_seed_clock
{
    struct duration_t
    {
        long long v;
        long long count() const { return v; }
    };
    struct time_point_t
    {
        long long v;
        duration_t time_since_epoch() const { return {v}; }
    };
    /* each thread that first touches the reference RNG consumes one seed */
    static time_point_t now() { return {g_sptNextSeed.fetch_add(1)}; }
};
} // namespace chrono
} // namespace std

#define random spt_ref_random
#define steady_clock spt_seed_clock
#define uint8_t ::std::uint8_t

#include <Math.hpp>
#include <Definitions.hpp>
#include <Globals.hpp>

/* ---- run-time render configuration (Globals.hpp:12-19) ----------------- */
struct SptRefConfig
{
    uint32_t width = 1440, height = 1440, samples = 100, bounces = 10, size = 1440 * 1440 * 3;
    unsigned char *data = nullptr;
    float *tap = nullptr; /* W*H*4 linear float colour as passed to io::WritePixel */
};
static SptRefConfig g_sptCfg;
#define g_width (g_sptCfg.width)
#define g_height (g_sptCfg.height)
#define g_samples (g_sptCfg.samples)
#define g_bounces (g_sptCfg.bounces)
#define g_size (g_sptCfg.size)
#define g_data (g_sptCfg.data)

#include <Random.hpp>

#ifdef SPT_REF_WIDE
#undef uint8_t
#define uint8_t ::std::uint32_t
#endif
#include <Collision.hpp>
#ifdef SPT_REF_WIDE
#undef uint8_t
#define uint8_t ::std::uint8_t
#endif

#include <IOHelpers.hpp>

namespace io
{
inline void SptWritePixelTap(uint32_t index, math::Vec4 color)
{
    if (g_sptCfg.tap)
    {
        float *p = g_sptCfg.tap + (index / 3) * 4;
        p[0] = color.xyzw[0];
        p[1] = color.xyzw[1];
        p[2] = color.xyzw[2];
        p[3] = color.xyzw[3];
    }
    WritePixel(index, color);
}
} // namespace io
#define WritePixel SptWritePixelTap

#include <SceneGenerators.hpp>
#include <SingleThreadPathTracer.hpp>
#include <TaskBasedPathTracer.hpp>

#undef WritePixel


