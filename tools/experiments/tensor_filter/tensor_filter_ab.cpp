// tensor_filter_ab.cpp -- smallest possible GPU check of an EXPERIMENTAL -DSPT_TENSOR_FILTER=1 build of the library (no
// Python, starts in a second): renders two synthetic scenes through the C ABI four times from the SAME library -- bounce kernel
// and camera kernel each on the tensor core (default of that build) or on the FFMA2 loop (SPT_TENSOR=0 / SPT_TENSOR_PRIMARY=0) --
// and compares the accumulators byte for byte against the all-FFMA2 render.
//   g++ -O2 -std=c++17 -I include -o build_ab/tensor_filter_ab tools/tensor_filter_ab.cpp -ldl
//   timeout 30 build_ab/tensor_filter_ab build_ab/tensor.so [scene.bin width height spp depth reps]   (tools/export_scene.py)
#include "spt_cuda.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <random>
#include <vector>

// modes: (bounce kernel, camera kernel) on the tensor core; the last one is the shipped path
static const char *const kModes[4][3] = {{"1", "1", "tensor+tensor"}, {"1", "0", "tensor+ffma2 "}, {"0", "1", "ffma2 +tensor"}, {"0", "0", "ffma2 +ffma2 "}};
static void set_mode(int m)
{
    setenv("SPT_TENSOR", kModes[m][0], 1);
    setenv("SPT_TENSOR_PRIMARY", kModes[m][1], 1);
}

#define LOAD(name) auto name = reinterpret_cast<decltype(&::name)>(dlsym(h, #name)); if (!name) { std::printf("missing %s\n", #name); return 2; }

int main(int argc, char **argv)
{
    if (argc < 2)
        return 2;
    void *h = dlopen(argv[1], RTLD_NOW);
    if (!h)
    {
        std::printf("dlopen: %s\n", dlerror());
        return 2;
    }
    LOAD(spt_cuda_create) LOAD(spt_cuda_destroy) LOAD(spt_cuda_set_scene) LOAD(spt_cuda_render) LOAD(spt_cuda_get_timing)
    LOAD(spt_cuda_last_error) LOAD(spt_cuda_abi_version) LOAD(spt_cuda_get_counters) LOAD(spt_cuda_reset_counters)
    if (spt_cuda_abi_version() != SPT_ABI_VERSION)
        return 2;
    SptContext *ctx = nullptr;
    if (spt_cuda_create(&ctx, 0))
    {
        std::printf("create: %s\n", spt_cuda_last_error());
        return 2;
    }
    int bad = 0;
    for (uint32_t nSmall : {100u, 300u}) // below and above 128 filter slots (the second exercises tensor + FFMA2 remainder)
    {
        std::mt19937 rng(7 + nSmall);
        std::uniform_real_distribution<float> uni(0.f, 1.f);
        uint32_t const n = nSmall + 1;
        std::vector<float> cen(n * 4, 0.f), rad(n), col(n * 4, 0.f), fuzz(n);
        std::vector<uint8_t> mat(n);
        cen[1] = -1000.f, rad[0] = 1000.f, mat[0] = 3, col[0] = col[1] = col[2] = 128.f, fuzz[0] = 0.f; // the ground, diffuse
        for (uint32_t i = 1; i < n; ++i)
        {
            cen[4 * i] = -9.f + 18.f * uni(rng), cen[4 * i + 1] = 0.3f + 2.5f * uni(rng), cen[4 * i + 2] = 3.f + 14.f * uni(rng);
            rad[i] = 0.2f + 0.5f * uni(rng);
            mat[i] = uint8_t(1 + i % 3); // mirror, glass, diffuse
            for (int k = 0; k < 3; ++k)
                col[4 * i + k] = 60.f + 190.f * uni(rng);
            fuzz[i] = 0.3f * uni(rng);
        }
        SptScene sc{n, cen.data(), rad.data(), mat.data(), col.data(), fuzz.data()};
        if (spt_cuda_set_scene(ctx, &sc))
        {
            std::printf("set_scene: %s\n", spt_cuda_last_error());
            return 2;
        }
        SptCamera cam{};
        cam.view[0] = cam.view[5] = cam.view[10] = 1.f;
        cam.eye[0] = 0.f, cam.eye[1] = 1.5f, cam.eye[2] = -4.f;
        cam.sky[0] = 128.f, cam.sky[1] = 178.f, cam.sky[2] = 255.f;
        SptParams prm{};
        prm.width = 200, prm.height = 100, prm.depth = 8, prm.pass_cap = 0, prm.sample_begin = 0, prm.sample_end = 6, prm.seed = 99;
        prm.variant = SPT_VARIANT_MEGAKERNEL, prm.flags = SPT_FLAG_COUNTERS;
        SptSegment seg{0, prm.height, 0, prm.width};
        std::vector<float> acc[4];
        std::vector<uint8_t> rgb(size_t(prm.width) * prm.height * 3);
        for (int pass = 0; pass < 4; ++pass)
        {
            set_mode(pass);
            acc[pass].assign(size_t(prm.width) * prm.height * 4, 0.f);
            spt_cuda_reset_counters(ctx);
            auto const t0 = std::chrono::steady_clock::now();
            int const rc = spt_cuda_render(ctx, &cam, &prm, seg, rgb.data(), acc[pass].data());
            double const ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            SptTiming tm{};
            SptCounters cn{};
            spt_cuda_get_timing(ctx, &tm);
            spt_cuda_get_counters(ctx, &cn);
            std::printf("n=%u %s rc=%d wall %.1f ms trace %.3f ms segments %llu exact %llu%s%s\n", n, kModes[pass][2], rc, ms,
                        tm.trace_ms, (unsigned long long)cn.segments, (unsigned long long)cn.exact, rc ? " : " : "", rc ? spt_cuda_last_error() : "");
            std::fflush(stdout);
            if (rc)
                return 1;
        }
        for (int pass = 0; pass < 3; ++pass)
        {
            size_t diff = 0;
            for (size_t i = 0; i < acc[3].size(); ++i)
                diff += std::memcmp(&acc[pass][i], &acc[3][i], 4) != 0;
            std::printf("n=%u %s vs shipped path: %zu of %zu words differ -> %s\n", n, kModes[pass][2], diff, acc[3].size(),
                        diff ? "MISMATCH" : "bit-identical");
            bad += diff != 0;
        }
    }
    // ---- optional timing A/B on a scene file (tools/export_scene.py): argv = lib scene.bin width height spp depth reps ----
    if (argc >= 8)
    {
        FILE *fp = std::fopen(argv[2], "rb");
        uint32_t n = 0;
        SptCamera cam{};
        if (!fp || std::fread(&n, 4, 1, fp) != 1 || std::fread(&cam, sizeof(cam), 1, fp) != 1)
            return 2;
        std::vector<float> cen(n * 4), rad(n), col(n * 4), fuzz(n);
        std::vector<uint8_t> mat(n);
        if (std::fread(cen.data(), 16, n, fp) != n || std::fread(rad.data(), 4, n, fp) != n || std::fread(col.data(), 16, n, fp) != n ||
            std::fread(fuzz.data(), 4, n, fp) != n || std::fread(mat.data(), 1, n, fp) != n)
            return 2;
        std::fclose(fp);
        SptScene sc{n, cen.data(), rad.data(), mat.data(), col.data(), fuzz.data()};
        if (spt_cuda_set_scene(ctx, &sc))
            return 2;
        SptParams prm{};
        prm.width = uint32_t(std::atoi(argv[3])), prm.height = uint32_t(std::atoi(argv[4])), prm.sample_end = uint32_t(std::atoi(argv[5]));
        prm.depth = uint32_t(std::atoi(argv[6])), prm.pass_cap = 10, prm.seed = 1, prm.variant = SPT_VARIANT_MEGAKERNEL;
        int const reps = std::atoi(argv[7]);
        SptSegment seg{0, prm.height, 0, prm.width};
        std::vector<float> acc[4];
        std::vector<uint8_t> rgb(size_t(prm.width) * prm.height * 3);
        double const samples = double(prm.width) * prm.height * prm.sample_end;
        for (int rep = 0; rep < reps; ++rep)
            for (int pass = 0; pass < 4; ++pass)
            {
                set_mode(pass);
                acc[pass].assign(size_t(prm.width) * prm.height * 4, 0.f);
                if (spt_cuda_render(ctx, &cam, &prm, seg, rgb.data(), acc[pass].data()))
                {
                    std::printf("render: %s\n", spt_cuda_last_error());
                    return 1;
                }
                SptTiming tm{};
                spt_cuda_get_timing(ctx, &tm);
                std::printf("rep %d %s n=%u %ux%u spp %u depth %u: trace %.3f ms = %.1f Msamples/s\n", rep, kModes[pass][2], n,
                            prm.width, prm.height, prm.sample_end, prm.depth, tm.trace_ms, samples / tm.trace_ms * 1e-3);
                std::fflush(stdout);
            }
        for (int pass = 0; pass < 3; ++pass)
        {
            size_t diff = 0;
            for (size_t i = 0; i < acc[3].size(); ++i)
                diff += std::memcmp(&acc[pass][i], &acc[3][i], 4) != 0;
            std::printf("scene file, %s vs shipped path: %zu of %zu words differ -> %s\n", kModes[pass][2], diff, acc[3].size(),
                        diff ? "MISMATCH" : "bit-identical");
            bad += diff != 0;
        }
    }
    spt_cuda_destroy(ctx);
    std::printf(bad ? "FAIL\n" : "OK\n");
    return bad ? 1 : 0;
}
