/*
 * oracle/ref_shim/ref_integration.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * The drop-in, demonstrated inside the reference's own translation unit: the UNMODIFIED reference headers
 * (scene generators, globals, io::SaveImage, the CPU tracers) plus include/CudaPathTracer.hpp, with the one-line
 * change INTEGRATION.md describes -- RenderJob's call (Renderer.hpp:245-252) goes to RenderSegmentCuda.
 * Renders the same scene with the reference's RenderSegment on the CPU and with RenderSegmentCuda on the GPU into
 * the reference's g_data, saves both with the reference's io::SaveImage path (stbi_write_bmp) and prints the
 * PSNR between the two 8-bit images.  Built by oracle/Makefile (target integration) into oracle/_ref/.
 *
 *   ref_integration <scene: 0 reference | 1 random> <width> <height> <spp> <depth> <out_prefix>
 */
#include "ref_prelude.hpp"

#include <CudaPathTracer.hpp>

#include <cstdio>
#include <iostream>

static void RunOnFreshThread(uint32_t seed, void (*fn)())
{
    g_sptNextSeed.store(seed);
    std::thread t(fn);
    t.join();
}

static RenderSegmentData g_fullImage;

int main(int argc, char **argv)
{
    if (argc < 7)
    {
        std::cerr << "usage: ref_integration scene width height spp depth out_prefix\n";
        return 2;
    }
    int const scene = atoi(argv[1]);
    g_sptCfg.width = uint32_t(atoi(argv[2]));
    g_sptCfg.height = uint32_t(atoi(argv[3]));
    g_sptCfg.samples = uint32_t(atoi(argv[4]));
    g_sptCfg.bounces = uint32_t(atoi(argv[5]));
    g_sptCfg.size = g_sptCfg.width * g_sptCfg.height * 3u;
    std::string const prefix = argv[6];
    std::vector<unsigned char> cpu(g_sptCfg.size), gpu(g_sptCfg.size);

    viewMatrix = Transpose(math::CreateCameraBasisMatrix(eyePos, lookAt, upDir)); // Renderer.hpp:321
    if (scene == 0)
        RunOnFreshThread(1, [] { InitSpheres(); });
    else
        RunOnFreshThread(1, [] { GenerateSpheres(); });
    g_fullImage = RenderSegmentData{0, g_sptCfg.height, 0, g_sptCfg.width};

    // the reference's own tracer on the CPU
    g_sptCfg.data = cpu.data();
    RunOnFreshThread(7, [] { RenderSegment(g_fullImage); });
    stbi_write_bmp((prefix + "_cpu.bmp").c_str(), int(g_sptCfg.width), int(g_sptCfg.height), 3, cpu.data());

    // the drop-in: same call shape, same globals, same g_data
    g_sptCfg.data = gpu.data();
    try
    {
        RenderSegmentCuda(g_fullImage);
    }
    catch (std::exception const &e)
    {
        std::cerr << e.what() << "\n";
        return 1;
    }
    stbi_write_bmp((prefix + "_cuda.bmp").c_str(), int(g_sptCfg.width), int(g_sptCfg.height), 3, gpu.data());
    spt_cuda::Shutdown();

    double mse = 0;
    for (uint32_t i = 0; i < g_sptCfg.size; ++i)
    {
        double const d = (double(cpu[i]) - double(gpu[i])) / 255.0;
        mse += d * d;
    }
    mse /= g_sptCfg.size;
    std::printf("spheres %u  psnr8bit %.2f dB\n", g_sphereNumber, mse > 0 ? 10.0 * std::log10(1.0 / mse) : 99.0);
    return 0;
}
