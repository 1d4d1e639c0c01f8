// spt_headless -- the reference's TracePaths() without a window: scene -> viewMatrix -> RenderSegmentCuda over the
// whole image -> io::SaveImage (BMP).  Stands in for Main.cpp:3-6 + Renderer.hpp:316-357 (MainLoop) with the
// GL/GLFW parts left out (no graphics stack on this box) and the compile-time constants of Globals.hpp:8-18
// turned into flags.
//
//   spt_headless [--scene reference|random|metalglass|lattice|<file.sptscene>] [--dump-scene file.sptscene] [--seed S] [--width W] [--height H]
//                [--spp N] [--depth D] [--tracer recursive|task|wavefront] [--tiles T] [--out file.bmp]
//                [--progressive P]   P passes over the samples on a persistent accumulator; prints k/P and rewrites
//                                    the image after each (the reference's live preview, Renderer.hpp:157-164)
//                [--pfm file.pfm]    also write the linear float image (needs --progressive >= 1)
//                [--hdr file.hdr]    the same image as Radiance RGBE (what stbi_write_hdr serialises)
//                [--repeat R]        render R times in this process and print one Duration line each (the first render of a
//                                    process also allocates the staging buffers and loads the kernels)
//                [--checkpoint file] resume from the accumulators in `file` when it was written for THIS render (size, depth,
//                                    seed, scene, camera, tracer), rewrite it after every pass
//                [--gpus N]          one render spread over CUDA devices 0..N-1 by the library (spt_cuda_create_multi)
//                [--devices a,b,..]  the same with an explicit list of CUDA ordinals (an ordinal may repeat: several internal
//                                    contexts on one GPU)
//                [--staging-mib M]   first-hit staging budget of split rendering per device (default 8192)
//                [--grid 1]          SPT_FLAG_GRID: closest hits through the library's uniform grid instead of the linear loop
//                                    (opt-in, same image; for long sphere lists such as --scene lattice)
#include "../../include/spt_host/Scene.hpp"
#include "../../include/spt_host/SceneFile.hpp"
#include "../../include/spt_host/SceneGenerators.hpp"
#include "../../include/spt_host/Image.hpp"
#include "../../include/CudaPathTracer.hpp"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

int main(int argc, char **argv)
{
    std::string scene = "reference", tracer = "recursive", out;
    std::string pfm, hdr, checkpoint, dumpScene, deviceList;
    uint32_t seed = 1, width = 1440, height = 1440, spp = 100, depth = 10, tiles = 1, progressive = 0, gpus = 1, stagingMib = 0, repeat = 1, grid = 0;
    for (int i = 1; i + 1 < argc; i += 2)
    {
        std::string const k = argv[i], v = argv[i + 1];
        if (k == "--scene") scene = v;
        else if (k == "--seed") seed = uint32_t(std::stoul(v));
        else if (k == "--width") width = uint32_t(std::stoul(v));
        else if (k == "--height") height = uint32_t(std::stoul(v));
        else if (k == "--spp") spp = uint32_t(std::stoul(v));
        else if (k == "--depth") depth = uint32_t(std::stoul(v));
        else if (k == "--tracer") tracer = v;
        else if (k == "--tiles") tiles = uint32_t(std::stoul(v));
        else if (k == "--out") out = v;
        else if (k == "--progressive") progressive = uint32_t(std::stoul(v));
        else if (k == "--pfm") pfm = v;
        else if (k == "--hdr") hdr = v;
        else if (k == "--grid") grid = uint32_t(std::stoul(v));
        else if (k == "--repeat") repeat = std::max<uint32_t>(1u, uint32_t(std::stoul(v)));
        else if (k == "--dump-scene") dumpScene = v;
        else if (k == "--checkpoint") checkpoint = v;
        else if (k == "--gpus") gpus = uint32_t(std::stoul(v));
        else if (k == "--devices") deviceList = v;
        else if (k == "--staging-mib") stagingMib = uint32_t(std::stoul(v));
        else { std::cerr << "unknown flag " << k << "\n"; return 2; }
    }
    SetRenderConfig(width, height, spp, depth);
    UpdateViewMatrix(); // Renderer.hpp:321
    if (scene == "reference") InitSpheres(seed);            // eSceneGenerator::REFERENCE
    else if (scene == "random") GenerateSpheres(seed);      // eSceneGenerator::RANDOM
    else if (scene == "metalglass") GenerateMetalGlassSpheres(seed);
    else if (scene == "lattice") GenerateLatticeSpheres(seed);
    else if (!io::LoadScene(scene.c_str())) { std::cerr << "unknown scene / not a scene file: " << scene << "\n"; return 2; }
    if (!dumpScene.empty() && !io::SaveScene(dumpScene.c_str())) { std::cerr << "cannot write " << dumpScene << "\n"; return 1; }
    spt_cuda::SetSeed(seed);
    spt_cuda::InvalidateScene();
    if (!deviceList.empty())
    {
        std::vector<int> devices;
        for (size_t pos = 0; pos <= deviceList.size();)
        {
            size_t const comma = std::min(deviceList.find(',', pos), deviceList.size());
            devices.push_back(std::stoi(deviceList.substr(pos, comma - pos)));
            pos = comma + 1;
        }
        spt_cuda::SetDevices(devices);
    }
    else if (gpus > 1)
    {
        std::vector<int> devices(gpus);
        for (uint32_t k = 0; k < gpus; ++k)
            devices[k] = int(k);
        spt_cuda::SetDevices(devices);
    }
    if (stagingMib)
        spt_cuda::SetStagingLimit(uint64_t(stagingMib) << 20);
    if (grid)
        spt_cuda::SetFlags(SPT_FLAG_GRID);

    if ((!pfm.empty() || !hdr.empty() || !checkpoint.empty()) && progressive == 0)
        progressive = 1;
    // context(s) and scene upload are set-up, as window and GL are in the reference (Renderer.hpp:316-333): not in Duration
    auto const tSetup = std::chrono::high_resolution_clock::now();
    try
    {
        spt_cuda::Acquire();
    }
    catch (std::exception const &e)
    {
        std::cerr << e.what() << "\n";
        return 1;
    }
    std::cout << "Setup: " << std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - tSetup).count()
              << " ms\n";
    for (uint32_t rep = 0; rep < repeat; ++rep)
    {
    auto const t0 = std::chrono::high_resolution_clock::now();
    try
    {
        if (progressive > 0)
        {
            uint32_t const passCap = tracer == "recursive" ? 0u : 10u;
            uint32_t const variant = tracer == "wavefront" ? SPT_VARIANT_WAVEFRONT : SPT_VARIANT_MEGAKERNEL;
            std::vector<float> accum(size_t(width) * height * 4u), resume;
            io::AccumHeader h{};
            uint32_t first = 0;
            uint64_t const identity = io::RenderIdentity(passCap); // scene, camera, tracer semantics
            if (!checkpoint.empty() && io::LoadAccum(checkpoint.c_str(), width, height, h, resume, identity) && h.depth == depth &&
                h.seed == seed && h.samplesDone <= spp)
            {
                first = h.samplesDone;
                std::cout << "Resuming at sample " << first << " of " << spp << std::endl;
            }
            else
                resume.clear();
            spt_cuda::RenderImageProgressiveCuda(
                progressive, passCap, variant,
                [&](uint32_t k, uint32_t n, uint32_t samplesDone) {
                    std::cout << k << "/" << n << "  (" << samplesDone << " spp)" << std::endl; // Renderer.hpp:285
                    if (n > 1 && !io::SaveImage(out.empty() ? nullptr : out.c_str()))               // the live preview, as a file
                        std::cerr << "cannot write image\n";
                    // after EVERY pass: an interrupted render resumes from the last completed one
                    if (!checkpoint.empty() && !io::SaveAccum(checkpoint.c_str(), width, height, samplesDone, depth, seed, accum.data(), identity))
                        std::cerr << "cannot write " << checkpoint << "\n";
                },
                first, resume.empty() ? nullptr : resume.data(), accum.data(), !checkpoint.empty());
            if (!pfm.empty() && !io::WritePfm(pfm.c_str(), width, height, accum.data()))
                std::cerr << "cannot write " << pfm << "\n";
            if (!hdr.empty() && !io::WriteHdr(hdr.c_str(), width, height, accum.data()))
                std::cerr << "cannot write " << hdr << "\n";
        }
        else
        {
        // tiles == 1: RenderImage's single {0, H, 0, W} call (Renderer.hpp:304-308); tiles > 1: the
        // threadCount x threadCount grid of RenderImageParallelMain (Renderer.hpp:264-274), one call per tile
        uint32_t const sw = (g_width + tiles - 1) / tiles, sh = (g_height + tiles - 1) / tiles;
        uint32_t done = 0;
        for (uint32_t j = 0; j < tiles; ++j)
            for (uint32_t i = 0; i < tiles; ++i)
            {
                // ceil-sized tiles: the last ones may be short or empty (e.g. width 10, 8 tiles): clamp BOTH ends
                RenderSegmentData const seg{std::min(sh * j, g_height), std::min(sh * (j + 1), g_height), std::min(sw * i, g_width),
                                            std::min(sw * (i + 1), g_width)};
                if (tiles > 1)
                    std::cout << ++done << "/" << tiles * tiles << std::endl; // Renderer.hpp:285
                if (tracer == "recursive") RenderSegmentCuda(seg);
                else if (tracer == "task") RenderSegmentTaskCuda(seg);
                else if (tracer == "wavefront") RenderSegmentCudaWavefront(seg);
                else { std::cerr << "unknown tracer " << tracer << "\n"; return 2; }
            }
        }
    }
    catch (std::exception const &e)
    {
        std::cerr << e.what() << "\n";
        return 1;
    }
    auto const dt = std::chrono::high_resolution_clock::now() - t0;
    std::cout << "Duration: " << std::chrono::duration_cast<std::chrono::milliseconds>(dt).count() << " ms\n"; // Renderer.hpp:297
    double const sec = std::chrono::duration<double>(dt).count();
    std::cout << "Spheres: " << g_sphereNumber << "  Msamples/s: " << double(width) * height * spp / sec / 1e6 << "\n";
    }
    if (!io::SaveImage(out.empty() ? nullptr : out.c_str()))
    {
        std::cerr << "cannot write image\n";
        return 1;
    }
    spt_cuda::Shutdown();
    return 0;
}
