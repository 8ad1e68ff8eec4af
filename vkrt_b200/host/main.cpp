// main.cpp — headless counterpart of the reference's src/main.cpp:5-18: build a Context and a Raytracer (or, with
// --renderer raster, the Rasterizer the reference keeps commented out at src/main.cpp:8), loop render() until it returns
// false; optionally dump frames and print per-run statistics as one JSON line.
#include "Context.hpp"
#include "Rasterizer.hpp"
#include "Raytracer.hpp"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

int main(int argc, char** argv)
{
    Context::Options options;
    options.width = 1280, options.height = 720; // BASELINE.json config 1
    uint32_t spp = 1;
    int textureLod = VKRT_TEXTURE_LOD_ZERO;
    bool rasterizer = false;
    std::string out;
    for (int i = 1; i < argc; ++i)
    {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--width") options.width = static_cast<uint32_t>(std::atoi(next()));
        else if (a == "--height") options.height = static_cast<uint32_t>(std::atoi(next()));
        else if (a == "--frames") options.frames = std::atoi(next());
        else if (a == "--spp") spp = static_cast<uint32_t>(std::atoi(next()));
        else if (a == "--device") options.device = std::atoi(next());
        else if (a == "--script") options.keyScript = next();
        else if (a == "--models") Model::setModelsFolder(next());
        else if (a == "--out") out = next();
        else if (a == "--comm-file") options.commFile = next();
        else if (a == "--renderer") rasterizer = std::string(next()) == "raster";
        else if (a == "--texture-lod") textureLod = std::string(next()) == "cones" ? VKRT_TEXTURE_LOD_RAY_CONES : VKRT_TEXTURE_LOD_ZERO;
        else
        {
            printf("usage: %s [--width W] [--height H] [--frames N] [--spp S] [--device D] [--script W:150,Z:150] [--models DIR] "
                   "[--out frame_prefix] [--comm-file PATH] [--texture-lod zero|cones] [--renderer raytrace|raster]\n", argv[0]);
            return a == "--help" ? 0 : 2;
        }
    }
    if (const char* r = std::getenv("RANK")) options.rank = std::atoi(r);
    if (const char* w = std::getenv("WORLD_SIZE")) options.worldSize = std::atoi(w);
    if (const char* l = std::getenv("LOCAL_RANK")) options.device = std::atoi(l);
    if (options.worldSize > 1 && options.commFile.empty()) options.commFile = "/tmp/vkrt_comm_id";

    Context context(options);
    if (rasterizer)
    {
        Rasterizer graphicsApp(context); // src/main.cpp:8
        const auto t0 = std::chrono::steady_clock::now();
        uint64_t frames = 0;
        while (graphicsApp.render())
        {
            ++frames;
            if (!out.empty())
            {
                char name[512];
                snprintf(name, sizeof(name), "%s%04llu.ppm", out.c_str(), static_cast<unsigned long long>(frames - 1));
                graphicsApp.saveFrame(name);
            }
        }
        vkrt_context_synchronize(context.getDevice());
        const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("{\"renderer\": \"raster\", \"frames\": %llu, \"seconds\": %.6f, \"triangles\": %llu, \"kernel_launches\": %llu}\n",
               static_cast<unsigned long long>(frames), seconds, static_cast<unsigned long long>(graphicsApp.getBlasInfo().triangleCount),
               static_cast<unsigned long long>(graphicsApp.getStats().kernelLaunches));
        return 0;
    }
    Raytracer graphicsApp(context);
    graphicsApp.setSamplesPerPixel(spp);
    graphicsApp.setTextureLod(textureLod);

    const auto t0 = std::chrono::steady_clock::now();
    bool running = true;
    uint64_t frames = 0;
    while (running)
    {
        running = graphicsApp.render();
        if (running)
        {
            ++frames;
            if (!out.empty() && options.rank == 0)
            {
                char name[512];
                snprintf(name, sizeof(name), "%s%04llu.ppm", out.c_str(), static_cast<unsigned long long>(frames - 1));
                graphicsApp.saveFrame(name);
            }
        }
    }
    vkrt_context_synchronize(context.getDevice());
    const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const VkrtStats s = graphicsApp.getStats();
    const VkrtBlasInfo& b = graphicsApp.getBlasInfo();
    if (options.rank == 0)
    {
        printf("{\"frames\": %llu, \"seconds\": %.6f, \"rays_primary\": %llu, \"rays_shadow\": %llu, \"rays_reflection\": %llu, "
               "\"mrays_per_s\": %.3f, \"blas_triangles\": %llu, \"blas_nodes\": %llu, \"blas_build_ms\": %.3f, \"kernel_launches\": %llu}\n",
               static_cast<unsigned long long>(frames), seconds, static_cast<unsigned long long>(s.raysPrimary),
               static_cast<unsigned long long>(s.raysShadow), static_cast<unsigned long long>(s.raysReflection),
               static_cast<double>(s.raysPrimary + s.raysShadow + s.raysReflection) / seconds / 1e6,
               static_cast<unsigned long long>(b.triangleCount), static_cast<unsigned long long>(b.nodeCount), b.buildMilliseconds,
               static_cast<unsigned long long>(s.kernelLaunches));
    }
    return 0;
}
