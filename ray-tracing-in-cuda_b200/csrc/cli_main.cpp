// cli_main.cpp — `parallel_compute [-f scene.json]`: the reference's command line (gpu-version/main.cu:448-526,
// jsonmain) on top of the C ABI.  Same contract as the reference binary that dna.py / blue.py / blue2.py shell out
// to (dna.py:107-112): default scene "sample_scene.json", writes ./main.ppm (gamma 2, P3 text) and the scene's
// "output_file" PNG (no gamma), exit code 0 on success and 99 on any failure (rtweekend.cuh:43-53).
// Extra flags (not in the reference): --spp N --width W --height H --depth D --device K --seed S --no-ppm --quiet.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "rt.h"

static int fail(rt_ctx *ctx, const char *what) {
    fprintf(stderr, "parallel_compute: %s: %s\n", what, rt_last_error(ctx));
    if (ctx) rt_destroy(ctx);
    return 99;
}

// --gpus N: the samples of the image are dealt to N devices in whole sample chunks (rt_group_*, one NCCL gather inside the library);
// the float image is bit-identical to the one-GPU run.  Replaces blue.py's "one process per GPU" (gpu-version/blue.py:23-32).
static int main_group(const std::string &file, int spp, int width, int height, int depth, int device, int gpus, int write_ppm, int quiet,
                      unsigned seed) {
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<int> devs;
    for (int i = 0; i < gpus; ++i) devs.push_back(device + i);
    rt_group *g = nullptr;
    if (rt_group_create(&g, devs.data(), gpus)) { fprintf(stderr, "parallel_compute: rt_group_create: %s\n", rt_last_error(nullptr)); return 99; }
    auto gfail = [&](const char *what) { fprintf(stderr, "parallel_compute: %s: %s\n", what, rt_group_last_error(g)); rt_group_destroy(g); return 99; };
    if (rt_group_scene_upload_json(g, file.c_str())) return gfail("rt_group_scene_upload_json");
    if ((spp > 0 || width > 0 || height > 0 || depth > 0) && rt_group_scene_override(g, width, height, spp, depth)) return gfail("rt_group_scene_override");
    rt_scene_info info;
    if (rt_scene_info_get(rt_group_ctx(g, 0), &info)) return gfail("rt_scene_info_get");
    rt_render_params prm{};
    prm.seed = seed;
    if (rt_group_render(g, &prm)) return gfail("rt_group_render");
    const size_t n = (size_t)info.width * info.height * 3;
    std::vector<unsigned char> png(n), ppm(n);
    rt_stats st{};
    if (rt_group_download(g, nullptr, 0, png.data(), info.samples_per_pixel, 0, &st)) return gfail("rt_group_download");
    if (write_ppm) {
        if (rt_group_download(g, nullptr, 0, ppm.data(), info.samples_per_pixel, 1, nullptr)) return gfail("rt_group_download");
        if (rt_write_ppm("main.ppm", ppm.data(), info.width, info.height)) { fprintf(stderr, "parallel_compute: cannot write main.ppm\n"); rt_group_destroy(g); return 99; }
    }
    if (rt_write_png(info.output_file, png.data(), info.width, info.height)) {
        fprintf(stderr, "parallel_compute: cannot write %s\n", info.output_file);
        rt_group_destroy(g);
        return 99;
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (!quiet)
        fprintf(stderr, "parallel_compute: %s %dx%d %d spp depth %d on %d GPUs: %lld rays, %.1f ms on device, %.3f s total -> %s\n", file.c_str(),
                info.width, info.height, info.samples_per_pixel, info.max_depth, gpus, (long long)st.rays, st.ms_total, sec, info.output_file);
    rt_group_destroy(g);
    return 0;
}

int main(int argc, char **argv) {
    std::string file = "sample_scene.json";            // main.cu:454
    int spp = 0, width = 0, height = 0, depth = 0, device = 0, write_ppm = 1, quiet = 0, gpus = 1;
    unsigned seed = 0x2023;
    for (int i = 1; i < argc; ++i) {
        auto next = [&](const char *flag) -> const char * {
            if (i + 1 >= argc) { fprintf(stderr, "parallel_compute: %s needs a value\n", flag); exit(99); }
            return argv[++i];
        };
        if (!strcmp(argv[i], "-f")) file = next("-f");                       // main.cu:455-460
        else if (!strcmp(argv[i], "--spp")) spp = atoi(next("--spp"));
        else if (!strcmp(argv[i], "--width")) width = atoi(next("--width"));
        else if (!strcmp(argv[i], "--height")) height = atoi(next("--height"));
        else if (!strcmp(argv[i], "--depth")) depth = atoi(next("--depth"));
        else if (!strcmp(argv[i], "--device")) device = atoi(next("--device"));
        else if (!strcmp(argv[i], "--gpus")) gpus = atoi(next("--gpus"));     // devices device .. device+gpus-1 render one image together
        else if (!strcmp(argv[i], "--seed")) seed = (unsigned)strtoul(next("--seed"), nullptr, 0);
        else if (!strcmp(argv[i], "--no-ppm")) write_ppm = 0;
        else if (!strcmp(argv[i], "--quiet")) quiet = 1;
        else { fprintf(stderr, "parallel_compute: unknown argument %s\n", argv[i]); return 99; }
    }
    const auto t0 = std::chrono::steady_clock::now();
    if (gpus > 1) return main_group(file, spp, width, height, depth, device, gpus, write_ppm, quiet, seed);
    rt_ctx *ctx = nullptr;
    if (rt_create(&ctx, device, nullptr)) return fail(nullptr, "rt_create");
    if (rt_scene_upload_json(ctx, file.c_str())) return fail(ctx, "rt_scene_upload_json");
    if ((spp > 0 || width > 0 || height > 0 || depth > 0) && rt_scene_override(ctx, width, height, spp, depth)) return fail(ctx, "rt_scene_override");
    rt_scene_info info;
    if (rt_scene_info_get(ctx, &info)) return fail(ctx, "rt_scene_info_get");
    rt_render_params prm{};
    prm.seed = seed;
    if (rt_render(ctx, &prm)) return fail(ctx, "rt_render");
    const size_t n = (size_t)info.width * info.height * 3;
    std::vector<unsigned char> png(n), ppm(n);
    rt_stats st{};
    if (rt_download(ctx, nullptr, 0, png.data(), info.samples_per_pixel, 0, &st)) return fail(ctx, "rt_download");
    if (write_ppm) {                                                         // output_image(..., "main.ppm"), main.cu:512
        if (rt_download(ctx, nullptr, 0, ppm.data(), info.samples_per_pixel, 1, nullptr)) return fail(ctx, "rt_download");
        if (rt_write_ppm("main.ppm", ppm.data(), info.width, info.height)) { fprintf(stderr, "parallel_compute: cannot write main.ppm\n"); rt_destroy(ctx); return 99; }
    }
    // write_image(..., data["output_file"]), main.cu:514 — the directory must exist, as in the reference
    if (rt_write_png(info.output_file, png.data(), info.width, info.height)) {
        fprintf(stderr, "parallel_compute: cannot write %s\n", info.output_file);
        rt_destroy(ctx);
        return 99;
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (!quiet)
        fprintf(stderr, "parallel_compute: %s %dx%d %d spp depth %d: %lld rays, %.1f ms on device, %.3f s total -> %s\n", file.c_str(),
                info.width, info.height, info.samples_per_pixel, info.max_depth, (long long)st.rays, st.ms_total, sec, info.output_file);
    rt_destroy(ctx);
    return 0;
}
