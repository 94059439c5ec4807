// akari-b200 — command-line driver: `akari-b200 scene.akari` renders a scene file through libakari_b200.so.
//
// Mirrors the reference's `akari [--gpu] scene.akari` (cmd/akari.cpp:41-101 -> SceneNode::render,
// core/nodes/scene.cpp:113-165): parse, commit (BVH build + upload), render with the scene's integrator node, write
// Scene.output as 8-bit sRGB PNG. Everything goes through the C ABI of include/akari_b200.h; there is no CPU path
// (`--gpu` is accepted for command-line compatibility and is always on).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <string>
#include <utility>
#include <vector>

#include "../../include/akari_b200.h"

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

#define CHECK(call)                                                                                                    \
    do {                                                                                                               \
        if ((call) != AKB_OK) {                                                                                        \
            fprintf(stderr, "akari-b200: %s\n", akb_last_error());                                                     \
            return 1;                                                                                                  \
        }                                                                                                              \
    } while (0)

int main(int argc, char **argv) {
    const char *input = nullptr, *pfm = nullptr, *output = nullptr;
    int device = -1, gpus = 1, spp = -1, max_depth = -1, bvh = AKB_BVH_DEFAULT, split = AKB_SPLIT_TILE;
    std::vector<std::pair<std::string, std::string>> opts;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> const char * {
            if (i + 1 >= argc) {
                fprintf(stderr, "akari-b200: option %s needs a value\n", a.c_str());
                exit(1);
            }
            return argv[++i];
        };
        if (a == "--gpu" || a == "-v" || a == "--verbose") continue;
        else if (a == "-d" || a == "--device") device = atoi(next());
        else if (a == "--gpus") gpus = atoi(next());
        else if (a == "--split") split = std::string(next()) == "sample" ? AKB_SPLIT_SAMPLE : AKB_SPLIT_TILE;
        else if (a == "--option") { // --option name=value -> akb_set_option
            const std::string kv = next();
            const size_t eq = kv.find('=');
            if (eq == std::string::npos) {
                fprintf(stderr, "akari-b200: --option wants name=value\n");
                return 1;
            }
            opts.emplace_back(kv.substr(0, eq), kv.substr(eq + 1));
        }
        else if (a == "--spp") spp = atoi(next());
        else if (a == "--max-depth") max_depth = atoi(next());
        else if (a == "--pfm") pfm = next();
        else if (a == "--bvh") {
            const std::string v = next();
            bvh = v == "host" ? AKB_BVH_HOST_SAH : (v == "device" ? AKB_BVH_DEVICE : AKB_BVH_DEFAULT);
        }
        else if (a == "-o" || a == "--output") output = next();
        else if (a == "-i" || a == "--input") input = next();
        else if (a == "-h" || a == "--help") input = nullptr, i = argc;
        else input = argv[i];
    }
    if (!input) {
        printf("usage: akari-b200 [--gpu] [-d device | --gpus N [--split tile|sample]] [--option name=value] [--spp N] [--max-depth N] [--bvh host|device] [-o out.png] [--pfm out.pfm] scene.akari\n");
        return 0;
    }
    // --gpus N: the frame is split over the first N GPUs of this box (scene replicated, one film collective);
    // -d D: the single device D
    for (const auto &kv : opts) CHECK(akb_set_option(kv.first.c_str(), kv.second.c_str()));
    if (device >= 0 && gpus <= 1) CHECK(akb_init_device(device));
    else CHECK(akb_init(gpus < 1 ? 1 : gpus));
    akb_scene_file *file = nullptr;
    CHECK(akb_scene_file_load(input, &file));
    akb_render_params params;
    CHECK(akb_scene_file_params(file, &params));
    if (spp >= 0) params.spp = spp;
    if (max_depth >= 0) params.max_depth = max_depth;
    params.n_gpus = gpus, params.split = split;
    akb_scene_desc desc_copy = *akb_scene_file_desc(file);
    desc_copy.bvh_build = (uint32_t)bvh;
    const akb_scene_desc *desc = &desc_copy;
    const int W = desc->camera.width, H = desc->camera.height;
    printf("preparing scene\n");
    double t = now_s();
    akb_scene *scene = nullptr;
    CHECK(akb_scene_create(desc, &scene));
    printf("building acceleration structure took (%.3fs): %llu triangles, %llu BVH8 nodes (%s build, %.3fs), %llu lights\n", now_s() - t,
           (unsigned long long)akb_scene_info(scene, 0), (unsigned long long)akb_scene_info(scene, 1),
           akb_scene_info(scene, 7) == AKB_BVH_DEVICE ? "device" : "host", akb_scene_info(scene, 6) * 1e-6,
           (unsigned long long)akb_scene_info(scene, 3));
    std::vector<float> film((size_t)W * H * 4);
    akb_render_stats st;
    t = now_s();
    CHECK(akb_render(scene, &params, film.data(), &st));
    printf("render done took (%.3fs): %llu samples, %llu + %llu rays, %.1f Msamples/s on %d GPU(s) (device time %.2f ms, film combine %.2f ms)\n", now_s() - t,
           (unsigned long long)st.n_samples, (unsigned long long)st.n_closest_rays, (unsigned long long)st.n_shadow_rays,
           st.ms_total > 0 ? st.n_samples / (st.ms_total * 1e3) : 0.0, st.n_gpus_used, st.ms_total, st.ms_combine);
    std::vector<uint8_t> rgb((size_t)W * H * 3);
    CHECK(akb_film_to_srgb8(film.data(), W, H, rgb.data()));
    const std::string out = output ? output : akb_scene_file_output(file);
    CHECK(akb_write_png(out.c_str(), rgb.data(), W, H));
    printf("written %s\n", out.c_str());
    if (pfm) CHECK(akb_write_pfm(pfm, film.data(), W, H));
    akb_scene_destroy(scene);
    akb_scene_file_destroy(file);
    akb_shutdown();
    return 0;
}
