// firefly_cli -- headless front end over the C ABI only (include/firefly.h): scene file in, image out.
// What the reference's main() does (kernel.cu:223-370) without the GLFW window: build the scene, commit, render, write.
//
//   firefly_cli SCENE.ffscene [--out FILE.png|.ppm|.exr|.pfm] [--assets DIR] [--spp N] [--bounces N] [--size W H]
//               [--integrator wavefront|megakernel|ascoded] [--builder lbvh|sah] [--nee 0|1] [--aa 0|1] [--seed N]
//               [--device N] [--repeat N] [--gpus N] [--split samples|tiles] [--reduce nccl|peer]
// --gpus N renders on devices 0..N-1 through the library's multi-GPU group (ff_group_create: scene replicated, samples or
// tiles split, one framebuffer reduce per frame over NVLink); the reference's main() has only a dead multi_gpu flag
// (utilities.h:484-487).
// Prints one JSON line with the scene / build / render figures on stdout; errors go to stderr, exit status 1.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/firefly.h"

static int die(const char* what) {
    fprintf(stderr, "firefly_cli: %s: %s\n", what, ff_last_error());
    return 1;
}

int main(int argc, char** argv) {
    if (argc < 2 || !strcmp(argv[1], "-h") || !strcmp(argv[1], "--help")) {
        fprintf(stderr, "usage: firefly_cli SCENE [--out FILE] [--assets DIR] [--spp N] [--bounces N] [--size W H] [--integrator NAME]\n"
                        "                   [--builder lbvh|sah] [--nee 0|1] [--aa 0|1] [--seed N] [--device N] [--repeat N]\n"
                        "                   [--gpus N] [--split samples|tiles] [--reduce nccl|peer]\n");
        return argc < 2 ? 1 : 0;
    }
    const char* scene_path = argv[1];
    const char *out_path = nullptr, *assets = nullptr;
    int spp = -1, bounces = -1, w = -1, h = -1, integrator = -1, builder = -1, nee = -1, aa = -1, device = 0, repeat = 1;
    int gpus = 0, split = FF_SPLIT_SAMPLES, reduce = FF_REDUCE_PEER;
    long long seed = -1;
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        auto need = [&](int n) { if (i + n >= argc) { fprintf(stderr, "firefly_cli: %s needs %d argument(s)\n", a.c_str(), n); exit(1); } };
        if (a == "--out") { need(1); out_path = argv[++i]; }
        else if (a == "--assets") { need(1); assets = argv[++i]; }
        else if (a == "--spp") { need(1); spp = atoi(argv[++i]); }
        else if (a == "--bounces") { need(1); bounces = atoi(argv[++i]); }
        else if (a == "--size") { need(2); w = atoi(argv[++i]); h = atoi(argv[++i]); }
        else if (a == "--nee") { need(1); nee = atoi(argv[++i]); }
        else if (a == "--aa") { need(1); aa = atoi(argv[++i]); }
        else if (a == "--seed") { need(1); seed = atoll(argv[++i]); }
        else if (a == "--device") { need(1); device = atoi(argv[++i]); }
        else if (a == "--repeat") { need(1); repeat = atoi(argv[++i]); }
        else if (a == "--gpus") { need(1); gpus = atoi(argv[++i]); if (gpus < 1 || gpus > 8) { fprintf(stderr, "firefly_cli: --gpus 1..8\n"); return 1; } }
        else if (a == "--split") { need(1); const std::string v = argv[++i]; split = v == "tiles" ? FF_SPLIT_TILES : FF_SPLIT_SAMPLES; }
        else if (a == "--reduce") { need(1); const std::string v = argv[++i]; reduce = v == "nccl" ? FF_REDUCE_NCCL : FF_REDUCE_PEER; }
        else if (a == "--integrator") {
            need(1);
            const std::string v = argv[++i];
            integrator = v == "wavefront" ? FF_INTEGRATOR_WAVEFRONT : v == "megakernel" ? FF_INTEGRATOR_MEGAKERNEL : v == "ascoded" ? FF_INTEGRATOR_AS_CODED : -2;
            if (integrator == -2) { fprintf(stderr, "firefly_cli: unknown integrator %s\n", v.c_str()); return 1; }
        } else if (a == "--builder") {
            need(1);
            const std::string v = argv[++i];
            builder = v == "lbvh" ? FF_BUILDER_LBVH : v == "sah" ? FF_BUILDER_SAH : -2;
            if (builder == -2) { fprintf(stderr, "firefly_cli: unknown builder %s\n", v.c_str()); return 1; }
        } else { fprintf(stderr, "firefly_cli: unknown option %s\n", a.c_str()); return 1; }
    }
    // one context per GPU: the group's own with --gpus, else a single one on --device (a group of one either way)
    ff_group* group = nullptr;
    ff_context* solo = nullptr;
    const int n_ctx = gpus > 0 ? gpus : 1;
    if (gpus > 0) {
        std::vector<int> ids(gpus);
        for (int i = 0; i < gpus; ++i) ids[i] = i;
        if (ff_group_create(ids.data(), gpus, &group) != FF_OK) return die("ff_group_create");
    } else {
        if (ff_context_create(&solo, device) != FF_OK) return die("ff_context_create");
        if (ff_group_create_rank(solo, nullptr, 0, 1, &group) != FF_OK) { die("ff_group_create_rank"); ff_context_destroy(solo); return 1; }
    }
    std::vector<ff_scene*> scenes(n_ctx, nullptr);
    ff_scene_file desc;
    ff_scene_info info;
    double build_ms = 0;
    auto cleanup = [&]() {
        for (ff_scene* s : scenes) ff_scene_destroy(s);
        ff_group_destroy(group);
        if (solo) ff_context_destroy(solo);
    };
    for (int i = 0; i < n_ctx; ++i) {   // scene + BVH replicated: every GPU loads and builds its own copy
        if (ff_scene_load_file(ff_group_context(group, i), scene_path, assets, &scenes[i], &desc) != FF_OK) { die("ff_scene_load_file"); cleanup(); return 1; }
        if (builder >= 0) desc.builder = builder;
        double ms = 0;
        if (ff_scene_commit(scenes[i], desc.builder, &ms) != FF_OK) { die("ff_scene_commit"); cleanup(); return 1; }
        if (i == 0) { build_ms = ms; ff_scene_get_info(scenes[0], &info); }
    }
    if (spp > 0) desc.params.spp = spp;
    if (bounces >= 0) desc.params.max_bounces = bounces;
    if (w > 0 && h > 0) { desc.camera.width = w; desc.camera.height = h; }
    if (integrator >= 0) desc.params.integrator = integrator;
    if (nee >= 0) desc.params.nee = nee;
    if (aa >= 0) desc.params.aa = aa;
    if (seed >= 0) desc.params.seed = (uint64_t)seed;
    const size_t npix = (size_t)desc.camera.width * desc.camera.height;
    std::vector<float> rgba(npix * 4);
    std::vector<ff_stats> st(n_ctx);
    memset(st.data(), 0, sizeof(ff_stats) * n_ctx);
    double reduce_ms = 0;
    const bool plain = (desc.params.integrator & 0xff) != FF_INTEGRATOR_WAVEFRONT && (desc.params.integrator & 0xff) != FF_INTEGRATOR_MEGAKERNEL;
    for (int r = 0; r < (repeat > 0 ? repeat : 1); ++r) {
        // as-coded frame / feature buffers are one primary ray per pixel: rendered on the first GPU
        const int rc = plain ? ff_render(scenes[0], &desc.camera, &desc.params, rgba.data(), nullptr, &st[0])
                             : ff_group_render(group, scenes.data(), &desc.camera, &desc.params, split, reduce, rgba.data(), nullptr, st.data(), &reduce_ms);
        if (rc != FF_OK) { die(plain ? "ff_render" : "ff_group_render"); cleanup(); return 1; }
    }
    int rc = 0;
    if (out_path && ff_write_image(out_path, rgba.data(), desc.camera.width, desc.camera.height) != FF_OK) rc = die("ff_write_image");
    double gpu_ms = 0, trace_ms = 0;
    unsigned long long rays = 0, shadow = 0, paths = 0;
    unsigned launches = 0;
    for (const ff_stats& q : st) {   // GPUs run side by side: times are the slowest GPU's, counts are sums
        gpu_ms = q.gpu_ms > gpu_ms ? q.gpu_ms : gpu_ms; trace_ms = q.trace_ms > trace_ms ? q.trace_ms : trace_ms;
        rays += q.rays; shadow += q.shadow_rays; paths += q.paths; launches += q.kernel_launches;
    }
    printf("{\"scene\": \"%s\", \"geometries\": %d, \"triangles\": %llu, \"bvh_nodes\": %llu, \"builder\": \"%s\", \"build_ms\": %.4f, "
           "\"width\": %d, \"height\": %d, \"spp\": %d, \"max_bounces\": %d, \"nee\": %d, \"integrator\": %d, \"gpus\": %d, \"gpu_ms\": %.4f, "
           "\"trace_ms\": %.4f, \"reduce_ms\": %.4f, \"rays\": %llu, \"shadow_rays\": %llu, \"paths\": %llu, \"mrays_per_s\": %.2f, "
           "\"msamples_per_s\": %.2f, \"kernel_launches\": %u, \"out\": \"%s\"}\n",
           scene_path, desc.n_geometries, (unsigned long long)info.n_triangles, (unsigned long long)info.n_nodes,
           desc.builder == FF_BUILDER_SAH ? "sah" : "lbvh", build_ms, desc.camera.width, desc.camera.height, desc.params.spp,
           desc.params.max_bounces, desc.params.nee, desc.params.integrator, n_ctx, gpu_ms, trace_ms, reduce_ms, rays, shadow, paths,
           gpu_ms > 0 ? (double)(rays + shadow) / gpu_ms / 1e3 : 0.0, gpu_ms > 0 ? (double)paths / gpu_ms / 1e3 : 0.0, launches,
           out_path ? out_path : "");
    cleanup();
    return rc;
}
