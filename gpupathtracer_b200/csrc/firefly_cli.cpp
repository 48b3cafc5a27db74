// firefly_cli -- headless front end over the C ABI only (include/firefly.h): scene file in, image out.
// What the reference's main() does (kernel.cu:223-370) without the GLFW window: build the scene, commit, render, write.
//
//   firefly_cli SCENE.ffscene [--out FILE.png|.ppm|.exr|.pfm] [--assets DIR] [--spp N] [--bounces N] [--size W H]
//               [--integrator wavefront|megakernel|ascoded] [--builder lbvh|sah] [--nee 0|1] [--aa 0|1] [--seed N]
//               [--device N] [--repeat N]
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
                        "                   [--builder lbvh|sah] [--nee 0|1] [--aa 0|1] [--seed N] [--device N] [--repeat N]\n");
        return argc < 2 ? 1 : 0;
    }
    const char* scene_path = argv[1];
    const char *out_path = nullptr, *assets = nullptr;
    int spp = -1, bounces = -1, w = -1, h = -1, integrator = -1, builder = -1, nee = -1, aa = -1, device = 0, repeat = 1;
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
    ff_context* ctx = nullptr;
    if (ff_context_create(&ctx, device) != FF_OK) return die("ff_context_create");
    ff_scene* scene = nullptr;
    ff_scene_file desc;
    if (ff_scene_load_file(ctx, scene_path, assets, &scene, &desc) != FF_OK) { die("ff_scene_load_file"); ff_context_destroy(ctx); return 1; }
    if (spp > 0) desc.params.spp = spp;
    if (bounces >= 0) desc.params.max_bounces = bounces;
    if (w > 0 && h > 0) { desc.camera.width = w; desc.camera.height = h; }
    if (integrator >= 0) desc.params.integrator = integrator;
    if (builder >= 0) desc.builder = builder;
    if (nee >= 0) desc.params.nee = nee;
    if (aa >= 0) desc.params.aa = aa;
    if (seed >= 0) desc.params.seed = (uint64_t)seed;
    double build_ms = 0;
    if (ff_scene_commit(scene, desc.builder, &build_ms) != FF_OK) { die("ff_scene_commit"); ff_scene_destroy(scene); ff_context_destroy(ctx); return 1; }
    ff_scene_info info;
    ff_scene_get_info(scene, &info);
    const size_t npix = (size_t)desc.camera.width * desc.camera.height;
    std::vector<float> rgba(npix * 4);
    ff_stats st;
    memset(&st, 0, sizeof(st));
    for (int r = 0; r < (repeat > 0 ? repeat : 1); ++r)
        if (ff_render(scene, &desc.camera, &desc.params, rgba.data(), nullptr, &st) != FF_OK) {
            die("ff_render"); ff_scene_destroy(scene); ff_context_destroy(ctx); return 1;
        }
    int rc = 0;
    if (out_path && ff_write_image(out_path, rgba.data(), desc.camera.width, desc.camera.height) != FF_OK) rc = die("ff_write_image");
    printf("{\"scene\": \"%s\", \"geometries\": %d, \"triangles\": %llu, \"bvh_nodes\": %llu, \"builder\": \"%s\", \"build_ms\": %.4f, "
           "\"width\": %d, \"height\": %d, \"spp\": %d, \"max_bounces\": %d, \"nee\": %d, \"integrator\": %d, \"gpu_ms\": %.4f, \"trace_ms\": %.4f, "
           "\"rays\": %llu, \"shadow_rays\": %llu, \"paths\": %llu, \"mrays_per_s\": %.2f, \"msamples_per_s\": %.2f, \"kernel_launches\": %u, "
           "\"out\": \"%s\"}\n",
           scene_path, desc.n_geometries, (unsigned long long)info.n_triangles, (unsigned long long)info.n_nodes,
           desc.builder == FF_BUILDER_SAH ? "sah" : "lbvh", build_ms, desc.camera.width, desc.camera.height, desc.params.spp,
           desc.params.max_bounces, desc.params.nee, desc.params.integrator, st.gpu_ms, st.trace_ms, (unsigned long long)st.rays,
           (unsigned long long)st.shadow_rays, (unsigned long long)st.paths,
           st.gpu_ms > 0 ? (double)(st.rays + st.shadow_rays) / st.gpu_ms / 1e3 : 0.0, st.gpu_ms > 0 ? (double)st.paths / st.gpu_ms / 1e3 : 0.0,
           st.kernel_launches, out_path ? out_path : "");
    ff_scene_destroy(scene);
    ff_context_destroy(ctx);
    return rc;
}
