/*
 * rt_cli.cpp — headless host for the ray-scene kernel.
 *
 * Replaces the reference's application shell for the hot path only: the GLFW
 * window + render loop (main.cpp:47-93), the wall clock that feeds the `time`
 * uniform (main.cpp:111-118, 213) and the fullscreen-quad blit (main.cpp:240-260)
 * become a loop over explicit times that calls the C ABI (include/rt_b200.h) and
 * dumps each frame as PPM / raw RGBA.  No window, no GL.
 *
 *   rt_cli [--builtin | --scene FILE] [--time T0] [--dt DT] [--frames N]
 *          [--width W] [--height H] [--depth D] [--spp S] [--strict] [--device I]
 *          [--out PREFIX] [--format ppm|raw|none] [--save-scene FILE] [--quiet]
 *
 * --builtin (default): the shader's own scene and orbit camera at each time
 *   (glsl:261-321, 334-364); time advances like the reference's clock, i.e.
 *   quantised to milliseconds: time = float(ms) / 1000.0f (main.cpp:115-117).
 * --scene FILE: an `rtscene 1` file (rt_scene_io.cpp); static scene, and its
 *   camera record (or the built-in orbit camera when it has none).
 * One JSON line per frame on stdout: time, kernel ms, rays, Mrays/s, file.
 * Exit status 0, or 1 with the library's error string on stderr (the reference
 * calls exit(EXIT_FAILURE) or prints and carries on, main.cpp:50-53, Utils.cpp:84-94).
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"

namespace {

struct Args {
    std::string scene_file, out_prefix = "frame", format = "ppm", save_scene;
    float t0 = 0.0f, dt = 1.0f / 60.0f;
    int frames = 1, width = 1280, height = 720;   // main.cpp:18-19 defaults
    int depth = -1, spp = 1, strict = 0, device = -1, quiet = 0;
};

int usage(const char* msg) {
    if (msg) fprintf(stderr, "rt_cli: %s\n", msg);
    fprintf(stderr,
            "usage: rt_cli [--builtin | --scene FILE] [--time T0] [--dt DT] [--frames N] [--width W] [--height H]\n"
            "              [--depth D] [--spp S] [--strict] [--device I] [--out PREFIX] [--format ppm|raw|none]\n"
            "              [--save-scene FILE] [--quiet]\n");
    return 2;
}

int die(const char* what) {
    fprintf(stderr, "rt_cli: %s: %s\n", what, rt_last_error_string());
    return 1;
}

}  // namespace

int main(int argc, char** argv) {
    Args a;
    for (int i = 1; i < argc; i++) {
        const std::string k = argv[i];
        auto need = [&](const char* name) -> const char* {
            if (i + 1 >= argc) { usage((std::string(name) + " needs a value").c_str()); exit(2); }
            return argv[++i];
        };
        if (k == "--builtin") a.scene_file.clear();
        else if (k == "--scene") a.scene_file = need("--scene");
        else if (k == "--time") a.t0 = strtof(need("--time"), nullptr);
        else if (k == "--dt") a.dt = strtof(need("--dt"), nullptr);
        else if (k == "--frames") a.frames = atoi(need("--frames"));
        else if (k == "--width") a.width = atoi(need("--width"));
        else if (k == "--height") a.height = atoi(need("--height"));
        else if (k == "--depth") a.depth = atoi(need("--depth"));
        else if (k == "--spp") a.spp = atoi(need("--spp"));
        else if (k == "--device") a.device = atoi(need("--device"));
        else if (k == "--out") a.out_prefix = need("--out");
        else if (k == "--format") a.format = need("--format");
        else if (k == "--save-scene") a.save_scene = need("--save-scene");
        else if (k == "--strict") a.strict = 1;
        else if (k == "--quiet") a.quiet = 1;
        else if (k == "--help" || k == "-h") { usage(nullptr); return 0; }
        else return usage(("unknown option " + k).c_str());
    }
    if (a.frames < 1 || a.width < 1 || a.height < 1 || a.spp < 1) return usage("frames, width, height and spp must be positive");
    if (a.format != "ppm" && a.format != "raw" && a.format != "none") return usage("format must be ppm, raw or none");

    rt_scene_file* sf = nullptr;
    if (!a.scene_file.empty() && rt_scene_load(a.scene_file.c_str(), &sf) != RT_OK) return die("rt_scene_load");

    rt_context* ctx = nullptr;
    if (rt_create(&ctx, a.device) != RT_OK) return die("rt_create");   // no CUDA device -> error, never a CPU fallback

    uint8_t* pixels = nullptr;
    const size_t bytes = (size_t)a.width * a.height * 4;
    if (rt_host_alloc(reinterpret_cast<void**>(&pixels), bytes) != RT_OK) return die("rt_host_alloc");

    int rc = 0;
    for (int k = 0; k < a.frames && rc == 0; k++) {
        // the reference's clock: milliseconds since start, as float seconds (main.cpp:111-118)
        const long long ms = (long long)((double)a.t0 * 1000.0 + (double)k * (double)a.dt * 1000.0 + 0.5);
        const float time = (float)ms / 1000.0f;
        rt_object objs[5];
        rt_light lights[3];
        rt_scene scene;
        rt_camera cam;
        if (sf) {
            scene = *rt_scene_file_scene(sf);
            const rt_camera* fc = rt_scene_file_camera(sf);
            if (fc) cam = *fc;
            else if (rt_builtin_camera(time, &cam) != RT_OK) { rc = die("rt_builtin_camera"); break; }
        } else {
            if (rt_builtin_scene(time, objs, lights, &scene) != RT_OK) { rc = die("rt_builtin_scene"); break; }
            if (rt_builtin_camera(time, &cam) != RT_OK) { rc = die("rt_builtin_camera"); break; }
        }
        if (a.depth >= 0) scene.max_depth = a.depth;
        if (k == 0 && !a.save_scene.empty() && rt_scene_save(a.save_scene.c_str(), &scene, &cam) != RT_OK) {
            rc = die("rt_scene_save");
            break;
        }
        rt_options opt;
        rt_options_default(&opt);
        rt_stats st;
        memset(&st, 0, sizeof(st));
        opt.spp = a.spp;
        opt.strict = a.strict;
        opt.stats = a.quiet ? nullptr : &st;   // the counters come from a second, instrumented launch
        if (rt_render_ex(ctx, &scene, &cam, pixels, a.width, a.height, &opt) != RT_OK) { rc = die("rt_render_ex"); break; }
        char path[4096] = "";
        if (a.format != "none") {
            snprintf(path, sizeof(path), "%s_%04d.%s", a.out_prefix.c_str(), k, a.format == "ppm" ? "ppm" : "rgba");
            const int wr = a.format == "ppm" ? rt_write_ppm(path, pixels, a.width, a.height)
                                             : rt_write_raw(path, pixels, a.width, a.height);
            if (wr != RT_OK) { rc = die("write frame"); break; }
        }
        if (!a.quiet)
            printf("{\"frame\": %d, \"time\": %.3f, \"width\": %d, \"height\": %d, \"depth\": %d, \"spp\": %d, "
                   "\"kernel_ms\": %.4f, \"total_ms\": %.4f, \"rays\": %llu, \"mrays_per_s\": %.1f, \"red_pixels\": %llu, "
                   "\"file\": \"%s\"}\n",
                   k, (double)time, a.width, a.height, scene.max_depth, a.spp, (double)st.kernel_ms, (double)st.total_ms,
                   (unsigned long long)st.rays, st.kernel_ms > 0.0f ? (double)st.rays / ((double)st.kernel_ms * 1e3) : 0.0,
                   (unsigned long long)st.red_pixels, path);
    }
    rt_host_free(pixels);
    rt_destroy(ctx);
    rt_scene_file_free(sf);
    return rc;
}
