// raytracer_main.cpp — `raytracer scene.txt` -> scene.ppm, the reference CLI
// (main.cpp:60-99, 604-656) around the CUDA render core.
//
//   parser output -> rt_upload_scene -> rt_render -> rt_format_ppm_device -> file
//   --gpus N : the same through rt_upload_scene_multi / rt_render_multi / rt_format_ppm_multi (N GPUs of
//              this node, one process; the file is byte-identical for every N)
//   --frames K : render the frame K times (the 3rd and later are CUDA-graph replays); --stats prints the last
//
// Exit status: the reference returns 0 on every path it survives and dies with
// SIGABRT (uncaught exception) on malformed input; we print the same messages
// and return 0 / 1 respectively.  There is no CPU fallback: without a CUDA
// device the render call fails and so does the program.
#include "rt.h"
#include "rt_host.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char* argv[])
{
    if (argc <= 1) {
        // main.cpp:653
        std::printf("Error: Incorrect number of arguments in input file. Please follow this formate: imsize width height\n");
        return 0;
    }
    bool verbose = false;
    int max_depth = 4, gpus = 1, frames = 1;
    const char* scene_path = nullptr;
    for (int i = 1; i < argc; i++) {
        if (!std::strcmp(argv[i], "--stats")) verbose = true;
        else if (!std::strcmp(argv[i], "--depth") && i + 1 < argc) max_depth = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--frames") && i + 1 < argc) frames = std::atoi(argv[++i]);
        else if (!scene_path) scene_path = argv[i];
    }
    if (!scene_path) return 0;

    double t0 = now_s();
    RtSceneDesc* desc = nullptr;
    int rc = rt_host_parse_scene_file(scene_path, nullptr, &desc);
    if (rc == RT_E_IO) { std::printf("%s\n", rt_host_last_error()); return 0; }   // main.cpp:567-568
    if (rc != RT_OK) {
        const char* msg = rt_host_last_error();
        if (!std::strncmp(msg, "Error: Requires command", 23)) { std::printf("%s\n", msg); return 0; }   // 574-602
        std::fprintf(stderr, "%s\n", msg);
        return 1;
    }
    double t1 = now_s();

    RtRenderParams params{};
    params.max_depth = max_depth; params.epsilon = 1.0e-3f; params.accel = RT_ACCEL_AUTO;
    params.rank = 0; params.world_size = 1; params.tile_rows = gpus > 1 ? 0 : 8;
    RtStats stats{};
    RtScene* scene = nullptr;
    RtMultiScene* multi = nullptr;
    const char* text = nullptr; size_t text_len = 0;
    double t2, t3;
    if (gpus > 1) {
        rc = rt_upload_scene_multi(desc, gpus, nullptr, &multi);
        if (rc != RT_OK) { std::fprintf(stderr, "rt_upload_scene_multi: %s\n", rt_last_error()); return 1; }
        t2 = now_s();
        for (int k = 0; k < (frames > 0 ? frames : 1); k++) {
            if (k == frames - 1) t2 = frames > 1 ? now_s() : t2;
            rc = rt_render_multi(multi, &params, &stats);
            if (rc != RT_OK) { std::fprintf(stderr, "rt_render_multi: %s\n", rt_last_error()); return 1; }
        }
        t3 = now_s();
        rc = rt_format_ppm_multi(multi, &text, &text_len);
        if (rc != RT_OK) { std::fprintf(stderr, "rt_format_ppm_multi: %s\n", rt_last_error()); return 1; }
    } else {
        rc = rt_upload_scene(desc, &scene);
        if (rc != RT_OK) { std::fprintf(stderr, "rt_upload_scene: %s\n", rt_last_error()); return 1; }
        t2 = now_s();
        for (int k = 0; k < (frames > 0 ? frames : 1); k++) {
            if (k == frames - 1) t2 = frames > 1 ? now_s() : t2;
            rc = rt_render(scene, &params, &stats);
            if (rc != RT_OK) { std::fprintf(stderr, "rt_render: %s\n", rt_last_error()); return 1; }
        }
        t3 = now_s();
        // the P3 text is formatted on the device (rt_format_ppm_device); the host only writes it out
        rc = rt_format_ppm_device(scene, nullptr, &text, &text_len);
        if (rc != RT_OK) { std::fprintf(stderr, "rt_format_ppm_device: %s\n", rt_last_error()); return 1; }
    }
    double t4 = now_s();

    char out_path[4096];
    if (rt_host_output_path(scene_path, out_path, sizeof out_path) != RT_OK) return 1;
    {
        FILE* f = std::fopen(out_path, "wb");
        if (!f) { std::printf("ERROR: failed to create ppm image\n"); return 0; }   // main.cpp:621
        const size_t put = std::fwrite(text, 1, text_len, f);
        std::fclose(f);
        if (put != text_len) { std::fprintf(stderr, "short write to %s\n", out_path); return 1; }
    }
    double t5 = now_s();

    if (verbose) {
        std::fprintf(stderr,
                     "gpus %d | parse %.3f ms | upload%s %.3f ms | render %.3f ms (device %.3f ms, %llu nodes, accel %s, %d graph replays) | "
                     "format+download %.3f ms | write %.3f ms\n",
                     gpus, (t1 - t0) * 1e3, frames > 1 ? "+earlier frames" : "", (t2 - t1) * 1e3, (t3 - t2) * 1e3, stats.render_ms,
                     (unsigned long long)stats.nodes_total, stats.accel_used == RT_ACCEL_LBVH ? "lbvh" : "brute", stats.graph_replays,
                     (t4 - t3) * 1e3, (t5 - t4) * 1e3);
    }
    if (multi) rt_free_scene_multi(multi);
    if (scene) rt_free_scene(scene);
    rt_host_free_scene_desc(desc);
    return 0;
}
