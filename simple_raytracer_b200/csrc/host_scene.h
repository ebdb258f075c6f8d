// host_scene.h — the state behind an RtScene handle (host side of the CUDA core).
#pragma once

#include "rt.h"
#include "scene_device.h"

#include <cuda_runtime.h>
#include <string>
#include <vector>

namespace rt {

void set_error(const std::string& msg);
bool cuda_ok(cudaError_t e, const char* what);   // false + rt_last_error on failure

#define RT_CUDA(call)                                     \
    do {                                                  \
        if (!::rt::cuda_ok((call), #call)) return RT_E_CUDA; \
    } while (0)

// Device allocation that frees itself.
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes) { o.p = nullptr; o.bytes = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept
    {
        if (this != &o) { release(); p = o.p; bytes = o.bytes; o.p = nullptr; o.bytes = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
    }
    // grows only; contents are NOT preserved
    bool reserve(size_t n)
    {
        if (n <= bytes) return true;
        release();
        if (!cuda_ok(cudaMalloc(&p, n ? n : 16), "cudaMalloc")) { p = nullptr; return false; }
        bytes = n;
        return true;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Wavefront work buffers: queues + node records per generation (scene_device.h).
struct Workspace {
    DevBuf level_mem[kMaxLevels];
    int capacity[kMaxLevels] = { 0 };
    DevBuf counters;
    DevBuf tiles;          // RGBA8 [local_tile][row][col]
    DevBuf primary_ids;
    DevBuf rgb24, flags8;  // scanline-order planes for rt_download_framebuffer
    int device = -1;
};

struct SceneHost {
    int device = 0;
    int width = 0, height = 0;
    int nf = 0, ns = 0, nl = 0, ntex = 0, nmat = 0;
    float bkg[3] = { 0, 0, 0 };
    float bkg_eta = 0;
    float scene_lo[3] = { 0, 0, 0 }, scene_hi[3] = { 0, 0, 0 };   // bounds of all primitives
    bool zero_mask_exit = false;        // all material / light values finite, exponents tame (shade.cuh)
    bool nonunit_directional = false;   // some directional light has |dir| != 1

    DevBuf tri, sph, face_attr, obj, mat, light, tex_desc, texels;
    DevBuf bvh_nodes, bvh_tri, bvh_sph;
    int bvh_n_nodes = 0, bvh_root = 0;
    bool bvh_built = false;

    CameraView cam{};
    Workspace ws;

    cudaStream_t stream = nullptr;      // owned
    cudaStream_t stream2 = nullptr;     // owned: the shadow -> colour chain of the LBVH pipeline (render.cu)
    cudaEvent_t ev_fork[kMaxLevels] = { nullptr }, ev_join = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t prof_ev[kMaxLevels + 2] = { nullptr };   // params.profile: boundaries of the generations
    cudaEvent_t prof_mid[kMaxLevels][2] = { { nullptr } };   // LBVH pipeline: after trace, after shadow
    bool prof_split = false;
    int prof_levels = 0;                                 // events recorded by the last frame (0 = none)
    Counters* h_counters = nullptr;     // pinned

    // last frame
    bool rendered = false;
    RtRenderParams last{};
    int last_local_rows = 0, last_local_pixels = 0, last_accel = 0, last_launches = 0;
    bool have_primary_ids = false;
};

// upload.cu
int build_scene(const RtSceneDesc* d, SceneHost* s);
void make_camera(const float eye[3], const float viewdir[3], const float updir[3], float hfov_deg, int width,
                 int height, CameraView* cam);
SceneView scene_view(const SceneHost* s);

// bvh.cu
int build_bvh(SceneHost* s);
int tuning_int(const char* name, int dflt, int lo, int hi);   // internal tuning knobs (environment)

// render.cu
struct TileLayout {
    int n_tiles_total, n_local_tiles, n_local_rows, n_local_pixels;
};
TileLayout tile_layout(int height, int width, int world, int tile_rows);
int enqueue_frame(SceneHost* s, const RtRenderParams& p, uchar4* tiles_out, cudaStream_t stream, int* launches);
int ensure_workspace(SceneHost* s, const RtRenderParams& p, const unsigned* demand /* per level or null */);
void stash_workspace(Workspace&& ws);

}  // namespace rt

struct RtScene {
    rt::SceneHost h;
};
