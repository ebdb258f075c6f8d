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

// devmem.cu: device blocks come from / go back to a process-wide cache
void* dev_alloc(size_t n, size_t* capacity);
void dev_free(void* p, size_t capacity);
size_t dev_release_cached();     // cudaFree of everything cached; returns the bytes released

// Device allocation that frees itself (into the cache: only release a buffer whose work is done).
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
        if (p) dev_free(p, bytes);
        p = nullptr; bytes = 0;
    }
    // grows only; contents are NOT preserved
    bool reserve(size_t n)
    {
        if (n <= bytes) return true;
        release();
        p = dev_alloc(n, &bytes);     // bytes = the block's capacity (>= n)
        if (!p) { bytes = 0; return false; }
        return true;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

constexpr int kMaxSubframes = 8;

// Pinned host allocation that frees itself.
struct PinnedBuf {
    void* p = nullptr;
    size_t bytes = 0;
    PinnedBuf() = default;
    PinnedBuf(const PinnedBuf&) = delete;
    PinnedBuf& operator=(const PinnedBuf&) = delete;
    PinnedBuf(PinnedBuf&& o) noexcept : p(o.p), bytes(o.bytes) { o.p = nullptr; o.bytes = 0; }
    PinnedBuf& operator=(PinnedBuf&& o) noexcept
    {
        if (this != &o) { release(); p = o.p; bytes = o.bytes; o.p = nullptr; o.bytes = 0; }
        return *this;
    }
    ~PinnedBuf() { release(); }
    void release()
    {
        if (p) cudaFreeHost(p);
        p = nullptr; bytes = 0;
    }
    bool reserve(size_t n)
    {
        if (n <= bytes) return true;
        release();
        if (!cuda_ok(cudaMallocHost(&p, n), "cudaMallocHost")) { p = nullptr; return false; }
        bytes = n;
        return true;
    }
};

// Streams and events of the LBVH pipeline's launch chains (render.cu): sub-frame j runs
// trace -> spawn on main[j] (highest stream priority: the critical path) and shadow -> colour
// on side[j] (lowest); both fork from and join the caller's stream.  Created on first use,
// travels with the Workspace.
struct ChainStreams {
    cudaStream_t main[kMaxSubframes] = { nullptr }, side[kMaxSubframes] = { nullptr };
    cudaEvent_t fork[kMaxSubframes] = { nullptr };
    cudaEvent_t side_done[kMaxSubframes] = { nullptr }, main_done[kMaxSubframes] = { nullptr }, begin = nullptr;
    int chains = 0;      // chains whose streams / events exist
    ChainStreams() = default;
    ChainStreams(const ChainStreams&) = delete;
    ChainStreams& operator=(const ChainStreams&) = delete;
    ChainStreams(ChainStreams&& o) noexcept { steal(o); }
    ChainStreams& operator=(ChainStreams&& o) noexcept
    {
        if (this != &o) { release(); steal(o); }
        return *this;
    }
    ~ChainStreams() { release(); }
    bool ensure(int n);  // render.cu
    void sync() const;
    void release();
private:
    void steal(ChainStreams& o)
    {
        for (int j = 0; j < kMaxSubframes; j++) {
            main[j] = o.main[j]; side[j] = o.side[j]; side_done[j] = o.side_done[j]; main_done[j] = o.main_done[j];
            o.main[j] = o.side[j] = nullptr; o.side_done[j] = o.main_done[j] = nullptr;
            fork[j] = o.fork[j]; o.fork[j] = nullptr;
        }
        begin = o.begin; o.begin = nullptr;
        chains = o.chains; o.chains = 0;
    }
};

// Wavefront work buffers: queues + node records per generation (scene_device.h).
struct Workspace {
    DevBuf level_mem[kMaxLevels];
    int capacity[kMaxLevels] = { 0 };
    DevBuf counters;
    DevBuf tiles;          // RGBA8 [local_tile][row][col]
    DevBuf primary_ids;
    DevBuf redo_list;      // shadow chains for the slow exact path: kRedoCapacity entries per sub-frame
    DevBuf rgb24, flags8;  // scanline-order planes for rt_download_framebuffer
    ChainStreams chains;
    PinnedBuf h_counters;  // one Counters per sub-frame
    // ppm_device.cu: per-block byte counts and offsets, the text on the device and in pinned host memory
    DevBuf ppm_blocks, ppm_offsets, ppm_text;
    PinnedBuf ppm_total_host, ppm_host;
    int device = -1;
};

// The captured launch sequence of a frame and the FrameParams it was captured for (render.cu: enqueue_frame).
struct FrameGraph {
    cudaGraphExec_t exec = nullptr;
    FrameParams ps[kMaxSubframes];
    int K = 0, accel = 0, launches = 0;
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
    DevBuf bvh8_nodes, bvh8_prims;      // the 8-wide tree of the any-hit walk (bvh.cu: collapse8_kernel)
    int bvh_n_nodes = 0, bvh_root = 0, bvh8_n_nodes = 0;
    int bvh_levels4 = 0, bvh_levels8 = 0;   // depth of the two trees (build_bvh checks them against the walk stacks)
    bool bvh_built = false;
    // occluder maps of the first kMaxShadowMaps lights (directional ones; bvh.cu: build_shadow_maps): all maps in one
    // buffer, smap[k].cells holds the map's OFFSET (in ints) + 1 as an integer (0 = no map) until scene_view() adds the base
    DevBuf smap_cells;
    struct SmapHost { size_t offset_plus_1 = 0; float ub[3] = { 0, 0, 0 }, vb[3] = { 0, 0, 0 }; float u0 = 0, v0 = 0, inv_cell = 0; float origin[3] = { 0, 0, 0 }; int R = 0, cube = 0; };
    SmapHost smap[kMaxShadowMaps];

    CameraView cam{};
    Workspace ws;

    cudaStream_t stream = nullptr;      // owned
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t prof_ev[kMaxLevels + 2] = { nullptr };   // params.profile: boundaries of the generations
    cudaEvent_t prof_mid[kMaxLevels][3] = { { nullptr } };   // LBVH pipeline: after trace + spawn, after shadow, after trace
    bool prof_split = false;
    int prof_levels = 0;                                 // events recorded by the last frame (0 = none)
    Counters* h_counters = nullptr;     // = ws.h_counters (pinned, one per sub-frame), set by ensure_workspace

    // last frame
    bool rendered = false;
    RtRenderParams last{};
    int last_local_rows = 0, last_local_pixels = 0, last_accel = 0, last_launches = 0;
    int last_subframes = 1;             // launch chains of the last frame (its queues were cut that many ways)
    bool have_primary_ids = false;
    FrameGraph graph;
    int graph_replays = 0;              // frames of this scene that were graph launches
    FrameParams last_ps[kMaxSubframes]; // the FrameParams of the last frame's launch chains (rt_download_hit_histograms)
};

// upload.cu
int build_scene(const RtSceneDesc* d, SceneHost* s);
void make_camera(const float eye[3], const float viewdir[3], const float updir[3], float hfov_deg, int width,
                 int height, CameraView* cam, bool normalised = false);
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
// Reads the counters of the last frame (stream must be idle): sums the sub-frames into `total`,
// returns the frame's kOverflow* bits (0 = complete); with kOverflowQueue `demand` says what to grow to.
unsigned frame_overflowed(const SceneHost* s, Counters* total, unsigned demand[kMaxLevels]);
int launch_hit_histograms(const SceneHost* s, unsigned long long* dev_hits, unsigned long long* dev_shadow, cudaStream_t st);
int overflow_error(unsigned bits);      // RT_OK, or RT_E_LIMIT + message for the bits no regrow cures
void stash_workspace(Workspace&& ws);

int launch_split_planes(SceneHost* s, cudaStream_t st, const uchar4* src = nullptr, int n_pixels = 0);   // src: null = the last frame's tiles
int launch_deinterleave(const SceneHost* s, const void* gathered, void* out, int world, int tile_rows, cudaStream_t st);

// api.cu
int checked_params(const RtRenderParams* in, RtRenderParams* out);     // defaults + validation of RtRenderParams

// ppm_device.cu
int format_ppm_device(SceneHost* s, const uchar4* px, const char** text, size_t* len, cudaStream_t st);

}  // namespace rt

struct RtScene {
    rt::SceneHost h;
};
