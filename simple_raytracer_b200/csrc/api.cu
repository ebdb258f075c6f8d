// api.cu — the C ABI of include/rt.h.
#include "bvh.cuh"
#include "host_scene.h"

#include <cstdio>
#include <cstring>
#include <new>

namespace rt {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

bool cuda_ok(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return true;
    g_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return false;
}


namespace {

RtRenderParams default_params()
{
    RtRenderParams p;
    std::memset(&p, 0, sizeof p);
    p.max_depth = 4; p.epsilon = 1.0e-3f; p.accel = RT_ACCEL_AUTO;
    p.rank = 0; p.world_size = 1; p.tile_rows = 8;
    return p;
}

int check_params(const RtRenderParams* in, RtRenderParams* out)
{
    *out = in ? *in : default_params();
    if (out->world_size == 0) out->world_size = 1;
    if (out->tile_rows == 0) out->tile_rows = 8;
    if (out->max_depth < 0 || out->max_depth > 5) { set_error("max_depth must be in 0..5"); return RT_E_INVALID; }
    if (out->world_size < 1 || out->rank < 0 || out->rank >= out->world_size) { set_error("bad rank / world_size"); return RT_E_INVALID; }
    if (out->tile_rows < 1) { set_error("tile_rows must be >= 1"); return RT_E_INVALID; }
    if (out->subframes < 0 || out->subframes > kMaxSubframes) { set_error("subframes must be in 0..8"); return RT_E_INVALID; }
    if (out->work_sharing < 0 || out->work_sharing > 2) { set_error("work_sharing must be 0, 1 or 2"); return RT_E_INVALID; }
    if (out->accel < RT_ACCEL_AUTO || out->accel > RT_ACCEL_LBVH) { set_error("bad accel"); return RT_E_INVALID; }   // (RT_ACCEL_LBVH8: probes only)
    return RT_OK;
}

// ---- parity probes ---------------------------------------------------------------
template <class Accel>
__global__ void probe_closest_kernel(SceneView S, Accel acc, int n, const float* __restrict__ o, const float* __restrict__ d,
                                     float tmin, int* __restrict__ out_id, float* __restrict__ out_t,
                                     float* __restrict__ out_n, float* __restrict__ out_b)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 ro = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    const Hit h = acc.closest(ro, rd, tmin);
    if (h.obj < 0) {
        out_id[i] = 0; out_t[i] = 0.0f;
        for (int k = 0; k < 3; k++) { out_n[3 * i + k] = 0.0f; out_b[3 * i + k] = 0.0f; }
        return;
    }
    const Surface sf = finish_hit(S, ro, rd, h);
    out_id[i] = S.obj[h.obj].z; out_t[i] = h.t;
    out_n[3 * i] = sf.n.x; out_n[3 * i + 1] = sf.n.y; out_n[3 * i + 2] = sf.n.z;
    const bool face = h.obj < S.nf;
    out_b[3 * i] = face ? sf.a : 0.0f; out_b[3 * i + 1] = face ? sf.b : 0.0f; out_b[3 * i + 2] = face ? sf.g : 0.0f;
}

template <class Accel>
__global__ void probe_shadow_kernel(SceneView S, Accel acc, int n, const float* __restrict__ o, const float* __restrict__ d,
                                    const int* __restrict__ self_order, float tmin, const float* __restrict__ tmax,
                                    float* __restrict__ out_mask, int* __restrict__ out_hits)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 ro = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    float mask = 1.0f;
    const int hits = acc.shadow(ro, rd, self_order[i], tmin, tmax ? tmax[i] : 0.0f, tmax != nullptr, mask);
    out_mask[i] = mask; out_hits[i] = hits;
}

// ... through the 8-wide tree of the any-hit kernel
__global__ void probe_shadow8_kernel(SceneView S, BvhAccel acc, int n, const float* __restrict__ o, const float* __restrict__ d,
                                     const int* __restrict__ self_order, float tmin, const float* __restrict__ tmax,
                                     float* __restrict__ out_mask, int* __restrict__ out_hits)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 ro = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    float mask = 1.0f;
    const int hits = acc.shadow(ro, rd, self_order[i], tmin, tmax ? tmax[i] : 0.0f, tmax != nullptr, mask, true);
    out_mask[i] = mask; out_hits[i] = hits;
}

__global__ void ids_to_order_kernel(SceneView S, int n, const int* __restrict__ ids, int* __restrict__ order)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // reference ids are unique per object; a linear search keeps the probe independent of any id layout
    const int want = ids[i];
    int found = -1;
    for (int k = 0; k < S.nf + S.ns; k++)
        if (S.obj[k].z == want) { found = k; break; }
    order[i] = found;
}

int probe_accel(SceneHost* s, int accel)
{
    if (accel == RT_ACCEL_AUTO) accel = (s->nf + s->ns) > 48 ? RT_ACCEL_LBVH : RT_ACCEL_BRUTE;
    if ((accel == RT_ACCEL_LBVH || accel == RT_ACCEL_LBVH8) && !s->bvh_built) {
        const int rc = build_bvh(s);
        if (rc != RT_OK) return rc;
    }
    return accel;
}

}  // namespace

int checked_params(const RtRenderParams* in, RtRenderParams* out) { return check_params(in, out); }

}  // namespace rt

using namespace rt;

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }

const char* rt_last_error(void) { return g_error.c_str(); }

int rt_upload_scene(const RtSceneDesc* desc, RtScene** out)
{
    if (!desc || !out) { set_error("rt_upload_scene: null argument"); return RT_E_INVALID; }
    *out = nullptr;
    int ndev = 0;
    if (!cuda_ok(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount") || ndev < 1) {
        if (ndev < 1 && g_error.empty()) set_error("no CUDA device");
        set_error("rt_upload_scene needs a CUDA device (there is no CPU fallback): " + g_error);
        return RT_E_CUDA;
    }
    RtScene* sc = new (std::nothrow) RtScene();
    if (!sc) { set_error("out of memory"); return RT_E_NOMEM; }
    SceneHost* s = &sc->h;
    int rc = RT_OK;
    if (!cuda_ok(cudaGetDevice(&s->device), "cudaGetDevice") ||
        !cuda_ok(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !cuda_ok(cudaEventCreate(&s->ev0), "cudaEventCreate") || !cuda_ok(cudaEventCreate(&s->ev1), "cudaEventCreate"))
        rc = RT_E_CUDA;
    if (rc == RT_OK) rc = build_scene(desc, s);
    if (rc != RT_OK) { rt_free_scene(sc); return rc; }
    *out = sc;
    return RT_OK;
}

void rt_free_scene(RtScene* scene)
{
    if (!scene) return;
    SceneHost* s = &scene->h;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    s->ws.chains.sync();
    if (s->graph.exec) { cudaGraphExecDestroy(s->graph.exec); s->graph.exec = nullptr; }
    if (s->ws.device >= 0) stash_workspace(std::move(s->ws));
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    for (int k = 0; k < kMaxLevels + 2; k++)
        if (s->prof_ev[k]) cudaEventDestroy(s->prof_ev[k]);
    for (int k = 0; k < kMaxLevels; k++)
        for (int j = 0; j < 3; j++)
            if (s->prof_mid[k][j]) cudaEventDestroy(s->prof_mid[k][j]);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete scene;
}

int rt_set_camera(RtScene* scene, const float eye[3], const float viewdir[3], const float updir[3], float hfov_deg)
{
    if (!scene || !eye || !viewdir || !updir) { set_error("rt_set_camera: null argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    make_camera(eye, viewdir, updir, hfov_deg, s->width, s->height, &s->cam);
    // BVH padding depends on how far ray origins can be from the scene: rebuild lazily
    s->bvh_built = false;
    return RT_OK;
}

size_t rt_tile_buffer_bytes(const RtScene* scene, int32_t world_size, int32_t tile_rows)
{
    if (!scene || world_size < 1 || tile_rows < 1) return 0;
    const TileLayout tl = tile_layout(scene->h.height, scene->h.width, world_size, tile_rows);
    return (size_t)tl.n_local_pixels * sizeof(uchar4);
}

int rt_tile_layout(int32_t height, int32_t width, int32_t world_size, int32_t tile_rows, int32_t out[4])
{
    if (height < 1 || width < 1 || world_size < 1 || tile_rows < 1 || !out) { set_error("rt_tile_layout: bad argument"); return RT_E_INVALID; }
    const TileLayout tl = tile_layout(height, width, world_size, tile_rows);
    out[0] = tl.n_tiles_total; out[1] = tl.n_local_tiles; out[2] = tl.n_local_rows; out[3] = tl.n_local_pixels;
    return RT_OK;
}

int rt_render_async(RtScene* scene, const RtRenderParams* params, void* dev_tiles, void* cuda_stream)
{
    if (!scene) { set_error("rt_render_async: null scene"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    RtRenderParams p;
    int rc = check_params(params, &p);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaSetDevice(s->device));
    if ((rc = ensure_workspace(s, p, nullptr)) != RT_OK) return rc;
    return enqueue_frame(s, p, (uchar4*)dev_tiles, cuda_stream ? (cudaStream_t)cuda_stream : s->stream, nullptr);
}

size_t rt_release_cached_memory(void)
{
    {
        Workspace drop;
        stash_workspace(std::move(drop));      // the retired workspace goes first
    }
    return dev_release_cached();
}

int rt_async_frame_status(RtScene* scene)
{
    if (!scene) { set_error("rt_async_frame_status: null scene"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    if (!s->rendered) { set_error("rt_async_frame_status before a frame"); return RT_E_STATE; }
    RT_CUDA(cudaSetDevice(s->device));
    RT_CUDA(cudaMemcpy(s->h_counters, s->ws.counters.p, sizeof(Counters) * s->last_subframes, cudaMemcpyDeviceToHost));
    unsigned demand[kMaxLevels];
    const unsigned bits = frame_overflowed(s, nullptr, demand);
    if (!bits) return RT_OK;
    int rc = overflow_error(bits);
    if (rc != RT_OK) return rc;
    rc = ensure_workspace(s, s->last, demand);
    if (rc != RT_OK) return rc;
    set_error("the last asynchronous frame overflowed a wavefront queue (queues grown: render it again)");
    return RT_E_NOMEM;
}

int rt_render(RtScene* scene, const RtRenderParams* params, RtStats* stats)
{
    if (!scene) { set_error("rt_render: null scene"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    RtRenderParams p;
    int rc = check_params(params, &p);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaSetDevice(s->device));
    if ((rc = ensure_workspace(s, p, nullptr)) != RT_OK) return rc;
    int regrows = 0, launches = 0;
    float ms = 0.0f;
    Counters total;
    for (;;) {
        RT_CUDA(cudaEventRecord(s->ev0, s->stream));
        if ((rc = enqueue_frame(s, p, nullptr, s->stream, &launches)) != RT_OK) return rc;
        RT_CUDA(cudaEventRecord(s->ev1, s->stream));
        RT_CUDA(cudaMemcpyAsync(s->h_counters, s->ws.counters.p, sizeof(Counters) * s->last_subframes,
                                cudaMemcpyDeviceToHost, s->stream));
        RT_CUDA(cudaStreamSynchronize(s->stream));
        RT_CUDA(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        unsigned demand[kMaxLevels];
        const unsigned bits = frame_overflowed(s, &total, demand);
        if ((rc = overflow_error(bits)) != RT_OK) return rc;
        if (!bits) break;
        if (++regrows > 16) { set_error("wavefront queues still overflow after 16 regrows"); return RT_E_NOMEM; }
        if ((rc = ensure_workspace(s, p, demand)) != RT_OK) return rc;
    }
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        const Counters& c = total;
        stats->render_ms = ms;
        for (int l = 0; l <= p.max_depth; l++) {
            stats->nodes_per_level[l] = c.nodes[l];
            stats->nodes_total += c.nodes[l];
            if (l >= 1) { stats->rays_reflect += c.rays_reflect[l]; stats->rays_transmit += c.rays[l] - c.rays_reflect[l]; }
        }
        // one primary ray per pixel of this rank's rows that exist in the image
        uint64_t rows = 0;
        const TileLayout tl = tile_layout(s->height, s->width, p.world_size, p.tile_rows);
        for (int r = 0; r < tl.n_local_rows; r++) {
            const int tile = r / p.tile_rows, in_tile = r % p.tile_rows;
            if ((tile * p.world_size + p.rank) * p.tile_rows + in_tile < s->height) rows++;
        }
        stats->rays_primary = rows * (uint64_t)s->width;
        stats->rays_shadow = stats->nodes_total * (uint64_t)s->nl;   // one shadow TraceRay per light per node (896 / 928)
        stats->bvh_nodes_visited = c.bvh_nodes; stats->prim_tests = c.prim_tests; stats->texel_fetches = c.texels;
        stats->shadow_boxes_tested = c.shadow_boxes; stats->shadow_prim_tests = c.shadow_prims;
        stats->shadow_rays_walked = (s->last_accel == RT_ACCEL_LBVH && p.instrument) ? c.shadow_walks : stats->rays_shadow;
        stats->shadow_chains_redone = c.redo_count;
        stats->shadow_map_hits = (int32_t)std::min<unsigned long long>(c.shadow_map_hits, 0x7fffffffull);
        stats->bvh_levels4 = s->last_accel == RT_ACCEL_LBVH ? s->bvh_levels4 : 0;
        stats->bvh_levels8 = s->last_accel == RT_ACCEL_LBVH ? s->bvh_levels8 : 0;
        stats->graph_replays = s->graph_replays;
        stats->accel_used = s->last_accel; stats->kernel_launches = launches; stats->queue_regrows = regrows;
        if (p.instrument && s->last_accel == RT_ACCEL_LBVH) {
            for (int l = 0; l <= p.max_depth; l++)
                for (int k = 0; k < 2; k++) {
                    const WalkClock& w = c.clock[k][l];
                    if (!w.warps) continue;
                    const double first = (double)~w.inv_first, empty = (double)~w.inv_empty, last = (double)w.last;
                    (k ? stats->shadow_span_ms : stats->trace_span_ms)[l] = (last - first) * 1e-6;
                    (k ? stats->shadow_drain_ms : stats->trace_drain_ms)[l] = (last - empty) * 1e-6;
                    (k ? stats->shadow_warp_drain_ms : stats->trace_warp_drain_ms)[l] = (double)w.drain_sum / w.warps * 1e-6;
                    (k ? stats->shadow_drain_steps : stats->trace_drain_steps)[l] = (double)w.drain_steps / w.warps;
                    (k ? stats->shadow_drain_lanes : stats->trace_drain_lanes)[l] =
                        w.drain_steps ? (double)w.drain_lane_steps / (double)w.drain_steps : 0.0;
                    if (!k) stats->trace_drain_steals[l] = (double)w.drain_steals / w.warps;
                }
        }
        if (s->prof_levels > 0) {
            float t = 0.0f;
            for (int l = 0; l < s->prof_levels; l++) {
                RT_CUDA(cudaEventElapsedTime(&t, s->prof_ev[l], s->prof_ev[l + 1]));
                stats->level_ms[l] = t;
                if (s->prof_split) {
                    RT_CUDA(cudaEventElapsedTime(&t, s->prof_ev[l], s->prof_mid[l][2]));      stats->trace_ms[l] = t;
                    RT_CUDA(cudaEventElapsedTime(&t, s->prof_mid[l][2], s->prof_mid[l][0])); stats->spawn_ms[l] = t;
                    RT_CUDA(cudaEventElapsedTime(&t, s->prof_mid[l][0], s->prof_mid[l][1])); stats->shadow_ms[l] = t;
                    RT_CUDA(cudaEventElapsedTime(&t, s->prof_mid[l][1], s->prof_ev[l + 1])); stats->shade_ms[l] = t;
                }
            }
            RT_CUDA(cudaEventElapsedTime(&t, s->prof_ev[s->prof_levels], s->prof_ev[s->prof_levels + 1]));
            stats->combine_ms = t;
        }
    }
    return RT_OK;
}

int rt_download_framebuffer(RtScene* scene, uint8_t* rgb, uint8_t* flags)
{
    if (!scene || !rgb) { set_error("rt_download_framebuffer: null argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    if (!s->rendered) { set_error("rt_download_framebuffer before rt_render"); return RT_E_STATE; }
    RT_CUDA(cudaSetDevice(s->device));
    int rc = launch_split_planes(s, s->stream);
    if (rc != RT_OK) return rc;
    const RtRenderParams& p = s->last;
    const size_t W = (size_t)s->width;
    if (p.world_size == 1) {
        RT_CUDA(cudaMemcpyAsync(rgb, s->ws.rgb24.p, W * s->height * 3, cudaMemcpyDeviceToHost, s->stream));
        if (flags) RT_CUDA(cudaMemcpyAsync(flags, s->ws.flags8.p, W * s->height, cudaMemcpyDeviceToHost, s->stream));
    } else {
        // local tile t holds global rows [(t*world + rank)*tile_rows, +tile_rows)
        const int n_tiles = s->last_local_rows / p.tile_rows;
        for (int t = 0; t < n_tiles; t++) {
            const int row0 = (t * p.world_size + p.rank) * p.tile_rows;
            if (row0 >= s->height) break;
            const int rows = std::min(p.tile_rows, s->height - row0);
            const size_t src = (size_t)t * p.tile_rows * W;
            RT_CUDA(cudaMemcpyAsync(rgb + (size_t)row0 * W * 3, s->ws.rgb24.as<unsigned char>() + src * 3, (size_t)rows * W * 3,
                                    cudaMemcpyDeviceToHost, s->stream));
            if (flags)
                RT_CUDA(cudaMemcpyAsync(flags + (size_t)row0 * W, s->ws.flags8.as<unsigned char>() + src, (size_t)rows * W,
                                        cudaMemcpyDeviceToHost, s->stream));
        }
    }
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

int rt_format_ppm_device(RtScene* scene, const void* dev_rgba, const char** text, size_t* len)
{
    if (!scene || !text || !len) { set_error("rt_format_ppm_device: null argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    RT_CUDA(cudaSetDevice(s->device));
    const uchar4* px = static_cast<const uchar4*>(dev_rgba);
    if (!px) {
        // the scene's own frame: tiles of a single-rank render ARE scanline order
        if (!s->rendered) { set_error("rt_format_ppm_device before rt_render"); return RT_E_STATE; }
        if (s->last.world_size != 1) { set_error("rt_format_ppm_device: the last frame is one rank's shard; pass the assembled frame"); return RT_E_STATE; }
        px = s->ws.tiles.as<uchar4>();
    }
    return format_ppm_device(s, px, text, len, s->stream);
}

int rt_deinterleave_tiles(const RtScene* scene, const void* dev_gathered, void* dev_rgba_out, int32_t world_size,
                          int32_t tile_rows, void* cuda_stream)
{
    if (!scene || !dev_gathered || !dev_rgba_out || world_size < 1 || tile_rows < 1) { set_error("rt_deinterleave_tiles: bad argument"); return RT_E_INVALID; }
    return launch_deinterleave(&scene->h, dev_gathered, dev_rgba_out, world_size, tile_rows,
                               cuda_stream ? (cudaStream_t)cuda_stream : scene->h.stream);
}

int rt_download_primary_ids(RtScene* scene, int32_t* ids)
{
    if (!scene || !ids) { set_error("rt_download_primary_ids: null argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    if (!s->rendered || !s->have_primary_ids || s->last.world_size != 1) {
        set_error("primary ids need a previous rt_render with instrument != 0 and world_size == 1");
        return RT_E_STATE;
    }
    RT_CUDA(cudaSetDevice(s->device));
    RT_CUDA(cudaMemcpy(ids, s->ws.primary_ids.p, (size_t)s->width * s->height * sizeof(int), cudaMemcpyDeviceToHost));
    return RT_OK;
}

int rt_download_hit_histograms(RtScene* scene, uint64_t* hits, uint64_t* shadow)
{
    if (!scene || !hits || !shadow) { set_error("rt_download_hit_histograms: null argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    if (!s->rendered || s->last_accel != RT_ACCEL_LBVH) {
        set_error("hit histograms need a previous frame of the LBVH pipeline (params.accel = RT_ACCEL_LBVH)");
        return RT_E_STATE;
    }
    RT_CUDA(cudaSetDevice(s->device));
    const size_t nh = (size_t)kMaxLevels * (size_t)(s->nf + s->ns), nsw = (size_t)kMaxLevels * (size_t)std::max(s->nl, 1) * 3;
    DevBuf dh, dsw;
    if (!dh.reserve(nh * 8) || !dsw.reserve(nsw * 8)) return RT_E_NOMEM;
    RT_CUDA(cudaMemsetAsync(dh.p, 0, nh * 8, s->stream));
    RT_CUDA(cudaMemsetAsync(dsw.p, 0, nsw * 8, s->stream));
    const int rc = launch_hit_histograms(s, dh.as<unsigned long long>(), dsw.as<unsigned long long>(), s->stream);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaMemcpyAsync(hits, dh.p, nh * 8, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaMemcpyAsync(shadow, dsw.p, (size_t)kMaxLevels * (size_t)s->nl * 3 * 8, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

int rt_trace_closest(RtScene* scene, int32_t accel, int32_t n, const float* origins, const float* dirs, float tmin,
                     int32_t* out_id, float* out_t, float* out_normal, float* out_bary)
{
    if (!scene || n < 0 || !origins || !dirs || !out_id || !out_t) { set_error("rt_trace_closest: bad argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    RT_CUDA(cudaSetDevice(s->device));
    accel = probe_accel(s, accel);
    if (accel < 0) return accel;
    if (n == 0) return RT_OK;
    const size_t N = (size_t)n;
    DevBuf o, d, id, t, nn, bb;
    if (!o.reserve(N * 12) || !d.reserve(N * 12) || !id.reserve(N * 4) || !t.reserve(N * 4) || !nn.reserve(N * 12) || !bb.reserve(N * 12)) return RT_E_NOMEM;
    RT_CUDA(cudaMemcpyAsync(o.p, origins, N * 12, cudaMemcpyHostToDevice, s->stream));
    RT_CUDA(cudaMemcpyAsync(d.p, dirs, N * 12, cudaMemcpyHostToDevice, s->stream));
    const SceneView S = scene_view(s);
    const int B = 128, G = (n + B - 1) / B;
    if (accel == RT_ACCEL_LBVH8) { set_error("rt_trace_closest: closest hits walk the 4-wide tree (RT_ACCEL_LBVH)"); return RT_E_INVALID; }
    if (accel == RT_ACCEL_LBVH) {
        BvhAccel a = make_bvh_accel<false>(S);
        probe_closest_kernel<<<G, B, 0, s->stream>>>(S, a, n, o.as<float>(), d.as<float>(), tmin, id.as<int>(), t.as<float>(), nn.as<float>(), bb.as<float>());
    } else {
        BruteAccel a; a.tri = S.tri; a.sph = S.sph; a.nf = S.nf; a.ns = S.ns;
        probe_closest_kernel<<<G, B, 0, s->stream>>>(S, a, n, o.as<float>(), d.as<float>(), tmin, id.as<int>(), t.as<float>(), nn.as<float>(), bb.as<float>());
    }
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(out_id, id.p, N * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaMemcpyAsync(out_t, t.p, N * 4, cudaMemcpyDeviceToHost, s->stream));
    if (out_normal) RT_CUDA(cudaMemcpyAsync(out_normal, nn.p, N * 12, cudaMemcpyDeviceToHost, s->stream));
    if (out_bary) RT_CUDA(cudaMemcpyAsync(out_bary, bb.p, N * 12, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

int rt_trace_shadow(RtScene* scene, int32_t accel, int32_t n, const float* origins, const float* dirs,
                    const int32_t* self_id, float tmin, const float* tmax, float* out_mask, int32_t* out_hits)
{
    if (!scene || n < 0 || !origins || !dirs || !self_id || !out_mask) { set_error("rt_trace_shadow: bad argument"); return RT_E_INVALID; }
    SceneHost* s = &scene->h;
    RT_CUDA(cudaSetDevice(s->device));
    accel = probe_accel(s, accel);
    if (accel < 0) return accel;
    if (n == 0) return RT_OK;
    const size_t N = (size_t)n;
    DevBuf o, d, sid, sord, tm, mask, hits;
    if (!o.reserve(N * 12) || !d.reserve(N * 12) || !sid.reserve(N * 4) || !sord.reserve(N * 4) || !tm.reserve(N * 4) || !mask.reserve(N * 4) || !hits.reserve(N * 4)) return RT_E_NOMEM;
    RT_CUDA(cudaMemcpyAsync(o.p, origins, N * 12, cudaMemcpyHostToDevice, s->stream));
    RT_CUDA(cudaMemcpyAsync(d.p, dirs, N * 12, cudaMemcpyHostToDevice, s->stream));
    RT_CUDA(cudaMemcpyAsync(sid.p, self_id, N * 4, cudaMemcpyHostToDevice, s->stream));
    if (tmax) RT_CUDA(cudaMemcpyAsync(tm.p, tmax, N * 4, cudaMemcpyHostToDevice, s->stream));
    const SceneView S = scene_view(s);
    const int B = 128, G = (n + B - 1) / B;
    ids_to_order_kernel<<<G, B, 0, s->stream>>>(S, n, sid.as<int>(), sord.as<int>());
    const float* tmp = tmax ? tm.as<float>() : nullptr;
    if (accel == RT_ACCEL_LBVH8 && S.bvh.n_nodes8 == 0) { set_error("rt_trace_shadow: the scene has no 8-wide tree (8 primitives or fewer)"); return RT_E_STATE; }
    if (accel == RT_ACCEL_LBVH || accel == RT_ACCEL_LBVH8) {
        BvhAccel a = make_bvh_accel<false>(S);
        a.count_all = true;
        if (accel == RT_ACCEL_LBVH8) probe_shadow8_kernel<<<G, B, 0, s->stream>>>(S, a, n, o.as<float>(), d.as<float>(), sord.as<int>(), tmin, tmp, mask.as<float>(), hits.as<int>());
        else probe_shadow_kernel<<<G, B, 0, s->stream>>>(S, a, n, o.as<float>(), d.as<float>(), sord.as<int>(), tmin, tmp, mask.as<float>(), hits.as<int>());
    } else {
        BruteAccel a; a.tri = S.tri; a.sph = S.sph; a.nf = S.nf; a.ns = S.ns;
        probe_shadow_kernel<<<G, B, 0, s->stream>>>(S, a, n, o.as<float>(), d.as<float>(), sord.as<int>(), tmin, tmp, mask.as<float>(), hits.as<int>());
    }
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(out_mask, mask.p, N * 4, cudaMemcpyDeviceToHost, s->stream));
    if (out_hits) RT_CUDA(cudaMemcpyAsync(out_hits, hits.p, N * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

}  /* extern "C" */
