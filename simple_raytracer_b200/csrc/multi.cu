// multi.cu — N GPUs behind the C ABI (SURVEY 8b "RtScene owns all device memory for all GPUs, NCCL
// communicator, streams", 8e): ONE process, one host thread, ncclCommInitAll.
//
//   rt_upload_scene_multi   the scene is uploaded, packed and its LBVH built ONCE, on the first device; the
//                           packed tables and both trees (position independent: indices and offsets, no
//                           pointers) then go to the other devices with one grouped ncclBroadcast per
//                           table over NVLink — instead of N copies of the description over PCIe and N
//                           redundant builds.
//   rt_render_multi         device r renders the interleaved row tiles t with t % N == r into its send
//                           buffer (the same enqueue_frame as rt_render_async, CUDA-graph replay included),
//                           one grouped ncclAllGather of the RGBA8 tiles, one de-interleave kernel on the
//                           first device.  Queues that overflow are grown and the frame redone, as in rt_render.
//   rt_download_framebuffer_multi / rt_format_ppm_multi / rt_free_scene_multi
//
// NCCL is resolved with dlopen at the first multi-GPU upload (libnccl.so.2; RT_NCCL_LIB overrides), so
// librt_b200.so itself has no NCCL dependency: a process that already carries an NCCL (torch's) keeps
// using its own, and single-GPU users need none.
#include "bvh.cuh"
#include "host_scene.h"

#include <nccl.h>
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

namespace rt {
namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

bool load_nccl()
{
    if (g_nccl.lib) return true;
    const char* names[] = { std::getenv("RT_NCCL_LIB"), "libnccl.so.2", "libnccl.so" };
    void* h = nullptr;
    for (const char* nm : names) {
        if (!nm || !*nm) continue;
        h = dlopen(nm, RTLD_NOW | RTLD_NOLOAD);            // one the process already carries, first
        if (!h) h = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
        if (h) break;
    }
    if (!h) { set_error(std::string("multi-GPU rendering needs NCCL: ") + (dlerror() ? dlerror() : "libnccl.so.2 not found")); return false; }
    NcclApi a;
    a.lib = h;
    a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(dlsym(h, "ncclCommInitAll"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(h, "ncclAllGather"));
    a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(dlsym(h, "ncclBroadcast"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(dlsym(h, "ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(dlsym(h, "ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    if (!a.CommInitAll || !a.CommDestroy || !a.AllGather || !a.Broadcast || !a.GroupStart || !a.GroupEnd || !a.GetErrorString) {
        set_error("the NCCL library lacks a required entry point");
        return false;
    }
    g_nccl = a;
    return true;
}

bool nccl_ok(ncclResult_t r, const char* what)
{
    if (r == ncclSuccess) return true;
    set_error(std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "NCCL error"));
    return false;
}
#define RT_NCCL(call) do { if (!nccl_ok((call), #call)) return RT_E_CUDA; } while (0)

// Communicators are kept for the life of the process, one set per device list: ncclCommInitAll costs
// about a second on an 8-GPU node, a scene upload a few milliseconds.  (One host thread drives a multi
// scene at a time — the handle is not thread-safe — so sharing them between scenes is safe.)
struct CommSet {
    std::vector<int> dev;
    std::vector<ncclComm_t> comm;
};
std::vector<CommSet>& comm_cache()
{
    static std::vector<CommSet>* c = new std::vector<CommSet>();     // leaked on purpose, like the device-memory cache
    return *c;
}

int get_comms(const std::vector<int>& dev, std::vector<ncclComm_t>* out)
{
    for (const CommSet& c : comm_cache())
        if (c.dev == dev) { *out = c.comm; return RT_OK; }
    CommSet c;
    c.dev = dev;
    c.comm.assign(dev.size(), nullptr);
    RT_NCCL(g_nccl.CommInitAll(c.comm.data(), (int)dev.size(), dev.data()));
    comm_cache().push_back(c);
    *out = c.comm;
    return RT_OK;
}

}  // namespace
}  // namespace rt

// All device state of one scene on N GPUs.
struct RtMultiScene {
    int n = 0;
    std::vector<int> dev;
    std::vector<RtScene*> sc;               // sc[0]: the uploaded scene; sc[1..]: replicas filled by broadcast
    std::vector<ncclComm_t> comm;           // empty when n == 1; owned by the process-wide cache (get_comms)
    std::vector<rt::DevBuf> send, gathered; // per device: its tiles / everybody's tiles
    rt::DevBuf frame;                       // first device: the assembled frame, RGBA8 in scanline order
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int tile_rows = 0;
    bool rendered = false;
};

namespace rt {
namespace {

// The device buffers that make up a resident scene (packed tables, textures, both trees).
DevBuf SceneHost::* const kSceneBuffers[] = {
    &SceneHost::tri, &SceneHost::sph, &SceneHost::face_attr, &SceneHost::obj, &SceneHost::mat, &SceneHost::light,
    &SceneHost::tex_desc, &SceneHost::texels, &SceneHost::bvh_nodes, &SceneHost::bvh_tri, &SceneHost::bvh_sph,
    &SceneHost::bvh8_nodes, &SceneHost::bvh8_prims, &SceneHost::smap_cells,
};

// An RtScene without contents on the current device (what rt_upload_scene creates before build_scene).
int new_replica(RtScene** out)
{
    RtScene* sc = new (std::nothrow) RtScene();
    if (!sc) { set_error("out of memory"); return RT_E_NOMEM; }
    SceneHost* s = &sc->h;
    if (!cuda_ok(cudaGetDevice(&s->device), "cudaGetDevice") ||
        !cuda_ok(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !cuda_ok(cudaEventCreate(&s->ev0), "cudaEventCreate") || !cuda_ok(cudaEventCreate(&s->ev1), "cudaEventCreate")) {
        rt_free_scene(sc);
        return RT_E_CUDA;
    }
    *out = sc;
    return RT_OK;
}

void copy_scalars(const SceneHost* a, SceneHost* b)
{
    b->width = a->width; b->height = a->height;
    b->nf = a->nf; b->ns = a->ns; b->nl = a->nl; b->ntex = a->ntex; b->nmat = a->nmat;
    std::memcpy(b->bkg, a->bkg, sizeof b->bkg); b->bkg_eta = a->bkg_eta;
    std::memcpy(b->scene_lo, a->scene_lo, sizeof b->scene_lo); std::memcpy(b->scene_hi, a->scene_hi, sizeof b->scene_hi);
    b->zero_mask_exit = a->zero_mask_exit; b->nonunit_directional = a->nonunit_directional;
    b->bvh_n_nodes = a->bvh_n_nodes; b->bvh_root = a->bvh_root; b->bvh8_n_nodes = a->bvh8_n_nodes;
    b->bvh_levels4 = a->bvh_levels4; b->bvh_levels8 = a->bvh_levels8; b->bvh_built = a->bvh_built;
    for (int k = 0; k < kMaxShadowMaps; k++) b->smap[k] = a->smap[k];
    b->cam = a->cam;
}

// Device 0's tables and trees to every replica: one grouped ncclBroadcast per buffer (in place on the root).
int replicate(RtMultiScene* m)
{
    if (m->n == 1) return RT_OK;
    SceneHost* root = &m->sc[0]->h;
    for (int i = 1; i < m->n; i++) {
        RT_CUDA(cudaSetDevice(m->dev[i]));
        SceneHost* r = &m->sc[i]->h;
        copy_scalars(root, r);
        for (auto mb : kSceneBuffers)
            if ((root->*mb).bytes && !(r->*mb).reserve((root->*mb).bytes)) return RT_E_NOMEM;
    }
    RT_CUDA(cudaSetDevice(m->dev[0]));
    RT_CUDA(cudaStreamSynchronize(root->stream));
    for (auto mb : kSceneBuffers) {
        const size_t bytes = (root->*mb).bytes;
        if (!bytes) continue;
        RT_NCCL(g_nccl.GroupStart());
        for (int i = 0; i < m->n; i++) {
            SceneHost* r = &m->sc[i]->h;
            RT_NCCL(g_nccl.Broadcast((r->*mb).p, (r->*mb).p, bytes, ncclUint8, 0, m->comm[i], r->stream));
        }
        RT_NCCL(g_nccl.GroupEnd());
    }
    for (int i = 0; i < m->n; i++) {
        RT_CUDA(cudaSetDevice(m->dev[i]));
        RT_CUDA(cudaStreamSynchronize(m->sc[i]->h.stream));
    }
    RT_CUDA(cudaSetDevice(m->dev[0]));
    return RT_OK;
}

}  // namespace
}  // namespace rt

using namespace rt;

extern "C" {

void rt_free_scene_multi(RtMultiScene* m)
{
    if (!m) return;
    for (int i = 0; i < (int)m->sc.size(); i++) {
        if (i < (int)m->dev.size()) cudaSetDevice(m->dev[i]);
        if (m->sc[i]) { cudaStreamSynchronize(m->sc[i]->h.stream); }
        if (i < (int)m->send.size()) m->send[i].release();
        if (i < (int)m->gathered.size()) m->gathered[i].release();
        if (i == 0) {
            m->frame.release();
            if (m->ev0) cudaEventDestroy(m->ev0);
            if (m->ev1) cudaEventDestroy(m->ev1);
        }
        rt_free_scene(m->sc[i]);
    }
    if (!m->dev.empty()) cudaSetDevice(m->dev[0]);
    delete m;
}

int rt_upload_scene_multi(const RtSceneDesc* desc, int32_t num_gpus, const int32_t* device_ids, RtMultiScene** out)
{
    if (!desc || !out || num_gpus < 1) { set_error("rt_upload_scene_multi: bad argument"); return RT_E_INVALID; }
    *out = nullptr;
    int ndev = 0;
    if (!cuda_ok(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount") || ndev < 1) {
        set_error("rt_upload_scene_multi needs a CUDA device (there is no CPU fallback): " + std::string(rt_last_error()));
        return RT_E_CUDA;
    }
    if (num_gpus > ndev || num_gpus > 16) { set_error("rt_upload_scene_multi: " + std::to_string(num_gpus) + " GPUs asked for, " + std::to_string(ndev) + " visible (at most 16)"); return RT_E_INVALID; }
    RtMultiScene* m = new (std::nothrow) RtMultiScene();
    if (!m) { set_error("out of memory"); return RT_E_NOMEM; }
    m->n = num_gpus;
    for (int i = 0; i < num_gpus; i++) {
        const int d = device_ids ? device_ids[i] : i;
        if (d < 0 || d >= ndev) { set_error("rt_upload_scene_multi: bad device id"); rt_free_scene_multi(m); return RT_E_INVALID; }
        m->dev.push_back(d);
    }
    m->sc.assign(num_gpus, nullptr);
    m->send.resize(num_gpus); m->gathered.resize(num_gpus);
    int rc = RT_OK;
    if (num_gpus > 1) {
        if (!load_nccl()) { rt_free_scene_multi(m); return RT_E_CUDA; }
        if ((rc = get_comms(m->dev, &m->comm)) != RT_OK) { rt_free_scene_multi(m); return rc; }
    }
    // the scene itself: once, on the first device (LBVH included when the scene is large)
    if (!cuda_ok(cudaSetDevice(m->dev[0]), "cudaSetDevice")) rc = RT_E_CUDA;
    if (rc == RT_OK) rc = rt_upload_scene(desc, &m->sc[0]);
    if (rc == RT_OK && m->sc[0]->h.nf + m->sc[0]->h.ns > 48) rc = build_bvh(&m->sc[0]->h);
    if (rc == RT_OK && (!cuda_ok(cudaEventCreate(&m->ev0), "cudaEventCreate") || !cuda_ok(cudaEventCreate(&m->ev1), "cudaEventCreate"))) rc = RT_E_CUDA;
    for (int i = 1; i < num_gpus && rc == RT_OK; i++) {
        if (!cuda_ok(cudaSetDevice(m->dev[i]), "cudaSetDevice")) { rc = RT_E_CUDA; break; }
        rc = new_replica(&m->sc[i]);
    }
    if (rc == RT_OK) rc = replicate(m);
    if (rc != RT_OK) { rt_free_scene_multi(m); return rc; }
    *out = m;
    return RT_OK;
}

int rt_num_gpus_multi(const RtMultiScene* m) { return m ? m->n : 0; }

int rt_render_multi(RtMultiScene* m, const RtRenderParams* params, RtStats* stats)
{
    if (!m) { set_error("rt_render_multi: null scene"); return RT_E_INVALID; }
    RtRenderParams base;
    int rc = checked_params(params, &base);
    if (rc != RT_OK) return rc;
    SceneHost* s0 = &m->sc[0]->h;
    const int N = m->n, W = s0->width, H = s0->height;
    if (!params || params->tile_rows == 0) base.tile_rows = (H % 10 == 0) ? 10 : 8;       // 2160 -> 216 tiles, 4320 -> 432
    base.world_size = N;
    const TileLayout tl = tile_layout(H, W, N, base.tile_rows);
    const size_t bytes = (size_t)tl.n_local_pixels * sizeof(uchar4);
    int regrows = 0, launches = 0;
    std::vector<Counters> totals(N);
    for (;;) {
        for (int i = 0; i < N; i++) {
            RT_CUDA(cudaSetDevice(m->dev[i]));
            RtRenderParams p = base; p.rank = i;
            if ((rc = ensure_workspace(&m->sc[i]->h, p, nullptr)) != RT_OK) return rc;
            if (!m->send[i].reserve(bytes) || (N > 1 && !m->gathered[i].reserve(bytes * N))) return RT_E_NOMEM;
        }
        RT_CUDA(cudaSetDevice(m->dev[0]));
        if (!m->frame.reserve((size_t)W * H * sizeof(uchar4))) return RT_E_NOMEM;
        RT_CUDA(cudaEventRecord(m->ev0, s0->stream));
        launches = 0;
        for (int i = 0; i < N; i++) {
            RT_CUDA(cudaSetDevice(m->dev[i]));
            RtRenderParams p = base; p.rank = i;
            int n = 0;
            if ((rc = enqueue_frame(&m->sc[i]->h, p, m->send[i].as<uchar4>(), m->sc[i]->h.stream, &n)) != RT_OK) return rc;
            launches += n;
        }
        if (N > 1) {
            RT_NCCL(g_nccl.GroupStart());
            for (int i = 0; i < N; i++)
                RT_NCCL(g_nccl.AllGather(m->send[i].p, m->gathered[i].p, bytes, ncclUint8, m->comm[i], m->sc[i]->h.stream));
            RT_NCCL(g_nccl.GroupEnd());
        }
        RT_CUDA(cudaSetDevice(m->dev[0]));
        if ((rc = launch_deinterleave(s0, N > 1 ? m->gathered[0].p : m->send[0].p, m->frame.p, N, base.tile_rows, s0->stream)) != RT_OK) return rc;
        RT_CUDA(cudaEventRecord(m->ev1, s0->stream));
        bool again = false;
        for (int i = 0; i < N; i++) {
            RT_CUDA(cudaSetDevice(m->dev[i]));
            SceneHost* s = &m->sc[i]->h;
            RT_CUDA(cudaMemcpyAsync(s->h_counters, s->ws.counters.p, sizeof(Counters) * s->last_subframes, cudaMemcpyDeviceToHost, s->stream));
            RT_CUDA(cudaStreamSynchronize(s->stream));
            unsigned demand[kMaxLevels];
            const unsigned bits = frame_overflowed(s, &totals[i], demand);
            if ((rc = overflow_error(bits)) != RT_OK) return rc;
            if (bits) {
                again = true;
                if ((rc = ensure_workspace(s, s->last, demand)) != RT_OK) return rc;
            }
        }
        if (!again) break;
        if (++regrows > 16) { set_error("wavefront queues still overflow after 16 regrows"); return RT_E_NOMEM; }
    }
    RT_CUDA(cudaSetDevice(m->dev[0]));
    m->tile_rows = base.tile_rows;
    m->rendered = true;
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        float ms = 0.0f;
        RT_CUDA(cudaEventElapsedTime(&ms, m->ev0, m->ev1));
        stats->render_ms = ms;
        for (int i = 0; i < N; i++) {
            const Counters& c = totals[i];
            for (int l = 0; l <= base.max_depth; l++) {
                stats->nodes_per_level[l] += c.nodes[l];
                stats->nodes_total += c.nodes[l];
                if (l >= 1) { stats->rays_reflect += c.rays_reflect[l]; stats->rays_transmit += c.rays[l] - c.rays_reflect[l]; }
            }
            stats->bvh_nodes_visited += c.bvh_nodes; stats->prim_tests += c.prim_tests; stats->texel_fetches += c.texels;
            stats->shadow_boxes_tested += c.shadow_boxes; stats->shadow_prim_tests += c.shadow_prims;
            stats->shadow_rays_walked += c.shadow_walks; stats->shadow_chains_redone += c.redo_count;
            stats->shadow_map_hits = (int32_t)std::min<unsigned long long>((unsigned long long)stats->shadow_map_hits + c.shadow_map_hits, 0x7fffffffull);
            stats->graph_replays += m->sc[i]->h.graph_replays;
        }
        stats->rays_primary = (uint64_t)W * (uint64_t)H;
        stats->rays_shadow = stats->nodes_total * (uint64_t)s0->nl;
        if (!(base.instrument && s0->last_accel == RT_ACCEL_LBVH)) stats->shadow_rays_walked = stats->rays_shadow;
        stats->accel_used = s0->last_accel; stats->kernel_launches = launches + 1; stats->queue_regrows = regrows;
        stats->bvh_levels4 = s0->last_accel == RT_ACCEL_LBVH ? s0->bvh_levels4 : 0;
        stats->bvh_levels8 = s0->last_accel == RT_ACCEL_LBVH ? s0->bvh_levels8 : 0;
    }
    return RT_OK;
}

int rt_download_framebuffer_multi(RtMultiScene* m, uint8_t* rgb, uint8_t* flags)
{
    if (!m || !rgb) { set_error("rt_download_framebuffer_multi: null argument"); return RT_E_INVALID; }
    if (!m->rendered) { set_error("rt_download_framebuffer_multi before rt_render_multi"); return RT_E_STATE; }
    SceneHost* s = &m->sc[0]->h;
    RT_CUDA(cudaSetDevice(m->dev[0]));
    const size_t n = (size_t)s->width * s->height;
    const int rc = launch_split_planes(s, s->stream, m->frame.as<uchar4>(), (int)n);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaMemcpyAsync(rgb, s->ws.rgb24.p, n * 3, cudaMemcpyDeviceToHost, s->stream));
    if (flags) RT_CUDA(cudaMemcpyAsync(flags, s->ws.flags8.p, n, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

int rt_format_ppm_multi(RtMultiScene* m, const char** text, size_t* len)
{
    if (!m || !text || !len) { set_error("rt_format_ppm_multi: null argument"); return RT_E_INVALID; }
    if (!m->rendered) { set_error("rt_format_ppm_multi before rt_render_multi"); return RT_E_STATE; }
    RT_CUDA(cudaSetDevice(m->dev[0]));
    return rt_format_ppm_device(m->sc[0], m->frame.p, text, len);
}

}  /* extern "C" */
