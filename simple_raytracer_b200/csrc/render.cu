// render.cu — frame scheduling: workspace sizing, the launch sequence of one frame
// (generations 0..D forward, D-1..0 backward), and the framebuffer plumbing.
#include "bvh.cuh"
#include "host_scene.h"
#include "wavefront_bvh.cuh"

#include <algorithm>
#include <cstring>
#include <mutex>

namespace rt {

namespace {

int g_sm_count[64] = { 0 };

int sm_count(int device)
{
    if (device < 0 || device >= 64) return 148;
    if (!g_sm_count[device]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
        g_sm_count[device] = n;
    }
    return g_sm_count[device];
}

struct BruteSmemFactory {
    static RT_DEV BruteAccel make(const SceneView& S, float4* smem) { return make_brute<true>(S, smem); }
    static RT_DEV void flush(const BruteAccel&, Counters*) {}
};
struct BruteGlobalFactory {
    static RT_DEV BruteAccel make(const SceneView& S, float4* smem) { return make_brute<false>(S, smem); }
    static RT_DEV void flush(const BruteAccel&, Counters*) {}
};
// scanline-order planes for rt_download_framebuffer: tiles RGBA8 -> rgb24 + flags8
__global__ void split_planes_kernel(const uchar4* __restrict__ tiles, int n, unsigned char* __restrict__ rgb,
                                    unsigned char* __restrict__ flags)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uchar4 p = tiles[i];
    rgb[3 * (size_t)i] = p.x; rgb[3 * (size_t)i + 1] = p.y; rgb[3 * (size_t)i + 2] = p.z;
    flags[i] = p.w;
}

// [rank][local_tile][row][col] -> [row][col] after the allgather
__global__ void deinterleave_kernel(const uchar4* __restrict__ gathered, uchar4* __restrict__ out, int width, int height,
                                    int world, int tile_rows, int n_local_rows)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)width * height) return;
    const int row = (int)(i / width), col = (int)(i - (size_t)row * width);
    const int tile = row / tile_rows, in_tile = row - tile * tile_rows;
    const int rank = tile % world, local_tile = tile / world;
    const size_t src = ((size_t)rank * n_local_rows + (size_t)local_tile * tile_rows + in_tile) * width + col;
    out[i] = gathered[src];
}

std::mutex g_cache_mutex;
Workspace g_ws_cache;   // one retired workspace, reused by the next scene on the same device
bool g_ws_cache_valid = false;

}  // namespace

TileLayout tile_layout(int height, int width, int world, int tile_rows)
{
    TileLayout t;
    t.n_tiles_total = (height + tile_rows - 1) / tile_rows;
    t.n_local_tiles = (t.n_tiles_total + world - 1) / world;   // equal per rank: allgather needs equal counts
    t.n_local_rows = t.n_local_tiles * tile_rows;
    t.n_local_pixels = t.n_local_rows * width;
    return t;
}

void stash_workspace(Workspace&& ws)
{
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    g_ws_cache = std::move(ws);
    g_ws_cache_valid = true;
}

// Bytes of one queue slot of generation `level` (fill_levels lays the arrays out).
static size_t slot_bytes(int level, bool bvh, int nl)
{
    // level 0 has no ray queue (primary rays are generated in place)
    size_t per_slot = (level == 0 ? 0 : 4 * 16) + 3 * 16 + 8;
    if (bvh) per_slot += 2 * 16 + 4 + 4 + (size_t)nl * 4;    // hp, hb, node_list, walk_list, sh (wavefront_bvh.cuh)
    return per_slot;
}

// One sub-frame's slice of a generation's memory: its arrays, the slice rounded up to 256 bytes.
static size_t slice_bytes(int level, size_t cap, bool bvh, int nl)
{
    return (cap * slot_bytes(level, bvh, nl) + 64 + 255) / 256 * 256;
}

// ... and the whole generation: enough for any split into 1..kMaxSubframes slices of cap / K slots
// (K slices of cap / K slots need at most cap * per_slot + K * (64 + 255) bytes).
static size_t level_bytes(int level, size_t cap, bool bvh, int nl)
{
    return cap * slot_bytes(level, bvh, nl) + (size_t)kMaxSubframes * (64 + 256);
}

// Tuning of the LBVH pipeline.  These are the measured optima on B200 (sweeps under
// profiles/); a product build has no other source for them.  A lab build (-DRT_TUNING_ENV) lets the
// environment override each one for sweeps (RT_<NAME>).
struct Tuning {
    int refill_lanes = 6;            // a warp fetches new rays when this many lanes are idle
    int inner_min_lanes = 8;         // closest hit: leave the inner-node loop below this many descending lanes
    int leaf_min_lanes = 8;          // ... and the leaf loop below this many leaf holders
    int sh_refill_lanes = 6, sh_inner_min_lanes = 6, sh_leaf_min_lanes = 6;   // the same for the any-hit kernel
    int grid_mult = 16;              // persistent grids: blocks per SM asked for (the hardware keeps 7-8 resident)
    int subframes = 0;               // launch chains per frame: 0 = 2 from subframe_min_kpixels on, 1 below
    int subframe_min_kpixels = 3000;
    int steal_max_krays = 1500;      // closest-hit launches of at most this many thousand slots share work in their tail
    int steal_shadow_max_kpixels = 2500;   // ... the any-hit launch of chains of at most this many thousand pixels
    int colour_merge_max_kpixels = 2500;   // ... whose colour halves then also run as one launch
};

static const Tuning& tuning()
{
    static const Tuning t = [] {
        Tuning v;
#ifdef RT_TUNING_ENV
        v.refill_lanes = tuning_int("RT_REFILL_LANES", v.refill_lanes, 1, 32);
        v.inner_min_lanes = tuning_int("RT_INNER_MIN_LANES", v.inner_min_lanes, 1, 32);
        v.leaf_min_lanes = tuning_int("RT_LEAF_MIN_LANES", v.leaf_min_lanes, 1, 32);
        v.sh_refill_lanes = tuning_int("RT_SHADOW_REFILL_LANES", v.sh_refill_lanes, 1, 32);
        v.sh_inner_min_lanes = tuning_int("RT_SHADOW_INNER_MIN_LANES", v.sh_inner_min_lanes, 1, 32);
        v.sh_leaf_min_lanes = tuning_int("RT_SHADOW_LEAF_MIN_LANES", v.sh_leaf_min_lanes, 1, 32);
        v.grid_mult = tuning_int("RT_GRID_MULT", v.grid_mult, 1, 64);
        v.subframes = tuning_int("RT_SUBFRAMES", v.subframes, 0, kMaxSubframes);
        v.subframe_min_kpixels = tuning_int("RT_SUBFRAME_MIN_KPIXELS", v.subframe_min_kpixels, 0, 1 << 30);
        v.steal_max_krays = tuning_int("RT_STEAL_MAX_KRAYS", v.steal_max_krays, 0, 1 << 30);
        v.steal_shadow_max_kpixels = tuning_int("RT_STEAL_SHADOW_MAX_KPIXELS", v.steal_shadow_max_kpixels, 0, 1 << 30);
        v.colour_merge_max_kpixels = tuning_int("RT_COLOUR_MERGE_MAX_KPIXELS", v.colour_merge_max_kpixels, 0, 1 << 30);
#endif
        return v;
    }();
    return t;
}

// Launch chains of a frame.  Two interleaved chains pay when one chain's kernel tails are a small
// part of its launches (many pixels per GPU: the other chain's blocks fill them); below that one
// chain is faster.  Profiled frames and the brute-force path run one serial chain.
static int pick_subframes(const RtRenderParams& p, int accel, const TileLayout& tl)
{
    if (accel != RT_ACCEL_LBVH || p.profile) return 1;
    const Tuning& T = tuning();
    int k = p.subframes ? p.subframes : T.subframes;
    if (k == 0) k = (long long)tl.n_local_pixels >= 1000LL * T.subframe_min_kpixels ? 2 : 1;
    return std::max(1, std::min(k, tl.n_local_tiles));
}

bool ChainStreams::ensure(int n)
{
    if (!begin && !cuda_ok(cudaEventCreateWithFlags(&begin, cudaEventDisableTiming), "cudaEventCreate")) return false;
    for (int j = chains; j < n; j++) {
        // the trace -> spawn chain is the critical path: highest priority; shadow -> colour lowest
        int lo = 0, hi = 0;
        if (!cuda_ok(cudaDeviceGetStreamPriorityRange(&lo, &hi), "cudaDeviceGetStreamPriorityRange")) return false;
        if (!cuda_ok(cudaStreamCreateWithPriority(&main[j], cudaStreamNonBlocking, hi), "cudaStreamCreate")) return false;
        if (!cuda_ok(cudaStreamCreateWithPriority(&side[j], cudaStreamNonBlocking, lo), "cudaStreamCreate")) return false;
        if (!cuda_ok(cudaEventCreateWithFlags(&fork[j], cudaEventDisableTiming), "cudaEventCreate") ||
            !cuda_ok(cudaEventCreateWithFlags(&side_done[j], cudaEventDisableTiming), "cudaEventCreate") ||
            !cuda_ok(cudaEventCreateWithFlags(&main_done[j], cudaEventDisableTiming), "cudaEventCreate")) return false;
        chains = j + 1;
    }
    return true;
}

void ChainStreams::sync() const
{
    for (int j = 0; j < chains; j++) {
        if (main[j]) cudaStreamSynchronize(main[j]);
        if (side[j]) cudaStreamSynchronize(side[j]);
    }
}

void ChainStreams::release()
{
    for (int j = 0; j < kMaxSubframes; j++) {
        if (main[j]) cudaStreamDestroy(main[j]);
        if (side[j]) cudaStreamDestroy(side[j]);
        if (side_done[j]) cudaEventDestroy(side_done[j]);
        if (main_done[j]) cudaEventDestroy(main_done[j]);
        if (fork[j]) cudaEventDestroy(fork[j]);
        main[j] = side[j] = nullptr; side_done[j] = main_done[j] = fork[j] = nullptr;
    }
    if (begin) cudaEventDestroy(begin);
    begin = nullptr;
    chains = 0;
}

static int pick_accel(const SceneHost* s, int requested)
{
    if (s->nf + s->ns == 0) return RT_ACCEL_BRUTE;
    // Non-finite material / light values (a scene file may say "nan"): the tree path's early-outs — a shadow walk ends
    // at the first opaque occluder, a chain at a carried product of 0, an occluder map settles a ray — all rest on
    // every shadow factor lying in [0,1].  The brute-force kernel multiplies every factor in the reference's order.
    if (!s->zero_mask_exit) return RT_ACCEL_BRUTE;
    if (requested == RT_ACCEL_BRUTE || requested == RT_ACCEL_LBVH) return requested;
    return (s->nf + s->ns) > 48 ? RT_ACCEL_LBVH : RT_ACCEL_BRUTE;
}

// Sizes the queues.  demand == nullptr: first guess from the pixel count (or what a
// retired workspace already holds); otherwise grow to the observed demand.
int ensure_workspace(SceneHost* s, const RtRenderParams& p, const unsigned* demand)
{
    Workspace& ws = s->ws;
    if (ws.device < 0) {
        std::lock_guard<std::mutex> lock(g_cache_mutex);
        if (g_ws_cache_valid && g_ws_cache.device == s->device) {
            ws = std::move(g_ws_cache);
            g_ws_cache_valid = false;
        }
        ws.device = s->device;
    }
    const TileLayout tl = tile_layout(s->height, s->width, p.world_size, p.tile_rows);
    const size_t npx = (size_t)tl.n_local_pixels;
    const bool bvh = pick_accel(s, p.accel) == RT_ACCEL_LBVH;
    if (!ws.h_counters.reserve(sizeof(Counters) * kMaxSubframes)) return RT_E_NOMEM;
    s->h_counters = static_cast<Counters*>(ws.h_counters.p);
    if (!ws.counters.reserve(sizeof(Counters) * kMaxSubframes) || !ws.tiles.reserve(npx * sizeof(uchar4))) return RT_E_NOMEM;
    if (p.instrument && !ws.primary_ids.reserve(npx * sizeof(int))) return RT_E_NOMEM;
    if (bvh && !ws.redo_list.reserve((size_t)kRedoCapacity * kMaxSubframes * sizeof(unsigned))) return RT_E_NOMEM;
    for (int l = 0; l <= p.max_depth; l++) {
        size_t want;
        // level 0: one slot per pixel; the sub-frames' virtual layouts pad up to one tile each
        if (l == 0) want = npx + (size_t)kMaxSubframes * p.tile_rows * s->width;
        else if (demand) want = std::max<size_t>((size_t)ws.capacity[l], (size_t)(demand[l] * 1.25) + 1024);
        else {
            // rays of generation l per pixel seen on the example scenes: <= ~1.2, 0.5, 0.25, ...
            static const double guess[kMaxLevels] = { 1.0, 1.25, 0.75, 0.5, 0.5, 0.5, 0.5, 0.5 };
            want = std::max<size_t>((size_t)ws.capacity[l], (size_t)(npx * guess[l]) + 1024);
        }
        if (want > 0x0ffffff0u) { set_error("wavefront queue would exceed 2^28 slots"); return RT_E_NOMEM; }
        want = std::max<size_t>(want, (size_t)ws.capacity[l]);
        if ((size_t)ws.capacity[l] < want || ws.level_mem[l].bytes < level_bytes(l, want, bvh, s->nl)) {
            if (!ws.level_mem[l].reserve(level_bytes(l, want, bvh, s->nl))) return RT_E_NOMEM;
            ws.capacity[l] = (int)want;
        }
    }
    return RT_OK;
}

// Capacity of one sub-frame's queue of generation l when the level memory is cut K ways.
static size_t chain_capacity(const SceneHost* s, int l, int K, size_t chain_pixels)
{
    return l == 0 ? chain_pixels : (size_t)s->ws.capacity[l] / (size_t)K;
}

static void fill_levels(const SceneHost* s, int max_depth, bool bvh, int K, int j, size_t chain_pixels, FrameParams* P)
{
    const Workspace& ws = s->ws;
    for (int l = 0; l < kMaxLevels; l++) {
        LevelBuffers& L = P->level[l];
        std::memset(&L, 0, sizeof L);
        if (l > max_depth) continue;
        const size_t cap = chain_capacity(s, l, K, chain_pixels);
        char* base = ws.level_mem[l].as<char>() + (size_t)j * slice_bytes(l, cap, bvh, s->nl);
        L.capacity = (int)cap;
        if (l > 0) {
            L.rq_o = (float4*)base; base += cap * 16;
            L.rq_d = (float4*)base; base += cap * 16;
            L.rq_m = (int4*)base;   base += cap * 16;
            L.rq_s = (int4*)base;   base += cap * 16;
        }
        L.nr_a = (float4*)base; base += cap * 16;
        L.nr_b = (float4*)base; base += cap * 16;
        L.res = (float4*)base;  base += cap * 16;
        if (bvh) {
            L.hp = (float4*)base; base += cap * 16;
            L.hb = (float4*)base; base += cap * 16;
        }
        // 8- and 4-byte arrays after all the 16-byte ones (cap may be odd)
        L.nr_c = (int2*)base;   base += cap * 8;
        if (bvh) {
            L.sh = (float*)base;  base += cap * 4 * (size_t)s->nl;
            L.node_list = (int*)base; base += cap * 4;
            L.walk_list = (int*)base;
        }
    }
}

template <class Factory>
static void launch_level(const FrameParams& P, int level, int grid, size_t smem, cudaStream_t st)
{
    level_kernel<Factory><<<grid, kBlock, smem, st>>>(P, level);
}

// Persistent grids: whole waves of resident blocks, never more blocks than there is work.
static int persistent_grid(long long work, int sms, int mult, int block = kBlock)
{
    return (int)std::max<long long>(1, std::min<long long>((work + block - 1) / block, (long long)sms * mult));
}

template <bool kCount>
static void launch_trace(const FrameParams& P, int level, bool sharing, int grid, cudaStream_t st)
{
    if (sharing) bvh_trace_kernel<kCount, true><<<grid, kWalkBlock, 0, st>>>(P, level);
    else bvh_trace_kernel<kCount, false><<<grid, kWalkBlock, 0, st>>>(P, level);
}

static bool any_shadow_map(const FrameParams& P)
{
    for (int k = 0; k < kMaxShadowMaps; k++) if (P.scene.smap[k].cells != nullptr) return true;
    return false;
}

template <bool kCount>
static void launch_shadow(const FrameParams& P, int first, int last, bool sharing, int grid, cudaStream_t st)
{
    // a sample of the nodes against the occluder maps (which of them pay in this frame), then light 0's map for every
    // node: what it settles never enters the walk kernel
    if (any_shadow_map(P)) bvh_shadow_probe_kernel<<<32, 256, 0, st>>>(P, first);
    if (P.settled_first) bvh_shadow_settle_kernel<kCount><<<std::max(1, std::min(grid, 148 * 16)), 256, 0, st>>>(P, first, last);
    if (sharing) bvh_shadow_kernel<kCount, true><<<grid, kWalkBlock, 0, st>>>(P, first, last);
    else bvh_shadow_kernel<kCount, false><<<grid, kWalkBlock, 0, st>>>(P, first, last);
    // the chains it gave up (a handful at most): a few blocks that find the list empty otherwise
    bvh_shadow_redo_kernel<kCount><<<8, kBlock, 0, st>>>(P);
}

// Enqueues one frame on `stream`.  The queues must already be sized (ensure_workspace).
//
// Brute force (scenes of at most 48 primitives): one fused level_kernel per generation, then the
// bottom-up combine passes.
//
// LBVH pipeline: the rank's tiles are dealt out to K sub-frames (tile t of the rank's buffer ->
// sub-frame t % K), each a complete, independent chain with its own queues and counters.  A chain
// runs in two phases:
//   1. trace(0) -> spawn(0) -> trace(1) -> ... -> trace(D) -> spawn(D)  (spawn(D) only lists the hits)
//      on the chain's main stream: the closest-hit rays do not depend on the shadow rays;
//   2. ONE any-hit launch over the nodes of all generations, the redo of what it gave up, the colour
//      halves (one launch for small chains, one per generation otherwise) on the chain's side stream,
//      then the bottom-up combine passes on the main stream again.
// A persistent walk kernel ends with a tail in which a few long rays keep a few warps busy; with
// two chains (large frames) the other chain's blocks fill the SMs during that tail.
// Instrumented frames (params.instrument) run the very same schedule with the counting builds of
// the walk kernels; profiled frames (params.profile) run one chain serially, shadow and colour
// kernels per generation, with events between the kernels.
int enqueue_frame(SceneHost* s, const RtRenderParams& p, uchar4* tiles_out, cudaStream_t stream, int* launches)
{
    const TileLayout tl = tile_layout(s->height, s->width, p.world_size, p.tile_rows);
    const int accel = pick_accel(s, p.accel);
    if (accel == RT_ACCEL_LBVH && !s->bvh_built) {
        const int rc = build_bvh(s);
        if (rc != RT_OK) return rc;
    }
    const bool lbvh = accel == RT_ACCEL_LBVH;
    const Tuning& T = tuning();
    const int K = pick_subframes(p, accel, tl);
    const bool overlap = lbvh && !p.profile;
    const bool count = p.instrument != 0;
    // sub-frame j is virtual rank (rank + world*j) of (world*K): the same interleaved-tile arithmetic
    const TileLayout vl = tile_layout(s->height, s->width, p.world_size * K, p.tile_rows);
    const int D = p.max_depth;

    static thread_local FrameParams PS[kMaxSubframes];
    for (int j = 0; j < K; j++) {
        FrameParams& P = PS[j];
        std::memset(&P, 0, sizeof P);
        { const SceneView sv = scene_view(s); std::memcpy(&P.scene, &sv, sizeof sv); }     // (bytewise: padding stays defined)
        P.cam = s->cam;
        fill_levels(s, D, lbvh, K, j, (size_t)vl.n_local_pixels, &P);
        P.counters = s->ws.counters.as<Counters>() + j;
        P.tiles = tiles_out ? tiles_out : s->ws.tiles.as<uchar4>();
        P.primary_ids = p.instrument ? s->ws.primary_ids.as<int>() : nullptr;
        P.max_depth = D; P.epsilon = p.epsilon;
        P.rank = p.rank + p.world_size * j; P.world = p.world_size * K; P.tile_rows = p.tile_rows;
        P.n_local_rows = vl.n_local_rows; P.n_local_pixels = vl.n_local_pixels;
        P.out_mul = K; P.out_add = j; P.out_tiles = tl.n_local_tiles;
        P.refill_lanes = T.refill_lanes; P.inner_min_lanes = T.inner_min_lanes; P.leaf_min_lanes = T.leaf_min_lanes;
        P.sh_refill_lanes = T.sh_refill_lanes; P.sh_inner_min_lanes = T.sh_inner_min_lanes; P.sh_leaf_min_lanes = T.sh_leaf_min_lanes;
        P.redo_list = lbvh ? s->ws.redo_list.as<unsigned>() + (size_t)j * kRedoCapacity : nullptr;
        P.settled_first = (lbvh && P.scene.smap[0].cells != nullptr) ? 1 : 0;
    }

    ChainStreams& cs = s->ws.chains;
    if (overlap && !cs.ensure(K)) return RT_E_CUDA;
    int n = 0;
    // The launch sequence of the frame on `stream` (directly, or under stream capture: see below).
    auto issue = [&](cudaStream_t stream) -> int {
        RT_CUDA(cudaMemsetAsync(s->ws.counters.p, 0, sizeof(Counters) * K, stream));
        const int sms = sm_count(s->device);
        const size_t smem = (size_t)(s->nf * kTriWords + s->ns * kSphWords) * sizeof(float4);
        const bool stage = accel == RT_ACCEL_BRUTE && smem <= 48 * 1024;
        s->prof_levels = 0;
        if (p.profile) {
            for (int k = 0; k < kMaxLevels + 2; k++)
                if (!s->prof_ev[k]) RT_CUDA(cudaEventCreate(&s->prof_ev[k]));
            for (int k = 0; k < kMaxLevels; k++)
                for (int j = 0; j < 3; j++)
                    if (!s->prof_mid[k][j]) RT_CUDA(cudaEventCreate(&s->prof_mid[k][j]));
        }
        s->prof_split = p.profile && lbvh;
        if (overlap) {
            RT_CUDA(cudaEventRecord(cs.begin, stream));        // the counters are cleared
            for (int j = 0; j < K; j++) RT_CUDA(cudaStreamWaitEvent(cs.main[j], cs.begin, 0));
        }
        const int steal = p.work_sharing;      // 0: by launch size, 1: always, 2: never

        if (!lbvh) {
            for (int l = 0; l <= D; l++) {
                if (p.profile) RT_CUDA(cudaEventRecord(s->prof_ev[l], stream));
                const FrameParams& P = PS[0];
                const int grid = persistent_grid(P.level[l].capacity, sms, T.grid_mult);
                if (stage) launch_level<BruteSmemFactory>(P, l, grid, smem, stream);
                else launch_level<BruteGlobalFactory>(P, l, grid, 0, stream);
                n++;
            }
        } else if (p.profile) {
            // one serial chain, every kernel on its own between events (bench.py's per-kernel times)
            const FrameParams& P = PS[0];
            for (int l = 0; l <= D; l++) {
                const long long work = P.level[l].capacity;
                const bool sharing = steal == 1 || (steal == 0 && work <= 1000LL * T.steal_max_krays);
                RT_CUDA(cudaEventRecord(s->prof_ev[l], stream));
                if (count) launch_trace<true>(P, l, sharing, persistent_grid(work, sms, T.grid_mult, kWalkBlock), stream);
                else launch_trace<false>(P, l, sharing, persistent_grid(work, sms, T.grid_mult, kWalkBlock), stream);
                RT_CUDA(cudaEventRecord(s->prof_mid[l][2], stream));
                bvh_spawn_kernel<<<persistent_grid(work, sms, T.grid_mult), kBlock, 0, stream>>>(P, l);
                n += 2;
                RT_CUDA(cudaEventRecord(s->prof_mid[l][0], stream));
                if (s->nl) {
                    const int sh_grid = persistent_grid(work * s->nl, sms, T.grid_mult, kWalkBlock);
                    if (count) launch_shadow<true>(P, l, l, sharing, sh_grid, stream);
                    else launch_shadow<false>(P, l, l, sharing, sh_grid, stream);
                    n += 2 + P.settled_first + (any_shadow_map(P) ? 1 : 0);
                }
                RT_CUDA(cudaEventRecord(s->prof_mid[l][1], stream));
                bvh_colour_kernel<<<persistent_grid(work, sms, T.grid_mult), kBlock, 0, stream>>>(P, l, l);
                n++;
            }
        } else {
            for (int j = 0; j < K; j++) {
                const FrameParams& P = PS[j];
                cudaStream_t ms = cs.main[j], side = cs.side[j];
                long long all_work = 0;
                for (int l = 0; l <= D; l++) {
                    const long long work = P.level[l].capacity;
                    all_work += work;
                    const bool sharing = steal == 1 || (steal == 0 && work <= 1000LL * T.steal_max_krays);
                    const int wgrid = persistent_grid(work, sms, T.grid_mult, kWalkBlock);
                    if (count) launch_trace<true>(P, l, sharing, wgrid, ms);
                    else launch_trace<false>(P, l, sharing, wgrid, ms);
                    bvh_spawn_kernel<<<persistent_grid(work, sms, T.grid_mult), kBlock, 0, ms>>>(P, l);
                    n += 2;
                }
                RT_CUDA(cudaEventRecord(cs.fork[j], ms));
                RT_CUDA(cudaStreamWaitEvent(side, cs.fork[j], 0));
                const long long chain_pixels = P.level[0].capacity;
                if (s->nl) {
                    const bool sharing = steal == 1 || (steal == 0 && chain_pixels <= 1000LL * T.steal_shadow_max_kpixels);
                    const int sh_grid = persistent_grid(all_work * s->nl, sms, T.grid_mult, kWalkBlock);
                    if (count) launch_shadow<true>(P, 0, D, sharing, sh_grid, side);
                    else launch_shadow<false>(P, 0, D, sharing, sh_grid, side);
                    n += 2 + P.settled_first + (any_shadow_map(P) ? 1 : 0);
                }
                if (chain_pixels <= 1000LL * T.colour_merge_max_kpixels) {
                    bvh_colour_kernel<<<persistent_grid(all_work, sms, T.grid_mult), kBlock, 0, side>>>(P, 0, D);
                    n++;
                } else for (int l = 0; l <= D; l++) {
                    bvh_colour_kernel<<<persistent_grid(P.level[l].capacity, sms, T.grid_mult), kBlock, 0, side>>>(P, l, l);
                    n++;
                }
            }
        }
        const int prof_n = D + 1;
        if (p.profile) RT_CUDA(cudaEventRecord(s->prof_ev[prof_n], stream));
        // bottom-up combine per chain, on the chain's main stream once its colour kernels are done
        for (int j = 0; j < K; j++) {
            const FrameParams& P = PS[j];
            cudaStream_t ms = overlap ? cs.main[j] : stream;
            if (overlap) {
                RT_CUDA(cudaEventRecord(cs.side_done[j], cs.side[j]));
                RT_CUDA(cudaStreamWaitEvent(ms, cs.side_done[j], 0));
            }
            for (int l = std::max(D - 1, 0); l >= 0; l--) {
                // (max_depth == 0: no backward pass, but generation 0 still has to be quantised)
                const long long work = P.level[l].capacity;
                const int grid = (int)std::max<long long>(1, std::min<long long>((work + 255) / 256, (long long)sms * 8));
                combine_kernel<<<grid, 256, 0, ms>>>(P, l);
                n++;
            }
            if (overlap) {
                RT_CUDA(cudaEventRecord(cs.main_done[j], ms));
                RT_CUDA(cudaStreamWaitEvent(stream, cs.main_done[j], 0));
            }
        }
        RT_CUDA(cudaGetLastError());
        if (p.profile) {
            RT_CUDA(cudaEventRecord(s->prof_ev[prof_n + 1], stream));
            s->prof_levels = prof_n;
        }
        if (tiles_out && tiles_out != s->ws.tiles.as<uchar4>())
            RT_CUDA(cudaMemcpyAsync(s->ws.tiles.p, tiles_out, (size_t)tl.n_local_pixels * sizeof(uchar4),
                                    cudaMemcpyDeviceToDevice, stream));
        return RT_OK;
    };

    // CUDA-graph replay (SURVEY 8b "captured in a CUDA Graph for replays", 8f-4): a frame is fully described by
    // the FrameParams its kernels receive, so a frame whose FrameParams equal the previous frame's (same
    // buffers, camera, sharding, parameters) is captured once — on its second sighting, so that a
    // load-render-free sequence never pays for an instantiation — and replayed from then on: one graph launch
    // instead of 11-30 kernel launches + event traffic.  Profiled / instrumented frames and frames enqueued on a
    // stream the caller is capturing are issued directly.
    FrameGraph& G = s->graph;
    if (!p.profile) { s->prof_levels = 0; s->prof_split = false; }
    bool graphable = !p.profile && !p.instrument;
    if (graphable) {
        cudaStreamCaptureStatus cst = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cst) != cudaSuccess || cst != cudaStreamCaptureStatusNone) { cudaGetLastError(); graphable = false; }
    }
    const bool same = graphable && G.K == K && G.accel == accel && std::memcmp(G.ps, PS, sizeof(FrameParams) * K) == 0;
    if (same && G.exec) {
        RT_CUDA(cudaGraphLaunch(G.exec, stream));
        n = G.launches;
        s->graph_replays++;
    } else if (same) {
        cudaGraph_t graph = nullptr;
        RT_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
        const int rc = issue(stream);
        const cudaError_t ce = cudaStreamEndCapture(stream, &graph);
        if (rc != RT_OK || ce != cudaSuccess || !graph) {
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            if (rc != RT_OK) return rc;
            set_error(std::string("cudaStreamEndCapture: ") + cudaGetErrorString(ce));
            return RT_E_CUDA;
        }
        const cudaError_t ci = cudaGraphInstantiate(&G.exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ci != cudaSuccess) { G.exec = nullptr; set_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ci)); return RT_E_CUDA; }
        G.launches = n;
        RT_CUDA(cudaGraphLaunch(G.exec, stream));
    } else {
        if (G.exec) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
        G.K = graphable ? K : 0; G.accel = accel;
        if (graphable) std::memcpy(G.ps, PS, sizeof(FrameParams) * K);
        const int rc = issue(stream);
        if (rc != RT_OK) return rc;
    }
    std::memcpy(s->last_ps, PS, sizeof(FrameParams) * K);
    s->rendered = true;
    s->last = p;
    s->last_local_rows = tl.n_local_rows; s->last_local_pixels = tl.n_local_pixels;
    s->last_accel = accel; s->last_launches = n + 1; s->last_subframes = K;
    s->have_primary_ids = p.instrument != 0;
    if (launches) *launches = n + 1;
    return RT_OK;
}

// Sums the sub-frames' counters of the last frame and checks every queue against the slice
// it had.  demand: what the whole generation would have needed had every slice been as full
// as the fullest one.  Returns the kOverflow* bits of the frame (0 = complete).
unsigned frame_overflowed(const SceneHost* s, Counters* total, unsigned demand[kMaxLevels])
{
    const int K = s->last_subframes;
    const RtRenderParams& p = s->last;
    Counters t;
    std::memset(&t, 0, sizeof t);
    unsigned overflow = 0;
    double factor = 1.0;
    unsigned fullest[kMaxLevels] = { 0 };
    for (int j = 0; j < K; j++) {
        const Counters& c = s->h_counters[j];
        overflow |= c.overflow;
        for (int l = 0; l < kMaxLevels; l++) {
            t.rays[l] += c.rays[l]; t.nodes[l] += c.nodes[l]; t.rays_reflect[l] += c.rays_reflect[l];
            fullest[l] = std::max(fullest[l], c.rays[l]);
            if (l >= 1 && l <= p.max_depth) {
                const unsigned cap = (unsigned)(s->ws.capacity[l] / K);
                if (c.rays[l] > cap) {
                    overflow |= kOverflowQueue;
                    // a truncated generation under-reports what the deeper ones need: scale those up with it
                    factor = std::max(factor, (double)c.rays[l] / (double)std::max(1u, cap));
                }
            }
        }
        t.bvh_nodes += c.bvh_nodes; t.prim_tests += c.prim_tests; t.texels += c.texels;
        t.shadow_boxes += c.shadow_boxes; t.shadow_prims += c.shadow_prims; t.shadow_walks += c.shadow_walks;
        t.shadow_map_hits += c.shadow_map_hits;
        t.redo_count += c.redo_count;
        if (j == 0) std::memcpy(t.clock, c.clock, sizeof t.clock);      // (the first chain's)
    }
    t.overflow = overflow;
    if (total) *total = t;
    if (demand)
        for (int l = 0; l < kMaxLevels; l++)
            demand[l] = (unsigned)std::min(2.0e8, (double)fullest[l] * K * factor + 64.0 * K);
    return overflow;
}

// What rt_render / rt_async_frame_status make of a frame's overflow bits other than kOverflowQueue.
int overflow_error(unsigned bits)
{
    if (bits & kOverflowStack) {
        set_error("a BVH traversal stack overflowed (degenerate tree): the frame is incomplete");
        return RT_E_LIMIT;
    }
    if (bits & kOverflowRedo) {
        set_error("more than 65536 shadow chains with over 12 transparent occluders on one ray in a launch chain: the frame is incomplete");
        return RT_E_LIMIT;
    }
    return RT_OK;
}

// Identity dump of the last LBVH frame (include/rt.h: rt_download_hit_histograms): per generation, which objects
// ShadeRay was invoked on, and in which class (== 0, in between, == 1) the carried shadow_mask was after each light.
__global__ void hit_histogram_kernel(const __grid_constant__ FrameParams P, int n_obj, unsigned long long* __restrict__ hits,
                                     unsigned long long* __restrict__ shadow)
{
    const int nl = P.scene.nl;
    for (int level = 0; level <= P.max_depth; level++) {
        const LevelBuffers& LB = P.level[level];
        const unsigned n = P.counters->nodes[level];
        for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
            const int obj = __float_as_int(LB.hp[LB.node_list[k]].w);
            atomicAdd(&hits[(size_t)level * n_obj + obj], 1ull);
            float mask = 1.0f;                                   // 788
            for (int li = 0; li < nl; li++) {
                if (mask != 0.0f) mask = LB.sh[(size_t)li * n + k];   // a chain stops once the mask is 0 (and 0 stays 0)
                atomicAdd(&shadow[((size_t)level * nl + li) * 3 + (mask == 0.0f ? 0 : mask == 1.0f ? 2 : 1)], 1ull);
            }
        }
    }
}

int launch_hit_histograms(const SceneHost* s, unsigned long long* dev_hits, unsigned long long* dev_shadow, cudaStream_t st)
{
    for (int j = 0; j < s->last_subframes; j++)
        hit_histogram_kernel<<<148 * 4, 256, 0, st>>>(s->last_ps[j], s->nf + s->ns, dev_hits, dev_shadow);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

int launch_split_planes(SceneHost* s, cudaStream_t st, const uchar4* src, int n_pixels)
{
    const int n = src ? n_pixels : s->last_local_pixels;
    if (!s->ws.rgb24.reserve((size_t)n * 3) || !s->ws.flags8.reserve((size_t)n)) return RT_E_NOMEM;
    split_planes_kernel<<<(n + 255) / 256, 256, 0, st>>>(src ? src : s->ws.tiles.as<uchar4>(), n, s->ws.rgb24.as<unsigned char>(),
                                                        s->ws.flags8.as<unsigned char>());
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

int launch_deinterleave(const SceneHost* s, const void* gathered, void* out, int world, int tile_rows, cudaStream_t st)
{
    const TileLayout tl = tile_layout(s->height, s->width, world, tile_rows);
    const size_t n = (size_t)s->width * s->height;
    deinterleave_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const uchar4*)gathered, (uchar4*)out, s->width,
                                                                    s->height, world, tile_rows, tl.n_local_rows);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

}  // namespace rt
