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

static size_t level_bytes(int level, size_t cap, bool bvh, int nl)
{
    // level 0 has no ray queue (primary rays are generated in place)
    size_t per_slot = (level == 0 ? 0 : 4 * 16) + 3 * 16 + 8;
    if (bvh) per_slot += 2 * 16 + 4 + (size_t)nl * 4;    // hp, hb, node_list, sh (wavefront_bvh.cuh)
    // slop: every sub-frame's slice starts on a 256-byte boundary
    return cap * per_slot + 512 * (kMaxSubframes + 1);
}

// Launch chains of a frame.  Sub-frames only pay when the per-kernel tails are a large part
// of the frame (few pixels per GPU); profiled / instrumented frames and the brute-force
// path run as one serial chain.
static int pick_subframes(const RtRenderParams& p, int accel, const TileLayout& tl)
{
    if (accel != RT_ACCEL_LBVH || p.profile || p.instrument) return 1;
    static const int overlap_knob = tuning_int("RT_OVERLAP", 1, 0, 1);
    if (!overlap_knob) return 1;
    static const int knob = tuning_int("RT_SUBFRAMES", 0, 0, kMaxSubframes);
    int k = p.subframes ? p.subframes : knob;
    // measured (tools/merge_sweep2.sh, two-phase frame): two interleaved chains pay from about 3 M pixels
    // per rank on (4K: 40.1 -> 39.3 ms whole, 21.2 -> 20.5 ms half), one chain below; see DESIGN.md 5.4
    if (k == 0) k = tl.n_local_pixels >= 3000000 ? 2 : 1;
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
        for (int l = 0; l < kMaxLevels; l++)
            if (!cuda_ok(cudaEventCreateWithFlags(&fork[j][l], cudaEventDisableTiming), "cudaEventCreate")) return false;
        if (!cuda_ok(cudaEventCreateWithFlags(&side_done[j], cudaEventDisableTiming), "cudaEventCreate") ||
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
        main[j] = side[j] = nullptr; side_done[j] = main_done[j] = nullptr;
        for (int l = 0; l < kMaxLevels; l++) {
            if (fork[j][l]) cudaEventDestroy(fork[j][l]);
            fork[j][l] = nullptr;
        }
    }
    if (begin) cudaEventDestroy(begin);
    begin = nullptr;
    chains = 0;
}

static int pick_accel(const SceneHost* s, int requested)
{
    if (s->nf + s->ns == 0) return RT_ACCEL_BRUTE;
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
    if (bvh && s->nl > 0 && !ws.hints.reserve((size_t)(s->nf + s->ns) * s->nl * sizeof(int))) return RT_E_NOMEM;
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
        if (want > 0x7ffffff0u) { set_error("wavefront queue would exceed 2^31 slots"); return RT_E_NOMEM; }
        want = std::max<size_t>(want, (size_t)ws.capacity[l]);
        if ((size_t)ws.capacity[l] < want || ws.level_mem[l].bytes < level_bytes(l, want, bvh, s->nl)) {
            if (!ws.level_mem[l].reserve(level_bytes(l, want, bvh, s->nl))) return RT_E_NOMEM;
            ws.capacity[l] = (int)want;
        }
    }
    if (bvh) {
        size_t entries = 0;
        for (int l = 1; l <= p.max_depth; l++) entries += (size_t)ws.capacity[l];
        if (!ws.work_order.reserve(std::max<size_t>(entries, 1) * sizeof(unsigned))) return RT_E_NOMEM;
        size_t hits = 0;
        for (int l = 0; l < p.max_depth; l++) hits += l == 0 ? npx + (size_t)kMaxSubframes * p.tile_rows * s->width : (size_t)ws.capacity[l];
        if (!ws.hit_order.reserve(std::max<size_t>(hits, 1) * sizeof(unsigned))) return RT_E_NOMEM;
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
        const size_t slice = (level_bytes(l, cap, bvh, s->nl) - 512 * kMaxSubframes + 255) / 256 * 256;
        char* base = ws.level_mem[l].as<char>() + (size_t)j * slice;
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
            L.node_list = (int*)base;
        }
    }
}

template <class Factory>
static void launch_level(const FrameParams& P, int level, int grid, size_t smem, cudaStream_t st)
{
    level_kernel<Factory><<<grid, kBlock, smem, st>>>(P, level);
}

// ... of a walk kernel (blocks of kWalkBlock threads; mult counts 128-thread blocks per SM)
static int walk_grid(long long work, int sms, int mult)
{
    return (int)std::max<long long>(1, std::min<long long>((work + kWalkBlock - 1) / kWalkBlock, (long long)sms * mult * kWalkScale));
}

static int persistent_grid(long long work, int sms, int mult)
{
    return (int)std::max<long long>(1, std::min<long long>((work + kBlock - 1) / kBlock, (long long)sms * mult));
}

// Enqueues one frame on `stream`.  The queues must already be sized (ensure_workspace).
//
// LBVH pipeline: the rank's tiles are dealt out to K sub-frames (tile t of the rank's buffer
// -> sub-frame t % K), each a complete, independent chain of generation launches with its own
// queues and counters: trace(g) -> spawn(g) -> trace(g+1) ... on one stream and
// shadow(g) -> colour(g) on a second one.  A persistent walk kernel ends with a tail in which
// a few long rays keep a few warps busy; with 2K streams the other chains' blocks fill the
// SMs during that tail (the frame of a 1/8 shard is otherwise half tails).
int enqueue_frame(SceneHost* s, const RtRenderParams& p, uchar4* tiles_out, cudaStream_t stream, int* launches)
{
    const TileLayout tl = tile_layout(s->height, s->width, p.world_size, p.tile_rows);
    const int accel = pick_accel(s, p.accel);
    if (accel == RT_ACCEL_LBVH && !s->bvh_built) {
        const int rc = build_bvh(s);
        if (rc != RT_OK) return rc;
    }
    const bool lbvh = accel == RT_ACCEL_LBVH;
    // Fused trace: every generation's closest-hit rays in ONE persistent launch that spawns as it goes
    // (wavefront_bvh.cuh: bvh_trace_all_kernel), then every generation's shadow chains in one launch
    static const int fused_env = tuning_int("RT_FUSED", 0, 0, 1);
    static const int spawn_every = tuning_int("RT_SPAWN_EVERY", 8, 2, 64);
    const bool fused = lbvh && (p.fused_trace == 1 || (p.fused_trace == 0 && fused_env)) && p.subframes <= 1;
    const int K = fused ? 1 : pick_subframes(p, accel, tl);
    static const int overlap_knob = tuning_int("RT_OVERLAP", 1, 0, 1);
    const bool overlap = lbvh && overlap_knob && !p.profile && !p.instrument && !fused;
    // sub-frame j is virtual rank (rank + world*j) of (world*K): the same interleaved-tile arithmetic
    const TileLayout vl = tile_layout(s->height, s->width, p.world_size * K, p.tile_rows);

    static const int refill = tuning_int("RT_REFILL_LANES", 6, 1, 32);
    static const int grid_mult = tuning_int("RT_GRID_MULT", 16, 1, 64);
    static const int inner_min = tuning_int("RT_INNER_MIN_LANES", 8, 1, 32);
    static const int leaf_min = tuning_int("RT_LEAF_MIN_LANES", 8, 1, 32);
    // ... of the any-hit kernel (0 = as the closest-hit kernel)
    static const int sh_refill = tuning_int("RT_SHADOW_REFILL_LANES", 0, 0, 32);
    static const int sh_inner_min = tuning_int("RT_SHADOW_INNER_MIN_LANES", 6, 0, 32);
    static const int sh_leaf_min = tuning_int("RT_SHADOW_LEAF_MIN_LANES", 6, 0, 32);
    static const int near_first = tuning_int("RT_SHADOW_NEAR_FIRST", 0, 0, 1);
    // Work sharing in the walk kernels' tails.  0: never, 1: always, 2: for launches of at most
    // RT_STEAL_MAX_KRAYS thousand queue slots.  Measured on the 4K stress frame (tools/steal_sweep.sh):
    // 1/8 shard 7.43 -> 7.00 ms, 1/4 shard 12.34 -> 12.22, whole frame 42.3 -> 44.5 (the sharing
    // build's main loop carries more state), hence the size rule.
    static const int steal_env = tuning_int("RT_STEAL", 2, 0, 2);
    static const int steal_shadow_env = tuning_int("RT_STEAL_SHADOW", 2, 0, 2);
    static const long long steal_max_work = 1000LL * tuning_int("RT_STEAL_MAX_KRAYS", 1500, 0, 2000000);
    static const long long steal_shadow_max_work = 1000LL * tuning_int("RT_STEAL_SHADOW_MAX_KRAYS", 2500, 0, 2000000);   // merged any-hit launch: by the chain's pixels
    const int steal = p.work_sharing == 1 ? 1 : p.work_sharing == 2 ? 0 : steal_env;
    const int steal_shadow = p.work_sharing == 1 ? 1 : p.work_sharing == 2 ? 0 : steal_shadow_env;
    // Any-hit blocks that end after a bounded number of jobs + a high-priority trace chain: a trace
    // launch that becomes ready gets SM slots at once instead of waiting behind persistent shadow blocks
    static const int shadow_budget = tuning_int("RT_SHADOW_JOB_BUDGET", 0, 0, 1 << 20);
    static const int shadow_mult = tuning_int("RT_SHADOW_GRID_MULT", 0, 0, 64);   // 0: same as RT_GRID_MULT
    // Occluder hints of the any-hit kernel (wavefront_bvh.cuh: ShadowJob) and the interleaved passes
    // that make them available in time
    // Generations from this one on share ONE any-hit launch after the last trace (fewer kernel tails,
    // less overlap with the trace chain); 99 = every generation has its own
    static const int merge_env = tuning_int("RT_SHADOW_MERGE_FROM", 0, 0, 99);
    static const int colour_merge = tuning_int("RT_COLOUR_MERGE", 1, 0, 1);   // ... and one launch for their colour halves
    static const long long merge_max_work = 1000LL * tuning_int("RT_SHADOW_MERGE_MAX_KRAYS", 2000000, 0, 2000000);
    static const int hint_env = tuning_int("RT_HINT", 0, 0, 1);
    static const int hint_passes = tuning_int("RT_HINT_PASSES", 16, 1, 256);
    static const int hint_group = tuning_int("RT_HINT_GROUP", 32, 1, 4096);
    static const int hint_min_chunk = 1024 * tuning_int("RT_HINT_MIN_CHUNK_K", 256, 1, 1 << 20);
    static const int hint_mode = tuning_int("RT_HINT_MODE", 3, 0, 3);
    static const int hint_keep = tuning_int("RT_HINT_KEEP", 0, 0, 1);      // experiments only: hints survive the frame
    const bool hints = lbvh && s->nl > 0 && (p.occluder_hints == 1 || (p.occluder_hints == 0 && hint_env));

    // L1 / shared-memory split of the walk kernels (percent of the maximum shared memory; -1 = the driver's choice)
    static const int carveout = tuning_int("RT_CARVEOUT", -1, -1, 100);
    static bool carveout_set = false;
    if (carveout >= 0 && !carveout_set) {
        carveout_set = true;
        cudaFuncSetAttribute(bvh_trace_kernel<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        cudaFuncSetAttribute(bvh_trace_kernel<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        cudaFuncSetAttribute(bvh_shadow_kernel<false, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        cudaFuncSetAttribute(bvh_shadow_kernel<false, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
    }
    static thread_local FrameParams PS[kMaxSubframes];
    for (int j = 0; j < K; j++) {
        FrameParams& P = PS[j];
        std::memset(&P, 0, sizeof P);
        P.scene = scene_view(s);
        P.cam = s->cam;
        fill_levels(s, p.max_depth, lbvh, K, j, (size_t)vl.n_local_pixels, &P);
        P.counters = s->ws.counters.as<Counters>() + j;
        P.tiles = tiles_out ? tiles_out : s->ws.tiles.as<uchar4>();
        P.primary_ids = p.instrument ? s->ws.primary_ids.as<int>() : nullptr;
        P.max_depth = p.max_depth; P.epsilon = p.epsilon;
        P.rank = p.rank + p.world_size * j; P.world = p.world_size * K; P.tile_rows = p.tile_rows;
        P.n_local_rows = vl.n_local_rows; P.n_local_pixels = vl.n_local_pixels;
        P.out_mul = K; P.out_add = j; P.out_tiles = tl.n_local_tiles;
        P.refill_lanes = refill; P.inner_min_lanes = inner_min; P.leaf_min_lanes = leaf_min;
        P.sh_refill_lanes = sh_refill ? sh_refill : refill;
        P.sh_inner_min_lanes = sh_inner_min ? sh_inner_min : inner_min;
        P.sh_leaf_min_lanes = sh_leaf_min ? sh_leaf_min : leaf_min;
        P.steal = 1;
        P.shadow_job_budget = overlap ? shadow_budget : 0;
        P.occluder_hint = hints ? s->ws.hints.as<int>() : nullptr;
        P.hint_prims = s->nf + s->ns;
        P.hint_passes = hint_passes; P.hint_group = hint_group; P.hint_min_chunk = hint_min_chunk;
        P.hint_mode = hint_mode;
        if (fused) {
            size_t entries = 0;
            for (int l = 1; l <= p.max_depth; l++) entries += (size_t)s->ws.capacity[l];
            P.work_order = s->ws.work_order.as<unsigned>();
            P.work_capacity = (unsigned)std::min<size_t>(entries, 0xffffffffu);
            size_t hits = 0;
            for (int l = 0; l < p.max_depth; l++) hits += (size_t)P.level[l].capacity;
            P.hit_order = s->ws.hit_order.as<unsigned>();
            P.hit_capacity = (unsigned)std::min<size_t>(hits, 0xffffffffu);
            P.spawn_every = spawn_every;
        }
    }

    RT_CUDA(cudaMemsetAsync(s->ws.counters.p, 0, sizeof(Counters) * K, stream));
    // every frame starts without hints: a frame never depends on an earlier one
    if (hints && !(hint_keep && s->rendered)) RT_CUDA(cudaMemsetAsync(s->ws.hints.p, 0xff, (size_t)(s->nf + s->ns) * s->nl * sizeof(int), stream));
    const int sms = sm_count(s->device);
    const size_t smem = (size_t)(s->nf * kTriWords + s->ns * kSphWords) * sizeof(float4);
    const bool stage = accel == RT_ACCEL_BRUTE && smem <= 48 * 1024;
    int n = 0;
    s->prof_levels = 0;
    if (p.profile) {
        for (int k = 0; k < kMaxLevels + 2; k++)
            if (!s->prof_ev[k]) RT_CUDA(cudaEventCreate(&s->prof_ev[k]));
        for (int k = 0; k < kMaxLevels; k++)
            for (int j = 0; j < 2; j++)
                if (!s->prof_mid[k][j]) RT_CUDA(cudaEventCreate(&s->prof_mid[k][j]));
    }
    s->prof_split = p.profile && lbvh;
    ChainStreams& cs = s->ws.chains;
    if (overlap) {
        if (!cs.ensure(K)) return RT_E_CUDA;
        RT_CUDA(cudaEventRecord(cs.begin, stream));        // the counters are cleared
        for (int j = 0; j < K; j++) RT_CUDA(cudaStreamWaitEvent(cs.main[j], cs.begin, 0));
    }
    if (fused) {
        const FrameParams& P = PS[0];
        if (P.work_capacity) RT_CUDA(cudaMemsetAsync(P.work_order, 0xff, (size_t)P.work_capacity * sizeof(unsigned), stream));
        if (P.hit_capacity) RT_CUDA(cudaMemsetAsync(P.hit_order, 0xff, (size_t)P.hit_capacity * sizeof(unsigned), stream));
        // exactly the blocks that are resident at once: a block only ends when the frame is traced
        static int resident[2] = { 0, 0 };
        const int ri = p.instrument ? 1 : 0;
        if (!resident[ri]) {
            int nb = 0;
            if (p.instrument) RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, bvh_trace_all_kernel<true>, kWalkBlock, 0));
            else RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, bvh_trace_all_kernel<false>, kWalkBlock, 0));
            resident[ri] = std::max(nb, 1);
        }
        // (at least one spawner warp: warp w spawns when w % spawn_every == spawn_every - 1)
        const int grid = std::max((int)std::min<long long>(((long long)P.n_local_pixels + kWalkBlock - 1) / kWalkBlock, (long long)sms * resident[ri]), (spawn_every * 32 + kWalkBlock - 1) / kWalkBlock);
        if (p.profile) RT_CUDA(cudaEventRecord(s->prof_ev[0], stream));
        if (p.instrument) bvh_trace_all_kernel<true><<<grid, kWalkBlock, 0, stream>>>(P);
        else bvh_trace_all_kernel<false><<<grid, kWalkBlock, 0, stream>>>(P);
        n++;
        if (p.profile) RT_CUDA(cudaEventRecord(s->prof_mid[0][0], stream));
        if (s->nl) {
            long long work = 0;
            for (int l = 0; l <= p.max_depth; l++) work += P.level[l].capacity;
            const int sh_grid = walk_grid(work * std::max(1, s->nl), sms, shadow_mult ? shadow_mult : grid_mult);
            if (p.instrument) { if (hints) bvh_shadow_kernel<true, false, false, true, true><<<sh_grid, kWalkBlock, 0, stream>>>(P, -1); else bvh_shadow_kernel<true, false, false, true><<<sh_grid, kWalkBlock, 0, stream>>>(P, -1); }
            else { if (hints) bvh_shadow_kernel<false, false, false, true, true><<<sh_grid, kWalkBlock, 0, stream>>>(P, -1); else bvh_shadow_kernel<false, false, false, true><<<sh_grid, kWalkBlock, 0, stream>>>(P, -1); }
            n++;
        }
        if (p.profile) RT_CUDA(cudaEventRecord(s->prof_mid[0][1], stream));
        for (int l = 0; l <= p.max_depth; l++) {
            const int cgrid = persistent_grid(P.level[l].capacity, sms, grid_mult);
            bvh_shade_kernel<true, false><<<cgrid, kBlock, 0, stream>>>(P, l);
            n++;
        }
    }
    for (int l = 0; l <= p.max_depth && !fused; l++) {
        if (p.profile) RT_CUDA(cudaEventRecord(s->prof_ev[l], stream));
        for (int j = 0; j < K; j++) {
            const FrameParams& P = PS[j];
            // persistent grid: whole waves of resident blocks; generation 0 has one slot per pixel
            const long long work = P.level[l].capacity;
            const int grid = persistent_grid(work, sms, grid_mult);
            const int wgrid = walk_grid(work, sms, grid_mult);
            if (!lbvh) {
                if (stage) launch_level<BruteSmemFactory>(P, l, grid, smem, stream);
                else launch_level<BruteGlobalFactory>(P, l, grid, 0, stream);
                n++;
                continue;
            }
            // trace: one thread per ray; shadow: one chain per node; colour / spawn: one thread per node
            cudaStream_t ms = overlap ? cs.main[j] : stream;
            int sh_grid = walk_grid(work * std::max(1, s->nl), sms, shadow_mult ? shadow_mult : grid_mult);
            if (overlap && shadow_budget > 0)   // bounded blocks: as many as the queue may need (surplus blocks end at once)
                sh_grid = (int)std::max<long long>(sh_grid, std::min<long long>((work + 4LL * shadow_budget - 1) / (4LL * shadow_budget) + 1, 1 << 20));
            const bool sharing = steal == 1 || (steal == 2 && work <= steal_max_work);
            if (p.instrument) {
                if (sharing) bvh_trace_kernel<true, true><<<wgrid, kWalkBlock, 0, ms>>>(P, l);
                else bvh_trace_kernel<true, false><<<wgrid, kWalkBlock, 0, ms>>>(P, l);
            } else {
                if (sharing) bvh_trace_kernel<false, true><<<wgrid, kWalkBlock, 0, ms>>>(P, l);
                else bvh_trace_kernel<false, false><<<wgrid, kWalkBlock, 0, ms>>>(P, l);
            }
            n++;
            // shadow -> colour depends only on this generation's hits, spawn -> next trace too:
            // two chains on two streams (profiled / instrumented frames stay serial)
            cudaStream_t side = ms;
            if (overlap) {
                RT_CUDA(cudaEventRecord(cs.fork[j][l], ms));
                RT_CUDA(cudaStreamWaitEvent(cs.side[j], cs.fork[j][l], 0));
                side = cs.side[j];
                // the critical chain first: spawn feeds the next generation's trace
                if (l < p.max_depth) { bvh_shade_kernel<false, true><<<grid, kBlock, 0, ms>>>(P, l); n++; }
            }
            if (p.profile) RT_CUDA(cudaEventRecord(s->prof_mid[l][0], stream));
            const bool merged = overlap && s->nl && l >= merge_env && (long long)P.level[0].capacity <= merge_max_work;
            if (merged) {
                // this generation's shadow chains wait for the merged launch after the last trace
                if (l == p.max_depth) {
                    long long mwork = 0;
                    for (int m = merge_env; m <= p.max_depth; m++) mwork += P.level[m].capacity;
                    const int mgrid = walk_grid(mwork * std::max(1, s->nl), sms, shadow_mult ? shadow_mult : grid_mult);
                    // (work sharing by the size of the chain's frame, like the per-generation launches: measured, 1/8 shard 6.09 -> 5.96 ms)
                    const bool msharing = steal_shadow == 1 || (steal_shadow == 2 && (long long)P.level[0].capacity <= steal_shadow_max_work);
                    if (msharing) bvh_shadow_kernel<false, false, true, true, false><<<mgrid, kWalkBlock, 0, side>>>(P, -merge_env - 1);
                    else bvh_shadow_kernel<false, false, false, true, false><<<mgrid, kWalkBlock, 0, side>>>(P, -merge_env - 1);
                    n++;
                    // (measured: whole 4K frame 37.5 -> 37.9 ms, 1/8 shard 5.92 -> 5.86 ms: small frames only)
                    if (colour_merge && (long long)P.level[0].capacity <= steal_shadow_max_work) {
                        bvh_shade_kernel<true, false, true><<<persistent_grid(mwork, sms, grid_mult), kBlock, 0, side>>>(P, -merge_env - 1);
                        n++;
                    } else for (int m = merge_env; m <= p.max_depth; m++) {
                        bvh_shade_kernel<true, false><<<persistent_grid(P.level[m].capacity, sms, grid_mult), kBlock, 0, side>>>(P, m);
                        n++;
                    }
                }
                continue;
            }
            if (s->nl) {
                const bool sh_sharing = steal_shadow == 1 || (steal_shadow == 2 && work <= steal_max_work);
                if (p.instrument) {
                    if (sh_sharing) { if (hints) bvh_shadow_kernel<true, false, true, false, true><<<sh_grid, kWalkBlock, 0, side>>>(P, l); else bvh_shadow_kernel<true, false, true><<<sh_grid, kWalkBlock, 0, side>>>(P, l); }
                    else { if (hints) bvh_shadow_kernel<true, false, false, false, true><<<sh_grid, kWalkBlock, 0, side>>>(P, l); else bvh_shadow_kernel<true, false, false><<<sh_grid, kWalkBlock, 0, side>>>(P, l); }
                } else if (near_first) bvh_shadow_kernel<false, true, false><<<sh_grid, kWalkBlock, 0, side>>>(P, l);
                else if (sh_sharing) { if (hints) bvh_shadow_kernel<false, false, true, false, true><<<sh_grid, kWalkBlock, 0, side>>>(P, l); else bvh_shadow_kernel<false, false, true><<<sh_grid, kWalkBlock, 0, side>>>(P, l); }
                else { if (hints) bvh_shadow_kernel<false, false, false, false, true><<<sh_grid, kWalkBlock, 0, side>>>(P, l); else bvh_shadow_kernel<false, false, false><<<sh_grid, kWalkBlock, 0, side>>>(P, l); }
                n++;
            }
            if (p.profile) RT_CUDA(cudaEventRecord(s->prof_mid[l][1], stream));
            if (overlap) bvh_shade_kernel<true, false><<<grid, kBlock, 0, side>>>(P, l);
            else bvh_shade_kernel<true, true><<<grid, kBlock, 0, ms>>>(P, l);
            n++;
        }
    }
    const int prof_n = fused ? 1 : p.max_depth + 1;      // a fused frame is profiled as one "generation"
    if (p.profile) RT_CUDA(cudaEventRecord(s->prof_ev[prof_n], stream));
    // bottom-up combine per chain, on the chain's main stream once its colour kernels are done
    for (int j = 0; j < K; j++) {
        const FrameParams& P = PS[j];
        cudaStream_t ms = overlap ? cs.main[j] : stream;
        if (overlap) {
            RT_CUDA(cudaEventRecord(cs.side_done[j], cs.side[j]));
            RT_CUDA(cudaStreamWaitEvent(ms, cs.side_done[j], 0));
        }
        for (int l = std::max(p.max_depth - 1, 0); l >= 0; l--) {
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
// as the fullest one.
bool frame_overflowed(const SceneHost* s, Counters* total, unsigned demand[kMaxLevels])
{
    const int K = s->last_subframes;
    const RtRenderParams& p = s->last;
    Counters t;
    std::memset(&t, 0, sizeof t);
    bool overflow = false;
    double factor = 1.0;
    unsigned fullest[kMaxLevels] = { 0 };
    for (int j = 0; j < K; j++) {
        const Counters& c = s->h_counters[j];
        overflow = overflow || c.overflow != 0;
        t.overflow |= c.overflow;
        for (int l = 0; l < kMaxLevels; l++) {
            t.rays[l] += c.rays[l]; t.nodes[l] += c.nodes[l]; t.rays_reflect[l] += c.rays_reflect[l];
            fullest[l] = std::max(fullest[l], c.rays[l]);
            if (l >= 1 && l <= p.max_depth) {
                const unsigned cap = (unsigned)(s->ws.capacity[l] / K);
                if (c.rays[l] > cap) {
                    overflow = true;
                    // a truncated generation under-reports what the deeper ones need: scale those up with it
                    factor = std::max(factor, (double)c.rays[l] / (double)std::max(1u, cap));
                }
            }
        }
        t.bvh_nodes += c.bvh_nodes; t.prim_tests += c.prim_tests; t.texels += c.texels;
        t.shadow_boxes += c.shadow_boxes; t.shadow_prims += c.shadow_prims; t.shadow_walks += c.shadow_walks;
        t.shadow_hint_hits += c.shadow_hint_hits;
        t.work.stalled |= c.work.stalled;
        if (j == 0) std::memcpy(t.clock, c.clock, sizeof t.clock);      // instrumented frames have one chain
    }
    if (total) *total = t;
    if (demand)
        for (int l = 0; l < kMaxLevels; l++)
            demand[l] = (unsigned)std::min(2.0e9, (double)fullest[l] * K * factor + 64.0 * K);
    return overflow;
}

int launch_split_planes(SceneHost* s, cudaStream_t st)
{
    const int n = s->last_local_pixels;
    if (!s->ws.rgb24.reserve((size_t)n * 3) || !s->ws.flags8.reserve((size_t)n)) return RT_E_NOMEM;
    split_planes_kernel<<<(n + 255) / 256, 256, 0, st>>>(s->ws.tiles.as<uchar4>(), n, s->ws.rgb24.as<unsigned char>(),
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
