// wavefront_bvh.cuh — the LBVH pipeline: one generation = three kernels, each with ONE
// kind of work per thread, so that the lanes of a warp stay in the same loop:
//
//   bvh_trace_kernel   one thread per queued ray: closest hit (tree walk).  Hits are
//                      appended to the generation's node list (one atomicAdd per warp),
//                      misses get their final record at once.
//   bvh_shadow_kernel  one thread per (node, light): the shadow ray of main.cpp:896 / 928
//                      as an any-hit walk that stops at the first opaque occluder.  Threads
//                      are light-major, so a warp holds neighbouring surface points and the
//                      SAME light: directional lights give 32 parallel rays.
//   bvh_shade_kernel   one thread per node: ShadeRay's arithmetic with the shadow products
//                      read from the records, spawn of the two child rays.
//
// The fused level_kernel (wavefront.cuh) does all three in one thread, which is right for
// the brute-force path (every lane walks the same primitive list in lock step) and wrong
// here: with nine independent tree walks per thread a warp ran 3.8 of 32 lanes on average
// (ncu, profiles/r1_level_fused_ncu.txt).
#pragma once

#include "bvh.cuh"
#include "wavefront.cuh"

// colour / spawn kernels: 8 blocks per SM (64 registers, a few spills) instead of the 5 the compiler
// picks on its own (96): they are latency-bound gathers and the extra warps pay for the spills
// (measured A/B, 15 frames each, twice: 4K frame 37.81 -> 37.44 ms; 6 blocks: 37.8)
#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 8
#endif
// closest-hit kernel: blocks per SM it is compiled for (0 = as it comes: 72 registers, 7 blocks)
#ifndef RT_TRACE_MIN_BLOCKS
#define RT_TRACE_MIN_BLOCKS 0
#endif
// fused trace kernel: blocks per SM it is compiled for (0 = the compiler's choice)
#ifndef RT_TRACE_ALL_MIN_BLOCKS
#define RT_TRACE_ALL_MIN_BLOCKS 7
#endif
// any-hit kernel: 8 blocks per SM (64 registers)
#ifndef RT_SHADOW_MIN_BLOCKS
#define RT_SHADOW_MIN_BLOCKS 8
#endif
namespace rt {

static_assert(kWalkBlock == 32 || kWalkBlock == 64 || kWalkBlock == 128, "walk kernels: 1, 2 or 4 warps per block");

template <bool kCount>
RT_DEV void flush_bvh_counters(const BvhAccelT<kCount>& a, Counters* c, bool shadow = false)
{
    if (!kCount) return;
    unsigned long long b = a.cnt_boxes, p = a.cnt_prims;
    for (int o = 16; o > 0; o >>= 1) {
        b += __shfl_down_sync(0xffffffffu, b, o);
        p += __shfl_down_sync(0xffffffffu, p, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&c->bvh_nodes, b); atomicAdd(&c->prim_tests, p);
        if (shadow) { atomicAdd(&c->shadow_boxes, b); atomicAdd(&c->shadow_prims, p); }
    }
}

// One closest-hit job of the persistent trace kernel: plain closest hit (primary, reflection,
// sphere-incidence transmission) or the face-incidence rule (main.cpp:992-1011).
template <bool kCount, bool kStealing>
struct TraceJob {
    static constexpr bool kOrdered = true;       // closest hit: near children first, the bound shrinks
    // min over (t, order) / OR of `aborted` are folds in any visiting order, so the subtrees of
    // one ray may be walked by several lanes.  Two builds of the kernel: the work-sharing one
    // ends sooner (the stragglers of the last rays are split up) but its main loop carries
    // more state; render.cu picks it for launches short enough for the tail to matter.
    static constexpr bool kSteal = kStealing;
    static constexpr bool kWideLoads = false;
    static constexpr bool kWide8 = false;
    const FrameParams& P;
    const BvhAccelT<kCount>& acc;
    int level;
    // per ray
    unsigned slot;
    V3 o, ray;
    Hit best; int best_order;
    int back;                 // >= 0: face rule with B = back (best = B's own hit, best.t = tB)
    bool skip_spheres, aborted, padding;

    RT_DEV bool begin(unsigned j, V3& wo, V3& wray, float& tmin, float& tmax)
    {
        slot = j;
        RayIn r;
        padding = !load_ray(P, level, slot, r);
        o = r.o; ray = r.ray; wo = o; wray = ray;
        tmin = level == 0 ? 0.0f : P.epsilon;                                   // 736 / 997 / 1117
        tmax = 3.402823466e+38f;
        best.t = 3.402823466e+38f; best.obj = -1; best.b = best.g = 0.0f; best_order = 0x7fffffff;
        aborted = false; back = -1;
        skip_spheres = acc.spheres_unbounded(ray);
        if (padding) return false;
        if (level > 0 && !r.reflect && r.inc < P.scene.nf && r.ns > 0) {
            // B's own hit by a direct test of its file-order record; the walk only looks for a reason to abort
            back = r.stack[r.ns - 1];
            if (back < acc.nf) {
                const float4* w = acc.tri + (size_t)back * kTriWords;
                float t, b, g;
                if (tri_distance(w[0], o, ray, t) && t > tmin && tri_inside(w[1], w[2], w[3], w[4], o, ray, t, b, g)) { best.t = t; best.obj = back; best.b = b; best.g = g; }
            } else {
                float tf, tn;
                const int n = sphere_roots(acc.sph[(size_t)(back - acc.nf) * kSphWords], o, ray, tf, tn);
                if (n >= 1 && tf > tmin && tf < best.t) { best.t = tf; best.obj = back; }
                if (n == 2 && tn > tmin && tn < best.t) { best.t = tn; best.obj = back; }
            }
        }
        return true;
    }

    RT_DEV float visit(const float4* w, float tmax)
    {
        const float tmin = level == 0 ? 0.0f : P.epsilon;
        const float4 w4 = ldp(w + 4);
        const int order = __float_as_int(w4.z);
        const bool is_tri = __float_as_int(w4.w) == 0;
        if (back >= 0) {
            // abort <=> a hit of another object with t > eps and (order < order(B) or t < tB)
            if (order == back) return tmax;
            const float limit = order < back ? 3.402823466e+38f : best.t;
            bool bad = false;
            if (is_tri) {
                float t, b, g;
                if (tri_distance(ldp(w), o, ray, t) && t > tmin && t < limit) bad = tri_inside(ldp(w + 1), ldp(w + 2), ldp(w + 3), w4, o, ray, t, b, g);
            } else {
                float tf, tn;
                const int n = sphere_roots(ldp(w), o, ray, tf, tn);
                bad = (n >= 1 && tf > tmin && tf < limit) || (n == 2 && tn > tmin && tn < limit);
            }
            if (bad) { aborted = true; return -1.0f; }
            return tmax;
        }
        if (is_tri) {
            float t;
            if (tri_distance(ldp(w), o, ray, t) && t > tmin && (t < best.t || (t == best.t && order < best_order))) {
                float b, g;
                if (tri_inside(ldp(w + 1), ldp(w + 2), ldp(w + 3), w4, o, ray, t, b, g)) { best.t = t; best.obj = order; best.b = b; best.g = g; best_order = order; }
            }
        } else if (!skip_spheres) {
            sphere_candidate(ldp(w), order, tmin);
        }
        return fminf(tmax, best.t);
    }

    RT_DEV bool next(V3&, V3&, float&, float&) { return false; }
    static constexpr bool kFirstLeaf = false;
    RT_DEV int first_leaf() { return 0; }

    // work sharing at the end of a launch (persistent_walk): a helper lane walks a subtree
    // for the ray of lane src with a copy of its state ...
    RT_DEV void adopt(int src, bool take, V3& wo, V3& wray)
    {
        constexpr unsigned kFull = 0xffffffffu;
        const float ox = __shfl_sync(kFull, o.x, src), oy = __shfl_sync(kFull, o.y, src), oz = __shfl_sync(kFull, o.z, src);
        const float rx = __shfl_sync(kFull, ray.x, src), ry = __shfl_sync(kFull, ray.y, src), rz = __shfl_sync(kFull, ray.z, src);
        const float bt = __shfl_sync(kFull, best.t, src), bb = __shfl_sync(kFull, best.b, src), bg = __shfl_sync(kFull, best.g, src);
        const int bo = __shfl_sync(kFull, best.obj, src), bord = __shfl_sync(kFull, best_order, src);
        const int bk = __shfl_sync(kFull, back, src);
        const int fl = __shfl_sync(kFull, (skip_spheres ? 1 : 0) | (padding ? 2 : 0), src);
        if (take) {
            o = mk3(ox, oy, oz); ray = mk3(rx, ry, rz);
            best.t = bt; best.b = bb; best.g = bg; best.obj = bo; best_order = bord;
            back = bk; skip_spheres = (fl & 1) != 0; padding = (fl & 2) != 0; aborted = false;
        }
        wo = o; wray = ray;
    }
    // ... and its result is folded into the owner's with the very comparison visit() uses
    RT_DEV void merge(int src, bool mine)
    {
        constexpr unsigned kFull = 0xffffffffu;
        const float bt = __shfl_sync(kFull, best.t, src), bb = __shfl_sync(kFull, best.b, src), bg = __shfl_sync(kFull, best.g, src);
        const int bo = __shfl_sync(kFull, best.obj, src), bord = __shfl_sync(kFull, best_order, src);
        const int ab = __shfl_sync(kFull, aborted ? 1 : 0, src);
        if (!mine) return;
        if (back >= 0) { aborted = aborted || ab != 0; return; }
        if (bo >= 0 && (bt < best.t || (bt == best.t && bord < best_order))) {
            best.t = bt; best.b = bb; best.g = bg; best.obj = bo; best_order = bord;
        }
    }

    // lanes walking subtrees of one ray: the nearest hit any of them has found bounds them all
    // (positive floats order like their bit patterns; FLT_MAX = nothing yet); a face-rule walk
    // has no bound, it ends when any lane has found the reason to abort
    RT_DEV float share(unsigned group, float tmax)
    {
        if (__reduce_or_sync(group, aborted ? 1u : 0u)) { aborted = true; return -1.0f; }
        return __uint_as_float(__reduce_min_sync(group, __float_as_uint(fmaxf(tmax, 0.0f))));
    }

    RT_DEV void sphere_candidate(const float4& w0, int order, float tmin)
    {
        float tf, tn;
        const int n = sphere_roots(w0, o, ray, tf, tn);
        if (n >= 1 && tf > tmin && (tf < best.t || (tf == best.t && order < best_order))) { best.t = tf; best.obj = order; best_order = order; }
        if (n == 2 && tn > tmin && (tn < best.t || (tn == best.t && order < best_order))) { best.t = tn; best.obj = order; best_order = order; }
    }

    RT_DEV void retire(bool fin)
    {
        const SceneView& S = P.scene;
        const LevelBuffers& LB = P.level[level];
        if (fin && !padding && back < 0 && skip_spheres) {
            const float tmin = level == 0 ? 0.0f : P.epsilon;
            for (int i = 0; i < acc.ns; i++) sphere_candidate(acc.sph[(size_t)i * kSphWords], acc.nf + i, tmin);
        }
        const bool hit = fin && !padding && !aborted && best.obj >= 0;
        // append to the node list: one atomic per warp
        const unsigned hm = __ballot_sync(0xffffffffu, hit);
        unsigned base = 0;
        if ((threadIdx.x & 31) == 0 && hm) base = atomicAdd(&P.counters->nodes[level], (unsigned)__popc(hm));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (hit) {
            const V3 p = vadd(o, vmulf(ray, best.t));                            // 1245 / 1303, same as finish_hit
            LB.hp[slot] = make_float4(p.x, p.y, p.z, __int_as_float(best.obj));
            LB.hb[slot] = make_float4(best.b, best.g, best.t, 0.0f);
            LB.node_list[base + __popc(hm & lanemask_lt())] = (int)slot;
            // nodes of the last generation are finished by the colour kernel itself
            if (level < P.max_depth) LB.res[slot] = make_float4(0.f, 0.f, 0.f, kResPending);
        } else if (fin) {
            write_miss(P, level, slot, padding, aborted);
        }
        if (P.primary_ids) {   // instrumented frame
            if (fin && !padding && level == 0) P.primary_ids[out_slot(P, slot)] = hit ? S.obj[best.obj].z : 0;
            const unsigned tm = __ballot_sync(0xffffffffu, hit && S.obj[best.obj].y >= 0);
            if ((threadIdx.x & 31) == 0 && tm) atomicAdd(&P.counters->texels, (unsigned long long)__popc(tm));
        }
    }
};

template <bool kCount, bool kStealing>
__global__ void __launch_bounds__(kWalkBlock, (RT_TRACE_MIN_BLOCKS && !kCount ? RT_TRACE_MIN_BLOCKS : (kStealing || kWalkScale > 1 ? 7 : 0)) * kWalkScale) bvh_trace_kernel(const __grid_constant__ FrameParams P, int level)
{
    const auto acc = make_bvh_accel<kCount>(P.scene);
    const unsigned count = level == 0 ? (unsigned)P.n_local_pixels
                                      : min(P.counters->rays[level], (unsigned)P.level[level].capacity);
    TraceJob<kCount, kStealing> job{ P, acc, level };
    acc.persistent_walk(count, &P.counters->fetch_trace[level], job, P.refill_lanes, P.inner_min_lanes, P.leaf_min_lanes,
                        kCount ? &P.counters->clock[0][level] : nullptr, P.steal != 0);
    flush_bvh_counters(acc, P.counters);
}

// A dummy shadow source for the spawn half (it never asks).
struct NoShadow {
    RT_DEV void operator()(int, V3, V3, int, float, float, bool, float&) const {}
};

// ---- the FUSED trace kernel -------------------------------------------------------------------
// Every generation's closest-hit rays in one persistent launch (BvhAccelT::persistent_walk_all).
// Walker warps walk rays of any generation; a ray that hits something below the last generation
// is appended to the hit list, where the kernel's SPAWNER warps (every spawn_every-th) take hits
// 32 at a time, run the spawn half of ShadeRay at full warp width and publish the children to
// the work list the walkers draw from.  (Spawning inside the walkers' retire step runs the
// double-precision Fresnel code with one or two lanes active: measured 38.6 ms against 21 + 2.2.)
RT_DEV void st_volatile(unsigned* p, unsigned v)
{
    asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
RT_DEV unsigned ld_volatile_u32(const unsigned* p)
{
    unsigned v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// `n` rays of this warp are through (children, if any, already published): the frame is traced
// when every published ray has retired.  Lane 0 only.
RT_DEV void count_retired(WorkQueue* q, unsigned n0, unsigned n)
{
    const unsigned r = atomicAdd(&q->retired, n) + n;
    if (r == n0 + ld_volatile_u32(&q->tail)) atomicExch(&q->done, 1u);
}

template <bool kCount>
struct TraceAllJob : TraceJob<kCount, false> {
    using Base = TraceJob<kCount, false>;
    WorkQueue* q;
    unsigned n0;

    RT_DEV bool begin(int gen, unsigned slot_, V3& wo, V3& wray, float& tmin, float& tmax)
    {
        this->level = gen;
        return Base::begin(slot_, wo, wray, tmin, tmax);
    }

    RT_DEV void retire(bool fin)
    {
        constexpr unsigned kFull = 0xffffffffu;
        const FrameParams& P = this->P;
        const SceneView& S = P.scene;
        const int level = this->level;
        const unsigned slot = this->slot;
        const unsigned lt = lanemask_lt();
        const bool lane0 = (threadIdx.x & 31) == 0;
        if (fin && !this->padding && this->back < 0 && this->skip_spheres) {
            const float tmin = level == 0 ? 0.0f : P.epsilon;
            for (int i = 0; i < this->acc.ns; i++) this->sphere_candidate(this->acc.sph[(size_t)i * kSphWords], this->acc.nf + i, tmin);
        }
        const bool hit = fin && !this->padding && !this->aborted && this->best.obj >= 0;
        // the node lists: one atomic per warp and generation present
        for (int L = 0; L <= P.max_depth; L++) {
            const unsigned hm = __ballot_sync(kFull, hit && level == L);
            if (!hm) continue;
            unsigned base = 0;
            if (lane0) base = atomicAdd(&P.counters->nodes[L], (unsigned)__popc(hm));
            base = __shfl_sync(kFull, base, 0);
            if (hit && level == L) {
                const LevelBuffers& LB = P.level[L];
                const V3 p = vadd(this->o, vmulf(this->ray, this->best.t));      // 1245 / 1303, same as finish_hit
                LB.hp[slot] = make_float4(p.x, p.y, p.z, __int_as_float(this->best.obj));
                LB.hb[slot] = make_float4(this->best.b, this->best.g, this->best.t, 0.0f);
                LB.node_list[base + __popc(hm & lt)] = (int)slot;
                if (L < P.max_depth) LB.res[slot] = make_float4(0.f, 0.f, 0.f, kResPending);
            }
        }
        if (fin && !hit) write_miss(P, level, slot, this->padding, this->aborted);
        if (P.primary_ids) {   // instrumented frame
            if (fin && !this->padding && level == 0) P.primary_ids[out_slot(P, slot)] = hit ? S.obj[this->best.obj].z : 0;
            const unsigned tm = __ballot_sync(kFull, hit && S.obj[this->best.obj].y >= 0);
            if (lane0 && tm) atomicAdd(&P.counters->texels, (unsigned long long)__popc(tm));
        }
        // hits that may have children go to the spawners; everything else is through
        const bool parent = hit && level < P.max_depth;
        const unsigned pm = __ballot_sync(kFull, parent);
        if (pm) {
            unsigned base = 0;
            if (lane0) base = atomicAdd(&q->hit_tail, (unsigned)__popc(pm));
            base = __shfl_sync(kFull, base, 0);
            __threadfence();                   // the hit records before the entries that announce them
            if (parent) st_volatile(P.hit_order + base + __popc(pm & lt), ((unsigned)level << 28) | slot);
        }
        const unsigned fm = __ballot_sync(kFull, fin && !parent);
        if (lane0 && fm) count_retired(q, n0, (unsigned)__popc(fm));
    }
};

// A spawner warp of the fused kernel.
RT_DEV void spawner_loop(const FrameParams& P, WorkQueue* q, unsigned n0)
{
    constexpr unsigned kFull = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    unsigned base = 0, pending = 0, polls = 0, spins = 0;
    for (;;) {
        if (!pending) {
            if (lane == 0) base = atomicAdd(&q->hit_head, 32u);
            base = __shfl_sync(kFull, base, 0);
            pending = kFull; polls = 0;
        }
        const unsigned e = base + lane;
        unsigned v = 0xffffffffu;
        if (((pending >> lane) & 1u) && e < P.hit_capacity) v = ld_volatile_u32(P.hit_order + e);
        const unsigned present = __ballot_sync(kFull, v != 0xffffffffu);
        if (present != 0u && (present == pending || ++polls > 4u)) {
            // the spawn half of ShadeRay for the hits that have arrived, then the children's entries
            const bool mine = ((present >> lane) & 1u) != 0u;
            const int level = mine ? (int)(v >> 28) : 0;
            const unsigned slot = v & 0x0fffffffu;
            RayIn r; r.ns = 0;
            Hit h; h.obj = -1; h.t = 0.f; h.b = h.g = 0.f;
            if (mine) {
                load_ray(P, level, slot, r);
                const float4 hp = __ldcg(P.level[level].hp + slot), hb = __ldcg(P.level[level].hb + slot);
                h.obj = __float_as_int(hp.w); h.b = hb.x; h.g = hb.y; h.t = hb.z;
            }
            int children[2] = { -1, -1 };
            const NoShadow none;
            shade_and_spawn<false, true, NoShadow, true>(P, none, level, slot, r, h, mine, children);
            const unsigned mt = __ballot_sync(kFull, children[0] >= 0), mr = __ballot_sync(kFull, children[1] >= 0);
            if (mt | mr) {
                unsigned wb = 0;
                if (lane == 0) wb = atomicAdd(&q->tail, (unsigned)(__popc(mt) + __popc(mr)));
                wb = __shfl_sync(kFull, wb, 0);
                __threadfence();               // the rays' records before the entries that announce them
                const unsigned tag = (unsigned)(level + 1) << 28;
                if (children[0] >= 0) st_volatile(P.work_order + wb + __popc(mt & lt), tag | (unsigned)children[0]);
                if (children[1] >= 0) st_volatile(P.work_order + wb + __popc(mt) + __popc(mr & lt), tag | (unsigned)children[1]);
            }
            __threadfence();                   // tail before retired
            __syncwarp();
            if (lane == 0) count_retired(q, n0, (unsigned)__popc(present));
            pending &= ~present;
            polls = 0; spins = 0;
            continue;
        }
        unsigned d = 0;
        if (lane == 0) d = ld_volatile_u32(&q->done);
        if (__shfl_sync(kFull, d, 0)) break;
        __nanosleep(300);
        if (++spins > (1u << 22)) {              // ~2 s without a hit: give up loudly
            if (lane == 0) atomicExch(&q->stalled, 1u);
            break;
        }
    }
}

template <bool kCount>
__global__ void __launch_bounds__(kWalkBlock, (kCount ? 0 : RT_TRACE_ALL_MIN_BLOCKS) * kWalkScale) bvh_trace_all_kernel(const __grid_constant__ FrameParams P)
{
    const unsigned n0 = (unsigned)P.n_local_pixels;
    const unsigned warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const auto acc = make_bvh_accel<kCount>(P.scene);
    if (warp % (unsigned)P.spawn_every == (unsigned)P.spawn_every - 1u) {
        spawner_loop(P, &P.counters->work, n0);
    } else {
        TraceAllJob<kCount> job{ { P, acc, 0 }, &P.counters->work, n0 };
        acc.persistent_walk_all(n0, &P.counters->work, P.work_order, P.work_capacity, job,
                                P.refill_lanes, P.inner_min_lanes, P.leaf_min_lanes);
    }
    flush_bvh_counters(acc, P.counters);
}

// The shadow rays (main.cpp:896 / 928) of ONE node as a chain: job k = node_list[k], the lane
// walks light 0, 1, ... in turn.  The reference never resets shadow_mask between lights
// (main.cpp:788) and every factor lies in [0,1], so once the running product
// is exactly 0 it stays 0 whatever the later lights meet: the chain stops there and the
// remaining records are never read (RecordedShadow stops at the same point).  In a dense
// scene that removes most shadow walks without changing a bit of the result.
// The launch plan of the any-hit kernel, the same for every thread: kept in shared memory so that
// it costs the walk loop no registers (the kernel sits at its 64-register limit).
struct ShadowPlan {
    int first_level;             // kAllLevels: the launch covers generations first_level .. max_depth
    unsigned total;              // jobs before padding: this generation's nodes, or those generations'
    unsigned pass_jobs, passes, group;
};
RT_DEV ShadowPlan& shadow_plan()
{
    __shared__ ShadowPlan plan;
    return plan;
}

constexpr int kFoundNone = -1;
constexpr int kFoundByHint = 0x7fffffff;

template <bool kCount, bool kNearFirst, bool kStealing, bool kAllLevels, bool kHints>
struct ShadowJob {
    // One light's walk is a fold in any order too: "an opaque occluder exists" is an OR and the
    // partial factors are collected and sorted by iteration order before they are multiplied
    // (shadow_end), so helpers only have to hand over their finds.  The chain over the lights
    // stays with the owner lane.
    static constexpr bool kSteal = kStealing;
    static constexpr bool kWideLoads = true;     // 256-bit node loads (bvh.cuh: test_node)
    static constexpr bool kWide8 = true;         // walks the 8-wide tree when the scene has one (bvh.cuh: step8)
    static constexpr bool kOrdered = kNearFirst; // any hit: order is only a heuristic (near boxes first finds occluders sooner?)
    static constexpr bool kFirstLeaf = kHints;   // first_leaf() may hand out a record to test before the walk
    const FrameParams& P;
    const BvhAccelT<kCount>& acc;
    // kAllLevels: ONE launch walks the nodes of every generation (after the fused trace kernel): job
    // j counts through generation 0's node list, then generation 1's, ...; `level` and `n_nodes`
    // are then per lane
    int level;
    unsigned n_nodes;
    // Occluder hints.  "Some opaque primitive blocks this shadow ray" is all the reference's product
    // needs to become exactly 0 (every factor is in [0,1]; main.cpp:909 / 946), and WHICH primitive
    // proves it does not matter.  Neighbouring surface points of one primitive are mostly blocked by
    // the same occluder, so the walk that finds one leaves its record index in
    // hint[light][surface primitive] and later jobs from that primitive test that record first
    // (one 80-byte exact test instead of a tree walk).  A stale or missing hint only costs the
    // test: the result never depends on it.  For hints to be THERE in time the node list is taken
    // in interleaved passes (perm): pass p handles the groups g with g % passes == p, so the
    // neighbours of a node were finished a whole pass earlier, not in the same warp.
    // per node
    unsigned k;
    int li;
    V3 p; int self;
    float mask;
    // >= 0 and < kFoundByHint: record index of the opaque occluder this walk ended at; kFoundByHint:
    // that was the hinted record; <= -2: -2 - (record index first_leaf() handed out), nothing found yet
    int found;
    unsigned long long walks = 0;      // instrumented variant: shadow rays actually walked by this thread
    unsigned long long hint_hits = 0;  // ... and settled by a hint
    typename BvhAccelT<kCount>::ShadowVisit v;

    // the table persistent_walk visits (read where needed: no register across the walk)
    RT_DEV const float4* records() const { return P.scene.bvh.n_nodes8 > 0 ? P.scene.bvh.prims8 : P.scene.bvh_tri; }

    // jobs of the launch and the job -> node permutation described above (one thread per block,
    // before the walk; the caller synchronises)
    RT_DEV static void make_plan(const FrameParams& P, int level)
    {
        ShadowPlan& pl = shadow_plan();
        unsigned total = 0;
        pl.first_level = kAllLevels ? -level - 1 : level;      // level = -(first generation) - 1
        if (kAllLevels) { for (int l = pl.first_level; l <= P.max_depth; l++) total += P.counters->nodes[l]; }
        else total = P.counters->nodes[level];
        pl.total = total; pl.passes = 1; pl.group = 1; pl.pass_jobs = total;
        if (kHints && P.hint_passes > 1) {
            pl.group = (unsigned)max(P.hint_group, 1);
            const unsigned groups = (total + pl.group - 1) / pl.group;
            pl.passes = min((unsigned)P.hint_passes, max(total / (unsigned)max(P.hint_min_chunk, 1), 1u));
            pl.pass_jobs = (groups + pl.passes - 1) / pl.passes * pl.group;
        }
    }
    RT_DEV static unsigned perm(unsigned j)
    {
        if (!kHints) return j;
        const ShadowPlan& pl = shadow_plan();
        if (pl.passes == 1) return j;
        const unsigned ps = j / pl.pass_jobs, r = j - ps * pl.pass_jobs;
        const unsigned g = r / pl.group, w = r - g * pl.group;
        return (g * pl.passes + ps) * pl.group + w;
    }

    RT_DEV void setup(V3& wo, V3& wray, float& tmin, float& tmax)
    {
        const float4 l0 = P.scene.light[3 * li];
        tmin = P.epsilon;
        if (l0.w == 0.0f) {
            wray = vneg(ld3(l0));                                                // 895: un-normalised ray
            tmax = 3.402823466e+38f;
            acc.shadow_begin(p, wray, self, tmin, 0.0f, false, v);
        } else {
            wray = vnorm(vsub(ld3(l0), p));                                      // 921
            const V3 dp = vsub(p, ld3(l0));
            tmax = fsqrt(vsum3(fmul(dp.x, dp.x), fmul(dp.y, dp.y), fmul(dp.z, dp.z)));   // 922
            acc.shadow_begin(p, wray, self, tmin, tmax, true, v);
        }
        wo = p;
        found = kFoundNone;
    }
    // The hinted record as a leaf code: persistent_walk tests it (through visit(), like any leaf)
    // before it opens the root.
    RT_DEV int first_leaf()
    {
        if (!kHints || v.count_all || !(P.hint_mode & 1)) return 0;
        const int h = __ldcg(P.occluder_hint + (size_t)li * P.hint_prims + self);
        if (h < 0) return 0;
        found = -2 - h;
        return ~(h << 3);
    }
    RT_DEV bool begin(unsigned j, V3& wo, V3& wray, float& tmin, float& tmax)
    {
        k = perm(j); li = 0; mask = 1.0f;                                        // 788
        if (kHints && k >= shadow_plan().total) { k = 0xffffffffu; return false; }   // padding of the last groups
        if (kAllLevels) {
            level = shadow_plan().first_level;
            n_nodes = P.counters->nodes[level];
            while (k >= n_nodes) { k -= n_nodes; level++; n_nodes = P.counters->nodes[level]; }
        }
        const float4 hp = P.level[level].hp[P.level[level].node_list[k]];
        p = mk3(hp.x, hp.y, hp.z);
        self = __float_as_int(hp.w);
        setup(wo, wray, tmin, tmax);
        return true;
    }
    RT_DEV float visit(const float4* w, float tmax)
    {
        const float r = v(w, tmax);
        if (kHints && r < 0.0f) {          // (count_all walks never stop; found stays unused there)
            const int idx = (int)((w - records()) / kTriWords);
            found = (found == -2 - idx) ? kFoundByHint : idx;
        }
        return r;
    }

    RT_DEV void adopt(int src, bool take, V3& wo, V3& wray)
    {
        constexpr unsigned kFull = 0xffffffffu;
        const float ox = __shfl_sync(kFull, v.o.x, src), oy = __shfl_sync(kFull, v.o.y, src), oz = __shfl_sync(kFull, v.o.z, src);
        const float rx = __shfl_sync(kFull, v.ray.x, src), ry = __shfl_sync(kFull, v.ray.y, src), rz = __shfl_sync(kFull, v.ray.z, src);
        const float t0 = __shfl_sync(kFull, v.tmin, src), t1 = __shfl_sync(kFull, v.tmax_hit, src);
        const int sf = __shfl_sync(kFull, v.self, src);
        const int fl = __shfl_sync(kFull, (v.use_tmax ? 1 : 0) | (v.skip_spheres ? 2 : 0) | (v.count_all ? 4 : 0), src);
        if (take) {
            v.o = mk3(ox, oy, oz); v.ray = mk3(rx, ry, rz); v.tmin = t0; v.tmax_hit = t1; v.self = sf;
            v.use_tmax = (fl & 1) != 0; v.skip_spheres = (fl & 2) != 0; v.count_all = (fl & 4) != 0;
            v.hits = 0; v.n_partial = 0; v.overflow_mask = 1.0f; v.opaque = false;
            found = kFoundNone;
        }
        wo = v.o; wray = v.ray;
    }
    RT_DEV void merge(int src, bool mine)
    {
        constexpr unsigned kFull = 0xffffffffu;
        const int op = __shfl_sync(kFull, v.opaque ? 1 : 0, src), h = __shfl_sync(kFull, v.hits, src);
        const int np = __shfl_sync(kFull, v.n_partial, src);
        const float ov = __shfl_sync(kFull, v.overflow_mask, src);
        const int fd = __shfl_sync(kFull, found, src);
        if (mine) {
            v.hits += h;
            v.opaque = v.opaque || op != 0;
            if (kHints && fd >= 0 && fd != kFoundByHint && found < 0) found = fd;
            if (ov != 1.0f) v.overflow_mask = clamp01(fmul(ov, v.overflow_mask));
        }
        for (int i = 0; i < np; i++) {
            // (constant indices after unrolling: the lists stay in registers)
            int po = 0; float pk = 0.0f;
            #pragma unroll
            for (int q = 0; q < BvhAccelT<kCount>::kMaxPartial; q++) if (q == i) { po = v.p_order[q]; pk = v.p_k[q]; }
            po = __shfl_sync(kFull, po, src); pk = __shfl_sync(kFull, pk, src);
            if (mine) { v.hits--; v.add(po, pk); }
        }
    }
    // an opaque occluder found by any lane ends the walk of this light for all of them
    RT_DEV float share(unsigned group, float tmax)
    {
        if (__reduce_or_sync(group, v.opaque ? 1u : 0u) && !v.count_all) return -1.0f;
        return tmax;
    }

    RT_DEV void retire(bool fin)
    {
        if (!fin || (kHints && k == 0xffffffffu)) return;
        acc.shadow_end(v);
        const bool by_hint = kHints && v.opaque && found == kFoundByHint;
        if (kCount) { if (by_hint) hint_hits++; else walks++; }
        // the carried product exactly as BvhAccelT::shadow applies it (main.cpp:909 / 946 in
        // iteration order); the record is the mask AFTER this light
        if (v.opaque) {
            mask = clamp01(fmul(0.0f, mask));
            if (kHints && found >= 0 && !by_hint && (P.hint_mode & 2))
                __stcg(P.occluder_hint + (size_t)li * P.hint_prims + self, found);
        } else {
            for (int i = 0; i < v.n_partial; i++) mask = clamp01(fmul(v.p_k[i], mask));
            if (v.overflow_mask != 1.0f) mask = clamp01(fmul(v.overflow_mask, mask));
        }
        P.level[level].sh[(size_t)li * n_nodes + k] = mask;
    }
    RT_DEV bool next(V3& wo, V3& wray, float& tmin, float& tmax)
    {
        if ((kHints && k == 0xffffffffu) || mask == 0.0f || li + 1 >= P.scene.nl) return false;
        li++;
        setup(wo, wray, tmin, tmax);
        return true;
    }
};

// level < 0 (kAllLevels): the nodes of generations -level - 1 .. max_depth in one launch
template <bool kCount, bool kNearFirst, bool kStealing, bool kAllLevels = false, bool kHints = false>
__global__ void __launch_bounds__(kWalkBlock, (kCount ? 0 : kHints ? RT_SHADOW_MIN_BLOCKS : 0) * kWalkScale) bvh_shadow_kernel(const __grid_constant__ FrameParams P, int level)
{
    using Job = ShadowJob<kCount, kNearFirst, kStealing, kAllLevels, kHints>;
    const auto acc = make_bvh_accel<kCount>(P.scene);
    if (threadIdx.x == 0) Job::make_plan(P, level);
    __syncthreads();
    const unsigned n_jobs = shadow_plan().pass_jobs * shadow_plan().passes;
    Job job{ P, acc, kAllLevels ? 0 : level, kAllLevels ? 0u : shadow_plan().total };
    const int cl = kAllLevels ? -level - 1 : level;
    acc.persistent_walk(n_jobs, &P.counters->fetch_shadow[cl], job, P.sh_refill_lanes, P.sh_inner_min_lanes, P.sh_leaf_min_lanes,
                        kCount ? &P.counters->clock[1][cl] : nullptr, P.steal != 0,
                        P.shadow_job_budget > 0 ? (unsigned)P.shadow_job_budget : 0xffffffffu);
    flush_bvh_counters(acc, P.counters, true);
    if (kCount) {
        unsigned long long w = job.walks, hh = job.hint_hits;
        for (int o = 16; o > 0; o >>= 1) { w += __shfl_down_sync(0xffffffffu, w, o); hh += __shfl_down_sync(0xffffffffu, hh, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(&P.counters->shadow_walks, w); atomicAdd(&P.counters->shadow_hint_hits, hh); }
    }
}

// Shadow source of the shade kernel: the carried shadow_mask after each light, as left by
// the node's chain in bvh_shadow_kernel.
struct RecordedShadow {
    const float* rec;           // sh + k: mask after light li at rec[li * n_nodes]
    unsigned n_nodes;
    RT_DEV void operator()(int li, V3, V3, int, float, float, bool, float& mask) const
    {
        if (mask == 0.0f) return;     // the chain stopped here: later records were never written (and 0 stays 0)
        mask = rec[(size_t)li * n_nodes];
    }
};

// kColour / kSpawn select the half of ShadeRay this launch does (wavefront.cuh:
// shade_and_spawn).  The spawn half depends only on the hits, so the next generation's trace
// kernel can start before this generation's shadow rays have been walked: the frame scheduler
// (render.cu) runs the chain trace -> spawn -> trace -> ... first and all generations' shadow
// chains in one any-hit launch after it, then the colour halves.
template <bool kColour, bool kSpawn>
RT_DEV void shade_level(const FrameParams& P, int level)
{
    const LevelBuffers& LB = P.level[level];
    const unsigned n_nodes = P.counters->nodes[level];
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned rounded = (n_nodes + 31u) & ~31u;
    for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < rounded; k += stride) {
        const bool hit = k < n_nodes;
        RayIn r; r.ns = 0;
        Hit h; h.obj = -1; h.t = 0.f; h.b = h.g = 0.f;
        unsigned slot = 0;
        if (hit) {
            slot = (unsigned)LB.node_list[k];
            load_ray(P, level, slot, r);
            const float4 hp = LB.hp[slot], hb = LB.hb[slot];
            h.obj = __float_as_int(hp.w); h.b = hb.x; h.g = hb.y; h.t = hb.z;
        }
        if (kColour) {
            const RecordedShadow shadow{ LB.sh + k, n_nodes };
            shade_and_spawn<kColour, kSpawn>(P, shadow, level, slot, r, h, hit);
        } else {
            const NoShadow none;
            shade_and_spawn<kColour, kSpawn>(P, none, level, slot, r, h, hit);
        }
    }
}

// kRange: generations -level - 1 .. max_depth one after the other in this launch (the colour halves
// after the merged any-hit launch: nothing orders them among themselves)
template <bool kColour, bool kSpawn, bool kRange = false>
__global__ void __launch_bounds__(kBlock, RT_SHADE_MIN_BLOCKS) bvh_shade_kernel(const __grid_constant__ FrameParams P, int level)
{
    if constexpr (kRange) {
        for (int l = -level - 1; l <= P.max_depth; l++) shade_level<kColour, kSpawn>(P, l);
    } else {
        shade_level<kColour, kSpawn>(P, level);
    }
}

}  // namespace rt
