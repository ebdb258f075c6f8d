// wavefront_bvh.cuh — the LBVH pipeline: every kernel has ONE kind of work per thread, so that
// the lanes of a warp stay in the same loop:
//
//   bvh_trace_kernel   one thread per queued ray: closest hit (persistent tree walk); the hit
//                      (or "none") is left in the slot's record.
//   bvh_spawn_kernel   one thread per slot, in slot order: lists the slots that hit something
//                      (the ShadeRay nodes of the generation) and runs the half of ShadeRay that
//                      depends only on the hit — medium state machine, Fresnel weights, the two
//                      child rays appended to the next generation's queue.
//   bvh_shadow_probe_kernel / bvh_shadow_settle_kernel   one thread per node, before the walks: light 0's
//                      occluder map guesses an opaque occluder, the reference's exact test confirms it, and the
//                      node's shadow chain is over without a walk; the rest are listed for the walk kernel.
//   bvh_shadow_kernel  one job per listed node: the shadow rays (main.cpp:896 / 928) to light 0, 1, ...
//                      as a chain of any-hit walks that stops as soon as the carried product is 0.
//   bvh_shadow_redo_kernel  the few chains the any-hit kernel had to give up (more partial
//                      occluders on one ray than its registers hold), by the exact slow path.
//   bvh_colour_kernel  one thread per node: texture, Blinn-Phong sum over the lights with the
//                      recorded shadow products.
//
// The fused level_kernel (wavefront.cuh) does all of this in one thread, which is right for
// the brute-force path (every lane walks the same primitive list in lock step) and wrong
// here: with nine independent tree walks per thread a warp ran 3.8 of 32 lanes on average
// (ncu, profiles/r1_level_fused_ncu.txt).
//
// Order.  Queues and node lists keep SLOT order (up to the order in which warps claim their 32
// slots): generation 0's slots are 8 x 4 pixel patches (pixel_of_slot), and the children of 32
// consecutive slots are consecutive in the next queue, so the 32 rays a warp of the walk kernels
// fetches together started close together — in every generation.  (Round 1 appended hits in the
// order the walks happened to finish: 18.7 of 32 lanes per instruction.)
#pragma once

#include "bvh.cuh"
#include "wavefront.cuh"

// colour / spawn kernels: 8 blocks per SM (64 registers, a few spills) instead of the 5 the compiler
// picks on its own (96): they are latency-bound gathers and the extra warps pay for the spills
// (measured A/B, 15 frames each, twice: 4K frame 37.81 -> 37.44 ms; 6 blocks: 37.8)
#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 8
#endif
// closest-hit kernel: blocks per SM it is compiled for (0 = the compiler's choice)
#ifndef RT_TRACE_MIN_BLOCKS
#define RT_TRACE_MIN_BLOCKS 7
#endif
// any-hit kernel: blocks per SM (0 = the compiler's choice; it lands on 64 registers = 8 blocks)
#ifndef RT_SHADOW_MIN_BLOCKS
#define RT_SHADOW_MIN_BLOCKS 0
#endif
namespace rt {

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
    // min over (t, order) / OR of `aborted` are folds in any visiting order, so the subtrees of
    // one ray may be walked by several lanes.  Two builds of the kernel: the work-sharing one
    // ends sooner (the stragglers of the last rays are split up) but its main loop carries
    // more state; render.cu picks it for launches short enough for the tail to matter.
    static constexpr bool kSteal = kStealing;
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
    RT_DEV bool walk_needed() const { return true; }

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
        if (hit) {
            const V3 p = vadd(o, vmulf(ray, best.t));                            // 1245 / 1303, same as finish_hit
            LB.hp[slot] = make_float4(p.x, p.y, p.z, __int_as_float(best.obj));
            LB.hb[slot] = make_float4(best.b, best.g, best.t, 0.0f);
            // nodes of the last generation are finished by the colour kernel itself
            if (level < P.max_depth) LB.res[slot] = make_float4(0.f, 0.f, 0.f, kResPending);
        } else if (fin) {
            LB.hp[slot] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));        // bvh_spawn_kernel lists the hits
            write_miss(P, level, slot, padding, aborted);
        }
        if (kCount && P.primary_ids) {   // instrumented frame
            if (fin && !padding && level == 0) {
                int lrow, col;
                P.primary_ids[pixel_of_slot(P, slot, lrow, col)] = hit ? S.obj[best.obj].z : 0;
            }
            const unsigned tm = __ballot_sync(0xffffffffu, hit && S.obj[best.obj].y >= 0);
            if ((threadIdx.x & 31) == 0 && tm) atomicAdd(&P.counters->texels, (unsigned long long)__popc(tm));
        }
    }
};


template <bool kCount, bool kStealing>
__global__ void __launch_bounds__(kWalkBlock, kCount ? 0 : RT_TRACE_MIN_BLOCKS) bvh_trace_kernel(const __grid_constant__ FrameParams P, int level)
{
    auto acc = make_bvh_accel<kCount>(P.scene);
    acc.overflow_flag = &P.counters->overflow;
    const unsigned count = level == 0 ? (unsigned)P.n_local_pixels
                                      : min(P.counters->rays[level], (unsigned)P.level[level].capacity);
    TraceJob<kCount, kStealing> job{ P, acc, level };
    acc.persistent_walk(count, &P.counters->fetch_trace[level], job, P.refill_lanes, P.inner_min_lanes, P.leaf_min_lanes,
                        kCount ? &P.counters->clock[0][level] : nullptr);
    flush_bvh_counters(acc, P.counters);
}

// A dummy shadow source for the spawn half (it never asks).
struct NoShadow {
    RT_DEV void operator()(int, V3, V3, int, float, float, bool, float&) const {}
};

// The ShadeRay nodes of generation `level` in slot order, and the spawn half of ShadeRay for each
// (nothing to spawn in the last generation: the list only).  One atomicAdd per warp for the list,
// one for the children.
__global__ void __launch_bounds__(kBlock, RT_SHADE_MIN_BLOCKS) bvh_spawn_kernel(const __grid_constant__ FrameParams P, int level)
{
    const LevelBuffers& LB = P.level[level];
    const unsigned count = level == 0 ? (unsigned)P.n_local_pixels : min(P.counters->rays[level], (unsigned)LB.capacity);
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned rounded = (count + 31u) & ~31u;
    const bool lane0 = (threadIdx.x & 31) == 0;
    const bool spawning = level < P.max_depth;
    for (unsigned slot = blockIdx.x * blockDim.x + threadIdx.x; slot < rounded; slot += stride) {
        RayIn r; r.ns = 0;
        Hit h; h.obj = -1; h.t = 0.f; h.b = h.g = 0.f;
        // every record of the slot is asked for at once (most slots of a dense scene hold a hit): one memory
        // round trip per slot instead of hit -> ray -> barycentrics one after the other
        if (slot < count) {
            h.obj = __float_as_int(__ldcg(&LB.hp[slot].w));
            if (spawning) {
                const float4 hb = __ldcg(LB.hb + slot);
                load_ray(P, level, slot, r);
                h.b = hb.x; h.g = hb.y; h.t = hb.z;
            }
        }
        const bool hit = h.obj >= 0;
        const unsigned hm = __ballot_sync(0xffffffffu, hit);
        if (!hm) continue;                            // warp-uniform
        unsigned base = 0;
        if (lane0) base = atomicAdd(&P.counters->nodes[level], (unsigned)__popc(hm));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (hit) LB.node_list[base + __popc(hm & lanemask_lt())] = (int)slot;
        if (!spawning) continue;
        const NoShadow none;
        shade_and_spawn<false, true>(P, none, level, slot, r, h, hit);
    }
}

// The shadow rays (main.cpp:896 / 928) of ONE node as a chain: job k = node_list[k], the lane
// walks light 0, 1, ... in turn.  The reference never resets shadow_mask between lights
// (main.cpp:788) and every factor lies in [0,1], so once the running product
// is exactly 0 it stays 0 whatever the later lights meet: the chain stops there and the
// remaining records are never read (RecordedShadow stops at the same point).  In a dense
// scene that removes most shadow walks without changing a bit of the result.
// One launch covers the nodes of generations first .. last (the frame scheduler passes all of
// them at once: one kernel tail per frame instead of one per generation): job j counts through
// generation `first`'s node list, then the next one's, ...
// The shadow ray of light li from p as ShadeRay casts it (main.cpp:895 / 921-922): direction, and the light's
// distance for a point light (hits beyond it do not count).
RT_DEV void shadow_ray_of_light(const FrameParams& P, int li, V3 p, V3& ray, float& tmax, bool& use_tmax)
{
    const float4 l0 = P.scene.light[3 * li];
    if (l0.w == 0.0f) {
        ray = vneg(ld3(l0));                                                 // 895: un-normalised ray
        tmax = 0.0f; use_tmax = false;
    } else {
        ray = vnorm(vsub(ld3(l0), p));                                       // 921
        const V3 dp = vsub(p, ld3(l0));
        tmax = fsqrt(vsum3(fmul(dp.x, dp.x), fmul(dp.y, dp.y), fmul(dp.z, dp.z)));   // 922
        use_tmax = true;
    }
}

// What the frame's sample said about light li's map (bvh_shadow_probe_kernel; no sample: use the map).
RT_DEV bool occluder_map_pays(const FrameParams& P, int li)
{
    const unsigned tries = P.counters->smap_tries[li], hits = P.counters->smap_hits[li];
    return tries == 0u || hits * 8u >= tries;
}

// The occluder map's guess for the shadow ray (p, ray) of light li (scene_device.h: ShadowMapView): the reference's
// own test of this ray against the opaque primitive that covers the ray's cell nearest to the light.  Accepted
// (-> true, v.opaque set): the product is exactly 0 whatever else the ray meets (every factor lies in [0,1]), no
// walk.  Not accepted: the attempt is forgotten (v as shadow_begin left it) and the ray has to be walked.
// v must come from shadow_begin(p, ray, self, eps, tmax, use_tmax).  kForce: whatever the frame's sample said
// (the sampling kernel itself).
template <bool kCount, bool kForce = false>
RT_DEV bool occluder_map_settles(const FrameParams& P, const BvhAccelT<kCount>& acc, int li, V3 p, V3 ray, int self,
                                 float tmax, bool use_tmax, typename BvhAccelT<kCount>::ShadowVisit& v)
{
    if (li >= kMaxShadowMaps || P.scene.smap[li].cells == nullptr || (!kForce && !occluder_map_pays(P, li))) return false;
    const ShadowMapView& m = P.scene.smap[li];
    const float R = (float)m.R;
    float fu, fv;
    size_t face = 0;
    if (m.cube) {
        const float d[3] = { p.x - m.origin[0], p.y - m.origin[1], p.z - m.origin[2] };
        int az = 0;
        if (fabsf(d[1]) > fabsf(d[az])) az = 1;
        if (fabsf(d[2]) > fabsf(d[az])) az = 2;
        const float z = fabsf(d[az]);
        if (!(z > 0.0f)) return false;
        face = (size_t)(2 * az + (d[az] < 0.0f ? 1 : 0)) * (size_t)(m.R * m.R);
        fu = (d[(az + 1) % 3] / z + 1.0f) * (0.5f * R);
        fv = (d[(az + 2) % 3] / z + 1.0f) * (0.5f * R);
    } else {
        fu = (p.x * m.ub[0] + p.y * m.ub[1] + p.z * m.ub[2] - m.u0) * m.inv_cell;
        fv = (p.x * m.vb[0] + p.y * m.vb[1] + p.z * m.vb[2] - m.v0) * m.inv_cell;
    }
    if (!(fu >= 0.0f && fv >= 0.0f && fu < R && fv < R)) return false;
    const int slot = __ldg(m.cells + face + (size_t)(int)fv * m.R + (int)fu);
    if (slot < 0) return false;
    const float4* pr = acc.nodes8 != nullptr ? acc.prims8 : acc.prims;
    if (kCount) acc.cnt_prims++;
    v(pr + (size_t)slot * kTriWords, 3.402823466e+38f);
    if (v.opaque) return true;
    acc.shadow_begin(p, ray, self, P.epsilon, tmax, use_tmax, v);
    return false;
}

// Does a map pay in THIS frame?  A lookup is one dependent gather per ray; where most rays reach their light (the lit
// side of a convex mesh seen from the light's side) it settles nothing and costs ~10 % of the any-hit time.  A few
// thousand nodes of the first generation, spread over the list, are tried against every map; the kernels that follow
// leave a map alone if it confirmed an occluder for fewer than one sampled node in eight.  (Whatever is decided,
// the frame is the same: a map only ever replaces a walk whose result it has proved.)
constexpr unsigned kProbeNodes = 8192;
__global__ void __launch_bounds__(256) bvh_shadow_probe_kernel(const __grid_constant__ FrameParams P, int first)
{
    auto acc = make_bvh_accel<false>(P.scene);
    const LevelBuffers& LB = P.level[first];
    const unsigned n = P.counters->nodes[first];
    const unsigned samples = min(n, kProbeNodes);
    const unsigned stride = samples ? n / samples : 1u;
    const int nl = min(P.scene.nl, kMaxShadowMaps);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < ((samples + 31u) & ~31u); i += gridDim.x * blockDim.x) {
        for (int li = 0; li < nl; li++) {
            if (P.scene.smap[li].cells == nullptr) continue;                     // (uniform)
            bool hit = false;
            if (i < samples) {
                const float4 hp = LB.hp[LB.node_list[i * stride]];
                const V3 p = mk3(hp.x, hp.y, hp.z);
                const int self = __float_as_int(hp.w);
                V3 ray; float tmax; bool use_tmax;
                shadow_ray_of_light(P, li, p, ray, tmax, use_tmax);
                BvhAccelT<false>::ShadowVisit v;
                acc.shadow_begin(p, ray, self, P.epsilon, tmax, use_tmax, v);
                hit = occluder_map_settles<false, true>(P, acc, li, p, ray, self, tmax, use_tmax, v);
            }
            const unsigned hm = __ballot_sync(0xffffffffu, hit), tm = __ballot_sync(0xffffffffu, i < samples);
            if ((threadIdx.x & 31) == 0) {
                atomicAdd(&P.counters->smap_tries[li], (unsigned)__popc(tm));
                if (hm) atomicAdd(&P.counters->smap_hits[li], (unsigned)__popc(hm));
            }
        }
    }
}

// Before the any-hit walks of a chain's nodes: one thread per node tries light 0's occluder map.  Settled nodes get
// their record (the carried product after light 0 is 0: the chain ends there) and never enter the persistent walk
// kernel; the others are listed (LevelBuffers::walk_list) as that kernel's jobs.  Launched only when light 0 has a map.
template <bool kCount>
__global__ void __launch_bounds__(256) bvh_shadow_settle_kernel(const __grid_constant__ FrameParams P, int first, int last)
{
    auto acc = make_bvh_accel<kCount>(P.scene);
    unsigned long long hits = 0;
    if (!occluder_map_pays(P, 0)) return;                // (the walk kernel then takes the node lists as they are)
    for (int level = first; level <= last; level++) {
        const LevelBuffers& LB = P.level[level];
        const unsigned n = P.counters->nodes[level];
        const unsigned rounded = (n + 31u) & ~31u;
        for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < rounded; k += gridDim.x * blockDim.x) {
            bool walk = false;
            if (k < n) {
                const float4 hp = LB.hp[LB.node_list[k]];
                const V3 p = mk3(hp.x, hp.y, hp.z);
                const int self = __float_as_int(hp.w);
                V3 ray; float tmax; bool use_tmax;
                shadow_ray_of_light(P, 0, p, ray, tmax, use_tmax);
                typename BvhAccelT<kCount>::ShadowVisit v;
                acc.shadow_begin(p, ray, self, P.epsilon, tmax, use_tmax, v);
                if (occluder_map_settles(P, acc, 0, p, ray, self, tmax, use_tmax, v)) {
                    LB.sh[k] = clamp01(fmul(0.0f, 1.0f));                    // 909 / 946 with shadow_mask == 1 (788)
                    if (kCount) hits++;
                } else {
                    walk = true;
                }
            }
            const unsigned wm = __ballot_sync(0xffffffffu, walk);
            if (wm) {
                unsigned base = 0;
                if ((threadIdx.x & 31) == 0) base = atomicAdd(&P.counters->walk_nodes[level], (unsigned)__popc(wm));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (walk) LB.walk_list[base + __popc(wm & lanemask_lt())] = (int)k;
            }
        }
    }
    if (kCount) {
        for (int o = 16; o > 0; o >>= 1) hits += __shfl_down_sync(0xffffffffu, hits, o);
        if ((threadIdx.x & 31) == 0 && hits) atomicAdd(&P.counters->shadow_map_hits, hits);
        flush_bvh_counters(acc, P.counters, true);
    }
}

template <bool kCount, bool kStealing>
struct ShadowJob {
    // One light's walk is a fold in any order too: "an opaque occluder exists" is an OR and the
    // partial factors are collected and sorted by iteration order before they are multiplied
    // (shadow_end), so helpers only have to hand over their finds.  The chain over the lights
    // stays with the owner lane.
    static constexpr bool kSteal = kStealing;
    static constexpr bool kWide8 = true;         // walks the 8-wide tree when the scene has one (bvh.cuh: step8)
    const FrameParams& P;
    const BvhAccelT<kCount>& acc;
    int first;
    // per node
    int level;
    unsigned n_nodes;
    unsigned k;
    int li;
    V3 p; int self;
    float mask;
    bool from_lists;                   // the jobs are the walk lists' (bvh_shadow_settle_kernel has dealt with light 0)
    bool settled;                      // this light's ray needs no walk: the occluder map's candidate is an opaque hit
    unsigned long long walks = 0;      // instrumented variant: shadow rays actually walked by this thread
    unsigned long long map_hits = 0;   // ... and those the occluder map settled
    typename BvhAccelT<kCount>::ShadowVisit v;

    RT_DEV void setup(V3& wo, V3& wray, float& tmin, float& tmax)
    {
        tmin = P.epsilon;
        float far; bool use_far;
        shadow_ray_of_light(P, li, p, wray, far, use_far);
        tmax = use_far ? far : 3.402823466e+38f;
        acc.shadow_begin(p, wray, self, tmin, far, use_far, v);
        // (light 0 was looked up by bvh_shadow_settle_kernel when the jobs come from the walk lists)
        // (with maps for light 0 only — the default — the pre-pass has asked it already and the walk kernel carries no lookup)
        settled = false;
        if (kMaxShadowMaps > 1)
            settled = !(li == 0 && from_lists) && occluder_map_settles(P, acc, li, p, wray, self, far, use_far, v);
        wo = p;
    }
    RT_DEV bool begin(unsigned j, V3& wo, V3& wray, float& tmin, float& tmax)
    {
        k = j; li = 0; mask = 1.0f;                                              // 788
        level = first;
        if (from_lists) {
            unsigned n_walk = P.counters->walk_nodes[level];
            while (k >= n_walk) { k -= n_walk; level++; n_walk = P.counters->walk_nodes[level]; }
            k = (unsigned)P.level[level].walk_list[k];
            n_nodes = P.counters->nodes[level];
        } else {
            n_nodes = P.counters->nodes[level];
            while (k >= n_nodes) { k -= n_nodes; level++; n_nodes = P.counters->nodes[level]; }
        }
        const float4 hp = P.level[level].hp[P.level[level].node_list[k]];
        p = mk3(hp.x, hp.y, hp.z);
        self = __float_as_int(hp.w);
        setup(wo, wray, tmin, tmax);
        return !settled;
    }
    RT_DEV bool walk_needed() const { return !settled; }
    RT_DEV float visit(const float4* w, float tmax) { return v(w, tmax); }

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
            v.hits = 0; v.n_partial = 0; v.overflowed = false; v.opaque = false;
        }
        wo = v.o; wray = v.ray;
    }
    RT_DEV void merge(int src, bool mine)
    {
        constexpr unsigned kFull = 0xffffffffu;
        const int op = __shfl_sync(kFull, (v.opaque ? 1 : 0) | (v.overflowed ? 2 : 0), src), h = __shfl_sync(kFull, v.hits, src);
        const int np = __shfl_sync(kFull, v.n_partial, src);
        if (mine) {
            v.hits += h;
            v.opaque = v.opaque || (op & 1) != 0;
            v.overflowed = v.overflowed || (op & 2) != 0;
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
    // an opaque occluder found by any lane ends the walk of this light for all of them (and so
    // does a list that is full: the node goes to the redo kernel)
    RT_DEV float share(unsigned group, float tmax)
    {
        if (__reduce_or_sync(group, (v.opaque || v.overflowed) ? 1u : 0u) && !v.count_all) return -1.0f;
        return tmax;
    }

    RT_DEV void retire(bool fin)
    {
        if (!fin) return;
        acc.shadow_end(v);
        if (kCount) { if (settled) map_hits++; else walks++; }
        if (v.overflowed && !v.opaque) {
            // more transparent occluders than the visitor holds: the whole chain of this node is
            // redone exactly by bvh_shadow_redo_kernel; this lane drops it
            const unsigned at = atomicAdd(&P.counters->redo_count, 1u);
            if (at < (unsigned)kRedoCapacity) P.redo_list[at] = ((unsigned)level << 28) | k;
            else atomicOr(&P.counters->overflow, kOverflowRedo);
            mask = 0.0f;                           // next() ends the chain
            return;
        }
        // the carried product exactly as BvhAccelT::shadow applies it (main.cpp:909 / 946 in
        // iteration order); the record is the mask AFTER this light
        if (v.opaque) {
            mask = clamp01(fmul(0.0f, mask));
        } else {
            for (int i = 0; i < v.n_partial; i++) mask = clamp01(fmul(v.p_k[i], mask));
        }
        P.level[level].sh[(size_t)li * n_nodes + k] = mask;
    }
    RT_DEV bool next(V3& wo, V3& wray, float& tmin, float& tmax)
    {
        if (mask == 0.0f || li + 1 >= P.scene.nl) return false;
        li++;
        setup(wo, wray, tmin, tmax);
        return true;
    }
};

template <bool kCount, bool kStealing>
__global__ void __launch_bounds__(kWalkBlock, kCount ? 0 : RT_SHADOW_MIN_BLOCKS) bvh_shadow_kernel(const __grid_constant__ FrameParams P, int first, int last)
{
    using Job = ShadowJob<kCount, kStealing>;
    auto acc = make_bvh_accel<kCount>(P.scene);
    acc.overflow_flag = &P.counters->overflow;
    const bool from_lists = P.settled_first && occluder_map_pays(P, 0);
    unsigned n_jobs = 0;
    for (int l = first; l <= last; l++) n_jobs += from_lists ? P.counters->walk_nodes[l] : P.counters->nodes[l];
    Job job{ P, acc, first };
    job.from_lists = from_lists;
    acc.persistent_walk(n_jobs, &P.counters->fetch_shadow[first], job, P.sh_refill_lanes, P.sh_inner_min_lanes, P.sh_leaf_min_lanes,
                        kCount ? &P.counters->clock[1][first] : nullptr);
    flush_bvh_counters(acc, P.counters, true);
    if (kCount) {
        unsigned long long w = job.walks, mh = job.map_hits;
        for (int o = 16; o > 0; o >>= 1) { w += __shfl_down_sync(0xffffffffu, w, o); mh += __shfl_down_sync(0xffffffffu, mh, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(&P.counters->shadow_walks, w); atomicAdd(&P.counters->shadow_map_hits, mh); }
    }
}

// The chains bvh_shadow_kernel gave up, one thread each: BvhAccelT::shadow has no limit on the
// number of partial occluders (it repeats the walk).  Almost always zero entries.
template <bool kCount>
__global__ void __launch_bounds__(kBlock) bvh_shadow_redo_kernel(const __grid_constant__ FrameParams P)
{
    const unsigned n = min(P.counters->redo_count, (unsigned)kRedoCapacity);
    auto acc = make_bvh_accel<kCount>(P.scene);
    acc.overflow_flag = &P.counters->overflow;
    unsigned long long walks = 0;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned e = P.redo_list[i];
        const int level = (int)(e >> 28);
        const unsigned k = e & 0x0fffffffu;
        const LevelBuffers& LB = P.level[level];
        const unsigned n_nodes = P.counters->nodes[level];
        const float4 hp = LB.hp[LB.node_list[k]];
        const V3 p = mk3(hp.x, hp.y, hp.z);
        const int self = __float_as_int(hp.w);
        float mask = 1.0f;                                                       // 788
        for (int li = 0; li < P.scene.nl; li++) {
            const float4 l0 = P.scene.light[3 * li];
            if (l0.w == 0.0f) {
                acc.shadow(p, vneg(ld3(l0)), self, P.epsilon, 0.0f, false, mask);
            } else {
                const V3 L = vnorm(vsub(ld3(l0), p));
                const V3 dp = vsub(p, ld3(l0));
                acc.shadow(p, L, self, P.epsilon, fsqrt(vsum3(fmul(dp.x, dp.x), fmul(dp.y, dp.y), fmul(dp.z, dp.z))), true, mask);
            }
            if (kCount) walks++;
            LB.sh[(size_t)li * n_nodes + k] = mask;
            if (mask == 0.0f) break;
        }
    }
    if (kCount) {
        for (int o = 16; o > 0; o >>= 1) walks += __shfl_down_sync(0xffffffffu, walks, o);
        if ((threadIdx.x & 31) == 0 && walks) atomicAdd(&P.counters->shadow_walks, walks);
        // (the chain's first walks were already counted by the any-hit kernel: the redo count is a
        // handful per frame at most and only the instrumented statistics see the difference)
        flush_bvh_counters(acc, P.counters, true);
    }
}

// Shadow source of the colour kernel: the carried shadow_mask after each light, as left by
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

// The colour half of ShadeRay (shade_colour) for the nodes of generations first .. last, one after
// the other in this launch (nothing orders them among themselves).
__global__ void __launch_bounds__(kBlock, RT_SHADE_MIN_BLOCKS) bvh_colour_kernel(const __grid_constant__ FrameParams P, int first, int last)
{
    for (int level = first; level <= last; level++) {
        const LevelBuffers& LB = P.level[level];
        const unsigned n_nodes = P.counters->nodes[level];
        const unsigned stride = gridDim.x * blockDim.x;
        for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < n_nodes; k += stride) {
            const unsigned slot = (unsigned)LB.node_list[k];
            RayIn r;
            load_ray(P, level, slot, r);
            const float4 hp = LB.hp[slot], hb = LB.hb[slot];
            Hit h; h.obj = __float_as_int(hp.w); h.b = hb.x; h.g = hb.y; h.t = hb.z;
            const RecordedShadow shadow{ LB.sh + k, n_nodes };
            shade_and_spawn<true, false>(P, shadow, level, slot, r, h, true);
        }
    }
}

}  // namespace rt
