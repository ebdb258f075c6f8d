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

namespace rt {

template <bool kCount>
RT_DEV void flush_bvh_counters(const BvhAccelT<kCount>& a, Counters* c)
{
    if (!kCount) return;
    unsigned long long b = a.cnt_boxes, p = a.cnt_prims;
    for (int o = 16; o > 0; o >>= 1) {
        b += __shfl_down_sync(0xffffffffu, b, o);
        p += __shfl_down_sync(0xffffffffu, p, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&c->bvh_nodes, b); atomicAdd(&c->prim_tests, p); }
}

template <bool kCount>
__global__ void __launch_bounds__(kBlock) bvh_trace_kernel(const __grid_constant__ FrameParams P, int level)
{
    const SceneView& S = P.scene;
    const LevelBuffers& LB = P.level[level];
    const auto acc = make_bvh_accel<kCount>(S);
    const unsigned count = level == 0 ? (unsigned)P.n_local_pixels : min(P.counters->rays[level], (unsigned)LB.capacity);
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned rounded = (count + 31u) & ~31u;
    for (unsigned slot = blockIdx.x * blockDim.x + threadIdx.x; slot < rounded; slot += stride) {
        const bool in_range = slot < count;
        RayIn r;
        const bool active = in_range && load_ray(P, level, slot, r);
        Hit h; h.obj = -1; h.t = 0.f; h.b = h.g = 0.f;
        bool aborted = false;
        if (active) h = trace_ray(P, acc, level, r, aborted);
        const bool hit = active && !aborted && h.obj >= 0;
        // append to the node list: one atomic per warp
        const unsigned hm = __ballot_sync(0xffffffffu, hit);
        unsigned base = 0;
        if ((threadIdx.x & 31) == 0 && hm) base = atomicAdd(&P.counters->nodes[level], (unsigned)__popc(hm));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (hit) {
            const V3 p = vadd(r.o, vmulf(r.ray, h.t));                           // 1245 / 1303, same as finish_hit
            LB.hp[slot] = make_float4(p.x, p.y, p.z, __int_as_float(h.obj));
            LB.hb[slot] = make_float4(h.b, h.g, h.t, 0.0f);
            LB.node_list[base + __popc(hm & lanemask_lt())] = (int)slot;
        } else if (in_range) {
            write_miss(P, level, slot, !active, aborted);
        }
        if (active && level == 0 && P.primary_ids) P.primary_ids[slot] = hit ? S.obj[h.obj].z : 0;
        if (P.primary_ids) {   // instrumented frame: one texel per shaded hit on a textured object
            const unsigned tm = __ballot_sync(0xffffffffu, hit && S.obj[h.obj].y >= 0);
            if ((threadIdx.x & 31) == 0 && tm) atomicAdd(&P.counters->texels, (unsigned long long)__popc(tm));
        }
    }
    flush_bvh_counters(acc, P.counters);
}

template <bool kCount>
__global__ void __launch_bounds__(kBlock) bvh_shadow_kernel(const __grid_constant__ FrameParams P, int level)
{
    const SceneView& S = P.scene;
    const LevelBuffers& LB = P.level[level];
    const auto acc = make_bvh_accel<kCount>(S);
    const unsigned n_nodes = P.counters->nodes[level];
    const unsigned long long total = (unsigned long long)n_nodes * (unsigned)S.nl;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned li = (unsigned)(i / n_nodes), k = (unsigned)(i - (unsigned long long)li * n_nodes);
        const float4 hp = LB.hp[LB.node_list[k]];
        const V3 p = mk3(hp.x, hp.y, hp.z);
        const int self = __float_as_int(hp.w);
        const float4 l0 = S.light[3 * li];
        float4 rec;
        if (l0.w == 0.0f) {
            rec = acc.shadow_factors(p, vneg(ld3(l0)), self, P.epsilon, 0.0f, false);       // 895-912: un-normalised ray
        } else {
            const V3 L = vnorm(vsub(ld3(l0), p));                                           // 921
            const V3 dp = vsub(p, ld3(l0));
            const float dist = fsqrt(vsum3(fmul(dp.x, dp.x), fmul(dp.y, dp.y), fmul(dp.z, dp.z)));   // 922
            rec = acc.shadow_factors(p, L, self, P.epsilon, dist, true);                    // 928-949
        }
        LB.sh[i] = rec;
    }
    flush_bvh_counters(acc, P.counters);
}

// Shadow source of the shade kernel: the records of bvh_shadow_kernel.
struct RecordedShadow {
    const float4* rec;          // sh + k: record of light li at rec[li * n_nodes]
    unsigned n_nodes;
    const BvhAccel& acc;        // only for the rare "more than four partial occluders" re-trace
    RT_DEV void operator()(int li, V3 p, V3 dir, int self, float eps, float tmax, bool use_tmax, float& mask) const
    {
        const float4 f = rec[(size_t)li * n_nodes];
        if (f.x < 0.0f) { acc.shadow(p, dir, self, eps, tmax, use_tmax, mask); return; }
        mask = clamp01(fmul(f.x, mask)); mask = clamp01(fmul(f.y, mask));                   // 909 / 946, in iteration order
        mask = clamp01(fmul(f.z, mask)); mask = clamp01(fmul(f.w, mask));
    }
};

__global__ void __launch_bounds__(kBlock) bvh_shade_kernel(const __grid_constant__ FrameParams P, int level)
{
    const SceneView& S = P.scene;
    const LevelBuffers& LB = P.level[level];
    const BvhAccel acc = make_bvh_accel<false>(S);
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
        const RecordedShadow shadow{ LB.sh + k, n_nodes, acc };
        shade_and_spawn(P, shadow, level, slot, r, h, hit);
    }
}

}  // namespace rt
