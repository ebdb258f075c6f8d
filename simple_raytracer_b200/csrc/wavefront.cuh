// wavefront.cuh — the per-generation kernels of the render.
//
//   level_kernel<Accel>(d) : for every ray queued for generation d (d = 0: one
//       primary ray per pixel, generated in place, main.cpp:718-728) find the
//       closest hit, run the medium state machine, shade the hit (shade.cuh) and
//       append the transmission / reflection rays it spawns to the queue of
//       generation d+1.  Queue slots are handed out with one atomicAdd per warp
//       (ballot + popc), so the queue of the next generation is dense: lanes of
//       later generations are all busy however sparse the glass pixels are.
//   combine_kernel(d)      : bottom-up, finishes the nodes of generation d from
//       the finished results of d+1 with the reference's clamped operators
//       (main.cpp:1083, 1087, 1194, 1198, 1206); generation 0 also quantises to
//       8 bits (760-762) and writes this rank's row tiles.
//
// Grids are persistent (a multiple of the SM count) and stride over the device-side
// queue count, so a frame needs no host synchronisation between generations.
#pragma once

#include "shade.cuh"

namespace rt {

constexpr int kBlock = 128;

RT_DEV unsigned lanemask_lt()
{
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Global row of local row r under interleaved row-tile sharding.
RT_DEV int global_row(const FrameParams& P, int local_row)
{
    const int tile = local_row / P.tile_rows, in_tile = local_row - tile * P.tile_rows;
    return (tile * P.world + P.rank) * P.tile_rows + in_tile;
}

// Pixel of level-0 slot `slot`.  Slots run tile by tile; inside a tile (tile_rows rows of the
// image) they run through blocks of 8 columns, row by row inside a block, so the 32 consecutive
// slots a warp takes are an 8 x 4 patch of pixels rather than a 32 x 1 strip: primary rays of a
// warp stay closer together in the tree, and so do the generations spawned from them (queues keep
// slot order).  The last block of a row is narrower when the width is not a multiple of 8.
// (A/B against scanline order, 4K stress frame: 37.63 -> 37.03 ms; primary-ray launch 5.50 -> 5.32 ms,
// lanes per instruction of that launch 18.7 -> 20.5; profiles/r2_ab_sort_tiles.txt.)
//   lrow, col : row of this sub-frame's tile buffer / image column
//   returns the index in the RANK's tile buffer [local_tile][row][col], -1 for the virtual
//   padding tiles of a sub-frame past the end of the buffer.
RT_DEV int pixel_of_slot(const FrameParams& P, unsigned slot, int& lrow, int& col)
{
    const unsigned W = (unsigned)P.cam.width, tr = (unsigned)P.tile_rows;
    const unsigned per_tile = tr * W, blk = 8u * tr;
    const unsigned vt = slot / per_tile, rem = slot - vt * per_tile;
    const unsigned cb = rem / blk, r2 = rem - cb * blk;
    unsigned rr, c;
    if (cb * 8u + 8u <= W) { rr = r2 >> 3; c = cb * 8u + (r2 & 7u); }
    else { const unsigned wl = W - cb * 8u; rr = r2 / wl; c = cb * 8u + (r2 - rr * wl); }
    lrow = (int)(vt * tr + rr); col = (int)c;
    const unsigned t = vt * (unsigned)P.out_mul + (unsigned)P.out_add;
    return t < (unsigned)P.out_tiles ? (int)(t * per_tile + rr * W + c) : -1;
}

// Shared-memory staging of the primitive tables for small scenes: every lane of a
// warp walks the same primitive at the same time, so the reads are broadcasts.
template <bool kStage>
RT_DEV BruteAccel make_brute(const SceneView& S, float4* smem)
{
    BruteAccel a;
    a.nf = S.nf; a.ns = S.ns;
    if (kStage) {
        const int ntw = S.nf * kTriWords, nsw = S.ns * kSphWords;
        for (int i = threadIdx.x; i < ntw; i += blockDim.x) smem[i] = S.tri[i];
        for (int i = threadIdx.x; i < nsw; i += blockDim.x) smem[ntw + i] = S.sph[i];
        __syncthreads();
        a.tri = smem; a.sph = smem + ntw;
    } else {
        a.tri = S.tri; a.sph = S.sph;
    }
    return a;
}

// The ray of one queue slot, with the medium state its parent call left.
struct RayIn {
    V3 o, ray;
    float n_i, n_t;        // indices of the PARENT call
    int inc;               // incidence object of the parent call (order), -1 for primary rays
    int ns;                // medium stack length
    bool reflect, exiting;
    int stack[kMaxStack];
};

// Level 0: generates the primary ray of local pixel `slot` (main.cpp:720, 728); deeper
// levels: reads the queue entry.  Returns false for the padding rows of the last tile.
RT_DEV bool load_ray(const FrameParams& P, int level, unsigned slot, RayIn& r)
{
    r.o = mk3(0.f, 0.f, 0.f); r.ray = mk3(0.f, 0.f, 0.f);
    r.n_i = r.n_t = 0.f; r.inc = -1; r.ns = 0; r.reflect = r.exiting = false;
    if (level == 0) {
        int lrow, col;
        pixel_of_slot(P, slot, lrow, col);
        const int row = global_row(P, lrow);
        if (row >= P.cam.height) return false;
        const V3 ul = mk3(P.cam.ul[0], P.cam.ul[1], P.cam.ul[2]);
        const V3 dh = mk3(P.cam.delta_h[0], P.cam.delta_h[1], P.cam.delta_h[2]);
        const V3 dv = mk3(P.cam.delta_v[0], P.cam.delta_v[1], P.cam.delta_v[2]);
        r.o = mk3(P.cam.eye[0], P.cam.eye[1], P.cam.eye[2]);
        const V3 p = vadd(vadd(ul, vmulf(dh, (float)col)), vmulf(dv, (float)row));   // 720
        r.ray = vnorm(vsub(p, r.o));                                                // 728
        return true;
    }
    // (queue entries are read once or twice per frame: ld.global.cg keeps them out of L1, which the tree needs)
    const LevelBuffers& LB = P.level[level];
    const float4 qo = __ldcg(LB.rq_o + slot), qd = __ldcg(LB.rq_d + slot);
    const int4 qm = __ldcg(LB.rq_m + slot), qs = __ldcg(LB.rq_s + slot);
    r.o = mk3(qo.x, qo.y, qo.z); r.n_i = qo.w;
    r.ray = mk3(qd.x, qd.y, qd.z); r.n_t = qd.w;
    r.inc = qm.y;
    r.reflect = (qm.z & kRayKindReflect) != 0;
    r.exiting = (qm.z & kRayStateExiting) != 0;
    r.ns = (qm.z >> kRayNStackShift) & 15;
    r.stack[0] = qm.w; r.stack[1] = qs.x; r.stack[2] = qs.y; r.stack[3] = qs.z; r.stack[4] = qs.w;
    return true;
}

// Closest hit of the slot's ray with the rule its kind asks for.
template <class Accel>
RT_DEV Hit trace_ray(const FrameParams& P, const Accel& acc, int level, const RayIn& r, bool& aborted)
{
    aborted = false;
    if (level == 0) return acc.closest(r.o, r.ray, 0.0f);                            // 729-742: t > 0.0f
    if (!r.reflect && r.inc < P.scene.nf && r.ns > 0)
        return acc.closest_face_rule(r.o, r.ray, P.epsilon, r.stack[r.ns - 1], aborted);   // 992-1011
    return acc.closest(r.o, r.ray, P.epsilon);                                       // 1113-1124
}

// The record of a slot whose ray found nothing to shade.
RT_DEV void write_miss(const FrameParams& P, int level, unsigned slot, bool padding, bool aborted)
{
    const SceneView& S = P.scene;
    float4 r;
    if (padding) r = make_float4(0.f, 0.f, 0.f, kResPrimaryMiss);
    else if (level == 0) r = make_float4(S.bkg[0], S.bkg[1], S.bkg[2], kResPrimaryMiss);   // 721
    else if (aborted) r = make_float4(0.f, 0.f, 0.f, kResZeroTerm);                         // SKIP_TRANS
    else r = make_float4(S.bkg[0], S.bkg[1], S.bkg[2], kResFinal);   // 1087 / 1198: bkg takes the child's place
    P.level[level].res[slot] = r;
}

// ShadeRay for the hit `h` of slot `slot` (main.cpp:783-1207 up to the recursive calls).
//   kColour: the node's base colour (shade_colour) -> nr_a, or the final result for the last
//            generation (whose nodes have no children and no combine pass of their own);
//   kSpawn : medium state machine, Fresnel weights (shade_spawn), the two child rays appended
//            to the queue of generation level+1 -> nr_b, nr_c.
// The fused brute-force kernel does both in one call, the LBVH pipeline in two kernels.
// Must be called by all 32 lanes of a warp (`hit` false for idle lanes): queue slots are
// handed out with one atomicAdd per warp, transmission rays first, then reflection rays, so
// the children of 32 consecutive slots are consecutive in the next generation's queue.
template <bool kColour, bool kSpawn, class ShadowSrc>
RT_DEV void shade_and_spawn(const FrameParams& P, const ShadowSrc& shadow, int level, unsigned slot, RayIn& r,
                            const Hit& h, bool hit)
{
    const SceneView& S = P.scene;
    const LevelBuffers& LB = P.level[level];
    Surface sf;
    if (hit) sf = finish_hit(S, r.o, r.ray, h);

    if (kColour && hit) {
        const Col base = shade_colour(S, shadow, r.ray, h.obj, sf, P.epsilon);
        if (level == P.max_depth) LB.res[slot] = make_float4(base.r, base.g, base.b, kResFinal);
        else LB.nr_a[slot] = make_float4(base.r, base.g, base.b, 0.0f);
    }
    if (!kSpawn) return;

    Spawn sp; sp.want_t = sp.want_r = false;
    float tw1 = 0.f, tw2 = 0.f, fr = 0.f, c_ni = 0.f, c_nt = 0.f;
    bool c_exiting = false;
    if (hit) {
        if (level == 0) {                                                        // 746-757
            c_ni = S.bkg_eta; c_nt = material_eta(S, h.obj); c_exiting = false;
            r.stack[0] = h.obj; r.ns = 1;
        } else {
            medium_update(S, r.reflect, r.exiting, r.inc, h.obj, r.n_i, r.n_t, r.stack, r.ns, c_ni, c_nt, c_exiting);
        }
        shade_spawn(S, r.ray, h.obj, sf, c_ni, c_nt, P.max_depth - level, tw1, tw2, fr, sp);
    }

    // ---- spawn: warp-aggregated slot allocation in the next generation's queue -----
    int child_t = -1, child_r = -1;
    const unsigned mt = __ballot_sync(0xffffffffu, sp.want_t);
    const unsigned mr = __ballot_sync(0xffffffffu, sp.want_r);
    if (mt | mr) {                                   // warp-uniform
        const LevelBuffers& NB = P.level[level + 1];
        const int nt = __popc(mt), nr = __popc(mr);
        unsigned basei = 0;
        if ((threadIdx.x & 31) == 0) {
            basei = atomicAdd(&P.counters->rays[level + 1], (unsigned)(nt + nr));
            if (nr) atomicAdd(&P.counters->rays_reflect[level + 1], (unsigned)nr);
        }
        basei = __shfl_sync(0xffffffffu, basei, 0);
        const unsigned lt = lanemask_lt();
        const int flags_common = (c_exiting ? kRayStateExiting : 0) | (r.ns << kRayNStackShift);
        if (sp.want_t) {
            const unsigned s = basei + (unsigned)__popc(mt & lt);
            if (s < (unsigned)NB.capacity) {
                child_t = (int)s;
                NB.rq_o[s] = make_float4(sf.p.x, sf.p.y, sf.p.z, c_ni);
                NB.rq_d[s] = make_float4(sp.T.x, sp.T.y, sp.T.z, c_nt);
                NB.rq_m[s] = make_int4((int)slot, h.obj, flags_common, r.stack[0]);
                NB.rq_s[s] = make_int4(r.stack[1], r.stack[2], r.stack[3], r.stack[4]);
            } else {
                atomicOr(&P.counters->overflow, kOverflowQueue);
            }
        }
        if (sp.want_r) {
            const unsigned s = basei + (unsigned)nt + (unsigned)__popc(mr & lt);
            if (s < (unsigned)NB.capacity) {
                child_r = (int)s;
                NB.rq_o[s] = make_float4(sf.p.x, sf.p.y, sf.p.z, c_ni);
                NB.rq_d[s] = make_float4(sp.R.x, sp.R.y, sp.R.z, c_nt);
                NB.rq_m[s] = make_int4((int)slot, h.obj, flags_common | kRayKindReflect, r.stack[0]);
                NB.rq_s[s] = make_int4(r.stack[1], r.stack[2], r.stack[3], r.stack[4]);
            } else {
                atomicOr(&P.counters->overflow, kOverflowQueue);
            }
        }
    }

    if (!hit || level == P.max_depth) return;
    // a node without children is finished by the combine pass like any other: x + 0 leaves a
    // clamped x unchanged
    LB.nr_b[slot] = make_float4(tw1, tw2, fr, 0.0f);
    LB.nr_c[slot] = make_int2(child_t, child_r);
    if (kColour) LB.res[slot] = make_float4(0.f, 0.f, 0.f, kResPending);   // (LBVH pipeline: the trace kernel wrote it)
}

// Fused path (brute force): trace + shadow rays + shade + spawn of one slot in one thread.
template <class Accel>
RT_DEV void process_slot(const FrameParams& P, const Accel& acc, int level, unsigned slot, bool in_range)
{
    const SceneView& S = P.scene;
    RayIn r;
    const bool active = in_range && load_ray(P, level, slot, r);
    Hit h; h.obj = -1; h.t = 0.f; h.b = h.g = 0.f;
    bool aborted = false;
    if (active) h = trace_ray(P, acc, level, r, aborted);
    const bool hit = active && !aborted && h.obj >= 0;

    // nodes[level] counts ShadeRay invocations: one atomic per warp
    {
        const unsigned hm = __ballot_sync(0xffffffffu, hit);
        if ((threadIdx.x & 31) == 0 && hm) atomicAdd(&P.counters->nodes[level], (unsigned)__popc(hm));
    }
    if (active && level == 0 && P.primary_ids) {
        int lrow, col;
        P.primary_ids[pixel_of_slot(P, slot, lrow, col)] = hit ? S.obj[h.obj].z : 0;
    }
    if (P.primary_ids) {   // instrumented frame: one texel per shaded hit on a textured object
        const unsigned tm = __ballot_sync(0xffffffffu, hit && S.obj[h.obj].y >= 0);
        if ((threadIdx.x & 31) == 0 && tm) atomicAdd(&P.counters->texels, (unsigned long long)__popc(tm));
    }
    const DirectShadow<Accel> direct{ acc };
    shade_and_spawn<true, true>(P, direct, level, slot, r, h, hit);
    if (in_range && !hit) write_miss(P, level, slot, !active, aborted);
}

// AccelFactory builds the per-block accelerator view (stages shared memory if any).
// blocks per SM the fused kernel is compiled for: 8 (64 registers, ~250 B of spills) against the compiler's own
// 96 registers / 5 blocks: test7 at 4K 1.67 -> 1.56 ms (6 blocks: 1.62); the kernel waits on dependent exact-rounding
// chains, more resident warps hide them (profiles/r2_level_kernel_test7_ncu_full.txt: issue 51 %, 28 % occupancy before)
#ifndef RT_LEVEL_MIN_BLOCKS
#define RT_LEVEL_MIN_BLOCKS 8
#endif
template <class AccelFactory>
__global__ void __launch_bounds__(kBlock, RT_LEVEL_MIN_BLOCKS) level_kernel(const __grid_constant__ FrameParams P, int level)
{
    extern __shared__ float4 smem[];
    const auto acc = AccelFactory::make(P.scene, smem);
    const unsigned count = level == 0 ? (unsigned)P.n_local_pixels
                                      : min(P.counters->rays[level], (unsigned)P.level[level].capacity);
    const unsigned stride = gridDim.x * blockDim.x;
    // all lanes of a warp run the same number of iterations (full-mask ballots inside)
    const unsigned rounded = (count + 31u) & ~31u;
    for (unsigned slot = blockIdx.x * blockDim.x + threadIdx.x; slot < rounded; slot += stride)
        process_slot(P, acc, level, slot, slot < count);
    AccelFactory::flush(acc, P.counters);
}

RT_DEV Col child_term(const LevelBuffers& NB, int child, bool& zero)
{
    const float4 r = NB.res[child];
    zero = (r.w == kResZeroTerm);
    return mkc(r.x, r.y, r.z);
}

__global__ void __launch_bounds__(256) combine_kernel(const __grid_constant__ FrameParams P, int level)
{
    const LevelBuffers& LB = P.level[level];
    const LevelBuffers& NB = P.level[level + 1 < kMaxLevels ? level + 1 : level];
    const unsigned count = level == 0 ? (unsigned)P.n_local_pixels
                                      : min(P.counters->rays[level], (unsigned)LB.capacity);
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned slot = blockIdx.x * blockDim.x + threadIdx.x; slot < count; slot += stride) {
        float4 r = LB.res[slot];
        if (r.w == kResPending) {
            const float4 a = LB.nr_a[slot], b = LB.nr_b[slot];      // base colour | tw1, tw2, F_r
            const int2 ch = LB.nr_c[slot];
            const int ct = ch.x, cr = ch.y;
            Col acc = mkc(a.x, a.y, a.z);
            Col tt = mkc(0.f, 0.f, 0.f), tr = mkc(0.f, 0.f, 0.f);
            if (ct >= 0) {
                bool zero;
                const Col c = child_term(NB, ct, zero);
                if (!zero) tt = cmulf(cmulf(c, b.x), b.y);                       // 1083 / 1087
            }
            if (cr >= 0) {
                bool zero;
                const Col c = child_term(NB, cr, zero);
                tr = cmulf(c, b.z);                                              // 1194 / 1198
            }
            acc = cadd(cadd(acc, tt), tr);                                       // 1206
            r = make_float4(acc.r, acc.g, acc.b, kResFinal);
            if (level != 0) LB.res[slot] = r;
        }
        if (level == 0) {
            // quantise, main.cpp:760-762: (int) map(c, 0, 1, 0, 255)
            const float c[3] = { r.x, r.y, r.z };
            unsigned char out[3]; unsigned flags = (r.w == kResPrimaryMiss) ? 8u : 0u;
            #pragma unroll
            for (int k = 0; k < 3; k++) {
                const int q = ref_f2i(mapf(c[k], 0.0f, 1.0f, 0.0f, 255.0f));
                if (q == (int)0x80000000) { flags |= 1u << k; out[k] = 0; }
                else out[k] = (unsigned char)min(max(q, 0), 255);
            }
            int lrow, col;
            const int dst = pixel_of_slot(P, slot, lrow, col);
            if (dst >= 0) P.tiles[dst] = make_uchar4(out[0], out[1], out[2], (unsigned char)flags);
        }
    }
}

}  // namespace rt
