// bvh.cuh — LBVH traversal that returns what the reference's brute-force loops
// return.
//
// The tree is 4-wide with quantised child boxes: the binary LBVH of bvh.cu is collapsed
// by skipping every other level, and each 64-byte node stores up to four children as
// 8-bit boxes in the node's own frame (as in Ylitie et al. 2017, "Efficient Incoherent
// Ray Traversal on GPUs Through Compressed Wide BVHs", 4- instead of 8-wide):
//   w0 = origin.xyz (fp32), ex | ey << 8 | ez << 16     plane = origin + q * 2^(e - 127)
//   w1 = child[0..3]   >= 0: index of an inner node; < 0: a leaf, ~child = (first << 3) |
//                      (count - 1): `count` consecutive slots of the Morton-ordered
//                      primitive array from slot `first`.  Unused slots hold an inverted
//                      box and the code of the never-hit dummy record (bvh.cu).
//   w2 = child 0 {lo.x, hi.x, lo.y, hi.y}{lo.z, hi.z, 0, 0}, child 1 likewise     (bytes)
//   w3 = child 2, child 3
// One fetch decides four subtrees: against the 2-wide fp32 node of the same size that halves
// the node visits per ray, i.e. the L1 gather traffic (the binding unit) and
// the length of the dependent-load chain that sets the tail latency of a launch.
//
// Leaf slots are 5-word records (kTriWords) for triangles AND spheres:
//   triangle: the record of scene_device.h, w4 = d12, kshadow, as_float(order), 0
//   sphere  : w0 = centre, radius ; w4 = 0, kshadow, as_float(order), as_float(1)
//
// Conservativeness.  Primitive boxes are padded at build time (bvh.cu) so that every point
// the reference's float arithmetic accepts lies inside.  Quantisation rounds lo down / hi up
// and then moves each plane one further quantum outward; the decode below is exact except
// for the rounding of the per-node constant c (<= half a quantum), so a decoded box always
// contains the padded box.  The box test only culls; exact-rounding tests (trace.cuh) run
// at the leaves, so a hit's t, b, g are bit-identical to the brute-force walk; ties in t
// are broken by the iteration order of the reference (smaller `order` wins).
#pragma once

#include "trace.cuh"

namespace rt {

RT_DEV float4 ldp(const float4* p) { return *p; }

// shared-memory part of the persistent walk's stack (entries per thread)
constexpr int kBvhStack = 96;     // local-memory part of the traversal stack (entries)
constexpr int kSmemStack = 24;    // shared-memory part (16 / 32 measured the same or slower)
// A walk never needs more than kStackNeed4(levels) / kStackNeed8(levels) entries on a tree of that many
// levels (an ordered 4-wide step leaves at most 3 entries behind per level; an 8-wide step the rest of
// the group it came from and the group of its inner children, plus at most 8 primitives on top):
// build_bvh refuses trees that could overflow the stacks, the walks raise kOverflowStack should it
// happen all the same.
constexpr int kStackNeed4(int levels) { return 3 * levels + 2; }
constexpr int kStackNeed8(int levels) { return 2 * levels + 10; }
constexpr int kWalkBlock = 128;   // block size of the persistent walk kernels (64 / 32 measured slower)
constexpr int kLeafMax = 4;       // scenes of <= kLeafMax primitives are a single node
constexpr int kLeafDefault = 1;   // primitives per leaf otherwise: with one triangle per leaf the parent's box test
                                  // rejects most of the 80-byte exact tests (measured: 32 -> 6.4 tests per ray)
constexpr int kWalkDone = 0x7fffffff;

// Ray-side constants of the quantised box test.
struct RayQ {
    V3 o, inv;               // origin; 1 / direction (safe_inv)
    unsigned sn[3], sf[3];   // PRMT selectors of the near / far plane byte per axis
};

template <bool kCount>
struct BvhAccelT {
    const float4* nodes;   // 4 words per node, layout above (bit patterns)
    const float4* prims;   // Morton order, kTriWords words per slot
    const float4* tri;     // file order (direct tests of one known object)
    const float4* sph;
    const uint4* nodes8;   // 8-wide tree of the any-hit walk (Node8 below), null when not built
    const float4* prims8;  // ... and its node-ordered primitive records
    int nf, ns;
    int root;              // node index; n_nodes == 0 means empty scene
    int n_nodes;
    bool count_all = false;   // shadow(): keep counting hits after an opaque occluder (parity probes)
    unsigned* overflow_flag = nullptr;   // Counters::overflow of the frame (null: probes), see kOverflow*
    // instrumented variant only (kCount): nodes visited / primitives tested by this thread
    mutable unsigned long long cnt_boxes = 0, cnt_prims = 0;

    RT_DEV static bool finite3(V3 v) { return isfinite(v.x) && isfinite(v.y) && isfinite(v.z); }

    // 1 / d for the box test.  A zero (or denormal-small) component is replaced by +-1e-18:
    // over any distance that matters the ray moves < 1e-15 along that axis, far inside the
    // box padding, and the plane distances stay finite and correctly ordered.
    RT_DEV static float safe_inv(float d)
    {
        return 1.0f / (fabsf(d) > 1.0e-18f ? d : copysignf(1.0e-18f, d));
    }

    RT_DEV static RayQ make_rayq(V3 o, V3 ray)
    {
        RayQ r;
        r.o = o;
        r.inv = mk3(safe_inv(ray.x), safe_inv(ray.y), safe_inv(ray.z));
        // result bytes [q, 0, 0, 0x4B] = the float 2^23 + q; q = byte 0/1 (x, z) or 2/3 (y) of the
        // child word: lo for the near plane of a ray going up the axis, hi otherwise
        const unsigned nx = signbit(r.inv.x) ? 1u : 0u, ny = signbit(r.inv.y) ? 1u : 0u, nz = signbit(r.inv.z) ? 1u : 0u;
        r.sn[0] = 0x7540u | nx;        r.sf[0] = 0x7540u | (nx ^ 1u);
        r.sn[1] = 0x7542u | ny;        r.sf[1] = 0x7542u | (ny ^ 1u);
        r.sn[2] = 0x7540u | nz;        r.sf[2] = 0x7540u | (nz ^ 1u);
        return r;
    }

    // Box test of the four children of one node.  t[k] = entry distance, mask bit k = hit.
    RT_DEV void test_node(const float4* __restrict__ nd, int node, const RayQ& r, float tmin, float tmax,
                          float t[4], int code[4], unsigned& mask) const
    {
        // (as two 256-bit loads the node costs half the L1 requests, but the 8-register alignment costs
        // this kernel, which sits at its register limit, more than that saves: measured)
        const uint4* w = reinterpret_cast<const uint4*>(nd) + (size_t)node * 4;
        const uint4 h = __ldg(w), c = __ldg(w + 1), b01 = __ldg(w + 2), b23 = __ldg(w + 3);
        if (kCount) cnt_boxes += 1;
        // plane distance = (origin + q * 2^e - o) * inv = (2^23 + q) * s + c,
        //   s = 2^e * inv,  c = (origin - o) * inv - 2^23 * s
        const float sx = __uint_as_float((h.w & 0xffu) << 23) * r.inv.x;
        const float sy = __uint_as_float(((h.w >> 8) & 0xffu) << 23) * r.inv.y;
        const float sz = __uint_as_float(((h.w >> 16) & 0xffu) << 23) * r.inv.z;
        const float cx = fmaf(-8388608.0f, sx, (__uint_as_float(h.x) - r.o.x) * r.inv.x);
        const float cy = fmaf(-8388608.0f, sy, (__uint_as_float(h.y) - r.o.y) * r.inv.y);
        const float cz = fmaf(-8388608.0f, sz, (__uint_as_float(h.z) - r.o.z) * r.inv.z);
        const unsigned a[4] = { b01.x, b01.z, b23.x, b23.z }, b[4] = { b01.y, b01.w, b23.y, b23.w };
        code[0] = (int)c.x; code[1] = (int)c.y; code[2] = (int)c.z; code[3] = (int)c.w;
        mask = 0;
        #pragma unroll
        for (int k = 0; k < 4; k++) {
            const float nx = fmaf(__uint_as_float(__byte_perm(a[k], 0x4B000000u, r.sn[0])), sx, cx);
            const float ny = fmaf(__uint_as_float(__byte_perm(a[k], 0x4B000000u, r.sn[1])), sy, cy);
            const float nz = fmaf(__uint_as_float(__byte_perm(b[k], 0x4B000000u, r.sn[2])), sz, cz);
            const float fx = fmaf(__uint_as_float(__byte_perm(a[k], 0x4B000000u, r.sf[0])), sx, cx);
            const float fy = fmaf(__uint_as_float(__byte_perm(a[k], 0x4B000000u, r.sf[1])), sy, cy);
            const float fz = fmaf(__uint_as_float(__byte_perm(b[k], 0x4B000000u, r.sf[2])), sz, cz);
            // fminf / fmaxf drop NaN (overflowing products of degenerate rays): conservative
            const float tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, tmin));
            const float tf = fminf(fminf(fx, fy), fminf(fz, tmax));
            t[k] = tn;
            if (tn <= tf) mask |= 1u << k;
        }
    }

    // ---- traversal stacks ---------------------------------------------------------------------
    // pop() -> kWalkDone when empty; push(v); push_sorted(n, c1, c2, c3): the 2nd .. 4th nearest of
    // n hit children, farthest first, so that c1 ends on top.  An entry that does not fit raises
    // kOverflowStack (the frame is then reported as failed, never silently wrong).
    struct LocalStack {
        int e[kBvhStack];
        int sp = 0;
        unsigned* flag;
        RT_DEV int pop() { return sp ? e[--sp] : kWalkDone; }
        RT_DEV void push(int v)
        {
            if (sp < kBvhStack) e[sp++] = v;
            else if (flag) atomicOr(flag, kOverflowStack);
        }
        RT_DEV void push_sorted(int n, int c1, int c2, int c3)
        {
            if (n > 3) push(c3);
            if (n > 2) push(c2);
            if (n > 1) push(c1);
        }
    };

    // The persistent walk's stack: the first kSmemStack entries live in shared memory, laid out
    // [entry][thread] so that a warp's pushes and pops never conflict whatever the lanes' depths are
    // (a local-memory stack costs one L1 wavefront per distinct depth, and L1 is what binds the
    // loop); deeper entries overflow into local memory.  bot: entries below it were given to
    // helper lanes (work sharing).
    struct SharedStack {
        int* col;                // &s_stack[0][threadIdx.x]
        int* deep;               // local-memory part
        int sp, bot;
        unsigned* flag;
        RT_DEV int pop()
        {
            if (sp == bot) return kWalkDone;
            --sp;
            return sp < kSmemStack ? col[sp * kWalkBlock] : deep[sp - kSmemStack];
        }
        RT_DEV void push(int v)
        {
            if (sp < kSmemStack) col[sp * kWalkBlock] = v;
            else if (sp < kSmemStack + kBvhStack) deep[sp - kSmemStack] = v;
            else { if (flag) atomicOr(flag, kOverflowStack); return; }
            sp++;
        }
        RT_DEV void push_sorted(int n, int c1, int c2, int c3)
        {
            if (sp + 3 <= kSmemStack) {
                // the common case without a branch per entry: predicated stores at fixed offsets
                // below the new top
                int* top = col + (sp + n - 1) * kWalkBlock;
                if (n > 1) top[-1 * kWalkBlock] = c1;
                if (n > 2) top[-2 * kWalkBlock] = c2;
                if (n > 3) top[-3 * kWalkBlock] = c3;
                sp += max(n, 1) - 1;
            } else {
                if (n > 3) push(c3);
                if (n > 2) push(c2);
                if (n > 1) push(c1);
            }
        }
    };

    RT_DEV static int select4(int c0, int c1, int c2, int c3, unsigned idx)
    {
        const int lo = (idx & 1u) ? c1 : c0, hi = (idx & 1u) ? c3 : c2;
        return (idx & 2u) ? hi : lo;
    }

    // One traversal step at inner node `node`: tests its four children, leaves all hit children
    // but the nearest on the stack (far to near) and returns the nearest (a node index, a leaf
    // code, or pop()'s result when nothing was hit).  Entry distances are >= tmin >= 0, so their
    // bit patterns order like unsigned integers: the child index rides in the two lowest
    // mantissa bits and the four keys go through a five-comparator network of integer min / max
    // (order is only a heuristic here: every accepted hit comes from the exact primitive test).
    // No branch depends on how many children were hit, so the lanes of a warp stay together
    // (A/B against round 1's float compare-and-swap network with a branch per push, 4K stress frame:
    // 37.50 -> 37.03 ms, 1/8 shard 5.98 -> 5.88 ms; profiles/r2_ab_sort_tiles.txt).
    template <class Stack>
    RT_DEV int step(const float4* __restrict__ nd, int node, const RayQ& r, float tmin, float tmax, Stack& st) const
    {
        float t[4]; int code[4]; unsigned mask;
        test_node(nd, node, r, tmin, tmax, t, code, mask);
        unsigned k[4];
        #pragma unroll
        for (int i = 0; i < 4; i++)
            k[i] = ((mask >> i) & 1u) ? ((__float_as_uint(t[i]) & ~3u) | (unsigned)i) : 0xffffffffu;
        auto cswap = [&](int i, int j) { const unsigned lo = min(k[i], k[j]), hi = max(k[i], k[j]); k[i] = lo; k[j] = hi; };
        cswap(0, 1); cswap(2, 3); cswap(0, 2); cswap(1, 3); cswap(1, 2);
        const int n = __popc(mask);
        const int c0 = select4(code[0], code[1], code[2], code[3], k[0]);
        const int c1 = select4(code[0], code[1], code[2], code[3], k[1]);
        const int c2 = select4(code[0], code[1], code[2], code[3], k[2]);
        const int c3 = select4(code[0], code[1], code[2], code[3], k[3]);
        st.push_sorted(n, c1, c2, c3);
        return n ? c0 : st.pop();
    }

    // ---- the 8-wide tree (any-hit walks) ------------------------------------------------------
    // Node8, 96 bytes (bvh.cu: collapse8_kernel):
    //   w0: frame origin x, y, z | ex, ey, ez, n_inner (bits 24..27)
    //   w1: child_base, prim_base, -, -
    //   w2 / w3 / w4: x / y / z planes, word j = lo, hi of slot 2j | lo, hi of slot 2j+1
    // Slots 0 .. n_inner-1 are inner nodes (node index child_base + slot), the next slots are
    // primitives (record prim_base + slot - n_inner), the rest are empty (lo = 255 > hi = 0).
    // Children are found by base + rank, so a stack entry can stand for ALL pending inner
    // children of a node: (child_base << 8) | mask of slots still to visit ("group"); >= 0.
    // Primitives travel as the usual leaf codes (< 0).
    template <class Stack>
    RT_DEV int step8(const uint4* __restrict__ nd8, int group, const RayQ& r, float tmin, float tmax, Stack& st) const
    {
        // next child of the group; what is left of the group goes back on the stack (caller)
        const unsigned mask = (unsigned)group & 0xffu;
        const unsigned slot = (unsigned)__ffs((int)mask) - 1u;
        const unsigned node = ((unsigned)group >> 8) + slot;
        const unsigned rest = mask & (mask - 1u);
        if (rest) st.push((int)(((unsigned)group & ~0xffu) | rest));
        const uint4* w = nd8 + (size_t)node * 6;
        uint4 h, c, px, py, pz;
        asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
            : "=r"(h.x), "=r"(h.y), "=r"(h.z), "=r"(h.w), "=r"(c.x), "=r"(c.y), "=r"(c.z), "=r"(c.w) : "l"(w));
        asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
            : "=r"(px.x), "=r"(px.y), "=r"(px.z), "=r"(px.w), "=r"(py.x), "=r"(py.y), "=r"(py.z), "=r"(py.w) : "l"(w + 2));
        pz = __ldg(w + 4);
        if (kCount) cnt_boxes += 1;
        const float sx = __uint_as_float((h.w & 0xffu) << 23) * r.inv.x;
        const float sy = __uint_as_float(((h.w >> 8) & 0xffu) << 23) * r.inv.y;
        const float sz = __uint_as_float(((h.w >> 16) & 0xffu) << 23) * r.inv.z;
        const float cx = fmaf(-8388608.0f, sx, (__uint_as_float(h.x) - r.o.x) * r.inv.x);
        const float cy = fmaf(-8388608.0f, sy, (__uint_as_float(h.y) - r.o.y) * r.inv.y);
        const float cz = fmaf(-8388608.0f, sz, (__uint_as_float(h.z) - r.o.z) * r.inv.z);
        const unsigned wx[4] = { px.x, px.y, px.z, px.w }, wy[4] = { py.x, py.y, py.z, py.w }, wz[4] = { pz.x, pz.y, pz.z, pz.w };
        // (lo, hi) of an even slot sit in bytes 0, 1 of its word, of an odd slot two further up
        const unsigned yn = r.sn[1] - 2u, yf = r.sf[1] - 2u;
        unsigned hit = 0;
        #pragma unroll
        for (int k = 0; k < 8; k++) {
            const unsigned up = (k & 1) ? 2u : 0u;
            const float nx = fmaf(__uint_as_float(__byte_perm(wx[k >> 1], 0x4B000000u, r.sn[0] + up)), sx, cx);
            const float ny = fmaf(__uint_as_float(__byte_perm(wy[k >> 1], 0x4B000000u, yn + up)), sy, cy);
            const float nz = fmaf(__uint_as_float(__byte_perm(wz[k >> 1], 0x4B000000u, r.sn[2] + up)), sz, cz);
            const float fx = fmaf(__uint_as_float(__byte_perm(wx[k >> 1], 0x4B000000u, r.sf[0] + up)), sx, cx);
            const float fy = fmaf(__uint_as_float(__byte_perm(wy[k >> 1], 0x4B000000u, yf + up)), sy, cy);
            const float fz = fmaf(__uint_as_float(__byte_perm(wz[k >> 1], 0x4B000000u, r.sf[2] + up)), sz, cz);
            const float tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, tmin));
            const float tf = fminf(fminf(fx, fy), fminf(fz, tmax));
            if (tn <= tf) hit |= 1u << k;
        }
        const unsigned n_inner = (h.w >> 24) & 15u;
        const unsigned inner = hit & ((1u << n_inner) - 1u);
        unsigned leaves = hit >> n_inner;
        if (inner) st.push((int)((c.x << 8) | inner));
        if (!leaves) return st.pop();
        // primitives on top of the stack (an any-hit walk wants them first); the last one does not
        // make the round trip through the stack
        for (;;) {
            const unsigned k = (unsigned)__ffs((int)leaves) - 1u;
            leaves &= leaves - 1u;
            const int leaf = ~(int)((c.y + k) << 3);
            if (!leaves) return leaf;
            st.push(leaf);
        }
    }

    // Generic traversal, "while-while" with one postponed leaf per lane (Aila & Laine 2009):
    // lanes keep descending inner nodes until every lane of the warp holds a leaf, then all
    // test primitives together; that keeps box tests and primitive tests in separate,
    // mostly converged loops.  `visit(slot_words, tcur)` tests one primitive and returns the
    // new upper bound for culling (or a negative value to stop the whole walk).  A postponed
    // leaf is visited late, so boxes may be tested against a stale (larger) bound: that only
    // costs work, never a hit.
    template <class Visit>
    RT_DEV void walk(V3 o, V3 ray, float tmin, float tmax, Visit& visit) const
    {
        if (n_nodes == 0 || !finite3(ray) || !finite3(o)) return;   // NaN/inf rays hit nothing in the reference either
        const RayQ rq = make_rayq(o, ray);
        constexpr int kDone = kWalkDone;
        LocalStack st;
        st.flag = overflow_flag;
        int node = root;      // >= 0: inner node index, < 0: leaf code, kDone: nothing left
        int leaf = 0;         // postponed leaf code (< 0), 0 = none
        while (node != kDone) {
            while ((unsigned)node < (unsigned)kDone) {
                node = step(nodes, node, rq, tmin, tmax, st);
                if (node < 0 && leaf == 0) {                       // postpone the first leaf, keep descending
                    leaf = node;
                    node = st.pop();
                }
                if (!__any_sync(__activemask(), leaf == 0)) break; // every lane has a leaf to test
            }
            while (leaf < 0) {
                const int code = ~leaf;
                const int first = code >> 3, count = (code & 7) + 1;
                for (int k = 0; k < count; k++) {
                    if (kCount) cnt_prims++;
                    tmax = visit(prims + (size_t)(first + k) * kTriWords, tmax);
                    if (tmax < 0.0f) return;
                }
                leaf = 0;
                if (node < 0) {                                    // the next node is a leaf too: take it now
                    leaf = node;
                    node = st.pop();
                }
            }
        }
    }

    // The same on the 8-wide tree (no ordering: any-hit walks and the parity probes of that tree).
    template <class Visit>
    RT_DEV void walk8(V3 o, V3 ray, float tmin, float tmax, Visit& visit) const
    {
        if (n_nodes == 0 || nodes8 == nullptr || !finite3(ray) || !finite3(o)) return;
        const RayQ rq = make_rayq(o, ray);
        LocalStack st;
        st.flag = overflow_flag;
        int node = 1;         // the group "slot 0 of base 0" = node 0
        while (node != kWalkDone) {
            if (node >= 0) { node = step8(nodes8, node, rq, tmin, tmax, st); continue; }
            if (kCount) cnt_prims++;
            tmax = visit(prims8 + (size_t)((~node) >> 3) * kTriWords, tmax);
            if (tmax < 0.0f) return;
            node = st.pop();
        }
    }

    // Persistent form of walk(): a warp keeps pulling jobs (rays) from a global counter and
    // replaces lanes whose ray has finished, so a warp does not idle behind its longest ray
    // (Aila & Laine's "persistent while-while" with dynamic fetch).  All 32 lanes must call it.
    //   job.begin(j, o, ray, tmin, tmax) -> false when job j needs no tree walk
    //   job.visit(words, tcur)           -> new culling bound, < 0 to stop this ray
    //   job.retire(bool finishing)       -> called by ALL lanes whenever some lane finishes
    //                                       (warp collectives allowed inside)
    //   job.next(o, ray, tmin, tmax)     -> true when the finished lane's job has another ray
    //   Job::kWide8                      -> walk the 8-wide tree (any hit, no ordering) when there is one
    //   Job::kSteal                      -> once the job queue is dry, lanes whose ray has finished take
    //                                       subtrees off the stacks of the lanes still walking (a kernel
    //                                       otherwise ends with ~0.3 ms in which a few long rays keep a few
    //                                       lanes of every warp busy).  Needs a result that is a fold over
    //                                       the primitives in any order:
    //   job.adopt(src, take, o, ray)     -> ALL lanes; lanes with `take` copy the ray state of lane src
    //   job.merge(src, mine)             -> ALL lanes; the lane with `mine` folds lane src's result into its own
    //   job.share(group, tmax)           -> the lanes (mask `group`, >= 2) walking one ray: returns the culling
    //                                       bound they should all use from now on, < 0 to stop
    RT_DEV static unsigned long long globaltimer()
    {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        return t;
    }

    template <class Job>
    RT_DEV void persistent_walk(unsigned total, unsigned* counter, Job& job, int refill_lanes, int inner_min = 1, int leaf_min = 1,
                                WalkClock* clock = nullptr) const
    {
        constexpr int kDone = kWalkDone;
        constexpr unsigned kFull = 0xffffffffu;
        // loop invariants in registers (the accelerator object itself may sit in local memory)
        const float4* __restrict__ const nd = nodes;
        const uint4* __restrict__ const nd8 = nodes8;
        const bool wide8 = Job::kWide8 && nodes8 != nullptr;
        const float4* __restrict__ const pr = wide8 ? prims8 : prims;
        const int root_ = wide8 ? 1 : root;      // 8-wide: the group "slot 0 of base 0" = node 0
        const bool have_tree = n_nodes != 0;
        const unsigned lane = threadIdx.x & 31u;
        const unsigned lt = (1u << lane) - 1u;
        __shared__ int s_stack[kSmemStack][kWalkBlock];
        int deep[kBvhStack];
        const int tid = threadIdx.x;
        SharedStack st;
        st.col = &s_stack[0][tid]; st.deep = deep; st.sp = 0; st.bot = 0; st.flag = overflow_flag;
        int node = kDone, leaf = 0;
        unsigned owner = lane;       // != lane: this lane walks a subtree for lane `owner`
        bool alive = false, pool_empty = false;
        RayQ rq = make_rayq(mk3(0.f, 0.f, 0.f), mk3(1.f, 1.f, 1.f));
        float tmin = 0.f, tmax = 0.f;
        unsigned long long t_empty = 0;
        bool worked = false;
        unsigned d_steps = 0, d_lanes = 0, d_steals = 0;
        if (kCount && clock && lane == 0) atomicMax(&clock->inv_first, ~globaltimer());
        auto start = [&](bool need_walk, V3 o, V3 ray) {
            st.sp = 0; st.bot = 0; leaf = 0; owner = lane;
            node = (need_walk && have_tree && finite3(ray) && finite3(o)) ? root_ : kDone;
            rq = make_rayq(o, ray);
        };
        for (;;) {
            const unsigned dead = __ballot_sync(kFull, !alive);
            if (!pool_empty && (__popc(dead) >= refill_lanes || dead == kFull)) {
                unsigned base = 0;
                if (lane == 0) base = atomicAdd(counter, (unsigned)__popc(dead));
                base = __shfl_sync(kFull, base, 0);
                if (base + (unsigned)__popc(dead) >= total) {
                    pool_empty = true;
                    if (kCount && clock) t_empty = globaltimer();
                }
                if (!alive) {
                    const unsigned j = base + (unsigned)__popc(dead & lt);
                    if (j < total) {
                        V3 o, ray;
                        const bool need_walk = job.begin(j, o, ray, tmin, tmax);
                        alive = true;
                        start(need_walk, o, ray);
                    }
                }
            }
            if constexpr (Job::kSteal) {
                if (pool_empty) {
                    const unsigned idle = __ballot_sync(kFull, !alive);
                    // (only the shared-memory part of a stack can be read by another lane)
                    const int avail = alive ? min(st.sp, kSmemStack) - st.bot : 0;
                    const unsigned donors = __ballot_sync(kFull, avail > 0);
                    if (idle != 0u && donors != 0u) {
                        // the r-th idle lane takes the lower half (the older entries: the farther,
                        // larger subtrees) of the r-th donor's stack, order kept
                        const int pairs = min(__popc(idle), __popc(donors));
                        const int rank = __popc((alive ? donors : idle) & lt);
                        const bool give = avail > 0 && rank < pairs;
                        const bool take = !alive && rank < pairs;
                        const int n_give = give ? (avail + 1) / 2 : 0;
                        const int src = take ? (int)__fns(donors, 0, rank + 1) : (int)lane;
                        const int from = __shfl_sync(kFull, st.bot, src), n_take = __shfl_sync(kFull, n_give, src);
                        const float tmin_s = __shfl_sync(kFull, tmin, src), tmax_s = __shfl_sync(kFull, tmax, src);
                        const unsigned owner_s = __shfl_sync(kFull, owner, src);
                        V3 o, ray;
                        job.adopt(src, take, o, ray);
                        if (give) st.bot += n_give;
                        if (kCount) d_steals += (unsigned)pairs;
                        __syncwarp();      // the donors' earlier stack writes are visible to the takers
                        if (take) {
                            alive = true; owner = owner_s;
                            tmin = tmin_s; tmax = tmax_s;
                            rq = make_rayq(o, ray);
                            const int col = tid - (int)lane + src;
                            for (int i = 0; i < n_take; i++) s_stack[i][tid] = s_stack[from + i][col];
                            st.sp = n_take; st.bot = 0; leaf = 0;
                            node = st.pop();
                            if (node < 0) { leaf = node; node = st.pop(); }
                        }
                        __syncwarp();
                    }
                    // lanes on the same ray share what bounds / ends the walk
                    const unsigned group = __match_any_sync(kFull, alive ? owner : 32u + lane);
                    if (group & (group - 1u)) {
                        tmax = job.share(group, tmax);
                        if (tmax < 0.0f) { node = kDone; leaf = 0; st.sp = st.bot; }
                    }
                }
            }
            if (!__any_sync(kFull, alive)) break;
            if (kCount) worked = true;
            // inner nodes, until every live lane holds a leaf (or has nothing left)
            for (;;) {
                if (kCount && pool_empty) {
                    d_steps++;
                    d_lanes += (unsigned)__popc(__ballot_sync(kFull, alive && (unsigned)node < (unsigned)kDone));
                }
                if (alive && (unsigned)node < (unsigned)kDone) {
                    if (Job::kWide8 && wide8) node = step8(nd8, node, rq, tmin, tmax, st);
                    else node = step(nd, node, rq, tmin, tmax, st);
                    if (node < 0 && leaf == 0) {
                        leaf = node;
                        node = st.pop();
                    }
                }
                if (__popc(__ballot_sync(kFull, alive && leaf == 0 && (unsigned)node < (unsigned)kDone)) < inner_min) break;
            }
            // pending leaves
            for (bool first = true;; first = false) {
                // always one pass when any lane holds a leaf (progress), further passes only while
                // enough lanes still have one
                const int holders = __popc(__ballot_sync(kFull, alive && leaf < 0));
                if (holders == 0 || (!first && holders < leaf_min)) break;
                if (alive && leaf < 0) {
                    const int code = ~leaf;
                    const int first_slot = code >> 3, count = (code & 7) + 1;
                    leaf = 0;
                    for (int k = 0; k < count; k++) {
                        if (kCount) cnt_prims++;
                        tmax = job.visit(pr + (size_t)(first_slot + k) * kTriWords, tmax);
                        if (tmax < 0.0f) { node = kDone; st.sp = st.bot; break; }
                    }
                    if (node < 0) {
                        leaf = node;
                        node = st.pop();
                    }
                }
            }
            bool fin = alive && node == kDone && leaf == 0;
            if constexpr (Job::kSteal) {
                if (pool_empty) {
                    // helpers that are done hand their result to the owner of the ray and go idle
                    unsigned hf = __ballot_sync(kFull, fin && owner != lane);
                    while (hf) {
                        const int src = __ffs(hf) - 1;
                        hf &= hf - 1u;
                        const unsigned to = __shfl_sync(kFull, owner, src);
                        job.merge(src, lane == to);
                    }
                    if (fin && owner != lane) { alive = false; owner = lane; fin = false; }
                    // an owner retires once nobody walks for it any more
                    const unsigned group = __match_any_sync(kFull, alive ? owner : 32u + lane);
                    if (fin && (group & ~(1u << lane)) != 0u) fin = false;
                }
            }
            if (__any_sync(kFull, fin)) {
                job.retire(fin);
                if (fin) {
                    // a job may be a chain of rays (the lights of one node): go on in this lane
                    V3 o, ray;
                    if (job.next(o, ray, tmin, tmax)) start(job.walk_needed(), o, ray);
                    else alive = false;
                }
            }
        }
        if (kCount && clock && lane == 0) {
            const unsigned long long t = globaltimer();
            atomicMax(&clock->last, t);
            if (worked) {
                atomicMax(&clock->inv_empty, ~t_empty);
                atomicAdd(&clock->drain_sum, t - t_empty);
                atomicAdd(&clock->warps, 1u);
                atomicAdd(&clock->drain_steps, (unsigned long long)d_steps);
                atomicAdd(&clock->drain_lane_steps, (unsigned long long)d_lanes);
                atomicAdd(&clock->drain_steals, (unsigned long long)d_steals);
            }
        }
    }

    struct ClosestVisit {
        V3 o, ray; float tmin; Hit best; int best_order; bool skip_spheres;
        RT_DEV float operator()(const float4* w, float tmax)
        {
            const float4 w4 = w[4];
            const int order = __float_as_int(w4.z);
            if (__float_as_int(w4.w) == 0) {
                float t;
                if (tri_distance(w[0], o, ray, t) && t > tmin && (t < best.t || (t == best.t && order < best_order))) {
                    float b, g;
                    if (tri_inside(w[1], w[2], w[3], w4, o, ray, t, b, g)) { best.t = t; best.obj = order; best.b = b; best.g = g; best_order = order; }
                }
            } else if (!skip_spheres) {
                float tf, tn;
                const int n = sphere_roots(w[0], o, ray, tf, tn);
                if (n >= 1 && tf > tmin && (tf < best.t || (tf == best.t && order < best_order))) { best.t = tf; best.obj = order; best_order = order; }
                if (n == 2 && tn > tmin && (tn < best.t || (tn == best.t && order < best_order))) { best.t = tn; best.obj = order; best_order = order; }
            }
            return fminf(tmax, best.t);
        }
    };

    // |ray|^2 far from 1: the sphere quadratic (A == 1, main.cpp:1232) accepts a region no box
    // bounds.  Such rays test every sphere directly and use the tree for triangles only.
    RT_DEV bool spheres_unbounded(V3 ray) const
    {
        const float s2 = ray.x * ray.x + ray.y * ray.y + ray.z * ray.z;
        return ns > 0 && !(fabsf(s2 - 1.0f) <= 1.0e-5f);   // bvh.cu pads sphere boxes for 1.2e-5
    }

    RT_DEV Hit closest(V3 o, V3 ray, float tmin) const
    {
        ClosestVisit v; v.o = o; v.ray = ray; v.tmin = tmin; v.skip_spheres = spheres_unbounded(ray);
        v.best.t = 3.402823466e+38f; v.best.obj = -1; v.best.b = v.best.g = 0.0f; v.best_order = 0x7fffffff;
        walk(o, ray, tmin, 3.402823466e+38f, v);
        if (v.skip_spheres) {
            for (int i = 0; i < ns; i++) {
                float tf, tn;
                const int n = sphere_roots(sph[(size_t)i * kSphWords], o, ray, tf, tn);
                const int order = nf + i;
                if (n >= 1 && tf > tmin && (tf < v.best.t || (tf == v.best.t && order < v.best_order))) { v.best.t = tf; v.best.obj = order; v.best_order = order; }
                if (n == 2 && tn > tmin && (tn < v.best.t || (tn == v.best.t && order < v.best_order))) { v.best.t = tn; v.best.obj = order; v.best_order = order; }
            }
        }
        return v.best;
    }

    // Face-incidence transmission (main.cpp:992-1011) in its order-free form:
    //   abort  <=>  some hit h of an object != B has t(h) > eps and
    //               ( order(h) < order(B)  or  t(h) < tB ),
    // tB = smallest t > eps among B's own hits on this ray (FLT_MAX if none).
    struct FaceRuleVisit {
        V3 o, ray; float tmin, tB; int back; bool aborted;
        RT_DEV float operator()(const float4* w, float tmax)
        {
            const float4 w4 = w[4];
            const int order = __float_as_int(w4.z);
            if (order == back) return tmax;
            const float limit = order < back ? 3.402823466e+38f : tB;
            bool bad = false;
            if (__float_as_int(w4.w) == 0) {
                float t, b, g;
                if (tri_distance(w[0], o, ray, t) && t > tmin && t < limit) bad = tri_inside(w[1], w[2], w[3], w4, o, ray, t, b, g);
            } else {
                float tf, tn;
                const int n = sphere_roots(w[0], o, ray, tf, tn);
                bad = (n >= 1 && tf > tmin && tf < limit) || (n == 2 && tn > tmin && tn < limit);
            }
            if (bad) { aborted = true; return -1.0f; }
            return tmax;
        }
    };

    RT_DEV Hit closest_face_rule(V3 o, V3 ray, float tmin, int back, bool& aborted) const
    {
        // B's own hit by a direct test of its file-order record
        Hit hb; hb.t = 3.402823466e+38f; hb.obj = -1; hb.b = hb.g = 0.0f;
        if (back < nf) {
            const float4* w = tri + (size_t)back * kTriWords;
            float t, b, g;
            if (tri_distance(w[0], o, ray, t) && t > tmin && tri_inside(w[1], w[2], w[3], w[4], o, ray, t, b, g)) { hb.t = t; hb.obj = back; hb.b = b; hb.g = g; }
        } else {
            float tf, tn;
            const int n = sphere_roots(sph[(size_t)(back - nf) * kSphWords], o, ray, tf, tn);
            if (n >= 1 && tf > tmin && tf < hb.t) { hb.t = tf; hb.obj = back; }
            if (n == 2 && tn > tmin && tn < hb.t) { hb.t = tn; hb.obj = back; }
        }
        FaceRuleVisit v; v.o = o; v.ray = ray; v.tmin = tmin; v.tB = hb.t; v.back = back; v.aborted = false;
        // objects before B in iteration order abort at ANY distance: no upper bound for culling
        walk(o, ray, tmin, 3.402823466e+38f, v);
        aborted = v.aborted;
        return hb;
    }

    // Shadow product.  An opaque occluder (k == 0) makes the product exactly 0 whatever
    // the order, so the walk stops there.  Partial factors are multiplied in the
    // reference's iteration order (float products are not associative): they are
    // collected, sorted by order, then applied.  The any-hit kernel's visitor holds kMaxPartial
    // of them in registers; a ray with more transparent occluders than that sets `overflowed`
    // and its node is redone by bvh_shadow_redo_kernel through shadow() below, which has no
    // limit — never a product in the wrong order.
    static constexpr int kMaxPartial = 12;
    struct ShadowVisit {
        V3 o, ray; float tmin, tmax_hit; bool use_tmax, skip_spheres, count_all; int self; int hits;
        int n_partial; int p_order[kMaxPartial]; float p_k[kMaxPartial];
        bool overflowed, opaque;
        RT_DEV void add(int order, float k)
        {
            hits++;
            if (k == 0.0f) { opaque = true; return; }
            if (n_partial < kMaxPartial) { p_order[n_partial] = order; p_k[n_partial] = k; n_partial++; }
            else overflowed = true;
        }
        RT_DEV float operator()(const float4* w, float tmax)
        {
            const float4 w4 = ldp(w + 4);
            const int order = __float_as_int(w4.z);
            if (order == self) return tmax;
            if (__float_as_int(w4.w) == 0) {
                float t, b, g;
                if (tri_distance(ldp(w), o, ray, t) && t > tmin && (!use_tmax || t < tmax_hit) &&
                    tri_inside(ldp(w + 1), ldp(w + 2), ldp(w + 3), w4, o, ray, t, b, g))
                    add(order, w4.y);
            } else if (!skip_spheres) {
                float tf, tn;
                const int n = sphere_roots(ldp(w), o, ray, tf, tn);
                if (n >= 1 && tf > tmin && (!use_tmax || tf < tmax_hit)) add(order, w4.y);
                if (n == 2 && tn > tmin && (!use_tmax || tn < tmax_hit)) add(order, w4.y);
            }
            return ((opaque || overflowed) && !count_all) ? -1.0f : tmax;
        }
    };

    RT_DEV void shadow_begin(V3 o, V3 ray, int self, float tmin, float tmax, bool use_tmax, ShadowVisit& v) const
    {
        v.o = o; v.ray = ray; v.tmin = tmin; v.tmax_hit = tmax; v.use_tmax = use_tmax; v.self = self;
        v.hits = 0; v.n_partial = 0; v.overflowed = false; v.opaque = false; v.count_all = count_all;
        // The sphere quadratic assumes a unit ray (A == 1, main.cpp:1232).  For a ray of
        // length s != 1 (un-normalised directional lights, main.cpp:895) the accepted set is
        // rho^2 <= r^2 + (s^2 - 1) p^2 — unbounded for s > 1, so no box bounds it:
        // such rays test every sphere directly and use the tree for the triangles only.
        v.skip_spheres = spheres_unbounded(ray);
    }

    // after the walk: spheres of rays no box bounds, then the sort by iteration order
    RT_DEV void shadow_end(ShadowVisit& v) const
    {
        const V3 o = v.o, ray = v.ray;
        const float tmin = v.tmin, tmax = v.tmax_hit;
        const bool use_tmax = v.use_tmax;
        const int self = v.self;
        if (v.skip_spheres && !((v.opaque || v.overflowed) && !count_all)) {
            for (int i = 0; i < ns; i++) {
                if (nf + i == self) continue;
                const float4* w = sph + (size_t)i * kSphWords;
                float tf, tn;
                const int n = sphere_roots(w[0], o, ray, tf, tn);
                if (n >= 1 && tf > tmin && (!use_tmax || tf < tmax)) v.add(nf + i, w[1].x);
                if (n == 2 && tn > tmin && (!use_tmax || tn < tmax)) v.add(nf + i, w[1].x);
                if ((v.opaque || v.overflowed) && !count_all) break;
            }
        }
        if (v.opaque) return;
        // insertion sort by order (stable: the two roots of one sphere share k)
        for (int i = 1; i < v.n_partial; i++) {
            const int ko = v.p_order[i]; const float kk = v.p_k[i];
            int j = i - 1;
            while (j >= 0 && v.p_order[j] > ko) { v.p_order[j + 1] = v.p_order[j]; v.p_k[j + 1] = v.p_k[j]; j--; }
            v.p_order[j + 1] = ko; v.p_k[j + 1] = kk;
        }
    }

    // The same product for ANY number of partial occluders, by one thread on its own (parity
    // probes, bvh_shadow_redo_kernel): the walk is repeated, and each pass keeps the kWindow
    // factors with the smallest iteration keys above the last one applied (key = 2 * order +
    // {0: triangle / far root, 1: near root}: unique per hit, and the two roots of a sphere
    // carry the same factor).  One pass for up to kWindow factors, ceil(n / kWindow) otherwise.
    static constexpr int kWindow = 12;
    struct ShadowWindow {
        V3 o, ray; float tmin, tmax_hit; bool use_tmax, skip_spheres, count_all, counting; int self; int hits;
        int last_key, n; int key[kWindow]; float kf[kWindow];
        bool more, opaque;
        RT_DEV void add(int order, int sub, float k)
        {
            if (counting) hits++;
            if (k == 0.0f) { opaque = true; return; }
            const int ky = 2 * order + sub;
            if (ky <= last_key) return;
            if (n == kWindow) {
                more = true;
                if (ky > key[kWindow - 1]) return;
                n--;                                   // the largest key leaves the window
            }
            int i = n;
            while (i > 0 && key[i - 1] > ky) { key[i] = key[i - 1]; kf[i] = kf[i - 1]; i--; }
            key[i] = ky; kf[i] = k; n++;
        }
        RT_DEV void sphere(const float4& w0, int order, float k)
        {
            float tf, tn;
            const int r = sphere_roots(w0, o, ray, tf, tn);
            if (r >= 1 && tf > tmin && (!use_tmax || tf < tmax_hit)) add(order, 0, k);
            if (r == 2 && tn > tmin && (!use_tmax || tn < tmax_hit)) add(order, 1, k);
        }
        RT_DEV float operator()(const float4* w, float tmax)
        {
            const float4 w4 = w[4];
            const int order = __float_as_int(w4.z);
            if (order == self) return tmax;
            if (__float_as_int(w4.w) == 0) {
                float t, b, g;
                if (tri_distance(w[0], o, ray, t) && t > tmin && (!use_tmax || t < tmax_hit) &&
                    tri_inside(w[1], w[2], w[3], w4, o, ray, t, b, g))
                    add(order, 0, w4.y);
            } else if (!skip_spheres) {
                sphere(w[0], order, w4.y);
            }
            return (opaque && !count_all) ? -1.0f : tmax;
        }
    };

    // wide8: walk the 8-wide tree (what the any-hit kernel walks) instead of the 4-wide one
    RT_DEV int shadow(V3 o, V3 ray, int self, float tmin, float tmax, bool use_tmax, float& mask, bool wide8 = false) const
    {
        ShadowWindow v;
        v.o = o; v.ray = ray; v.tmin = tmin; v.tmax_hit = tmax; v.use_tmax = use_tmax; v.self = self;
        v.count_all = count_all; v.skip_spheres = spheres_unbounded(ray);
        v.hits = 0; v.last_key = -1; v.counting = true;
        for (;;) {
            v.n = 0; v.more = false; v.opaque = false;
            if (wide8 && nodes8 != nullptr) walk8(o, ray, tmin, use_tmax ? tmax : 3.402823466e+38f, v);
            else walk(o, ray, tmin, use_tmax ? tmax : 3.402823466e+38f, v);
            if (v.skip_spheres && !(v.opaque && !count_all))
                for (int i = 0; i < ns; i++) {
                    if (nf + i == self) continue;
                    const float4* w = sph + (size_t)i * kSphWords;
                    v.sphere(w[0], nf + i, w[1].x);
                    if (v.opaque && !count_all) break;
                }
            if (v.opaque) { mask = clamp01(fmul(0.0f, mask)); return v.hits; }
            for (int i = 0; i < v.n; i++) mask = clamp01(fmul(v.kf[i], mask));
            if (!v.more) return v.hits;
            v.last_key = v.key[v.n - 1];
            v.counting = false;
        }
    }
};

using BvhAccel = BvhAccelT<false>;

template <bool kCount>
__host__ __device__ __forceinline__ BvhAccelT<kCount> make_bvh_accel(const SceneView& S)
{
    BvhAccelT<kCount> a;
    a.nodes = S.bvh.nodes; a.prims = S.bvh_tri; a.tri = S.tri; a.sph = S.sph;
    a.nodes8 = S.bvh.n_nodes8 > 0 ? S.bvh.nodes8 : nullptr; a.prims8 = S.bvh.prims8;
    a.nf = S.nf; a.ns = S.ns; a.root = S.bvh.root; a.n_nodes = S.bvh.n_nodes; a.count_all = false;
    a.overflow_flag = nullptr;
    return a;
}

}  // namespace rt
