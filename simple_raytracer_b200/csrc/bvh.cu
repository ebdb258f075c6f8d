// bvh.cu — on-device LBVH build: padded primitive bounds -> 30-bit Morton codes ->
// LSD radix sort (4 x 8 bits, hand-written) -> Karras hierarchy -> bottom-up refit
// -> traversal nodes with both child boxes + primitive records in Morton order.
//
// The reference has no acceleration structure (TraceRay is O(N) per ray,
// main.cpp:1215-1407); this one only decides WHICH primitives get the reference's
// exact tests.  Bounds are padded so that every hit the reference's float
// arithmetic accepts lies inside:
//   triangle: the accepted point o + ray*t passes the float barycentric test, i.e.
//             lies within rounding error of the triangle: pad = 1e-4 * extent +
//             8e-6 * (largest coordinate magnitude in play; this also covers the FFMA slab test);
//   sphere  : B*B - 4*C >= 0 in mixed precision with a ray that is unit only up to
//             rounding accepts rho^2 <= r^2 + c*eps*D^2 (D = distance from the ray
//             origin to the centre): r_eff^2 = r^2 + 1.2e-5 * Dmax^2
//             (rays with | |ray|^2 - 1 | > 1e-5 never use the tree for spheres, bvh.cuh).
#include "bvh.cuh"
#include "host_scene.h"
#include "smap_span.h"

#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace rt {

int tuning_int(const char* name, int dflt, int lo, int hi)
{
    const char* v = std::getenv(name);
    if (!v || !*v) return dflt;
    const int x = std::atoi(v);
    return x < lo ? lo : (x > hi ? hi : x);
}

namespace {

struct BuildBounds {
    float lo[3], hi[3];      // bounds of all primitives and the eye
    float scale;             // largest |coordinate|
};

RT_DEV unsigned expand_bits10(unsigned v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// one thread per primitive (order index): padded box + Morton code of its centre
__global__ void prim_bounds_kernel(const float4* __restrict__ tri, const float4* __restrict__ sph, int nf, int ns,
                                   BuildBounds bb, float4* __restrict__ box_lo, float4* __restrict__ box_hi,
                                   unsigned* __restrict__ keys, unsigned* __restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nf + ns) return;
    float lo[3], hi[3];
    if (i < nf) {
        const float4* w = tri + (size_t)i * kTriWords;
        const float4 v0 = w[1], e1 = w[2], e2 = w[3];
        const float p[3][3] = { { v0.x, v0.y, v0.z }, { v0.x + e1.x, v0.y + e1.y, v0.z + e1.z }, { v0.x + e2.x, v0.y + e2.y, v0.z + e2.z } };
        float ext = 0.0f;
        for (int c = 0; c < 3; c++) {
            lo[c] = fminf(p[0][c], fminf(p[1][c], p[2][c]));
            hi[c] = fmaxf(p[0][c], fmaxf(p[1][c], p[2][c]));
            ext = fmaxf(ext, hi[c] - lo[c]);
        }
        const float pad = 1.0e-4f * ext + 8.0e-6f * bb.scale;
        for (int c = 0; c < 3; c++) { lo[c] -= pad; hi[c] += pad; }
    } else {
        const float4 s = sph[(size_t)(i - nf) * kSphWords];
        const float c3[3] = { s.x, s.y, s.z };
        float d2 = 0.0f;
        for (int c = 0; c < 3; c++) {
            const float d = fmaxf(fabsf(c3[c] - bb.lo[c]), fabsf(c3[c] - bb.hi[c]));
            d2 += d * d;
        }
        const float r = fabsf(s.w);
        const float reff = sqrtf(r * r + 1.2e-5f * d2) * 1.00001f + 8.0e-6f * bb.scale;
        for (int c = 0; c < 3; c++) { lo[c] = c3[c] - reff; hi[c] = c3[c] + reff; }
    }
    // NaN / inf primitives (degenerate input) never hit anything: give them an empty box
    bool ok = true;
    for (int c = 0; c < 3; c++) ok = ok && isfinite(lo[c]) && isfinite(hi[c]);
    if (!ok) { for (int c = 0; c < 3; c++) { lo[c] = FLT_MAX; hi[c] = -FLT_MAX; } }
    box_lo[i] = make_float4(lo[0], lo[1], lo[2], 0.0f);
    box_hi[i] = make_float4(hi[0], hi[1], hi[2], 0.0f);
    unsigned code = 0;
    if (ok) {
        unsigned q[3];
        for (int c = 0; c < 3; c++) {
            const float extent = bb.hi[c] - bb.lo[c];
            float u = extent > 0.0f ? (0.5f * (lo[c] + hi[c]) - bb.lo[c]) / extent : 0.5f;
            u = fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
            q[c] = (unsigned)u;
        }
        code = (expand_bits10(q[0]) << 2) | (expand_bits10(q[1]) << 1) | expand_bits10(q[2]);
    }
    keys[i] = code;
    vals[i] = (unsigned)i;
}

// ---- LSD radix sort, 8 bits per pass -------------------------------------------
constexpr int kSortThreads = 256;
constexpr int kSortRounds = 8;                       // items per thread
constexpr int kSortTile = kSortThreads * kSortRounds;

__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(const unsigned* __restrict__ keys, int n, int shift,
                                                                  unsigned* __restrict__ hist, int nblocks)
{
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * kSortTile;
    for (int r = 0; r < kSortRounds; r++) {
        const int i = base + r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];   // digit-major: a scan gives global offsets
}

// Exclusive scan of the digit-major histogram in two small steps (a single block over all
// 256 * nblocks counters took 0.12 ms per pass for 1 M keys): one block per digit scans its row
// of per-block counts in place and leaves the row total; one block then scans the 256 totals.
// The scatter kernel adds the two.
__global__ void __launch_bounds__(256) scan_rows_kernel(unsigned* __restrict__ hist, int nblocks, unsigned* __restrict__ totals)
{
    __shared__ unsigned warp_sums[8];
    __shared__ unsigned carry;
    unsigned* row = hist + (size_t)blockIdx.x * nblocks;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 256) {
        const int i = base + threadIdx.x;
        const unsigned v = i < nblocks ? row[i] : 0u;
        unsigned x = v;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if ((threadIdx.x & 31) >= o) x += y;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = x;
        __syncthreads();
        unsigned before = 0;
        for (int w = 0; w < (int)(threadIdx.x >> 5); w++) before += warp_sums[w];
        const unsigned c = carry;
        if (i < nblocks) row[i] = c + before + x - v;
        __syncthreads();
        if (threadIdx.x == 255) carry = c + before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

__global__ void __launch_bounds__(256) scan_totals_kernel(unsigned* __restrict__ totals)
{
    __shared__ unsigned warp_sums[8];
    const unsigned v = totals[threadIdx.x];
    unsigned x = v;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
        if ((threadIdx.x & 31) >= o) x += y;
    }
    if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = x;
    __syncthreads();
    unsigned before = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) before += warp_sums[w];
    totals[threadIdx.x] = before + x - v;
}

// stable scatter: ranks inside a block come from warp match + per-warp digit counts
__global__ void __launch_bounds__(kSortThreads) radix_scatter_kernel(const unsigned* __restrict__ keys_in,
                                                                     const unsigned* __restrict__ vals_in, int n, int shift,
                                                                     const unsigned* __restrict__ offsets,
                                                                     const unsigned* __restrict__ digit_offsets, int nblocks,
                                                                     unsigned* __restrict__ keys_out, unsigned* __restrict__ vals_out)
{
    constexpr int kWarps = kSortThreads / 32;
    __shared__ unsigned digit_base[256];
    __shared__ unsigned warp_cnt[kWarps][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    digit_base[threadIdx.x] = offsets[threadIdx.x * nblocks + blockIdx.x] + digit_offsets[threadIdx.x];
    for (int w = 0; w < kWarps; w++) warp_cnt[w][threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * kSortTile;
    for (int r = 0; r < kSortRounds; r++) {
        const int i = base + r * kSortThreads + threadIdx.x;
        const bool valid = i < n;
        unsigned key = 0, val = 0, digit = 256;   // 256: invalid lanes match only each other
        if (valid) { key = keys_in[i]; val = vals_in[i]; digit = (key >> shift) & 255u; }
        const unsigned peers = __match_any_sync(0xffffffffu, digit);
        const unsigned rank = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank == 0) warp_cnt[warp][digit] = __popc(peers);
        __syncthreads();
        if (valid) {
            unsigned off = digit_base[digit] + rank;
            for (int w = 0; w < warp; w++) off += warp_cnt[w][digit];
            keys_out[off] = key; vals_out[off] = val;
        }
        __syncthreads();
        unsigned total = 0;
        for (int w = 0; w < kWarps; w++) { total += warp_cnt[w][threadIdx.x]; warp_cnt[w][threadIdx.x] = 0; }
        digit_base[threadIdx.x] += total;
        __syncthreads();
    }
}

// ---- Karras 2012 -------------------------------------------------------------------
RT_DEV int delta(const unsigned* keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const unsigned a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((unsigned)i ^ (unsigned)j);   // duplicate codes: fall back to the index
    return __clz(a ^ b);
}

// internal node i: children + covered range; child < 0 encodes leaf ~child
__global__ void hierarchy_kernel(const unsigned* __restrict__ keys, int n, int2* __restrict__ children,
                                 int2* __restrict__ range, int* __restrict__ parent_internal, int* __restrict__ parent_leaf)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? ~gamma : gamma;
    const int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(left, right);
    range[i] = make_int2(lo, hi);
    if (left < 0) parent_leaf[gamma] = i; else parent_internal[gamma] = i;
    if (right < 0) parent_leaf[gamma + 1] = i; else parent_internal[gamma + 1] = i;
    if (i == 0) parent_internal[0] = -1;
}

RT_DEV void box_union(float4& lo, float4& hi, const float4 a_lo, const float4 a_hi)
{
    lo.x = fminf(lo.x, a_lo.x); lo.y = fminf(lo.y, a_lo.y); lo.z = fminf(lo.z, a_lo.z);
    hi.x = fmaxf(hi.x, a_hi.x); hi.y = fmaxf(hi.y, a_hi.y); hi.z = fmaxf(hi.z, a_hi.z);
}

// one thread per leaf climbs; the second arrival at a node computes its box
__global__ void refit_kernel(int n, const unsigned* __restrict__ sorted_vals, const float4* __restrict__ prim_lo,
                             const float4* __restrict__ prim_hi, const int2* __restrict__ children,
                             const int* __restrict__ parent_internal, const int* __restrict__ parent_leaf,
                             float4* __restrict__ node_lo, float4* __restrict__ node_hi, unsigned* __restrict__ arrivals)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = parent_leaf[k];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&arrivals[node], 1u) == 0u) return;   // first child to arrive: the sibling finishes the node
        const int2 ch = children[node];
        float4 lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
        const int c2[2] = { ch.x, ch.y };
        for (int c = 0; c < 2; c++) {
            if (c2[c] < 0) {
                const unsigned p = sorted_vals[~c2[c]];
                box_union(lo, hi, prim_lo[p], prim_hi[p]);
            } else {
                box_union(lo, hi, __ldcg(&node_lo[c2[c]]), __ldcg(&node_hi[c2[c]]));
            }
        }
        node_lo[node] = lo; node_hi[node] = hi;
        node = parent_internal[node];
    }
}

// ---- 4-wide quantised traversal nodes (layout: bvh.cuh) ---------------------------------
// Writes one node from up to four (box, code) entries.  Per axis: quantum q = 2^e with e the
// smallest exponent that spans the node box in 250 steps (but never below 64 ulp of the
// coordinates, so that fp32 rounding in the decode stays far below one quantum); origin =
// node.lo - 2q; child lo -> floor(..) - 1, child hi -> ceil(..) + 1: one quantum of slack on
// both sides of every plane, which absorbs the <= 1/2 quantum rounding of the decode constant.
__device__ void write_node4(float4* out, int n_e, const float4* lo, const float4* hi, const int* code, int dummy_code)
{
    float nlo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, nhi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (int k = 0; k < n_e; k++) {
        const float l[3] = { lo[k].x, lo[k].y, lo[k].z }, h[3] = { hi[k].x, hi[k].y, hi[k].z };
        for (int c = 0; c < 3; c++) { nlo[c] = fminf(nlo[c], l[c]); nhi[c] = fmaxf(nhi[c], h[c]); }
    }
    float origin[3], q[3];
    unsigned eb[3];
    for (int c = 0; c < 3; c++) {
        float ext = nhi[c] - nlo[c];
        if (!(ext >= 0.0f) || !isfinite(ext)) ext = 0.0f;
        int e = -120;
        if (ext > 0.0f) { int k; frexpf(ext / 250.0f, &k); e = k; }          // ext / 250 < 2^k
        const float maxabs = fmaxf(fabsf(nlo[c]), fabsf(nhi[c]));
        if (maxabs > 0.0f && isfinite(maxabs)) e = max(e, ilogbf(maxabs) - 17);   // >= 64 ulp
        e = min(max(e, -120), 120);
        q[c] = ldexpf(1.0f, e);
        eb[c] = (unsigned)(e + 127);
        origin[c] = nlo[c] - 2.0f * q[c];
    }
    unsigned bytes[4][2];
    int codes[4];
    for (int k = 0; k < 4; k++) {
        if (k < n_e) {
            const float l[3] = { lo[k].x, lo[k].y, lo[k].z }, h[3] = { hi[k].x, hi[k].y, hi[k].z };
            unsigned ql[3], qh[3];
            for (int c = 0; c < 3; c++) {
                const float fl = floorf((l[c] - origin[c]) / q[c]) - 1.0f;
                const float fh = ceilf((h[c] - origin[c]) / q[c]) + 1.0f;
                // NaN / inf (degenerate primitives carry an inverted infinite box) clamp to the range ends
                ql[c] = (unsigned)fminf(fmaxf(fl, 0.0f), 255.0f);
                qh[c] = (unsigned)fminf(fmaxf(fh, 0.0f), 255.0f);
            }
            bytes[k][0] = ql[0] | (qh[0] << 8) | (ql[1] << 16) | (qh[1] << 24);
            bytes[k][1] = ql[2] | (qh[2] << 8);
            codes[k] = code[k];
        } else {
            bytes[k][0] = 0x00ff00ffu;      // lo = 255, hi = 0 on every axis: never hit ...
            bytes[k][1] = 0x000000ffu;
            codes[k] = dummy_code;          // ... and harmless if a degenerate ray gets through anyway
        }
    }
    uint4* w = reinterpret_cast<uint4*>(out);
    w[0] = make_uint4(__float_as_uint(origin[0]), __float_as_uint(origin[1]), __float_as_uint(origin[2]),
                      eb[0] | (eb[1] << 8) | (eb[2] << 16));
    w[1] = make_uint4((unsigned)codes[0], (unsigned)codes[1], (unsigned)codes[2], (unsigned)codes[3]);
    w[2] = make_uint4(bytes[0][0], bytes[0][1], bytes[1][0], bytes[1][1]);
    w[3] = make_uint4(bytes[2][0], bytes[2][1], bytes[3][0], bytes[3][1]);
}

// scenes of <= kLeafMax (4) primitives: one node with one leaf per primitive
__global__ void emit_tiny_kernel(int n, const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi,
                                 const unsigned* __restrict__ sorted_vals, float4* __restrict__ nodes)
{
    if (blockIdx.x || threadIdx.x) return;
    float4 lo[4], hi[4];
    int code[4];
    for (int k = 0; k < n; k++) {
        const unsigned p = sorted_vals[k];
        lo[k] = prim_lo[p]; hi[k] = prim_hi[p]; code[k] = ~(k << 3);
    }
    write_node4(nodes, n, lo, hi, code, ~(n << 3));
}

// primitive records in Morton order (spheres widened to the triangle record size)
__global__ void emit_prims_kernel(int n, int nf, const unsigned* __restrict__ sorted_vals, const float4* __restrict__ tri,
                                  const float4* __restrict__ sph, float4* __restrict__ out)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n) return;
    float4* o = out + (size_t)k * kTriWords;
    if (k == n) {   // the never-hit dummy record: a triangle with a zero normal, `dem == 0` rejects it (main.cpp:1297)
        o[0] = o[1] = o[2] = o[3] = make_float4(0.f, 0.f, 0.f, 0.f);
        o[4] = make_float4(0.f, 1.0f, __int_as_float(0x7fffffff), __int_as_float(0));
        return;
    }
    const int p = (int)sorted_vals[k];
    if (p < nf) {
        const float4* w = tri + (size_t)p * kTriWords;
        for (int q = 0; q < kTriWords; q++) o[q] = w[q];
    } else {
        const float4* w = sph + (size_t)(p - nf) * kSphWords;
        o[0] = w[0];
        o[1] = o[2] = o[3] = make_float4(0.f, 0.f, 0.f, 0.f);
        o[4] = make_float4(0.0f, w[1].x, w[1].y, __int_as_float(1));
    }
}


// ---- 8-wide compressed nodes for the any-hit walk (layout: bvh.cuh, Node8) ------------------
// Breadth-first collapse of the binary tree, one launch per level: a thread takes one binary
// node of the current level's queue, opens its children greedily by surface area until it
// holds 8 entries (or only primitives), reserves CONSECUTIVE slots for its inner entries in
// the next level's queue and for its primitives in the node-ordered primitive table (one
// atomicAdd each), so that a child is found by base + rank and no per-child index is stored.

__device__ float box_area(float4 lo, float4 hi)
{
    const float dx = fmaxf(hi.x - lo.x, 0.0f), dy = fmaxf(hi.y - lo.y, 0.0f), dz = fmaxf(hi.z - lo.z, 0.0f);
    return dx * dy + dy * dz + dz * dx;
}

// The 4-wide tree by greedy collapse: one thread per 4-wide node of the current level takes its
// binary node, opens the entry with the largest surface area until it holds four, writes the node
// at the binary node's own index and queues its inner entries for the next level.  (Fuller nodes
// and fewer levels than collapsing every other level of the binary tree: closest-hit kernels 21.3
// -> 21.1 ms on the 4K stress frame, round 1.)
// counts[l] = nodes of level l (level 0: the binary root, not stored): the host launches a fixed
// number of levels back to back without reading anything back.
constexpr int kCollapseLevels = 128;     // entries of the level-count arrays
__global__ void collapse4_kernel(int n, int level, unsigned* __restrict__ counts, const int* __restrict__ queue_cur,
                                 int* __restrict__ queue_next, const unsigned* __restrict__ sorted_vals,
                                 const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi,
                                 const int2* __restrict__ children, const float4* __restrict__ node_lo,
                                 const float4* __restrict__ node_hi, float4* __restrict__ nodes)
{
    const unsigned cur_count = level == 0 ? 1u : counts[level];
    unsigned* const next_count = counts + level + 1;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < cur_count; t += gridDim.x * blockDim.x) {
    const int root = level == 0 ? 0 : queue_cur[t];
    int ent[4];
    int n_e = 2;
    { const int2 ch = children[root]; ent[0] = ch.x; ent[1] = ch.y; }
    while (n_e < 4) {
        int best = -1; float best_area = -1.0f;
        for (int k = 0; k < n_e; k++) {
            if (ent[k] < 0) continue;
            const float a = box_area(node_lo[ent[k]], node_hi[ent[k]]);
            if (a > best_area || best < 0) { best_area = a; best = k; }
        }
        if (best < 0) break;
        const int2 ch = children[ent[best]];
        ent[best] = ch.x; ent[n_e++] = ch.y;
    }
    float4 lo[4], hi[4];
    int code[4];
    unsigned n_in = 0;
    for (int k = 0; k < n_e; k++) {
        if (ent[k] < 0) {
            const int slot = ~ent[k];
            const unsigned p = sorted_vals[slot];
            lo[k] = prim_lo[p]; hi[k] = prim_hi[p]; code[k] = ~(slot << 3);
        } else {
            lo[k] = node_lo[ent[k]]; hi[k] = node_hi[ent[k]]; code[k] = ent[k];
            n_in++;
        }
    }
    if (n_in) {
        unsigned at = atomicAdd(next_count, n_in);
        for (int k = 0; k < n_e; k++) if (ent[k] >= 0) queue_next[at++] = ent[k];
    }
    write_node4(nodes + (size_t)root * 4, n_e, lo, hi, code, ~(n << 3));
    }
}

// (level counts on the device as in collapse4_kernel; level l's nodes start at the sum of the earlier counts)
__global__ void collapse8_kernel(int n_prims, int nf, int level, unsigned* __restrict__ counts, unsigned* __restrict__ prim_count,
                                 const int* __restrict__ queue_cur, int* __restrict__ queue_next,
                                 const unsigned* __restrict__ sorted_vals,
                                 const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi,
                                 const int2* __restrict__ children, const float4* __restrict__ node_lo,
                                 const float4* __restrict__ node_hi, const float4* __restrict__ tri,
                                 const float4* __restrict__ sph, uint4* __restrict__ nodes8, float4* __restrict__ prims8)
{
    const unsigned cur_count = level == 0 ? 1u : counts[level];
    unsigned cur_offset = 0;
    for (int l = 0; l < level; l++) cur_offset += l == 0 ? 1u : counts[l];
    const unsigned next_offset = cur_offset + cur_count;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < cur_count; t += gridDim.x * blockDim.x) {
    const int root = level == 0 ? 0 : queue_cur[t];
    int ent[8];
    int n_e = 2;
    { const int2 ch = children[root]; ent[0] = ch.x; ent[1] = ch.y; }
    while (n_e < 8) {
        int best = -1; float best_area = -1.0f;
        for (int k = 0; k < n_e; k++) {
            if (ent[k] < 0) continue;
            const float a = box_area(node_lo[ent[k]], node_hi[ent[k]]);
            if (a > best_area || best < 0) { best_area = a; best = k; }
        }
        if (best < 0) break;
        const int2 ch = children[ent[best]];
        ent[best] = ch.x; ent[n_e++] = ch.y;
    }
    // slots: inner entries first, then the primitives (bvh.cuh: a child is base + slot, a primitive
    // prim_base + slot - n_inner)
    {
        int tmp[8], m = 0;
        for (int k = 0; k < n_e; k++) if (ent[k] >= 0) tmp[m++] = ent[k];
        for (int k = 0; k < n_e; k++) if (ent[k] < 0) tmp[m++] = ent[k];
        for (int k = 0; k < n_e; k++) ent[k] = tmp[k];
    }
    float4 lo[8], hi[8];
    float nlo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, nhi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    unsigned imask = 0, lmask = 0;
    for (int k = 0; k < n_e; k++) {
        if (ent[k] < 0) { const unsigned p = sorted_vals[~ent[k]]; lo[k] = prim_lo[p]; hi[k] = prim_hi[p]; lmask |= 1u << k; }
        else { lo[k] = node_lo[ent[k]]; hi[k] = node_hi[ent[k]]; imask |= 1u << k; }
        const float l[3] = { lo[k].x, lo[k].y, lo[k].z }, h[3] = { hi[k].x, hi[k].y, hi[k].z };
        for (int c = 0; c < 3; c++) { nlo[c] = fminf(nlo[c], l[c]); nhi[c] = fmaxf(nhi[c], h[c]); }
    }
    const unsigned n_in = (unsigned)__popc(imask), n_leaf = (unsigned)__popc(lmask);
    const unsigned child_base = next_offset + (n_in ? atomicAdd(counts + level + 1, n_in) : 0u);
    const unsigned prim_base = n_leaf ? atomicAdd(prim_count, n_leaf) : 0u;
    // the frame: as write_node4
    float origin[3], q[3];
    unsigned eb[3];
    for (int c = 0; c < 3; c++) {
        float ext = nhi[c] - nlo[c];
        if (!(ext >= 0.0f) || !isfinite(ext)) ext = 0.0f;
        int e = -120;
        if (ext > 0.0f) { int k; frexpf(ext / 250.0f, &k); e = k; }
        const float maxabs = fmaxf(fabsf(nlo[c]), fabsf(nhi[c]));
        if (maxabs > 0.0f && isfinite(maxabs)) e = max(e, ilogbf(maxabs) - 17);
        e = min(max(e, -120), 120);
        q[c] = ldexpf(1.0f, e);
        eb[c] = (unsigned)(e + 127);
        origin[c] = nlo[c] - 2.0f * q[c];
    }
    unsigned planes[3][4] = { { 0, 0, 0, 0 }, { 0, 0, 0, 0 }, { 0, 0, 0, 0 } };   // [axis][child pair]
    unsigned rank_in = 0, rank_leaf = 0;
    for (int k = 0; k < 8; k++) {
        unsigned ql[3] = { 255u, 255u, 255u }, qh[3] = { 0u, 0u, 0u };            // empty slot: never hit
        if (k < n_e) {
            const float l[3] = { lo[k].x, lo[k].y, lo[k].z }, h[3] = { hi[k].x, hi[k].y, hi[k].z };
            for (int c = 0; c < 3; c++) {
                const float fl = floorf((l[c] - origin[c]) / q[c]) - 1.0f;
                const float fh = ceilf((h[c] - origin[c]) / q[c]) + 1.0f;
                ql[c] = (unsigned)fminf(fmaxf(fl, 0.0f), 255.0f);
                qh[c] = (unsigned)fminf(fmaxf(fh, 0.0f), 255.0f);
            }
            if (ent[k] >= 0) {
                queue_next[child_base - next_offset + rank_in] = ent[k];
                rank_in++;
            } else {
                // the primitive's record, node-ordered (spheres widened to the triangle record size)
                const int p = (int)sorted_vals[~ent[k]];
                float4* o = prims8 + (size_t)(prim_base + rank_leaf) * kTriWords;
                if (p < nf) {
                    const float4* w = tri + (size_t)p * kTriWords;
                    for (int qd = 0; qd < kTriWords; qd++) o[qd] = w[qd];
                } else {
                    const float4* w = sph + (size_t)(p - nf) * kSphWords;
                    o[0] = w[0];
                    o[1] = o[2] = o[3] = make_float4(0.f, 0.f, 0.f, 0.f);
                    o[4] = make_float4(0.0f, w[1].x, w[1].y, __int_as_float(1));
                }
                rank_leaf++;
            }
        }
        for (int c = 0; c < 3; c++) planes[c][k >> 1] |= (ql[c] | (qh[c] << 8)) << ((k & 1) * 16);
    }
    uint4* w = nodes8 + (size_t)(cur_offset + t) * 6;
    w[0] = make_uint4(__float_as_uint(origin[0]), __float_as_uint(origin[1]), __float_as_uint(origin[2]),
                      eb[0] | (eb[1] << 8) | (eb[2] << 16) | (n_in << 24));
    w[1] = make_uint4(child_base, prim_base, lmask, imask);
    w[2] = make_uint4(planes[0][0], planes[0][1], planes[0][2], planes[0][3]);
    w[3] = make_uint4(planes[1][0], planes[1][1], planes[1][2], planes[1][3]);
    w[4] = make_uint4(planes[2][0], planes[2][1], planes[2][2], planes[2][3]);
    w[5] = make_uint4(0u, 0u, 0u, 0u);
    }
    (void)n_prims;
}

// ---- occluder maps (scene_device.h: ShadowMapView) ------------------------------------------------------------
//
// An id buffer seen from the light — orthographic for a directional light, the six 90-degree faces of a cube around
// a point light: one thread per primitive slot of the array the any-hit walk reads rasterises an OPAQUE primitive
// (shadow factor 1 - opacity == 0) over the cells whose centres it covers and keeps, per cell, the one nearest to
// the light (64-bit atomicMin on depth | slot).  Plain float arithmetic: the map is a guess that the exact test
// confirms or not, so a triangle that reaches behind a cube face's plane is simply left out of that face.
// Primitives covering more than kSmapThreadCells cells go to a list that a second kernel rasterises one block each.
struct SmapBuild {
    float dir[3], ub[3], vb[3];          // directional: light travel direction, basis of its plane
    float u0, v0, inv_cell, cell;
    float origin[3];                     // point: position
    int R;
    int cube;
};
constexpr int kSmapThreadCells = 1024;
// ... and those big ones share a budget of cell tests per map: a scene of many map-filling primitives (walls behind
// walls) would otherwise cost primitives x cells.  Over the budget, each of them is rasterised on every stride-th
// row and column only — holes in a guess cost nothing but guesses.
constexpr unsigned long long kSmapBigBudget = 1ull << 30;

RT_DEV unsigned long long smap_key(float depth, int slot)
{
    unsigned b = __float_as_uint(depth);
    b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);            // floats in unsigned order
    return ((unsigned long long)b << 32) | (unsigned)slot;
}

// A primitive as one view (the light's plane, or one cube face) sees it, in cell units (cell i has its centre at i + 0.5).
struct SmapPrim {
    bool ok, is_tri;
    float x[3], y[3], d[3];      // triangle: projected vertices, and what orders hits along a cell's ray (below)
    float c[3], r;               // sphere: centre in the view's frame (x, y, depth axis), radius, world units
    float pl[4];                 // point light, triangle: plane n . q = pl[3] in the face's frame (q relative to the light)
    int x0, y0, wx, wy;          // the cells whose centres its bounding box holds
};

// view: -1 = the directional light's plane, 0..5 = cube face (axis = view / 2, negative side if view & 1)
RT_DEV SmapPrim smap_setup(const float4* w, const SmapBuild& m, int view)
{
    SmapPrim q;
    q.ok = false;
    const float4 w4 = w[4];
    if (w4.y != 0.0f) return q;                                  // not opaque (or NaN)
    q.is_tri = __float_as_int(w4.w) == 0;
    const float R = (float)m.R;
    float lo_x, hi_x, lo_y, hi_y;
    if (view < 0) {
        if (q.is_tri) {
            const float4 v0 = w[1], e1 = w[2], e2 = w[3];
            const float p[3][3] = { { v0.x, v0.y, v0.z }, { v0.x + e1.x, v0.y + e1.y, v0.z + e1.z }, { v0.x + e2.x, v0.y + e2.y, v0.z + e2.z } };
            for (int k = 0; k < 3; k++) {
                q.x[k] = (p[k][0] * m.ub[0] + p[k][1] * m.ub[1] + p[k][2] * m.ub[2] - m.u0) * m.inv_cell;
                q.y[k] = (p[k][0] * m.vb[0] + p[k][1] * m.vb[1] + p[k][2] * m.vb[2] - m.v0) * m.inv_cell;
                q.d[k] = p[k][0] * m.dir[0] + p[k][1] * m.dir[1] + p[k][2] * m.dir[2];
            }
        } else {
            const float4 c = w[0];
            q.x[0] = (c.x * m.ub[0] + c.y * m.ub[1] + c.z * m.ub[2] - m.u0) * m.inv_cell;
            q.y[0] = (c.x * m.vb[0] + c.y * m.vb[1] + c.z * m.vb[2] - m.v0) * m.inv_cell;
            q.d[0] = c.x * m.dir[0] + c.y * m.dir[1] + c.z * m.dir[2];
            q.r = fabsf(c.w);
            const float rc = q.r * m.inv_cell;
            q.x[1] = q.x[0] - rc; q.x[2] = q.x[0] + rc; q.y[1] = q.y[0] - rc; q.y[2] = q.y[0] + rc;
        }
    } else {
        // the face's frame: z = the face's axis (pointing away from the light), x / y the other two axes; a point
        // (x, y, z) projects to (x / z, y / z) in [-1, 1]^2, cell coordinate (x / z + 1) * R / 2
        const int az = view >> 1, ax = (az + 1) % 3, ay = (az + 2) % 3;
        const float sgn = (view & 1) ? -1.0f : 1.0f;
        const float half = 0.5f * R;
        if (q.is_tri) {
            const float4 v0 = w[1], e1 = w[2], e2 = w[3];
            const float p[3][3] = { { v0.x, v0.y, v0.z }, { v0.x + e1.x, v0.y + e1.y, v0.z + e1.z }, { v0.x + e2.x, v0.y + e2.y, v0.z + e2.z } };
            float f[3][3];
            for (int k = 0; k < 3; k++) {
                f[k][0] = p[k][ax] - m.origin[ax]; f[k][1] = p[k][ay] - m.origin[ay]; f[k][2] = sgn * (p[k][az] - m.origin[az]);
                if (!(f[k][2] > 1.0e-6f * (fabsf(f[k][0]) + fabsf(f[k][1])) && f[k][2] > 0.0f)) return q;   // reaches behind the face's plane
                q.x[k] = (f[k][0] / f[k][2] + 1.0f) * half;
                q.y[k] = (f[k][1] / f[k][2] + 1.0f) * half;
            }
            // plane through the three points (light at the origin): a cell's ray (cx, cy, 1) * t meets it at
            // t = pl[3] / (n . (cx, cy, 1)); t orders the hits along that ray
            const float a[3] = { f[1][0] - f[0][0], f[1][1] - f[0][1], f[1][2] - f[0][2] };
            const float b[3] = { f[2][0] - f[0][0], f[2][1] - f[0][1], f[2][2] - f[0][2] };
            q.pl[0] = a[1] * b[2] - a[2] * b[1]; q.pl[1] = a[2] * b[0] - a[0] * b[2]; q.pl[2] = a[0] * b[1] - a[1] * b[0];
            q.pl[3] = q.pl[0] * f[0][0] + q.pl[1] * f[0][1] + q.pl[2] * f[0][2];
        } else {
            const float4 c = w[0];
            const float cc[3] = { c.x, c.y, c.z };
            q.c[0] = cc[ax] - m.origin[ax]; q.c[1] = cc[ay] - m.origin[ay]; q.c[2] = sgn * (cc[az] - m.origin[az]);
            q.r = fabsf(c.w);
            const float zn = q.c[2] - q.r, zf = q.c[2] + q.r;
            if (!(zn > 1.0e-6f * (fabsf(q.c[0]) + fabsf(q.c[1]) + q.r) && zn > 0.0f)) return q;            // too close to the face's plane
            // bounds of x / z over the sphere's bounding box (z in [zn, zf], both positive)
            const float xl = q.c[0] - q.r, xh = q.c[0] + q.r, yl = q.c[1] - q.r, yh = q.c[1] + q.r;
            q.x[1] = (fminf(xl / zn, xl / zf) + 1.0f) * half; q.x[2] = (fmaxf(xh / zn, xh / zf) + 1.0f) * half;
            q.y[1] = (fminf(yl / zn, yl / zf) + 1.0f) * half; q.y[2] = (fmaxf(yh / zn, yh / zf) + 1.0f) * half;
        }
    }
    if (q.is_tri) {
        lo_x = fminf(q.x[0], fminf(q.x[1], q.x[2])); hi_x = fmaxf(q.x[0], fmaxf(q.x[1], q.x[2]));
        lo_y = fminf(q.y[0], fminf(q.y[1], q.y[2])); hi_y = fmaxf(q.y[0], fmaxf(q.y[1], q.y[2]));
    } else {
        lo_x = q.x[1]; hi_x = q.x[2]; lo_y = q.y[1]; hi_y = q.y[2];
    }
    if (!(isfinite(lo_x) && isfinite(hi_x) && isfinite(lo_y) && isfinite(hi_y))) return q;
    const int x0 = (int)fmaxf(ceilf(lo_x - 0.5f), 0.0f), x1 = (int)fminf(floorf(hi_x - 0.5f), R - 1.0f);
    const int y0 = (int)fmaxf(ceilf(lo_y - 0.5f), 0.0f), y1 = (int)fminf(floorf(hi_y - 0.5f), R - 1.0f);
    if (x1 < x0 || y1 < y0) return q;
    q.x0 = x0; q.y0 = y0; q.wx = x1 - x0 + 1; q.wy = y1 - y0 + 1;
    q.ok = true;
    return q;
}

// (step: every step-th row and column of the primitive's cells, 1 = all of them)
RT_DEV void smap_raster(const SmapPrim& q, int slot, const SmapBuild& m, int view, unsigned long long* cells, int start, int stride,
                        int step = 1)
{
    const int nx = (q.wx + step - 1) / step, ny = (q.wy + step - 1) / step;
    const int total = nx * ny;
    float den = 0.0f;
    if (q.is_tri) {
        den = (q.x[1] - q.x[0]) * (q.y[2] - q.y[0]) - (q.x[2] - q.x[0]) * (q.y[1] - q.y[0]);
        if (!(fabsf(den) > 1.0e-12f)) return;                   // edge-on
    }
    unsigned long long* face = cells + (view < 0 ? 0 : (size_t)view * m.R * m.R);
    const float to_dir = 2.0f / (float)m.R;
    for (int i = start; i < total; i += stride) {
        const int ix = q.x0 + (i % nx) * step, iy = q.y0 + (i / nx) * step;
        const float cx = (float)ix + 0.5f, cy = (float)iy + 0.5f;
        float depth;
        if (q.is_tri) {
            // (a projective map keeps lines straight: inside the projected triangle <=> the cell's ray hits it)
            const float b1 = ((cx - q.x[0]) * (q.y[2] - q.y[0]) - (q.x[2] - q.x[0]) * (cy - q.y[0])) / den;
            const float b2 = ((q.x[1] - q.x[0]) * (cy - q.y[0]) - (cx - q.x[0]) * (q.y[1] - q.y[0])) / den;
            const float b0 = 1.0f - b1 - b2;
            if (!(b0 >= 0.0f && b1 >= 0.0f && b2 >= 0.0f)) continue;
            if (view < 0) depth = b0 * q.d[0] + b1 * q.d[1] + b2 * q.d[2];
            else depth = q.pl[3] / (q.pl[0] * (cx * to_dir - 1.0f) + q.pl[1] * (cy * to_dir - 1.0f) + q.pl[2]);
        } else if (view < 0) {
            const float dx = (cx - q.x[0]) * m.cell, dy = (cy - q.y[0]) * m.cell;
            const float h2 = q.r * q.r - (dx * dx + dy * dy);
            if (!(h2 >= 0.0f)) continue;
            depth = q.d[0] - sqrtf(h2);
        } else {
            // ray (dx, dy, 1) * t against the sphere: the nearer root
            const float dx = cx * to_dir - 1.0f, dy = cy * to_dir - 1.0f;
            const float dd = dx * dx + dy * dy + 1.0f;
            const float bq = dx * q.c[0] + dy * q.c[1] + q.c[2];
            const float cq = q.c[0] * q.c[0] + q.c[1] * q.c[1] + q.c[2] * q.c[2] - q.r * q.r;
            const float disc = bq * bq - dd * cq;
            if (!(disc >= 0.0f)) continue;
            depth = (bq - sqrtf(disc)) / dd;
        }
        if (!(isfinite(depth)) || (view >= 0 && !(depth > 0.0f))) continue;
        // (most primitives of a dense scene lie behind what the cell already holds: look before the atomic)
        unsigned long long* cell = face + (size_t)iy * m.R + ix;
        const unsigned long long key = smap_key(depth, slot);
        if (key < *(volatile unsigned long long*)cell) atomicMin(cell, key);
    }
}

// A triangle by one thread, row by row: the span of cells inside it (smap_span.h — checked against the per-cell test
// of smap_raster on the CPU, tests/test_smap_span.py) instead of a test per cell of the bounding box, the barycentrics
// stepped along the row: half the cells of a bounding box lie outside the triangle, and the index arithmetic of the
// generic loop costs more than its test (1 M triangles into 4096 x 4096 cells: 1.9 -> 1.25 ms).
RT_DEV void smap_raster_rows(const SmapPrim& q, int slot, const SmapBuild& m, int view, unsigned long long* cells)
{
    const float den = (q.x[1] - q.x[0]) * (q.y[2] - q.y[0]) - (q.x[2] - q.x[0]) * (q.y[1] - q.y[0]);
    if (!(fabsf(den) > 1.0e-12f)) return;                       // edge-on
    const float inv_den = 1.0f / den;
    unsigned long long* face = cells + (view < 0 ? 0 : (size_t)view * m.R * m.R);
    const float to_dir = 2.0f / (float)m.R;
    for (int iy = q.y0; iy < q.y0 + q.wy; iy++) {
        const float cy = (float)iy + 0.5f;
        const SmapRow row = smap_row(q.x, q.y, inv_den, cy);
        int first, last;
        if (!smap_span(row, q.x0, q.wx, first, last)) continue;
        unsigned long long* cell = face + (size_t)iy * m.R + first;
        float cx = (float)first + 0.5f;
        for (int ix = first; ix <= last; ix++, cell++, cx += 1.0f) {
            float depth;
            if (view < 0) {
                const float b1 = row.a1 * cx + row.c1, b2 = row.a2 * cx + row.c2;
                depth = (1.0f - b1 - b2) * q.d[0] + b1 * q.d[1] + b2 * q.d[2];
            } else {
                depth = q.pl[3] / (q.pl[0] * (cx * to_dir - 1.0f) + q.pl[1] * (cy * to_dir - 1.0f) + q.pl[2]);
                if (!(depth > 0.0f)) continue;
            }
            if (!isfinite(depth)) continue;
            const unsigned long long key = smap_key(depth, slot);
            if (key < *(volatile unsigned long long*)cell) atomicMin(cell, key);
        }
    }
}

// n_big[0] = entries of `big`, n_big[1..2] = the cells their bounding boxes hold (64 bits)
__global__ void smap_prims_kernel(const float4* __restrict__ prims, int n, SmapBuild m, unsigned long long* __restrict__ cells,
                                  int* __restrict__ big, unsigned* __restrict__ n_big)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    for (int view = m.cube ? 0 : -1; view < (m.cube ? 6 : 0); view++) {
        const SmapPrim q = smap_setup(prims + (size_t)slot * kTriWords, m, view);
        if (!q.ok) continue;
        if ((long long)q.wx * q.wy > kSmapThreadCells) {
            big[atomicAdd(n_big, 1u)] = slot * 8 + (view + 1);
            atomicAdd(reinterpret_cast<unsigned long long*>(n_big + 2), (unsigned long long)q.wx * (unsigned long long)q.wy);
            continue;
        }
        if (q.is_tri) smap_raster_rows(q, slot, m, view, cells);
        else smap_raster(q, slot, m, view, cells, 0, 1);
    }
}

__global__ void smap_big_kernel(const float4* __restrict__ prims, SmapBuild m, unsigned long long* __restrict__ cells,
                                const int* __restrict__ big, const unsigned* __restrict__ n_big)
{
    const unsigned n = *n_big;
    const unsigned long long work = *reinterpret_cast<const unsigned long long*>(n_big + 2);
    int step = 1;
    while ((work / ((unsigned long long)step * step)) > kSmapBigBudget) step *= 2;
    for (unsigned b = blockIdx.x; b < n; b += gridDim.x) {
        const int slot = big[b] >> 3, view = (big[b] & 7) - 1;
        const SmapPrim q = smap_setup(prims + (size_t)slot * kTriWords, m, view);
        if (q.ok) smap_raster(q, slot, m, view, cells, (int)threadIdx.x, (int)blockDim.x, step);
    }
}

__global__ void smap_finish_kernel(const unsigned long long* __restrict__ cells, size_t n_cells, int* __restrict__ out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cells) return;
    const unsigned long long k = cells[i];
    out[i] = k == ~0ull ? -1 : (int)(unsigned)(k & 0xffffffffull);
}

// Maps for the first kMaxShadowMaps lights, over the primitive array the any-hit walk reads (prims: n records of
// kTriWords words).  Lights whose direction / position is not finite, and scenes without finite bounds, get none.
static int build_shadow_maps(SceneHost* s, const float4* prims, int n, const BuildBounds& bb, cudaStream_t st)
{
    for (int k = 0; k < kMaxShadowMaps; k++) s->smap[k] = SceneHost::SmapHost();
    const int nl = std::min(s->nl, kMaxShadowMaps);
    if (nl <= 0 || n <= 0 || n >= (1 << 28)) return RT_OK;
    for (int c = 0; c < 3; c++) if (!std::isfinite(bb.lo[c]) || !std::isfinite(bb.hi[c])) return RT_OK;
    float4 lights[kMaxShadowMaps * 3];
    RT_CUDA(cudaMemcpyAsync(lights, s->light.p, (size_t)nl * 3 * sizeof(float4), cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    // four times as many cells per axis as the square root of the primitive count: cells a good deal smaller than
    // the typical primitive, so that a point in a cell is usually behind the primitive that covers the cell's centre.
    // Measured on the 1 M-triangle stress scene at 4K (tools/r2_job18.sh): shadow rays left to walk 7.7 / 4.9 / 3.3 /
    // 2.4 M of 23 M with 1024 / 2048 / 4096 / 8192 cells per axis, frame 31.1 / 29.4 / 28.4 / 27.9 ms; 8192 costs
    // 12 ms more to build (four maps of 268 MB), 4096 nothing measurable.  A cube face gets half of that per axis.
    int R = 64;
    double per_sqrt = 4.0;
    int r_max = 4096;
#ifdef RT_TUNING_ENV
    if (const char* e = std::getenv("RT_SMAP_PER_SQRT")) per_sqrt = std::atof(e);
    if (const char* e = std::getenv("RT_SMAP_MAX_RES")) r_max = std::atoi(e);
#endif
    while (R < r_max && (double)R < per_sqrt * std::sqrt((double)n)) R *= 2;
    // Light 0 decides most chains (94 % of the stress scene's end there): the full resolution for it, a quarter per
    // axis for the others — a map of 4096 x 4096 cells takes 1.4 ms to build over 1 M triangles, which a map that is
    // asked for one node in sixteen does not earn back within a frame.  A cube face gets half of its light's figure.
    struct Want { int light; bool cube; int R; size_t cells; size_t offset; };
    Want want[kMaxShadowMaps];
    int n_maps = 0;
    size_t total = 0, largest = 0;
    for (int k = 0; k < nl; k++) {
        const float4 l0 = lights[3 * k];
        if (!std::isfinite(l0.x) || !std::isfinite(l0.y) || !std::isfinite(l0.z)) continue;
        const bool cube = l0.w != 0.0f;
        if (!cube && !((double)l0.x * l0.x + (double)l0.y * l0.y + (double)l0.z * l0.z > 1.0e-60)) continue;
        const int Rk = std::max(64, k == 0 ? R : R / 4);
        const int Rm = cube ? std::max(32, Rk / 2) : Rk;
        const size_t cells = cube ? (size_t)6 * Rm * Rm : (size_t)Rm * Rm;
        want[n_maps++] = Want{ k, cube, Rm, cells, total };
        total += cells;
        largest = std::max(largest, cells);
    }
    if (n_maps == 0) return RT_OK;
    // (building the maps on a stream of their own while the first frame's closest-hit kernels run was tried: the two
    // share the SMs and the frame ends no sooner)
    DevBuf keys, big, n_big;
    if (!s->smap_cells.reserve(total * sizeof(int)) || !keys.reserve(largest * 8) || !big.reserve((size_t)n * 6 * 4) || !n_big.reserve(16))
        return RT_E_NOMEM;
    for (int j = 0; j < n_maps; j++) {
        const int k = want[j].light;
        const float4 l0 = lights[3 * k];
        SmapBuild m;
        std::memset(&m, 0, sizeof m);
        m.cube = want[j].cube ? 1 : 0;
        if (m.cube) {
            m.origin[0] = l0.x; m.origin[1] = l0.y; m.origin[2] = l0.z;
            m.R = want[j].R;
        } else {
            const double len = std::sqrt((double)l0.x * l0.x + (double)l0.y * l0.y + (double)l0.z * l0.z);
            const double d[3] = { l0.x / len, l0.y / len, l0.z / len };
            // an axis the direction is least aligned with, then an orthonormal pair
            int a = 0;
            if (std::fabs(d[1]) < std::fabs(d[a])) a = 1;
            if (std::fabs(d[2]) < std::fabs(d[a])) a = 2;
            double ax[3] = { 0, 0, 0 }; ax[a] = 1.0;
            double u[3] = { d[1] * ax[2] - d[2] * ax[1], d[2] * ax[0] - d[0] * ax[2], d[0] * ax[1] - d[1] * ax[0] };
            const double ul = std::sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
            for (int c = 0; c < 3; c++) u[c] /= ul;
            const double v[3] = { d[1] * u[2] - d[2] * u[1], d[2] * u[0] - d[0] * u[2], d[0] * u[1] - d[1] * u[0] };
            double umin = 1e300, umax = -1e300, vmin = 1e300, vmax = -1e300;
            for (int corner = 0; corner < 8; corner++) {
                double pu = 0, pv = 0;
                for (int c = 0; c < 3; c++) {
                    const double x = (corner >> c) & 1 ? bb.hi[c] : bb.lo[c];
                    pu += x * u[c]; pv += x * v[c];
                }
                umin = std::min(umin, pu); umax = std::max(umax, pu); vmin = std::min(vmin, pv); vmax = std::max(vmax, pv);
            }
            const double extent = std::max(umax - umin, vmax - vmin) * 1.0001 + 1.0e-20;
            if (!std::isfinite(extent) || !(extent > 0.0)) continue;
            for (int c = 0; c < 3; c++) { m.dir[c] = (float)d[c]; m.ub[c] = (float)u[c]; m.vb[c] = (float)v[c]; }
            const int Rm = want[j].R;
            m.cell = (float)(extent / Rm); m.inv_cell = (float)(Rm / extent);
            m.u0 = (float)(0.5 * (umin + umax) - 0.5 * extent); m.v0 = (float)(0.5 * (vmin + vmax) - 0.5 * extent);
            m.R = Rm;
            if (!std::isfinite(m.inv_cell) || !std::isfinite(m.u0) || !std::isfinite(m.v0)) continue;
        }
        RT_CUDA(cudaMemsetAsync(keys.p, 0xff, want[j].cells * 8, st));
        RT_CUDA(cudaMemsetAsync(n_big.p, 0, 16, st));
        smap_prims_kernel<<<(n + 255) / 256, 256, 0, st>>>(prims, n, m, keys.as<unsigned long long>(), big.as<int>(), n_big.as<unsigned>());
        smap_big_kernel<<<1024, 256, 0, st>>>(prims, m, keys.as<unsigned long long>(), big.as<int>(), n_big.as<unsigned>());
        smap_finish_kernel<<<(unsigned)((want[j].cells + 255) / 256), 256, 0, st>>>(keys.as<unsigned long long>(), want[j].cells,
                                                                                   s->smap_cells.as<int>() + want[j].offset);
        RT_CUDA(cudaGetLastError());
        SceneHost::SmapHost& h = s->smap[k];
        h.offset_plus_1 = want[j].offset + 1;
        for (int c = 0; c < 3; c++) { h.ub[c] = m.ub[c]; h.vb[c] = m.vb[c]; h.origin[c] = m.origin[c]; }
        h.u0 = m.u0; h.v0 = m.v0; h.inv_cell = m.inv_cell; h.R = m.R; h.cube = m.cube;
    }
    RT_CUDA(cudaStreamSynchronize(st));      // (the temporaries go out of scope)
    return RT_OK;
}

}  // namespace

int build_bvh(SceneHost* s)
{
    const int n = s->nf + s->ns;
    if (n <= 0) { s->bvh_built = true; s->bvh_n_nodes = 0; return RT_OK; }
    cudaStream_t st = s->stream;
    BuildBounds bb;
    float scale = 0.0f;
    for (int c = 0; c < 3; c++) {
        bb.lo[c] = std::fmin(s->scene_lo[c], s->cam.eye[c]);
        bb.hi[c] = std::fmax(s->scene_hi[c], s->cam.eye[c]);
        scale = std::fmax(scale, std::fmax(std::fabs(bb.lo[c]), std::fabs(bb.hi[c])));
    }
    bb.scale = scale;

    const int n_internal = n > kLeafMax ? n - 1 : 1;
    const int nblocks = (n + kSortTile - 1) / kSortTile;
    DevBuf box_lo, box_hi, keys_a, keys_b, vals_a, vals_b, hist, children, range, par_i, par_l, node_lo, node_hi, arrivals;
    const size_t N = (size_t)n;
    if (!box_lo.reserve(N * 16) || !box_hi.reserve(N * 16) || !keys_a.reserve(N * 4) || !keys_b.reserve(N * 4) ||
        !vals_a.reserve(N * 4) || !vals_b.reserve(N * 4) || !hist.reserve((size_t)256 * (nblocks + 1) * 4) ||
        !children.reserve(N * 8) || !range.reserve(N * 8) || !par_i.reserve(N * 4) || !par_l.reserve(N * 4) ||
        !node_lo.reserve(N * 16) || !node_hi.reserve(N * 16) || !arrivals.reserve(N * 4))
        return RT_E_NOMEM;
    if (!s->bvh_nodes.reserve((size_t)n_internal * 64) || !s->bvh_tri.reserve((N + 1) * kTriWords * 16)) return RT_E_NOMEM;

    const int B = 256, G = (n + B - 1) / B;
    prim_bounds_kernel<<<G, B, 0, st>>>(s->tri.as<float4>(), s->sph.as<float4>(), s->nf, s->ns, bb, box_lo.as<float4>(),
                                        box_hi.as<float4>(), keys_a.as<unsigned>(), vals_a.as<unsigned>());
    unsigned *kin = keys_a.as<unsigned>(), *kout = keys_b.as<unsigned>(), *vin = vals_a.as<unsigned>(), *vout = vals_b.as<unsigned>();
    for (int shift = 0; shift < 32; shift += 8) {
        radix_hist_kernel<<<nblocks, kSortThreads, 0, st>>>(kin, n, shift, hist.as<unsigned>(), nblocks);
        scan_rows_kernel<<<256, 256, 0, st>>>(hist.as<unsigned>(), nblocks, hist.as<unsigned>() + (size_t)256 * nblocks);
        scan_totals_kernel<<<1, 256, 0, st>>>(hist.as<unsigned>() + (size_t)256 * nblocks);
        radix_scatter_kernel<<<nblocks, kSortThreads, 0, st>>>(kin, vin, n, shift, hist.as<unsigned>(),
                                                               hist.as<unsigned>() + (size_t)256 * nblocks, nblocks, kout, vout);
        std::swap(kin, kout); std::swap(vin, vout);
    }
    // 4 passes: the sorted data is back in keys_a / vals_a
    int levels4 = 1, levels8 = 0;
    if (n > kLeafMax) {
        RT_CUDA(cudaMemsetAsync(arrivals.p, 0, N * 4, st));
        hierarchy_kernel<<<G, B, 0, st>>>(kin, n, children.as<int2>(), range.as<int2>(), par_i.as<int>(), par_l.as<int>());
        refit_kernel<<<G, B, 0, st>>>(n, vin, box_lo.as<float4>(), box_hi.as<float4>(), children.as<int2>(), par_i.as<int>(),
                                      par_l.as<int>(), node_lo.as<float4>(), node_hi.as<float4>(), arrivals.as<unsigned>());
        {
            DevBuf queue_a, queue_b, counts;
            PinnedBuf h_counts;
            const size_t cbytes = (size_t)kCollapseLevels * sizeof(unsigned);
            if (!queue_a.reserve(N * 4) || !queue_b.reserve(N * 4) || !counts.reserve(cbytes) || !h_counts.reserve(cbytes)) return RT_E_NOMEM;
            RT_CUDA(cudaMemsetAsync(counts.p, 0, cbytes, st));
            int* qc = queue_a.as<int>(); int* qn = queue_b.as<int>();
            const int cgrid = std::min((n + B - 1) / B, 148 * 8);
            const unsigned* hc = static_cast<const unsigned*>(h_counts.p);
            // a batch of levels without a word from the device, then a look at where it stands
            int level = 0;
            for (int batch = 24;; batch = 8) {
                for (int k = 0; k < batch && level < kCollapseLevels - 1; k++, level++) {
                    collapse4_kernel<<<cgrid, B, 0, st>>>(n, level, counts.as<unsigned>(), qc, qn, vin, box_lo.as<float4>(),
                                                         box_hi.as<float4>(), children.as<int2>(), node_lo.as<float4>(),
                                                         node_hi.as<float4>(), s->bvh_nodes.as<float4>());
                    std::swap(qc, qn);
                }
                RT_CUDA(cudaGetLastError());
                RT_CUDA(cudaMemcpyAsync(h_counts.p, counts.p, cbytes, cudaMemcpyDeviceToHost, st));
                RT_CUDA(cudaStreamSynchronize(st));
                if (hc[level] == 0) break;
                if (level >= kCollapseLevels - 1) { set_error("collapse4: tree deeper than 127 levels"); return RT_E_LIMIT; }
            }
            // levels of the tree = the root's + those that received nodes
            levels4 = 1;
            while (levels4 < level && hc[levels4] != 0) levels4++;
        }
    } else {
        emit_tiny_kernel<<<1, 32, 0, st>>>(n, box_lo.as<float4>(), box_hi.as<float4>(), vin, s->bvh_nodes.as<float4>());
    }
    emit_prims_kernel<<<(n + 1 + B - 1) / B, B, 0, st>>>(n, s->nf, vin, s->tri.as<float4>(), s->sph.as<float4>(), s->bvh_tri.as<float4>());
    RT_CUDA(cudaGetLastError());
    // the 8-wide tree of the any-hit walk (scenes with a binary tree worth collapsing)
    s->bvh8_n_nodes = 0;
    if (n > 8) {
        DevBuf queue_a, queue_b, counts;
        PinnedBuf h_counts;
        const size_t cbytes = (size_t)(kCollapseLevels + 1) * sizeof(unsigned);      // [kCollapseLevels] = primitives placed
        if (!queue_a.reserve(N * 4) || !queue_b.reserve(N * 4) || !counts.reserve(cbytes) || !h_counts.reserve(cbytes) ||
            !s->bvh8_nodes.reserve((size_t)n_internal * 96) || !s->bvh8_prims.reserve((N + 1) * kTriWords * 16))
            return RT_E_NOMEM;
        RT_CUDA(cudaMemsetAsync(counts.p, 0, cbytes, st));
        int* qc = queue_a.as<int>(); int* qn = queue_b.as<int>();
        const int cgrid = std::min((n + B - 1) / B, 148 * 8);
        const unsigned* hc = static_cast<const unsigned*>(h_counts.p);
        int level = 0;
        for (int batch = 16;; batch = 8) {
            for (int k = 0; k < batch && level < kCollapseLevels - 1; k++, level++) {
                collapse8_kernel<<<cgrid, B, 0, st>>>(n, s->nf, level, counts.as<unsigned>(), counts.as<unsigned>() + kCollapseLevels,
                                                     qc, qn, vin, box_lo.as<float4>(), box_hi.as<float4>(), children.as<int2>(),
                                                     node_lo.as<float4>(), node_hi.as<float4>(), s->tri.as<float4>(),
                                                     s->sph.as<float4>(), s->bvh8_nodes.as<uint4>(), s->bvh8_prims.as<float4>());
                std::swap(qc, qn);
            }
            RT_CUDA(cudaGetLastError());
            RT_CUDA(cudaMemcpyAsync(h_counts.p, counts.p, cbytes, cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaStreamSynchronize(st));
            if (hc[level] == 0) break;
            if (level >= kCollapseLevels - 1) { set_error("collapse8: tree deeper than 127 levels"); return RT_E_LIMIT; }
        }
        levels8 = 1;
        while (levels8 < level && hc[levels8] != 0) levels8++;
        if (hc[kCollapseLevels] != (unsigned)n) { set_error("collapse8: primitive count mismatch"); return RT_E_CUDA; }
        unsigned total = 1;
        for (int l = 1; l < level; l++) total += hc[l];
        s->bvh8_n_nodes = (int)total;
    }
    RT_CUDA(cudaStreamSynchronize(st));
    // The walks' stacks hold kSmemStack + kBvhStack entries (and the parity probes' kBvhStack): a tree
    // that could need more is refused here rather than walked with a chance of a dropped subtree.
    // (Karras trees over 30-bit Morton codes + index tie-breaks are at most ~50 binary levels deep;
    // the stress scene's trees have about a dozen levels.)
    s->bvh_levels4 = levels4; s->bvh_levels8 = levels8;
    if (kStackNeed4(levels4) > kBvhStack || kStackNeed8(levels8) > kSmemStack + kBvhStack) {
        set_error("LBVH too deep for the traversal stacks (" + std::to_string(levels4) + " / " + std::to_string(levels8) +
                  " levels): degenerate primitive distribution");
        return RT_E_LIMIT;
    }
    s->bvh_n_nodes = n_internal; s->bvh_root = 0; s->bvh_built = true;
    // the any-hit walk reads the 8-wide tree's primitive records when there is one (bvh.cuh: make_bvh_accel)
    return build_shadow_maps(s, s->bvh8_n_nodes > 0 ? s->bvh8_prims.as<float4>() : s->bvh_tri.as<float4>(), n, bb, st);
}

}  // namespace rt
