// scene_device.h — how a scene lives in HBM, and the wavefront work buffers.
//
// The reference keeps a pointer-linked, std::map-indexed AoS scene
// (src/definitions.h:249-311).  Here every table is a flat array of 16-byte
// words.  Hot intersection data and cold shading attributes are split into
// separate arrays; inside the intersection tables one primitive's words are
// contiguous (80 B per triangle, 32 B per sphere) because a BVH leaf visit is a
// single thread reading one whole primitive: contiguity makes that 3 sectors
// instead of 5 cache lines.  In the brute-force path all lanes of a warp read the
// same primitive at the same time (a broadcast), from a shared-memory copy.
//
// Objects are named by their position in the reference's TraceRay iteration
// order ("order"): faces 0..nf-1 in file order, then spheres nf..nf+ns-1
// (std::map key order "face" < "sphere", main.cpp:1218).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rt {

constexpr int kTriWords = 5;     // float4 words per triangle intersection record
constexpr int kSphWords = 2;     // float4 words per sphere intersection record
constexpr int kFaceAttrWords = 4;
constexpr int kMaxLevels = 8;    // generations of ShadeRay: max_depth + 1 <= 8
constexpr int kMaxStack = 8;     // medium stack entries carried per ray

// Triangle intersection record (ray-independent constants hoisted, bit-identical
// to recomputing them per ray: main.cpp:1280-1282, 1301, 1361-1366):
//   w0 = surface_normal.xyz, D            D   = -(n . v0)
//   w1 = v0.xyz,             det          det = d11*d22 - d12*d12
//   w2 = e1.xyz,             d11
//   w3 = e2.xyz,             d22
//   w4 = d12, kshadow, as_float(order), 0      kshadow = (float)(1.0 - opacity)
// Sphere intersection record:
//   w0 = center.xyz, radius
//   w1 = kshadow, as_float(order), 0, 0
//
// Face shading attributes (4 words): normalised vertex normals and clamped-later
// texture coordinates:
//   a0 = vnn0.xyz, vt0.x   a1 = vnn1.xyz, vt0.y   a2 = vnn2.xyz, vt1.x   a3 = vt1.y, vt2.x, vt2.y, 0
// Object table (one int4 per order): material index, texture index (-1), reference id, flags
constexpr int kObjSmooth = 1;

// Material (3 words): {od.rgb, ka} {os.rgb, kd} {ks, n, opacity, eta}
// Light    (3 words): {xyz, w} {rgb, 0} {L.xyz for directional lights = -norm(dir), 0}
// Texture descriptor (int4): texel offset, width, height, 0 ; texels are RGBA8 (A unused)

struct BvhView {
    const float4* nodes;     // 4 words per node, see bvh.cuh
    int n_nodes;
    int root;
    const uint4* nodes8;     // 6 words per node (Node8, bvh.cuh); n_nodes8 == 0: not built
    const float4* prims8;    // primitive records in the order the 8-wide nodes reference them
    int n_nodes8;
};

// Occluder maps (bvh.cu: build_shadow_maps): an id buffer seen from the light — orthographic for a directional
// light, the six faces of a cube around a point light — holding per cell the OPAQUE primitive nearest to the light
// that covers the cell's centre, as its slot in the primitive array the any-hit walk reads.  A shadow ray first runs the reference's exact test against the
// primitive of its cell (ShadowJob::setup): if that accepts, the carried product is exactly 0 and the walk is
// skipped; if not, the ray is walked as before.  The map only guesses; nothing it holds can change a result.
#ifndef RT_MAX_SHADOW_MAPS
#define RT_MAX_SHADOW_MAPS 1
#endif
constexpr int kMaxShadowMaps = RT_MAX_SHADOW_MAPS;      // lights 0 .. kMaxShadowMaps - 1 may have one (why 1)
struct ShadowMapView {
    const int* cells;                  // [R * R] (directional) or [6][R * R] (point), -1 = none; nullptr = no map
    float ub[3], vb[3];                // directional: orthonormal basis of the light's plane
    float u0, v0, inv_cell;
    float origin[3];                   // point: the light's position; face f = 2 * axis + (that coordinate of p - origin < 0)
    int R;
    int cube;                          // 0: directional, 1: point
};

struct SceneView {
    int nf, ns, nl, ntex;
    const float4* tri;        // nf * kTriWords, file order
    const float4* sph;        // ns * kSphWords, file order
    const float4* face_attr;  // nf * kFaceAttrWords
    const int4*   obj;        // nf + ns
    const float4* mat;        // n_materials * 3
    const float4* light;      // nl * 3
    const int4*   tex_desc;   // ntex
    const uchar4* texels;
    // LBVH over all primitives (bvh.cuh); prim records re-ordered along the Morton curve
    BvhView bvh;
    const float4* bvh_tri;    // triangles in leaf order (same 5-word records)
    const float4* bvh_sph;    // spheres in leaf order
    float bkg[3];
    float bkg_eta;
    int zero_mask_exit;       // shade_node may leave the light loop once shadow_mask == 0 (see there)
    ShadowMapView smap[kMaxShadowMaps];
};

struct CameraView {          // main.cpp:677-710, computed on the host in rt_upload_scene / rt_set_camera
    float eye[3];
    float ul[3];
    float delta_h[3];
    float delta_v[3];
    int width, height;
};

// Result status of a ray / node slot (stored in the w component of `res`)
constexpr float kResFinal = 0.0f;     // rgb is what the parent multiplies by its weights
constexpr float kResZeroTerm = 1.0f;  // `goto SKIP_TRANS`: the parent's term stays (0,0,0)
constexpr float kResPending = 2.0f;   // node shaded, children outstanding: combine pass finishes it
constexpr float kResPrimaryMiss = 3.0f;

// One generation of the wavefront.  Slot i of level d is both "ray i of the queue"
// and "the ShadeRay node that ray produced, if it hit anything".
struct LevelBuffers {
    // ray queue written by level d-1 (unused for level 0: primary rays are generated)
    float4* rq_o;   // origin.xyz, incidence_refraction_index of the PARENT call
    float4* rq_d;   // direction.xyz, transmission_refraction_index of the PARENT call
    int4*   rq_m;   // parent slot, incidence object (order), flags, stack[0]
    int4*   rq_s;   // stack[1..4]
    // node record written by the shade step, read by the combine pass
    float4* nr_a;   // base.rgb (= diffuse*ka + tmp_specular), 0
    float4* nr_b;   // tw1 = (float)(1.0 - F_t), tw2 = (float)(1.0 - opacity), F_r, 0
    int2*   nr_c;   // slots of the transmission / reflection child in generation d+1 (-1 = none)
    float4* res;    // rgb, status
    // LBVH pipeline only (wavefront_bvh.cuh): the hit of slot i, the list of slots that hit
    // something (= the ShadeRay nodes of this generation, in slot order up to the order in which
    // warps of the spawn kernel claim their 32 slots), and one shadow record per (node, light),
    // light-major: sh[li * n_nodes + k] for node_list[k] = the carried shadow_mask after light li
    float4* hp;     // hit point.xyz, as_float(object order); w = as_float(-1) for a slot without a hit
    float4* hb;     // b, g, t, 0
    int*    node_list;
    int*    walk_list;  // the nodes (indices into node_list) whose shadow chain has to be walked: all of them, or those
                        // bvh_shadow_settle_kernel could not settle with light 0's occluder map
    float*  sh;
    int capacity;
};

// flags word of rq_m.z
constexpr int kRayKindReflect = 1;   // bit0: 0 = transmission ray, 1 = reflection ray
constexpr int kRayStateExiting = 2;  // bit1: previous_ray_state == EXITING
constexpr int kRayNStackShift = 2;   // bits 2..5: medium stack length

// Instrumented walk kernels: %globaltimer stamps (ns).  first / empty hold ~t so that a zeroed
// record means "unset" and atomicMax finds the minimum.
struct WalkClock {
    unsigned long long inv_first, inv_empty, last, drain_sum;
    unsigned long long drain_steps, drain_lane_steps, drain_steals;   // inner-loop passes after the queue ran dry, lanes active in them
    unsigned int warps, pad;
};

// Bits of Counters::overflow: the frame is incomplete and rt_render must not hand it out.
constexpr unsigned kOverflowQueue = 1u;    // a wavefront queue was too small: queues are grown, the frame redone
constexpr unsigned kOverflowStack = 2u;    // a traversal stack overflowed (build_bvh checks the depth: never expected)
constexpr unsigned kOverflowRedo = 4u;     // more shadow chains to redo than the redo list holds

constexpr int kRedoCapacity = 1 << 16;     // nodes per launch chain whose shadow chain may go through the slow exact path

struct Counters {
    unsigned int rays[kMaxLevels];        // rays queued for level d (level 0: unused)
    unsigned int nodes[kMaxLevels];       // hits shaded at level d
    unsigned int rays_reflect[kMaxLevels];
    unsigned int overflow;                // kOverflow* bits
    unsigned int redo_count;              // entries of FrameParams::redo_list ((generation << 28) | node index)
    unsigned int fetch_trace[kMaxLevels];   // LBVH pipeline: next job of the persistent trace / shadow kernels
    unsigned int fetch_shadow[kMaxLevels];
    unsigned int walk_nodes[kMaxLevels];    // entries of LevelBuffers::walk_list
    // bvh_shadow_probe_kernel: a sample of this frame's nodes against each occluder map (tests made, occluders
    // confirmed): a map that settles fewer than one ray in eight is left alone for the rest of the frame
    unsigned int smap_tries[kMaxShadowMaps], smap_hits[kMaxShadowMaps];
    unsigned long long bvh_nodes, prim_tests, texels;   // instrumented variant only (all walks)
    unsigned long long shadow_boxes, shadow_prims;      // ... of which in the shadow kernels
    unsigned long long shadow_walks;                    // instrumented LBVH frames: shadow rays not elided by the mask == 0 early-out
    unsigned long long shadow_map_hits;                 // ... shadow rays settled by the occluder map's candidate (no walk)
    WalkClock clock[2][kMaxLevels];                     // instrumented LBVH frames: [0] trace, [1] shadow kernels
};

struct FrameParams {
    SceneView scene;
    CameraView cam;
    LevelBuffers level[kMaxLevels];
    Counters* counters;
    uchar4* tiles;        // this rank's tile buffer [local_tile][row][col], A = flags
    int* primary_ids;     // instrumented variant: reference id of the primary hit per local pixel
    int max_depth;
    float epsilon;
    int rank, world, tile_rows;
    int n_local_rows;     // rows of the tile buffer (padded to whole tiles)
    int n_local_pixels;   // n_local_rows * width
    // Sub-frames (render.cu): a rank's tiles are dealt out to out_mul independent launch
    // chains; for chain j, `rank`/`world` above are the VIRTUAL ones (rank + world*j, world*out_mul)
    // and virtual local tile vt is tile vt*out_mul + out_add of the rank's tile buffer.
    int out_mul, out_add, out_tiles;
    int refill_lanes;     // persistent kernels: fetch new rays when this many lanes of a warp are idle
    int inner_min_lanes;  // ... leave the inner-node loop when fewer lanes than this can still descend
    int leaf_min_lanes;   // ... leave the leaf loop when fewer lanes than this hold a leaf
    int sh_refill_lanes, sh_inner_min_lanes, sh_leaf_min_lanes;   // the same three for the any-hit kernel
    // Shadow chains the any-hit kernel could not finish exactly (more than BvhAccelT::kMaxPartial
    // transparent occluders on one ray): redone by bvh_shadow_redo_kernel.  kRedoCapacity entries.
    unsigned int* redo_list;
    int settled_first;    // light 0 has an occluder map: bvh_shadow_settle_kernel runs before the any-hit kernel, whose
                          // jobs then come from the walk lists (unless the frame's probe found that the map does not pay)
};

}  // namespace rt
