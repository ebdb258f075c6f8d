/*
 * rt_oracle.h — CPU oracle for the render hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's create_view_window_and_ray_trace /
 * ShadeRay / TraceRay (main.cpp:670-1407) taking the same RtSceneDesc the CUDA
 * library takes.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product never does.
 *
 * Parity status: PINNED.  tests/test_oracle_pinned.py checks this restatement
 *  (1) bit-for-bit against the reference's own code compiled from
 *      /root/reference (oracle/_ref/libref.so, built by oracle/Makefile) on
 *      every runnable example scene and on generated branch-coverage scenes,
 *  (2) against fixtures of the reference's output committed under
 *      tests/golden/ (made by tests/golden/make_fixtures.py), and
 *  (3) against the two screenshots of the reference repo that are bit-exact
 *      with its current code (Test1.png, demo_image.png).
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include "../include/rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct RtoParams {
    int32_t max_depth;   /* 4 in the reference (main.cpp:100) */
    float   epsilon;     /* 1.0e-3f (main.cpp:101)            */
} RtoParams;

typedef struct RtoStats {
    uint64_t nodes_total;
    uint64_t nodes_per_level[8];
    uint64_t rays_primary, rays_shadow, rays_transmit, rays_reflect;
    uint64_t prim_tests;      /* ray x primitive tests (brute force) */
    uint64_t texel_fetches;
} RtoStats;

#define RTO_NAN_VALUE INT32_MIN   /* static_cast<int>(NaN) on x86-64 */

/* Renders n_pixels pixels. pixel_index[k] = row*width + col, or NULL for the
 * whole frame in scanline order (n_pixels must then be width*height).
 *   out_raw : n_pixels*3 int32 — exactly the ints the reference stores into
 *             matt(i,j,c) (main.cpp:760-762); RTO_NAN_VALUE for NaN channels.
 *   out_primary_id : n_pixels int32 or NULL — reference id of the primary hit,
 *             0 = miss.
 * Returns 0, or a negative RT_E_* code. */
int rto_render(const RtSceneDesc* desc, const RtoParams* params,
               int64_t n_pixels, const int64_t* pixel_index,
               int32_t* out_raw, int32_t* out_primary_id, RtoStats* stats);

/* Primary rays of the given pixels (main.cpp:677-728): n*3 floats each. */
/* Identity dumps for the next rto_render calls (NULL, NULL to stop): hits[8][n_faces + n_spheres] counts the
 * ShadeRay invocations per generation and object (order: faces in file order, then spheres), shadow[8][n_lights][3]
 * the carried shadow_mask after each light per generation (0: == 0, 1: in between, 2: == 1). */
void rto_set_histograms(uint64_t* hits, uint64_t* shadow);

int rto_primary_rays(const RtSceneDesc* desc, int64_t n_pixels, const int64_t* pixel_index,
                     float* out_origins, float* out_dirs);

/* The closest-hit loop of main.cpp:1109-1124 over TraceRay's result. */
int rto_trace_closest(const RtSceneDesc* desc, int32_t n,
                      const float* origins, const float* dirs, float tmin,
                      int32_t* out_id, float* out_t,
                      float* out_normal, float* out_bary);

/* The shadow loops of main.cpp:896-912 (tmax == NULL) / 928-949. */
int rto_trace_shadow(const RtSceneDesc* desc, int32_t n,
                     const float* origins, const float* dirs,
                     const int32_t* self_id, float tmin, const float* tmax,
                     float* out_mask, int32_t* out_hits);

/* Splits raw ints into the 8-bit frame + flag layout of rt_download_framebuffer. */
void rto_raw_to_rgb8(int64_t n_pixels, const int32_t* raw,
                     const int32_t* primary_id_or_null,
                     uint8_t* rgb, uint8_t* flags);

#ifdef __cplusplus
}
#endif
#endif
