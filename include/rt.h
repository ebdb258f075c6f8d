/*
 * rt.h — C ABI of the B200 render core for simple-raytracer scenes.
 *
 * This is the drop-in boundary for the reference's one hot path. The reference
 * has no FFI of its own: the seam is the single call
 *
 *     Mat3D matt = create_view_window_and_ray_trace(view_origin,
 *                      view_direction.norm(), view_up.norm(), fov_h,
 *                      height, width, background_color);      (main.cpp:607)
 *
 * whose callee reads everything else from the mutable global `Globals
 * environment` (main.cpp:58, src/definitions.h:304-311).  The structs below
 * carry exactly that state as plain-old-data, the functions replace that call.
 *
 * Conventions
 *  - plain C, no exceptions; every int-returning function returns RT_OK (0) or
 *    a negative RT_E_* code and leaves a message in rt_last_error().
 *  - all pointers inside RtSceneDesc are HOST pointers, caller-owned, and may
 *    be released as soon as rt_upload_scene() returns.
 *  - one RtScene lives on the CUDA device that was current when it was
 *    uploaded.  A handle is not thread-safe; distinct handles are independent.
 *  - there is no CPU fallback: without a CUDA device every call fails with
 *    RT_E_CUDA.
 */
#ifndef RT_B200_RT_H
#define RT_B200_RT_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 4

/* ---- status codes ------------------------------------------------------ */
#define RT_OK            0
#define RT_E_INVALID    -1   /* bad argument / malformed scene description   */
#define RT_E_CUDA       -2   /* CUDA runtime error (see rt_last_error)        */
#define RT_E_NOMEM      -3   /* host or device allocation failed              */
#define RT_E_STATE      -4   /* call order (e.g. download before render)      */
#define RT_E_PARSE      -5   /* scene / texture file could not be parsed      */
#define RT_E_IO         -6   /* file could not be opened / written            */
#define RT_E_LIMIT      -7   /* a stated limit of the implementation was hit: the
                                frame (or the LBVH) is refused, never returned wrong */

/* ---- scene description (what `environment` + main()'s locals hold) ------ */

/* Phong material, src/definitions.h:249-254, filled by `mtlcolor`
 * (main.cpp:272-308).  opacity is already clamped to [0,1]; opacity = eta = 1
 * when the scene line had no alpha/eta pair. */
typedef struct RtMaterial {
    float od[3];     /* diffuse colour   Od  */
    float os[3];     /* specular colour  Os  */
    float ka, kd, ks;
    float n;         /* specular exponent */
    float opacity;   /* alpha */
    float eta;       /* refraction index */
} RtMaterial;

/* 8-bit RGB texture, row-major [y][x][3]; the reference keeps it as
 * Mat3D[x][y][c] of size_t (src/definitions.h:197-237, utility.h:59-139). */
typedef struct RtTexture {
    int32_t width, height;
    const uint8_t* rgb;
} RtTexture;

/* src/definitions.h:265-270 + the SceneObjectInfo it points to (256-263). */
typedef struct RtSphere {
    float   center[3];
    float   radius;
    int32_t material;   /* index into RtSceneDesc.materials               */
    int32_t texture;    /* index into RtSceneDesc.textures, -1 = none     */
    int32_t id;         /* reference object id: 1-based creation order over
                           spheres and faces together (main.cpp:329, 476)  */
} RtSphere;

/* src/definitions.h:272-281.  The surface normal is NOT passed: the library
 * recomputes it with the expression of main.cpp:537-539. */
typedef struct RtFace {
    float   v[3][3];    /* vertex positions                                */
    float   vn[3][3];   /* per-vertex normals (raw, un-normalised)         */
    float   vt[3][2];   /* per-vertex texture coordinates                  */
    int32_t smooth;     /* Face::smooth_shading                            */
    int32_t material;
    int32_t texture;    /* -1 = none                                       */
    int32_t id;
} RtFace;

/* src/definitions.h:283-289; w == 0 -> directional with direction xyz
 * (un-normalised!), otherwise point light at xyz (main.cpp:383-396). */
typedef struct RtLight {
    float xyz[3];
    float w;
    float rgb[3];
} RtLight;

typedef struct RtSceneDesc {
    int32_t width, height;        /* imsize                                  */
    float   eye[3];
    float   viewdir[3];           /* raw; normalised inside like main.cpp:607 */
    float   updir[3];             /* raw; normalised inside like main.cpp:607 */
    float   hfov_deg;
    float   bkg[3];               /* NOT clamped (main.cpp:721)               */
    float   bkg_eta;              /* 0.0f when bkgcolor had no 4th argument
                                     (main.cpp:260-264, 751)                  */
    int32_t n_materials;  const RtMaterial* materials;
    int32_t n_textures;   const RtTexture*  textures;
    int32_t n_spheres;    const RtSphere*   spheres;   /* file order */
    int32_t n_faces;      const RtFace*     faces;     /* file order */
    int32_t n_lights;     const RtLight*    lights;    /* file order */
    int32_t camera_normalised;    /* 0: viewdir / updir are the scene file's raw vectors and are normalised
                                     inside (what main.cpp:607 does at the call site); 1: they are what
                                     create_view_window_and_ray_trace RECEIVES, i.e. already normalised,
                                     and are used bit for bit (the drop-in binding) */
} RtSceneDesc;

/* ---- render parameters (environment.other, main.cpp:100-101) ------------ */
#define RT_ACCEL_AUTO   0   /* brute force for small scenes, LBVH otherwise */
#define RT_ACCEL_BRUTE  1
#define RT_ACCEL_LBVH   2   /* (a scene with a non-finite material or light value is rendered by brute force all the
                               same: RtStats.accel_used says which) */
#define RT_ACCEL_LBVH8  3   /* rt_trace_shadow only: the 8-wide tree the any-hit kernel walks */

typedef struct RtRenderParams {
    int32_t max_depth;     /* environment.other["recursion_depth"], default 4 */
    float   epsilon;       /* environment.other["epsilon"], default 1.0e-3f   */
    int32_t accel;         /* RT_ACCEL_*                                      */
    int32_t rank;          /* row-tile sharding: this process renders tiles   */
    int32_t world_size;    /*   t with t % world_size == rank; default 0 / 1  */
    int32_t tile_rows;     /* rows per tile (default 8)                       */
    int32_t instrument;    /* 1: count BVH work / texels, keep primary hit ids: the
                              same launch schedule with the counting builds of the
                              walk kernels (slower; never use for timing)       */
    int32_t profile;       /* 1: one serial launch chain with CUDA events around every
                              kernel; rt_render reports RtStats.level_ms / trace_ms /
                              shadow_ms / shade_ms / combine_ms                 */
    int32_t subframes;     /* LBVH pipeline: independent launch chains the rank's
                              tiles are dealt out to (1..8; 0 = choose).  Never
                              changes a pixel, only how kernel tails overlap. */
    int32_t work_sharing;  /* LBVH walk kernels: once a launch's ray queue is dry,
                              idle lanes take subtrees off the stacks of the lanes
                              still walking.  0 = for short launches (default),
                              1 = always, 2 = never.  Never changes a pixel.   */
} RtRenderParams;

/* Filled by rt_render when a non-NULL pointer is passed.  The ray counts
 * follow the reference's unit "one TraceRay call" (main.cpp:729, 896, 928,
 * 992, 1113); they are only counted when params.instrument != 0, the node
 * counts always. */
typedef struct RtStats {
    double   render_ms;          /* device time of this rt_render (CUDA events) */
    uint64_t nodes_total;        /* ShadeRay invocations                       */
    uint64_t nodes_per_level[8]; /* by generation (0 = primary hits)           */
    uint64_t rays_primary, rays_shadow, rays_transmit, rays_reflect;
    uint64_t bvh_nodes_visited, prim_tests;
    uint64_t texel_fetches;
    int32_t  accel_used;         /* RT_ACCEL_BRUTE or RT_ACCEL_LBVH            */
    int32_t  kernel_launches;    /* launches of our kernels in this render     */
    int32_t  queue_regrows;      /* wavefront queues re-sized and frame redone */
    int32_t  reserved;
    double   level_ms[8];        /* params.profile: device time of the trace+shade+spawn
                                    kernel of each generation                   */
    double   combine_ms;         /* params.profile: all bottom-up combine passes  */
    uint64_t shadow_rays_walked; /* instrument, LBVH: shadow rays actually traced.  rays_shadow counts
                                    the reference's TraceRay calls (one per light per node); the
                                    LBVH pipeline skips the walks whose result cannot matter because
                                    the carried shadow_mask is already exactly 0.  Equal
                                    to rays_shadow on the brute-force path. */
    /* LBVH pipeline detail.  params.profile: device time per generation of the three kernels;
       params.instrument: the part of bvh_nodes_visited / prim_tests done by the shadow kernels. */
    double   trace_ms[8], shadow_ms[8], shade_ms[8];
    uint64_t shadow_boxes_tested, shadow_prim_tests;
    /* params.instrument, LBVH: how each persistent walk kernel spent its time (device
       %globaltimer): span = first warp's start .. last warp's end; drain = from the moment
       the job queue ran dry (no lane can be refilled any more) to the last warp's end;
       warp_drain = the same, averaged over warps. */
    double   trace_span_ms[8], trace_drain_ms[8], trace_warp_drain_ms[8];
    double   shadow_span_ms[8], shadow_drain_ms[8], shadow_warp_drain_ms[8];
    /* ... and, per warp on average, the inner-node passes it still made after the queue ran
       dry, the lanes active in such a pass, and the subtrees handed to idle lanes. */
    double   trace_drain_steps[8], trace_drain_lanes[8], trace_drain_steals[8];
    double   shadow_drain_steps[8], shadow_drain_lanes[8];
    uint64_t shadow_chains_redone; /* LBVH: nodes whose shadow chain met more than 12 transparent occluders on
                                    one ray and went through the exact slow path (bvh_shadow_redo_kernel) */
    int32_t  bvh_levels4, bvh_levels8; /* LBVH: depth of the closest-hit (4-wide) and any-hit (8-wide) trees */
    int32_t  graph_replays;      /* frames of this scene so far that were issued as ONE CUDA-graph launch: a frame
                                    whose buffers, camera, sharding and parameters equal the previous frame's is
                                    captured on its second sighting and replayed afterwards */
    int32_t  shadow_map_hits;    /* instrumented LBVH frames: shadow rays settled by a directional light's occluder map
                                    (the exact test against the map's candidate found an opaque occluder: no walk);
                                    saturates at INT32_MAX */
    double   spawn_ms[8];        /* params.profile, LBVH: bvh_spawn_kernel of each generation (trace_ms is the walk alone) */
} RtStats;

typedef struct RtScene RtScene;   /* opaque: all device state of one scene */

/* ---- the three calls named by the north star ---------------------------- */

/* Packs the scene into structure-of-arrays device buffers, uploads textures,
 * builds the LBVH when the scene is large.  Replaces the parser's writes into
 * `environment` (main.cpp:328-552). */
int rt_upload_scene(const RtSceneDesc* desc, RtScene** out);

/* Renders this rank's row tiles.  Replaces create_view_window_and_ray_trace
 * (main.cpp:670-767) and everything below it (ShadeRay 783-1207, TraceRay
 * 1215-1407).  Synchronous.  `params` may be NULL (defaults). */
int rt_render(RtScene* scene, const RtRenderParams* params, RtStats* stats_or_null);

/* Copies the last rendered frame to host memory.
 *   rgb   : height*width*3 bytes, row-major, top row first (= matt(i,j,c),
 *           main.cpp:760-762).
 *   flags : height*width bytes or NULL. bit0..2: channel r/g/b was NaN (the
 *           reference prints 18446744071562067968 there); bit3:
 *           primary ray missed (pixel = un-clamped bkgcolor); rgb holds the
 *           value clamped to 0..255 in that case.
 * With world_size > 1 only the rows of this rank's tiles are written; other
 * rows of the host buffers are left untouched. */
int rt_download_framebuffer(RtScene* scene, uint8_t* rgb, uint8_t* flags);

void        rt_free_scene(RtScene* scene);
const char* rt_last_error(void);
int         rt_abi_version(void);

/* ---- device-side entry points used by the multi-GPU / re-render plumbing - */

/* Number of bytes of this rank's tile buffer ([local_tile][row][col] RGBA8,
 * A = flags) for the given sharding. */
size_t rt_tile_buffer_bytes(const RtScene* scene, int32_t world_size, int32_t tile_rows);

/* The sharding arithmetic itself (pure host code, no CUDA call): rows are cut into
 * tiles of tile_rows rows, tile t belongs to rank t % world_size, every rank holds
 * ceil(n_tiles / world_size) tiles (padded so that an allgather has equal counts).
 *   out[0] = tiles in the image, out[1] = tiles per rank, out[2] = rows per rank
 *   (padded), out[3] = pixels per rank (padded). */
int rt_tile_layout(int32_t height, int32_t width, int32_t world_size,
                   int32_t tile_rows, int32_t out[4]);

/* Like rt_render, but asynchronous on `cuda_stream` (a cudaStream_t) and with the RGBA8 tile buffer
 * written to caller-provided DEVICE memory (e.g. the send buffer of an NCCL allgather).
 *   cuda_stream : NULL = the scene's OWN non-blocking stream (the one rt_render uses) — NOT the legacy
 *                 default stream: work a caller enqueues on the default stream is not ordered against
 *                 the frame; pass your stream explicitly when something of yours follows the frame.
 *   dev_tiles   : may be NULL to render into the scene's own buffer only.
 * A frame equal to the previous one (same buffers, camera, sharding, parameters) is replayed as one
 * CUDA-graph launch from its third sighting on; a stream the caller is capturing is enqueued into directly. */
int rt_render_async(RtScene* scene, const RtRenderParams* params,
                    void* dev_tiles, void* cuda_stream);

/* rt_render_async cannot re-run a frame whose wavefront queues turned out too small
 * (rt_render does; the reference's recursion has no such limit).  Call this once the
 * frame's stream has been synchronised: RT_OK = the frame is complete; RT_E_NOMEM = a
 * queue overflowed and the frame is truncated — the queues have been grown, render it
 * again.  A synchronous rt_render of the same view beforehand sizes the queues exactly. */
int rt_async_frame_status(RtScene* scene);

/* Device buffers of freed scenes (tables, LBVH build temporaries, wavefront queues) are
 * kept in a process-wide cache and handed to the next rt_upload_scene / rt_render, so a
 * load-render-free loop does not pay cudaMalloc / cudaFree (and cudaFree's device
 * synchronisation) per scene.  This returns everything cached to the driver; the
 * result is the number of bytes released. */
size_t rt_release_cached_memory(void);

/* The reference's output file, formatted on the device (replaces the writer loop
 * main.cpp:628-641): "P3 \n" "W H \n" "255 \n" and one "r g b \n" line per
 * pixel, NaN channels as 18446744071562067968 and primary-miss pixels as the un-clamped
 * (int)(bkgcolor*255), byte for byte what rt_host_format_ppm produces from the downloaded
 * frame.  dev_rgba: width*height RGBA8 in scanline order with the flags in A, in DEVICE
 * memory (the output of rt_deinterleave_tiles), or NULL for the frame of this scene's last
 * single-rank rt_render.  *text points at pinned host memory owned by the scene's
 * workspace: valid until the next call on this scene or rt_free_scene; *len excludes the
 * terminating NUL. */
int rt_format_ppm_device(RtScene* scene, const void* dev_rgba, const char** text, size_t* len);

/* rt_format_ppm_device without a single allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Needs no scene handle: the caller passes
 * the frame and receives the text, both in DEVICE memory of the current device.
 *   dev_rgba      : width*height RGBA8 in scanline order, A = flags (as rt_deinterleave_tiles leaves it,
 *                   or the tile buffer of a world_size == 1 rt_render_async)
 *   bkg           : RtSceneDesc.bkg (primary-miss pixels print the un-clamped (int)(bkg*255))
 *   dev_text      : receives "P3 \n" "W H \n" "255 \n" + one "r g b \n" line per pixel, byte for byte
 *                   what rt_format_ppm_device / rt_host_format_ppm produce; text_capacity bytes.
 *                   41 + 64 * width * height always suffices; 12 * width * height + 64 does when
 *                   no channel is NaN and bkg is in range
 *   dev_len       : receives the file's length in bytes (8 bytes, device).  *dev_len > text_capacity
 *                   means the buffer was too small: whole 256-pixel blocks are missing from the text
 * Enqueues three kernels on cuda_stream (NULL = the legacy default stream) and returns; it calls no
 * cudaMalloc / cudaFree / allocating routine and does not synchronise.  LIMIT: width * height <=
 * 33,177,600 pixels (7680 x 4320), RT_E_LIMIT beyond: the per-block scratch is a static __device__
 * array of 1.56 MB that every GPU loading this library holds until the process exits.  All calls on
 * a device SHARE that array and it keeps the last call's contents (each call rewrites every entry
 * it reads, nothing needs clearing): calls that overlap on the GPU — different streams without an
 * ordering between them — race.  Calls on one stream, or ordered by events, are safe. */
int rt_format_ppm_noalloc(const void* dev_rgba, int32_t width, int32_t height, const float bkg[3],
                          char* dev_text, size_t text_capacity, unsigned long long* dev_len,
                          void* cuda_stream);

/* After an allgather of every rank's tile buffer ([rank][local_tile][row][col]
 * RGBA8, device memory) puts the pixels in scanline order:
 * dev_rgba_out[height][width] RGBA8. Asynchronous on cuda_stream. */
int rt_deinterleave_tiles(const RtScene* scene, const void* dev_gathered,
                          void* dev_rgba_out, int32_t world_size,
                          int32_t tile_rows, void* cuda_stream);

/* ---- N GPUs of one node behind the same three calls ------------------------------
 * One process, one host thread: the handle owns the scene replicas of all devices, their streams and
 * the NCCL communicators (ncclCommInitAll; NCCL is dlopen'ed at the first multi-GPU upload, RT_NCCL_LIB
 * overrides the library name).  The scene is uploaded, packed and its LBVH built once on the first
 * device and broadcast to the others over NVLink; a frame is rendered as interleaved row tiles (tile t
 * on device t % N), exchanged with ONE ncclAllGather of the RGBA8 tiles and de-interleaved on the first
 * device.  The frame is byte-identical for every N (main.cpp:718-763 has no cross-pixel state).
 *   device_ids : num_gpus CUDA device ordinals, or NULL for 0 .. num_gpus-1
 *   params     : as rt_render; rank / world_size are ignored, tile_rows 0 = 10 when height % 10 == 0, else 8
 *   stats      : sums over the devices; render_ms = first device's stream, enqueue to de-interleaved frame */
typedef struct RtMultiScene RtMultiScene;
int  rt_upload_scene_multi(const RtSceneDesc* desc, int32_t num_gpus, const int32_t* device_ids, RtMultiScene** out);
int  rt_render_multi(RtMultiScene* scene, const RtRenderParams* params, RtStats* stats_or_null);
int  rt_download_framebuffer_multi(RtMultiScene* scene, uint8_t* rgb, uint8_t* flags);
int  rt_format_ppm_multi(RtMultiScene* scene, const char** text, size_t* len);   /* as rt_format_ppm_device */
int  rt_num_gpus_multi(const RtMultiScene* scene);
void rt_free_scene_multi(RtMultiScene* scene);

/* Re-render API: replace the camera without re-uploading. */
int rt_set_camera(RtScene* scene, const float eye[3], const float viewdir[3],
                  const float updir[3], float hfov_deg);

/* ---- parity probes (the reference's TraceRay + closest-hit loops) -------- */

/* Closest hit of n rays against the scene (HOST arrays): the loop of
 * main.cpp:730-742 / 1109-1124 applied to TraceRay (1215-1407) with the tie
 * rule "faces before spheres, file order, strict <" and t > tmin.
 *   origins, dirs : n*3 floats
 *   out_id        : reference object id of the hit, 0 = miss
 *   out_t         : hit distance (bit pattern of the reference's float)
 *   out_normal    : n*3 floats, may be NULL
 *   out_bary      : n*3 floats (a,b,g of a face hit), may be NULL */
int rt_trace_closest(RtScene* scene, int32_t accel, int32_t n,
                     const float* origins, const float* dirs, float tmin,
                     int32_t* out_id, float* out_t,
                     float* out_normal, float* out_bary);

/* Shadow attenuation of n rays (HOST arrays): product of (1 - opacity) over
 * all hits with tmin < t (< tmax[i] when tmax != NULL) of all objects whose id
 * differs from self_id[i] (main.cpp:896-912, 928-949).  out_hits = number of
 * qualifying hits. */
int rt_trace_shadow(RtScene* scene, int32_t accel, int32_t n,
                    const float* origins, const float* dirs,
                    const int32_t* self_id, float tmin, const float* tmax,
                    float* out_mask, int32_t* out_hits);

/* Identity of the SECONDARY hits and of shadow occlusion (north star: "hit primitive IDs and shadow occlusion
 * must be identical"; SURVEY 8c "oracle B": hit ids at main.cpp:745 / 1008 / 1120, occlusion at 909 / 946), of the
 * last frame of the LBVH pipeline (this rank's tiles):
 *   hits  [8][n_faces + n_spheres]  ShadeRay invocations of generation g on object o (order: faces in file
 *                                   order, then spheres)
 *   shadow[8][n_lights][3]          nodes of generation g whose carried shadow_mask after light l was exactly 0 /
 *                                   in (0,1) / exactly 1
 * Compared with the same histograms of the CPU oracle in tests/test_gpu_parity.py. */
int rt_download_hit_histograms(RtScene* scene, uint64_t* hits, uint64_t* shadow);

/* Primary-hit object id per pixel of the last render (height*width int32,
 * 0 = miss); available when the last render had instrument != 0. */
int rt_download_primary_ids(RtScene* scene, int32_t* ids);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_RT_H */
