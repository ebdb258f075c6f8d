/*
 * rt_oracle.c — CPU oracle for the render hot path.  TEST INFRASTRUCTURE ONLY
 * (see rt_oracle.h for who may load it and how its parity is pinned).
 *
 * Plain-C restatement of the reference algorithm.  Every function cites the
 * reference lines it follows (paths relative to /root/reference).  Float /
 * double usage follows SURVEY Appendix A2; build with
 *     gcc -O2 -ffp-contract=off            (no -march, no -ffast-math)
 * so that no FMA contraction changes the rounding.
 *
 * Ray-independent quantities (triangle edges, plane constant, Gram matrix,
 * normalised vertex normals, r^2) are hoisted into the prepared scene; they are
 * pure functions of the primitive, so the hoist is bit-identical to the
 * reference recomputing them per ray (main.cpp:1280-1282, 1301, 1361-1366).
 */
#include "rt_oracle.h"

#include <math.h>
#include <float.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846   /* src/config.h:11 */
#endif
#define VIEW_DIST 5.0                  /* src/config.h:8  `#define d 5.0` */

typedef struct { float x, y, z; } V3;
typedef struct { float r, g, b; } Col;

/* ---- Vector3, src/definitions.h:18-139 ---------------------------------- */
static inline V3 v3(float x, float y, float z) { V3 r = { x, y, z }; return r; }
static inline V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 vmulf(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 vdivf(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
/* sum(): x + y + z, left to right (definitions.h:24-27) */
static inline float vsum(V3 a) { return a.x + a.y + a.z; }
/* dot(): (a*b).sum() (definitions.h:29-32) */
static inline float vdot(V3 a, V3 b) { return vsum(v3(a.x * b.x, a.y * b.y, a.z * b.z)); }
/* mag(): std::sqrt(float) of square().sum() (definitions.h:34-37) */
static inline float vmag(V3 a) { return sqrtf(vsum(v3(a.x * a.x, a.y * a.y, a.z * a.z))); }
/* norm(): three divisions by mag() (definitions.h:57-60) */
static inline V3 vnorm(V3 a) { return vdivf(a, vmag(a)); }
static inline V3 vcross(V3 a, V3 b)   /* definitions.h:48-55 */
{
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

/* ---- Color, src/definitions.h:141-195: every op clamps, NaN passes ------- */
static inline float clamp01(float v) { return (v < 0.0f) ? 0.0f : (1.0f < v) ? 1.0f : v; }
static inline float clampf(float v, float lo, float hi) { return (v < lo) ? lo : (hi < v) ? hi : v; }
static inline Col cmul(Col a, Col o) { Col r = { clamp01(o.r * a.r), clamp01(o.g * a.g), clamp01(o.b * a.b) }; return r; }
static inline Col cmulf(Col a, float o) { Col r = { clamp01(o * a.r), clamp01(o * a.g), clamp01(o * a.b) }; return r; }
static inline Col cadd(Col a, Col o) { Col r = { clamp01(o.r + a.r), clamp01(o.g + a.g), clamp01(o.b + a.b) }; return r; }

/* map(), src/utility.h:23-26, all float */
static inline float mapf(float x, float in_min, float in_max, float out_min, float out_max)
{
    return (x - in_min) * (out_max - out_min) / (in_max - in_min) + out_min;
}

/* ---- prepared scene ------------------------------------------------------ */
typedef struct {
    V3 n;                 /* surface_normal, main.cpp:537-539                 */
    V3 v0, e1, e2;        /* main.cpp:1280-1281                               */
    float D;              /* -normal.dot(vertex[0]), main.cpp:1301            */
    float d11, d12, d22, det;   /* main.cpp:1361-1366                         */
    V3 vnn[3];            /* vertex_normal[k].norm(), main.cpp:1382-1384      */
    float vt[3][2];
    int smooth, mat, tex, id;
} OFace;

typedef struct {
    V3 c; float r;
    double r2;            /* pow(radius, 2.0) in double, main.cpp:1234        */
    int mat, tex, id;
} OSphere;

typedef struct {
    const RtSceneDesc* d;
    int nf, ns;
    OFace* f; OSphere* s;
    float eps; int max_depth;
    RtoStats* st;
    RtoStats st_local;
} OScene;

typedef struct { float t; V3 p, n, bary; } OHit;

/* Objects are named by their position in TraceRay's iteration order:
 * faces 0..nf-1 (file order) then spheres nf..nf+ns-1 (file order) — the
 * std::map<std::string,...> key order "face" < "sphere" (main.cpp:1218,
 * definitions.h:305). */
static inline int obj_is_sphere(const OScene* S, int o) { return o >= S->nf; }
static inline int obj_id(const OScene* S, int o) { return o < S->nf ? S->f[o].id : S->s[o - S->nf].id; }
static inline int obj_tex(const OScene* S, int o) { return o < S->nf ? S->f[o].tex : S->s[o - S->nf].tex; }
static inline const RtMaterial* obj_mat(const OScene* S, int o)
{
    return &S->d->materials[o < S->nf ? S->f[o].mat : S->s[o - S->nf].mat];
}

static int prepare(OScene* S, const RtSceneDesc* d, const RtoParams* p, RtoStats* st)
{
    memset(S, 0, sizeof *S);
    if (!d || d->n_faces < 0 || d->n_spheres < 0 || d->n_lights < 0) return RT_E_INVALID;
    S->d = d; S->nf = d->n_faces; S->ns = d->n_spheres;
    S->eps = p ? p->epsilon : 1.0e-3f;
    S->max_depth = p ? p->max_depth : 4;
    S->st = st ? st : &S->st_local;
    memset(S->st, 0, sizeof *S->st);
    S->f = (OFace*)malloc(sizeof(OFace) * (size_t)(S->nf ? S->nf : 1));
    S->s = (OSphere*)malloc(sizeof(OSphere) * (size_t)(S->ns ? S->ns : 1));
    if (!S->f || !S->s) { free(S->f); free(S->s); return RT_E_NOMEM; }
    for (int i = 0; i < S->nf; i++) {
        const RtFace* in = &d->faces[i]; OFace* o = &S->f[i];
        V3 p0 = v3(in->v[0][0], in->v[0][1], in->v[0][2]);
        V3 p1 = v3(in->v[1][0], in->v[1][1], in->v[1][2]);
        V3 p2 = v3(in->v[2][0], in->v[2][1], in->v[2][2]);
        o->v0 = p0; o->e1 = vsub(p1, p0); o->e2 = vsub(p2, p0);
        o->n = vnorm(vcross(o->e1, o->e2));                 /* main.cpp:537-539 */
        o->D = -vdot(o->n, p0);                             /* main.cpp:1301 */
        o->d11 = vdot(o->e1, o->e1); o->d12 = vdot(o->e1, o->e2); o->d22 = vdot(o->e2, o->e2);
        o->det = o->d11 * o->d22 - o->d12 * o->d12;         /* main.cpp:1366 */
        for (int k = 0; k < 3; k++) {
            o->vnn[k] = vnorm(v3(in->vn[k][0], in->vn[k][1], in->vn[k][2]));
            o->vt[k][0] = in->vt[k][0]; o->vt[k][1] = in->vt[k][1];
        }
        o->smooth = in->smooth; o->mat = in->material; o->tex = in->texture; o->id = in->id;
        if (o->mat < 0 || o->mat >= d->n_materials || o->tex >= d->n_textures) return RT_E_INVALID;
    }
    for (int i = 0; i < S->ns; i++) {
        const RtSphere* in = &d->spheres[i]; OSphere* o = &S->s[i];
        o->c = v3(in->center[0], in->center[1], in->center[2]); o->r = in->radius;
        o->r2 = pow((double)in->radius, 2.0);               /* main.cpp:1234 */
        o->mat = in->material; o->tex = in->texture; o->id = in->id;
        if (o->mat < 0 || o->mat >= d->n_materials || o->tex >= d->n_textures) return RT_E_INVALID;
    }
    return RT_OK;
}

static void unprepare(OScene* S) { free(S->f); free(S->s); S->f = NULL; S->s = NULL; }

/* ---- TraceRay, main.cpp:1215-1407 ----------------------------------------
 * The reference returns a vector of every intersection; here each one is handed
 * to a visitor in the same order (faces in file order, then per sphere the far
 * root before the near root).  A visitor returning non-zero stops the walk —
 * used only for the `goto SKIP_TRANS` rule, which leaves the loop. */
typedef int (*visit_fn)(void* ctx, int obj, const OHit* h);

static void trace_ray(const OScene* S, V3 o, V3 ray, visit_fn visit, void* ctx)
{
    S->st->prim_tests += (uint64_t)(S->nf + S->ns);
    /* faces, main.cpp:1275-1402 */
    for (int i = 0; i < S->nf; i++) {
        const OFace* f = &S->f[i];
        float dem = vdot(f->n, ray);                                 /* 1296 */
        if (dem == 0.0f) continue;                                   /* 1297 */
        float distance = -(vdot(f->n, o) + f->D) / dem;              /* 1302 */
        V3 p = vadd(o, vmulf(ray, distance));                        /* 1303 */
        V3 ep = vsub(p, f->v0);                                      /* 1360 */
        float d1p = vdot(f->e1, ep), d2p = vdot(f->e2, ep);          /* 1364-1365 */
        if (f->det != 0.0f) {                                        /* 1367 */
            float b = (f->d22 * d1p - f->d12 * d2p) / f->det;        /* 1368 */
            float g = (f->d11 * d2p - f->d12 * d1p) / f->det;        /* 1369 */
            float a = 1.0f - (b + g);                                /* 1370 */
            if (((0.0f < a) && (a < 1.0f)) && ((0.0f < b) && (b < 1.0f)) && ((0.0f < g) && (g < 1.0f))) {
                OHit h; h.t = distance; h.p = p; h.bary = v3(a, b, g);
                if (f->smooth)                                       /* 1380-1385 */
                    h.n = vnorm(vadd(vadd(vmulf(f->vnn[0], a), vmulf(f->vnn[1], b)), vmulf(f->vnn[2], g)));
                else
                    h.n = f->n;
                if (visit(ctx, i, &h)) return;
            }
        }
    }
    /* spheres, main.cpp:1219-1274 */
    for (int i = 0; i < S->ns; i++) {
        const OSphere* s = &S->s[i];
        V3 dir = vsub(o, s->c);                                      /* 1225 */
        float A = 1.0f;                                              /* 1232 */
        float B = (float)(2.0 * (double)vsum(v3(ray.x * dir.x, ray.y * dir.y, ray.z * dir.z)));   /* 1233 */
        float C = (float)((double)vsum(v3(dir.x * dir.x, dir.y * dir.y, dir.z * dir.z)) - s->r2); /* 1234 */
        float det = (float)(pow((double)B, 2.0) - (4.0 * (double)A * (double)C));                  /* 1242 */
        OHit h; h.bary = v3(0.0f, 0.0f, 0.0f);
        int obj = S->nf + i;
        if (!signbit(det)) {                                         /* 1243 */
            float t1 = (float)((double)(-B + sqrtf(det)) / (2.0 * (double)A));   /* 1244 */
            h.t = t1; h.p = vadd(o, vmulf(ray, t1));
            h.n = vnorm(vdivf(vsub(h.p, s->c), s->r));               /* 1249 */
            if (visit(ctx, obj, &h)) return;
            float t2 = (float)((double)(-B - sqrtf(det)) / (2.0 * (double)A));   /* 1252 */
            h.t = t2; h.p = vadd(o, vmulf(ray, t2));
            h.n = vnorm(vdivf(vsub(h.p, s->c), s->r));               /* 1257 */
            if (visit(ctx, obj, &h)) return;
        } else if (det == 0) {                                       /* 1260: only -0.0f */
            float t = (float)((double)(-B) / 2.0);                   /* 1261 */
            h.t = t; h.p = vadd(o, vmulf(ray, t));
            h.n = vnorm(vdivf(vsub(h.p, s->c), s->r));               /* 1266 */
            if (visit(ctx, obj, &h)) return;
        }
    }
}

/* closest-hit loops: main.cpp:730-742 (tmin = 0), 1109-1124 (tmin = eps) */
typedef struct { float tmin, best; int obj; OHit hit; } ClosestCtx;
static int visit_closest(void* c, int obj, const OHit* h)
{
    ClosestCtx* x = (ClosestCtx*)c;
    if (h->t > x->tmin && h->t < x->best) { x->best = h->t; x->obj = obj; x->hit = *h; }
    return 0;
}

/* transmission closest hit with the face rule, main.cpp:992-1011 */
typedef struct { float tmin, best; int obj; OHit hit; int face_rule; int back_id; const OScene* S; int aborted; } TransCtx;
static int visit_trans(void* c, int obj, const OHit* h)
{
    TransCtx* x = (TransCtx*)c;
    if (h->t > x->tmin && h->t < x->best) {
        if (x->face_rule && obj_id(x->S, obj) != x->back_id) { x->aborted = 1; return 1; }   /* 1000-1002 */
        x->best = h->t; x->obj = obj; x->hit = *h;
    }
    return 0;
}

/* shadow loops, main.cpp:898-912 (directional) and 930-949 (point) */
typedef struct { const OScene* S; int self_id; float tmin, tmax; int use_tmax; float mask; int hits; } ShadowCtx;
static int visit_shadow(void* c, int obj, const OHit* h)
{
    ShadowCtx* x = (ShadowCtx*)c;
    if (obj_id(x->S, obj) == x->self_id) return 0;                   /* 900, 935 */
    if (h->t > x->tmin && (!x->use_tmax || h->t < x->tmax)) {        /* 907, 944 */
        float k = (float)(1.0 - (double)obj_mat(x->S, obj)->opacity);   /* 909, 946 */
        x->mask = clamp01(k * x->mask);
        x->hits++;
    }
    return 0;
}

/* ---- ShadeRay, main.cpp:783-1207 ------------------------------------------ */
#define STATE_ENTERING 0
#define STATE_EXITING  1
#define MAX_STACK 16   /* the medium stack grows by at most one entry per generation */

static int in_stack(const int* stack, int n, int obj)   /* utility.h:6-12 */
{
    for (int i = 0; i < n; i++) if (stack[i] == obj) return 1;
    return 0;
}

static Col texel(const OScene* S, int tex, int x, int y)
{
    const RtTexture* T = &S->d->textures[tex];
    /* static_cast<int>(NaN) would index out of range in the reference (UB);
     * keep the oracle memory-safe. */
    if (x < 0) x = 0;
    if (x >= T->width) x = T->width - 1;
    if (y < 0) y = 0;
    if (y >= T->height) y = T->height - 1;
    const uint8_t* px = T->rgb + ((size_t)y * (size_t)T->width + (size_t)x) * 3;
    S->st->texel_fetches++;
    Col c = { mapf((float)px[0], 0.0f, 255.0f, 0.0f, 1.0f),            /* 823-825, 858-860 */
              mapf((float)px[1], 0.0f, 255.0f, 0.0f, 1.0f),
              mapf((float)px[2], 0.0f, 255.0f, 0.0f, 1.0f) };
    return c;
}

/* Identity dumps (SURVEY 8c "oracle B"): what ShadeRay was called FOR, beyond the pixels it produced.
 *   hits  [level][object order]      ShadeRay invocations of generation `level` on that object — the hit ids
 *                                    chosen at main.cpp:745 (level 0), 1008 (transmission), 1120 (reflection)
 *   shadow[level][light][class]      the carried shadow_mask after that light's TraceRay (main.cpp:909 / 946):
 *                                    class 0: exactly 0 (an opaque occluder so far), 1: in (0,1), 2: exactly 1
 * Filled when rto_set_histograms() was given arrays (caller-sized: 8 x n_objects, 8 x n_lights x 3). */
static uint64_t* g_hist_hits = NULL;
static uint64_t* g_hist_shadow = NULL;
void rto_set_histograms(uint64_t* hits, uint64_t* shadow) { g_hist_hits = hits; g_hist_shadow = shadow; }

static Col shade_ray(const OScene* S, V3 incidence_ray, int obj, const OHit* hit,
                     float n_i, float n_t, const int* stack, int nstack,
                     int ray_state, int depth, int level, Col bkg)
{
    const RtMaterial* m = obj_mat(S, obj);
    const int is_sphere = obj_is_sphere(S, obj);
    const int self_id = obj_id(S, obj);
    V3 N = hit->n;                                                   /* 785 */
    V3 I = vmulf(incidence_ray, -1.0f);                              /* 786 */
    Col tmp_specular = { 0.0f, 0.0f, 0.0f };
    float shadow_mask = 1.0f;                                        /* 788: r=g=b always */
    Col diffuse;
    Col tmp_reflection = { 0.0f, 0.0f, 0.0f };
    Col tmp_transparency = { 0.0f, 0.0f, 0.0f };
    float cosi = vdot(N, I);                                         /* 794 */
    const float bkg_eta = S->d->bkg_eta;

    S->st->nodes_total++;
    if (level < 8) S->st->nodes_per_level[level]++;
    if (g_hist_hits && level < 8) g_hist_hits[(size_t)level * (size_t)(S->nf + S->ns) + (size_t)obj]++;

    /* diffuse colour / texture, main.cpp:800-867 */
    int tex = obj_tex(S, obj);
    if (tex >= 0) {
        const RtTexture* T = &S->d->textures[tex];
        float width = (float)T->width, height = (float)T->height;
        if (is_sphere) {
            float v = (float)(acos((double)N.z) / M_PI);             /* 805 */
            float phi = (float)atan2((double)N.y, (double)N.x);      /* 806 */
            float u = mapf(phi, (float)-M_PI, (float)M_PI, 0.0f, 1.0f);   /* 808 */
            v = clampf(v, 0.0f, 1.0f); u = clampf(u, 0.0f, 1.0f);    /* 813-814 */
            int i = (int)clampf((float)round(((double)height - 1.0) * (double)v), 0.0f, (float)((double)height - 1.0));  /* 817 */
            int j = (int)clampf((float)round(((double)width - 1.0) * (double)u), 0.0f, (float)((double)width - 1.0));    /* 818 */
            diffuse = texel(S, tex, j, i);                           /* get(j, i, c): x = j, y = i */
        } else {
            const OFace* f = &S->f[obj];
            float a = hit->bary.x, b = hit->bary.y, g = hit->bary.z; /* face->barycentric_cords of this ray, 1372 */
            float u = (a * clampf(f->vt[0][0], 0.0f, 1.0f)) + (b * clampf(f->vt[1][0], 0.0f, 1.0f)) + (g * clampf(f->vt[2][0], 0.0f, 1.0f)); /* 834-837 */
            float v = (a * clampf(f->vt[0][1], 0.0f, 1.0f)) + (b * clampf(f->vt[1][1], 0.0f, 1.0f)) + (g * clampf(f->vt[2][1], 0.0f, 1.0f)); /* 838-841 */
            v = clampf(v, 0.0f, 1.0f); u = clampf(u, 0.0f, 1.0f);    /* 844-845 */
            int i = (int)clampf((float)round((double)((width - 1.0f) * u)), 0.0f, (float)((double)width - 1.0));   /* 852 */
            int j = (int)clampf((float)round((double)((height - 1.0f) * v)), 0.0f, (float)((double)height - 1.0)); /* 853 */
            diffuse = texel(S, tex, i, j);                           /* get(i, j, c): x = i, y = j */
        }
    } else {
        diffuse.r = m->od[0]; diffuse.g = m->od[1]; diffuse.b = m->od[2];   /* 866 */
    }

    if (cosi < 0.0 && is_sphere) {                                   /* 869-872: spheres only */
        N = vmulf(N, -1.0f);
        cosi = vdot(N, I);
    }

    /* lights and hard shadows, main.cpp:878-959; shadow_mask is NOT reset per light */
    Col specular_col = { m->os[0], m->os[1], m->os[2] };
    for (int li = 0; li < S->d->n_lights; li++) {
        const RtLight* lt = &S->d->lights[li];
        V3 lxyz = v3(lt->xyz[0], lt->xyz[1], lt->xyz[2]);
        V3 L;
        ShadowCtx sc; sc.S = S; sc.self_id = self_id; sc.tmin = S->eps; sc.mask = shadow_mask; sc.hits = 0;
        if (lt->w == 0) {
            L = vmulf(vnorm(lxyz), -1.0f);                           /* 887 */
            V3 ray = vmulf(lxyz, -1.0f);                             /* 895: NOT normalised */
            sc.use_tmax = 0; sc.tmax = 0.0f;
            S->st->rays_shadow++;
            trace_ray(S, hit->p, ray, visit_shadow, &sc);            /* 896 */
        } else {
            L = vnorm(vsub(lxyz, hit->p));                           /* 921 */
            V3 dp = vsub(hit->p, lxyz);
            float distance_to_light = sqrtf(vsum(v3(dp.x * dp.x, dp.y * dp.y, dp.z * dp.z)));   /* 922 */
            sc.use_tmax = 1; sc.tmax = distance_to_light;
            S->st->rays_shadow++;
            trace_ray(S, hit->p, L, visit_shadow, &sc);              /* 928 */
        }
        shadow_mask = sc.mask;
        if (g_hist_shadow && level < 8)
            g_hist_shadow[((size_t)level * (size_t)S->d->n_lights + (size_t)li) * 3 + (shadow_mask == 0.0f ? 0 : shadow_mask == 1.0f ? 2 : 1)]++;
        V3 H = vnorm(vadd(L, I));                                    /* 952 */
        float ndl = vdot(N, L), ndh = vdot(N, H);
        float mdl = (0.0f < ndl) ? ndl : 0.0f;                       /* std::max(0.0f, x): NaN -> 0 */
        float mdh = (0.0f < ndh) ? ndh : 0.0f;
        Col diffuse_component = cmulf(cmulf(diffuse, m->kd), mdl);   /* 953 */
        Col specular_component = cmulf(cmulf(specular_col, m->ks), powf(mdh, m->n));   /* 954 */
        Col lc = { lt->rgb[0], lt->rgb[1], lt->rgb[2] };
        Col sm = { shadow_mask, shadow_mask, shadow_mask };
        tmp_specular = cadd(tmp_specular, cmul(cmul(lc, sm), cadd(diffuse_component, specular_component)));   /* 955-958 */
    }

    float snells_ratio = n_i / n_t;                                  /* 961 */
    float critical_angle = asinf(n_t / n_i);                         /* 962 */
    float incidence_angle = acosf(cosi);                             /* 963 */
    int tir = (critical_angle < incidence_angle) && ((double)incidence_angle < (90.0 * M_PI / 180.0));   /* 964 */
    float F_0 = powf((n_t - n_i) / (n_t + n_i), 2.0f);               /* 965 */
    float F = (float)((double)F_0 + (1.0 - (double)F_0) * (double)powf((float)(1.0 - (double)cosi), 5.0f));   /* 966 */

    /* transmission, main.cpp:976-1089 */
    if (depth > 0 && !tir && m->opacity < 1.0 && m->eta > 0) {
        float root = sqrtf((float)(1.0 - ((double)powf(snells_ratio, 2.0f) * (1.0 - (double)powf(cosi, 2.0f)))));   /* 981-983 */
        V3 T = vadd(vmulf(vmulf(N, -1.0f), root), vmulf(vsub(vmulf(N, cosi), I), snells_ratio));   /* 979-987 */
        TransCtx tc; tc.tmin = S->eps; tc.best = FLT_MAX; tc.obj = -1; tc.S = S; tc.aborted = 0;
        tc.face_rule = (nstack > 0 && !is_sphere);                   /* 1000 */
        tc.back_id = nstack > 0 ? obj_id(S, stack[nstack - 1]) : 0;
        S->st->rays_transmit++;
        trace_ray(S, hit->p, T, visit_trans, &tc);                   /* 992 */
        if (!tc.aborted) {
            float w1 = (float)(1.0 - (double)F);                     /* 1083 / 1087 */
            float w2 = (float)(1.0 - (double)m->opacity);
            if (tc.obj >= 0) {
                int hobj = tc.obj;
                float new_ni, new_nt; int ns2[MAX_STACK]; int n2 = nstack; int new_state;
                memcpy(ns2, stack, sizeof(int) * (size_t)nstack);
                if (ray_state == STATE_ENTERING) {                   /* 1024 */
                    if (obj_id(S, hobj) == self_id) {                /* 1026 */
                        new_state = STATE_EXITING;
                        /* back()/pop_back() on an empty vector is UB in the reference;
                         * unreachable on sane scenes, kept memory-safe here */
                        new_ni = n2 > 0 ? obj_mat(S, ns2[n2 - 1])->eta : bkg_eta;   /* 1028 */
                        if (n2 > 0) n2--;                            /* 1029 */
                        new_nt = n2 > 0 ? obj_mat(S, ns2[n2 - 1])->eta : bkg_eta;   /* 1030 */
                        if (n2 > 0) n2--;                            /* 1031 */
                    } else {
                        new_state = STATE_ENTERING;                  /* 1035-1038 */
                        new_ni = n_t; new_nt = obj_mat(S, hobj)->eta;
                        ns2[n2++] = hobj;
                    }
                } else {                                             /* 1043 */
                    if (n2 > 0) {
                        if (!in_stack(ns2, n2, hobj)) {              /* 1047-1051 */
                            new_state = STATE_ENTERING;
                            new_ni = n_t; new_nt = obj_mat(S, hobj)->eta;
                            ns2[n2++] = hobj;
                        } else {                                     /* 1054-1058 */
                            new_state = STATE_EXITING;
                            new_ni = n_t; new_nt = obj_mat(S, ns2[n2 - 1])->eta;
                            n2--;
                        }
                    } else {                                         /* 1062-1066 */
                        new_state = STATE_ENTERING;
                        new_ni = bkg_eta; new_nt = obj_mat(S, hobj)->eta;
                        ns2[0] = hobj; n2 = 1;
                    }
                }
                Col child = shade_ray(S, T, hobj, &tc.hit, new_ni, new_nt, ns2, n2, new_state, depth - 1, level + 1, bkg);   /* 1072 */
                tmp_transparency = cmulf(cmulf(child, w1), w2);      /* 1083 */
            } else {
                tmp_transparency = cmulf(cmulf(bkg, w1), w2);        /* 1087 */
            }
        }
    }
    /* SKIP_TRANS: (1091) */

    /* reflection, main.cpp:1103-1200 */
    F_0 = powf((m->eta - 1) / (m->eta + 1), 2.0f);                   /* 1103 */
    F = (float)((double)F_0 + (1.0 - (double)F_0) * (double)powf((float)(1.0 - (double)cosi), 5.0f));   /* 1104 */
    if (depth > 0 && F != 0.0 && m->ks > 0.0) {                      /* 1105 */
        V3 R = vsub(vmulf(N, (float)(2.0 * (double)cosi)), I);       /* 1108 */
        ClosestCtx cc; cc.tmin = S->eps; cc.best = FLT_MAX; cc.obj = -1;
        S->st->rays_reflect++;
        trace_ray(S, hit->p, R, visit_closest, &cc);                 /* 1113 */
        if (cc.obj >= 0) {
            int hobj = cc.obj;
            float new_ni, new_nt; int ns2[MAX_STACK]; int n2 = nstack; int new_state;
            memcpy(ns2, stack, sizeof(int) * (size_t)nstack);
            if (ray_state == STATE_ENTERING) {                       /* 1137 */
                if (n2 > 0) {
                    if (!in_stack(ns2, n2, hobj)) {                  /* 1141-1145 */
                        new_state = STATE_ENTERING;
                        new_ni = n_i; new_nt = obj_mat(S, hobj)->eta;
                        ns2[n2++] = obj;                             /* pushes the INCIDENCE object */
                    } else {                                         /* 1148-1152 */
                        new_state = STATE_ENTERING;
                        new_ni = n_i; new_nt = obj_mat(S, ns2[n2 - 1])->eta;
                        n2--;
                    }
                } else {                                             /* 1156-1160 */
                    new_state = STATE_ENTERING;
                    new_ni = n_i; new_nt = obj_mat(S, hobj)->eta;
                    ns2[0] = hobj; n2 = 1;
                }
            } else {                                                 /* 1166 */
                if (obj_id(S, hobj) == self_id) {                    /* 1168-1171 */
                    new_state = STATE_EXITING;
                    new_ni = n_i; new_nt = n_t;
                } else {                                             /* 1174-1178 */
                    new_state = STATE_ENTERING;
                    new_ni = n_i; new_nt = obj_mat(S, hobj)->eta;
                    ns2[n2++] = hobj;
                }
            }
            Col child = shade_ray(S, R, hobj, &cc.hit, new_ni, new_nt, ns2, n2, new_state, depth - 1, level + 1, bkg);   /* 1184 */
            tmp_reflection = cmulf(child, F);                        /* 1194 */
        } else {
            tmp_reflection = cmulf(bkg, F);                          /* 1198 */
        }
    }

    /* main.cpp:1206 */
    return cadd(cadd(cadd(cmulf(diffuse, m->ka), tmp_specular), tmp_transparency), tmp_reflection);
}

/* ---- create_view_window_and_ray_trace, main.cpp:670-767 -------------------- */
typedef struct { V3 eye, ul, delta_h, delta_v; } OCamera;

static void make_camera(const RtSceneDesc* d, OCamera* cam)
{
    V3 view_origin = v3(d->eye[0], d->eye[1], d->eye[2]);
    /* the caller normalises both (main.cpp:607); camera_normalised: they already are what the callee receives */
    V3 view_direction = v3(d->viewdir[0], d->viewdir[1], d->viewdir[2]);
    V3 view_up = v3(d->updir[0], d->updir[1], d->updir[2]);
    if (!d->camera_normalised) { view_direction = vnorm(view_direction); view_up = vnorm(view_up); }
    float res_h = (float)d->height, res_w = (float)d->width;         /* passed as float, 670 */
    V3 u = vnorm(vcross(view_direction, view_up));                   /* 677 */
    V3 v = vcross(u, view_direction);                                /* 684 */
    float aspect_ratio = res_w / res_h;                              /* 689 */
    float w = (float)((double)2.0f * VIEW_DIST * tan((0.5 * (double)d->hfov_deg) * M_PI / (double)180.0f));   /* 690 */
    float h = w / aspect_ratio;                                      /* 691 */
    V3 n = view_direction;
    V3 nd = vmulf(n, (float)VIEW_DIST);
    V3 ul = vadd(vsub(vadd(view_origin, nd), vmulf(u, w / 2.0f)), vmulf(v, h / 2.0f));   /* 697 */
    V3 ur = vadd(vadd(vadd(view_origin, nd), vmulf(u, w / 2.0f)), vmulf(v, h / 2.0f));   /* 698 */
    V3 ll = vsub(vsub(vadd(view_origin, nd), vmulf(u, w / 2.0f)), vmulf(v, h / 2.0f));   /* 699 */
    cam->eye = view_origin; cam->ul = ul;
    cam->delta_h = vdivf(vsub(ur, ul), res_w - 1.0f);                /* 709 */
    cam->delta_v = vdivf(vsub(ll, ul), res_h - 1.0f);                /* 710 */
}

static void render_pixel(const OScene* S, const OCamera* cam, int i, int j, int32_t out[3], int32_t* primary_id)
{
    const RtSceneDesc* d = S->d;
    Col bkg = { d->bkg[0], d->bkg[1], d->bkg[2] };
    V3 p = vadd(vadd(cam->ul, vmulf(cam->delta_h, (float)j)), vmulf(cam->delta_v, (float)i));   /* 720 */
    Col pixel_color = bkg;                                           /* 721 */
    V3 ray = vnorm(vsub(p, cam->eye));                               /* 728 */
    ClosestCtx cc; cc.tmin = 0.0f; cc.best = FLT_MAX; cc.obj = -1;   /* 736: distance > 0.0f */
    S->st->rays_primary++;
    trace_ray(S, cam->eye, ray, visit_closest, &cc);                 /* 729 */
    if (cc.obj >= 0) {
        int stack[MAX_STACK]; stack[0] = cc.obj;                     /* 746 */
        pixel_color = shade_ray(S, ray, cc.obj, &cc.hit, d->bkg_eta, obj_mat(S, cc.obj)->eta,
                                stack, 1, STATE_ENTERING, S->max_depth, 0, bkg);   /* 747-757 */
    }
    if (primary_id) *primary_id = cc.obj >= 0 ? obj_id(S, cc.obj) : 0;
    float c[3] = { pixel_color.r, pixel_color.g, pixel_color.b };
    for (int k = 0; k < 3; k++) {
        float q = mapf(c[k], 0.0f, 1.0f, 0.0f, 255.0f);              /* 760-762 */
        /* static_cast<int>(NaN / out of range) is cvttss2si -> INT_MIN on x86-64 */
        out[k] = (q != q || q >= 2147483648.0f || q < -2147483648.0f) ? RTO_NAN_VALUE : (int32_t)q;
    }
}

int rto_render(const RtSceneDesc* desc, const RtoParams* params, int64_t n_pixels,
               const int64_t* pixel_index, int32_t* out_raw, int32_t* out_primary_id, RtoStats* stats)
{
    OScene S; OCamera cam;
    int rc = prepare(&S, desc, params, stats);
    if (rc != RT_OK) { unprepare(&S); return rc; }
    if (desc->width < 2 || desc->height < 2 || !out_raw) { unprepare(&S); return RT_E_INVALID; }
    if (!pixel_index && n_pixels != (int64_t)desc->width * desc->height) { unprepare(&S); return RT_E_INVALID; }
    if (S.max_depth + 2 > MAX_STACK) { unprepare(&S); return RT_E_INVALID; }
    make_camera(desc, &cam);
    for (int64_t k = 0; k < n_pixels; k++) {
        int64_t idx = pixel_index ? pixel_index[k] : k;
        int i = (int)(idx / desc->width), j = (int)(idx % desc->width);
        render_pixel(&S, &cam, i, j, out_raw + 3 * k, out_primary_id ? out_primary_id + k : NULL);
    }
    unprepare(&S);
    return RT_OK;
}

/* The primary rays of n pixels (main.cpp:720, 728): origin = eye, direction = (p - eye).norm(). */
int rto_primary_rays(const RtSceneDesc* desc, int64_t n_pixels, const int64_t* pixel_index,
                     float* out_origins, float* out_dirs)
{
    OCamera cam;
    if (!desc || desc->width < 2 || desc->height < 2 || !pixel_index || !out_origins || !out_dirs) return RT_E_INVALID;
    make_camera(desc, &cam);
    for (int64_t k = 0; k < n_pixels; k++) {
        int i = (int)(pixel_index[k] / desc->width), j = (int)(pixel_index[k] % desc->width);
        V3 p = vadd(vadd(cam.ul, vmulf(cam.delta_h, (float)j)), vmulf(cam.delta_v, (float)i));   /* 720 */
        V3 ray = vnorm(vsub(p, cam.eye));                                                        /* 728 */
        out_origins[3 * k] = cam.eye.x; out_origins[3 * k + 1] = cam.eye.y; out_origins[3 * k + 2] = cam.eye.z;
        out_dirs[3 * k] = ray.x; out_dirs[3 * k + 1] = ray.y; out_dirs[3 * k + 2] = ray.z;
    }
    return RT_OK;
}

int rto_trace_closest(const RtSceneDesc* desc, int32_t n, const float* origins, const float* dirs,
                      float tmin, int32_t* out_id, float* out_t, float* out_normal, float* out_bary)
{
    OScene S;
    int rc = prepare(&S, desc, NULL, NULL);
    if (rc != RT_OK) { unprepare(&S); return rc; }
    for (int32_t k = 0; k < n; k++) {
        ClosestCtx cc; cc.tmin = tmin; cc.best = FLT_MAX; cc.obj = -1;
        memset(&cc.hit, 0, sizeof cc.hit);
        trace_ray(&S, v3(origins[3 * k], origins[3 * k + 1], origins[3 * k + 2]),
                  v3(dirs[3 * k], dirs[3 * k + 1], dirs[3 * k + 2]), visit_closest, &cc);
        out_id[k] = cc.obj >= 0 ? obj_id(&S, cc.obj) : 0;
        out_t[k] = cc.obj >= 0 ? cc.hit.t : 0.0f;
        if (out_normal) { out_normal[3 * k] = cc.hit.n.x; out_normal[3 * k + 1] = cc.hit.n.y; out_normal[3 * k + 2] = cc.hit.n.z; }
        if (out_bary) { out_bary[3 * k] = cc.hit.bary.x; out_bary[3 * k + 1] = cc.hit.bary.y; out_bary[3 * k + 2] = cc.hit.bary.z; }
    }
    unprepare(&S);
    return RT_OK;
}

int rto_trace_shadow(const RtSceneDesc* desc, int32_t n, const float* origins, const float* dirs,
                     const int32_t* self_id, float tmin, const float* tmax, float* out_mask, int32_t* out_hits)
{
    OScene S;
    int rc = prepare(&S, desc, NULL, NULL);
    if (rc != RT_OK) { unprepare(&S); return rc; }
    for (int32_t k = 0; k < n; k++) {
        ShadowCtx sc; sc.S = &S; sc.self_id = self_id ? self_id[k] : 0; sc.tmin = tmin;
        sc.use_tmax = tmax != NULL; sc.tmax = tmax ? tmax[k] : 0.0f; sc.mask = 1.0f; sc.hits = 0;
        trace_ray(&S, v3(origins[3 * k], origins[3 * k + 1], origins[3 * k + 2]),
                  v3(dirs[3 * k], dirs[3 * k + 1], dirs[3 * k + 2]), visit_shadow, &sc);
        out_mask[k] = sc.mask;
        if (out_hits) out_hits[k] = sc.hits;
    }
    unprepare(&S);
    return RT_OK;
}

void rto_raw_to_rgb8(int64_t n_pixels, const int32_t* raw, const int32_t* primary_id, uint8_t* rgb, uint8_t* flags)
{
    for (int64_t k = 0; k < n_pixels; k++) {
        uint8_t fl = 0;
        for (int c = 0; c < 3; c++) {
            int32_t v = raw[3 * k + c];
            if (v == RTO_NAN_VALUE) { fl |= (uint8_t)(1u << c); v = 0; }
            rgb[3 * k + c] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
        if (primary_id && primary_id[k] == 0) fl |= 8u;
        if (flags) flags[k] = fl;
    }
}
