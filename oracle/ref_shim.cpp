/*
 * ref_shim.cpp — builds the REFERENCE ITSELF into oracle/_ref/libref.so.
 * TEST INFRASTRUCTURE ONLY; never linked or loaded by the product.
 *
 * The reference translation unit is compiled where it lies
 * (-I/root/reference, `#include "main.cpp"`), unmodified; no reference source
 * is copied into this repository.  Its `main` is renamed so that the shim can
 * (a) fill the global `Globals environment` (main.cpp:58) from an RtSceneDesc
 *     exactly as the parser would (main.cpp:328-552) and call the reference's
 *     own create_view_window_and_ray_trace / ShadeRay / TraceRay, and
 * (b) still run the whole reference program (ref_main) for parser parity.
 *
 * Build: see oracle/Makefile (g++ -std=c++20 -O2 -ffp-contract=off).
 */
#include "../include/rt.h"

#include <chrono>
#include <cstring>
#include <limits>

#define main reference_main
#include "main.cpp"            /* /root/reference/main.cpp via -I */
#undef main
#undef d                       /* src/config.h:8 `#define d 5.0` poisons identifiers */

namespace {

std::vector<Texture*> g_textures;

void shim_reset()
{
    /* the reference never frees anything (SURVEY 8b "Ownership"); neither do we */
    environment.scene_object_infos.clear();
    environment.faces.clear();
    environment.spheres.clear();
    environment.scene_lights.clear();
    environment.commands.clear();
    environment.other.clear();
    g_textures.clear();
}

Material shim_material(const RtMaterial& m)
{
    Material out;
    out.diffuse = { m.od[0], m.od[1], m.od[2] };
    out.specular = { m.os[0], m.os[1], m.os[2] };
    out.ka = m.ka; out.kd = m.kd; out.ks = m.ks; out.n = m.n;
    out.opacity = m.opacity; out.refraction_index = m.eta;
    return out;
}

Vector3 shim_v3(const float* p) { return Vector3{ p[0], p[1], p[2] }; }

int shim_load(const RtSceneDesc* s, int max_depth, float eps)
{
    shim_reset();
    environment.other["recursion_depth"] = (float)max_depth;       /* main.cpp:100 */
    environment.other["epsilon"] = eps;                             /* main.cpp:101 */
    /* main.cpp:261-264: the key exists only when bkgcolor had a 4th argument;
     * an absent key reads as 0.0f, so writing 0.0f is equivalent. */
    environment.other["bkg_refraction_index"] = s->bkg_eta;

    for (int t = 0; t < s->n_textures; t++) {                      /* utility.h:115-131 */
        const RtTexture& in = s->textures[t];
        Texture* tex = new Texture((float)in.width, (float)in.height);
        for (int y = 0; y < in.height; y++)
            for (int x = 0; x < in.width; x++)
                for (int c = 0; c < 3; c++)
                    tex->image->set(x, y, c, in.rgb[((size_t)y * in.width + x) * 3 + c]);
        g_textures.push_back(tex);
    }
    for (int i = 0; i < s->n_spheres; i++) {                       /* main.cpp:328-377 */
        const RtSphere& in = s->spheres[i];
        Sphere* sp = new Sphere();
        SceneObjectInfo* info = new SceneObjectInfo();
        info->id = (unsigned)in.id; info->type = "sphere";
        sp->radius = in.radius; sp->center = shim_v3(in.center);
        info->material = shim_material(s->materials[in.material]);
        info->has_texture = in.texture >= 0;
        info->texture = in.texture >= 0 ? g_textures[in.texture] : nullptr;
        sp->object_info = info;
        environment.scene_object_infos["sphere"].push_back(info);
        environment.spheres[info->id] = sp;
    }
    for (int i = 0; i < s->n_faces; i++) {                         /* main.cpp:470-552 */
        const RtFace& in = s->faces[i];
        Face* tri = new Face();
        SceneObjectInfo* info = new SceneObjectInfo();
        info->id = (unsigned)in.id; info->type = "face";
        for (int k = 0; k < 3; k++) {
            tri->vertex[k] = shim_v3(in.v[k]);
            tri->vertex_normal[k] = shim_v3(in.vn[k]);
            tri->texture_coords[k] = Point{ in.vt[k][0], in.vt[k][1] };
        }
        tri->smooth_shading = in.smooth != 0;
        info->material = shim_material(s->materials[in.material]);
        info->has_texture = in.texture >= 0;
        info->texture = in.texture >= 0 ? g_textures[in.texture] : nullptr;
        Vector3 e1 = tri->vertex[1] - tri->vertex[0];               /* main.cpp:537-539 */
        Vector3 e2 = tri->vertex[2] - tri->vertex[0];
        tri->surface_normal = e1.cross(e2).norm();
        tri->object_info = info;
        environment.faces[info->id] = tri;
        environment.scene_object_infos["face"].push_back(info);
    }
    for (int i = 0; i < s->n_lights; i++) {                        /* main.cpp:378-411 */
        const RtLight& in = s->lights[i];
        Light l;
        l.position = Vector3{ 0, 0, 0 }; l.direction = Vector3{ 0, 0, 0 };
        l.w = in.w;
        if (l.w == 0) l.direction = shim_v3(in.xyz); else l.position = shim_v3(in.xyz);
        l.color = { in.rgb[0], in.rgb[1], in.rgb[2] };
        environment.scene_lights.push_back(l);
    }
    return 0;
}

int32_t shim_to_raw(size_t v)
{
    /* matt holds size_t((int)x): NaN became 0xFFFFFFFF80000000 */
    return (int32_t)(int64_t)v;
}

}  // namespace

extern "C" {

/* Whole frame through the reference's own create_view_window_and_ray_trace. */
int ref_render(const RtSceneDesc* s, int max_depth, float eps, int32_t* out_raw, double* render_seconds)
{
    shim_load(s, max_depth, eps);
    Vector3 vo = shim_v3(s->eye), vd = shim_v3(s->viewdir), vu = shim_v3(s->updir);
    Color bkg = { s->bkg[0], s->bkg[1], s->bkg[2] };
    auto t0 = std::chrono::steady_clock::now();
    Mat3D matt = create_view_window_and_ray_trace(vo, vd.norm(), vu.norm(), s->hfov_deg,
                                                  s->height, s->width, bkg);   /* main.cpp:607 */
    auto t1 = std::chrono::steady_clock::now();
    if (render_seconds) *render_seconds = std::chrono::duration<double>(t1 - t0).count();
    for (int i = 0; i < s->height; i++)
        for (int j = 0; j < s->width; j++)
            for (int c = 0; c < 3; c++)
                out_raw[((size_t)i * s->width + j) * 3 + c] = shim_to_raw(matt(i, j, c));
    return 0;
}

/* A subset of pixels: the per-pixel body of main.cpp:720-762 around the
 * reference's own TraceRay / ShadeRay.  Used to time the reference on scenes
 * whose full frame would take hours, and to dump primary hit ids.  The camera
 * lines restate main.cpp:677-710 (checked against ref_render by the tests). */
int ref_render_pixels(const RtSceneDesc* s, int max_depth, float eps, int64_t n, const int64_t* pixel_index,
                      int32_t* out_raw, int32_t* out_primary_id, double* render_seconds)
{
    shim_load(s, max_depth, eps);
    Vector3 view_origin = shim_v3(s->eye);
    Vector3 view_direction = shim_v3(s->viewdir).norm(), view_up = shim_v3(s->updir).norm();
    Color background_color = { s->bkg[0], s->bkg[1], s->bkg[2] };
    float res_h = (float)s->height, res_w = (float)s->width, fov_h = s->hfov_deg;
    Vector3 u = (view_direction).cross(view_up).norm();
    Vector3 v = u.cross(view_direction);
    float aspect_ratio = res_w / res_h;
    float w = 2.0f * 5.0 * tan((0.5 * fov_h) * M_PI / 180.0f);
    float h = w / aspect_ratio;
    Vector3 n_ = view_direction;
    Vector3 ul = view_origin + n_ * 5.0 - u * (w / 2.0f) + v * (h / 2.0f);
    Vector3 ur = view_origin + n_ * 5.0 + u * (w / 2.0f) + v * (h / 2.0f);
    Vector3 ll = view_origin + n_ * 5.0 - u * (w / 2.0f) - v * (h / 2.0f);
    Vector3 delta_h = (ur - ul) / (res_w - 1.0f);
    Vector3 delta_v = (ll - ul) / (res_h - 1.0f);
    auto t0 = std::chrono::steady_clock::now();
    for (int64_t k = 0; k < n; k++) {
        int i = (int)(pixel_index[k] / s->width), j = (int)(pixel_index[k] % s->width);
        Vector3 p = ul + (delta_h * static_cast<float>(j)) + (delta_v * static_cast<float>(i));
        Color pixel_color = background_color;
        float min_distance = std::numeric_limits<float>::max();
        Vector3 ray = (p - view_origin).norm();
        std::vector<ObjectIntersections> results = TraceRay(view_origin, ray);
        Intersection min_intersection;
        SceneObjectInfo* hit_object = nullptr;
        for (auto& oi : results)
            for (auto& is : oi.intersections)
                if (is.distance > 0.0f && is.distance < min_distance) {
                    min_distance = is.distance; hit_object = oi.object_info; min_intersection = is;
                }
        if (hit_object != nullptr) {
            std::vector<SceneObjectInfo*> stack = { hit_object };
            pixel_color = ShadeRay(ray, hit_object, min_intersection, environment.other["bkg_refraction_index"],
                                   hit_object->material.refraction_index, stack, RayState::ENTERING,
                                   environment.other["recursion_depth"], background_color);
        }
        if (out_primary_id) out_primary_id[k] = hit_object ? (int32_t)hit_object->id : 0;
        size_t q[3] = { (size_t) static_cast<int>(map(pixel_color.r, 0, 1.0, MIN_PIXEL_VALUE, MAX_PIXEL_VALUE)),
                        (size_t) static_cast<int>(map(pixel_color.g, 0, 1.0, MIN_PIXEL_VALUE, MAX_PIXEL_VALUE)),
                        (size_t) static_cast<int>(map(pixel_color.b, 0, 1.0, MIN_PIXEL_VALUE, MAX_PIXEL_VALUE)) };
        for (int c = 0; c < 3; c++) out_raw[3 * k + c] = shim_to_raw(q[c]);
    }
    auto t1 = std::chrono::steady_clock::now();
    if (render_seconds) *render_seconds = std::chrono::duration<double>(t1 - t0).count();
    return 0;
}

/* The reference's TraceRay + the closest-hit loop of main.cpp:1109-1124. */
int ref_trace_closest(const RtSceneDesc* s, int32_t n, const float* origins, const float* dirs, float tmin,
                      int32_t* out_id, float* out_t, float* out_normal, float* out_bary)
{
    shim_load(s, 4, 1.0e-3f);
    for (int32_t k = 0; k < n; k++) {
        Vector3 o = shim_v3(origins + 3 * k), r = shim_v3(dirs + 3 * k);
        float min_distance = std::numeric_limits<float>::max();
        Intersection best{}; SceneObjectInfo* obj = nullptr;
        Vector3 bary{ 0, 0, 0 };
        for (auto& [object, intersections] : TraceRay(o, r))
            for (auto& is : intersections)
                if (is.distance > tmin && is.distance < min_distance) {
                    min_distance = is.distance; obj = object; best = is;
                }
        /* face->barycentric_cords holds the values of the last TraceRay on this ray (main.cpp:1372) */
        if (obj && obj->type == "face") bary = environment.faces[obj->id]->barycentric_cords;
        out_id[k] = obj ? (int32_t)obj->id : 0;
        out_t[k] = obj ? best.distance : 0.0f;
        if (out_normal) { out_normal[3 * k] = best.normal.x; out_normal[3 * k + 1] = best.normal.y; out_normal[3 * k + 2] = best.normal.z; }
        if (out_bary) { out_bary[3 * k] = bary.x; out_bary[3 * k + 1] = bary.y; out_bary[3 * k + 2] = bary.z; }
    }
    return 0;
}

/* The reference's TraceRay + the shadow loops of main.cpp:898-912 / 930-949. */
int ref_trace_shadow(const RtSceneDesc* s, int32_t n, const float* origins, const float* dirs,
                     const int32_t* self_id, float tmin, const float* tmax, float* out_mask, int32_t* out_hits)
{
    shim_load(s, 4, 1.0e-3f);
    for (int32_t k = 0; k < n; k++) {
        Vector3 o = shim_v3(origins + 3 * k), r = shim_v3(dirs + 3 * k);
        Color shadow_mask = { 1.0, 1.0, 1.0 };
        int hits = 0;
        for (auto [object, intersections] : TraceRay(o, r)) {
            if (self_id && (int32_t)object->id == self_id[k]) continue;
            for (auto is : intersections)
                if (is.distance > tmin && (!tmax || is.distance < tmax[k])) {
                    shadow_mask = shadow_mask * (1.0 - object->material.opacity);
                    hits++;
                }
        }
        out_mask[k] = shadow_mask.r;
        if (out_hits) out_hits[k] = hits;
    }
    return 0;
}

/* The whole reference program: `raytracer scene.txt` -> scene.ppm next to it. */
int ref_main(int argc, char** argv)
{
    shim_reset();
    return reference_main(argc, argv);
}

}  /* extern "C" */
