// shade.cuh — one ShadeRay invocation (main.cpp:783-1207) as a wavefront node:
// texture lookup, per-light hard shadows + Blinn-Phong, Fresnel terms, and the two
// rays it spawns.  The recursion of the reference becomes: this function leaves a
// node record (base colour + the weights of its two child terms) and up to two
// rays for the next generation; combine.cuh folds the children back bottom-up
// with the reference's clamped operators.
#pragma once

#include "trace.cuh"

namespace rt {

struct Material {
    Col od, os;
    float ka, kd, ks, n, opacity, eta;
};

RT_DEV Material load_material(const SceneView& S, int m)
{
    const float4 a = S.mat[3 * m], b = S.mat[3 * m + 1], c = S.mat[3 * m + 2];
    Material r;
    r.od = mkc(a.x, a.y, a.z); r.ka = a.w;
    r.os = mkc(b.x, b.y, b.z); r.kd = b.w;
    r.ks = c.x; r.n = c.y; r.opacity = c.z; r.eta = c.w;
    return r;
}

RT_DEV float material_eta(const SceneView& S, int obj) { return S.mat[3 * S.obj[obj].x + 2].w; }

// Nearest-texel fetch, texel/255 through map() (main.cpp:823-825, 858-860).
RT_DEV Col fetch_texel(const SceneView& S, const int4 td, int x, int y)
{
    x = min(max(x, 0), td.y - 1);   // static_cast<int>(NaN) indexes out of range in the reference (UB)
    y = min(max(y, 0), td.z - 1);
    const uchar4 px = S.texels[(size_t)td.x + (size_t)y * (size_t)td.y + (size_t)x];
    return mkc(mapf((float)px.x, 0.0f, 255.0f, 0.0f, 1.0f), mapf((float)px.y, 0.0f, 255.0f, 0.0f, 1.0f),
               mapf((float)px.z, 0.0f, 255.0f, 0.0f, 1.0f));
}

// Diffuse colour of the hit: material Od or the texture (main.cpp:800-867).
RT_DEV Col diffuse_colour(const SceneView& S, const int4 oi, const Material& m, int obj, const Surface& sf)
{
    if (oi.y < 0) return m.od;                                                   // 866
    const int4 td = S.tex_desc[oi.y];
    const float width = (float)td.y, height = (float)td.z;
    if (obj >= S.nf) {
        // sphere: lat-long from the UNFLIPPED normal, double acos/atan2/round   805-818
        float v = (float)(acos((double)sf.n.z) / 3.14159265358979323846);
        const float phi = (float)atan2((double)sf.n.y, (double)sf.n.x);
        const float pif = 3.14159274101257324f;                                  // (float)M_PI
        float u = mapf(phi, -pif, pif, 0.0f, 1.0f);                              // 808
        v = clampf(v, 0.0f, 1.0f); u = clampf(u, 0.0f, 1.0f);                    // 813-814
        const int i = ref_f2i(clampf((float)round(dmul(dsub((double)height, 1.0), (double)v)), 0.0f, (float)((double)height - 1.0)));
        const int j = ref_f2i(clampf((float)round(dmul(dsub((double)width, 1.0), (double)u)), 0.0f, (float)((double)width - 1.0)));
        return fetch_texel(S, td, j, i);                                         // get(j, i, c)
    }
    // face: barycentric blend of the clamped vertex UVs, float products        834-853
    const float4* fa = S.face_attr + (size_t)obj * kFaceAttrWords;
    const float4 a0 = fa[0], a1 = fa[1], a2 = fa[2], a3 = fa[3];
    const float t0x = clampf(a0.w, 0.0f, 1.0f), t0y = clampf(a1.w, 0.0f, 1.0f);
    const float t1x = clampf(a2.w, 0.0f, 1.0f), t1y = clampf(a3.x, 0.0f, 1.0f);
    const float t2x = clampf(a3.y, 0.0f, 1.0f), t2y = clampf(a3.z, 0.0f, 1.0f);
    float u = fadd(fadd(fmul(sf.a, t0x), fmul(sf.b, t1x)), fmul(sf.g, t2x));
    float v = fadd(fadd(fmul(sf.a, t0y), fmul(sf.b, t1y)), fmul(sf.g, t2y));
    v = clampf(v, 0.0f, 1.0f); u = clampf(u, 0.0f, 1.0f);                        // 844-845
    const int i = ref_f2i(clampf((float)round((double)fmul(fsub(width, 1.0f), u)), 0.0f, (float)((double)width - 1.0)));
    const int j = ref_f2i(clampf((float)round((double)fmul(fsub(height, 1.0f), v)), 0.0f, (float)((double)height - 1.0)));
    return fetch_texel(S, td, i, j);                                             // get(i, j, c)
}

// Schlick term shared by both Fresnel evaluations (main.cpp:966, 1104):
// F = F_0 + (1.0 - F_0) * powf(1.0 - cos, 5.0), the sum and product in double.
RT_DEV float schlick(float F0, float cosi)
{
    const float base = (float)dsub(1.0, (double)cosi);
    return (float)dadd((double)F0, dmul(dsub(1.0, (double)F0), (double)ref_pow5f(base)));
}

// What a node hands to the next generation.
struct Spawn {
    bool want_t, want_r;
    V3 T, R;
};

// Where the shadow product of one light comes from.  DirectShadow traces the shadow ray
// in this thread (brute-force path: all lanes walk the same primitive list);
// RecordedShadow (wavefront_bvh.cuh) reads what the shadow kernel left.
template <class Accel>
struct DirectShadow {
    const Accel& acc;
    RT_DEV void operator()(int, V3 p, V3 dir, int self, float eps, float tmax, bool use_tmax, float& mask) const
    {
        acc.shadow(p, dir, self, eps, tmax, use_tmax, mask);
    }
};

// ShadeRay before its recursive calls, in two independent halves (the LBVH pipeline runs them
// as separate kernels so that the next generation's rays do not wait for this one's shadows):
//
//   shade_colour : texture, per-light shadows + Blinn-Phong -> diffuse*ka + tmp_specular
//                  (main.cpp:800-958 and the first two terms of 1206)
//   shade_spawn  : Fresnel terms, TIR test, the transmission and reflection rays and the
//                  weights of their terms (main.cpp:961-987, 1083/1087, 1103-1108)
//
// Both flip the normal of a sphere hit from inside the same way (main.cpp:869-872).
template <class ShadowSrc>
RT_DEV Col shade_colour(const SceneView& S, const ShadowSrc& shadow, V3 incidence_ray, int obj, const Surface& sf, float eps)
{
    const int4 oi = S.obj[obj];
    const Material m = load_material(S, oi.x);
    const bool is_sphere = obj >= S.nf;
    V3 N = sf.n;                                                                 // 785
    const V3 I = vneg(incidence_ray);                                            // 786
    float cosi = vdot(N, I);                                                     // 794
    const Col diffuse = diffuse_colour(S, oi, m, obj, sf);
    if (cosi < 0.0f && is_sphere) { N = vneg(N); cosi = vdot(N, I); }            // 869-872

    Col tmp_specular = mkc(0.0f, 0.0f, 0.0f);
    float shadow_mask = 1.0f;                                                    // 788, never reset (A5.1)
    for (int li = 0; li < S.nl; li++) {                                          // 878
        // Once the carried mask is exactly 0 it stays 0 (every factor is in [0,1]) and each
        // remaining light adds light.color * 0 * (finite) = +0 to a clamped sum: nothing
        // left to do.  Only taken when upload found all material / light values finite and
        // the exponents small enough that powf cannot overflow (S.zero_mask_exit).
        if (S.zero_mask_exit && shadow_mask == 0.0f) break;
        const float4 l0 = S.light[3 * li], l1 = S.light[3 * li + 1];
        V3 L;
        if (l0.w == 0.0f) {
            L = ld3(S.light[3 * li + 2]);                                        // 887 (hoisted)
            shadow(li, sf.p, vneg(ld3(l0)), obj, eps, 0.0f, false, shadow_mask);  // 895-912: un-normalised ray
        } else {
            L = vnorm(vsub(ld3(l0), sf.p));                                      // 921
            const V3 dp = vsub(sf.p, ld3(l0));
            const float dist = fsqrt(vsum3(fmul(dp.x, dp.x), fmul(dp.y, dp.y), fmul(dp.z, dp.z)));   // 922
            shadow(li, sf.p, L, obj, eps, dist, true, shadow_mask);               // 928-949
        }
        const V3 H = vnorm(vadd(L, I));                                          // 952
        const float mdl = max0(vdot(N, L)), mdh = max0(vdot(N, H));
        const Col dc = cmulf(cmulf(diffuse, m.kd), mdl);                         // 953
        const Col sc = cmulf(cmulf(m.os, m.ks), ref_powf(mdh, m.n));             // 954
        const Col lc = mkc(l1.x, l1.y, l1.z);
        const Col sm = mkc(shadow_mask, shadow_mask, shadow_mask);
        tmp_specular = cadd(tmp_specular, cmul(cmul(lc, sm), cadd(dc, sc)));     // 955-958
    }
    return cadd(cmulf(diffuse, m.ka), tmp_specular);                             // first two terms of 1206
}

RT_DEV void shade_spawn(const SceneView& S, V3 incidence_ray, int obj, const Surface& sf, float n_i, float n_t,
                        int depth_left, float& tw1, float& tw2, float& fr, Spawn& sp)
{
    const int4 oi = S.obj[obj];
    const Material m = load_material(S, oi.x);
    const bool is_sphere = obj >= S.nf;
    V3 N = sf.n;                                                                 // 785
    const V3 I = vneg(incidence_ray);                                            // 786
    float cosi = vdot(N, I);                                                     // 794
    if (cosi < 0.0f && is_sphere) { N = vneg(N); cosi = vdot(N, I); }            // 869-872

    const float snell = fdiv(n_i, n_t);                                          // 961
    // 962-964: critical = asinf(n_t / n_i); incidence = acosf(cos); tir = critical < incidence && incidence < pi/2.
    // Two double libm calls per node, decided without them unless the node sits within 1e-4 of a boundary:
    // asin(x) < acos(c)  <=>  x < 0, or x^2 + c^2 < 1 (x in [0,1], c in [0,1]); NaN (|x| > 1, c > 1) compares false;
    // the two float roundings move the angles by < 1.3e-7, |x^2 + c^2 - 1| >= 1e-4 keeps them >= 1e-4 apart.
    const float xq = fdiv(n_t, n_i);
    bool tir;
    {
        const float sq = fmaf(xq, xq, fmaf(cosi, cosi, -1.0f));
        const bool c_clear = cosi > 1.0e-4f && cosi <= 1.0f;            // incidence < pi/2 for sure, and defined
        if (!(cosi >= -1.0f && cosi <= 1.0f) || !(xq >= -1.0f && xq <= 1.0f)) tir = false;       // an angle is NaN
        else if (cosi < -1.0e-4f) tir = false;                                                    // incidence > pi/2
        else if (c_clear && xq < -1.0e-4f) tir = true;
        else if (c_clear && xq > 1.0e-4f && fabsf(sq) >= 1.0e-4f) tir = sq < 0.0f;
        else {
            const float critical = ref_asinf(xq);                                // 962
            const float incidence = ref_acosf(cosi);                             // 963
            tir = (critical < incidence) && ((double)incidence < (90.0 * 3.14159265358979323846 / 180.0));   // 964
        }
    }
    const float F0t = ref_pow2f(fdiv(fsub(n_t, n_i), fadd(n_t, n_i)));           // 965
    const float Ft = schlick(F0t, cosi);                                         // 966

    sp.want_t = depth_left > 0 && !tir && m.opacity < 1.0f && m.eta > 0.0f;      // 976
    if (sp.want_t) {
        // T = (-N) * sqrtf(1 - snell^2 (1 - cos^2)) + ((N*cos) - I) * snell     979-987
        const double inner = dsub(1.0, dmul((double)ref_pow2f(snell), dsub(1.0, (double)ref_pow2f(cosi))));
        const float root = fsqrt((float)inner);
        sp.T = vadd(vmulf(vneg(N), root), vmulf(vsub(vmulf(N, cosi), I), snell));
    }
    tw1 = (float)dsub(1.0, (double)Ft);                                          // 1083 / 1087
    tw2 = (float)dsub(1.0, (double)m.opacity);

    const float F0r = ref_pow2f(fdiv(fsub(m.eta, 1.0f), fadd(m.eta, 1.0f)));     // 1103
    fr = schlick(F0r, cosi);                                                     // 1104
    sp.want_r = depth_left > 0 && fr != 0.0f && m.ks > 0.0f;                     // 1105
    if (sp.want_r) sp.R = vsub(vmulf(N, fmul(2.0f, cosi)), I);                   // 1108
}

// Medium-stack state machine run when a spawned ray has found its hit
// (main.cpp:1021-1070 transmission, 1134-1182 reflection).
// `inc` = incidence object of the parent call, (n_i, n_t, exiting) = the parent's
// indices and ray state, stack/ns = the parent's stack, updated in place.
RT_DEV void medium_update(const SceneView& S, bool reflect, bool exiting, int inc, int hit,
                          float n_i, float n_t, int* stack, int& ns, float& new_ni, float& new_nt, bool& new_exiting)
{
    auto in_stack = [&](int o) { for (int k = 0; k < ns; k++) if (stack[k] == o) return true; return false; };
    if (!reflect) {
        if (!exiting) {                                                          // 1024
            if (hit == inc) {                                                    // 1026-1031
                new_exiting = true;
                new_ni = ns > 0 ? material_eta(S, stack[ns - 1]) : S.bkg_eta;    // back() on empty is UB upstream
                if (ns > 0) ns--;
                new_nt = ns > 0 ? material_eta(S, stack[ns - 1]) : S.bkg_eta;
                if (ns > 0) ns--;
            } else {                                                             // 1034-1038
                new_exiting = false; new_ni = n_t; new_nt = material_eta(S, hit);
                stack[ns++] = hit;
            }
        } else if (ns > 0) {                                                     // 1044
            if (!in_stack(hit)) {                                                // 1047-1051
                new_exiting = false; new_ni = n_t; new_nt = material_eta(S, hit);
                stack[ns++] = hit;
            } else {                                                             // 1054-1058
                new_exiting = true; new_ni = n_t; new_nt = material_eta(S, stack[ns - 1]);
                ns--;
            }
        } else {                                                                 // 1062-1066
            new_exiting = false; new_ni = S.bkg_eta; new_nt = material_eta(S, hit);
            stack[0] = hit; ns = 1;
        }
    } else {
        if (!exiting) {                                                          // 1137
            if (ns > 0) {
                if (!in_stack(hit)) {                                            // 1141-1145: pushes the INCIDENCE object
                    new_exiting = false; new_ni = n_i; new_nt = material_eta(S, hit);
                    stack[ns++] = inc;
                } else {                                                         // 1148-1152
                    new_exiting = false; new_ni = n_i; new_nt = material_eta(S, stack[ns - 1]);
                    ns--;
                }
            } else {                                                             // 1156-1160
                new_exiting = false; new_ni = n_i; new_nt = material_eta(S, hit);
                stack[0] = hit; ns = 1;
            }
        } else if (hit == inc) {                                                 // 1168-1171
            new_exiting = true; new_ni = n_i; new_nt = n_t;
        } else {                                                                 // 1174-1178
            new_exiting = false; new_ni = n_i; new_nt = material_eta(S, hit);
            stack[ns++] = hit;
        }
    }
}

}  // namespace rt
