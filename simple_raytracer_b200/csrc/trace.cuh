// trace.cuh — ray/primitive tests and the brute-force scene walk.
//
// Restates TraceRay (main.cpp:1215-1407) and the three loops its callers run over
// its result: closest hit (730-742, 1109-1124), transmission closest hit with the
// face rule (992-1011) and the shadow product (898-912, 930-949).  The reference
// materialises every intersection in a vector and then scans it; here each scan is
// folded into the walk, visiting primitives in the same order (faces in file
// order, then spheres; per sphere the far root before the near root) so that the
// strict-'<' tie rule and the order-dependent face rule come out identically.
#pragma once

#include "device_math.cuh"
#include "scene_device.h"

namespace rt {

struct Hit {
    float t;
    int obj;      // order index, -1 = miss
    float b, g;   // barycentrics of a face hit (a = 1 - (b + g)), main.cpp:1368-1370
};

RT_DEV V3 ld3(const float4& q) { return mk3(q.x, q.y, q.z); }

// Ray-plane distance of a triangle, main.cpp:1296-1302.  false when dem == 0.
RT_DEV bool tri_distance(const float4& w0, V3 o, V3 ray, float& t)
{
    const V3 n = ld3(w0);
    const float dem = vdot(n, ray);                              // 1296
    if (dem == 0.0f) return false;                               // 1297
    t = fdiv(-fadd(vdot(n, o), w0.w), dem);                      // 1302 (D hoisted, 1301)
    return true;
}

// Barycentric inside test at distance t, main.cpp:1303, 1358-1378 (strict).
RT_DEV bool tri_inside(const float4& w1, const float4& w2, const float4& w3, const float4& w4,
                       V3 o, V3 ray, float t, float& b, float& g)
{
    const float det = w1.w;
    if (!(det != 0.0f)) return false;                            // 1367
    const V3 p = vadd(o, vmulf(ray, t));                         // 1303
    const V3 ep = vsub(p, ld3(w1));                              // 1360
    const float d1p = vdot(ld3(w2), ep);                         // 1364
    const float d2p = vdot(ld3(w3), ep);                         // 1365
    const float d11 = w2.w, d22 = w3.w, d12 = w4.x;
    b = fdiv(fsub(fmul(d22, d1p), fmul(d12, d2p)), det);         // 1368
    g = fdiv(fsub(fmul(d11, d2p), fmul(d12, d1p)), det);         // 1369
    const float a = fsub(1.0f, fadd(b, g));                      // 1370
    return ((0.0f < a) && (a < 1.0f)) && ((0.0f < b) && (b < 1.0f)) && ((0.0f < g) && (g < 1.0f));   // 1378
}

// Sphere quadratic with A == 1 regardless of |ray| (main.cpp:1225-1268).
// Returns the number of roots the reference pushes (0, 1 or 2); t_far is pushed first.
RT_DEV int sphere_roots(const float4& w0, V3 o, V3 ray, float& t_far, float& t_near)
{
    const V3 dir = vsub(o, ld3(w0));                                             // 1225
    const float dot = vsum3(fmul(ray.x, dir.x), fmul(ray.y, dir.y), fmul(ray.z, dir.z));
    const float B = fmul(2.0f, dot);                                             // 1233: exact doubling
    const float ss = vsum3(fmul(dir.x, dir.x), fmul(dir.y, dir.y), fmul(dir.z, dir.z));
    {
        // Screening (culls only): most tests are clear misses, and the exact discriminant below costs six
        // float<->double conversions and three FP64 operations.  e approximates B^2 - 4C within a few ulp of
        // m = B^2 + 4(ss + r^2); when e < -1e-6 m (16 times that error) the exact det is negative and not
        // -0: no root.  NaN / inf operands fail the comparison and take the exact path.
        const float rr = w0.w * w0.w;
        const float e = fmaf(B, B, -4.0f * (ss - rr)), m = fmaf(B, B, 4.0f * (ss + rr));
        if (e < -1.0e-6f * m) return 0;
    }
    const double r = (double)w0.w;
    const float C = (float)dsub((double)ss, dmul(r, r));                         // 1234: pow(r, 2.0) is exact
    const float det = (float)dsub(dmul((double)B, (double)B), dmul(4.0, (double)C));   // 1242
    if (!signbit(det)) {                                                         // 1243
        const float sq = fsqrt(det);
        t_far = fmul(fadd(-B, sq), 0.5f);                                        // 1244: / 2.0 is exact
        t_near = fmul(fsub(-B, sq), 0.5f);                                       // 1252
        return 2;
    }
    if (det == 0.0f) {                                                           // 1260: only -0.0f gets here
        t_far = t_near = fmul(-B, 0.5f);                                         // 1261
        return 1;
    }
    return 0;
}

// Hit point and shading normal of an accepted hit (main.cpp:1245-1257, 1380-1393).
struct Surface {
    V3 p, n;
    float a, b, g;
};

RT_DEV Surface finish_hit(const SceneView& S, V3 o, V3 ray, const Hit& h)
{
    Surface s;
    s.p = vadd(o, vmulf(ray, h.t));
    s.b = h.b; s.g = h.g; s.a = fsub(1.0f, fadd(h.b, h.g));
    if (h.obj < S.nf) {
        const int4 oi = S.obj[h.obj];
        if (oi.w & kObjSmooth) {
            const float4* fa = S.face_attr + (size_t)h.obj * kFaceAttrWords;
            const float4 a0 = fa[0], a1 = fa[1], a2 = fa[2];
            // (vnn0 * a + vnn1 * b) + vnn2 * g, then norm()            1381-1385
            s.n = vnorm(vadd(vadd(vmulf(ld3(a0), s.a), vmulf(ld3(a1), s.b)), vmulf(ld3(a2), s.g)));
        } else {
            s.n = ld3(S.tri[(size_t)h.obj * kTriWords]);                       // 1387
        }
    } else {
        const float4 w0 = S.sph[(size_t)(h.obj - S.nf) * kSphWords];
        s.n = vnorm(vdivf(vsub(s.p, ld3(w0)), w0.w));                          // 1249 / 1257
    }
    return s;
}

// ---- brute force over primitive tables (global memory or a shared-memory copy) --
struct BruteAccel {
    const float4* tri;
    const float4* sph;
    int nf, ns;

    // closest hit with t > tmin, strict '<', first in iteration order wins ties
    RT_DEV Hit closest(V3 o, V3 ray, float tmin) const
    {
        Hit best; best.t = 3.402823466e+38f; best.obj = -1; best.b = 0.0f; best.g = 0.0f;   // FLT_MAX, 722
        for (int i = 0; i < nf; i++) {
            const float4* w = tri + (size_t)i * kTriWords;
            float t;
            if (!tri_distance(w[0], o, ray, t)) continue;
            if (!(t > tmin && t < best.t)) continue;                           // 736 / 1117
            float b, g;
            if (tri_inside(w[1], w[2], w[3], w[4], o, ray, t, b, g)) { best.t = t; best.obj = i; best.b = b; best.g = g; }
        }
        for (int i = 0; i < ns; i++) {
            float tf, tn;
            const int n = sphere_roots(sph[(size_t)i * kSphWords], o, ray, tf, tn);
            if (n >= 1 && tf > tmin && tf < best.t) { best.t = tf; best.obj = nf + i; }
            if (n == 2 && tn > tmin && tn < best.t) { best.t = tn; best.obj = nf + i; }
        }
        return best;
    }

    // Transmission closest hit for FACE incidence (main.cpp:992-1011): the walk is
    // abandoned (`goto SKIP_TRANS`) the first time a candidate that improves the
    // running minimum belongs to an object other than `back` (= stack.back()).
    RT_DEV Hit closest_face_rule(V3 o, V3 ray, float tmin, int back, bool& aborted) const
    {
        Hit best; best.t = 3.402823466e+38f; best.obj = -1; best.b = 0.0f; best.g = 0.0f;
        aborted = false;
        for (int i = 0; i < nf; i++) {
            const float4* w = tri + (size_t)i * kTriWords;
            float t;
            if (!tri_distance(w[0], o, ray, t)) continue;
            if (!(t > tmin && t < best.t)) continue;
            float b, g;
            if (tri_inside(w[1], w[2], w[3], w[4], o, ray, t, b, g)) {
                if (i != back) { aborted = true; return best; }                // 1000-1002
                best.t = t; best.obj = i; best.b = b; best.g = g;
            }
        }
        for (int i = 0; i < ns; i++) {
            float tf, tn;
            const int n = sphere_roots(sph[(size_t)i * kSphWords], o, ray, tf, tn);
            if (n >= 1 && tf > tmin && tf < best.t) {
                if (nf + i != back) { aborted = true; return best; }
                best.t = tf; best.obj = nf + i;
            }
            if (n == 2 && tn > tmin && tn < best.t) {
                if (nf + i != back) { aborted = true; return best; }
                best.t = tn; best.obj = nf + i;
            }
        }
        return best;
    }

    // Shadow product (main.cpp:898-912, 930-949): every intersection of every
    // object other than `self` with tmin < t (< tmax for point lights) multiplies the mask by
    // (float)(1.0 - opacity), clamped, in iteration order.  Returns the hit count.
    RT_DEV int shadow(V3 o, V3 ray, int self, float tmin, float tmax, bool use_tmax, float& mask) const
    {
        int hits = 0;
        for (int i = 0; i < nf; i++) {
            if (i == self) continue;                                           // 900 / 935
            const float4* w = tri + (size_t)i * kTriWords;
            float t;
            if (!tri_distance(w[0], o, ray, t)) continue;
            if (!(t > tmin && (!use_tmax || t < tmax))) continue;              // 907 / 944
            float b, g;
            if (tri_inside(w[1], w[2], w[3], w[4], o, ray, t, b, g)) { mask = clamp01(fmul(w[4].y, mask)); hits++; }
        }
        for (int i = 0; i < ns; i++) {
            if (nf + i == self) continue;
            const float4* w = sph + (size_t)i * kSphWords;
            float tf, tn;
            const int n = sphere_roots(w[0], o, ray, tf, tn);
            const float k = w[1].x;
            if (n >= 1 && tf > tmin && (!use_tmax || tf < tmax)) { mask = clamp01(fmul(k, mask)); hits++; }
            if (n == 2 && tn > tmin && (!use_tmax || tn < tmax)) { mask = clamp01(fmul(k, mask)); hits++; }
        }
        return hits;
    }
};

}  // namespace rt
