// smap_span.h — the cells of one row of an occluder map (bvh.cu: build_shadow_maps) whose centres a projected
// triangle covers, as a span [first, last] instead of a test per cell of the triangle's bounding box.
//
// Host and device: tests/test_smap_span.py compiles this header with g++ and compares the spans with the per-cell
// barycentric test on random triangles.  Plain float arithmetic; a cell on an edge may fall on either side — the map
// is a guess that the exact test confirms or not (scene_device.h: ShadowMapView).
#pragma once

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#include <cmath>
#define RT_HD inline
#endif

namespace rt {

// Barycentrics b1, b2 of the point (cx, cy) in the triangle (x[0..2], y[0..2]) are linear in cx for a fixed row:
// b_k = a_k * cx + c_k(cy), b0 = 1 - b1 - b2.  inv_den = 1 / ((x1-x0)(y2-y0) - (x2-x0)(y1-y0)).
struct SmapRow {
    float a1, c1, a2, c2;      // b1 = a1 * cx + c1, b2 = a2 * cx + c2 on this row
};

RT_HD SmapRow smap_row(const float x[3], const float y[3], float inv_den, float cy)
{
    SmapRow r;
    r.a1 = (y[2] - y[0]) * inv_den;
    r.c1 = (-x[0] * (y[2] - y[0]) - (x[2] - x[0]) * (cy - y[0])) * inv_den;
    r.a2 = -(y[1] - y[0]) * inv_den;
    r.c2 = ((x[1] - x[0]) * (cy - y[0]) + x[0] * (y[1] - y[0])) * inv_den;
    return r;
}

// a * cx + c >= 0 narrows [lo, hi]; false: nothing left.
RT_HD bool smap_clip(float a, float c, float& lo, float& hi)
{
    if (a > 0.0f) { const float t = -c / a; if (t > lo) lo = t; }
    else if (a < 0.0f) { const float t = -c / a; if (t < hi) hi = t; }
    else if (!(c >= 0.0f)) return false;
    return lo <= hi;
}

// Cells x0 .. x0 + wx - 1 of the row (centres at i + 0.5): the span [first, last] inside the triangle; false: none.
RT_HD bool smap_span(const SmapRow& r, int x0, int wx, int& first, int& last)
{
    float lo = (float)x0 + 0.5f, hi = (float)(x0 + wx - 1) + 0.5f;
    if (!smap_clip(r.a1, r.c1, lo, hi)) return false;
    if (!smap_clip(r.a2, r.c2, lo, hi)) return false;
    if (!smap_clip(-(r.a1 + r.a2), 1.0f - r.c1 - r.c2, lo, hi)) return false;
    // (NaN bounds fail the comparisons above or the one below)
    const float f = ceilf(lo - 0.5f), l = floorf(hi - 0.5f);
    if (!(f <= l)) return false;
    first = (int)f < x0 ? x0 : (int)f;
    last = (int)l > x0 + wx - 1 ? x0 + wx - 1 : (int)l;
    return first <= last;
}

}  // namespace rt
