// CPU check of csrc/smap_span.h (tests/test_smap_span.py): for random triangles, the spans of smap_span() against
// the per-cell barycentric test the map builder used before.  Prints "cells <in both> <span only> <test only>".
#include "smap_span.h"

#include <cstdio>
#include <cstdlib>
#include <random>

int main(int argc, char** argv)
{
    const unsigned seed = argc > 1 ? (unsigned)std::atoi(argv[1]) : 1u;
    const int n_tris = argc > 2 ? std::atoi(argv[2]) : 20000;
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> pos(-4.0f, 68.0f), size(0.05f, 30.0f), unit(-1.0f, 1.0f);
    const int R = 64;
    long long both = 0, span_only = 0, test_only = 0, far_off = 0;
    for (int t = 0; t < n_tris; t++) {
        float x[3], y[3];
        const float cx0 = pos(rng), cy0 = pos(rng), s = size(rng);
        for (int k = 0; k < 3; k++) { x[k] = cx0 + s * unit(rng); y[k] = cy0 + s * unit(rng); }
        const float den = (x[1] - x[0]) * (y[2] - y[0]) - (x[2] - x[0]) * (y[1] - y[0]);
        if (!(std::fabs(den) > 1.0e-12f)) continue;
        float lo_x = std::fmin(x[0], std::fmin(x[1], x[2])), hi_x = std::fmax(x[0], std::fmax(x[1], x[2]));
        float lo_y = std::fmin(y[0], std::fmin(y[1], y[2])), hi_y = std::fmax(y[0], std::fmax(y[1], y[2]));
        const int x0 = (int)std::fmax(std::ceil(lo_x - 0.5f), 0.0f), x1 = (int)std::fmin(std::floor(hi_x - 0.5f), R - 1.0f);
        const int y0 = (int)std::fmax(std::ceil(lo_y - 0.5f), 0.0f), y1 = (int)std::fmin(std::floor(hi_y - 0.5f), R - 1.0f);
        if (x1 < x0 || y1 < y0) continue;
        const float inv_den = 1.0f / den;
        for (int iy = y0; iy <= y1; iy++) {
            const float cy = (float)iy + 0.5f;
            int first = 0, last = -1;
            const rt::SmapRow row = rt::smap_row(x, y, inv_den, cy);
            const bool any = rt::smap_span(row, x0, x1 - x0 + 1, first, last);
            if (any && (first < x0 || last > x1 || first > last)) { std::printf("span out of range\n"); return 1; }
            for (int ix = x0; ix <= x1; ix++) {
                const float cx = (float)ix + 0.5f;
                const float b1 = ((cx - x[0]) * (y[2] - y[0]) - (x[2] - x[0]) * (cy - y[0])) / den;
                const float b2 = ((x[1] - x[0]) * (cy - y[0]) - (cx - x[0]) * (y[1] - y[0])) / den;
                const float b0 = 1.0f - b1 - b2;
                const bool in_test = b0 >= 0.0f && b1 >= 0.0f && b2 >= 0.0f;
                const bool in_span = any && ix >= first && ix <= last;
                if (in_test && in_span) both++;
                else if (in_span) {
                    span_only++;
                    // a disagreement is a cell whose centre lies on an edge, within rounding
                    if (std::fmin(b0, std::fmin(b1, b2)) < -1.0e-3f) far_off++;
                } else if (in_test) {
                    test_only++;
                    if (std::fmin(b0, std::fmin(b1, b2)) > 1.0e-3f) far_off++;
                }
            }
        }
    }
    std::printf("cells %lld %lld %lld far_off %lld\n", both, span_only, test_only, far_off);
    return 0;
}
