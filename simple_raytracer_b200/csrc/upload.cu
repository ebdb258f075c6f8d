// upload.cu — rt_upload_scene: raw RtSceneDesc arrays -> packed device tables.
//
// The raw description is copied to the device as it is (one cudaMemcpy per
// array) and packed THERE: one thread per primitive computes the ray-independent
// constants the reference recomputes for every ray (main.cpp:1280-1282, 1301,
// 1361-1366, 1382-1384) with the exact-rounding operators of device_math.cuh, and
// writes them with 128-bit stores.  Textures are widened RGB8 -> RGBA8 on the
// device so that a texel is one aligned 32-bit load.
#include "device_math.cuh"
#include "host_scene.h"

#include <cmath>
#include <cstring>

namespace rt {

namespace {

// What the packing kernels report back: bounds of all primitives (for the BVH padding) and the
// first primitive with a material / texture index out of range.  Floats are kept as order-
// preserving ints so that atomicMin / atomicMax work on them.
struct PackReport {
    int lo[3], hi[3];
    int bad_face, bad_sphere;
};

RT_DEV int ordered(float f) { const int i = __float_as_int(f); return i ^ ((i >> 31) & 0x7fffffff); }
inline float unordered(int i) { i ^= (i >> 31) & 0x7fffffff; float f; std::memcpy(&f, &i, 4); return f; }

// warp-reduced bounds -> one atomic pair per axis per warp; NaN coordinates are ignored (fminf / fmaxf)
RT_DEV void report_bounds(PackReport* rep, bool valid, const float lo[3], const float hi[3])
{
    for (int c = 0; c < 3; c++) {
        float l = valid ? lo[c] : 3.0e38f, h = valid ? hi[c] : -3.0e38f;
        for (int o = 16; o > 0; o >>= 1) {
            l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
            h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
        }
        if ((threadIdx.x & 31) == 0) { atomicMin(&rep->lo[c], ordered(l)); atomicMax(&rep->hi[c], ordered(h)); }
    }
}

__global__ void pack_faces_kernel(const RtFace* __restrict__ in, const RtMaterial* __restrict__ mats, int nf, int nmat,
                                  int ntex, float4* __restrict__ tri, float4* __restrict__ attr, int4* __restrict__ obj,
                                  PackReport* __restrict__ rep)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < nf;
    RtFace f;
    if (valid) f = in[i];
    {
        float lo[3], hi[3];
        for (int c = 0; c < 3; c++) {
            lo[c] = valid ? fminf(f.v[0][c], fminf(f.v[1][c], f.v[2][c])) : 0.f;
            hi[c] = valid ? fmaxf(f.v[0][c], fmaxf(f.v[1][c], f.v[2][c])) : 0.f;
        }
        report_bounds(rep, valid, lo, hi);
    }
    if (!valid) return;
    if (f.material < 0 || f.material >= nmat || f.texture < -1 || f.texture >= ntex) {
        atomicMin(&rep->bad_face, i);
        f.material = 0; f.texture = -1;      // keep the kernel in bounds; the upload is rejected anyway
        if (nmat == 0) return;
    }
    const V3 p0 = mk3(f.v[0][0], f.v[0][1], f.v[0][2]);
    const V3 p1 = mk3(f.v[1][0], f.v[1][1], f.v[1][2]);
    const V3 p2 = mk3(f.v[2][0], f.v[2][1], f.v[2][2]);
    const V3 e1 = vsub(p1, p0), e2 = vsub(p2, p0);                 // main.cpp:537-538 / 1280-1281
    const V3 n = vnorm(vcross(e1, e2));                            // 539
    const float D = -vdot(n, p0);                                  // 1301
    const float d11 = vdot(e1, e1), d12 = vdot(e1, e2), d22 = vdot(e2, e2);   // 1361-1363
    const float det = fsub(fmul(d11, d22), fmul(d12, d12));        // 1366
    const float kshadow = (float)dsub(1.0, (double)mats[f.material].opacity);   // 909 / 946
    float4* w = tri + (size_t)i * kTriWords;
    w[0] = make_float4(n.x, n.y, n.z, D);
    w[1] = make_float4(p0.x, p0.y, p0.z, det);
    w[2] = make_float4(e1.x, e1.y, e1.z, d11);
    w[3] = make_float4(e2.x, e2.y, e2.z, d22);
    w[4] = make_float4(d12, kshadow, __int_as_float(i), 0.0f);
    V3 vn[3];
    for (int k = 0; k < 3; k++) vn[k] = vnorm(mk3(f.vn[k][0], f.vn[k][1], f.vn[k][2]));   // 1382-1384
    float4* a = attr + (size_t)i * kFaceAttrWords;
    a[0] = make_float4(vn[0].x, vn[0].y, vn[0].z, f.vt[0][0]);
    a[1] = make_float4(vn[1].x, vn[1].y, vn[1].z, f.vt[0][1]);
    a[2] = make_float4(vn[2].x, vn[2].y, vn[2].z, f.vt[1][0]);
    a[3] = make_float4(f.vt[1][1], f.vt[2][0], f.vt[2][1], 0.0f);
    obj[i] = make_int4(f.material, f.texture, f.id, f.smooth ? kObjSmooth : 0);
}

__global__ void pack_spheres_kernel(const RtSphere* __restrict__ in, const RtMaterial* __restrict__ mats, int ns, int nf,
                                    int nmat, int ntex, float4* __restrict__ sph, int4* __restrict__ obj,
                                    PackReport* __restrict__ rep)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < ns;
    RtSphere s;
    if (valid) s = in[i];
    {
        float lo[3], hi[3];
        for (int c = 0; c < 3; c++) {
            const float r = valid ? fabsf(s.radius) : 0.f;
            lo[c] = valid ? s.center[c] - r : 0.f;
            hi[c] = valid ? s.center[c] + r : 0.f;
        }
        report_bounds(rep, valid, lo, hi);
    }
    if (!valid) return;
    if (s.material < 0 || s.material >= nmat || s.texture < -1 || s.texture >= ntex) {
        atomicMin(&rep->bad_sphere, i);
        s.material = 0; s.texture = -1;
        if (nmat == 0) return;
    }
    const float kshadow = (float)dsub(1.0, (double)mats[s.material].opacity);
    sph[(size_t)i * kSphWords] = make_float4(s.center[0], s.center[1], s.center[2], s.radius);
    sph[(size_t)i * kSphWords + 1] = make_float4(kshadow, __int_as_float(nf + i), 0.0f, 0.0f);
    obj[nf + i] = make_int4(s.material, s.texture, s.id, 0);
}

__global__ void pack_misc_kernel(const RtMaterial* __restrict__ mats, int nm, const RtLight* __restrict__ lights, int nl,
                                 float4* __restrict__ mat, float4* __restrict__ light)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nm) {
        const RtMaterial m = mats[i];
        mat[3 * i] = make_float4(m.od[0], m.od[1], m.od[2], m.ka);
        mat[3 * i + 1] = make_float4(m.os[0], m.os[1], m.os[2], m.kd);
        mat[3 * i + 2] = make_float4(m.ks, m.n, m.opacity, m.eta);
    }
    if (i < nl) {
        const RtLight l = lights[i];
        const V3 dir = mk3(l.xyz[0], l.xyz[1], l.xyz[2]);
        const V3 L = vneg(vnorm(dir));                             // main.cpp:887, hoisted
        light[3 * i] = make_float4(l.xyz[0], l.xyz[1], l.xyz[2], l.w);
        light[3 * i + 1] = make_float4(l.rgb[0], l.rgb[1], l.rgb[2], 0.0f);
        light[3 * i + 2] = make_float4(L.x, L.y, L.z, 0.0f);
    }
}

// RGB8 -> RGBA8: each thread converts 4 texels = 12 input bytes (three 32-bit
// loads) into one 128-bit store.
__global__ void widen_texels_kernel(const uint32_t* __restrict__ rgb, size_t n_texels, uchar4* __restrict__ out)
{
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // group of 4 texels
    const size_t first = q * 4;
    if (first >= n_texels) return;
    if (first + 4 <= n_texels) {
        const uint32_t a = rgb[q * 3], b = rgb[q * 3 + 1], c = rgb[q * 3 + 2];
        uint4 o;
        o.x = a & 0x00ffffffu;
        o.y = (a >> 24) | ((b & 0x0000ffffu) << 8);
        o.z = (b >> 16) | ((c & 0x000000ffu) << 16);
        o.w = c >> 8;
        reinterpret_cast<uint4*>(out)[q] = o;
    } else {
        const unsigned char* bytes = reinterpret_cast<const unsigned char*>(rgb);
        for (size_t t = first; t < n_texels; t++)
            out[t] = make_uchar4(bytes[3 * t], bytes[3 * t + 1], bytes[3 * t + 2], 0);
    }
}

template <class T>
int upload_array(DevBuf& buf, const T* host, size_t n, cudaStream_t st)
{
    if (!buf.reserve(n * sizeof(T) + 16)) return RT_E_NOMEM;
    if (n) RT_CUDA(cudaMemcpyAsync(buf.p, host, n * sizeof(T), cudaMemcpyHostToDevice, st));
    return RT_OK;
}

}  // namespace

// main.cpp:677-710 on the host: the only double `tan` of the frame (line 690) is
// glibc's, exactly as in the reference.  Host float arithmetic is IEEE single with
// no contraction (-ffp-contract=off, no -march).
void make_camera(const float eye[3], const float viewdir[3], const float updir[3], float hfov_deg, int width,
                 int height, CameraView* cam, bool normalised)
{
    struct H3 { float x, y, z; };
    auto add = [](H3 a, H3 b) { return H3{ a.x + b.x, a.y + b.y, a.z + b.z }; };
    auto sub = [](H3 a, H3 b) { return H3{ a.x - b.x, a.y - b.y, a.z - b.z }; };
    auto mulf = [](H3 a, float s) { return H3{ a.x * s, a.y * s, a.z * s }; };
    auto divf = [](H3 a, float s) { return H3{ a.x / s, a.y / s, a.z / s }; };
    auto norm = [&](H3 a) {
        volatile float xx = a.x * a.x, yy = a.y * a.y, zz = a.z * a.z;
        volatile float s = xx + yy;
        s = s + zz;
        return divf(a, std::sqrt((float)s));
    };
    auto cross = [](H3 a, H3 b) {
        volatile float x1 = a.y * b.z, x2 = a.z * b.y, y1 = a.z * b.x, y2 = a.x * b.z, z1 = a.x * b.y, z2 = a.y * b.x;
        return H3{ x1 - x2, y1 - y2, z1 - z2 };
    };
    const H3 origin{ eye[0], eye[1], eye[2] };
    const H3 vd_raw{ viewdir[0], viewdir[1], viewdir[2] }, vu_raw{ updir[0], updir[1], updir[2] };
    const H3 vd = normalised ? vd_raw : norm(vd_raw);                // main.cpp:607
    const H3 vu = normalised ? vu_raw : norm(vu_raw);
    const float res_h = (float)height, res_w = (float)width;
    const H3 u = norm(cross(vd, vu));                                // 677
    const H3 v = cross(u, vd);                                       // 684
    const float aspect = res_w / res_h;                              // 689
    const float w = (float)((double)2.0f * 5.0 * std::tan((0.5 * (double)hfov_deg) * 3.14159265358979323846 / (double)180.0f));   // 690
    const float h = w / aspect;                                      // 691
    const H3 nd = mulf(vd, 5.0f);
    const H3 ul = add(sub(add(origin, nd), mulf(u, w / 2.0f)), mulf(v, h / 2.0f));   // 697
    const H3 ur = add(add(add(origin, nd), mulf(u, w / 2.0f)), mulf(v, h / 2.0f));   // 698
    const H3 ll = sub(sub(add(origin, nd), mulf(u, w / 2.0f)), mulf(v, h / 2.0f));   // 699
    const H3 dh = divf(sub(ur, ul), res_w - 1.0f);                   // 709
    const H3 dv = divf(sub(ll, ul), res_h - 1.0f);                   // 710
    cam->eye[0] = origin.x; cam->eye[1] = origin.y; cam->eye[2] = origin.z;
    cam->ul[0] = ul.x; cam->ul[1] = ul.y; cam->ul[2] = ul.z;
    cam->delta_h[0] = dh.x; cam->delta_h[1] = dh.y; cam->delta_h[2] = dh.z;
    cam->delta_v[0] = dv.x; cam->delta_v[1] = dv.y; cam->delta_v[2] = dv.z;
    cam->width = width; cam->height = height;
}

int build_scene(const RtSceneDesc* d, SceneHost* s)
{
    if (d->width < 2 || d->height < 2) { set_error("imsize must be > 1 in both dimensions (main.cpp:242)"); return RT_E_INVALID; }
    if (d->n_faces < 0 || d->n_spheres < 0 || d->n_lights < 0 || d->n_materials < 0 || d->n_textures < 0) {
        set_error("negative count in RtSceneDesc");
        return RT_E_INVALID;
    }
    if ((d->n_faces && !d->faces) || (d->n_spheres && !d->spheres) || (d->n_lights && !d->lights) ||
        (d->n_materials && !d->materials) || (d->n_textures && !d->textures)) {
        set_error("null array in RtSceneDesc");
        return RT_E_INVALID;
    }
    s->width = d->width; s->height = d->height;
    s->nf = d->n_faces; s->ns = d->n_spheres; s->nl = d->n_lights; s->ntex = d->n_textures; s->nmat = d->n_materials;
    for (int k = 0; k < 3; k++) s->bkg[k] = d->bkg[k];
    s->bkg_eta = d->bkg_eta;
    s->nonunit_directional = false;
    for (int i = 0; i < d->n_lights; i++) {
        const RtLight& l = d->lights[i];
        if (l.w == 0.0f) {
            const double s2 = (double)l.xyz[0] * l.xyz[0] + (double)l.xyz[1] * l.xyz[1] + (double)l.xyz[2] * l.xyz[2];
            if (std::fabs(s2 - 1.0) > 1.0e-5) s->nonunit_directional = true;
        }
    }
    bool tame = true;
    for (int i = 0; i < d->n_materials; i++) {
        const RtMaterial& m = d->materials[i];
        const float v[12] = { m.od[0], m.od[1], m.od[2], m.os[0], m.os[1], m.os[2], m.ka, m.kd, m.ks, m.n, m.opacity, m.eta };
        for (float x : v) tame = tame && std::isfinite(x);
        tame = tame && std::fabs(m.n) <= 1.0e6f;
    }
    for (int i = 0; i < d->n_lights; i++)
        for (int c = 0; c < 3; c++) tame = tame && std::isfinite(d->lights[i].rgb[c]) && std::isfinite(d->lights[i].xyz[c]);
    s->zero_mask_exit = tame;
    make_camera(d->eye, d->viewdir, d->updir, d->hfov_deg, d->width, d->height, &s->cam, d->camera_normalised != 0);

    cudaStream_t st = s->stream;
    DevBuf raw_faces, raw_spheres, raw_mats, raw_lights, raw_tex;
    int rc;
    if ((rc = upload_array(raw_faces, d->faces, (size_t)d->n_faces, st)) != RT_OK) return rc;
    if ((rc = upload_array(raw_spheres, d->spheres, (size_t)d->n_spheres, st)) != RT_OK) return rc;
    if ((rc = upload_array(raw_mats, d->materials, (size_t)d->n_materials, st)) != RT_OK) return rc;
    if ((rc = upload_array(raw_lights, d->lights, (size_t)d->n_lights, st)) != RT_OK) return rc;

    const size_t nobj = (size_t)s->nf + (size_t)s->ns;
    if (!s->tri.reserve(sizeof(float4) * kTriWords * (size_t)s->nf + 16) ||
        !s->sph.reserve(sizeof(float4) * kSphWords * (size_t)s->ns + 16) ||
        !s->face_attr.reserve(sizeof(float4) * kFaceAttrWords * (size_t)s->nf + 16) ||
        !s->obj.reserve(sizeof(int4) * nobj + 16) || !s->mat.reserve(sizeof(float4) * 3 * (size_t)s->nmat + 16) ||
        !s->light.reserve(sizeof(float4) * 3 * (size_t)s->nl + 16))
        return RT_E_NOMEM;

    DevBuf report;
    if (!report.reserve(sizeof(PackReport))) return RT_E_NOMEM;
    {
        PackReport init;
        for (int c = 0; c < 3; c++) { init.lo[c] = 0x7f7fffff /* ordered(FLT_MAX) */; init.hi[c] = (int)0x80800000 /* ordered(-FLT_MAX) */; }
        init.bad_face = 0x7fffffff; init.bad_sphere = 0x7fffffff;
        RT_CUDA(cudaMemcpyAsync(report.p, &init, sizeof init, cudaMemcpyHostToDevice, st));
    }
    const int B = 256;
    if (s->nf)
        pack_faces_kernel<<<(s->nf + B - 1) / B, B, 0, st>>>(raw_faces.as<RtFace>(), raw_mats.as<RtMaterial>(), s->nf, s->nmat,
                                                            s->ntex, s->tri.as<float4>(), s->face_attr.as<float4>(),
                                                            s->obj.as<int4>(), report.as<PackReport>());
    if (s->ns)
        pack_spheres_kernel<<<(s->ns + B - 1) / B, B, 0, st>>>(raw_spheres.as<RtSphere>(), raw_mats.as<RtMaterial>(), s->ns,
                                                              s->nf, s->nmat, s->ntex, s->sph.as<float4>(), s->obj.as<int4>(),
                                                              report.as<PackReport>());
    const int nmisc = s->nmat > s->nl ? s->nmat : s->nl;
    if (nmisc)
        pack_misc_kernel<<<(nmisc + B - 1) / B, B, 0, st>>>(raw_mats.as<RtMaterial>(), s->nmat, raw_lights.as<RtLight>(),
                                                           s->nl, s->mat.as<float4>(), s->light.as<float4>());
    RT_CUDA(cudaGetLastError());

    // textures: one staging upload per texture, widened on the device
    std::vector<int4> desc((size_t)s->ntex);
    size_t total = 0;
    for (int t = 0; t < s->ntex; t++) {
        const RtTexture& T = d->textures[t];
        if (T.width < 1 || T.height < 1 || !T.rgb) { set_error("texture " + std::to_string(t) + ": bad size / null pixels"); return RT_E_INVALID; }
        total = (total + 3) & ~(size_t)3;   // keep every texture 16-byte aligned in the RGBA8 pool
        if (total + (size_t)T.width * T.height > 0x7fffffffu) { set_error("texture pool exceeds 2^31 texels"); return RT_E_INVALID; }
        desc[t] = make_int4((int)total, T.width, T.height, 0);
        total += (size_t)T.width * (size_t)T.height;
    }
    if (!s->texels.reserve(total * sizeof(uchar4) + 16) || !s->tex_desc.reserve(sizeof(int4) * (size_t)s->ntex + 16)) return RT_E_NOMEM;
    if (s->ntex) RT_CUDA(cudaMemcpyAsync(s->tex_desc.p, desc.data(), sizeof(int4) * desc.size(), cudaMemcpyHostToDevice, st));
    for (int t = 0; t < s->ntex; t++) {
        const RtTexture& T = d->textures[t];
        const size_t n = (size_t)T.width * (size_t)T.height;
        if (!raw_tex.reserve(n * 3 + 16)) return RT_E_NOMEM;
        RT_CUDA(cudaMemcpyAsync(raw_tex.p, T.rgb, n * 3, cudaMemcpyHostToDevice, st));
        const size_t groups = (n + 3) / 4;
        widen_texels_kernel<<<(unsigned)((groups + B - 1) / B), B, 0, st>>>(raw_tex.as<uint32_t>(), n,
                                                                           s->texels.as<uchar4>() + desc[t].x);
        RT_CUDA(cudaGetLastError());
        // raw_tex is reused by the next texture: stream order keeps this safe
    }

    PackReport h_rep;
    RT_CUDA(cudaMemcpyAsync(&h_rep, report.p, sizeof h_rep, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));   // raw staging buffers die here
    if (h_rep.bad_face < d->n_faces) {
        set_error("face " + std::to_string(h_rep.bad_face) + ": material / texture index out of range");
        return RT_E_INVALID;
    }
    if (h_rep.bad_sphere < d->n_spheres) {
        set_error("sphere " + std::to_string(h_rep.bad_sphere) + ": material / texture index out of range");
        return RT_E_INVALID;
    }
    // bounds of all primitives (for the BVH padding)
    for (int c = 0; c < 3; c++) { s->scene_lo[c] = unordered(h_rep.lo[c]); s->scene_hi[c] = unordered(h_rep.hi[c]); }
    return RT_OK;
}

SceneView scene_view(const SceneHost* s)
{
    SceneView v;
    std::memset(&v, 0, sizeof v);      // padding too: enqueue_frame compares FrameParams bytewise (graph replay)
    v.nf = s->nf; v.ns = s->ns; v.nl = s->nl; v.ntex = s->ntex;
    v.tri = s->tri.as<float4>(); v.sph = s->sph.as<float4>(); v.face_attr = s->face_attr.as<float4>();
    v.obj = s->obj.as<int4>(); v.mat = s->mat.as<float4>(); v.light = s->light.as<float4>();
    v.tex_desc = s->tex_desc.as<int4>(); v.texels = s->texels.as<uchar4>();
    v.bvh.nodes = s->bvh_nodes.as<float4>(); v.bvh.n_nodes = s->bvh_n_nodes; v.bvh.root = s->bvh_root;
    v.bvh_tri = s->bvh_tri.as<float4>(); v.bvh_sph = s->bvh_sph.as<float4>();
    v.bvh.nodes8 = s->bvh8_nodes.as<uint4>(); v.bvh.prims8 = s->bvh8_prims.as<float4>(); v.bvh.n_nodes8 = s->bvh8_n_nodes;
    for (int k = 0; k < 3; k++) v.bkg[k] = s->bkg[k];
    v.bkg_eta = s->bkg_eta;
    v.zero_mask_exit = s->zero_mask_exit ? 1 : 0;
    for (int k = 0; k < kMaxShadowMaps; k++) {
        const SceneHost::SmapHost& m = s->smap[k];
        if (!m.offset_plus_1 || !s->smap_cells.p) continue;
        ShadowMapView& d = v.smap[k];
        d.cells = s->smap_cells.as<int>() + (m.offset_plus_1 - 1);
        for (int c = 0; c < 3; c++) { d.ub[c] = m.ub[c]; d.vb[c] = m.vb[c]; d.origin[c] = m.origin[c]; }
        d.u0 = m.u0; d.v0 = m.v0; d.inv_cell = m.inv_cell; d.R = m.R; d.cube = m.cube;
    }
    return v;
}

}  // namespace rt
