// scene_parser.cpp — the scene-file grammar of the reference, re-stated.
//
// Follows main.cpp:88-602 of the reference.  The grammar
// and its quirks are the contract; the implementation is a flat token switch
// writing straight into an RtSceneDesc instead of the reference's pointer-linked
// maps.  Where the reference aborts (uncaught exception -> std::terminate) or
// prints and returns without an image, we return RT_E_PARSE with a message.
#include "host_common.h"

#include <cerrno>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string_view>
#include <unordered_map>

namespace rthost {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

void OwnedScene::seal()
{
    for (size_t i = 0; i < textures.size(); i++) textures[i].rgb = texture_pixels[i].data();
    desc.n_materials = (int32_t)materials.size(); desc.materials = materials.data();
    desc.n_textures = (int32_t)textures.size();   desc.textures = textures.data();
    desc.n_spheres = (int32_t)spheres.size();     desc.spheres = spheres.data();
    desc.n_faces = (int32_t)faces.size();         desc.faces = faces.data();
    desc.n_lights = (int32_t)lights.size();       desc.lights = lights.data();
}

bool parse_float(const std::string& tok, float* out)
{
    // std::stof == strtof + "no conversion" / ERANGE checks
    const char* s = tok.c_str();
    char* end = nullptr;
    errno = 0;
    float v = std::strtof(s, &end);
    if (end == s || errno == ERANGE) return false;
    *out = v;
    return true;
}

bool parse_int(const std::string& tok, int* out)
{
    const char* s = tok.c_str();
    char* end = nullptr;
    errno = 0;
    long v = std::strtol(s, &end, 10);
    if (end == s || errno == ERANGE || v < INT32_MIN || v > INT32_MAX) return false;
    *out = (int)v;
    return true;
}

namespace {

struct Vec3 { float x = 0, y = 0, z = 0; };
struct Vec2 { float x = 0, y = 0; };

struct ParseError { std::string msg; };

[[noreturn]] void fail(const std::string& msg) { throw ParseError{ msg }; }

// The reference keeps v / vn / vt in std::maps keyed by a running 1-based count; a
// missing key reads as zeros (std::map::operator[], main.cpp:489-511).  The keys are
// dense, so a vector with slot 0 unused is the same mapping.
template <class T>
T lookup(const std::vector<T>& m, unsigned key)
{
    return (key >= 1 && (size_t)key < m.size()) ? m[key] : T{};
}

using Tok = std::string_view;

// The arguments of one command: a window into the line's token list.
struct Args {
    const Tok* p;
    size_t n;
    size_t size() const { return n; }
    const Tok& operator[](size_t i) const { return p[i]; }
};

// std::stof on a token, fast: plain decimal tokens go through std::from_chars (correctly
// rounded, like glibc's strtof); anything else — leading blanks or '+', hex floats, inf /
// nan, trailing junk such as a '\r', results that strtof flags with ERANGE (overflow,
// subnormals) — takes the strtof path so every corner keeps std::stof's behaviour.
bool parse_float_tok(Tok t, float* out)
{
    // from_chars refuses what only strtof takes (leading blanks, '+', hex) and stops before
    // trailing junk: all of that shows as "did not consume the whole token"
    float v;
    const char* const e = t.data() + t.size();
    const auto r = std::from_chars(t.data(), e, v);
    if (r.ec == std::errc() && r.ptr == e) {
        if (std::isnormal(v)) { *out = v; return true; }
        if (v == 0.0f) {
            // exact zeros ("0", "-0.000") are fine; a non-zero decimal that ROUNDS to zero is
            // ERANGE for strtof -> slow path decides
            bool zero_digits = true;
            for (char c : t) { if (c == 'e' || c == 'E') break; if (c >= '1' && c <= '9') { zero_digits = false; break; } }
            if (zero_digits) { *out = v; return true; }
        }
    }
    return parse_float(std::string(t), out);
}

float need_float(const Args& a, size_t i, const char* what)
{
    float v;
    if (i >= a.size() || !parse_float_tok(a[i], &v)) fail(std::string("ERROR: Invalid args for ") + what + ". Please verify.");
    return v;
}

int need_int(const Args& a, size_t i, const char* what)
{
    int v;
    if (i >= a.size() || !parse_int(std::string(a[i]), &v)) fail(std::string("ERROR: Invalid ") + what + ". Please verify.");
    return v;
}

// sscanf("%d") on [p, e): optional sign + digits (tokens cannot start with a blank here).
// false when there is no digit.  Values beyond int take the slow sscanf path (ok = false).
bool scan_int(const char*& p, const char* e, int* out, bool* exotic)
{
    const char* q = p;
    bool neg = false;
    if (q < e && (*q == '+' || *q == '-')) { neg = (*q == '-'); q++; }
    if (q >= e || *q < '0' || *q > '9') return false;
    long long v = 0;
    while (q < e && *q >= '0' && *q <= '9') {
        v = v * 10 + (*q - '0');
        if (v > 0x7fffffffLL) { *exotic = true; return false; }
        q++;
    }
    *out = (int)(neg ? -v : v);
    p = q;
    return true;
}

struct Parser {
    OwnedScene* out;
    std::string texture_dir;

    std::vector<Vec3> vertices{ Vec3{} }, normals{ Vec3{} };     // slot 0 unused (keys are 1-based)
    std::vector<Vec2> texcoords{ Vec2{} };
    unsigned obj_id = 0;
    // The reference reads the file again at every `texture` command (main.cpp:318); the same
    // path gives the same texels, so later commands reuse the first copy.
    std::unordered_map<std::string, int> texture_by_path;
    std::unordered_map<std::string, int> material_by_value;
    int current_material = -1, current_texture = -1;
    bool use_texture = false, has_material = false;
    bool seen_imsize = false, seen_eye = false, seen_viewdir = false, seen_updir = false, seen_hfov = false,
         seen_bkg = false;

    void material_and_texture(int32_t* mat, int32_t* tex)
    {
        // main.cpp:351-365 / 521-534: objects capture the CURRENT material and texture
        if (use_texture) {
            if (!has_material || current_texture < 0) fail("ERROR: Must define a 'mtlcolor' and 'texture'. Please verify.");
            *tex = current_texture;
        } else {
            if (!has_material) fail("ERROR: Must define a 'mtlcolor'. Please verify.");
            *tex = -1;
        }
        *mat = current_material;
    }

    void command(Tok cmd, const Args& a)
    {
        RtSceneDesc& d = out->desc;
        if (cmd == "eye") {                                   // main.cpp:143-162
            for (int k = 0; k < 3; k++) d.eye[k] = need_float(a, k, "'eye' command");
            seen_eye = true;
        } else if (cmd == "viewdir") {                        // 163-181
            for (int k = 0; k < 3; k++) d.viewdir[k] = need_float(a, k, "'viewdir' command");
            seen_viewdir = true;
        } else if (cmd == "updir") {                          // 182-200
            for (int k = 0; k < 3; k++) d.updir[k] = need_float(a, k, "'updir' command");
            seen_updir = true;
        } else if (cmd == "hfov") {                           // 201-215
            d.hfov_deg = need_float(a, 0, "'hfov' command");
            seen_hfov = true;
        } else if (cmd == "imsize") {                         // 216-246: height first, then width
            int h = need_int(a, 1, "image dimensions");
            int w = need_int(a, 0, "image dimensions");
            if (h <= 1 || w <= 1) fail("ERROR: Invalid image dimensions. Please verify.");
            d.width = w; d.height = h;
            seen_imsize = true;
        } else if (cmd == "bkgcolor") {                       // 247-271
            for (int k = 0; k < 3; k++) d.bkg[k] = need_float(a, k, "'bkgcolor' command");
            if (a.size() > 3) d.bkg_eta = need_float(a, 3, "'bkgcolor' command");
            seen_bkg = true;
        } else if (cmd == "mtlcolor") {                       // 272-308
            use_texture = false;
            RtMaterial m{};
            const char* what = "'material'";
            for (int k = 0; k < 3; k++) m.od[k] = need_float(a, k, what);
            for (int k = 0; k < 3; k++) m.os[k] = need_float(a, 3 + k, what);
            m.ka = need_float(a, 6, what); m.kd = need_float(a, 7, what);
            m.ks = need_float(a, 8, what); m.n = need_float(a, 9, what);
            if (a.size() == 12) {                             // alpha/eta only with EXACTLY 12 args
                float al = need_float(a, 10, what);
                m.opacity = (al < 0.0f) ? 0.0f : (1.0f < al) ? 1.0f : al;   // std::clamp<float>
                m.eta = need_float(a, 11, what);
            } else {
                m.opacity = 1.0f; m.eta = 1.0f;
            }
            // (a scene that restates the material before every object would otherwise carry one
            // record per object: identical values share a record)
            const std::string key(reinterpret_cast<const char*>(&m), sizeof m);
            const auto known = material_by_value.find(key);
            if (known != material_by_value.end()) {
                current_material = known->second;
            } else {
                out->materials.push_back(m);
                current_material = (int)out->materials.size() - 1;
                material_by_value.emplace(key, current_material);
            }
            has_material = true;
        } else if (cmd == "texture") {                        // 309-327
            use_texture = true;
            std::string path(a[0]);
            if (!texture_dir.empty() && !path.empty() && path[0] != '/') path = texture_dir + "/" + path;
            const auto known = texture_by_path.find(path);
            if (known != texture_by_path.end()) {
                current_texture = known->second;
            } else {
                RtTexture t{};
                std::vector<uint8_t> px;
                if (read_texture_p3(path, &t.width, &t.height, &px) != RT_OK)
                    fail("ERROR: Issue reading 'texture' from ppm. Please verify. (" + g_error + ")");
                out->textures.push_back(t);
                out->texture_pixels.push_back(std::move(px));
                current_texture = (int)out->textures.size() - 1;
                texture_by_path.emplace(path, current_texture);
            }
        } else if (cmd == "sphere") {                         // 328-377
            obj_id++;
            RtSphere s{};
            s.id = (int32_t)obj_id;
            s.radius = need_float(a, 3, "'sphere' object");
            for (int k = 0; k < 3; k++) s.center[k] = need_float(a, k, "'sphere' object");
            material_and_texture(&s.material, &s.texture);
            out->spheres.push_back(s);
        } else if (cmd == "light") {                          // 378-411
            RtLight l{};
            l.w = need_float(a, 3, "'light' command");
            for (int k = 0; k < 3; k++) l.xyz[k] = need_float(a, k, "'light' command");
            for (int k = 0; k < 3; k++) l.rgb[k] = need_float(a, 4 + k, "'light' command");
            out->lights.push_back(l);
        }
        // v / vn / vt / f never get here: geometry() takes them
    }

    // The four geometry commands (main.cpp:412-552), which are all but a few lines of a
    // large scene file, on string views: no per-line allocations.
    bool geometry(Tok cmd, const Args& a, Tok line)
    {
        if (cmd == "v") {                                     // 412-430
            Vec3 p{ need_float(a, 0, "vertex"), need_float(a, 1, "vertex"), need_float(a, 2, "vertex") };
            vertices.push_back(p);
            return true;
        }
        if (cmd == "vn") {                                    // 431-449
            Vec3 p{ need_float(a, 0, "vertex normal"), need_float(a, 1, "vertex normal"), need_float(a, 2, "vertex normal") };
            normals.push_back(p);
            return true;
        }
        if (cmd == "vt") {                                    // 450-469
            Vec2 p{ need_float(a, 0, "texture coordinate"), need_float(a, 1, "texture coordinate") };
            texcoords.push_back(p);
            return true;
        }
        if (cmd != "f") return false;
        (void)line;
        obj_id++;                                             // 470-552
        RtFace f{};                                           // new Face() value-initialises: zeros
        f.id = (int32_t)obj_id;
        if (a.size() < 3) fail("ERROR: Invalid args for 'f' (face) object. Please verify.");
        for (int i = 0; i < 3; i++) {
            int v = 0, t = 0, n = 0;
            bool set_n = false, set_t = false;
            // Per-vertex format sniffing in the reference's order (487-517): "%d/%d/%d" must
            // convert 3, else "%d//%d" 2, else "%d/%d" 2, else "%d" 1.  Written out: a spec that
            // breaks off midway degrades to the longest form that still converts.
            const char* p = a[i].data();
            const char* e = p + a[i].size();
            bool exotic = false;
            int form = -1;                                    // 0: v, 1: v/t, 2: v//n, 3: v/t/n
            if (scan_int(p, e, &v, &exotic)) {
                form = 0;
                if (p < e && *p == '/') {
                    const char* q = p + 1;
                    if (q < e && *q == '/') {
                        q++;
                        if (scan_int(q, e, &n, &exotic)) form = 2;
                    } else if (scan_int(q, e, &t, &exotic)) {
                        form = 1;
                        if (q < e && *q == '/') { q++; if (scan_int(q, e, &n, &exotic)) form = 3; }
                    }
                }
            }
            if (exotic || (form < 0 && !a[i].empty() && (a[i][0] == ' ' || a[i][0] == '\t' || a[i][0] == '\r' ||
                                                         a[i][0] == '\v' || a[i][0] == '\f'))) {
                // out-of-int indices or leading blanks (sscanf skips them): ask sscanf itself
                const std::string z(a[i]);
                const char* s = z.c_str();
                v = t = n = 0;
                if (std::sscanf(s, "%d/%d/%d", &v, &t, &n) == 3) form = 3;
                else if (std::sscanf(s, "%d//%d", &v, &n) == 2) form = 2;
                else if (std::sscanf(s, "%d/%d", &v, &t) == 2) form = 1;
                else if (std::sscanf(s, "%d", &v) == 1) form = 0;
                else form = -1;
            }
            if (form < 0) fail("ERROR: Invalid args for 'f' object. Please verify.");
            f.smooth = (form >= 2) ? 1 : 0;
            set_n = form >= 2; set_t = (form == 1 || form == 3);
            const Vec3 pv = lookup(vertices, (unsigned)v);
            f.v[i][0] = pv.x; f.v[i][1] = pv.y; f.v[i][2] = pv.z;
            if (set_n) { const Vec3 nn = lookup(normals, (unsigned)n); f.vn[i][0] = nn.x; f.vn[i][1] = nn.y; f.vn[i][2] = nn.z; }
            if (set_t) { const Vec2 tc = lookup(texcoords, (unsigned)t); f.vt[i][0] = tc.x; f.vt[i][1] = tc.y; }
        }
        // has_texture is decided by the parser state, not by the face format (523-534)
        material_and_texture(&f.material, &f.texture);
        out->faces.push_back(f);
        return true;
    }

    void run(const char* data, size_t size)
    {
        std::vector<Tok> toks;
        const char* p = data;
        const char* const end = data + size;
        while (p < end) {                                     // main.cpp:106: std::getline on '\n'
            const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
            const char* le = nl ? nl : end;
            const Tok line(p, (size_t)(le - p));
            p = nl ? nl + 1 : end;
            toks.clear();
            const char* q = line.data();
            while (q < le) {                                  // 112: split on ' ' only
                const char* te = q;
                while (te < le && *te != ' ') te++;
                // 114: `!del.empty() || del.at(0) != '#'` -> an empty token throws
                // std::out_of_range outside every try block: the reference aborts.
                if (te == q)
                    fail("basic_string::at: empty token (double or leading space) in line '" + std::string(line) + "'");
                toks.emplace_back(q, (size_t)(te - q));
                q = te < le ? te + 1 : le;
            }
            if (toks.size() < 2) continue;                    // 119 / 136: no command or no arguments -> ignored
            const Tok cmd = toks[0];
            const Args args{ toks.data() + 1, toks.size() - 1 };
            if (geometry(cmd, args, line)) continue;
            command(cmd, args);                               // unknown commands fall through silently
        }
        // required commands, in the reference's order (574-602)
        if (!seen_imsize) fail("Error: Requires command 'imsize'");
        if (!seen_eye) fail("Error: Requires command 'eye'");
        if (!seen_viewdir) fail("Error: Requires command 'viewdir'");
        if (!seen_updir) fail("Error: Requires command 'updir'");
        if (!seen_hfov) fail("Error: Requires command 'hfov'");
        if (!seen_bkg) fail("Error: Requires command 'bkgcolor'");
    }
};

}  // namespace
}  // namespace rthost

using namespace rthost;

extern "C" int rt_host_parse_scene_file(const char* path, const char* texture_dir, RtSceneDesc** out)
{
    if (!path || !out) { set_error("rt_host_parse_scene_file: null argument"); return RT_E_INVALID; }
    *out = nullptr;
    std::vector<char> text;
    {
        FILE* fp = std::fopen(path, "rb");
        if (!fp) {                                            // main.cpp:567
            set_error(std::string("ERROR: Issue reading input file '") + path + "'. Please verify path.");
            return RT_E_IO;
        }
        std::fseek(fp, 0, SEEK_END);
        const long n = std::ftell(fp);
        std::fseek(fp, 0, SEEK_SET);
        bool ok = n >= 0;
        if (ok) {
            try { text.resize((size_t)n); } catch (const std::exception&) { ok = false; }
        }
        if (ok && n > 0) ok = std::fread(text.data(), 1, (size_t)n, fp) == (size_t)n;
        std::fclose(fp);
        if (!ok) {
            set_error(std::string("ERROR: Issue reading input file '") + path + "'. Please verify path.");
            return RT_E_IO;
        }
    }
    OwnedScene* scene = new (std::nothrow) OwnedScene();
    if (!scene) { set_error("out of memory"); return RT_E_NOMEM; }
    Parser p;
    p.out = scene;
    if (texture_dir) p.texture_dir = texture_dir;
    try {
        p.run(text.data(), text.size());
    } catch (const ParseError& e) {
        set_error(e.msg);
        delete scene;
        return RT_E_PARSE;
    } catch (const std::exception& e) {
        set_error(e.what());
        delete scene;
        return RT_E_NOMEM;
    }
    scene->seal();
    *out = &scene->desc;
    return RT_OK;
}

extern "C" void rt_host_free_scene_desc(RtSceneDesc* desc)
{
    delete reinterpret_cast<OwnedScene*>(desc);
}

extern "C" const char* rt_host_last_error(void) { return g_error.c_str(); }
