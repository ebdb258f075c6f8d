// scene_parser.cpp — the scene-file grammar of the reference, re-stated.
//
// Follows main.cpp:88-602 of the reference.  The grammar
// and its quirks are the contract; the implementation is a flat token switch
// writing straight into an RtSceneDesc instead of the reference's pointer-linked
// maps.  Where the reference aborts (uncaught exception -> std::terminate) or
// prints and returns without an image, we return RT_E_PARSE with a message.
#include "host_common.h"

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
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

float need_float(const std::vector<std::string>& a, size_t i, const char* what)
{
    float v;
    if (i >= a.size() || !parse_float(a[i], &v)) fail(std::string("ERROR: Invalid args for ") + what + ". Please verify.");
    return v;
}

int need_int(const std::vector<std::string>& a, size_t i, const char* what)
{
    int v;
    if (i >= a.size() || !parse_int(a[i], &v)) fail(std::string("ERROR: Invalid ") + what + ". Please verify.");
    return v;
}

// 1-based maps where a missing key reads as zeros (std::map::operator[],
// main.cpp:489-511)
template <class T>
T lookup(const std::unordered_map<unsigned, T>& m, unsigned key)
{
    auto it = m.find(key);
    return it == m.end() ? T{} : it->second;
}

struct Parser {
    OwnedScene* out;
    std::string texture_dir;

    std::unordered_map<unsigned, Vec3> vertices, normals;
    std::unordered_map<unsigned, Vec2> texcoords;
    unsigned obj_id = 0, n_v = 0, n_vn = 0, n_vt = 0;
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

    void command(const std::string& cmd, const std::vector<std::string>& a)
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
            out->materials.push_back(m);
            current_material = (int)out->materials.size() - 1;
            has_material = true;
        } else if (cmd == "texture") {                        // 309-327
            use_texture = true;
            std::string path = a[0];
            if (!texture_dir.empty() && !path.empty() && path[0] != '/') path = texture_dir + "/" + path;
            RtTexture t{};
            std::vector<uint8_t> px;
            if (read_texture_p3(path, &t.width, &t.height, &px) != RT_OK)
                fail("ERROR: Issue reading 'texture' from ppm. Please verify. (" + g_error + ")");
            out->textures.push_back(t);
            out->texture_pixels.push_back(std::move(px));
            current_texture = (int)out->textures.size() - 1;
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
        } else if (cmd == "v") {                              // 412-430
            n_v++;
            Vec3 p{ need_float(a, 0, "vertex"), need_float(a, 1, "vertex"), need_float(a, 2, "vertex") };
            vertices[n_v] = p;
        } else if (cmd == "vn") {                             // 431-449
            n_vn++;
            Vec3 p{ need_float(a, 0, "vertex normal"), need_float(a, 1, "vertex normal"), need_float(a, 2, "vertex normal") };
            normals[n_vn] = p;
        } else if (cmd == "vt") {                             // 450-469
            Vec2 p{ need_float(a, 0, "texture coordinate"), need_float(a, 1, "texture coordinate") };
            n_vt++;
            texcoords[n_vt] = p;
        } else if (cmd == "f") {                              // 470-552
            obj_id++;
            RtFace f{};                                       // new Face() value-initialises: zeros
            f.id = (int32_t)obj_id;
            if (a.size() < 3) fail("ERROR: Invalid args for 'f' (face) object. Please verify.");
            for (int i = 0; i < 3; i++) {
                unsigned t = 0, n = 0, v = 0;
                const char* s = a[i].c_str();
                Vec3 p{}, nn{}; Vec2 tc{};
                bool set_n = false, set_t = false;
                // per-vertex format sniffing, in the reference's order (487-517)
                if (std::sscanf(s, "%d/%d/%d", (int*)&v, (int*)&t, (int*)&n) == 3) {
                    f.smooth = 1; set_n = set_t = true;
                } else if (std::sscanf(s, "%d//%d", (int*)&v, (int*)&n) == 2) {
                    f.smooth = 1; set_n = true;
                } else if (std::sscanf(s, "%d/%d", (int*)&v, (int*)&t) == 2) {
                    f.smooth = 0; set_t = true;
                } else if (std::sscanf(s, "%d", (int*)&v) == 1) {
                    f.smooth = 0;
                } else {
                    fail("ERROR: Invalid args for 'f' object. Please verify.");
                }
                p = lookup(vertices, v);
                f.v[i][0] = p.x; f.v[i][1] = p.y; f.v[i][2] = p.z;
                if (set_n) { nn = lookup(normals, n); f.vn[i][0] = nn.x; f.vn[i][1] = nn.y; f.vn[i][2] = nn.z; }
                if (set_t) { tc = lookup(texcoords, t); f.vt[i][0] = tc.x; f.vt[i][1] = tc.y; }
            }
            // has_texture is decided by the parser state, not by the face format (523-534)
            material_and_texture(&f.material, &f.texture);
            out->faces.push_back(f);
        }
    }

    void run(std::istream& in)
    {
        std::string line;
        while (std::getline(in, line)) {                      // main.cpp:106
            std::vector<std::string> args;
            std::istringstream ss(line);
            std::string tok;
            while (std::getline(ss, tok, ' ')) {              // 112: split on ' ' only
                // 114: `!del.empty() || del.at(0) != '#'` -> an empty token throws
                // std::out_of_range outside every try block: the reference aborts.
                if (tok.empty())
                    fail("basic_string::at: empty token (double or leading space) in line '" + line + "'");
                args.push_back(tok);
            }
            if (args.empty()) continue;                       // 119
            std::string cmd = args[0];
            args.erase(args.begin());
            if (args.empty()) continue;                       // 136: no arguments -> ignored
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
    std::ifstream in(path);
    if (!in.is_open()) {                                      // main.cpp:567
        set_error(std::string("ERROR: Issue reading input file '") + path + "'. Please verify path.");
        return RT_E_IO;
    }
    OwnedScene* scene = new (std::nothrow) OwnedScene();
    if (!scene) { set_error("out of memory"); return RT_E_NOMEM; }
    Parser p;
    p.out = scene;
    if (texture_dir) p.texture_dir = texture_dir;
    try {
        p.run(in);
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
