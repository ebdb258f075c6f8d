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
#include <algorithm>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <chrono>
#include <string_view>
#include <thread>
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

// ---- the parse in three passes ------------------------------------------------------------
//
// The reference reads the file line by line into global state (main.cpp:106-565).  Almost all
// of a large scene file is v / vn / vt / f lines, whose meaning depends on earlier lines only
// through (a) how many v / vn / vt lines came before and (b) the last `mtlcolor` / `texture`
// command.  So the file is cut into chunks at line ends and
//   pass 1 (one thread per chunk): tokenises and converts every line of the chunk, keeping
//          vertex data in chunk-local arrays, objects with their RAW indices plus the local
//          vertex counts and the chunk-local last material / texture command at that line;
//   pass 2 (sequential, touches only per-chunk summaries): global vertex offsets, material and
//          texture numbering in file order, the parser state entering each chunk, camera and
//          light commands in order;
//   pass 3 (one thread per chunk): resolves every object against the global tables.
// Errors abort the reference at the first offending line, so every pass records the position
// of what it finds and the earliest one is reported.  One chunk = the sequential parse.

struct Pos {
    size_t chunk = ~(size_t)0, line = 0;
    bool valid() const { return chunk != ~(size_t)0; }
    bool before(const Pos& o) const { return !o.valid() || (valid() && (chunk < o.chunk || (chunk == o.chunk && line < o.line))); }
};

struct StateRef {          // the last mtlcolor / texture command of THIS chunk before an object (-1: none)
    int mtl = -1, tex = -1;
    bool tex_later = false; // the texture command came after the mtlcolor command
};

struct RawSphere { RtSphere s; StateRef st; size_t line; };
struct RawFace {
    int idx[3][3];         // per vertex: v, t, n as written
    uint8_t form[3];       // 0: v, 1: v/t, 2: v//n, 3: v/t/n
    uint32_t nv, nvn, nvt; // v / vn / vt lines of this chunk before the face
    StateRef st;
    size_t line;
    uint32_t local_obj;    // objects (spheres + faces) of this chunk before it
};

enum MiscKind : uint8_t { kEye, kViewdir, kUpdir, kHfov, kImsize, kBkg3, kBkg4, kLight };
struct Misc { MiscKind kind; float f[7]; int i[2]; };

struct Chunk {
    const char* begin = nullptr;
    const char* end = nullptr;
    std::vector<Vec3> v, vn;
    std::vector<Vec2> vt;
    std::vector<RtMaterial> mtl_unique;          // distinct records in order of first appearance
    std::vector<int> mtl_cmd;                    // per mtlcolor command: index into mtl_unique
    std::vector<std::string> tex_unique;         // distinct paths in order of first appearance
    std::vector<int> tex_cmd;                    // per texture command: index into tex_unique
    std::vector<size_t> tex_first_line;          // per distinct path: line of its first command
    std::vector<RawSphere> spheres;
    std::vector<RawFace> faces;
    std::vector<Misc> misc;                      // camera commands and lights, in order
    uint32_t n_obj = 0;
    bool out_use_texture = false;                // use_texture after the chunk's last mtlcolor / texture command
    Pos err_pos; std::string err;                // first failing line of pass 1 (parsing stops there)
    // filled by pass 2
    size_t v0 = 0, vn0 = 0, vt0 = 0, sphere0 = 0, face0 = 0; uint32_t obj0 = 0;
    std::vector<int> mtl_global, tex_global;     // per distinct record / path
    int in_material = -1, in_texture = -1; bool in_use_texture = false, in_has_material = false;
};

struct ChunkParser {
    Chunk& c;
    size_t chunk_index;
    const std::string& texture_dir;
    std::unordered_map<std::string, int> mtl_by_value, tex_by_path;
    int last_mtl = -1, last_tex = -1; bool tex_later = false;
    size_t line_no = 0;

    StateRef state() const { StateRef r; r.mtl = last_mtl; r.tex = last_tex; r.tex_later = tex_later; return r; }

    void command(Tok cmd, const Args& a)
    {
        Misc m{};
        if (cmd == "eye") {                                   // main.cpp:143-162
            m.kind = kEye; for (int k = 0; k < 3; k++) m.f[k] = need_float(a, k, "'eye' command");
            c.misc.push_back(m);
        } else if (cmd == "viewdir") {                        // 163-181
            m.kind = kViewdir; for (int k = 0; k < 3; k++) m.f[k] = need_float(a, k, "'viewdir' command");
            c.misc.push_back(m);
        } else if (cmd == "updir") {                          // 182-200
            m.kind = kUpdir; for (int k = 0; k < 3; k++) m.f[k] = need_float(a, k, "'updir' command");
            c.misc.push_back(m);
        } else if (cmd == "hfov") {                           // 201-215
            m.kind = kHfov; m.f[0] = need_float(a, 0, "'hfov' command");
            c.misc.push_back(m);
        } else if (cmd == "imsize") {                         // 216-246: height first, then width
            m.kind = kImsize;
            m.i[1] = need_int(a, 1, "image dimensions");
            m.i[0] = need_int(a, 0, "image dimensions");
            if (m.i[1] <= 1 || m.i[0] <= 1) fail("ERROR: Invalid image dimensions. Please verify.");
            c.misc.push_back(m);
        } else if (cmd == "bkgcolor") {                       // 247-271
            m.kind = kBkg3; for (int k = 0; k < 3; k++) m.f[k] = need_float(a, k, "'bkgcolor' command");
            if (a.size() > 3) { m.kind = kBkg4; m.f[3] = need_float(a, 3, "'bkgcolor' command"); }
            c.misc.push_back(m);
        } else if (cmd == "light") {                          // 378-411
            m.kind = kLight;
            m.f[3] = need_float(a, 3, "'light' command");
            for (int k = 0; k < 3; k++) m.f[k] = need_float(a, k, "'light' command");
            for (int k = 0; k < 3; k++) m.f[4 + k] = need_float(a, 4 + k, "'light' command");
            c.misc.push_back(m);
        } else if (cmd == "mtlcolor") {                       // 272-308
            RtMaterial r{};
            const char* what = "'material'";
            for (int k = 0; k < 3; k++) r.od[k] = need_float(a, k, what);
            for (int k = 0; k < 3; k++) r.os[k] = need_float(a, 3 + k, what);
            r.ka = need_float(a, 6, what); r.kd = need_float(a, 7, what);
            r.ks = need_float(a, 8, what); r.n = need_float(a, 9, what);
            if (a.size() == 12) {                             // alpha/eta only with EXACTLY 12 args
                const float al = need_float(a, 10, what);
                r.opacity = (al < 0.0f) ? 0.0f : (1.0f < al) ? 1.0f : al;   // std::clamp<float>
                r.eta = need_float(a, 11, what);
            } else {
                r.opacity = 1.0f; r.eta = 1.0f;
            }
            // a scene that restates the material before every object would otherwise carry one
            // record per object: identical values share a record
            const std::string key(reinterpret_cast<const char*>(&r), sizeof r);
            auto it = mtl_by_value.find(key);
            if (it == mtl_by_value.end()) {
                c.mtl_unique.push_back(r);
                it = mtl_by_value.emplace(key, (int)c.mtl_unique.size() - 1).first;
            }
            c.mtl_cmd.push_back(it->second);
            last_mtl = (int)c.mtl_cmd.size() - 1;
            tex_later = false;                                // use_texture = false
        } else if (cmd == "texture") {                        // 309-327
            // The reference reads the file again at every `texture` command (main.cpp:318); the
            // same path gives the same texels, so it is read once (pass 2).
            std::string path(a[0]);
            if (!texture_dir.empty() && !path.empty() && path[0] != '/') path = texture_dir + "/" + path;
            auto it = tex_by_path.find(path);
            if (it == tex_by_path.end()) {
                c.tex_unique.push_back(path);
                c.tex_first_line.push_back(line_no);
                it = tex_by_path.emplace(path, (int)c.tex_unique.size() - 1).first;
            }
            c.tex_cmd.push_back(it->second);
            last_tex = (int)c.tex_cmd.size() - 1;
            tex_later = true;                                 // use_texture = true
        } else if (cmd == "sphere") {                         // 328-377
            RawSphere r{};
            r.s.id = (int32_t)(++c.n_obj);                    // chunk-local for now
            r.s.radius = need_float(a, 3, "'sphere' object");
            for (int k = 0; k < 3; k++) r.s.center[k] = need_float(a, k, "'sphere' object");
            r.st = state(); r.line = line_no;
            c.spheres.push_back(r);
        }
        // unknown commands fall through silently; v / vn / vt / f never get here
    }

    // The four geometry commands (main.cpp:412-552), which are all but a few lines of a large
    // scene file.
    bool geometry(Tok cmd, const Args& a)
    {
        if (cmd == "v") {                                     // 412-430
            c.v.push_back(Vec3{ need_float(a, 0, "vertex"), need_float(a, 1, "vertex"), need_float(a, 2, "vertex") });
            return true;
        }
        if (cmd == "vn") {                                    // 431-449
            c.vn.push_back(Vec3{ need_float(a, 0, "vertex normal"), need_float(a, 1, "vertex normal"), need_float(a, 2, "vertex normal") });
            return true;
        }
        if (cmd == "vt") {                                    // 450-469
            c.vt.push_back(Vec2{ need_float(a, 0, "texture coordinate"), need_float(a, 1, "texture coordinate") });
            return true;
        }
        if (cmd != "f") return false;
        RawFace f{};                                          // 470-552
        f.local_obj = c.n_obj++;
        if (a.size() < 3) fail("ERROR: Invalid args for 'f' (face) object. Please verify.");
        for (int i = 0; i < 3; i++) {
            int v = 0, t = 0, n = 0;
            // Per-vertex format sniffing in the reference's order (487-517): "%d/%d/%d" must
            // convert 3, else "%d//%d" 2, else "%d/%d" 2, else "%d" 1.  Written out: a spec that
            // breaks off midway degrades to the longest form that still converts.
            const char* p = a[i].data();
            const char* e = p + a[i].size();
            bool exotic = false;
            int form = -1;
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
            if (exotic || (form < 0 && !a[i].empty() && (a[i][0] == '\t' || a[i][0] == '\r' || a[i][0] == '\v' || a[i][0] == '\f'))) {
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
            f.idx[i][0] = v; f.idx[i][1] = t; f.idx[i][2] = n; f.form[i] = (uint8_t)form;
        }
        f.nv = (uint32_t)c.v.size(); f.nvn = (uint32_t)c.vn.size(); f.nvt = (uint32_t)c.vt.size();
        f.st = state(); f.line = line_no;
        c.faces.push_back(f);
        return true;
    }

    void run()
    {
        std::vector<Tok> toks;
        const char* p = c.begin;
        const char* const end = c.end;
        try {
            for (; p < end; line_no++) {                      // main.cpp:106: std::getline on '\n'
                const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
                const char* le = nl ? nl : end;
                const Tok line(p, (size_t)(le - p));
                p = nl ? nl + 1 : end;
                toks.clear();
                const char* q = line.data();
                while (q < le) {                              // 112: split on ' ' only
                    const char* te = q;
                    while (te < le && *te != ' ') te++;
                    // 114: `!del.empty() || del.at(0) != '#'` -> an empty token throws
                    // std::out_of_range outside every try block: the reference aborts.
                    if (te == q)
                        fail("basic_string::at: empty token (double or leading space) in line '" + std::string(line) + "'");
                    toks.emplace_back(q, (size_t)(te - q));
                    q = te < le ? te + 1 : le;
                }
                if (toks.size() < 2) continue;                // 119 / 136: no command or no arguments -> ignored
                const Tok cmd = toks[0];
                const Args args{ toks.data() + 1, toks.size() - 1 };
                if (!geometry(cmd, args)) command(cmd, args);
            }
        } catch (const ParseError& e) {
            c.err_pos.chunk = chunk_index; c.err_pos.line = line_no;
            c.err = e.msg;
        }
        c.out_use_texture = tex_later;
    }
};

struct SceneBuilder {
    OwnedScene* out;
    std::string texture_dir;
    std::vector<Chunk> chunks;
    std::vector<Vec3> v, vn;       // global tables, slot 0 unused (the reference's maps are 1-based)
    std::vector<Vec2> vt;
    Pos err_pos; std::string err;  // earliest error found so far

    void note(const Pos& p, const std::string& msg)
    {
        if (p.before(err_pos)) { err_pos = p; err = msg; }
    }

    // objects capture the CURRENT material and texture (main.cpp:351-365 / 521-534)
    static const char* resolve_state(const Chunk& c, const StateRef& st, int32_t* mat, int32_t* tex)
    {
        const bool has_material = st.mtl >= 0 || c.in_has_material;
        // use_texture: set by `texture`, cleared by `mtlcolor`; whichever came last decides
        const bool use_texture = (st.mtl < 0 && st.tex < 0) ? c.in_use_texture : (st.tex >= 0 && (st.mtl < 0 || st.tex_later));
        const int material = st.mtl >= 0 ? c.mtl_global[c.mtl_cmd[st.mtl]] : c.in_material;
        const int texture = st.tex >= 0 ? c.tex_global[c.tex_cmd[st.tex]] : c.in_texture;
        if (use_texture) {
            if (!has_material || texture < 0) return "ERROR: Must define a 'mtlcolor' and 'texture'. Please verify.";
            *tex = texture;
        } else {
            if (!has_material) return "ERROR: Must define a 'mtlcolor'. Please verify.";
            *tex = -1;
        }
        *mat = material;
        return nullptr;
    }

    // pass 2: everything that needs the chunks in order, on per-chunk summaries only
    void link(bool* seen /* imsize, eye, viewdir, updir, hfov, bkg */)
    {
        RtSceneDesc& d = out->desc;
        std::unordered_map<std::string, int> mtl_by_value, tex_by_path;
        size_t nv = 1, nvn = 1, nvt = 1, ns = 0, nf = 0; uint32_t nobj = 0;
        int cur_mtl = -1, cur_tex = -1; bool use_tex = false, has_mtl = false;
        for (size_t k = 0; k < chunks.size(); k++) {
            Chunk& c = chunks[k];
            c.v0 = nv; c.vn0 = nvn; c.vt0 = nvt; c.sphere0 = ns; c.face0 = nf; c.obj0 = nobj;
            nv += c.v.size(); nvn += c.vn.size(); nvt += c.vt.size();
            ns += c.spheres.size(); nf += c.faces.size(); nobj += c.n_obj;
            c.in_material = cur_mtl; c.in_texture = cur_tex; c.in_use_texture = use_tex; c.in_has_material = has_mtl;
            c.mtl_global.resize(c.mtl_unique.size());
            for (size_t i = 0; i < c.mtl_unique.size(); i++) {
                const std::string key(reinterpret_cast<const char*>(&c.mtl_unique[i]), sizeof(RtMaterial));
                auto it = mtl_by_value.find(key);
                if (it == mtl_by_value.end()) {
                    out->materials.push_back(c.mtl_unique[i]);
                    it = mtl_by_value.emplace(key, (int)out->materials.size() - 1).first;
                }
                c.mtl_global[i] = it->second;
            }
            c.tex_global.assign(c.tex_unique.size(), -1);
            for (size_t i = 0; i < c.tex_unique.size(); i++) {
                auto it = tex_by_path.find(c.tex_unique[i]);
                if (it == tex_by_path.end()) {
                    Pos at; at.chunk = k; at.line = c.tex_first_line[i];
                    if (!at.before(err_pos)) continue;        // the parse has already failed before this line
                    RtTexture t{};
                    std::vector<uint8_t> px;
                    if (read_texture_p3(c.tex_unique[i], &t.width, &t.height, &px) != RT_OK) {
                        note(at, "ERROR: Issue reading 'texture' from ppm. Please verify. (" + g_error + ")");
                        continue;
                    }
                    out->textures.push_back(t);
                    out->texture_pixels.push_back(std::move(px));
                    it = tex_by_path.emplace(c.tex_unique[i], (int)out->textures.size() - 1).first;
                }
                c.tex_global[i] = it->second;
            }
            if (!c.mtl_cmd.empty()) { cur_mtl = c.mtl_global[c.mtl_cmd.back()]; has_mtl = true; }
            if (!c.tex_cmd.empty()) cur_tex = c.tex_global[c.tex_cmd.back()];
            // use_texture after the chunk: decided by its last mtlcolor / texture command
            if (!c.mtl_cmd.empty() || !c.tex_cmd.empty()) {
                // (ChunkParser kept the order of the two in tex_later at every step; the final value
                // is what the last object-independent state would be)
                use_tex = c.out_use_texture;
            }
            for (const Misc& m : c.misc) {
                switch (m.kind) {
                case kEye: for (int i = 0; i < 3; i++) d.eye[i] = m.f[i]; seen[1] = true; break;
                case kViewdir: for (int i = 0; i < 3; i++) d.viewdir[i] = m.f[i]; seen[2] = true; break;
                case kUpdir: for (int i = 0; i < 3; i++) d.updir[i] = m.f[i]; seen[3] = true; break;
                case kHfov: d.hfov_deg = m.f[0]; seen[4] = true; break;
                case kImsize: d.width = m.i[0]; d.height = m.i[1]; seen[0] = true; break;
                case kBkg4: d.bkg_eta = m.f[3]; /* fall through */
                case kBkg3: for (int i = 0; i < 3; i++) d.bkg[i] = m.f[i]; seen[5] = true; break;
                case kLight: {
                    RtLight l{};
                    l.w = m.f[3];
                    for (int i = 0; i < 3; i++) { l.xyz[i] = m.f[i]; l.rgb[i] = m.f[4 + i]; }
                    out->lights.push_back(l);
                    break; }
                }
            }
            if (c.err_pos.valid()) { note(c.err_pos, c.err); break; }   // nothing after the failing line exists
        }
        v.assign(nv, Vec3{}); vn.assign(nvn, Vec3{}); vt.assign(nvt, Vec2{});
        out->spheres.resize(ns); out->faces.resize(nf);
    }

    // pass 3 for one chunk
    void resolve(size_t k, Pos* first_err, std::string* first_msg)
    {
        const Chunk& c = chunks[k];
        auto fail_at = [&](size_t line, const char* msg) {
            Pos p; p.chunk = k; p.line = line;
            if (p.before(*first_err)) { *first_err = p; *first_msg = msg; }
        };
        for (size_t i = 0; i < c.spheres.size(); i++) {
            RtSphere s = c.spheres[i].s;
            s.id += (int32_t)c.obj0;
            if (const char* msg = resolve_state(c, c.spheres[i].st, &s.material, &s.texture)) fail_at(c.spheres[i].line, msg);
            out->spheres[c.sphere0 + i] = s;
        }
        for (size_t i = 0; i < c.faces.size(); i++) {
            const RawFace& r = c.faces[i];
            RtFace f{};                                       // new Face() value-initialises: zeros
            f.id = (int32_t)(c.obj0 + r.local_obj + 1);
            for (int j = 0; j < 3; j++) {
                const int form = r.form[j];
                f.smooth = (form >= 2) ? 1 : 0;               // the last vertex decides, like the reference's loop
                // a key the file has not defined YET reads as zeros (std::map::operator[])
                const unsigned iv = (unsigned)r.idx[j][0], it = (unsigned)r.idx[j][1], in = (unsigned)r.idx[j][2];
                const Vec3 pv = (iv >= 1 && (size_t)iv < c.v0 + r.nv) ? v[iv] : Vec3{};
                f.v[j][0] = pv.x; f.v[j][1] = pv.y; f.v[j][2] = pv.z;
                if (form >= 2) {
                    const Vec3 nn = (in >= 1 && (size_t)in < c.vn0 + r.nvn) ? vn[in] : Vec3{};
                    f.vn[j][0] = nn.x; f.vn[j][1] = nn.y; f.vn[j][2] = nn.z;
                }
                if (form == 1 || form == 3) {
                    const Vec2 tc = (it >= 1 && (size_t)it < c.vt0 + r.nvt) ? vt[it] : Vec2{};
                    f.vt[j][0] = tc.x; f.vt[j][1] = tc.y;
                }
            }
            // has_texture is decided by the parser state, not by the face format (523-534)
            if (const char* msg = resolve_state(c, r.st, &f.material, &f.texture)) fail_at(r.line, msg);
            out->faces[c.face0 + i] = f;
        }
    }
};

// Runs fn(k) for k in [0, n) on up to `threads` threads.
template <class Fn>
void parallel_for(size_t n, unsigned threads, Fn fn)
{
    if (threads <= 1 || n <= 1) { for (size_t k = 0; k < n; k++) fn(k); return; }
    std::atomic<size_t> next{ 0 };
    auto work = [&]() { for (size_t k; (k = next.fetch_add(1)) < n;) fn(k); };
    std::vector<std::thread> pool;
    const unsigned extra = (unsigned)std::min<size_t>(threads, n) - 1;
    for (unsigned t = 0; t < extra; t++) pool.emplace_back(work);
    work();
    for (std::thread& t : pool) t.join();
}

unsigned env_unsigned(const char* name, unsigned dflt)
{
    const char* e = std::getenv(name);
    if (!e || !*e) return dflt;
    const long v = std::strtol(e, nullptr, 10);
    return v > 0 ? (unsigned)v : dflt;
}

// The whole parse of a file image.  Throws ParseError with the reference's message.
void parse_text(const char* data, size_t size, const std::string& texture_dir, OwnedScene* out)
{
    // RT_HOST_THREADS / RT_HOST_CHUNK_BYTES: for tests (tiny chunks on tiny files) and tuning
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned threads = std::min(env_unsigned("RT_HOST_THREADS", std::min(hw, 32u)), 256u);
    const size_t chunk_bytes = std::max<size_t>(1, env_unsigned("RT_HOST_CHUNK_BYTES", 4u << 20));
    size_t n_chunks = std::max<size_t>(1, std::min<size_t>((size + chunk_bytes - 1) / chunk_bytes, (size_t)threads * 4));
    if (threads == 1) n_chunks = 1;

    SceneBuilder b;
    b.out = out; b.texture_dir = texture_dir;
    b.chunks.resize(n_chunks);
    const char* const end = data + size;
    const char* p = data;
    size_t used = 0;
    for (size_t k = 0; k < n_chunks && p < end; k++) {
        const char* q = (k + 1 == n_chunks) ? end : std::min(end, data + (size / n_chunks) * (k + 1));
        if (q < p) q = p;
        if (q < end) {                                        // cut after a line end
            const char* nl = (const char*)std::memchr(q, '\n', (size_t)(end - q));
            q = nl ? nl + 1 : end;
        }
        b.chunks[used].begin = p; b.chunks[used].end = q;
        used++;
        p = q;
    }
    b.chunks.resize(std::max<size_t>(used, 1));
    if (used == 0) { b.chunks[0].begin = b.chunks[0].end = data; }

    const bool verbose = std::getenv("RT_HOST_VERBOSE") != nullptr;
    auto now = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    parallel_for(b.chunks.size(), threads, [&](size_t k) {
        ChunkParser cp{ b.chunks[k], k, b.texture_dir, {}, {} };
        cp.run();
    });
    const double t1 = now();
    bool seen[6] = { false, false, false, false, false, false };
    b.link(seen);
    const double t2 = now();
    // chunks after a failing one are ignored: the reference never got there
    size_t live = b.chunks.size();
    for (size_t k = 0; k < b.chunks.size(); k++) if (b.chunks[k].err_pos.valid()) { live = k + 1; break; }
    parallel_for(live, threads, [&](size_t k) {
        const Chunk& c = b.chunks[k];
        if (!c.v.empty()) std::memcpy(&b.v[c.v0], c.v.data(), c.v.size() * sizeof(Vec3));
        if (!c.vn.empty()) std::memcpy(&b.vn[c.vn0], c.vn.data(), c.vn.size() * sizeof(Vec3));
        if (!c.vt.empty()) std::memcpy(&b.vt[c.vt0], c.vt.data(), c.vt.size() * sizeof(Vec2));
    });
    std::vector<Pos> epos(live); std::vector<std::string> emsg(live);
    parallel_for(live, threads, [&](size_t k) { b.resolve(k, &epos[k], &emsg[k]); });
    for (size_t k = 0; k < live; k++) if (epos[k].valid()) b.note(epos[k], emsg[k]);
    if (verbose)
        std::fprintf(stderr, "parse_text: %zu chunks, %u threads: tokenise %.1f ms, link %.1f ms, resolve %.1f ms\n",
                     b.chunks.size(), threads, t1 - t0, t2 - t1, now() - t2);
    if (b.err_pos.valid()) fail(b.err);
    // required commands, in the reference's order (574-602)
    if (!seen[0]) fail("Error: Requires command 'imsize'");
    if (!seen[1]) fail("Error: Requires command 'eye'");
    if (!seen[2]) fail("Error: Requires command 'viewdir'");
    if (!seen[3]) fail("Error: Requires command 'updir'");
    if (!seen[4]) fail("Error: Requires command 'hfov'");
    if (!seen[5]) fail("Error: Requires command 'bkgcolor'");
}

}  // namespace
}  // namespace rthost

using namespace rthost;

extern "C" int rt_host_parse_scene_file(const char* path, const char* texture_dir, RtSceneDesc** out)
{
    if (!path || !out) { set_error("rt_host_parse_scene_file: null argument"); return RT_E_INVALID; }
    *out = nullptr;
    // the file image: mapped (no copy, no zero-fill of a 250 MB buffer), read() as the fallback
    const char* text_data = nullptr; size_t text_size = 0;
    void* mapped = nullptr; std::vector<char> text;
    {
        const int fd = ::open(path, O_RDONLY);
        struct stat st;
        if (fd < 0 || ::fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) {                      // main.cpp:567
            if (fd >= 0) ::close(fd);
            set_error(std::string("ERROR: Issue reading input file '") + path + "'. Please verify path.");
            return RT_E_IO;
        }
        text_size = (size_t)st.st_size;
        if (text_size > 0) {
            mapped = ::mmap(nullptr, text_size, PROT_READ, MAP_PRIVATE, fd, 0);
            if (mapped == MAP_FAILED) {
                mapped = nullptr;
                bool ok = true;
                try { text.resize(text_size); } catch (const std::exception&) { ok = false; }
                size_t got = 0;
                while (ok && got < text_size) {
                    const ssize_t r = ::read(fd, text.data() + got, text_size - got);
                    if (r <= 0) ok = false; else got += (size_t)r;
                }
                if (!ok) {
                    ::close(fd);
                    set_error(std::string("ERROR: Issue reading input file '") + path + "'. Please verify path.");
                    return RT_E_IO;
                }
                text_data = text.data();
            } else {
                text_data = static_cast<const char*>(mapped);
            }
        }
        ::close(fd);
    }
    struct Unmap { void* p; size_t n; ~Unmap() { if (p) ::munmap(p, n); } } unmap{ mapped, text_size };
    OwnedScene* scene = new (std::nothrow) OwnedScene();
    if (!scene) { set_error("out of memory"); return RT_E_NOMEM; }
    try {
        parse_text(text_data, text_size, texture_dir ? std::string(texture_dir) : std::string(), scene);
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
