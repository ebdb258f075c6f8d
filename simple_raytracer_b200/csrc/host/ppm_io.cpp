// ppm_io.cpp — P3 texture loader and P3 frame writer of the host front-end.
//
//   loader: src/utility.h:59-139 of the reference
//   writer: main.cpp:613-650
//
// Both keep the reference's observable format rules but work on flat buffers:
// one pass over the file bytes for the loader (the reference builds a
// std::vector<std::string> of every token), one table-driven pass for the
// writer (the reference calls std::to_string and flushes with std::endl per
// pixel).
#include "host_common.h"

#include <algorithm>
#include <atomic>
#include <cctype>
#include <chrono>
#include <thread>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace rthost {

namespace {

bool read_file(const std::string& path, std::vector<char>* data)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return false; }
    data->resize((size_t)n);
    size_t got = n ? std::fread(data->data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == (size_t)n;
}

// std::stoi on the token [b, e): skip leading isspace, optional sign, decimal
// digits, trailing junk ignored.  false when there is no digit.
bool stoi_token(const char* b, const char* e, long* out)
{
    while (b < e && std::isspace((unsigned char)*b)) b++;
    bool neg = false;
    if (b < e && (*b == '+' || *b == '-')) { neg = (*b == '-'); b++; }
    if (b >= e || *b < '0' || *b > '9') return false;
    long v = 0;
    while (b < e && *b >= '0' && *b <= '9') {
        v = v * 10 + (*b - '0');
        if (v > 0x7fffffffL) return false;   // std::out_of_range in the reference
        b++;
    }
    *out = neg ? -v : v;
    return true;
}

}  // namespace

namespace {

// One chunk of the pixel part of a P3 file (utility.h:70-112): lines starting with '#' are
// skipped, an empty line aborts (line.at(0) throws), tokens are split on ' ' only, empty
// tokens are dropped, every other token goes through std::stoi.  Values are appended to
// `out` up to the first token that does not convert, whose index and message are kept — the
// caller decides whether that token is ever reached (the reference splits the WHOLE file into
// tokens first, where an empty line is fatal wherever it is, and then converts only the first
// width*height*3 of them).
struct BodyChunk {
    const char* begin = nullptr;
    const char* end = nullptr;
    bool at_line_start = true;
    std::vector<uint8_t> out;
    const char* err_at = nullptr;      // a token that does not convert / is out of range
    size_t err_token = 0;              // ... and its index among this chunk's value tokens
    const char* err_msg = nullptr;
    const char* empty_line_at = nullptr;
};

void parse_body_chunk(BodyChunk& c)
{
    const char* p = c.begin;
    const char* const end = c.end;
    bool line_start = c.at_line_start;
    c.out.reserve((size_t)(end - p) / 3);
    while (p < end) {
        if (line_start) {
            if (*p == '\n') { c.empty_line_at = p; return; }                          // utility.h:77 on ""
            if (*p == '#') {                                                        // comment line
                const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
                p = nl ? nl + 1 : end;
                continue;
            }
            line_start = false;
        }
        const char ch = *p;
        if (ch == ' ') { p++; continue; }                                           // utility.h:85: empty tokens dropped
        if (ch == '\n') { p++; line_start = true; continue; }
        if (c.err_at) {                                                             // only empty lines matter from here on
            const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
            p = nl ? nl : end;
            continue;
        }
        // a token: the common case is 1-3 digits
        const char* t = p;
        unsigned v = 0;
        while (p < end && (unsigned)(*p - '0') <= 9u && p - t < 4) { v = v * 10u + (unsigned)(*p - '0'); p++; }
        if (p > t && (p == end || *p == ' ' || *p == '\n') ) {
            if (v > 255u) { c.err_at = t; c.err_token = c.out.size(); c.err_msg = "value outside 0..255"; continue; }
            c.out.push_back((uint8_t)v);
            continue;
        }
        // anything else: std::stoi's rules (leading blanks such as '\r' or '\t', sign, trailing junk)
        const char* te = t;
        while (te < end && *te != ' ' && *te != '\n') te++;
        long lv;
        p = te;
        if (!stoi_token(t, te, &lv)) { c.err_at = t; c.err_token = c.out.size(); c.err_msg = "stoi"; continue; }
        if (lv < 0 || lv > 255) { c.err_at = t; c.err_token = c.out.size(); c.err_msg = "value outside 0..255"; continue; }
        c.out.push_back((uint8_t)lv);
    }
}

unsigned host_threads()
{
    unsigned n = std::max(1u, std::thread::hardware_concurrency());
    if (const char* e = std::getenv("RT_HOST_THREADS")) { const long v = std::strtol(e, nullptr, 10); if (v > 0) n = (unsigned)v; }
    return std::min(n, 64u);
}

}  // namespace

int read_texture_p3(const std::string& path, int32_t* width, int32_t* height, std::vector<uint8_t>* rgb)
{
    const bool verbose = std::getenv("RT_HOST_VERBOSE") != nullptr;
    auto now = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = now();
    std::vector<char> data;
    if (!read_file(path, &data)) {
        // the reference does not notice (width/height stay uninitialised, utility.h:62-67,115)
        set_error("cannot open texture '" + path + "'");
        return RT_E_IO;
    }
    const char* p = data.data();
    const char* end = p + data.size();
    long w = -1, h = -1;
    size_t token_no = 0, need = 0;
    const char* body = nullptr;        // first byte after the "255" token
    bool body_at_line_start = false;
    // header: "P3", width, height, "255" — the first four non-empty tokens outside comment lines
    while (p < end && !body) {
        const char* eol = (const char*)std::memchr(p, '\n', (size_t)(end - p));
        const char* line_end = eol ? eol : end;
        if (line_end == p) {                                  // utility.h:77 line.at(0) on "" throws
            set_error("texture '" + path + "': empty line (basic_string::at)");
            return RT_E_PARSE;
        }
        if (*p != '#') {                                      // utility.h:77 comment lines
            const char* t = p;
            while (t < line_end && !body) {
                const char* te = (const char*)std::memchr(t, ' ', (size_t)(line_end - t));
                if (!te) te = line_end;
                if (te > t) {                                 // utility.h:85 non-empty tokens only
                    token_no++;
                    if (token_no == 1) {
                        if (!(te - t == 2 && t[0] == 'P' && t[1] == '3')) {
                            set_error("Only supports PPM 'P3' file format.");
                            return RT_E_PARSE;
                        }
                    } else if (token_no == 2) {
                        if (!stoi_token(t, te, &w)) { set_error("texture '" + path + "': bad width"); return RT_E_PARSE; }
                    } else if (token_no == 3) {
                        if (!stoi_token(t, te, &h)) { set_error("texture '" + path + "': bad height"); return RT_E_PARSE; }
                        if (w <= 0 || h <= 0 || (double)w * (double)h > 1.0e9) {
                            set_error("texture '" + path + "': bad dimensions");
                            return RT_E_PARSE;
                        }
                        need = (size_t)w * (size_t)h * 3;
                    } else {
                        if (!(te - t == 3 && t[0] == '2' && t[1] == '5' && t[2] == '5')) {
                            set_error("PPM pixel value must be between 0 - 255 .");
                            return RT_E_PARSE;
                        }
                        body = te;
                    }
                }
                t = te < line_end ? te + 1 : line_end;
            }
        }
        if (!body) p = eol ? eol + 1 : end;
    }
    if (!body) { set_error("texture '" + path + "': too few tokens"); return RT_E_PARSE; }

    // pixel values: chunks cut after line ends, one thread each
    const double t_read = now();
    const unsigned threads = host_threads();
    const size_t body_bytes = (size_t)(end - body);
    size_t n_chunks = std::max<size_t>(1, std::min<size_t>(body_bytes / (256u << 10), (size_t)threads * 4));
    if (const char* e = std::getenv("RT_HOST_CHUNK_BYTES")) {
        const long cb = std::strtol(e, nullptr, 10);
        if (cb > 0) n_chunks = std::max<size_t>(1, std::min<size_t>(body_bytes / (size_t)cb + 1, 4096));
    }
    if (threads == 1) n_chunks = 1;
    std::vector<BodyChunk> chunks;
    {
        const char* q = body;
        for (size_t k = 0; k < n_chunks && q < end; k++) {
            const char* r = (k + 1 == n_chunks) ? end : std::min(end, body + (body_bytes / n_chunks) * (k + 1));
            if (r < q) r = q;
            if (r < end) {
                const char* nl = (const char*)std::memchr(r, '\n', (size_t)(end - r));
                r = nl ? nl + 1 : end;
            }
            BodyChunk c;
            c.begin = q; c.end = r; c.at_line_start = (k > 0) || body_at_line_start;
            chunks.push_back(std::move(c));
            q = r;
        }
    }
    {
        std::atomic<size_t> next{ 0 };
        auto work = [&]() { for (size_t k; (k = next.fetch_add(1)) < chunks.size();) parse_body_chunk(chunks[k]); };
        std::vector<std::thread> pool;
        const unsigned extra = chunks.size() > 1 ? (unsigned)std::min<size_t>(threads, chunks.size()) - 1 : 0;
        for (unsigned t = 0; t < extra; t++) pool.emplace_back(work);
        work();
        for (std::thread& t : pool) t.join();
    }
    const double t_parsed = now();
    for (const BodyChunk& c : chunks)
        if (c.empty_line_at) {
            set_error("texture '" + path + "': empty line (basic_string::at)");
            return RT_E_PARSE;
        }
    // in file order: a token that does not convert matters only among the first `need`
    rgb->assign(need, 0);
    size_t have = 0;
    for (const BodyChunk& c : chunks) {
        if (have >= need) break;
        if (c.err_at && have + c.err_token < need) {
            set_error("texture '" + path + "': " + c.err_msg);
            return RT_E_PARSE;
        }
        const size_t take = std::min(c.out.size(), need - have);
        if (take) std::memcpy(rgb->data() + have, c.out.data(), take);
        have += take;
        if (c.err_at) break;           // values after the bad token of this chunk were not collected
    }
    if (have < need) {
        set_error("texture '" + path + "': too few tokens");
        return RT_E_PARSE;
    }
    *width = (int32_t)w; *height = (int32_t)h;
    if (verbose)
        std::fprintf(stderr, "read_texture_p3: %zu chunks, %u threads: read %.1f ms, parse %.1f ms, gather %.1f ms\n",
                     chunks.size(), threads, t_read - t_begin, t_parsed - t_read, now() - t_parsed);
    return RT_OK;
}

namespace {

struct DigitTable {
    char text[256][4];
    uint8_t len[256];
    DigitTable()
    {
        for (int v = 0; v < 256; v++) len[v] = (uint8_t)std::snprintf(text[v], 4, "%d", v);
    }
};

// static_cast<int>(map(c, 0, 1, 0, 255)) widened to size_t (main.cpp:760-762)
unsigned long long quantise_like_reference(float c)
{
    float q = (c - 0.0f) * (255.0f - 0.0f) / (1.0f - 0.0f) + 0.0f;
    int v = (q != q || q >= 2147483648.0f || q < -2147483648.0f) ? INT32_MIN : (int)q;
    return (unsigned long long)(long long)v;
}

size_t format_ppm(int32_t width, int32_t height, const uint8_t* rgb, const uint8_t* flags, const float bkg[3],
                  char* out)
{
    static const DigitTable table;
    static const char kNaN[] = "18446744071562067968";   // (size_t)(int)NaN on x86-64
    char bkg_text[3][24]; size_t bkg_len[3] = { 0, 0, 0 };
    if (bkg)
        for (int c = 0; c < 3; c++) bkg_len[c] = (size_t)std::snprintf(bkg_text[c], 24, "%llu", quantise_like_reference(bkg[c]));
    char* o = out;
    o += std::sprintf(o, "P3 \n%d %d \n255 \n", width, height);   // main.cpp:628-630
    const size_t n = (size_t)width * (size_t)height;
    for (size_t k = 0; k < n; k++) {
        const uint8_t fl = flags ? flags[k] : 0;
        for (int c = 0; c < 3; c++) {
            if (fl & (1u << c)) {
                std::memcpy(o, kNaN, sizeof(kNaN) - 1); o += sizeof(kNaN) - 1;
            } else if ((fl & 8u) && bkg) {
                std::memcpy(o, bkg_text[c], bkg_len[c]); o += bkg_len[c];
            } else {
                const uint8_t v = rgb[3 * k + c];
                std::memcpy(o, table.text[v], 4); o += table.len[v];
            }
            *o++ = ' ';
        }
        *o++ = '\n';                                              // main.cpp:639-641
    }
    return (size_t)(o - out);
}

size_t ppm_capacity(int32_t width, int32_t height)
{
    return 64 + (size_t)width * (size_t)height * (3 * 21 + 1) + 8;
}

}  // namespace
}  // namespace rthost

using namespace rthost;

extern "C" int rt_host_read_texture_p3(const char* path, int32_t* width, int32_t* height, uint8_t** rgb)
{
    if (!path || !width || !height || !rgb) { set_error("rt_host_read_texture_p3: null argument"); return RT_E_INVALID; }
    std::vector<uint8_t> px;
    int rc = read_texture_p3(path, width, height, &px);
    if (rc != RT_OK) return rc;
    *rgb = (uint8_t*)std::malloc(px.size() ? px.size() : 1);
    if (!*rgb) { set_error("out of memory"); return RT_E_NOMEM; }
    std::memcpy(*rgb, px.data(), px.size());
    return RT_OK;
}

extern "C" void rt_host_free(void* p) { std::free(p); }

extern "C" int rt_host_format_ppm(int32_t width, int32_t height, const uint8_t* rgb, const uint8_t* flags,
                                  const float bkg[3], char** out_text, size_t* out_len)
{
    if (width < 1 || height < 1 || !rgb || !out_text || !out_len) { set_error("rt_host_format_ppm: bad argument"); return RT_E_INVALID; }
    // worst case only when flags are present; otherwise 4 chars per channel
    size_t cap = flags ? ppm_capacity(width, height) : 64 + (size_t)width * (size_t)height * 13 + 8;
    char* buf = (char*)std::malloc(cap);
    if (!buf) { set_error("out of memory"); return RT_E_NOMEM; }
    *out_len = format_ppm(width, height, rgb, flags, bkg, buf);
    *out_text = buf;
    return RT_OK;
}

extern "C" int rt_host_write_ppm(const char* path, int32_t width, int32_t height, const uint8_t* rgb,
                                 const uint8_t* flags, const float bkg[3])
{
    if (!path) { set_error("rt_host_write_ppm: null path"); return RT_E_INVALID; }
    char* text = nullptr; size_t len = 0;
    int rc = rt_host_format_ppm(width, height, rgb, flags, bkg, &text, &len);
    if (rc != RT_OK) return rc;
    FILE* f = std::fopen(path, "wb");
    if (!f) { std::free(text); set_error("ERROR: failed to create ppm image"); return RT_E_IO; }   // main.cpp:621
    size_t put = std::fwrite(text, 1, len, f);
    std::fclose(f);
    std::free(text);
    if (put != len) { set_error("short write"); return RT_E_IO; }
    return RT_OK;
}

extern "C" int rt_host_output_path(const char* scene_path, char* buf, size_t buf_len)
{
    if (!scene_path || !buf) return RT_E_INVALID;
    std::string p(scene_path);
    size_t dot = p.rfind('.');                                   // utility.h:34-41
    if (dot != std::string::npos) p.resize(dot);
    p += ".ppm";                                                 // main.cpp:616
    if (p.size() + 1 > buf_len) return RT_E_INVALID;
    std::memcpy(buf, p.c_str(), p.size() + 1);
    return RT_OK;
}
