// ppm_device.cu — the P3 text of a frame, formatted on the device.
//
// The reference writes its image as ASCII (main.cpp:628-641): "P3 \n", "W H \n", "255 \n",
// then one "r g b \n" line per pixel, every value through std::to_string and every line
// flushed with std::endl — at 4K that is ~97 MB of text and seconds of wall time, i.e. the
// whole run time once the render takes milliseconds.  The frame is already on the GPU as
// RGBA8 (A = flags), so the text is produced there: per-pixel line lengths, a two-level
// exclusive scan, then every 256-pixel block composes its lines in shared memory and stores
// them coalesced.  The host only copies the finished bytes (pinned, D2H) and writes the file.
// Byte-identical with the CPU formatter rt_host_format_ppm (tests/test_gpu_parity.py).
#include "host_scene.h"

#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

namespace rt {

namespace {

constexpr int kPpmBlock = 256;
constexpr int kTokMax = 20;            // "18446744071562067968"
constexpr int kLineMax = 3 * (kTokMax + 1) + 1;

struct PpmTokens {
    char nan[kTokMax];                 // (size_t)(int)NaN on x86-64
    char bkg[3][kTokMax];              // (size_t)(int)(bkg[c] * 255.0f): a primary miss prints the un-clamped value
    int bkg_len[3];
};

__device__ __forceinline__ int token_len(unsigned v, unsigned flags, int c, const PpmTokens& T)
{
    if (flags & (1u << c)) return kTokMax;
    if (flags & 8u) return T.bkg_len[c];
    return v >= 100u ? 3 : v >= 10u ? 2 : 1;
}

__device__ __forceinline__ int line_len(uchar4 p, const PpmTokens& T)
{
    return token_len(p.x, p.w, 0, T) + token_len(p.y, p.w, 1, T) + token_len(p.z, p.w, 2, T) + 4;   // 3 blanks + '\n'
}

__device__ __forceinline__ char* put_token(char* o, unsigned v, unsigned flags, int c, const PpmTokens& T)
{
    if (flags & (1u << c)) {
        for (int i = 0; i < kTokMax; i++) o[i] = T.nan[i];
        o += kTokMax;
    } else if (flags & 8u) {
        for (int i = 0; i < T.bkg_len[c]; i++) o[i] = T.bkg[c][i];
        o += T.bkg_len[c];
    } else {
        if (v >= 100u) *o++ = (char)('0' + v / 100u);
        if (v >= 10u) *o++ = (char)('0' + (v / 10u) % 10u);
        *o++ = (char)('0' + v % 10u);
    }
    *o++ = ' ';
    return o;
}

// exclusive scan of one value per thread over the block; total in *sum
__device__ __forceinline__ unsigned block_exclusive_scan(unsigned v, unsigned* sum)
{
    __shared__ unsigned warp_tot[kPpmBlock / 32];
    const unsigned lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    unsigned inc = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31u) warp_tot[w] = inc;
    __syncthreads();
    unsigned base = 0, total = 0;
    #pragma unroll
    for (int k = 0; k < kPpmBlock / 32; k++) {
        if ((unsigned)k < w) base += warp_tot[k];
        total += warp_tot[k];
    }
    __syncthreads();
    *sum = total;
    return base + inc - v;
}

__global__ void __launch_bounds__(kPpmBlock) ppm_len_kernel(const uchar4* __restrict__ px, unsigned n, PpmTokens T,
                                                           unsigned* __restrict__ block_bytes)
{
    const unsigned k = blockIdx.x * kPpmBlock + threadIdx.x;
    const unsigned len = k < n ? (unsigned)line_len(px[k], T) : 0u;
    unsigned total;
    block_exclusive_scan(len, &total);
    if (threadIdx.x == 0) block_bytes[blockIdx.x] = total;
}

// one block: exclusive scan of the per-block byte counts (64-bit: an 8K frame of NaN pixels exceeds 2^32)
__global__ void __launch_bounds__(1024) ppm_scan_kernel(const unsigned* __restrict__ block_bytes, unsigned n_blocks,
                                                       unsigned long long* __restrict__ block_offset,
                                                       unsigned long long* __restrict__ total_bytes)
{
    __shared__ unsigned long long part[1024];
    const unsigned per = (n_blocks + 1023u) / 1024u;
    const unsigned b0 = threadIdx.x * per, b1 = min(b0 + per, n_blocks);
    unsigned long long s = 0;
    for (unsigned b = b0; b < b1; b++) s += block_bytes[b];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; i++) { const unsigned long long t = part[i]; part[i] = run; run += t; }
        *total_bytes = run;
    }
    __syncthreads();
    unsigned long long run = part[threadIdx.x];
    for (unsigned b = b0; b < b1; b++) { block_offset[b] = run; run += block_bytes[b]; }
}

__global__ void __launch_bounds__(kPpmBlock) ppm_write_kernel(const uchar4* __restrict__ px, unsigned n, PpmTokens T,
                                                             const unsigned long long* __restrict__ block_offset,
                                                             char* __restrict__ out)
{
    __shared__ char text[kPpmBlock * kLineMax];
    const unsigned k = blockIdx.x * kPpmBlock + threadIdx.x;
    uchar4 p = make_uchar4(0, 0, 0, 0);
    unsigned len = 0;
    if (k < n) { p = px[k]; len = (unsigned)line_len(p, T); }
    unsigned total;
    const unsigned at = block_exclusive_scan(len, &total);
    if (k < n) {
        char* o = text + at;
        o = put_token(o, p.x, p.w, 0, T);
        o = put_token(o, p.y, p.w, 1, T);
        o = put_token(o, p.z, p.w, 2, T);
        *o = '\n';
    }
    __syncthreads();
    // coalesced copy-out: bytes up to the first 16-byte boundary of the destination, then 16-byte words
    char* dst = out + block_offset[blockIdx.x];
    const unsigned head = min(total, (unsigned)((16u - (unsigned)((size_t)dst & 15u)) & 15u));
    for (unsigned i = threadIdx.x; i < head; i += kPpmBlock) dst[i] = text[i];
    const unsigned words = (total - head) / 16u;
    for (unsigned i = threadIdx.x; i < words; i += kPpmBlock) {
        const char* s = text + head + i * 16u;     // shared side may be misaligned: assemble from bytes
        uint4 w;
        unsigned* wp = reinterpret_cast<unsigned*>(&w);
        #pragma unroll
        for (int q = 0; q < 4; q++)
            wp[q] = (unsigned)(unsigned char)s[4 * q] | ((unsigned)(unsigned char)s[4 * q + 1] << 8) |
                    ((unsigned)(unsigned char)s[4 * q + 2] << 16) | ((unsigned)(unsigned char)s[4 * q + 3] << 24);
        reinterpret_cast<uint4*>(dst + head)[i] = w;
    }
    for (unsigned i = head + words * 16u + threadIdx.x; i < total; i += kPpmBlock) dst[i] = text[i];
}

// static_cast<int>(map(c, 0, 1, 0, 255)) widened to size_t (main.cpp:760-762), as in ppm_io.cpp
unsigned long long quantise_like_reference(float c)
{
    const float q = (c - 0.0f) * (255.0f - 0.0f) / (1.0f - 0.0f) + 0.0f;
    const int v = (q != q || q >= 2147483648.0f || q < -2147483648.0f) ? INT32_MIN : (int)q;
    return (unsigned long long)(long long)v;
}

}  // namespace

// Formats `px` (width*height RGBA8, scanline order, A = flags) into s->ws.ppm_host.
int format_ppm_device(SceneHost* s, const uchar4* px, const char** text, size_t* len, cudaStream_t st)
{
    Workspace& ws = s->ws;
    const size_t n = (size_t)s->width * (size_t)s->height;
    if (n >= 0xffffffffull) { set_error("frame too large for the device formatter"); return RT_E_INVALID; }
    const unsigned n_blocks = (unsigned)((n + kPpmBlock - 1) / kPpmBlock);
    PpmTokens T;
    std::memset(&T, 0, sizeof T);
    std::memcpy(T.nan, "18446744071562067968", kTokMax);
    for (int c = 0; c < 3; c++) {
        char tmp[32];
        T.bkg_len[c] = std::snprintf(tmp, sizeof tmp, "%llu", quantise_like_reference(s->bkg[c]));
        std::memcpy(T.bkg[c], tmp, (size_t)T.bkg_len[c]);
    }
    char header[64];
    const int hl = std::snprintf(header, sizeof header, "P3 \n%d %d \n255 \n", s->width, s->height);   // main.cpp:628-630

    if (!ws.ppm_blocks.reserve((size_t)n_blocks * 4) || !ws.ppm_offsets.reserve(((size_t)n_blocks + 1) * 8) ||
        !ws.ppm_total_host.reserve(8))
        return RT_E_NOMEM;
    unsigned long long* d_total = ws.ppm_offsets.as<unsigned long long>() + n_blocks;
    ppm_len_kernel<<<n_blocks, kPpmBlock, 0, st>>>(px, (unsigned)n, T, ws.ppm_blocks.as<unsigned>());
    ppm_scan_kernel<<<1, 1024, 0, st>>>(ws.ppm_blocks.as<unsigned>(), n_blocks, ws.ppm_offsets.as<unsigned long long>(), d_total);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(ws.ppm_total_host.p, d_total, 8, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    const size_t body = (size_t)*static_cast<unsigned long long*>(ws.ppm_total_host.p);
    if (!ws.ppm_text.reserve(body + 16) || !ws.ppm_host.reserve((size_t)hl + body + 1)) return RT_E_NOMEM;
    ppm_write_kernel<<<n_blocks, kPpmBlock, 0, st>>>(px, (unsigned)n, T, ws.ppm_offsets.as<unsigned long long>(),
                                                    ws.ppm_text.as<char>());
    RT_CUDA(cudaGetLastError());
    char* host = static_cast<char*>(ws.ppm_host.p);
    std::memcpy(host, header, (size_t)hl);
    RT_CUDA(cudaMemcpyAsync(host + hl, ws.ppm_text.p, body, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    host[hl + body] = '\0';
    *text = host;
    *len = (size_t)hl + body;
    return RT_OK;
}

}  // namespace rt

// ---- rt_format_ppm_noalloc: the same text without a single allocation ---------------------------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// The formatter's scratch — one byte count and one offset per 256-pixel block — lives in a
// file-scope __device__ array sized for the largest frame BASELINE.json names (7680 x 4320);
// every GPU that loads the library holds these 1.56 MB until the process exits.  All calls on a
// device share the array: calls that overlap on the GPU (different streams) race.  Nothing in it
// has to start at zero — the length kernel rewrites every count the later kernels read, so a
// call never sees what the last call left.
namespace rt {
namespace {

constexpr unsigned kNoAllocMaxPixels = 7680u * 4320u;
constexpr unsigned kNoAllocMaxBlocks = (kNoAllocMaxPixels + kPpmBlock - 1) / kPpmBlock;

struct PpmScratch {
    unsigned long long offset[kNoAllocMaxBlocks + 1];   // [n_blocks] = bytes of all pixel lines
    unsigned bytes[kNoAllocMaxBlocks];
};
static __device__ __align__(256) PpmScratch g_ppm_scratch;

// In host code `g_ppm_scratch` is a stand-in, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
std::mutex g_scratch_mutex;
PpmScratch* g_scratch_addr[64] = { nullptr };

int scratch_address(PpmScratch** out)
{
    int dev = -1;
    RT_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("rt_format_ppm_noalloc: device index out of range"); return RT_E_CUDA; }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        RT_CUDA(cudaGetSymbolAddress(&p, g_ppm_scratch));
        g_scratch_addr[dev] = static_cast<PpmScratch*>(p);
    }
    *out = g_scratch_addr[dev];
    return RT_OK;
}

struct PpmHeader {
    char text[64];
    int len;
};

// Variant of ppm_write_kernel: the body starts behind the header (which block 0 writes), nothing is
// stored past `capacity`, and the file's length goes to *len_out (device memory).
__global__ void __launch_bounds__(kPpmBlock) ppm_write_noalloc_kernel(const uchar4* __restrict__ px, unsigned n, PpmTokens T,
                                                                     PpmHeader H, const PpmScratch* __restrict__ scratch,
                                                                     unsigned n_blocks, char* __restrict__ out,
                                                                     unsigned long long capacity,
                                                                     unsigned long long* __restrict__ len_out)
{
    __shared__ char text[kPpmBlock * kLineMax];
    const unsigned k = blockIdx.x * kPpmBlock + threadIdx.x;
    uchar4 p = make_uchar4(0, 0, 0, 0);
    unsigned len = 0;
    if (k < n) { p = px[k]; len = (unsigned)line_len(p, T); }
    unsigned total;
    const unsigned at = block_exclusive_scan(len, &total);
    if (k < n) {
        char* o = text + at;
        o = put_token(o, p.x, p.w, 0, T);
        o = put_token(o, p.y, p.w, 1, T);
        o = put_token(o, p.z, p.w, 2, T);
        *o = '\n';
    }
    __syncthreads();
    const unsigned long long begin = (unsigned long long)H.len + scratch->offset[blockIdx.x];
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) *len_out = (unsigned long long)H.len + scratch->offset[n_blocks];
        if ((unsigned long long)H.len <= capacity)
            for (unsigned i = threadIdx.x; i < (unsigned)H.len; i += kPpmBlock) out[i] = H.text[i];
    }
    if (begin + total > capacity) return;          // the caller sees *len_out > capacity: the text is incomplete
    char* dst = out + begin;
    for (unsigned i = threadIdx.x; i < total; i += kPpmBlock) dst[i] = text[i];
}

}  // namespace
}  // namespace rt

extern "C" int rt_format_ppm_noalloc(const void* dev_rgba, int32_t width, int32_t height, const float bkg[3],
                                     char* dev_text, size_t text_capacity, unsigned long long* dev_len,
                                     void* cuda_stream)
{
    using namespace rt;
    if (!dev_rgba || !bkg || !dev_text || !dev_len || width < 1 || height < 1) {
        set_error("rt_format_ppm_noalloc: bad argument");
        return RT_E_INVALID;
    }
    const unsigned long long n = (unsigned long long)width * (unsigned long long)height;
    if (n > kNoAllocMaxPixels) {
        set_error("rt_format_ppm_noalloc: " + std::to_string(n) + " pixels exceed the built-in scratch (33177600 pixels = 7680 x 4320); use rt_format_ppm_device");
        return RT_E_LIMIT;
    }
    PpmScratch* scratch = nullptr;
    const int rc = scratch_address(&scratch);
    if (rc != RT_OK) return rc;
    const unsigned n_blocks = (unsigned)((n + kPpmBlock - 1) / kPpmBlock);
    PpmTokens T;
    std::memset(&T, 0, sizeof T);
    std::memcpy(T.nan, "18446744071562067968", kTokMax);
    for (int c = 0; c < 3; c++) {
        char tmp[32];
        T.bkg_len[c] = std::snprintf(tmp, sizeof tmp, "%llu", quantise_like_reference(bkg[c]));
        std::memcpy(T.bkg[c], tmp, (size_t)T.bkg_len[c]);
    }
    PpmHeader H;
    std::memset(&H, 0, sizeof H);
    H.len = std::snprintf(H.text, sizeof H.text, "P3 \n%d %d \n255 \n", width, height);   // main.cpp:628-630
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const uchar4* px = static_cast<const uchar4*>(dev_rgba);
    ppm_len_kernel<<<n_blocks, kPpmBlock, 0, st>>>(px, (unsigned)n, T, scratch->bytes);
    ppm_scan_kernel<<<1, 1024, 0, st>>>(scratch->bytes, n_blocks, scratch->offset, scratch->offset + n_blocks);
    ppm_write_noalloc_kernel<<<n_blocks, kPpmBlock, 0, st>>>(px, (unsigned)n, T, H, scratch, n_blocks, dev_text,
                                                            (unsigned long long)text_capacity, dev_len);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}
