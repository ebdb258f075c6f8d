// devmem.cu — process-wide cache of device allocations.
//
// rt_upload_scene / rt_free_scene allocate and free ~40 device buffers per scene (tables,
// staging, LBVH build temporaries).  cudaMalloc / cudaFree of 10..100 MB buffers cost
// 0.1..several ms each and cudaFree synchronises the device, which made the end-to-end step
// (upload + render + download per frame) spend more host time in the allocator than in the
// LBVH build.  Freed blocks are kept here and handed out again (best fit, at most 25 % larger
// than asked); rt_release_cached_memory() returns them to the driver.
//
// Stream safety: a DevBuf is only released when the work that used it has been synchronised
// (build_scene, build_bvh and the probes synchronise their stream before their temporaries
// die; rt_free_scene synchronises every stream of the scene), so a cached block is idle.
#include "host_scene.h"

#include <map>
#include <mutex>

namespace rt {

namespace {

struct Cache {
    std::mutex mu;
    std::multimap<size_t, void*> free_blocks[16];   // per device: capacity -> block
    size_t cached_bytes = 0;
};

Cache& cache()
{
    static Cache* c = new Cache();   // leaked on purpose: no cudaFree after the runtime is gone
    return *c;
}

size_t round_up(size_t n)
{
    const size_t q = n >= (1u << 20) ? (size_t)1 << 20 : 512;
    return (n + q - 1) / q * q;
}

}  // namespace

void* dev_alloc(size_t n, size_t* capacity)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) dev = 0;
    const size_t want = round_up(n ? n : 16);
    Cache& c = cache();
    {
        std::lock_guard<std::mutex> lock(c.mu);
        auto it = c.free_blocks[dev].lower_bound(want);
        if (it != c.free_blocks[dev].end() && it->first <= want + want / 4) {
            void* p = it->second;
            *capacity = it->first;
            c.cached_bytes -= it->first;
            c.free_blocks[dev].erase(it);
            return p;
        }
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        dev_release_cached();          // give the driver its memory back and try once more
        e = cudaMalloc(&p, want);
    }
    if (!cuda_ok(e, "cudaMalloc")) return nullptr;
    *capacity = want;
    return p;
}

void dev_free(void* p, size_t capacity)
{
    if (!p) return;
    cudaPointerAttributes attr;
    int dev = 0;
    if (cudaPointerGetAttributes(&attr, p) == cudaSuccess) dev = attr.device;
    else cudaGetLastError();
    if (dev < 0 || dev >= 16) { cudaFree(p); return; }
    Cache& c = cache();
    std::lock_guard<std::mutex> lock(c.mu);
    c.free_blocks[dev].emplace(capacity, p);
    c.cached_bytes += capacity;
}

size_t dev_release_cached()
{
    Cache& c = cache();
    std::lock_guard<std::mutex> lock(c.mu);
    int cur = 0;
    cudaGetDevice(&cur);
    size_t freed = 0;
    for (int d = 0; d < 16; d++) {
        if (c.free_blocks[d].empty()) continue;
        cudaSetDevice(d);
        for (auto& kv : c.free_blocks[d]) { cudaFree(kv.second); freed += kv.first; }
        c.free_blocks[d].clear();
    }
    cudaSetDevice(cur);
    c.cached_bytes = 0;
    return freed;
}

}  // namespace rt
