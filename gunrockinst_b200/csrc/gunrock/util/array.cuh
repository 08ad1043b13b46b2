// array.cuh -- owning device (and pinned host) buffers.
//
// Replaces the reference's raw cudaMalloc/cudaFree pairs scattered through
// Problem::Init / Enactor::Enact (e.g. app/problem_base.cuh:279-303,
// app/bfs/bfs_enactor.cuh:254-258): every buffer is allocated once, outside
// any timed region, and grown geometrically if a later Reset needs more.
//
// Behind both sits a block cache: a released buffer is kept (per device, per size class) and
// handed to the next request of that class, so a caller that goes through the reference's
// one-shot entry points (gunrock_bfs_func & co.: upload, run, download, tear down -- every
// call) pays cudaMalloc / cudaFree / cudaHostAlloc once per size, not once per call.  On a
// scale-22 graph those calls were 9 of the 23 ms of gunrock_bfs_func.  GRB_POOL=0 turns the
// cache off; grb_release_cached_memory() empties it.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>
#include <vector>
#include "error.cuh"

namespace gunrock {
namespace util {

class BlockCache {
  public:
    static BlockCache &Device() { static BlockCache *c = new BlockCache(false); return *c; }   // never destroyed:
    static BlockCache &Pinned() { static BlockCache *c = new BlockCache(true);  return *c; }   // outlives the runtime

    // bytes -> size class: at most 12.5 % above the request, so few distinct classes exist
    static size_t Class(size_t bytes)
    {
        if (bytes <= 512) return 512;
        int top = 63;
        while (!((bytes >> top) & 1)) --top;
        const size_t step = (size_t)1 << (top - 3);
        return (bytes + step - 1) & ~(step - 1);
    }

    cudaError_t Take(size_t bytes, void **out, size_t *got)
    {
        const size_t cls = Class(bytes);
        int dev = 0;
        if (!pinned_) cudaGetDevice(&dev);
        if (enabled_) {
            std::lock_guard<std::mutex> lock(mu_);
            auto it = free_.find(std::make_pair(dev, cls));
            if (it != free_.end() && !it->second.empty()) {
                *out = it->second.back();
                it->second.pop_back();
                cached_ -= cls;
                held_[dev] -= cls;
                *got = cls;
                return cudaSuccess;
            }
        }
        cudaError_t e = Raw(out, cls);
        if (e == cudaErrorMemoryAllocation) {       // give the cache back to the driver and retry once
            cudaGetLastError();
            Trim();
            e = Raw(out, cls);
        }
        *got = cls;
        return e;
    }

    // Callers hand a block back only after the work that used it has completed (every Release() in this
    // library sits behind a stream / device synchronisation of its owner): blocks carry no stream or event,
    // so a block given back while a kernel still uses it would be a use-after-free for the next taker.
    void Give(void *p, size_t cls, int dev)
    {
        if (!p) return;
        if (enabled_) {
            std::lock_guard<std::mutex> lock(mu_);
            const int key = pinned_ ? 0 : dev;
            if (held_[key] + cls <= Limit(key)) {
                free_[std::make_pair(key, cls)].push_back(p);
                held_[key] += cls;
                cached_ += cls;
                return;
            }
        }
        Free(p, dev);
    }

    void Trim()
    {
        std::lock_guard<std::mutex> lock(mu_);
        for (auto &kv : free_)
            for (void *p : kv.second) Free(p, kv.first.first);
        free_.clear();
        held_.clear();
        cached_ = 0;
    }
    size_t CachedBytes() { std::lock_guard<std::mutex> lock(mu_); return cached_; }

  private:
    explicit BlockCache(bool pinned) : pinned_(pinned)
    {
        const char *e = getenv("GRB_POOL");
        enabled_ = !(e && e[0] == '0');
    }
    cudaError_t Raw(void **out, size_t bytes)
    {
        return pinned_ ? cudaHostAlloc(out, bytes, cudaHostAllocDefault) : cudaMalloc(out, bytes);
    }
    void Free(void *p, int dev)
    {
        if (pinned_) { cudaFreeHost(p); return; }
        int cur = 0;
        cudaGetDevice(&cur);
        if (cur != dev) cudaSetDevice(dev);
        cudaFree(p);
        if (cur != dev) cudaSetDevice(cur);
    }
    // most bytes the cache keeps per device: GRB_POOL_LIMIT_MB, else a quarter of THAT device's memory (other
    // allocators in the process -- torch, NCCL, the caller -- need the rest); 1 GB of pinned host memory
    size_t Limit(int dev)
    {
        auto it = limit_.find(dev);
        if (it != limit_.end()) return it->second;
        size_t lim;
        const char *e = getenv("GRB_POOL_LIMIT_MB");
        if (e && atoll(e) > 0) lim = (size_t)atoll(e) << 20;
        else if (pinned_) lim = (size_t)1 << 30;
        else {
            int cur = 0;
            cudaGetDevice(&cur);
            if (cur != dev) cudaSetDevice(dev);
            size_t free_b = 0, total_b = 0;
            lim = (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && total_b) ? total_b / 4 : ((size_t)8 << 30);
            if (cur != dev) cudaSetDevice(cur);
        }
        limit_[dev] = lim;
        return lim;
    }
    bool   pinned_, enabled_ = true;
    size_t cached_ = 0;
    std::map<int, size_t> limit_, held_;
    std::mutex mu_;
    std::map<std::pair<int, size_t>, std::vector<void *>> free_;
};

template <typename T>
struct DeviceArray {
    T     *ptr  = nullptr;
    size_t size = 0;   // elements in use
    size_t cap  = 0;   // elements allocated
    size_t block_ = 0; // bytes of the cached block behind ptr
    int    device_ = 0;
    static const size_t kPad = 16;

    DeviceArray() = default;
    DeviceArray(const DeviceArray &) = delete;
    DeviceArray &operator=(const DeviceArray &) = delete;
    ~DeviceArray() { Release(); }

    // make room for n elements (contents are NOT preserved when growing)
    cudaError_t Allocate(size_t n)
    {
        if (n > cap) {
            if (borrowed_) return cudaErrorInvalidValue;
            Release();
            // kPad bytes beyond the last element stay readable: kernels fetch aligned 16-byte groups
            size_t bytes = (n ? n : 1) * sizeof(T) + kPad, got = 0;
            cudaGetDevice(&device_);
            GR_TRY(BlockCache::Device().Take(bytes, (void **)&ptr, &got));
            block_ = got;
            cap = (got - kPad) / sizeof(T);
        }
        size = n;
        return cudaSuccess;
    }
    void Release()
    {
        if (ptr && !borrowed_) BlockCache::Device().Give(ptr, block_, device_);
        ptr = nullptr; size = cap = 0; block_ = 0; borrowed_ = false;
    }
    // point at n elements somebody else owns (a mapped host buffer): never freed, never grown
    void Borrow(T *device_visible, size_t n)
    {
        Release();
        ptr = device_visible; size = cap = n; borrowed_ = true;
    }
    bool borrowed_ = false;
    cudaError_t Zero(cudaStream_t s) { return size ? cudaMemsetAsync(ptr, 0, size * sizeof(T), s) : cudaSuccess; }
    cudaError_t FromHost(const T *h, size_t n, cudaStream_t s)
    {
        GR_TRY(Allocate(n));
        if (n) GR_TRY(cudaMemcpyAsync(ptr, h, n * sizeof(T), cudaMemcpyHostToDevice, s));
        return cudaSuccess;
    }
    cudaError_t ToHost(T *h, size_t n, cudaStream_t s) const
    {
        // (cudaMemcpyDefault: a borrowed array may alias mapped host memory)
        if (n) GR_TRY(cudaMemcpyAsync(h, ptr, n * sizeof(T), borrowed_ ? cudaMemcpyDefault : cudaMemcpyDeviceToHost, s));
        return cudaSuccess;
    }
};

// page-locked host buffer the device result words are copied into
template <typename T>
struct PinnedArray {
    T     *ptr  = nullptr;
    size_t size = 0;
    PinnedArray() = default;
    PinnedArray(const PinnedArray &) = delete;
    PinnedArray &operator=(const PinnedArray &) = delete;
    ~PinnedArray() { Release(); }
    cudaError_t Allocate(size_t n)
    {
        if (n > size) {
            Release();
            size_t got = 0;
            GR_TRY(BlockCache::Pinned().Take((n ? n : 1) * sizeof(T), (void **)&ptr, &got));
            block_ = got;
            size = n;
        }
        return cudaSuccess;
    }
    void Release() { if (ptr) BlockCache::Pinned().Give(ptr, block_, 0); ptr = nullptr; size = 0; block_ = 0; }
    size_t block_ = 0;
};

}  // namespace util
}  // namespace gunrock
