// array.cuh -- owning device (and pinned host) buffers.
//
// Replaces the reference's raw cudaMalloc/cudaFree pairs scattered through
// Problem::Init / Enactor::Enact (e.g. app/problem_base.cuh:279-303,
// app/bfs/bfs_enactor.cuh:254-258): every buffer is allocated once, outside
// any timed region, and grown geometrically if a later Reset needs more.
#pragma once
#include <cstddef>
#include <cstdint>
#include "error.cuh"

namespace gunrock {
namespace util {

template <typename T>
struct DeviceArray {
    T     *ptr  = nullptr;
    size_t size = 0;   // elements in use
    size_t cap  = 0;   // elements allocated

    DeviceArray() = default;
    DeviceArray(const DeviceArray &) = delete;
    DeviceArray &operator=(const DeviceArray &) = delete;
    ~DeviceArray() { Release(); }

    // make room for n elements (contents are NOT preserved when growing)
    cudaError_t Allocate(size_t n)
    {
        if (n > cap) {
            Release();
            size_t bytes = (n ? n : 1) * sizeof(T);
            GR_TRY(cudaMalloc((void **)&ptr, bytes));
            cap = n ? n : 1;
        }
        size = n;
        return cudaSuccess;
    }
    void Release()
    {
        if (ptr) cudaFree(ptr);
        ptr = nullptr; size = cap = 0;
    }
    cudaError_t Zero(cudaStream_t s) { return size ? cudaMemsetAsync(ptr, 0, size * sizeof(T), s) : cudaSuccess; }
    cudaError_t FromHost(const T *h, size_t n, cudaStream_t s)
    {
        GR_TRY(Allocate(n));
        if (n) GR_TRY(cudaMemcpyAsync(ptr, h, n * sizeof(T), cudaMemcpyHostToDevice, s));
        return cudaSuccess;
    }
    cudaError_t ToHost(T *h, size_t n, cudaStream_t s) const
    {
        if (n) GR_TRY(cudaMemcpyAsync(h, ptr, n * sizeof(T), cudaMemcpyDeviceToHost, s));
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
            GR_TRY(cudaHostAlloc((void **)&ptr, (n ? n : 1) * sizeof(T), cudaHostAllocDefault));
            size = n;
        }
        return cudaSuccess;
    }
    void Release() { if (ptr) cudaFreeHost(ptr); ptr = nullptr; size = 0; }
};

}  // namespace util
}  // namespace gunrock
