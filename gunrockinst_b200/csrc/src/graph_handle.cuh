// graph_handle.cuh -- the object behind grb_graph_t: a resident CSR with either
// 32-bit offsets (what gunrock.h callers pass, SIZET_INT) or 64-bit offsets
// (graphs past 2^31 entries, i.e. R-MAT scale 27 on one or two devices).
#pragma once
#include <cstdint>
#include "../gunrock/graph/csr.cuh"

typedef gunrock::graph::DeviceCsr<int, int>       Csr32;
typedef gunrock::graph::DeviceCsr<int, long long> Csr64;

struct grb_graph_s {
    int    offset_bytes = 4;
    Csr32 *g32 = nullptr;
    Csr64 *g64 = nullptr;
    cudaStream_t stream = 0;
    ~grb_graph_s()
    {
        delete g32; delete g64;
        if (stream) cudaStreamDestroy(stream);
    }
    long long nodes() const { return g32 ? g32->nodes : g64->nodes; }
    long long local_nodes() const { return g32 ? g32->local_nodes : g64->local_nodes; }
    long long edges() const { return g32 ? g32->edges : g64->edges; }
    int device() const { return g32 ? g32->device : g64->device; }
};

// every extern "C" entry point funnels errors through this
#define GRB_TRY(call)                                                                   \
    do {                                                                                \
        cudaError_t grb_e_ = ::gunrock::util::GRError((call), #call, __FILE__, __LINE__); \
        if (grb_e_ != cudaSuccess) return (int)grb_e_;                                  \
    } while (0)

// select the device or report why the library cannot run (no CPU fallback)
inline int grb_select_device(int device)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        fprintf(stderr, "libgunrock (B200 build): no CUDA device available (%s); "
                        "this library has no CPU path\n", cudaGetErrorString(e));
        return (int)(e != cudaSuccess ? e : cudaErrorNoDevice);
    }
    if (device < 0 || device >= count) {
        fprintf(stderr, "libgunrock (B200 build): device %d out of range (%d present)\n", device, count);
        return (int)cudaErrorInvalidDevice;
    }
    GRB_TRY(cudaSetDevice(device));
    return 0;
}
