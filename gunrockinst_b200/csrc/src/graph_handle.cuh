// graph_handle.cuh -- the object behind grb_graph_t: a resident CSR with either
// 32-bit offsets (what gunrock.h callers pass, SIZET_INT) or 64-bit offsets
// (graphs past 2^31 entries, i.e. R-MAT scale 27 on one or two devices).
#pragma once
#include <cstdint>
#include "../gunrock/graph/build.cuh"

typedef gunrock::graph::DeviceCsr<int, int>       Csr32;
typedef gunrock::graph::DeviceCsr<int, long long> Csr64;

struct grb_graph_s {
    int    offset_bytes = 4;
    Csr32 *g32 = nullptr;
    Csr64 *g64 = nullptr;
    cudaStream_t stream = 0;
    // degree-ordered twin (graph/build.cuh), built on first use by a BFS context that asks for it
    Csr32 *twin32 = nullptr;
    Csr64 *twin64 = nullptr;
    gunrock::util::DeviceArray<int> to_new, to_old;     // caller's id -> twin id and back
    ~grb_graph_s()
    {
        delete twin32; delete twin64;
        delete g32; delete g64;
        if (stream) cudaStreamDestroy(stream);
    }
    bool has_twin() const { return twin32 || twin64; }
    cudaError_t EnsureTwin()
    {
        if (has_twin()) {       // a transpose attached to the graph after the twin was built
            if (twin32 && !twin32->HasInverse() && g32->HasInverse()) return gunrock::graph::BuildTranspose(*twin32, stream);
            if (twin64 && !twin64->HasInverse() && g64->HasInverse()) return gunrock::graph::BuildTranspose(*twin64, stream);
            return cudaSuccess;
        }
        cudaError_t e;
        if (g32) { twin32 = new Csr32(); e = gunrock::graph::BuildDegreeOrdered(*g32, *twin32, to_new, to_old, stream); }
        else     { twin64 = new Csr64(); e = gunrock::graph::BuildDegreeOrdered(*g64, *twin64, to_new, to_old, stream); }
        if (e != cudaSuccess) { delete twin32; delete twin64; twin32 = nullptr; twin64 = nullptr; }
        return e;
    }
    // transpose of a directed graph (the pull direction's inverse graph)
    cudaError_t EnsureTranspose()
    {
        if (g32) return (g32->symmetric || g32->csc_offsets.size) ? cudaSuccess : gunrock::graph::BuildTranspose(*g32, stream);
        return (g64->symmetric || g64->csc_offsets.size) ? cudaSuccess : gunrock::graph::BuildTranspose(*g64, stream);
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
