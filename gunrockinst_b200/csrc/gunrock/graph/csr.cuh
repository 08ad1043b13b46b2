// csr.cuh -- device-resident CSR graph slice.
//
// Counterpart of the reference's ProblemBase::GraphSlice device arrays
// (gunrock/app/problem_base.cuh:73-140) and of its 1D striping rule
// owner(v) = v % num_gpus, local row = v / num_gpus (:186-210), which the
// reference declares but never implements (:336-338 "TODO: multiple GPU").
//
// HBM layout (all arrays 256-byte aligned by cudaMalloc):
//   row_offsets  SizeT   [local_nodes + 1]   CSR offsets of the rows this device owns
//   col_indices  VertexId[edges]             GLOBAL neighbour ids, rows back to back
//   edge_values  uint32  [edges]             optional weights, same order as col_indices
// For an undirected (symmetric) graph the CSC is the CSR itself, so the pull
// direction walks the same arrays; a directed graph may attach a separate
// transpose through `csc_offsets / csc_indices`.
#pragma once
#include <cstdint>
#include "../util/array.cuh"

namespace gunrock {
namespace graph {

// 1D striped vertex partition over 2^log_parts devices.
// The reference's (unimplemented) rule is owner = v % P, local row = v / P
// (problem_base.cuh:186-210).  The local-row rule is kept; the owner is the low
// bits XOR a multiplicative hash of the local row.  Reason: R-MAT with the
// reference's parameters and no vertex relabelling puts probability a+b = 0.75 on
// every id bit being 0, so plain v % 8 would hand device 0 about 42 % of all
// edges; XOR-ing with a well-mixed function of the other bits evens that out
// while staying a bijection (v <-> (owner, local)) computed in two instructions.
struct Partition {
    int log_parts = 0;      // number of devices = 1 << log_parts
    int rank      = 0;      // this device

    __host__ __device__ __forceinline__ int  Parts() const { return 1 << log_parts; }
    __host__ __device__ __forceinline__ unsigned Mix(long long local) const
    {
        return log_parts ? (((unsigned)local * 0x9E3779B1u) >> (32 - log_parts)) : 0u;
    }
    // (log_parts is warp-uniform: the single-device tests below cost one uniform branch and
    //  spare the per-edge hash arithmetic on the path every headline number is measured on)
    __host__ __device__ __forceinline__ int  Owner(long long v) const
    {
        if (log_parts == 0) return 0;
        return (int)(((unsigned)v & (unsigned)(Parts() - 1)) ^ Mix(v >> log_parts));
    }
    __host__ __device__ __forceinline__ bool Mine(long long v) const { return log_parts == 0 || Owner(v) == rank; }
    __host__ __device__ __forceinline__ long long Local(long long v) const { return v >> log_parts; }
    __host__ __device__ __forceinline__ long long Global(long long local) const
    {
        if (log_parts == 0) return local;
        return (local << log_parts) | (long long)((unsigned)rank ^ Mix(local));
    }
};

// non-owning view handed to kernels (passed by value in the parameter bank)
template <typename VertexId, typename SizeT>
struct CsrView {
    const SizeT    *row_offsets;
    const VertexId *col_indices;
    const unsigned *edge_values;
    long long       nodes;          // GLOBAL vertex count
    long long       local_nodes;    // rows held by this device
    long long       edges;          // entries held by this device
    Partition       part;
};

template <typename _VertexId, typename _SizeT>
struct DeviceCsr {
    typedef _VertexId VertexId;
    typedef _SizeT    SizeT;

    util::DeviceArray<SizeT>    row_offsets;
    util::DeviceArray<VertexId> col_indices;
    util::DeviceArray<unsigned> edge_values;
    // optional transpose for directed graphs (pull direction); empty => symmetric
    util::DeviceArray<SizeT>    csc_offsets;
    util::DeviceArray<VertexId> csc_indices;

    long long nodes = 0, local_nodes = 0, edges = 0;
    bool      symmetric = false;     // CSR doubles as CSC
    Partition part;
    int       device = 0;

    CsrView<VertexId, SizeT> View() const
    {
        CsrView<VertexId, SizeT> v;
        v.row_offsets = row_offsets.ptr;
        v.col_indices = col_indices.ptr;
        v.edge_values = edge_values.size ? edge_values.ptr : nullptr;
        v.nodes = nodes; v.local_nodes = local_nodes; v.edges = edges; v.part = part;
        return v;
    }
    // the arrays the pull direction walks
    CsrView<VertexId, SizeT> InverseView() const
    {
        CsrView<VertexId, SizeT> v = View();
        if (!symmetric && csc_offsets.size) {
            v.row_offsets = csc_offsets.ptr;
            v.col_indices = csc_indices.ptr;
            v.edge_values = nullptr;
        }
        return v;
    }
    bool HasInverse() const { return symmetric || csc_offsets.size != 0; }

    // upload a whole (unpartitioned) host CSR.  stream_from_host: only the offsets are uploaded, the
    // column ids (and weights) stay in the caller's memory, page-locked and mapped, and the kernels
    // read them over PCIe as they go -- the reference's stream_from_host (gunrock/app/
    // problem_base.cuh:252-276: cudaHostAlloc mapped + cudaHostGetDevicePointer) for graphs that do
    // not fit the device or are searched once.  The caller keeps the arrays alive and unchanged.
    template <typename HostSizeT>
    cudaError_t FromHost(long long n, long long m, const HostSizeT *h_offsets,
                         const VertexId *h_cols, const unsigned *h_weights, cudaStream_t s,
                         bool stream_from_host = false);
    bool  streams_from_host = false;
    void *registered_[2] = {nullptr, nullptr};      // host ranges this object page-locked
    DeviceCsr() = default;
    DeviceCsr(const DeviceCsr &) = delete;
    DeviceCsr &operator=(const DeviceCsr &) = delete;
    ~DeviceCsr()
    {
        col_indices.Release(); edge_values.Release();
        for (void *r : registered_) if (r) cudaHostUnregister(r);
    }
    // device-visible alias of a host buffer; page-locks it when the caller has not
    cudaError_t MapHost(const void *h, size_t bytes, void **dev, int slot)
    {
        if (cudaHostGetDevicePointer(dev, const_cast<void *>(h), 0) == cudaSuccess) return cudaSuccess;
        (void)cudaGetLastError();                   // not page-locked yet
        GR_TRY(cudaHostRegister(const_cast<void *>(h), bytes, cudaHostRegisterMapped));
        registered_[slot] = const_cast<void *>(h);
        return cudaHostGetDevicePointer(dev, const_cast<void *>(h), 0);
    }
};

template <typename VertexId, typename SizeT>
template <typename HostSizeT>
cudaError_t DeviceCsr<VertexId, SizeT>::FromHost(long long n, long long m, const HostSizeT *h_offsets,
                                                 const VertexId *h_cols, const unsigned *h_weights,
                                                 cudaStream_t s, bool stream_from_host)
{
    nodes = local_nodes = n; edges = m; part = Partition();
    if (sizeof(HostSizeT) == sizeof(SizeT)) {
        GR_TRY(row_offsets.FromHost((const SizeT *)h_offsets, (size_t)n + 1, s));
    } else {
        // widen / narrow on the host, then upload
        SizeT *tmp = (SizeT *)malloc(sizeof(SizeT) * ((size_t)n + 1));
        for (long long i = 0; i <= n; ++i) tmp[i] = (SizeT)h_offsets[i];
        cudaError_t e = row_offsets.FromHost(tmp, (size_t)n + 1, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        free(tmp);
        GR_TRY(e);
    }
    streams_from_host = stream_from_host && m > 0;
    if (streams_from_host) {
        void *d = nullptr;
        GR_TRY(MapHost(h_cols, sizeof(VertexId) * (size_t)m, &d, 0));
        col_indices.Borrow((VertexId *)d, (size_t)m);
        if (h_weights) {
            GR_TRY(MapHost(h_weights, sizeof(unsigned) * (size_t)m, &d, 1));
            edge_values.Borrow((unsigned *)d, (size_t)m);
        } else edge_values.size = 0;
        return cudaSuccess;
    }
    GR_TRY(col_indices.FromHost(h_cols, (size_t)m, s));
    if (h_weights) GR_TRY(edge_values.FromHost(h_weights, (size_t)m, s));
    else edge_values.size = 0;
    return cudaSuccess;
}

}  // namespace graph
}  // namespace gunrock
