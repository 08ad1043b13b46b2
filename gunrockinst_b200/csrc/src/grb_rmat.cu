// grb_rmat.cu -- grb_graph_rmat: seeded R-MAT graphs built entirely on the device
// (include/gunrock_b200.h; generator and canonicalisation in gunrock/graphio/rmat.cuh).
#include <cstdio>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/graphio/rmat.cuh"

extern "C" int grb_graph_rmat(int device, int scale, int64_t edges, uint64_t seed, int undirected,
                              double a, double b, double c, double d,
                              int log_parts, int rank, uint64_t weight_seed,
                              int force_offset_bytes, grb_graph_t *out)
{
    if (!out || scale < 1 || scale > 30 || edges < 0 || log_parts < 0 || log_parts > 4 || log_parts >= scale ||
        rank < 0 || rank >= (1 << log_parts))
        return (int)cudaErrorInvalidValue;
    *out = nullptr;
    int rc = grb_select_device(device);
    if (rc) return rc;
    // upper bound on the entries this device can hold decides the offset width
    const unsigned long long bound = (undirected ? 2ull : 1ull) * (unsigned long long)edges;
    // (a device's share of a hashed partition stays well within 1.5x the mean)
    const unsigned long long share = log_parts ? (bound >> log_parts) + (bound >> (log_parts + 1)) : bound;
    int ob = force_offset_bytes ? force_offset_bytes : (share >= 0x7fffffffull ? 8 : 4);
    if (ob != 4 && ob != 8) return (int)cudaErrorInvalidValue;
    grb_graph_s *h = new grb_graph_s();
    h->offset_bytes = ob;
    gunrock::graph::Partition part;
    part.log_parts = log_parts; part.rank = rank;
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) {
        if (ob == 4) {
            h->g32 = new Csr32(); h->g32->device = device;
            e = gunrock::graphio::BuildRmatCsr<int>(scale, (unsigned long long)edges, seed, undirected != 0,
                                                    a, b, c, d, part, weight_seed, *h->g32, nullptr, h->stream);
        } else {
            h->g64 = new Csr64(); h->g64->device = device;
            e = gunrock::graphio::BuildRmatCsr<long long>(scale, (unsigned long long)edges, seed, undirected != 0,
                                                          a, b, c, d, part, weight_seed, *h->g64, nullptr, h->stream);
        }
    }
    if (e != cudaSuccess) { gunrock::util::GRError(e, "grb_graph_rmat", __FILE__, __LINE__); delete h; return (int)e; }
    *out = h;
    return 0;
}
