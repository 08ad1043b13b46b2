// rmat.cuh -- seeded R-MAT edge generator and COO -> CSR canonicalisation, on the device.
//
// Counterpart of the reference's host-side graphio::BuildRmatGraph
// (gunrock/graphio/rmat.cuh:27-92), ChoosePartition / VaryParams
// (gunrock/graphio/utils.cuh:47-130) and Csr::FromCoo (gunrock/csr.cuh:247-340):
// same recursive quadrant choice with a, b, c, d perturbed by +-5 % and
// renormalised after every level, same "undirected => add the reverse tuple",
// same canonical CSR (sorted by (row, col), duplicates and self loops dropped).
// The reference draws from unseeded libc rand() one edge after another (about
// 10^11 calls at scale 25) and is not reproducible; here every draw is a
// counter-based hash of (seed, edge, level, slot), so all edges are generated in
// parallel and oracle/oracle.c (orc_rmat_edge) produces the identical list on the
// CPU.  Arithmetic is IEEE double without contraction (explicit _rn intrinsics).
//
// With a partition only tuples whose row this device owns are kept, and rows are
// stored by LOCAL index, so each device builds its own slice without ever
// holding the whole graph.
#pragma once
#include "../graph/build.cuh"

namespace gunrock {
namespace graphio {

__host__ __device__ __forceinline__ unsigned long long Mix64(unsigned long long z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

struct RmatParams {
    int scale;
    unsigned long long edges;         // generated (directed) tuples
    unsigned long long seed_mix;      // Mix64(seed + 0x632BE59BD9B4E019)
    double a, b, c, d;
    int undirected;
    graph::Partition part;
};

__device__ __forceinline__ double U01(unsigned bits) { return __dmul_rn((double)bits, 1.0 / 4294967296.0); }

__device__ __forceinline__ void RmatEdge(const RmatParams &p, unsigned long long e, int &u_out, int &v_out)
{
    double a = p.a, b = p.b, c = p.c, d = p.d;
    int u = 0, v = 0;
    const unsigned long long ebase = e * 0x9E3779B97F4A7C15ULL;
    for (int level = 0; level < p.scale; ++level) {
        const int step = 1 << (p.scale - 1 - level);
        const unsigned long long c0 = ebase + ((unsigned long long)level << 8);
        const unsigned long long r0 = Mix64(p.seed_mix ^ c0);
        const unsigned long long r1 = Mix64(p.seed_mix ^ (c0 + 1ull));
        const unsigned long long r2 = Mix64(p.seed_mix ^ (c0 + 2ull));
        const double x = U01((unsigned)(r0 >> 32));
        const double ab = __dadd_rn(a, b), abc = __dadd_rn(ab, c), abcd = __dadd_rn(abc, d);
        if (x < a) { }
        else if (a < x && x < ab) { v += step; }
        else if (ab < x && x < abc) { u += step; }
        else if (abc < x && x < abcd) { u += step; v += step; }
        const double va = __dmul_rn(__dmul_rn(a, 0.05), U01((unsigned)(r1 >> 32)));
        const double vb = __dmul_rn(__dmul_rn(b, 0.05), U01((unsigned)r1));
        const double vc = __dmul_rn(__dmul_rn(c, 0.05), U01((unsigned)(r2 >> 32)));
        const double vd = __dmul_rn(__dmul_rn(d, 0.05), U01((unsigned)r2));
        a = (r0 & 1) ? __dadd_rn(a, va) : __dsub_rn(a, va);
        b = (r0 & 2) ? __dadd_rn(b, vb) : __dsub_rn(b, vb);
        c = (r0 & 4) ? __dadd_rn(c, vc) : __dsub_rn(c, vc);
        d = (r0 & 8) ? __dadd_rn(d, vd) : __dsub_rn(d, vd);
        const double S = __dadd_rn(__dadd_rn(__dadd_rn(a, b), c), d);
        const double inv = __ddiv_rn(1.0, S);
        a = __dmul_rn(a, inv); b = __dmul_rn(b, inv); c = __dmul_rn(c, inv); d = __dmul_rn(d, inv);
    }
    u_out = u; v_out = v;
}

// sort key of tuple (row, col): (local row << scale) | col
__device__ __forceinline__ unsigned long long TupleKey(const RmatParams &p, int row, int col)
{
    return ((unsigned long long)p.part.Local(row) << p.scale) | (unsigned long long)(unsigned)col;
}

// COUNT_ONLY: how many tuples this device keeps; otherwise append them (order is
// irrelevant: they are sorted next)
template <bool COUNT_ONLY>
__global__ void __launch_bounds__(256)
RmatGenerateKernel(RmatParams p, unsigned long long *keys, unsigned long long cap, unsigned long long *counter)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    unsigned long long mine = 0;
    const unsigned long long rounds = (p.edges + stride - 1) / stride;
    for (unsigned long long r = 0; r < rounds; ++r) {
        const unsigned long long e = r * stride + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
        int u = 0, v = 0;
        bool fwd = false, rev = false;
        if (e < p.edges) {
            RmatEdge(p, e, u, v);
            fwd = p.part.Mine(u);
            rev = p.undirected && p.part.Mine(v);
        }
        if (COUNT_ONLY) { mine += (fwd ? 1 : 0) + (rev ? 1 : 0); continue; }
        const unsigned mf = __ballot_sync(0xffffffffu, fwd), mr = __ballot_sync(0xffffffffu, rev);
        const unsigned total = __popc(mf) + __popc(mr);
        if (total == 0) continue;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(counter, (unsigned long long)total);
        base = __shfl_sync(0xffffffffu, base, 0);
        const unsigned below = (1u << lane) - 1;
        if (fwd) { const unsigned long long at = base + __popc(mf & below); if (at < cap) keys[at] = TupleKey(p, u, v); }
        if (rev) { const unsigned long long at = base + __popc(mf) + __popc(mr & below); if (at < cap) keys[at] = TupleKey(p, v, u); }
    }
    if (COUNT_ONLY) {
        for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if (lane == 0 && mine) atomicAdd(counter, mine);
    }
}

// symmetric weights in [1, 64] (both directions of an edge hash the same pair;
// the reference's loader gives both directions the file's value, market.cuh:173-181)
static __global__ void EdgeWeightsKernel(const int *krow, const int *col, unsigned long long m, graph::Partition part,
                                  unsigned long long seed, unsigned *w)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride) {
        const unsigned long long u = (unsigned long long)part.Global(krow[j]), v = (unsigned long long)(unsigned)col[j];
        const unsigned long long lo = u < v ? u : v, hi = u < v ? v : u;
        w[j] = 1u + (unsigned)(Mix64(seed ^ (hi << 32 | lo)) & 63ull);
    }
}

// Builds the CSR slice of `part` in `out` (row_offsets / col_indices / edge_values are
// allocated here).  Returns cudaErrorInvalidValue if the slice needs 64-bit offsets and
// SizeT is 32 bits wide.
template <typename SizeT>
cudaError_t BuildRmatCsr(int scale, unsigned long long edges, unsigned long long seed, bool undirected,
                         double a, double b, double c, double d, graph::Partition part,
                         unsigned long long weight_seed, graph::DeviceCsr<int, SizeT> &out,
                         unsigned long long *kept_out, cudaStream_t s)
{
    RmatParams p;
    p.scale = scale; p.edges = edges; p.seed_mix = Mix64(seed + 0x632BE59BD9B4E019ULL);
    p.a = a; p.b = b; p.c = c; p.d = d; p.undirected = undirected ? 1 : 0; p.part = part;
    const long long nodes = 1ll << scale, local_nodes = nodes >> part.log_parts;

    util::DeviceArray<unsigned long long> counter;
    GR_TRY(counter.Allocate(2));
    GR_TRY(counter.Zero(s));
    unsigned long long tuples = 0;
    const unsigned gen_grid = 148 * 16;
    if (part.log_parts == 0) {
        tuples = undirected ? 2 * edges : edges;
    } else {
        RmatGenerateKernel<true><<<gen_grid, 256, 0, s>>>(p, nullptr, 0, counter.ptr);
        GR_TRY(cudaGetLastError());
        GR_TRY(cudaMemcpyAsync(&tuples, counter.ptr, sizeof(tuples), cudaMemcpyDeviceToHost, s));
        GR_TRY(cudaStreamSynchronize(s));
        GR_TRY(counter.Zero(s));
    }
    util::DeviceArray<unsigned long long> keys_a, keys_b;
    GR_TRY(keys_a.Allocate((size_t)tuples + 1));
    GR_TRY(keys_b.Allocate((size_t)tuples + 1));
    if (tuples) {
        RmatGenerateKernel<false><<<gen_grid, 256, 0, s>>>(p, keys_a.ptr, tuples, counter.ptr);
        GR_TRY(cudaGetLastError());
    }
    util::RadixSortScratch sort_scratch;
    unsigned long long *sorted = keys_a.ptr;
    const int key_bits = scale + (scale - part.log_parts);
    GR_TRY(util::RadixSort64(keys_a.ptr, keys_b.ptr, tuples, key_bits, sort_scratch, &sorted, s));
    unsigned long long *spare = (sorted == keys_a.ptr) ? keys_b.ptr : keys_a.ptr;

    // duplicates / self loops out, columns + rows compacted (graph/build.cuh)
    unsigned long long m = 0;
    int *krow = nullptr;                                    // the idle sort buffer holds the kept rows
    out.nodes = nodes; out.local_nodes = local_nodes; out.part = part;
    out.symmetric = undirected;
    GR_TRY(graph::SortedKeysToCsr<SizeT>(sorted, spare, tuples, scale, part, true, local_nodes, out.row_offsets,
                                         out.col_indices, &m, &krow, s));
    out.edges = (long long)m;
    if (weight_seed) {
        GR_TRY(out.edge_values.Allocate((size_t)m));
        if (m) {
            EdgeWeightsKernel<<<util::FillGrid(m), 256, 0, s>>>(krow, out.col_indices.ptr, m, part, weight_seed,
                                                               out.edge_values.ptr);
            GR_TRY(cudaGetLastError());
        }
    } else {
        out.edge_values.size = 0;
    }
    GR_TRY(cudaStreamSynchronize(s));
    if (kept_out) *kept_out = m;
    return cudaSuccess;
}

}  // namespace graphio
}  // namespace gunrock
