// probe_micro.cu -- microbenchmark behind DESIGN.md's "what bounds a top-down level":
// a stream of neighbour ids (4 B each, R-MAT distributed) is read once and every id is looked
// up in a 1-bit-per-vertex visited map.  Variants:
//   stream      read the ids only (the HBM bound)
//   g32         coalesced 4-byte id loads, lookup in the global map (round-1 ExpandKernel's shape)
//   g128        16-byte id loads, lookup in the global map
//   smem<K>     16-byte id loads; ids < K are looked up in a copy of the map's head held in shared
//               memory (K bits), the others in the global map
//   tma<K>      as smem<K>, the id tiles arrive through cp.async.bulk (TMA 1-D) into a 4-stage ring
// each on the ids as generated ("natural": hot vertices = ids with few one-bits) and after
// relabelling the vertices by descending probability ("ordered": hot vertices = low ids).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o build/probe_micro tools/probe_micro.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ unsigned long long splitmix(unsigned long long x)
{
    x += 0x9e3779b97f4a7c15ull; x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull; return x ^ (x >> 31);
}

// id with every bit set with probability 1/4 (R-MAT a+c = 0.75 for the destination bit)
__global__ void GenKernel(int *ids, unsigned long long count, int scale, const int *perm)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        unsigned long long r0 = splitmix(i * 2 + 1), r1 = splitmix(i * 2 + 2);
        unsigned v = (unsigned)(r0 & r1) & ((1u << scale) - 1);     // AND of two fair bits: P(1) = 1/4
        ids[i] = perm ? perm[v] : (int)v;
    }
}

__device__ __forceinline__ int4 LdStream4(const int4 *p)
{
    int4 v;
    asm volatile("ld.global.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ int LdStream(const int *p)
{
    int v; asm volatile("ld.global.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}

__global__ void __launch_bounds__(256, 8) StreamKernel(const int4 *ids, unsigned long long n4, unsigned long long *out)
{
    unsigned acc = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        int4 v = LdStream4(ids + i);
        acc += (unsigned)(v.x ^ v.y ^ v.z ^ v.w);
    }
    if (acc == 0x12345678u) atomicAdd(out, 1ull);
}

__global__ void __launch_bounds__(256, 8) G32Kernel(const int *ids, unsigned long long n, const unsigned *map, unsigned long long *out)
{
    unsigned hits = 0;
    const unsigned long long tile = 1024, tiles = n / tile;
    for (unsigned long long t = blockIdx.x; t < tiles; t += gridDim.x) {
        int v[4]; unsigned w[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) v[r] = LdStream(ids + t * tile + r * 256 + threadIdx.x);
#pragma unroll
        for (int r = 0; r < 4; ++r) w[r] = __ldg(map + ((unsigned)v[r] >> 5));
#pragma unroll
        for (int r = 0; r < 4; ++r) hits += (w[r] >> (v[r] & 31)) & 1u;
    }
    hits = __reduce_add_sync(0xffffffffu, hits);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, (unsigned long long)hits);
}

__global__ void __launch_bounds__(256, 8) G128Kernel(const int4 *ids, unsigned long long n4, const unsigned *map, unsigned long long *out)
{
    unsigned hits = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        int4 q = LdStream4(ids + i);
        int v[4] = {q.x, q.y, q.z, q.w}; unsigned w[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) w[r] = __ldg(map + ((unsigned)v[r] >> 5));
#pragma unroll
        for (int r = 0; r < 4; ++r) hits += (w[r] >> (v[r] & 31)) & 1u;
    }
    hits = __reduce_add_sync(0xffffffffu, hits);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, (unsigned long long)hits);
}

// ids < hot_bits are looked up in shared memory
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1) SmemKernel(const int4 *ids, unsigned long long n4, const unsigned *map,
                                                         unsigned hot_bits, unsigned long long *out)
{
    extern __shared__ unsigned s_map[];
    for (unsigned w = threadIdx.x; w < hot_bits / 32; w += THREADS) s_map[w] = map[w];
    __syncthreads();
    unsigned hits = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * THREADS;
    for (unsigned long long i = (unsigned long long)blockIdx.x * THREADS + threadIdx.x; i < n4; i += stride) {
        int4 q = LdStream4(ids + i);
        int v[4] = {q.x, q.y, q.z, q.w}; unsigned w[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) w[r] = ((unsigned)v[r] < hot_bits) ? s_map[(unsigned)v[r] >> 5] : __ldg(map + ((unsigned)v[r] >> 5));
#pragma unroll
        for (int r = 0; r < 4; ++r) hits += (w[r] >> (v[r] & 31)) & 1u;
    }
    hits = __reduce_add_sync(0xffffffffu, hits);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, (unsigned long long)hits);
}

// ---- TMA 1-D ring ---------------------------------------------------------------------------------
__device__ __forceinline__ unsigned SmemAddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void MbarInit(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(SmemAddr(bar)), "r"(count));
}
__device__ __forceinline__ void MbarExpectTx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(SmemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void MbarArrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(SmemAddr(bar)) : "memory");
}
__device__ __forceinline__ void MbarWait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" :: "r"(SmemAddr(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void BulkLoad(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(SmemAddr(dst)), "l"(src), "r"(bytes), "r"(SmemAddr(bar)) : "memory");
}

template <int THREADS, int STAGES, int TILE>     // TILE ids per stage
__global__ void __launch_bounds__(THREADS, 1) TmaKernel(const int *ids, unsigned long long n, const unsigned *map,
                                                        unsigned hot_bits, unsigned long long *out)
{
    extern __shared__ __align__(128) unsigned char s_raw[];
    int *s_tile = reinterpret_cast<int *>(s_raw);                                   // [STAGES][TILE]
    unsigned long long *full  = reinterpret_cast<unsigned long long *>(s_raw + sizeof(int) * STAGES * TILE);
    unsigned long long *empty = full + STAGES;
    unsigned *s_map = reinterpret_cast<unsigned *>(empty + STAGES);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { MbarInit(full + s, 1); MbarInit(empty + s, THREADS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (unsigned w = threadIdx.x; w < hot_bits / 32; w += THREADS) s_map[w] = map[w];
    __syncthreads();
    const unsigned long long tiles = n / TILE;
    const unsigned long long my_tiles = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    // prologue: the producer (thread 0) fills the ring
    if (threadIdx.x == 0) {
        for (unsigned long long k = 0; k < STAGES && k < my_tiles; ++k) {
            MbarExpectTx(full + k, TILE * 4);
            BulkLoad(s_tile + k * TILE, ids + (blockIdx.x + k * gridDim.x) * TILE, TILE * 4, full + k);
        }
    }
    unsigned hits = 0;
    for (unsigned long long k = 0; k < my_tiles; ++k) {
        const int s = (int)(k % STAGES);
        const unsigned phase = (unsigned)((k / STAGES) & 1);
        MbarWait(full + s, phase);
        const int4 *t4 = reinterpret_cast<const int4 *>(s_tile + s * TILE);
#pragma unroll
        for (int j = 0; j < TILE / 4 / THREADS; ++j) {
            int4 q = t4[j * THREADS + threadIdx.x];
            int v[4] = {q.x, q.y, q.z, q.w}; unsigned w[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) w[r] = ((unsigned)v[r] < hot_bits) ? s_map[(unsigned)v[r] >> 5] : __ldg(map + ((unsigned)v[r] >> 5));
#pragma unroll
            for (int r = 0; r < 4; ++r) hits += (w[r] >> (v[r] & 31)) & 1u;
        }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) MbarArrive(empty + s);
        if (threadIdx.x == 0 && k + STAGES < my_tiles) {       // refill the slot once every warp has left it
            MbarWait(empty + s, phase);
            MbarExpectTx(full + s, TILE * 4);
            BulkLoad(s_tile + s * TILE, ids + (blockIdx.x + (k + STAGES) * gridDim.x) * TILE, TILE * 4, full + s);
        }
    }
    hits = __reduce_add_sync(0xffffffffu, hits);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, (unsigned long long)hits);
}

static float TimeIt(void (*launch)(void *), void *ctx, int reps)
{
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    launch(ctx); launch(ctx);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(a)); launch(ctx); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); best = std::min(best, ms);
    }
    CK(cudaGetLastError());
    return best;
}

struct Ctx { const int *ids; unsigned long long n; const unsigned *map; unsigned hot_bits; unsigned long long *out; int sms; };

int main(int argc, char **argv)
{
    const int scale = argc > 1 ? atoi(argv[1]) : 25;
    const unsigned long long n = 1ull << (argc > 2 ? atoi(argv[2]) : 28);
    int sms = 148; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const unsigned nv = 1u << scale;
    // relabel by descending probability = ascending popcount
    std::vector<int> perm(nv);
    {
        std::vector<unsigned> cnt(33, 0);
        for (unsigned v = 0; v < nv; ++v) ++cnt[__builtin_popcount(v)];
        std::vector<unsigned> start(34, 0);
        for (int k = 0; k < 33; ++k) start[k + 1] = start[k] + cnt[k];
        for (unsigned v = 0; v < nv; ++v) perm[v] = (int)start[__builtin_popcount(v)]++;
    }
    int *d_perm, *d_ids; unsigned *d_map; unsigned long long *d_out;
    CK(cudaMalloc(&d_perm, sizeof(int) * nv)); CK(cudaMemcpy(d_perm, perm.data(), sizeof(int) * nv, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_ids, sizeof(int) * n));
    CK(cudaMalloc(&d_map, nv / 8 + 256 * 1024)); CK(cudaMemset(d_map, 0xff, nv / 8 + 256 * 1024));
    CK(cudaMalloc(&d_out, 8)); CK(cudaMemset(d_out, 0, 8));
    CK(cudaFuncSetAttribute(SmemKernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    CK(cudaFuncSetAttribute(SmemKernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    CK(cudaFuncSetAttribute((TmaKernel<1024, 4, 4096>), cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    CK(cudaFuncSetAttribute((TmaKernel<512, 4, 2048>), cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));

    printf("{\"scale\": %d, \"ids\": %llu, \"sms\": %d, \"results\": [\n", scale, n, sms);
    bool first = true;
    for (int ordered = 0; ordered < 2; ++ordered) {
        GenKernel<<<sms * 8, 256>>>(d_ids, n, scale, ordered ? d_perm : nullptr);
        CK(cudaDeviceSynchronize());
        Ctx c = {d_ids, n, d_map, 0, d_out, sms};
        auto report = [&](const char *name, unsigned hot, float ms) {
            printf("%s {\"order\": \"%s\", \"variant\": \"%s\", \"hot_bits\": %u, \"ms\": %.4f, \"gids_per_s\": %.1f, \"col_gbs\": %.1f}",
                   first ? "" : ",\n", ordered ? "ordered" : "natural", name, hot, ms, n / ms * 1e-6, 4.0 * n / ms * 1e-6);
            first = false; fflush(stdout);
        };
        report("stream", 0, TimeIt([](void *p) { Ctx *c = (Ctx *)p; StreamKernel<<<c->sms * 8, 256>>>((const int4 *)c->ids, c->n / 4, c->out); }, &c, 5));
        report("g32", 0, TimeIt([](void *p) { Ctx *c = (Ctx *)p; G32Kernel<<<c->sms * 8, 256>>>(c->ids, c->n, c->map, c->out); }, &c, 5));
        report("g128", 0, TimeIt([](void *p) { Ctx *c = (Ctx *)p; G128Kernel<<<c->sms * 8, 256>>>((const int4 *)c->ids, c->n / 4, c->map, c->out); }, &c, 5));
        const unsigned hots[] = {1u << 18, 1u << 19, 1u << 20, 3u << 19, 7u << 18};
        for (unsigned hot : hots) {
            if (hot > nv) continue;
            c.hot_bits = hot;
            report("smem1024", hot, TimeIt([](void *p) { Ctx *c = (Ctx *)p; SmemKernel<1024><<<c->sms, 1024, c->hot_bits / 8>>>((const int4 *)c->ids, c->n / 4, c->map, c->hot_bits, c->out); }, &c, 5));
            report("smem512", hot, TimeIt([](void *p) { Ctx *c = (Ctx *)p; SmemKernel<512><<<c->sms, 512, c->hot_bits / 8>>>((const int4 *)c->ids, c->n / 4, c->map, c->hot_bits, c->out); }, &c, 5));
            if (hot / 8 + 4 * 4096 * 4 + 128 <= 224 * 1024)
                report("tma1024", hot, TimeIt([](void *p) { Ctx *c = (Ctx *)p; TmaKernel<1024, 4, 4096><<<c->sms, 1024, c->hot_bits / 8 + 4 * 4096 * 4 + 128>>>(c->ids, c->n, c->map, c->hot_bits, c->out); }, &c, 5));
            if (hot / 8 + 4 * 2048 * 4 + 128 <= 224 * 1024)
                report("tma512", hot, TimeIt([](void *p) { Ctx *c = (Ctx *)p; TmaKernel<512, 4, 2048><<<c->sms, 512, c->hot_bits / 8 + 4 * 2048 * 4 + 128>>>(c->ids, c->n, c->map, c->hot_bits, c->out); }, &c, 5));
        }
    }
    printf("\n]}\n");
    return 0;
}
