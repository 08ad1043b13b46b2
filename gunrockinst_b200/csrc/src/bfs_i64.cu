// bfs_i64.cu -- BFS instantiations for 64-bit edge offsets (graphs past 2^31 entries;
// the reference's SizeT is int only, gunrock.h:33-36).
#include "bfs_runner.cuh"

BfsRunnerBase *MakeBfsRunner64(Csr64 *g, const grb_bfs_config *cfg, cudaStream_t s, cudaError_t *err,
                               const int *to_new, const int *to_old)
{
    const bool pred = cfg && cfg->mark_pred;
    if (pred) return MakeOne<BfsRunner<long long, true, false>>(g, cfg, s, err, to_new, to_old);
    return MakeOne<BfsRunner<long long, false, false>>(g, cfg, s, err, to_new, to_old);
}
