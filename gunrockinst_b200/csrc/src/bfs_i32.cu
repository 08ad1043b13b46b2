// bfs_i32.cu -- BFS instantiations for 32-bit edge offsets (the gunrock.h type triple
// <int,int,int>, gunrock/app/bfs/bfs_app.cu:248-352).
#include "bfs_runner.cuh"

BfsRunnerBase *MakeBfsRunner32(Csr32 *g, const grb_bfs_config *cfg, cudaStream_t s, cudaError_t *err,
                               const int *to_new, const int *to_old)
{
    const bool pred = cfg && cfg->mark_pred, idem = cfg && cfg->idempotence;
    // idempotent + predecessors: the reference stores the PREDECESSOR in labels[] in that
    // combination and skips validating it (tests/bfs/test_bfs.cu:472-482); served by the
    // exact (non-idempotent) path so that depths and parents are both well defined
    if (pred)  return MakeOne<BfsRunner<int, true, false>>(g, cfg, s, err, to_new, to_old);
    if (idem)  return MakeOne<BfsRunner<int, false, true>>(g, cfg, s, err, to_new, to_old);
    return MakeOne<BfsRunner<int, false, false>>(g, cfg, s, err, to_new, to_old);
}
