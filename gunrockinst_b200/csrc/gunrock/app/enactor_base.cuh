// enactor_base.cuh -- state every primitive's Enactor shares.
//
// Counterpart of the reference's EnactorBase / EnactorStats / FrontierAttribute
// (gunrock/app/enactor_base.cuh:36-190).  The reference sizes its grids from a
// 2010-era device table (util/cuda_properties.cuh:57-66) and keeps a mapped
// pinned `done` flag plus a throttle event; here the device is queried once,
// the operators run persistent grids that are multiples of the SM count, and
// the host learns about a level from one 64-byte counter slot.
#pragma once
#include "../oprtr/advance/kernel.cuh"
#include "../oprtr/filter/kernel.cuh"
#include "../oprtr/filter/stream_kernel.cuh"

namespace gunrock {
namespace app {

struct EnactorStats {
    long long          iteration      = 0;
    unsigned long long total_queued   = 0;    // sum of frontier lengths
    unsigned long long edges_inspected = 0;   // neighbour entries the advances looked at
    unsigned long long algo_bytes     = 0;    // algorithmic HBM bytes (DESIGN.md formulas)
    int                push_levels = 0, pull_levels = 0;
    int                launches = 0;           // kernels launched by Enact
};

struct FrontierAttribute {
    unsigned long long queue_length = 0;
    long long          queue_index  = 0;
    int                selector     = 0;
};

class EnactorBase {
  public:
    oprtr::advance::DeviceInfo dev;
    oprtr::WorkProgress        work_progress;
    EnactorStats               enactor_stats;
    FrontierAttribute          frontier_attribute;
    cudaStream_t               stream = 0;
    bool                       DEBUG  = false;

    cudaError_t SetupBase(int device, cudaStream_t s)
    {
        stream = s;
        GR_TRY(dev.Query(device));
        return work_progress.Setup(s);
    }
    cudaError_t ResetBase()
    {
        enactor_stats = EnactorStats();
        frontier_attribute = FrontierAttribute();
        return work_progress.Reset(stream);
    }
};

}  // namespace app
}  // namespace gunrock
