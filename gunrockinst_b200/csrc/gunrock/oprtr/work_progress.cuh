// work_progress.cuh -- device-side frontier bookkeeping shared by the operators.
//
// Counterpart of the reference's util::CtaWorkProgress(Lifetime)
// (gunrock/util/cta_work_progress.cuh): there, 4 rotating queue-length
// counters indexed queue_index & 3, with each kernel zeroing counter
// queue_index + 2 and the host reading one back with a 4-byte cudaMemcpy after
// every kernel.  Here the same rotation is kept (so advance -> filter can be
// chained without a host round trip), but a slot carries everything the
// enactor wants to know about a level, and the host fetches a whole 64-byte
// slot once per level into pinned memory.
#pragma once
#include "../util/array.cuh"

namespace gunrock {
namespace oprtr {

struct alignas(64) LevelCounters {
    unsigned long long out_len;     // items written to the output frontier
    unsigned long long out_degree;  // sum of out-degrees of those items (if tracked)
    unsigned long long edges;       // neighbour entries inspected by the advance
    unsigned long long scan_total;  // LB advance: sum of degrees of the INPUT frontier
    unsigned long long aux0;        // primitive specific (PR: vertices still moving, ...)
    unsigned long long aux1;        // primitive specific
    unsigned long long overflow;    // != 0: output frontier capacity exceeded
    unsigned long long reserved;
};

struct WorkProgress {
    static const int kSlots = 4;
    util::DeviceArray<LevelCounters> d_slots;
    util::PinnedArray<LevelCounters> h_slot;
    // two sets (by level parity) of chunk-ticket counters for the pull sweep, each counter on its own line
    static const int kTicketWords = 16 * 16;
    util::DeviceArray<unsigned long long> d_tickets;

    cudaError_t Setup(cudaStream_t s)
    {
        GR_TRY(d_slots.Allocate(kSlots));
        GR_TRY(h_slot.Allocate(kSlots));
        GR_TRY(d_tickets.Allocate(2 * kTicketWords));
        GR_TRY(d_tickets.Zero(s));
        return d_slots.Zero(s);
    }
    unsigned long long *Tickets(long long level) const { return d_tickets.ptr + (level & 1) * kTicketWords; }
    __host__ LevelCounters *Slot(long long queue_index) const { return d_slots.ptr + (queue_index & (kSlots - 1)); }

    // host view of the slot a kernel with this queue_index accumulated into
    cudaError_t Fetch(long long queue_index, LevelCounters &out, cudaStream_t s)
    {
        GR_TRY(cudaMemcpyAsync(h_slot.ptr, Slot(queue_index), sizeof(LevelCounters), cudaMemcpyDeviceToHost, s));
        GR_TRY(cudaStreamSynchronize(s));
        out = h_slot.ptr[0];
        return cudaSuccess;
    }
    cudaError_t Reset(cudaStream_t s) { return d_slots.Zero(s); }
};

// called by ONE thread of a kernel working on `queue_index`: clears the slot two
// steps ahead so it is clean when a later kernel starts accumulating into it
__device__ __forceinline__ void ClearFutureSlot(LevelCounters *slots, long long queue_index)
{
    LevelCounters *f = slots + ((queue_index + 2) & (WorkProgress::kSlots - 1));
    f->out_len = 0; f->out_degree = 0; f->edges = 0; f->scan_total = 0;
    f->aux0 = 0; f->aux1 = 0; f->overflow = 0;
}

}  // namespace oprtr
}  // namespace gunrock
