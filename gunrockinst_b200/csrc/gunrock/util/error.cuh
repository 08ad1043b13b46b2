// error.cuh -- CUDA error reporting for the library.
//
// Behaviour follows the reference's GRError (gunrock/util/error_utils.cu:25-37):
// a failing call prints "[file, line] message (CUDA error N: text)" to stderr
// and the error code is handed back to the caller, which unwinds; nothing
// throws across the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>

namespace gunrock {
namespace util {

inline cudaError_t GRError(cudaError_t err, const char *msg, const char *file, int line)
{
    if (err != cudaSuccess) {
        fprintf(stderr, "[%s, %d] %s (CUDA error %d: %s)\n", file, line, msg,
                (int)err, cudaGetErrorString(err));
        fflush(stderr);
    }
    return err;
}

}  // namespace util
}  // namespace gunrock

// evaluate `call`; on failure report it and make the enclosing function return the error
#define GR_TRY(call)                                                                   \
    do {                                                                               \
        cudaError_t gr_try_err_ = ::gunrock::util::GRError((call), #call, __FILE__, __LINE__); \
        if (gr_try_err_ != cudaSuccess) return gr_try_err_;                            \
    } while (0)
