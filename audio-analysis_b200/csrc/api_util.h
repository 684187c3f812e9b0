// Host-side helpers shared by the C-ABI translation units (error string, CUDA status mapping).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include "../../include/cacophony_b200.h"

namespace cac {
int fail(int code, const std::string& msg);   // records the calling thread's last error, returns code

template <class T>
inline int ensure_cap(T** ptr, size_t* cap, size_t bytes) {
    if (bytes <= *cap) return CAC_OK;
    if (*ptr) cudaFree(*ptr);
    *ptr = nullptr;
    *cap = 0;
    cudaError_t e = cudaMalloc((void**)ptr, bytes);
    if (e != cudaSuccess) return fail(CAC_ERR_ALLOC, std::string("cudaMalloc: ") + cudaGetErrorString(e));
    *cap = bytes;
    return CAC_OK;
}
}  // namespace cac

#define CUDA_TRY(expr)                                                                              \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess)                                                                      \
            return cac::fail(CAC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));     \
    } while (0)
