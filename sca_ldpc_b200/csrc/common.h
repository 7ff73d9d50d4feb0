// Shared host-side helpers of libscaldpc (error reporting, CUDA checks).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <string>

#include "../../include/scaldpc.h"

namespace scaldpc {

// Thread-local last-error string behind scaldpc_last_error().
std::string &last_error();
int fail(int code, const char *fmt, ...);

#define SCALDPC_CUDA(expr)                                                                         \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return ::scaldpc::fail(SCALDPC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,               \
                                   cudaGetErrorString(_e), __FILE__, __LINE__);                    \
    } while (0)

constexpr int kLanes = 32;  // trials per tile == warp width; lane = trial % 32

}  // namespace scaldpc
