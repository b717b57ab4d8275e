// Host-side plumbing shared by the translation units of libcst_reward.so.
#pragma once

#include <cuda_runtime.h>

#include "cst_reward.h"

// record the message returned by cst_last_error() and hand `code` back
int cst_fail(int code, const char* fmt, ...);
// one more kernel launched by this library (cst_kernel_launches)
void cst_count_launch(void);

#define CST_CUDA_TRY(expr)                                                                  \
    do {                                                                                    \
        cudaError_t e_ = (expr);                                                            \
        if (e_ != cudaSuccess)                                                              \
            return cst_fail(CST_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));    \
    } while (0)

// make `dev` current for the scope, restore the previous device afterwards
struct CstDeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit CstDeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) return;
        ok = (prev == dev) || (cudaSetDevice(dev) == cudaSuccess);
    }
    ~CstDeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
