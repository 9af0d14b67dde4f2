// Shared helpers for the deepore_b200 CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/deepore_b200.h"

namespace dp {

void set_error(const char *fmt, ...);

#define DP_CUDA_CHECK(expr)                                                                    \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            dp::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return DP_ECUDA;                                                                   \
        }                                                                                      \
    } while (0)

#define DP_REQUIRE(cond, msg)                                   \
    do {                                                        \
        if (!(cond)) {                                          \
            dp::set_error("%s:%d: %s", __FILE__, __LINE__, msg); \
            return DP_EINVAL;                                   \
        }                                                       \
    } while (0)

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

}  // namespace dp
