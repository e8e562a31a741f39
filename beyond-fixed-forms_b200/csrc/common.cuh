// Shared helpers of libbff.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "bff.h"

namespace bff {

void set_error(const char *fmt, ...);

#define BFF_REQUIRE(cond, ...)                 \
    do {                                       \
        if (!(cond)) {                         \
            ::bff::set_error(__VA_ARGS__);     \
            return BFF_EINVAL;                 \
        }                                      \
    } while (0)

#define BFF_CUDA(call)                                                                       \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            ::bff::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return BFF_ECUDA;                                                                \
        }                                                                                    \
    } while (0)

#define BFF_LAUNCH_CHECK(name)                                                               \
    do {                                                                                     \
        cudaError_t e_ = cudaGetLastError();                                                 \
        if (e_ != cudaSuccess) {                                                             \
            ::bff::set_error("launch of %s failed: %s", name, cudaGetErrorString(e_));       \
            return BFF_ECUDA;                                                                \
        }                                                                                    \
    } while (0)

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

constexpr int kNumSMs = 148;  // B200

}  // namespace bff
