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

// Grow-only bump arena (device or pinned host). reset() at the start of a call recycles all blocks;
// pointers stay valid until the next reset because blocks are never moved, only added.
struct Arena {
    struct Block { unsigned char *p; size_t cap; };
    Block blocks[32];
    int n_blocks = 0, cur = 0;
    size_t off = 0;
    bool pinned_host = false;
    void reset() { cur = 0; off = 0; }
    void *take(size_t bytes);          // nullptr on allocation failure (error text set)
    void release();
    template <typename T> T *get(size_t count) { return reinterpret_cast<T *>(take(count * sizeof(T))); }
};

}  // namespace bff

// A context owns one stream and grow-only workspaces used by the host-buffer and pipeline entry points.
struct bff_ctx {
    int device;
    cudaStream_t stream;
    unsigned char *ws;      // bff_*_host staging (single block)
    size_t ws_bytes;
    bff::Arena dev;         // pipeline scratch on the device
    bff::Arena pin;         // pipeline scratch in pinned host memory
    // side streams of the batched pipelines (created on first use): job q runs on lane q % n_lanes between a fork from
    // and a join into the caller's stream, so the small kernels of independent jobs overlap
    static constexpr int kMaxLanes = 8;
    cudaStream_t lane[kMaxLanes];
    cudaEvent_t lane_done[kMaxLanes];
    cudaEvent_t fork_ev;
    int n_lanes;
};
