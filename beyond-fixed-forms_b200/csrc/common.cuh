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
constexpr int32_t kMaxGridY = 65535;       // rows per launch of the kernels that index rows by blockIdx.y: the entry points loop over row blocks

constexpr int kNumSMs = 148;  // B200

// Grow-only bump arena (device or pinned host). reset() at the start of a call recycles all blocks;
// pointers stay valid until the next reset because blocks are never moved, only added.
struct Arena {
    struct Block { unsigned char *p; size_t cap; };
    Block blocks[32];
    int n_blocks = 0, cur = 0;
    size_t off = 0;
    bool pinned_host = false;
    unsigned generation = 0;           // bumped whenever reserve() moved the block (cached CUDA graphs hold its pointers)
    void reset();                      // recycle; several blocks are first merged into one (after a device synchronisation)
    void *take(size_t bytes);          // nullptr on allocation failure (error text set)
    void release();
    // One contiguous block of at least `bytes` (the device-resident pipelines carve all their scratch from it with a
    // layout computed up front). Growing synchronises the device, frees everything and allocates anew, so no pointer
    // handed out earlier survives a growth; returns BFF_OK / BFF_ENOMEM.
    int reserve(size_t bytes);
    unsigned char *base() const { return n_blocks ? blocks[0].p : nullptr; }
    template <typename T> T *get(size_t count) { return reinterpret_cast<T *>(take(count * sizeof(T))); }
};

// internal launchers shared between translation units (kernels are only launchable from the file that defines them)
int launch_merge_adjacency(const int32_t *inter, const int32_t *pop, const int32_t *label, int n, int stride_words, float thr,
                           uint32_t *adj, cudaStream_t s);
int launch_set_i32(int32_t *dst, int32_t value, cudaStream_t s);
// Sum (op 0) / max (op 1) of int32 buffers over the ranks of `comm` on `s`; a null communicator is a single GPU: no-op.
int comm_allreduce_i32(bff_comm *comm, int32_t *buf, size_t count, int op, cudaStream_t s);
static inline int comm_allreduce_sum_i32(bff_comm *comm, int32_t *buf, size_t count, cudaStream_t s) {
    return comm ? comm_allreduce_i32(comm, buf, count, 0, s) : BFF_OK;
}

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
    // side stream of the device-resident pipelines (created on first use): the vote-threshold chain does not depend on
    // the intersection chain until the point filter is applied, so it runs next to it
    cudaStream_t side;
    cudaEvent_t side_fork, side_join;
    int has_side;
};
