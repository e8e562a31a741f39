// Vote statistics for the point filter of tools/projection_2d_to_3d.py:512-583:
// max of the count vectors, the 2-D (masked_count, viewed_count) histogram from which the host
// derives detected_ratio.unique() / masked_counts.unique(), and the per-point keep predicate.
#include <mutex>

#include "common.cuh"

namespace bff {

__global__ void __launch_bounds__(256)
counts_max_kernel(const int32_t *__restrict__ mc, const int32_t *__restrict__ vc, long long n, int32_t *__restrict__ out2) {
    int m0 = 0, m1 = 0;
    // 16-byte loads over the aligned body, scalar tail; ONE atomic pair per CTA (all CTAs hit the same two words)
    const long long n4 = ((reinterpret_cast<uintptr_t>(mc) | reinterpret_cast<uintptr_t>(vc)) & 15) == 0 ? n / 4 : 0;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
        const int4 a = __ldg(reinterpret_cast<const int4 *>(mc) + i);
        m0 = max(m0, max(max(a.x, a.y), max(a.z, a.w)));
        if (vc != nullptr) {
            const int4 b = __ldg(reinterpret_cast<const int4 *>(vc) + i);
            m1 = max(m1, max(max(b.x, b.y), max(b.z, b.w)));
        }
    }
    for (long long i = n4 * 4 + (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        m0 = max(m0, __ldg(mc + i));
        if (vc != nullptr) m1 = max(m1, __ldg(vc + i));
    }
    for (int o = 16; o > 0; o >>= 1) {
        m0 = max(m0, __shfl_down_sync(0xffffffffu, m0, o));
        m1 = max(m1, __shfl_down_sync(0xffffffffu, m1, o));
    }
    __shared__ int s0[8], s1[8];
    if ((threadIdx.x & 31) == 0) { s0[threadIdx.x >> 5] = m0; s1[threadIdx.x >> 5] = m1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { m0 = max(m0, s0[w]); m1 = max(m1, s1[w]); }
        if (m0 > 0) atomicMax(out2, m0);
        if (m1 > 0) atomicMax(out2 + 1, m1);
    }
}

// Warp-aggregated histogram: lanes that hit the same bin elect one leader which adds the group
// size once (__match_any_sync), first into a CTA-private shared-memory histogram when the bins
// fit, then the non-empty shared bins are flushed with one global atomic each.
constexpr int kHistSmemBins = 8192;

__global__ void __launch_bounds__(256)
vote_hist_kernel(const int32_t *__restrict__ mc, const int32_t *__restrict__ vc, long long n, int mc_bins, int vc_bins,
                 uint32_t *__restrict__ hist, int use_smem) {
    __shared__ uint32_t s_hist[kHistSmemBins];
    const int bins = mc_bins * vc_bins;
    if (use_smem) {
        for (int i = threadIdx.x; i < bins; i += 256) s_hist[i] = 0;
        __syncthreads();
    }
    const long long stride = (long long)gridDim.x * 256;
    const long long n_round = (n + 31) / 32 * 32;   // keep warps converged for __match_any_sync
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n_round; i += stride) {
        int bin = -1;
        if (i < n) {
            int a = __ldg(mc + i), b = vc ? __ldg(vc + i) : 0;
            a = min(max(a, 0), mc_bins - 1);
            b = min(max(b, 0), vc_bins - 1);
            bin = a * vc_bins + b;
        }
        const unsigned peers = __match_any_sync(0xffffffffu, bin);
        const int leader = __ffs(peers) - 1;
        if (bin >= 0 && (int)(threadIdx.x & 31) == leader) {
            if (use_smem) atomicAdd(&s_hist[bin], (uint32_t)__popc(peers));
            else atomicAdd(hist + bin, (uint32_t)__popc(peers));
        }
    }
    if (use_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < bins; i += 256) {
            const uint32_t v = s_hist[i];
            if (v) atomicAdd(hist + i, v);
        }
    }
}

// keep bit per point; float32 arithmetic identical to torch's masked_counts / (viewed_counts + 1)
// on float32 tensors (projection_2d_to_3d.py:571) followed by the strict < compare (:578).
__global__ void __launch_bounds__(256)
point_keep_kernel(const int32_t *__restrict__ mc, const int32_t *__restrict__ vc, long long n, long long words, int mode,
                  float thr, uint32_t *__restrict__ keep) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    bool k = false;
    if (i < n) {
        const int m = __ldg(mc + i);
        if (m > 0) {
            const float fm = __int2float_rn(m);
            if (mode == 0) {
                const float r = __fdiv_rn(fm, __fadd_rn(__int2float_rn(__ldg(vc + i)), 1.0f));
                k = !(r < thr);
            } else if (mode == 1) {
                k = !(fm < thr);
            } else {
                k = true;
            }
        }
    }
    const unsigned w = __ballot_sync(0xffffffffu, k);
    if ((threadIdx.x & 31) == 0 && (i >> 5) < words) keep[i >> 5] = w;
}

}  // namespace bff

using namespace bff;

extern "C" int bff_counts_max(const int32_t *mc, const int32_t *vc, int64_t n, int32_t *out2, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n >= 0 && out2, "bff_counts_max: bad arguments");
    BFF_CUDA(cudaMemsetAsync(out2, 0, 8, stream));
    if (n == 0) return BFF_OK;
    BFF_REQUIRE(mc, "bff_counts_max: null counts");
    long long blocks = ceil_div64(n, 256 * 4);
    if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
    counts_max_kernel<<<(unsigned)blocks, 256, 0, stream>>>(mc, vc, n, out2);
    BFF_LAUNCH_CHECK("counts_max_kernel");
    return BFF_OK;
}

extern "C" int bff_vote_hist(const int32_t *mc, const int32_t *vc, int64_t n, int32_t mc_bins, int32_t vc_bins,
                             uint32_t *hist, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n >= 0 && mc_bins >= 1 && vc_bins >= 1 && hist, "bff_vote_hist: bad arguments");
    BFF_REQUIRE((long long)mc_bins * vc_bins < (1LL << 31), "bff_vote_hist: too many bins");
    BFF_REQUIRE(vc != nullptr || vc_bins == 1, "bff_vote_hist: vc_bins must be 1 without viewed counts");
    BFF_CUDA(cudaMemsetAsync(hist, 0, (size_t)mc_bins * vc_bins * 4, stream));
    if (n == 0) return BFF_OK;
    BFF_REQUIRE(mc, "bff_vote_hist: null counts");
    long long blocks = ceil_div64(n, 256);
    if (blocks > kNumSMs * 4) blocks = kNumSMs * 4;
    const int use_smem = (long long)mc_bins * vc_bins <= kHistSmemBins;
    vote_hist_kernel<<<(unsigned)blocks, 256, 0, stream>>>(mc, vc, n, mc_bins, vc_bins, hist, use_smem);
    BFF_LAUNCH_CHECK("vote_hist_kernel");
    return BFF_OK;
}

// ---------------------------------------------------------------------------------------------------
// bff_vote_unique: the distinct values the reference's unique() would see (projection_2d_to_3d.py:512-517 masked_counts,
// :571-575 detected_ratio = mc / (vc + 1) in float32), from the non-empty bins of the vote histogram: every non-empty bin
// inserts the bit pattern of its value into an open-addressing set (atomicCAS, linear probing; values are finite and
// >= +0, so equal bits <=> equal values), a second kernel compacts the set. Order of `values` is unspecified.
namespace bff {
__global__ void __launch_bounds__(256)
vote_unique_insert_kernel(const uint32_t *__restrict__ hist, int mc_bins, int vc_bins, int mode, uint32_t *__restrict__ table, uint32_t cap_mask,
                          int hash_shift) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= (long long)mc_bins * vc_bins || hist[i] == 0u) return;
    const int a = (int)(i / vc_bins), b = (int)(i % vc_bins);
    const float val = mode == 0 ? __fdiv_rn((float)a, __fadd_rn((float)b, 1.0f)) : (float)a;
    const uint32_t bits = __float_as_uint(val);
    // Fibonacci hashing: the TOP bits of the product. (Ratios of small integers have all-zero low mantissa bits, and the low
    // bits of a product depend on the low bits of the factors only: masking them instead piles every value into a few slots.)
    uint32_t slot = (bits * 2654435761u) >> hash_shift;
    while (true) {
        uint32_t old = *reinterpret_cast<volatile uint32_t *>(table + slot);      // many bins share a value (0 / (b + 1) = 0 ...):
        if (old == bits) return;                                                   // look before the atomic
        if (old == 0xffffffffu) {
            old = atomicCAS(table + slot, 0xffffffffu, bits);
            if (old == 0xffffffffu || old == bits) return;
        }
        slot = (slot + 1u) & cap_mask;
    }
}
__global__ void __launch_bounds__(256)
vote_unique_compact_kernel(const uint32_t *__restrict__ table, uint32_t cap, float *__restrict__ values, int32_t *__restrict__ count) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= cap) return;
    const uint32_t v = table[i];
    if (v != 0xffffffffu) values[atomicAdd(count, 1)] = __uint_as_float(v);
}
}  // namespace bff

extern "C" int64_t bff_vote_unique_table_size(int32_t mc_bins, int32_t vc_bins) {
    int64_t cap = 1024;
    while (cap < (int64_t)mc_bins * vc_bins * 2) cap <<= 1;
    return cap;
}

extern "C" int bff_vote_unique(const uint32_t *hist, int32_t mc_bins, int32_t vc_bins, int32_t mode, uint32_t *table, int64_t table_size,
                               float *values, int32_t *count, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(hist && table && values && count && mc_bins >= 1 && vc_bins >= 1 && (mode == 0 || mode == 1), "bff_vote_unique: bad arguments");
    BFF_REQUIRE(table_size >= bff_vote_unique_table_size(mc_bins, vc_bins) && (table_size & (table_size - 1)) == 0 && table_size <= (1LL << 31),
                "bff_vote_unique: table_size must be a power of two >= bff_vote_unique_table_size()");
    const long long bins = (long long)mc_bins * vc_bins;
    BFF_CUDA(cudaMemsetAsync(table, 0xff, (size_t)table_size * 4, stream));
    BFF_CUDA(cudaMemsetAsync(count, 0, 4, stream));
    vote_unique_insert_kernel<<<(unsigned)ceil_div64(bins, 256), 256, 0, stream>>>(hist, mc_bins, vc_bins, mode, table, (uint32_t)(table_size - 1),
                                                                                     32 - (int)(63 - __builtin_clzll((unsigned long long)table_size)));
    vote_unique_compact_kernel<<<(unsigned)ceil_div64(table_size, 256), 256, 0, stream>>>(table, (uint32_t)table_size, values, count);
    BFF_LAUNCH_CHECK("vote_unique kernels");
    return BFF_OK;
}

// ---------------------------------------------------------------------------------------------------
// bff_vote_unique_static: bff_vote_unique for callers that cannot allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc). No counterpart in the reference. The hash-set
// table -- the scratch the caller of bff_vote_unique has to provide -- lives in a file-scope device array of
// BFF_VOTE_UNIQUE_STATIC_TABLE_WORDS words (1 MiB per GPU that loads the library, held until exit). In host code the
// array's NAME is only a host-side stand-in; its device address is resolved once per device, with that device current,
// by cudaGetSymbolAddress and handed to the same two kernels bff_vote_unique launches. Nothing here calls cudaMalloc,
// cudaFree or an allocating library routine. All calls on a device share the one table (it keeps the previous call's
// contents, so each call re-initialises the part it uses on its stream): calls that overlap on the GPU race.
namespace bff {
constexpr long long kStaticTableWords = BFF_VOTE_UNIQUE_STATIC_TABLE_WORDS;
static_assert((kStaticTableWords & (kStaticTableWords - 1)) == 0 && kStaticTableWords == 2LL * BFF_VOTE_UNIQUE_STATIC_MAX_BINS,
              "the static table is a power of two, twice the largest accepted bin count");
static __device__ __align__(256) uint32_t g_unique_table_static[kStaticTableWords];
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static uint32_t *g_static_table_addr[kStaticMaxDevices] = {};

// device address of g_unique_table_static on the CURRENT device (resolved on that device's first call, then cached)
static int static_table_address(uint32_t **out) {
    int dev = -1;
    BFF_CUDA(cudaGetDevice(&dev));
    BFF_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "bff_vote_unique_static: device index %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (g_static_table_addr[dev] == nullptr) {
        void *addr = nullptr;
        BFF_CUDA(cudaGetSymbolAddress(&addr, g_unique_table_static));
        BFF_REQUIRE(addr != nullptr, "bff_vote_unique_static: cudaGetSymbolAddress returned a null address");
        g_static_table_addr[dev] = static_cast<uint32_t *>(addr);
    }
    *out = g_static_table_addr[dev];
    return BFF_OK;
}
}  // namespace bff

extern "C" int bff_vote_unique_static(const uint32_t *hist, int32_t mc_bins, int32_t vc_bins, int32_t mode, float *values, int32_t *count,
                                      void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(hist && values && count && mc_bins >= 1 && vc_bins >= 1 && (mode == 0 || mode == 1), "bff_vote_unique_static: bad arguments");
    const long long bins = (long long)mc_bins * vc_bins;
    BFF_REQUIRE(bins <= BFF_VOTE_UNIQUE_STATIC_MAX_BINS,
                "bff_vote_unique_static: %d x %d = %lld histogram bins exceed the static scratch table's limit of %d bins "
                "(use bff_vote_unique with a caller-provided table)", mc_bins, vc_bins, bins, BFF_VOTE_UNIQUE_STATIC_MAX_BINS);
    const int64_t table_size = bff_vote_unique_table_size(mc_bins, vc_bins);       // power of two >= 2 * bins, <= the static table
    BFF_REQUIRE(table_size <= bff::kStaticTableWords, "bff_vote_unique_static: table size %lld beyond the static table", (long long)table_size);
    uint32_t *table = nullptr;
    const int rc = bff::static_table_address(&table);
    if (rc != BFF_OK) return rc;
    // the table keeps the last call's contents: the slots this call probes start empty (0xffffffff), the count at zero
    BFF_CUDA(cudaMemsetAsync(table, 0xff, (size_t)table_size * 4, stream));
    BFF_CUDA(cudaMemsetAsync(count, 0, 4, stream));
    bff::vote_unique_insert_kernel<<<(unsigned)ceil_div64(bins, 256), 256, 0, stream>>>(hist, mc_bins, vc_bins, mode, table, (uint32_t)(table_size - 1),
                                                                                          32 - (int)(63 - __builtin_clzll((unsigned long long)table_size)));
    bff::vote_unique_compact_kernel<<<(unsigned)ceil_div64(table_size, 256), 256, 0, stream>>>(table, (uint32_t)table_size, values, count);
    BFF_LAUNCH_CHECK("vote_unique kernels (static table)");
    return BFF_OK;
}

extern "C" int bff_point_keep(const int32_t *mc, const int32_t *vc, int64_t n, int32_t mode, float thr,
                              uint32_t *keep, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n >= 0 && mode >= 0 && mode <= 2, "bff_point_keep: bad arguments");
    if (n == 0) return BFF_OK;
    BFF_REQUIRE(mc && keep && (mode != 0 || vc), "bff_point_keep: null pointer");
    const long long words = ceil_div64(n, 32);
    point_keep_kernel<<<(unsigned)ceil_div64(words * 32, 256), 256, 0, stream>>>(mc, vc, n, words, mode, thr, keep);
    BFF_LAUNCH_CHECK("point_keep_kernel");
    return BFF_OK;
}


// ---------------------------------------------------------------------------------------------------
// bff_labels_to_rows: per-point instance labels -> one packed row per listed instance id. The evaluator
// (evaluation/eval/scannetv2_inst_eval.py:348-359) tests `gts == gt_inst["instance_id"]` once per
// (prediction, ground-truth instance) pair; here every instance becomes a packed row once, and all
// pred x gt intersections are one bff_mask_iou call.
namespace bff {
constexpr int kLabelIdsPerPass = 1024;
__global__ void __launch_bounds__(256)
labels_to_rows_kernel(const long long *__restrict__ labels, long long n, const long long *__restrict__ ids, int n_ids,
                      long long words, uint32_t *__restrict__ rows) {
    __shared__ long long s_ids[kLabelIdsPerPass];
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    const long long lab = i < n ? __ldg(labels + i) : 0;
    const bool valid = i < n;
    for (int base = 0; base < n_ids; base += kLabelIdsPerPass) {
        const int cnt = min(kLabelIdsPerPass, n_ids - base);
        __syncthreads();
        for (int k = threadIdx.x; k < cnt; k += 256) s_ids[k] = __ldg(ids + base + k);
        __syncthreads();
        for (int k = 0; k < cnt; ++k) {
            const unsigned w = __ballot_sync(0xffffffffu, valid && lab == s_ids[k]);
            if ((threadIdx.x & 31) == 0 && valid) rows[(long long)(base + k) * words + (i >> 5)] = w;
        }
    }
}
}  // namespace bff

extern "C" int bff_labels_to_rows(const int64_t *labels, int64_t n, const int64_t *ids, int32_t n_ids, uint32_t *rows, void *stream_) {
    BFF_REQUIRE(n >= 0 && n_ids >= 0, "bff_labels_to_rows: negative size");
    if (n == 0 || n_ids == 0) return BFF_OK;
    BFF_REQUIRE(labels && ids && rows, "bff_labels_to_rows: null pointer");
    bff::labels_to_rows_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, (cudaStream_t)stream_>>>(
        (const long long *)labels, n, (const long long *)ids, n_ids, ceil_div64(n, 32), rows);
    BFF_LAUNCH_CHECK("labels_to_rows_kernel");
    return BFF_OK;
}
