// Packed point-mask kernels: all-pairs / rectangular intersection + IoU, segmented OR,
// sequential overlap resolution, AND + popcount, bit pack / unpack, RLE -> bits.
//
// A packed row is Wd = ceil(N/32) uint32 words; every kernel here is integer work on those
// words (AND / OR / POPC), bit-exact by construction. Reference lines replaced are cited at
// each entry point and in include/bff.h.
#include <algorithm>
#include <stdlib.h>

#include <atomic>

#include "common.cuh"
#include "iou_umma.cuh"

namespace bff {

// Row counts from which the tcgen05 (int8 tensor core) intersection beats the POPC kernel: the tensor path
// works on 128 x 128 row tiles, so small operands (the 200 x 7 refinement calls) stay on the CUDA cores.
// bff_set_iou_path (or BFF_IOU_PATH=skinny|popc|umma, read once) forces one path: the tests cross-check them.
static std::atomic<int> g_iou_path{-1};        // 0 automatic, 1 few-column kernel, 2 POPC tiles, 3 tcgen05
static int iou_path() {
    int p = g_iou_path.load(std::memory_order_relaxed);
    if (p < 0) {
        const char *e = getenv("BFF_IOU_PATH");
        p = !e ? 0 : e[0] == 's' ? 1 : e[0] == 'p' ? 2 : e[0] == 'u' ? 3 : 0;
        g_iou_path.store(p, std::memory_order_relaxed);
    }
    return p;
}
static int iou_use_umma(int n_a, int n_b, long long words) {
    const int p = iou_path();
    if (p == 2) return 0;
    if (p == 3) return 1;
    return n_a >= 96 && n_b >= 96 && words >= 64;
}

// ------------------------------------------------------------------ row popcounts
__global__ void __launch_bounds__(256) row_popcount_kernel(const uint32_t *__restrict__ rows, long long words,
                                                          int32_t *__restrict__ pop) {
    const uint32_t *r = rows + (long long)blockIdx.x * words;
    int s = 0;
    for (long long w = threadIdx.x; w < words; w += 256) s += __popc(__ldg(r + w));
    __shared__ int red[8];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < 8; ++i) t += red[i];
        pop[blockIdx.x] = t;
    }
}

// ------------------------------------------------------------------ intersections
// CTA tile: 64 rows of A x 64 rows of B, one K-slice of the words; 256 threads, each owning a
// 4 x 4 block of pairs (rows ti + 16*i of the A tile, tj + 16*j of the B tile, so that the 16
// tj-lanes of a warp hit 16 distinct banks and the two ti values broadcast).
constexpr int kTile = 64;
constexpr int kKW = 32;   // words staged per step

__global__ void __launch_bounds__(256)
intersect_kernel(const uint32_t *__restrict__ a, int n_a, const uint32_t *__restrict__ b, int n_b,
                 long long words, long long words_per_slice, int symmetric, int32_t *__restrict__ inter) {
    __shared__ uint32_t sa[kTile][kKW + 1];
    __shared__ uint32_t sb[kTile][kKW + 1];
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (symmetric && bj < bi) return;   // lower triangle is mirrored by the upper one
    const int tid = threadIdx.x;
    const int ti = tid >> 4, tj = tid & 15;
    const long long w_begin = (long long)blockIdx.z * words_per_slice;
    const long long w_end = min(words, w_begin + words_per_slice);
    int acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0;

    for (long long w0 = w_begin; w0 < w_end; w0 += kKW) {
        // stage 64 x 32 words of each operand: thread -> (row = tid/32 + 8*r, word = tid%32)
        const int lw = tid & 31;
        const bool w_ok = (w0 + lw) < w_end;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int row = (tid >> 5) + 8 * r;
            const int ga = bi * kTile + row, gb = bj * kTile + row;
            sa[row][lw] = (w_ok && ga < n_a) ? __ldg(a + (long long)ga * words + w0 + lw) : 0u;
            sb[row][lw] = (w_ok && gb < n_b) ? __ldg(b + (long long)gb * words + w0 + lw) : 0u;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < kKW; ++k) {
            uint32_t va[4], vb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) va[i] = sa[ti + 16 * i][k];
#pragma unroll
            for (int j = 0; j < 4; ++j) vb[j] = sb[tj + 16 * j][k];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] += __popc(va[i] & vb[j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ga = bi * kTile + ti + 16 * i;
        if (ga >= n_a) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gb = bj * kTile + tj + 16 * j;
            if (gb >= n_b || acc[i][j] == 0) continue;
            atomicAdd(inter + (long long)ga * n_b + gb, acc[i][j]);
            if (symmetric && bj != bi) atomicAdd(inter + (long long)gb * n_b + ga, acc[i][j]);
        }
    }
}

// ------------------------------------------------------------------ intersections, few columns
// The refinement calls are (200 stage-1 rows) x (a handful of stage-2 rows) and (few) x (few): a 64 x 64 tile would compute
// mostly padding. Here a warp owns ONE row of A; the B rows of the current 256-word chunk sit in shared memory and every
// lane keeps NB running sums, so A streams through exactly once. The row popcounts of both operands fall out of the same
// pass (no separate row_popcount launches): pop_a from the A words a warp reads anyway, pop_b from the staged B words in the
// first row block. inter / pop_a / pop_b must be zeroed; partial sums of the K-slices are added atomically.
constexpr int kSkinnyChunk = 256;      // words of every B row staged per step
constexpr int kSkinnyRows = 8;         // A rows (= warps) per CTA

template <int NB>
__global__ void __launch_bounds__(kSkinnyRows * 32)
intersect_skinny_kernel(const uint32_t *__restrict__ a, int n_a, const uint32_t *__restrict__ b, int n_b, long long words,
                        long long words_per_slice, int32_t *__restrict__ inter, int32_t *__restrict__ pop_a, int32_t *__restrict__ pop_b) {
    __shared__ uint32_t sb[NB][kSkinnyChunk];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * kSkinnyRows + warp;
    const bool row_ok = row < n_a;
    const uint32_t *ar = a + (long long)(row_ok ? row : 0) * words;
    const long long w_begin = (long long)blockIdx.y * words_per_slice;
    const long long w_end = min(words, w_begin + words_per_slice);
    int acc[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) acc[j] = 0;
    int acc_a = 0, acc_b = 0;                       // acc_b: thread t < NB * 8 ... see below (B popcount of row t / 8)
    for (long long w0 = w_begin; w0 < w_end; w0 += kSkinnyChunk) {
        const int cw = (int)min((long long)kSkinnyChunk, w_end - w0);
        __syncthreads();
        for (int i = threadIdx.x; i < NB * kSkinnyChunk; i += kSkinnyRows * 32) {
            const int j = i / kSkinnyChunk, k = i % kSkinnyChunk;
            const uint32_t v = (j < n_b && k < cw) ? __ldg(b + (long long)j * words + w0 + k) : 0u;
            sb[j][k] = v;
        }
        __syncthreads();
        if (pop_b != nullptr && blockIdx.x == 0) {   // B popcounts: once per K-slice, by the first row block
            for (int i = threadIdx.x; i < NB * kSkinnyChunk; i += kSkinnyRows * 32) {
                const int j = i / kSkinnyChunk;      // 256 consecutive i share j: a warp's 32 values belong to one row
                int c = __popc(sb[j][i % kSkinnyChunk]);
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                if (lane == 0 && c && j < n_b) atomicAdd(pop_b + j, c);
            }
        }
        if (row_ok) {
#pragma unroll 2
            for (int k = lane; k < cw; k += 32) {
                const uint32_t va = __ldg(ar + w0 + k);
                acc_a += __popc(va);
#pragma unroll
                for (int j = 0; j < NB; ++j) acc[j] += __popc(va & sb[j][k]);
            }
        }
    }
    (void)acc_b;
    if (!row_ok) return;
    for (int o = 16; o > 0; o >>= 1) acc_a += __shfl_xor_sync(0xffffffffu, acc_a, o);
    if (lane == 0 && pop_a != nullptr && acc_a) atomicAdd(pop_a + row, acc_a);
#pragma unroll
    for (int j = 0; j < NB; ++j) {
        int v = acc[j];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && j < n_b && v) atomicAdd(inter + (long long)row * n_b + j, v);
    }
}

// IoU exactly as calculate_iou (projection_2d_to_3d.py:158-165) / calculate_iou_between_stages
// (refinement.py:82-90): union = (f32(sA) + f32(sB)) - f32(I); iou = f32(I) / union; 0/0 = NaN.
__global__ void iou_finalize_kernel(const int32_t *__restrict__ inter, const int32_t *__restrict__ pa,
                                    const int32_t *__restrict__ pb, int n_a, int n_b, float *__restrict__ iou) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_a * n_b) return;
    const int r = (int)(i / n_b), c = (int)(i % n_b);
    const float fi = __int2float_rn(inter[i]);
    const float un = __fsub_rn(__fadd_rn(__int2float_rn(pa[r]), __int2float_rn(pb[c])), fi);
    iou[i] = __fdiv_rn(fi, un);
}

// ------------------------------------------------------------------ segmented OR
__global__ void __launch_bounds__(256)
rows_or_reduce_kernel(const uint32_t *__restrict__ rows, long long words,
                      const int32_t *__restrict__ idx, const int32_t *__restrict__ seg,
                      uint32_t *__restrict__ out) {
    // the group's member list goes through shared memory so that the row loads do not wait for an index load each,
    // and eight of them are in flight per thread
    __shared__ int32_t s_idx[256];
    const int g = blockIdx.y;
    const long long w = (long long)blockIdx.x * 256 + threadIdx.x;
    const bool live = w < words;
    const int k0 = seg[g], k1 = seg[g + 1];
    uint32_t v = 0;
    for (int base = k0; base < k1; base += 256) {
        const int cnt = min(256, k1 - base);
        __syncthreads();
        if ((int)threadIdx.x < cnt) s_idx[threadIdx.x] = __ldg(idx + base + threadIdx.x);
        __syncthreads();
        if (live) {
            int k = 0;
            for (; k + 8 <= cnt; k += 8) {
                uint32_t t[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) t[u] = __ldg(rows + (long long)s_idx[k + u] * words + w);
#pragma unroll
                for (int u = 0; u < 8; ++u) v |= t[u];
            }
            for (; k < cnt; ++k) v |= __ldg(rows + (long long)s_idx[k] * words + w);
        }
    }
    if (live) out[(long long)g * words + w] = v;
}

// ------------------------------------------------------------------ overlap resolution
// Every thread owns one word column and replays the host-built program rows[dst] &= ~rows[src]
// in order (projection_2d_to_3d.py:295-299). Columns are independent because the update is
// element-wise. Rows are cached in shared memory when they fit.
template <bool OR_INTO>
__device__ __forceinline__ uint32_t row_step(uint32_t dst, uint32_t src) { return OR_INTO ? (dst | src) : (dst & ~src); }

template <bool OR_INTO>
__global__ void __launch_bounds__(128)
row_program_kernel(uint32_t *__restrict__ rows, int n_rows, long long words,
                   const int32_t *__restrict__ prog, int n_steps, int use_smem) {
    extern __shared__ uint32_t s_col[];   // n_rows x 128
    const long long w = (long long)blockIdx.x * 128 + threadIdx.x;
    if (w >= words) return;
    if (use_smem) {
        for (int r = 0; r < n_rows; ++r) s_col[r * 128 + threadIdx.x] = rows[(long long)r * words + w];
        for (int k = 0; k < n_steps; ++k) {
            const int dst = __ldg(prog + 2 * k), src = __ldg(prog + 2 * k + 1);
            s_col[dst * 128 + threadIdx.x] = row_step<OR_INTO>(s_col[dst * 128 + threadIdx.x], s_col[src * 128 + threadIdx.x]);
        }
        for (int r = 0; r < n_rows; ++r) rows[(long long)r * words + w] = s_col[r * 128 + threadIdx.x];
    } else {
        for (int k = 0; k < n_steps; ++k) {
            const int dst = __ldg(prog + 2 * k), src = __ldg(prog + 2 * k + 1);
            rows[(long long)dst * words + w] = row_step<OR_INTO>(rows[(long long)dst * words + w], rows[(long long)src * words + w]);
        }
    }
}

// ------------------------------------------------------------------ AND keep + popcount
constexpr int kAndPopWordsPerBlock = 2048;        // a row is cut into pieces so that a handful of rows still fills the GPU
__global__ void __launch_bounds__(256)
rows_and_popcount_kernel(uint32_t *__restrict__ rows, long long words, const uint32_t *__restrict__ keep,
                         int32_t *__restrict__ pop) {
    uint32_t *r = rows + (long long)blockIdx.y * words;
    const long long w0 = (long long)blockIdx.x * kAndPopWordsPerBlock;
    const long long w1 = min(words, w0 + kAndPopWordsPerBlock);
    int s = 0;
    for (long long w = w0 + threadIdx.x; w < w1; w += 256) {
        uint32_t v = r[w];
        if (keep != nullptr) {
            v &= __ldg(keep + w);
            r[w] = v;
        }
        s += __popc(v);
    }
    __shared__ int red[8];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0 && pop != nullptr) {
        int t = 0;
        for (int i = 0; i < 8; ++i) t += red[i];
        if (t) atomicAdd(pop + blockIdx.y, t);          // pop is zeroed by the host wrapper
    }
}

// ------------------------------------------------------------------ pack / unpack
__global__ void __launch_bounds__(256)
pack_bits_kernel(const uint8_t *__restrict__ dense, long long n, long long words, uint32_t *__restrict__ packed) {
    const int row = blockIdx.y;
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;   // element index (warp = one word)
    const bool bit = (i < n) && (__ldg(dense + (long long)row * n + i) != 0);
    const unsigned w = __ballot_sync(0xffffffffu, bit);
    if ((threadIdx.x & 31) == 0 && (i >> 5) < words) packed[(long long)row * words + (i >> 5)] = w;
}

__global__ void __launch_bounds__(256)
unpack_bits_kernel(const uint32_t *__restrict__ packed, long long n, long long words, uint8_t *__restrict__ dense) {
    const int row = blockIdx.y;
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    dense[(long long)row * n + i] = (uint8_t)((__ldg(packed + (long long)row * words + (i >> 5)) >> (i & 31)) & 1u);
}

// The (rows, n) byte matrix as ONE flat array: a thread writes 16 consecutive bytes with one store. A chunk inside one
// row takes its 16 bits from at most two packed words; the few chunks that straddle a row end go byte by byte.
__device__ __forceinline__ uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

__global__ void __launch_bounds__(256)
unpack_bits_wide_kernel(const uint32_t *__restrict__ packed, long long n, long long words, long long total, uint4 *__restrict__ dense16) {
    const long long j0 = ((long long)blockIdx.x * 256 + threadIdx.x) * 16;
    if (j0 >= total) return;
    const long long row = j0 / n, i0 = j0 - row * n;
    if (i0 + 16 <= n && j0 + 16 <= total) {
        const uint32_t *src = packed + row * words + (i0 >> 5);
        const int sh = (int)(i0 & 31);
        const uint32_t lo = __ldg(src);
        const uint32_t hi = sh > 16 ? __ldg(src + 1) : 0u;      // sh > 16: bits sh..sh+15 reach into the next word (i0 + 16 <= n keeps it inside the row)
        const uint32_t v = __funnelshift_r(lo, hi, sh) & 0xffffu;
        dense16[j0 >> 4] = make_uint4(spread4(v & 15u), spread4((v >> 4) & 15u), spread4((v >> 8) & 15u), spread4(v >> 12));
        return;
    }
    uint8_t *out = reinterpret_cast<uint8_t *>(dense16);
    for (long long j = j0; j < j0 + 16 && j < total; ++j) {
        const long long r = j / n, i = j - r * n;
        out[j] = (uint8_t)((__ldg(packed + r * words + (i >> 5)) >> (i & 31)) & 1u);
    }
}

// ------------------------------------------------------------------ RLE -> bits
// One warp per run [lo, hi) of row r: lanes stride over the touched words and OR their part in.
__global__ void __launch_bounds__(256)
rle_to_bits_kernel(const int32_t *__restrict__ runs, long long n_runs, long long words, uint32_t *__restrict__ packed) {
    const long long run = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5;
    if (run >= n_runs) return;
    const int lane = threadIdx.x & 31;
    const int r = __ldg(runs + 3 * run);
    const long long lo = __ldg(runs + 3 * run + 1), hi = __ldg(runs + 3 * run + 2);
    if (hi <= lo) return;
    const long long w_first = lo >> 5, w_last = (hi - 1) >> 5;
    uint32_t *row = packed + (long long)r * words;
    for (long long w = w_first + lane; w <= w_last; w += 32) {
        uint32_t m = 0xffffffffu;
        if (w == w_first) m &= 0xffffffffu << (lo & 31);
        if (w == w_last) m &= 0xffffffffu >> (31 - ((hi - 1) & 31));
        atomicOr(row + w, m);
    }
}

// RLE -> bits straight from the wire format (rle_encode_decode.py:35-61, refinement.py:26-39): counts holds the masks'
// (1-based start, run length) int64 pairs back to back, pair_offset[r] .. pair_offset[r + 1] are row r's pairs. One warp
// per pair: the reference's arithmetic -- both numbers cast to int32, start made 0-based, end = int32(start + length),
// then Python's slice rule `mask[lo:hi] = 1` (negative bounds count from the end, everything clipped to [0, L]) -- and
// the touched words are OR-ed in. The row of a pair is found by bisection over pair_offset.
__global__ void __launch_bounds__(256)
rle_pairs_to_bits_kernel(const long long *__restrict__ counts, const long long *__restrict__ pair_offset, int n_rows, long long n_pairs,
                         long long length, long long words, uint32_t *__restrict__ packed) {
    __shared__ int s_row0;
    const long long pair0 = ((long long)blockIdx.x * 256) >> 5;          // first pair of this CTA (8 consecutive pairs, one per warp)
    if (pair0 >= n_pairs) return;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x < 32) {
        // row of the CTA's first pair = largest r with pair_offset[r] <= pair0, by a 32-ary search (two rounds up to 1024
        // rows): the lanes probe evenly spaced offsets at once instead of ten dependent loads of a bisection per warp
        int lo_r = 0, hi_r = n_rows;
        while (hi_r - lo_r > 1) {
            const int step = (hi_r - lo_r + 31) / 32;
            const int idx = lo_r + (lane + 1) * step;
            const bool le = idx < hi_r && __ldg(pair_offset + idx) <= pair0;
            const int cnt = __popc(__ballot_sync(0xffffffffu, le));         // monotone: the probes <= pair0 are a prefix of the lanes
            const int new_lo = lo_r + cnt * step;
            hi_r = min(hi_r, new_lo + step);
            lo_r = new_lo;
        }
        if (lane == 0) s_row0 = lo_r;
    }
    __syncthreads();
    const long long pair = pair0 + (threadIdx.x >> 5);
    if (pair >= n_pairs) return;
    int lo_r = s_row0;
    while (lo_r + 1 < n_rows && __ldg(pair_offset + lo_r + 1) <= pair) ++lo_r;   // the CTA's pairs span a row or two
    const long long start = (long long)(int)__ldg(counts + 2 * pair) - 1;
    const long long num = (long long)(int)__ldg(counts + 2 * pair + 1);
    const long long end = (long long)(int)(start + num);
    long long lo = start < 0 ? max(start + length, 0LL) : min(start, length);
    long long hi = end < 0 ? max(end + length, 0LL) : min(end, length);
    if (hi <= lo) return;
    const long long w_first = lo >> 5, w_last = (hi - 1) >> 5;
    uint32_t *row = packed + (long long)lo_r * words;
    for (long long w = w_first + lane; w <= w_last; w += 32) {
        uint32_t m = 0xffffffffu;
        if (w == w_first) m &= 0xffffffffu << (lo & 31);
        if (w == w_last) m &= 0xffffffffu >> (31 - ((hi - 1) & 31));
        atomicOr(row + w, m);
    }
}

}  // namespace bff

using namespace bff;

extern "C" int bff_rle_pairs_to_bits(const int64_t *counts, const int64_t *pair_offset, int32_t n_rows, int64_t n_pairs, int64_t length,
                                     uint32_t *packed, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_rows >= 0 && n_pairs >= 0 && length >= 0, "bff_rle_pairs_to_bits: negative size");
    const long long words = ceil_div64(length, 32);
    if (n_rows == 0 || words == 0) return BFF_OK;
    BFF_REQUIRE(packed, "bff_rle_pairs_to_bits: null output");
    BFF_CUDA(cudaMemsetAsync(packed, 0, (size_t)n_rows * words * 4, stream));
    if (n_pairs == 0) return BFF_OK;
    BFF_REQUIRE(counts && pair_offset, "bff_rle_pairs_to_bits: null input");
    BFF_REQUIRE(ceil_div64(n_pairs * 32, 256) <= 0x7fffffffLL, "bff_rle_pairs_to_bits: too many runs for one launch");
    rle_pairs_to_bits_kernel<<<(unsigned)ceil_div64(n_pairs * 32, 256), 256, 0, stream>>>((const long long *)counts, (const long long *)pair_offset,
                                                                                           n_rows, n_pairs, length, words, packed);
    BFF_LAUNCH_CHECK("rle_pairs_to_bits_kernel");
    return BFF_OK;
}

extern "C" int bff_set_iou_path(int32_t path) {
    BFF_REQUIRE(path >= 0 && path <= 3, "bff_set_iou_path: 0 automatic, 1 few-column, 2 POPC tiles, 3 tcgen05");
    g_iou_path.store(path, std::memory_order_relaxed);
    return BFF_OK;
}

extern "C" int bff_mask_iou(const uint32_t *a, int32_t n_a, const uint32_t *b, int32_t n_b, int64_t words,
                            int32_t *inter, int32_t *pop_a, int32_t *pop_b, float *iou, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_a >= 0 && n_b >= 0 && words >= 0, "bff_mask_iou: negative size");
    if (n_a == 0 || n_b == 0) return BFF_OK;
    BFF_REQUIRE(a && b, "bff_mask_iou: null rows");
    BFF_REQUIRE(!iou || (inter && pop_a && pop_b), "bff_mask_iou: iou needs inter, pop_a and pop_b buffers");
    // few columns (the refinement shapes): one pass that also yields both popcount vectors
    if (inter && words > 0 && n_b <= 32 && iou_path() <= 1) {
        const bool same_pop = pop_a != nullptr && pop_a == pop_b;           // symmetric call: one vector serves both
        BFF_CUDA(cudaMemsetAsync(inter, 0, (size_t)n_a * n_b * 4, stream));
        if (pop_a) BFF_CUDA(cudaMemsetAsync(pop_a, 0, (size_t)n_a * 4, stream));
        if (pop_b && !same_pop) BFF_CUDA(cudaMemsetAsync(pop_b, 0, (size_t)n_b * 4, stream));
        const long long row_blocks = ceil_div64(n_a, kSkinnyRows);
        long long slices = ceil_div64((long long)kNumSMs * 4, row_blocks);
        const long long max_slices = ceil_div64(words, kSkinnyChunk);
        if (slices > max_slices) slices = max_slices;
        if (slices < 1) slices = 1;
        if (slices > 65535) slices = 65535;
        const long long wps = ceil_div64(ceil_div64(words, slices), kSkinnyChunk) * kSkinnyChunk;
        slices = ceil_div64(words, wps);
        dim3 grid((unsigned)row_blocks, (unsigned)slices);
        int32_t *pb = same_pop ? nullptr : pop_b;
        if (n_b <= 8) intersect_skinny_kernel<8><<<grid, kSkinnyRows * 32, 0, stream>>>(a, n_a, b, n_b, words, wps, inter, pop_a, pb);
        else if (n_b <= 16) intersect_skinny_kernel<16><<<grid, kSkinnyRows * 32, 0, stream>>>(a, n_a, b, n_b, words, wps, inter, pop_a, pb);
        else intersect_skinny_kernel<32><<<grid, kSkinnyRows * 32, 0, stream>>>(a, n_a, b, n_b, words, wps, inter, pop_a, pb);
        if (iou) {
            const long long tot = (long long)n_a * n_b;
            iou_finalize_kernel<<<(unsigned)ceil_div64(tot, 256), 256, 0, stream>>>(inter, pop_a, pop_b, n_a, n_b, iou);
        }
        BFF_LAUNCH_CHECK("bff_mask_iou kernels (few columns)");
        return BFF_OK;
    }
    const int symmetric = (a == b && n_a == n_b) ? 1 : 0;
    // the tensor-core kernel counts the rows' bits while it expands them: no separate popcount launches on that path
    const bool umma = inter && words > 0 && iou_use_umma(n_a, n_b, words);
    BFF_REQUIRE(symmetric || pop_a == nullptr || pop_a != pop_b, "bff_mask_iou: one popcount buffer for two different operands");
    int32_t *upa = nullptr, *upb = nullptr;             // what the tensor-core kernel accumulates into
    if (!umma) {
        if (pop_a) {
            if (words > 0) row_popcount_kernel<<<n_a, 256, 0, stream>>>(a, words, pop_a);
            else BFF_CUDA(cudaMemsetAsync(pop_a, 0, (size_t)n_a * 4, stream));
        }
        if (pop_b && pop_b != pop_a) {
            if (words > 0) row_popcount_kernel<<<n_b, 256, 0, stream>>>(b, words, pop_b);
            else BFF_CUDA(cudaMemsetAsync(pop_b, 0, (size_t)n_b * 4, stream));
        }
    } else {
        upa = pop_a ? pop_a : (symmetric ? pop_b : nullptr);
        upb = symmetric ? nullptr : pop_b;
        if (upa) BFF_CUDA(cudaMemsetAsync(upa, 0, (size_t)n_a * 4, stream));
        if (upb) BFF_CUDA(cudaMemsetAsync(upb, 0, (size_t)n_b * 4, stream));
    }
    if (inter) {
        BFF_CUDA(cudaMemsetAsync(inter, 0, (size_t)n_a * n_b * 4, stream));
        if (words > 0) {
            const int ta = (n_a + kTile - 1) / kTile, tb = (n_b + kTile - 1) / kTile;
            // K-slices: enough CTAs to fill the machine a few times, slices a multiple of kKW words
            long long tiles = (long long)ta * tb;
            if (symmetric) tiles = (long long)ta * (ta + 1) / 2;
            long long slices = ceil_div64((long long)kNumSMs * 8, tiles);
            const long long max_slices = ceil_div64(words, kKW);
            if (slices > max_slices) slices = max_slices;
            if (slices < 1) slices = 1;
            if (slices > 65535) slices = 65535;
            long long wps = ceil_div64(ceil_div64(words, slices), kKW) * kKW;
            slices = ceil_div64(words, wps);
            if (umma) {
                BFF_CUDA(umma::launch_intersect(a, n_a, b, n_b, words, symmetric, inter, upa, upb, kNumSMs, stream));
                if (symmetric && pop_a && pop_b && pop_b != pop_a)
                    BFF_CUDA(cudaMemcpyAsync(pop_b, pop_a, (size_t)n_a * 4, cudaMemcpyDeviceToDevice, stream));
            } else {
                dim3 grid((unsigned)tb, (unsigned)ta, (unsigned)slices);
                intersect_kernel<<<grid, 256, 0, stream>>>(a, n_a, b, n_b, words, wps, symmetric, inter);
            }
        }
    }
    if (iou) {
        const long long tot = (long long)n_a * n_b;
        iou_finalize_kernel<<<(unsigned)ceil_div64(tot, 256), 256, 0, stream>>>(inter, pop_a, pop_b, n_a, n_b, iou);
    }
    BFF_LAUNCH_CHECK("bff_mask_iou kernels");
    return BFF_OK;
}

extern "C" int bff_rows_or_reduce(const uint32_t *rows, int64_t words, const int32_t *idx, const int32_t *seg,
                                  int32_t n_groups, uint32_t *out, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_groups >= 0 && words >= 0, "bff_rows_or_reduce: negative size");
    if (n_groups == 0 || words == 0) return BFF_OK;
    BFF_REQUIRE(rows && idx && seg && out, "bff_rows_or_reduce: null pointer");
    for (int32_t g0 = 0; g0 < n_groups; g0 += kMaxGridY) {      // blockIdx.y = group: at most 65535 per launch
        dim3 grid((unsigned)ceil_div64(words, 256), (unsigned)std::min(n_groups - g0, kMaxGridY));
        rows_or_reduce_kernel<<<grid, 256, 0, stream>>>(rows, words, idx, seg + g0, out + (size_t)g0 * words);
        BFF_LAUNCH_CHECK("rows_or_reduce_kernel");
    }
    return BFF_OK;
}

template <bool OR_INTO>
static int run_row_program(const char *name, uint32_t *rows, int32_t n_rows, int64_t words, const int32_t *prog,
                           int32_t n_steps, cudaStream_t stream) {
    BFF_REQUIRE(n_rows >= 0 && words >= 0 && n_steps >= 0, "%s: negative size", name);
    if (n_rows == 0 || words == 0 || n_steps == 0) return BFF_OK;
    BFF_REQUIRE(rows && prog, "%s: null pointer", name);
    const size_t smem = (size_t)n_rows * 128 * 4;
    // the shared-memory column cache pays off when the program revisits rows; it must also fit
    const int use_smem = smem <= 160 * 1024 && n_steps >= n_rows;
    if (use_smem)
        BFF_CUDA(cudaFuncSetAttribute((const void *)row_program_kernel<OR_INTO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    row_program_kernel<OR_INTO><<<(unsigned)ceil_div64(words, 128), 128, use_smem ? smem : 0, stream>>>(rows, n_rows, words, prog, n_steps, use_smem);
    BFF_LAUNCH_CHECK(name);
    return BFF_OK;
}

extern "C" int bff_overlap_resolve(uint32_t *rows, int32_t n_rows, int64_t words, const int32_t *prog,
                                   int32_t n_steps, void *stream_) {
    return run_row_program<false>("bff_overlap_resolve", rows, n_rows, words, prog, n_steps, (cudaStream_t)stream_);
}

extern "C" int bff_rows_or_into(uint32_t *rows, int32_t n_rows, int64_t words, const int32_t *prog,
                                int32_t n_steps, void *stream_) {
    return run_row_program<true>("bff_rows_or_into", rows, n_rows, words, prog, n_steps, (cudaStream_t)stream_);
}

extern "C" int bff_rows_and_popcount(uint32_t *rows, int32_t n_rows, int64_t words, const uint32_t *keep,
                                     int32_t *pop, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_rows >= 0 && words >= 0, "bff_rows_and_popcount: negative size");
    if (n_rows == 0) return BFF_OK;
    if (words == 0) {
        if (pop) BFF_CUDA(cudaMemsetAsync(pop, 0, (size_t)n_rows * 4, stream));
        return BFF_OK;
    }
    BFF_REQUIRE(rows, "bff_rows_and_popcount: null rows");
    if (pop) BFF_CUDA(cudaMemsetAsync(pop, 0, (size_t)n_rows * 4, stream));
    for (int32_t r0 = 0; r0 < n_rows; r0 += kMaxGridY) {
        dim3 grid((unsigned)ceil_div64(words, kAndPopWordsPerBlock), (unsigned)std::min(n_rows - r0, kMaxGridY));
        rows_and_popcount_kernel<<<grid, 256, 0, stream>>>(rows + (size_t)r0 * words, words, keep, pop ? pop + r0 : nullptr);
        BFF_LAUNCH_CHECK("rows_and_popcount_kernel");
    }
    return BFF_OK;
}

extern "C" int bff_pack_bits(const uint8_t *dense, int32_t n_rows, int64_t n, uint32_t *packed, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_rows >= 0 && n >= 0, "bff_pack_bits: negative size");
    if (n_rows == 0 || n == 0) return BFF_OK;
    BFF_REQUIRE(dense && packed, "bff_pack_bits: null pointer");
    const long long words = ceil_div64(n, 32);
    for (int32_t r0 = 0; r0 < n_rows; r0 += kMaxGridY) {
        dim3 grid((unsigned)ceil_div64(words * 32, 256), (unsigned)std::min(n_rows - r0, kMaxGridY));
        pack_bits_kernel<<<grid, 256, 0, stream>>>(dense + (size_t)r0 * n, n, words, packed + (size_t)r0 * words);
        BFF_LAUNCH_CHECK("pack_bits_kernel");
    }
    return BFF_OK;
}

extern "C" int bff_unpack_bits(const uint32_t *packed, int32_t n_rows, int64_t n, uint8_t *dense, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_rows >= 0 && n >= 0, "bff_unpack_bits: negative size");
    if (n_rows == 0 || n == 0) return BFF_OK;
    BFF_REQUIRE(dense && packed, "bff_unpack_bits: null pointer");
    const long long words = ceil_div64(n, 32);
    const long long total = (long long)n_rows * n;
    if ((reinterpret_cast<uintptr_t>(dense) & 15) == 0 && n >= 16 && ceil_div64(total, 16 * 256) <= 0x7fffffffLL) {
        unpack_bits_wide_kernel<<<(unsigned)ceil_div64(total, 16 * 256), 256, 0, stream>>>(packed, n, words, total, reinterpret_cast<uint4 *>(dense));
    } else {
        for (int32_t r0 = 0; r0 < n_rows; r0 += kMaxGridY) {
            dim3 grid((unsigned)ceil_div64(n, 256), (unsigned)std::min(n_rows - r0, kMaxGridY));
            unpack_bits_kernel<<<grid, 256, 0, stream>>>(packed + (size_t)r0 * words, n, words, dense + (size_t)r0 * n);
        }
    }
    BFF_LAUNCH_CHECK("unpack_bits_kernel");
    return BFF_OK;
}

extern "C" int bff_rle_to_bits(const int32_t *runs, int64_t n_runs, int32_t n_rows, int64_t words,
                               uint32_t *packed, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    BFF_REQUIRE(n_runs >= 0 && n_rows >= 0 && words >= 0, "bff_rle_to_bits: negative size");
    if (n_rows == 0 || words == 0) return BFF_OK;
    BFF_REQUIRE(packed, "bff_rle_to_bits: null output");
    BFF_CUDA(cudaMemsetAsync(packed, 0, (size_t)n_rows * words * 4, stream));
    if (n_runs == 0) return BFF_OK;
    BFF_REQUIRE(runs, "bff_rle_to_bits: null runs");
    rle_to_bits_kernel<<<(unsigned)ceil_div64(n_runs * 32, 256), 256, 0, stream>>>(runs, n_runs, words, packed);
    BFF_LAUNCH_CHECK("rle_to_bits_kernel");
    return BFF_OK;
}


// ---------------------------------------------------------------------------------------------------
// bff_rows_permute: out bit p of every row = in bit src_index[p]. Used to bring rows lifted on a spatially re-ordered
// copy of the cloud (bff_b200/lift.py: Morton order, so that a warp's 32 points are neighbours and the warp-level
// culling of lift_kernel bites) back to the cloud's own point order, which every output of the path is indexed by.
namespace bff {
constexpr int kPermuteRowsPerBlock = 16;
__global__ void __launch_bounds__(256)
rows_permute_kernel(const uint32_t *__restrict__ in, int n_rows, long long n, long long words, const int32_t *__restrict__ src,
                    uint32_t *__restrict__ out) {
    const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
    const bool valid = p < n;
    const int s = valid ? __ldg(src + p) : 0;
    const long long sw = s >> 5;
    const int sb = s & 31;
    const int r0 = blockIdx.y * kPermuteRowsPerBlock, r1 = min(n_rows, r0 + kPermuteRowsPerBlock);
    for (int r = r0; r < r1; ++r) {
        const bool bit = valid && ((__ldg(in + (long long)r * words + sw) >> sb) & 1u);
        const unsigned w = __ballot_sync(0xffffffffu, bit);
        if ((threadIdx.x & 31) == 0 && valid) out[(long long)r * words + (p >> 5)] = w;
    }
}
}  // namespace bff

extern "C" int bff_rows_permute(const uint32_t *rows_in, int32_t n_rows, int64_t n_points, const int32_t *src_index,
                                uint32_t *rows_out, void *stream_) {
    BFF_REQUIRE(n_rows >= 0 && n_points >= 0, "bff_rows_permute: negative size");
    if (n_rows == 0 || n_points == 0) return BFF_OK;
    BFF_REQUIRE(rows_in && src_index && rows_out && rows_in != rows_out, "bff_rows_permute: null or aliased pointer");
    BFF_REQUIRE(n_points < (1LL << 31), "bff_rows_permute: too many points for int32 indices");
    const long long words = ceil_div64(n_points, 32);
    const int32_t rows_per_launch = kMaxGridY * bff::kPermuteRowsPerBlock;
    for (int32_t r0 = 0; r0 < n_rows; r0 += rows_per_launch) {
        const int32_t nr = std::min(n_rows - r0, rows_per_launch);
        dim3 grid((unsigned)ceil_div64(n_points, 256), (unsigned)ceil_div64(nr, bff::kPermuteRowsPerBlock));
        bff::rows_permute_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(rows_in + (size_t)r0 * words, nr, n_points, words, src_index,
                                                                         rows_out + (size_t)r0 * words);
        BFF_LAUNCH_CHECK("rows_permute_kernel");
    }
    return BFF_OK;
}
