// Mask intersections on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM).
//
// inter[i][j] = popcount(A_i & B_j) is the (nA x N) . (N x nB) product of 0/1 matrices. The rows stay bit-packed
// in HBM (N/8 bytes per row); each CTA expands the 128-bit slices it needs to uint8 0/1 values on the fly --
// the B rows into shared memory in the canonical K-major "interleave" (no-swizzle) UMMA layout, the A rows
// straight into tensor memory (tcgen05.st; lane = row, 4 K-bytes per 32-bit column) -- and one thread issues
// 128x128x32 integer MMAs ([d_tmem], [a_tmem], b_desc) while the other threads expand the next slice. The int32 sums are
// exact (<= N < 2^31), so the result is bit-identical to the POPC kernel (north_star: "tensor cores only if an
// int8 mask-product formulation measurably wins" -- it does, see profiles/).
//
// Layout of one operand sub-tile (128 rows x 32 K-bytes, 4 KB) as the descriptor states it:
//   byte(r, k) at (r / 8) * SBO + (k / 16) * LBO + (r % 8) * 16 + (k % 16),   LBO = 128, SBO = 256
// A stage holds 4 such sub-tiles (128 K-bits) of B in shared memory and 32 TMEM columns of A; 4 stages form the
// producer/MMA ring (BFF_UMMA_A_TMEM=0 builds the earlier form with both operands in shared memory, 3 stages).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bff {
namespace umma {

constexpr int kTileM = 128;               // rows of A per CTA
constexpr int kTileN = 128;               // rows of B per CTA (= TMEM columns)
constexpr int kStageWords = 4;            // 32-bit words of every row per stage (= 4 MMAs of K = 32)
constexpr int kSubTileBytes = 128 * 32;   // one operand, one MMA
constexpr int kOperandStageBytes = kStageWords * kSubTileBytes;   // 16 KB
constexpr int kStageBytes = 2 * kOperandStageBytes;               // A + B = 32 KB
#ifndef BFF_UMMA_A_TMEM
#define BFF_UMMA_A_TMEM 1
#endif
// A operand in tensor memory (tcgen05.mma [d], [a_tmem], b_desc): at M = N = 128 an int8 MMA reading BOTH operands from
// shared memory consumes 8 KB per 64 clocks = the whole 128 B/clk of the SM's shared memory, leaving nothing for the
// producers' stores (ncu: tensor pipe 32 % busy, shared-memory wavefronts the limiter). With A expanded straight into
// TMEM by tcgen05.st only B crosses shared memory (64 B/clk read + 64 B/clk written).
constexpr bool kATmem = BFF_UMMA_A_TMEM != 0;
constexpr int kStages = kATmem ? 4 : 3;
constexpr int kSmemStageBytes = kATmem ? kOperandStageBytes : kStageBytes;      // what one ring slot holds in shared memory
constexpr int kSmemNeeded = kStages * kSmemStageBytes + 1024;                    // + alignment slack
// TMEM columns: 128 accumulators (+ 4 slots x 32 columns of A bytes, 4 per column) -> 256 per CTA, two CTAs fill the SM's 512;
// the shared-memory request is padded so that a third CTA (which could only spin in tcgen05.alloc) is never resident
constexpr int kSmemBytes = kATmem ? (kSmemNeeded > 77 * 1024 ? kSmemNeeded : 77 * 1024) : kSmemNeeded;
constexpr int kTmemCols = kATmem ? 256 : 128;
constexpr int kATmemCol0 = 128;           // first column of the A ring
constexpr int kAStageCols = kStageWords * 8;
constexpr int kProducerWarps = 8;
constexpr int kPrefetch = 3;              // stages of packed words held in registers ahead of the expansion
constexpr int kThreads = (kProducerWarps + 1) * 32;               // + the MMA-issuing warp

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout, version 1)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);            // start address, 16-byte units
    d |= (uint64_t)((128u >> 4) & 0x3fff) << 16;       // leading byte offset: next 16-byte K chunk
    d |= (uint64_t)((256u >> 4) & 0x3fff) << 32;       // stride byte offset: next group of 8 rows
    d |= (uint64_t)1 << 46;                            // descriptor version (Blackwell)
    return d;                                          // base offset 0, layout type 0 = no swizzle
}

// instruction descriptor for kind::i8: D = S32, A = B = UINT8, both K-major, M = 128, N = 128
__device__ __forceinline__ uint32_t make_idesc() {
    uint32_t d = 0;
    d |= 2u << 4;                       // c_format: S32
    d |= 0u << 7;                       // a_format: unsigned 8 bit
    d |= 0u << 10;                      // b_format: unsigned 8 bit
    d |= (uint32_t)(kTileN >> 3) << 17; // n_dim
    d |= (uint32_t)(kTileM >> 4) << 24; // m_dim
    return d;
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    }
}

// Expansion of one packed word into 32 bytes of 0/1, four K-bytes at a time: bytes (v >> p) & 0x01010101 hold bits p, p + 8,
// p + 16, p + 24 of the word. That is NOT the natural bit order -- and it does not have to be: the dot product over K is
// invariant under any permutation of K that A and B share, and both operands are expanded by this same function. Two
// instructions per four K-bytes (shift, mask) instead of four (nibble extract, multiply by 0x00204081, mask).
__device__ __forceinline__ uint32_t plane_to_bytes(uint32_t v, int p) { return (v >> p) & 0x01010101u; }

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

// grid: x = tile column (B), y = tile row (A), z = K slice. `inter` must be zeroed; partial sums are added.
// Warps 0..7 (256 threads) are PRODUCERS: thread t expands row t of the stage (0..127 = A -> TMEM lanes,
// 128..255 = B -> shared memory) from packed bits to 0/1 bytes, running up to kStages ahead. Warp 8 is the MMA ISSUER: one lane
// waits for a full stage, issues its 4 MMAs and commits them to the stage's `empty` barrier.
__global__ void __launch_bounds__(kThreads)
intersect_umma_kernel(const uint32_t *__restrict__ a, int n_a, const uint32_t *__restrict__ b, int n_b, long long words,
                      long long words_per_slice, int symmetric, int32_t *__restrict__ inter, int32_t *__restrict__ pop_a,
                      int32_t *__restrict__ pop_b) {
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (symmetric && bj < bi) return;                 // mirrored from the upper triangle
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages], done_bar;
    __shared__ uint32_t tmem_base_slot;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long w_begin = (long long)blockIdx.z * words_per_slice;
    const long long w_end = min(words, w_begin + words_per_slice);
    const int n_stages = (int)((w_end - w_begin + kStageWords - 1) / kStageWords);

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], kProducerWarps); mbar_init(&empty_bar[s], 1); }
        mbar_init(&done_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {                                    // TMEM: 128 columns x 128 lanes of int32 accumulators (+ the A ring)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&tmem_base_slot)), "r"((uint32_t)kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp < kProducerWarps) {
        // ---------------- producers
        const int op = tid >> 7, r = tid & 127;
        const int grow = (op == 0 ? bi * kTileM : bj * kTileN) + r;
        const bool row_ok = grow < (op == 0 ? n_a : n_b);
        const uint32_t *src = (op == 0 ? a : b) + (long long)grow * words;
        const uint32_t row_off = (uint32_t)((r >> 3) * 256 + (r & 7) * 16);
        // Row popcounts ride along: a producer thread sees every word of its row in this K slice anyway. Each row must be
        // counted by exactly one tile column / row: the A rows by the CTAs of the first tile column they meet (bj == 0, or
        // the diagonal tile of a symmetric call), the B rows by the first tile row (bi == 0; symmetric: pop_b == pop_a).
        int32_t *pop_out = nullptr;
        if (op == 0 && pop_a != nullptr && (symmetric ? bj == bi : bj == 0)) pop_out = pop_a;
        if (op == 1 && pop_b != nullptr && !symmetric && bi == 0) pop_out = pop_b;
        int pop_acc = 0;
        // the packed words of the next kPrefetch stages travel in registers: one stage of expansion is shorter than a trip
        // to L2, so a single stage of look-ahead left the producers waiting for their loads
        uint32_t pre[kPrefetch][kStageWords];
#pragma unroll
        for (int u = 0; u < kPrefetch; ++u)
#pragma unroll
            for (int k = 0; k < kStageWords; ++k) {
                const long long w = w_begin + (long long)u * kStageWords + k;
                pre[u][k] = (row_ok && w < w_end) ? __ldg(src + w) : 0u;
            }
        for (int s0 = 0; s0 < n_stages; s0 += kPrefetch) {
#pragma unroll
          for (int u = 0; u < kPrefetch; ++u) {
            const int s = s0 + u;
            if (s >= n_stages) break;
            const int buf = s % kStages;
            uint32_t wv[kStageWords];
#pragma unroll
            for (int k = 0; k < kStageWords; ++k) { wv[k] = pre[u][k]; pop_acc += __popc(wv[k]); }
            const long long w1 = w_begin + (long long)(s + kPrefetch) * kStageWords;      // this slot's next occupant
#pragma unroll
            for (int k = 0; k < kStageWords; ++k) pre[u][k] = (row_ok && (w1 + k) < w_end) ? __ldg(src + w1 + k) : 0u;
            // the MMAs that read this buffer kStages stages ago must have completed
            if (s >= kStages) mbar_wait(&empty_bar[buf], (uint32_t)(((s / kStages) - 1) & 1));
            if (kATmem && op == 0) {
                // row r of A = TMEM lane r (this warp owns lanes 32 warp .. 32 warp + 31); column c of the slot holds
                // K bytes 4c .. 4c+3 of the stage, i.e. nibble c % 8 of word c / 8
                uint32_t e[kAStageCols];
#pragma unroll
                for (int k = 0; k < kStageWords; ++k)
#pragma unroll
                    for (int q = 0; q < 8; ++q) e[k * 8 + q] = plane_to_bytes(wv[k], q);
                const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(kATmemCol0 + buf * kAStageCols);
                static_assert(kAStageCols == 32, "tcgen05.st below writes 32 columns");
                asm volatile(
                    "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                    "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
                    "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
                    ::"r"(taddr), "r"(e[0]), "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]),
                      "r"(e[8]), "r"(e[9]), "r"(e[10]), "r"(e[11]), "r"(e[12]), "r"(e[13]), "r"(e[14]), "r"(e[15]),
                      "r"(e[16]), "r"(e[17]), "r"(e[18]), "r"(e[19]), "r"(e[20]), "r"(e[21]), "r"(e[22]), "r"(e[23]),
                      "r"(e[24]), "r"(e[25]), "r"(e[26]), "r"(e[27]), "r"(e[28]), "r"(e[29]), "r"(e[30]), "r"(e[31])
                    : "memory");
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&full_bar[buf]);
                continue;
            }
            unsigned char *dst = smem + buf * kSmemStageBytes + (kATmem ? 0 : op * kOperandStageBytes) + row_off;
#pragma unroll
            for (int k = 0; k < kStageWords; ++k) {          // word k = the K = 32 slice of MMA k
#pragma unroll
                for (int h = 0; h < 2; ++h) {                // planes 4 h .. 4 h + 3 -> one 16-byte K chunk (same order as A)
                    uint4 o;
                    o.x = plane_to_bytes(wv[k], 4 * h);
                    o.y = plane_to_bytes(wv[k], 4 * h + 1);
                    o.z = plane_to_bytes(wv[k], 4 * h + 2);
                    o.w = plane_to_bytes(wv[k], 4 * h + 3);
                    *reinterpret_cast<uint4 *>(dst + k * kSubTileBytes + h * 128) = o;
                }
            }
            // generic-proxy writes -> visible to the tensor core (async proxy); one arrival per warp
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&full_bar[buf]);
          }
        }
        if (pop_out != nullptr && row_ok && pop_acc) atomicAdd(pop_out + grow, pop_acc);
    } else if (lane == 0) {
        // ---------------- MMA issuer (one thread)
        const uint32_t idesc = make_idesc();
        for (int s = 0; s < n_stages; ++s) {
            const int buf = s % kStages;
            mbar_wait(&full_bar[buf], (uint32_t)((s / kStages) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_base = smem_addr(smem + buf * kSmemStageBytes);
            const uint32_t b_base = kATmem ? a_base : a_base + kOperandStageBytes;
#pragma unroll
            for (int k = 0; k < kStageWords; ++k) {
                const uint64_t db = make_desc(b_base + k * kSubTileBytes);
                const uint32_t accumulate = (s > 0 || k > 0) ? 1u : 0u;
                if (kATmem) {
                    const uint32_t a_tmem = tmem_base + (uint32_t)(kATmemCol0 + buf * kAStageCols + k * 8);
                    asm volatile(
                        "{\n\t.reg .pred p;\n\t"
                        "setp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                        ::"r"(tmem_base), "r"(a_tmem), "l"(db), "r"(idesc), "r"(accumulate)
                        : "memory");
                } else {
                    const uint64_t da = make_desc(a_base + k * kSubTileBytes);
                    asm volatile(
                        "{\n\t.reg .pred p;\n\t"
                        "setp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                        ::"r"(tmem_base), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
                        : "memory");
                }
            }
            // arrives on the barrier when the MMAs issued so far have finished reading shared memory
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&empty_bar[buf])) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&done_bar)) : "memory");
    }

    // ---- epilogue (producer warps): all MMAs done, TMEM -> registers -> global
    if (warp < kProducerWarps && n_stages > 0) {
        mbar_wait(&done_bar, 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int quad = warp & 3;                         // a warp may only touch TMEM lanes 32*(warp%4) .. +31
        const int col0 = (warp >> 2) * 64;                 // warps 0-3: columns 0..63, warps 4-7: 64..127
        const int ga = bi * kTileM + quad * 32 + lane;
#pragma unroll
        for (int c = 0; c < 64; c += 32) {
            uint32_t v[32];
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(col0 + c);
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                  "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                  "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (ga < n_a) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int gb = bj * kTileN + col0 + c + j;
                    if (gb < n_b && v[j] != 0u) {
                        atomicAdd(inter + (long long)ga * n_b + gb, (int)v[j]);
                        if (symmetric && bi != bj) atomicAdd(inter + (long long)gb * n_b + ga, (int)v[j]);
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols) : "memory");
}

// Enqueue the tensor-core intersection. Returns a cudaError_t-compatible int (0 = ok).
// pop_a / pop_b (may be NULL; symmetric: pop_b ignored) must be zeroed like `inter`: the K slices add their parts.
inline cudaError_t launch_intersect(const uint32_t *a, int n_a, const uint32_t *b, int n_b, long long words, int symmetric,
                                    int32_t *inter, int32_t *pop_a, int32_t *pop_b, int num_sms, cudaStream_t stream) {
    const int ta = (n_a + kTileM - 1) / kTileM, tb = (n_b + kTileN - 1) / kTileN;
    long long tiles = symmetric ? (long long)ta * (ta + 1) / 2 : (long long)ta * tb;
    // Two CTAs fit an SM (TMEM columns, shared memory), i.e. 2 * num_sms slots per wave. The K slices are chosen so that the
    // LAST wave is as full as possible: few tiles (c2: 15) -> as many slices as fill the slots once (rounding up instead --
    // 300 CTAs on 296 slots -- doubled the kernel time); many tiles (c4: 630 on 296 slots = 2.13 waves) -> the slice count
    // in 1..16 with the best wave efficiency (7 slices: 4410 CTAs = 14.9 waves), as long as a slice keeps >= 256 words.
    const long long slots = (long long)num_sms * 2;
    long long slices = slots / tiles;
    if (slices < 1) {
        double best = 0.0;
        slices = 1;
        for (long long s = 1; s <= 16 && words / s >= 256; ++s) {
            const long long ctas = tiles * s;
            const double eff = (double)ctas / (double)(((ctas + slots - 1) / slots) * slots);
            if (eff > best + 1e-9) { best = eff; slices = s; }
        }
    }
    const long long max_slices = (words + 4 * kStageWords - 1) / (4 * kStageWords);
    if (slices > max_slices) slices = max_slices;
    if (slices < 1) slices = 1;
    if (slices > 65535) slices = 65535;
    long long wps = (((words + slices - 1) / slices) + kStageWords - 1) / kStageWords * kStageWords;
    slices = (words + wps - 1) / wps;
    cudaError_t e = cudaFuncSetAttribute((const void *)intersect_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)tb, (unsigned)ta, (unsigned)slices);
    intersect_umma_kernel<<<grid, kThreads, kSmemBytes, stream>>>(a, n_a, b, n_b, words, wps, symmetric, inter, pop_a, pop_b);
    return cudaGetLastError();
}

}  // namespace umma
}  // namespace bff
