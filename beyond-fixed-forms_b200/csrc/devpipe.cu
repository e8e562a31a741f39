// Device-resident pipelines (include/bff.h: bff_stage2_device, bff_refine_device, bff_refine_select_device).
//
// pipeline.cu takes the reference's small decisions (which masks merge, which vote threshold, which pairs overlap,
// which stage-1 mask matches) in C++ on pinned copies of tiny matrices: six stream synchronisations per scene, and at
// c2 the device idles for half of the step. Here every one of those decisions is a small kernel (mostly one CTA) that
// leaves its counts in a device result block; the bulk kernels that follow read the counts through pointers and are
// launched with grids sized by the upper bounds. A call therefore only ENQUEUES work: no synchronisation, no host
// round trip, capturable in a CUDA graph. Every decision kernel cites the reference lines it mirrors and is compared
// bit for bit with the host version in tests/test_devpipe_gpu.py.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "common.cuh"

namespace bff {

constexpr int kDecT = 1024;          // threads of a decision kernel (one CTA)
constexpr int kLoopRows = 64;        // grid.y of the bulk kernels that loop over a device-side row count

// ------------------------------------------------------------------------------------------------ helpers
// Exclusive prefix sum of one int per thread over the whole CTA (kDecT threads, all must call). s_scan: int[34].
__device__ __forceinline__ int block_excl_scan(int v, int *s_scan, int *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();                                  // s_scan may still be read from a previous call
    if (lane == 31) s_scan[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const int w = lane < (int)(blockDim.x >> 5) ? s_scan[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        s_scan[lane] = winc - w;
        if (lane == 31) s_scan[32] = winc;
    }
    __syncthreads();
    *total = s_scan[32];
    return s_scan[warp] + inc - v;
}

// float32 IoU exactly as calculate_iou / calculate_iou_between_stages (projection_2d_to_3d.py:158-165,
// refinement.py:82-90): union = (f32(sA) + f32(sB)) - f32(I); 0/0 = NaN
__device__ __forceinline__ float iou_f32(int inter, int pa, int pb) {
    const float fi = __int2float_rn(inter);
    return __fdiv_rn(fi, __fsub_rn(__fadd_rn(__int2float_rn(pa), __int2float_rn(pb)), fi));
}

// ------------------------------------------------------------------------------------------------ components
// find_unconnected_subgraphs_tensor + the group bookkeeping of merge_masks (projection_2d_to_3d.py:190-274) on the BIT
// adjacency of the merge graph. Union-find in shared memory: every set bit (i, j > i) is an edge whose two trees are
// joined by hooking the LARGER root under the smaller one (atomicCAS on the root's parent slot; a failed CAS means the
// root was hooked meanwhile and the search continues from its new parent), so a root is always the smallest node of its
// tree and, when all edges are in, of its component = the node at which the reference emits the component. One pass
// over the edges, independent of the graph's diameter (a first version propagated minimum labels along the edges and
// took 135 us for the elongated components of c2 -- an object seen from many frames is a long chain of overlapping
// partial views). Kept groups (>= min_agg members, and the root has an edge at all: a node without any edge is an empty
// lifted row, which the reference turns into []) are numbered in root order, members ascending; conf = sequential
// float32 sum in index order / len (:220-224). scratch: 2 n int32 (member counts per root, keep flags) + n x stride_words
// (member bit rows of the intermediate labels).
__device__ __forceinline__ int uf_find(volatile int *par, int x) {
    int r = x, p;
    while ((p = par[r]) != r) r = p;
    if (x != r && par[x] != r) par[x] = r;      // parents only ever point to smaller nodes of the same tree: any ancestor is valid
    return r;
}

__global__ void __launch_bounds__(kDecT)
components_kernel(const uint32_t *__restrict__ adj_g, int n, int stride_words, int min_agg, const float *__restrict__ conf, int cap,
                  int32_t *__restrict__ res, int32_t *__restrict__ scratch, int stage_graph) {
    extern __shared__ int s_lab_raw[];
    const uint32_t *adj = adj_g;
    volatile int *lab = s_lab_raw;
    __shared__ int s_scan[34];
    const int tid = threadIdx.x;
    int32_t *cnt = scratch, *flag = scratch + n;
    int32_t *group_first = res + BFF_S2R_HEADER, *group_size = group_first + cap;
    float *group_conf = reinterpret_cast<float *>(group_size + cap);
    int32_t *seg = res + BFF_S2R_HEADER + 7 * (size_t)cap, *members = seg + cap + 1;
    float *s_conf = reinterpret_cast<float *>(s_lab_raw + n);      // the confidences, staged once: the member walk below must not
    for (int i = tid; i < n; i += kDecT) {                          // wait for a global load per member (a 100-member group cost 35 us)
        lab[i] = i;
        cnt[i] = 0;
        s_conf[i] = __ldg(conf + i);
    }
    // small graphs (c2: 525 x 18 words): the adjacency and the member rows live in shared memory too, so that none of the
    // passes below waits for global memory; large graphs read them from global memory / the scratch block
    const long long items = (long long)n * stride_words;
    uint32_t *mem = reinterpret_cast<uint32_t *>(scratch + 2 * (size_t)n);
    if (stage_graph) {
        uint32_t *s_adj = reinterpret_cast<uint32_t *>(s_lab_raw + 2 * (size_t)n);
        for (long long it = tid; it < items; it += kDecT) s_adj[it] = __ldg(adj_g + it);
        mem = s_adj + items;
        adj = s_adj;
    }
    __syncthreads();
    // hook(u, v): join the trees of u and v
    auto hook = [&](int u, int v) {
        int ru = uf_find(lab, u), rv = uf_find(lab, v);
        while (ru != rv) {
            const int hi = max(ru, rv), lo = min(ru, rv);
            const int old = atomicCAS(s_lab_raw + hi, hi, lo);
            if (old == hi) break;
            ru = uf_find(lab, old);
            rv = uf_find(lab, lo);
        }
    };
    // every tree a star: pointer jumping until nothing moves (a hooked chain of d nodes takes log2 d rounds)
    auto flatten = [&]() {
        while (true) {
            int moved = 0;
            for (int i = tid; i < n; i += kDecT) {
                const int p = lab[i], gp = lab[p];
                if (gp != p) { lab[i] = gp; moved = 1; }
            }
            if (!__syncthreads_or(moved)) break;
        }
    };
    // pass 1: every node points at its smallest neighbour (or itself): one read per node, no atomics, no search. A clique
    // collapses onto its smallest member at once, an elongated component (an object seen from many frames is a chain of
    // overlapping partial views) becomes a shallow forest.
    for (int i = tid; i < n; i += kDecT) {
        const uint32_t *row = adj + (size_t)i * stride_words;
        int best = i;
        for (int w = 0; w <= (i >> 5); ++w) {
            const uint32_t bits = row[w];
            if (bits) {
                best = min(i, w * 32 + __ffs(bits) - 1);
                break;
            }
        }
        lab[i] = best;
    }
    __syncthreads();
    flatten();
    // member bit rows of the labels so far: bit i of row lab[i]
    for (long long it = tid; it < items; it += kDecT) mem[it] = 0u;
    __syncthreads();
    for (int i = tid; i < n; i += kDecT) atomicOr(mem + (size_t)lab[i] * stride_words + (i >> 5), 1u << (i & 31));
    __syncthreads();
    // pass 2: a word of row i may only hold members of i's own tree; whatever else it holds is an edge into another tree
    // and is joined exactly (root search + hook). One AND per word instead of one label read per edge: the merge graph of
    // c2 has 30k edges in dense cliques, and walking them bit by bit (lanes with full words holding up their warps) took
    // 55k warp instructions, hooking all of them 90 us.
    for (long long it = tid; it < items; it += kDecT) {
        const int i = (int)(it / stride_words), w = (int)(it % stride_words);
        uint32_t bits = adj[it] & ~mem[(size_t)lab[i] * stride_words + w];
        while (bits) {
            const int j = w * 32 + __ffs(bits) - 1;
            bits &= bits - 1;
            hook(i, j);
        }
    }
    __syncthreads();
    flatten();
    for (int i = tid; i < n; i += kDecT) atomicAdd(&cnt[lab[i]], 1);
    __syncthreads();
    // kept roots in ascending order: thread t owns the nodes [t per, (t + 1) per)
    const int per = (n + kDecT - 1) / kDecT;
    const int lo = min(n, tid * per), hi = min(n, lo + per);
    int k_local = 0, m_local = 0;
    for (int r = lo; r < hi; ++r) {
        int keep = 0;
        if (lab[r] == r && cnt[r] >= min_agg) {
            const uint32_t *row = adj + (size_t)r * stride_words;
            uint32_t any = 0;
            for (int w = 0; w < stride_words; ++w) any |= row[w];
            keep = any != 0u;
        }
        flag[r] = keep;
        if (keep) { ++k_local; m_local += cnt[r]; }
    }
    int g_total, m_total;
    int gbase = block_excl_scan(k_local, s_scan, &g_total);
    int mbase = block_excl_scan(m_local, s_scan, &m_total);
    for (int r = lo; r < hi; ++r) {
        if (!flag[r]) continue;
        if (gbase < cap) {
            group_first[gbase] = r;
            group_size[gbase] = cnt[r];
            seg[gbase] = mbase;
        }
        ++gbase;
        mbase += cnt[r];
    }
    const int g = min(g_total, cap);
    if (tid == 0) {
        res[BFF_S2R_N_GROUPS] = g;
        if (g_total > cap) atomicOr(&res[BFF_S2R_ERROR], 2);
        else seg[g] = m_total;
    }
    __syncthreads();
    const int *root_of = s_lab_raw;                                 // final now: plain (pipelined) shared-memory reads
    for (int gi = tid; gi < g; gi += kDecT) {
        const int r = group_first[gi];
        int pos = seg[gi];
        float acc = 0.0f;
#pragma unroll 4
        for (int i = r; i < n; ++i)
            if (root_of[i] == r) {
                members[pos++] = i;
                acc = __fadd_rn(acc, s_conf[i]);
            }
        group_conf[gi] = __fdiv_rn(acc, __int2float_rn(group_size[gi]));
    }
}

// ------------------------------------------------------------------------------------------------ bulk kernels, device-side counts
// out[g] = OR of rows[idx[k]], k in [seg[g], seg[g+1]) for g < *n_groups (merge_masks :214-223, refinement.py:269)
__global__ void __launch_bounds__(256)
rows_or_reduce_dc_kernel(const uint32_t *__restrict__ rows, long long words, const int32_t *__restrict__ idx, const int32_t *__restrict__ seg,
                         const int32_t *__restrict__ n_groups, uint32_t *__restrict__ out) {
    __shared__ int32_t s_idx[256];
    const long long w = (long long)blockIdx.x * 256 + threadIdx.x;
    const bool live = w < words;
    const int ng = *n_groups;
    for (int g = blockIdx.y; g < ng; g += gridDim.y) {
        const int k0 = seg[g], k1 = seg[g + 1];
        uint32_t v = 0;
        for (int base = k0; base < k1; base += 256) {
            const int cnt = min(256, k1 - base);
            __syncthreads();
            if ((int)threadIdx.x < cnt) s_idx[threadIdx.x] = idx[base + threadIdx.x];
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
}

// out[k] = rows[idx[k]] for k < *n (n_dev) or k < n_host
__global__ void __launch_bounds__(256)
rows_gather_dc_kernel(const uint32_t *__restrict__ rows, long long words, const int32_t *__restrict__ idx, const int32_t *__restrict__ n_dev,
                      int n_host, uint32_t *__restrict__ out) {
    const long long w = (long long)blockIdx.x * 256 + threadIdx.x;
    if (w >= words) return;
    const int cnt = n_dev ? *n_dev : n_host;
    for (int k = blockIdx.y; k < cnt; k += gridDim.y) out[(long long)k * words + w] = rows[(long long)idx[k] * words + w];
}

// out[k] = (src[k] < n1 ? s1 : s2)[...] for k < *n: the rows of the saved masks (refinement.py:340-392)
__global__ void __launch_bounds__(256)
rows_gather2_dc_kernel(const uint32_t *__restrict__ s1, int n1, const uint32_t *__restrict__ s2, long long words, const int32_t *__restrict__ src,
                       const int32_t *__restrict__ n, uint32_t *__restrict__ out) {
    const long long w = (long long)blockIdx.x * 256 + threadIdx.x;
    if (w >= words) return;
    const int cnt = *n;
    for (int k = blockIdx.y; k < cnt; k += gridDim.y) {
        const int s = src[k];
        out[(long long)k * words + w] = s < n1 ? s1[(long long)s * words + w] : s2[(long long)(s - n1) * words + w];
    }
}

// Intersections of a (n_a rows) with b (n_b rows), either count on the device (pointer) or on the host; one warp per row
// of a, the b rows of the current 256-word chunk in shared memory, b taken 32 rows at a time. Both popcount vectors come
// out of the same pass. inter (ld columns per row) / pop_a / pop_b must be zeroed; K-slice partial sums are added.
constexpr int kDcChunk = 256;
constexpr int kDcRows = 8;

template <int NB>
__device__ __forceinline__ void skinny_block(const uint32_t *__restrict__ a, int n_a, int row, const uint32_t *__restrict__ b, int nb,
                                             long long words, long long w_begin, long long w_end, int ld, int col0,
                                             int32_t *__restrict__ inter, int32_t *__restrict__ pop_a, int32_t *__restrict__ pop_b,
                                             uint32_t (*sb)[kDcChunk]) {
    const int lane = threadIdx.x & 31;
    const bool row_ok = row < n_a;
    const uint32_t *ar = a + (long long)(row_ok ? row : 0) * words;
    int acc[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) acc[j] = 0;
    int acc_a = 0;
    for (long long w0 = w_begin; w0 < w_end; w0 += kDcChunk) {
        const int cw = (int)min((long long)kDcChunk, w_end - w0);
        __syncthreads();
        for (int i = threadIdx.x; i < NB * kDcChunk; i += kDcRows * 32) {
            const int j = i / kDcChunk, k = i % kDcChunk;
            sb[j][k] = (j < nb && k < cw) ? __ldg(b + (long long)j * words + w0 + k) : 0u;
        }
        __syncthreads();
        if (pop_b != nullptr) {
            for (int i = threadIdx.x; i < NB * kDcChunk; i += kDcRows * 32) {
                const int j = i / kDcChunk;          // 256 consecutive i share j: a warp's 32 values belong to one row
                int c = __popc(sb[j][i % kDcChunk]);
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                if (lane == 0 && c && j < nb) atomicAdd(pop_b + col0 + j, c);
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
    if (!row_ok) return;
    if (pop_a != nullptr) {
        for (int o = 16; o > 0; o >>= 1) acc_a += __shfl_xor_sync(0xffffffffu, acc_a, o);
        if (lane == 0 && acc_a) atomicAdd(pop_a + row, acc_a);
    }
#pragma unroll
    for (int j = 0; j < NB; ++j) {
        int v = acc[j];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && j < nb && v) atomicAdd(inter + (long long)row * ld + col0 + j, v);
    }
}

__global__ void __launch_bounds__(kDcRows * 32)
intersect_dc_kernel(const uint32_t *__restrict__ a, const int32_t *__restrict__ n_a_dev, int n_a_host, const uint32_t *__restrict__ b,
                    const int32_t *__restrict__ n_b_dev, int n_b_host, long long words, long long words_per_slice, int ld,
                    int32_t *__restrict__ inter, int32_t *__restrict__ pop_a, int32_t *__restrict__ pop_b) {
    __shared__ uint32_t sb[32][kDcChunk];
    const int n_a = n_a_dev ? *n_a_dev : n_a_host, n_b = n_b_dev ? *n_b_dev : n_b_host;
    const long long w_begin = (long long)blockIdx.y * words_per_slice;
    const long long w_end = min(words, w_begin + words_per_slice);
    const int warp = threadIdx.x >> 5;
    for (int row0 = blockIdx.x * kDcRows; row0 < n_a; row0 += gridDim.x * kDcRows) {
        const int row = row0 + warp;
        for (int col0 = 0; col0 < n_b; col0 += 32) {
            const int nb = min(32, n_b - col0);
            const uint32_t *bb = b + (long long)col0 * words;
            int32_t *pa = col0 == 0 ? pop_a : nullptr;                        // a's popcounts: once, with the first column block
            int32_t *pb = row0 == 0 ? pop_b : nullptr;                        // b's popcounts: once, by the first row block
            if (nb <= 8) skinny_block<8>(a, n_a, row, bb, nb, words, w_begin, w_end, ld, col0, inter, pa, pb, sb);
            else if (nb <= 16) skinny_block<16>(a, n_a, row, bb, nb, words, w_begin, w_end, ld, col0, inter, pa, pb, sb);
            else skinny_block<32>(a, n_a, row, bb, nb, words, w_begin, w_end, ld, col0, inter, pa, pb, sb);
        }
    }
}

// rows[dst] = OR_INTO ? rows[dst] | rows[src] : rows[dst] & ~rows[src] for the *n_steps (dst, src) pairs of prog, in
// order; every thread owns one word column (solve_overlapping :295-299, refinement.py:248). With at most kProgRows rows
// (n_rows_dev given) the column is loaded into shared memory once, the program replayed there and the column written
// back: every step through global memory instead is a dependent load-modify-store chain.
constexpr int kProgRows = 64;
template <bool OR_INTO>
__global__ void __launch_bounds__(128)
row_program_dc_kernel(uint32_t *__restrict__ rows, long long words, const int32_t *__restrict__ prog, const int32_t *__restrict__ n_steps,
                      const int32_t *__restrict__ n_rows_dev) {
    __shared__ uint32_t s_col[kProgRows][128];
    const long long w = (long long)blockIdx.x * 128 + threadIdx.x;
    const int n = *n_steps;
    if (w >= words || n == 0) return;
    const int nr = n_rows_dev ? *n_rows_dev : kProgRows + 1;
    if (nr <= kProgRows) {
        for (int r = 0; r < nr; ++r) s_col[r][threadIdx.x] = rows[(long long)r * words + w];
        for (int k = 0; k < n; ++k) {
            const int dst = __ldg(prog + 2 * k), src = __ldg(prog + 2 * k + 1);
            const uint32_t d = s_col[dst][threadIdx.x], s = s_col[src][threadIdx.x];
            s_col[dst][threadIdx.x] = OR_INTO ? (d | s) : (d & ~s);
        }
        for (int r = 0; r < nr; ++r) rows[(long long)r * words + w] = s_col[r][threadIdx.x];
        return;
    }
    for (int k = 0; k < n; ++k) {
        const int dst = prog[2 * k], src = prog[2 * k + 1];
        const uint32_t d = rows[(long long)dst * words + w], s = rows[(long long)src * words + w];
        rows[(long long)dst * words + w] = OR_INTO ? (d | s) : (d & ~s);
    }
}

// rows[r] &= keep (in place) and pop[r] += popcount, r < *n_rows (projection_2d_to_3d.py:594-596); pop zeroed by the caller
constexpr int kAndPopWords = 2048;
__global__ void __launch_bounds__(256)
rows_and_popcount_dc_kernel(uint32_t *__restrict__ rows, long long words, const uint32_t *__restrict__ keep, const int32_t *__restrict__ n_rows,
                            int32_t *__restrict__ pop) {
    __shared__ int red[8];
    const long long w0 = (long long)blockIdx.x * kAndPopWords;
    const long long w1 = min(words, w0 + kAndPopWords);
    const int nr = *n_rows;
    for (int row = blockIdx.y; row < nr; row += gridDim.y) {
        uint32_t *r = rows + (long long)row * words;
        int s = 0;
        for (long long w = w0 + threadIdx.x; w < w1; w += 256) {
            const uint32_t v = r[w] & __ldg(keep + w);
            r[w] = v;
            s += __popc(v);
        }
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int i = 0; i < 8; ++i) t += red[i];
            if (t) atomicAdd(pop + row, t);
        }
    }
}

// ------------------------------------------------------------------------------------------------ solve_overlapping's schedule
// Pairs i < k that overlap in the INCOMING rows, in (i, k) order; the mask merged from MORE members keeps the shared
// points, ties: i loses (projection_2d_to_3d.py:285-299). prog[2 s] = row that loses the bits of row prog[2 s + 1].
__global__ void __launch_bounds__(kDecT)
overlap_program_kernel(const int32_t *__restrict__ ginter, int ld, int cap, int32_t *__restrict__ res, int32_t *__restrict__ prog) {
    __shared__ int s_scan[34];
    const int g = res[BFF_S2R_N_GROUPS];
    const int32_t *size = res + BFF_S2R_HEADER + cap;
    const int per = (g + kDecT - 1) / kDecT;
    const int lo = min(g, (int)threadIdx.x * per), hi = min(g, lo + per);
    int c = 0;
    for (int i = lo; i < hi; ++i)
        for (int k = i + 1; k < g; ++k) c += ginter[(long long)i * ld + k] > 0;
    int total;
    long long pos = block_excl_scan(c, s_scan, &total);
    for (int i = lo; i < hi; ++i)
        for (int k = i + 1; k < g; ++k)
            if (ginter[(long long)i * ld + k] > 0) {
                const bool i_wins = size[i] > size[k];
                prog[2 * pos] = i_wins ? k : i;
                prog[2 * pos + 1] = i_wins ? i : k;
                ++pos;
            }
    if (threadIdx.x == 0) res[BFF_S2R_N_OVERLAP] = total;
}

// ------------------------------------------------------------------------------------------------ vote threshold
// The reference takes unique() of N float32 values and picks element floor(p * len) (:512-517 masked_counts,
// :571-576 detected_ratio = mc / (vc + 1)). Every distinct value comes from a non-empty bin of the (mc, vc) histogram, so:
// maxima -> histogram -> the distinct float32 values of the non-empty bins in an open-addressing set -> a radix select
// of the floor(p * len)-th smallest (non-negative floats order like their bit patterns). All sizes stay on the device.
constexpr uint32_t kEmptySlot = 0xffffffffu;
constexpr int kHistSmem = 8192;

__device__ __forceinline__ void vote_layout(const int32_t *d_max, int mode, int &mb, int &vb, long long &bins, long long &tsize) {
    mb = d_max[0] + 1;
    vb = mode == 0 ? d_max[1] + 1 : 1;
    bins = (long long)mb * vb;
    tsize = 1024;
    while (tsize < 2 * bins) tsize <<= 1;
}

__global__ void __launch_bounds__(256)
vote_prepare_kernel(const int32_t *__restrict__ d_max, int mode, long long hist_cap, long long table_cap, uint32_t *__restrict__ hist,
                    uint32_t *__restrict__ table, int32_t *__restrict__ n_values, int32_t *__restrict__ res) {
    int mb, vb;
    long long bins, tsize;
    vote_layout(d_max, mode, mb, vb, bins, tsize);
    if (bins > hist_cap || tsize > table_cap) {
        if (blockIdx.x == 0 && threadIdx.x == 0) { atomicOr(&res[BFF_S2R_ERROR], 2); *n_values = 0; }
        return;
    }
    const long long stride = (long long)gridDim.x * 256;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < tsize; i += stride) {
        if (i < bins) hist[i] = 0u;
        table[i] = kEmptySlot;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { res[BFF_S2R_MC_MAX] = d_max[0]; res[BFF_S2R_VC_MAX] = d_max[1]; *n_values = 0; }
}

__global__ void __launch_bounds__(256)
vote_hist_dc_kernel(const int32_t *__restrict__ mc, const int32_t *__restrict__ vc, long long n, const int32_t *__restrict__ d_max, int mode,
                    long long hist_cap, long long table_cap, uint32_t *__restrict__ hist) {
    __shared__ uint32_t s_hist[kHistSmem];
    int mb, vb;
    long long bins, tsize;
    vote_layout(d_max, mode, mb, vb, bins, tsize);
    if (bins > hist_cap || tsize > table_cap) return;
    const bool use_smem = bins <= kHistSmem;
    if (use_smem) {
        for (int i = threadIdx.x; i < (int)bins; i += 256) s_hist[i] = 0;
        __syncthreads();
    }
    const long long stride = (long long)gridDim.x * 256;
    const long long n_round = (n + 31) / 32 * 32;     // keep warps converged for __match_any_sync
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n_round; i += stride) {
        int bin = -1;
        if (i < n) {
            const int a = min(max(__ldg(mc + i), 0), mb - 1);
            const int b = mode == 0 ? min(max(__ldg(vc + i), 0), vb - 1) : 0;
            bin = a * vb + b;
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
        for (int i = threadIdx.x; i < (int)bins; i += 256) {
            const uint32_t v = s_hist[i];
            if (v) atomicAdd(hist + i, v);
        }
    }
}

__global__ void __launch_bounds__(256)
vote_unique_insert_dc_kernel(const uint32_t *__restrict__ hist, const int32_t *__restrict__ d_max, int mode, long long hist_cap, long long table_cap,
                             uint32_t *__restrict__ table) {
    int mb, vb;
    long long bins, tsize;
    vote_layout(d_max, mode, mb, vb, bins, tsize);
    if (bins > hist_cap || tsize > table_cap) return;
    const uint32_t cap_mask = (uint32_t)(tsize - 1);
    const int hash_shift = 32 - (63 - __clzll(tsize));
    const long long stride = (long long)gridDim.x * 256;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < bins; i += stride) {
        if (hist[i] == 0u) continue;
        const int a = (int)(i / vb), b = (int)(i % vb);
        const float val = mode == 0 ? __fdiv_rn((float)a, __fadd_rn((float)b, 1.0f)) : (float)a;
        const uint32_t bits = __float_as_uint(val);
        uint32_t slot = (bits * 2654435761u) >> hash_shift;      // Fibonacci hashing: the TOP bits of the product
        while (true) {
            uint32_t old = *reinterpret_cast<volatile uint32_t *>(table + slot);
            if (old == bits) break;
            if (old == kEmptySlot) {
                old = atomicCAS(table + slot, kEmptySlot, bits);
                if (old == kEmptySlot || old == bits) break;
            }
            slot = (slot + 1u) & cap_mask;
        }
    }
}

// the occupied slots of the set, compacted (order irrelevant: the select below ranks them)
__global__ void __launch_bounds__(256)
vote_compact_dc_kernel(const uint32_t *__restrict__ table, const int32_t *__restrict__ d_max, int mode, long long hist_cap, long long table_cap,
                       uint32_t *__restrict__ values, int32_t *__restrict__ n_values) {
    int mb, vb;
    long long bins, tsize;
    vote_layout(d_max, mode, mb, vb, bins, tsize);
    if (bins > hist_cap || tsize > table_cap) return;
    const long long stride = (long long)gridDim.x * 256;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < tsize; i += stride) {
        const uint32_t v = table[i];
        if (v != kEmptySlot) values[atomicAdd(n_values, 1)] = v;
    }
}

// u[floor(p * len(u))] of the sorted distinct values (:516-517, :575-576; p a Python float: double arithmetic), by an
// 8-bit radix select over the compacted bit patterns. One CTA.
__global__ void __launch_bounds__(kDecT)
vote_select_kernel(const uint32_t *__restrict__ values, const int32_t *__restrict__ n_values, double fraction, float *__restrict__ d_thr,
                   int32_t *__restrict__ res) {
    __shared__ int s_hist[256];
    __shared__ uint32_t s_prefix;
    __shared__ long long s_k;
    const int n_u = *n_values;
    const double pos = floor(fraction * (double)n_u);
    if (!(pos >= 0.0 && pos < (double)n_u)) {                 // the reference raises IndexError here (also: a scratch bound was hit)
        if (threadIdx.x == 0) {
            atomicOr(&res[BFF_S2R_ERROR], 1);
            res[BFF_S2R_N_UNIQUE] = n_u;
            *d_thr = 0.0f;
        }
        return;
    }
    if (threadIdx.x == 0) { s_prefix = 0u; s_k = (long long)pos; }
    uint32_t mask = 0u;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += kDecT) s_hist[i] = 0;
        __syncthreads();
        const uint32_t prefix = s_prefix;
        for (int i = threadIdx.x; i < n_u; i += kDecT) {
            const uint32_t v = values[i];
            if ((v & mask) == prefix) atomicAdd(&s_hist[(v >> shift) & 255u], 1);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            // the bucket that holds the k-th value: lane l owns buckets 8 l .. 8 l + 7
            const int lane = threadIdx.x;
            int own = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) own += s_hist[lane * 8 + b];
            int inc = own;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            const long long k = s_k;
            const unsigned holds = __ballot_sync(0xffffffffu, k < inc);     // inclusive sums grow with the lane: first lane past k
            const int owner = holds ? __ffs(holds) - 1 : 31;
            if (lane == owner) {
                long long kk = k - (inc - own);
                int b = lane * 8;
                while (b < lane * 8 + 7 && kk >= s_hist[b]) { kk -= s_hist[b]; ++b; }
                s_k = kk;
                s_prefix = prefix | ((uint32_t)b << shift);
            }
        }
        mask |= 255u << shift;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *d_thr = __uint_as_float(s_prefix);
        res[BFF_S2R_FILTER_VALUE] = (int32_t)s_prefix;
        res[BFF_S2R_N_UNIQUE] = n_u;
    }
}

// keep bit per point of the slice, thr read from the device (:571-583)
__global__ void __launch_bounds__(256)
point_keep_dc_kernel(const int32_t *__restrict__ mc, const int32_t *__restrict__ vc, long long n, long long words, int mode,
                     const float *__restrict__ d_thr, uint32_t *__restrict__ keep) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    bool k = false;
    if (i < n) {
        const int m = __ldg(mc + i);
        if (m > 0) {
            const float fm = __int2float_rn(m);
            if (mode == 0) k = !(__fdiv_rn(fm, __fadd_rn(__int2float_rn(__ldg(vc + i)), 1.0f)) < *d_thr);
            else if (mode == 1) k = !(fm < *d_thr);
            else k = true;
        }
    }
    const unsigned w = __ballot_sync(0xffffffffu, k);
    if ((threadIdx.x & 31) == 0 && (i >> 5) < words) keep[i >> 5] = w;
}

// ------------------------------------------------------------------------------------------------ row filter
struct RowFilter {
    int small_is_float, ratio_is_float;
    long long small_i, ratio_i;
    float small_f, ratio_f;
};

// keep rows with after > remove_small_masks and after > remove_filtered_masks * before, torch's promotion rules
// (:601-623): int64 tensor vs Python int compares as integers, vs Python float in float32
__global__ void __launch_bounds__(kDecT)
row_filter_kernel(int cap, RowFilter f, int32_t *__restrict__ res) {
    __shared__ int s_scan[34];
    const int g = res[BFF_S2R_N_GROUPS];
    const float *group_conf = reinterpret_cast<const float *>(res + BFF_S2R_HEADER + 2 * (size_t)cap);
    const int32_t *before = res + BFF_S2R_HEADER + 3 * (size_t)cap, *after = before + cap;
    int32_t *selected = res + BFF_S2R_HEADER + 5 * (size_t)cap;
    float *selected_conf = reinterpret_cast<float *>(selected + cap);
    const int per = (g + kDecT - 1) / kDecT;
    const int lo = min(g, (int)threadIdx.x * per), hi = min(g, lo + per);
    auto pass = [&](int k) {
        const int a = after[k], b = before[k];
        const bool big = f.small_is_float ? (__int2float_rn(a) > f.small_f) : ((long long)a > f.small_i);
        const bool kept = f.ratio_is_float ? (__int2float_rn(a) > __fmul_rn(f.ratio_f, __int2float_rn(b))) : ((long long)a > f.ratio_i * (long long)b);
        return big && kept;
    };
    int c = 0;
    for (int k = lo; k < hi; ++k) c += pass(k);
    int total;
    int pos = block_excl_scan(c, s_scan, &total);
    for (int k = lo; k < hi; ++k)
        if (pass(k)) {
            selected[pos] = k;
            selected_conf[pos] = group_conf[k];
            ++pos;
        }
    if (threadIdx.x == 0) res[BFF_S2R_N_SELECTED] = total;
}

// ------------------------------------------------------------------------------------------------ refinement decisions
// torch.argmax over the stage-1 axis for every stage-2 column (refinement.py:211, :288): first maximum, NaN counts as
// the maximum (first NaN wins). One warp per column; IoU computed on the fly from the integer intersections.
struct ArgBest { float v; int i; };
__device__ __forceinline__ bool arg_better(ArgBest a, ArgBest b) {
    if (b.i == 0x7fffffff) return true;
    if (a.i == 0x7fffffff) return false;
    const bool na = isnan(a.v), nb = isnan(b.v);
    if (na || nb) return (na && nb) ? a.i < b.i : na;
    if (a.v != b.v) return a.v > b.v;
    return a.i < b.i;
}

__global__ void __launch_bounds__(256)
argmax_cols_kernel(const int32_t *__restrict__ inter, int ld, const int32_t *__restrict__ pa, const int32_t *__restrict__ pb, int n_a,
                   const int32_t *__restrict__ n_b_dev, int32_t *__restrict__ arg, float *__restrict__ val) {
    const int lane = threadIdx.x & 31;
    const int n_b = *n_b_dev;
    for (int k = blockIdx.x * 8 + (threadIdx.x >> 5); k < n_b; k += gridDim.x * 8) {
        ArgBest best = {0.0f, 0x7fffffff};
        const int pbk = pb[k];
        for (int i = lane; i < n_a; i += 32) {
            const ArgBest x = {iou_f32(inter[(long long)i * ld + k], pa[i], pbk), i};
            if (arg_better(x, best)) best = x;
        }
        for (int o = 16; o > 0; o >>= 1) {
            ArgBest other;
            other.v = __shfl_xor_sync(0xffffffffu, best.v, o);
            other.i = __shfl_xor_sync(0xffffffffu, best.i, o);
            if (arg_better(other, best)) best = other;
        }
        if (lane == 0) {
            arg[k] = best.i;
            if (val != nullptr) val[k] = best.v;
        }
    }
}

struct RefineLayout {
    int n1, cap2;
    __host__ __device__ size_t sims() const { return BFF_RFR_HEADER; }
    __host__ __device__ size_t first_match() const { return sims() + 2 * (size_t)cap2; }
    __host__ __device__ size_t best() const { return first_match() + cap2; }
    __host__ __device__ size_t match() const { return best() + cap2; }
    __host__ __device__ size_t match_iou() const { return match() + cap2; }
    __host__ __device__ size_t conf_out() const { return match_iou() + cap2; }
    __host__ __device__ size_t idx2() const { return conf_out() + cap2; }
    __host__ __device__ size_t seg2() const { return idx2() + cap2; }
    __host__ __device__ size_t other() const { return seg2() + cap2 + 1; }
    __host__ __device__ size_t final_src() const { return other() + n1; }
    __host__ __device__ size_t final_conf() const { return final_src() + n1 + cap2; }
};

// The sequential merge loop of pass 1 (refinement.py:230-249) and the merging of stage-2 masks that share a stage-1
// match (:258-281). Row i of the loop runs on warp 0 (redirects set by earlier rows decide whether row i is skipped, so
// rows are sequential; the scan over k inside a row is 32 wide and appends its ORs in k order). Then, with
// best = best_match_after_iou_check: masks whose match nobody shares keep their order; every match shared by several
// masks yields ONE merged mask (members ascending, float32 conf = sequential sum / count) and the merged masks follow
// in ascending order of the match, because the reference appends them while walking torch.unique(best).
// scratch: n1 int32 (how many stage-2 masks chose stage-1 mask u). dynamic shared memory: cap2 int32 + kMergeStage
// int32; with m <= kMergeDim the m x m intersections, the popcounts and the first matches are staged there, so that the
// sequential loop of warp 0 does not wait for a chain of global loads per row.
constexpr int kMergeDim = 96;
constexpr int kMergeStage = kMergeDim * kMergeDim + 2 * kMergeDim;
__global__ void __launch_bounds__(kDecT)
refine_merge_kernel(RefineLayout L, const int32_t *__restrict__ n2_dev, const int32_t *__restrict__ ginter_g, int ld_g, const int32_t *__restrict__ gpop_g,
                    float thr, const float *__restrict__ s2_conf, int32_t *__restrict__ res, int32_t *__restrict__ or_prog,
                    int32_t *__restrict__ scratch) {
    extern __shared__ int s_redirect[];
    __shared__ int s_scan[34];
    __shared__ int s_nor;
    const int tid = threadIdx.x, lane = tid & 31;
    const int m = *n2_dev;
    const int32_t *ginter = ginter_g, *gpop = gpop_g, *mm = res + L.first_match();
    int ld = ld_g;
    if (m <= kMergeDim) {
        int *st = s_redirect + L.cap2;
        for (int i = tid; i < m * m; i += kDecT) st[i] = ginter_g[(long long)(i / m) * ld_g + (i % m)];
        for (int i = tid; i < m; i += kDecT) { st[m * m + i] = gpop_g[i]; st[m * m + m + i] = mm[i]; }
        ginter = st; gpop = st + m * m; mm = st + m * m + m;
        ld = m;
    }
    int32_t *best = res + L.best();
    float *conf_out = reinterpret_cast<float *>(res + L.conf_out());
    int32_t *idx2 = res + L.idx2(), *seg2 = res + L.seg2();
    int32_t *cnt1 = scratch;
    for (int u = tid; u < L.n1; u += kDecT) cnt1[u] = 0;
    for (int k = tid; k < m; k += kDecT) s_redirect[k] = -1;
    __syncthreads();
    if (tid < 32) {
        int n_or = 0;
        for (int i = 0; i < m; ++i) {
            const int r = s_redirect[i];
            if (r != -1) {
                if (lane == 0) best[i] = mm[r];
                continue;
            }
            const int mi = mm[i];
            if (lane == 0) best[i] = mi;
            const int pi = gpop[i];
            for (int k0 = 0; k0 < m; k0 += 32) {
                const int k = k0 + lane;
                const bool hit = k < m && k != i && iou_f32(ginter[(long long)i * ld + k], pi, gpop[k]) > thr;     // diagonal set to 0 (:219)
                const unsigned hits = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    s_redirect[k] = i;
                    const int pos = n_or + __popc(hits & ((1u << lane) - 1u));
                    or_prog[2 * (long long)pos] = mi;
                    or_prog[2 * (long long)pos + 1] = mm[k];
                }
                n_or += __popc(hits);
            }
            __syncwarp();
        }
        if (lane == 0) s_nor = n_or;
    }
    __syncthreads();
    for (int k = tid; k < m; k += kDecT) atomicAdd(&cnt1[best[k]], 1);
    __syncthreads();
    // masks that keep their match to themselves, in order
    const int per = (m + kDecT - 1) / kDecT;
    const int lo = min(m, tid * per), hi = min(m, lo + per);
    int c = 0;
    for (int k = lo; k < hi; ++k) c += cnt1[best[k]] == 1;
    int n_single;
    int pos = block_excl_scan(c, s_scan, &n_single);
    for (int k = lo; k < hi; ++k)
        if (cnt1[best[k]] == 1) {
            idx2[pos] = k;
            seg2[pos] = pos;
            conf_out[pos] = s2_conf[k];
            ++pos;
        }
    // shared matches in ascending order of the stage-1 index
    const int per1 = (L.n1 + kDecT - 1) / kDecT;
    const int lo1 = min(L.n1, tid * per1), hi1 = min(L.n1, lo1 + per1);
    int cg = 0, cm = 0;
    for (int u = lo1; u < hi1; ++u)
        if (cnt1[u] > 1) { ++cg; cm += cnt1[u]; }
    int n_multi, n_multi_members;
    int gpos = block_excl_scan(cg, s_scan, &n_multi);
    int mpos = block_excl_scan(cm, s_scan, &n_multi_members);
    for (int u = lo1; u < hi1; ++u) {
        if (cnt1[u] <= 1) continue;
        const int out = n_single + gpos, first = n_single + mpos;
        seg2[out] = first;
        int p = first;
        float sum = 0.0f;
        for (int k = 0; k < m; ++k)
            if (best[k] == u) {
                idx2[p++] = k;
                sum = __fadd_rn(sum, s2_conf[k]);
            }
        conf_out[out] = __fdiv_rn(sum, __int2float_rn(cnt1[u]));
        ++gpos;
        mpos += cnt1[u];
    }
    if (tid == 0) {
        const int m_out = n_single + n_multi;
        seg2[m_out] = n_single + n_multi_members;
        res[BFF_RFR_N2_IN] = m;
        res[BFF_RFR_N2_OUT] = m_out;
        res[BFF_RFR_N_OR] = s_nor;
    }
}

// sims of the matched labels (refinement.py:297-302) and the unmatched stage-1 masks of the class (:293).
// scratch: n1 int32 flags.
__global__ void __launch_bounds__(kDecT)
refine_match_info_kernel(RefineLayout L, const uint8_t *__restrict__ is_query, const double *__restrict__ sim, int32_t *__restrict__ res,
                         int32_t *__restrict__ scratch) {
    __shared__ int s_scan[34];
    const int tid = threadIdx.x;
    const int m = res[BFF_RFR_N2_OUT];
    const int32_t *match = res + L.match();
    double *sims = reinterpret_cast<double *>(res + L.sims());
    int32_t *other = res + L.other();
    int32_t *matched = scratch;
    for (int i = tid; i < L.n1; i += kDecT) matched[i] = 0;
    __syncthreads();
    for (int k = tid; k < m; k += kDecT) {
        matched[match[k]] = 1;
        sims[k] = sim[match[k]];
    }
    __syncthreads();
    const int per = (L.n1 + kDecT - 1) / kDecT;
    const int lo = min(L.n1, tid * per), hi = min(L.n1, lo + per);
    int c = 0;
    for (int i = lo; i < hi; ++i) c += is_query[i] && !matched[i];
    int total;
    int pos = block_excl_scan(c, s_scan, &total);
    for (int i = lo; i < hi; ++i)
        if (is_query[i] && !matched[i]) other[pos++] = i;
    if (tid == 0) res[BFF_RFR_N_OTHER] = total;
}

// Pass 2 (refinement.py:317-412). mode 1: threshold = sorted(set(sims))[int(len * percentile)] over this scene's sims
// (0 when nothing matched); mode 2: the caller's threshold. Then per matched pair: IoU above refiment_iou_thres ->
// the stage-1 mask if the labels are similar enough, else nothing; IoU not above -> the stage-2 mask; confidences are
// the stage-2 ones; the class's unmatched stage-1 masks come first with confidence 0.5.
__global__ void __launch_bounds__(kDecT)
refine_select_kernel(RefineLayout L, int mode, double percentile, double given_thres, float iou_thres, int32_t *__restrict__ res) {
    extern __shared__ unsigned char s_first[];       // cap2 flags
    __shared__ int s_scan[34];
    __shared__ double s_thr;
    const int tid = threadIdx.x;
    const int m = res[BFF_RFR_N2_OUT], n_other = res[BFF_RFR_N_OTHER];
    const int32_t *match = res + L.match(), *other = res + L.other();
    const float *match_iou = reinterpret_cast<const float *>(res + L.match_iou());
    const float *conf_out = reinterpret_cast<const float *>(res + L.conf_out());
    const double *sims = reinterpret_cast<const double *>(res + L.sims());
    int32_t *final_src = res + L.final_src();
    float *final_conf = reinterpret_cast<float *>(res + L.final_conf());
    if (tid == 0) s_thr = mode == 2 ? given_thres : 0.0;
    __syncthreads();
    if (mode == 1 && m > 0) {
        // distinct values = first occurrences; the wanted one has exactly `pos` distinct smaller values
        int first = 0;
        for (int k = tid; k < m; k += kDecT) {
            bool f = true;
            for (int j = 0; j < k; ++j) f = f && sims[j] != sims[k];
            s_first[k] = f;
            first += f;
        }
        int n_u;
        block_excl_scan(first, s_scan, &n_u);           // (its barriers also publish s_first)
        const int pos = (int)((double)n_u * percentile);
        if (pos < 0 || pos >= n_u) {
            if (tid == 0) atomicOr(&res[BFF_RFR_ERROR], 1);
        } else {
            for (int k = tid; k < m; k += kDecT) {
                if (!s_first[k]) continue;
                int below = 0;
                for (int j = 0; j < m; ++j) below += s_first[j] && sims[j] < sims[k];
                if (below == pos) s_thr = sims[k];
            }
        }
        __syncthreads();
    }
    const double thr = s_thr;
    for (int i = tid; i < n_other; i += kDecT) {
        final_src[i] = other[i];
        final_conf[i] = 0.5f;
    }
    const int per = (m + kDecT - 1) / kDecT;
    const int lo = min(m, tid * per), hi = min(m, lo + per);
    auto pick = [&](int k) -> int {          // 0 nothing, 1 stage-1 mask, 2 stage-2 mask
        if (match_iou[k] > iou_thres) return sims[k] < thr ? 0 : 1;
        return 2;
    };
    int c = 0;
    for (int k = lo; k < hi; ++k) c += pick(k) != 0;
    int total;
    int pos = n_other + block_excl_scan(c, s_scan, &total);
    for (int k = lo; k < hi; ++k) {
        const int p = pick(k);
        if (p == 0) continue;
        final_src[pos] = p == 1 ? match[k] : L.n1 + k;
        final_conf[pos] = conf_out[k];
        ++pos;
    }
    if (tid == 0) {
        res[BFF_RFR_N_FINAL] = n_other + total;
        const unsigned long long bits = (unsigned long long)__double_as_longlong(thr);
        res[BFF_RFR_SIM_THRES] = (int32_t)(bits & 0xffffffffu);
        res[BFF_RFR_SIM_THRES + 1] = (int32_t)(bits >> 32);
    }
}

// ------------------------------------------------------------------------------------------------ column slices
__global__ void __launch_bounds__(256)
rows_slice_kernel(const uint32_t *__restrict__ rows, long long words, long long word0, long long slice_words, uint32_t *__restrict__ out) {
    const long long w = (long long)blockIdx.x * 256 + threadIdx.x;
    if (w >= slice_words) return;
    const long long src = word0 + w;
    out[(long long)blockIdx.y * slice_words + w] = src < words ? __ldg(rows + (long long)blockIdx.y * words + src) : 0u;
}

__global__ void __launch_bounds__(256)
rows_unslice_kernel(const uint32_t *__restrict__ slices, int n_rows, long long slice_words, long long words, uint32_t *__restrict__ out) {
    const long long w = (long long)blockIdx.x * 256 + threadIdx.x;
    if (w >= words) return;
    const long long p = w / slice_words;
    out[(long long)blockIdx.y * words + w] = __ldg(slices + (p * n_rows + blockIdx.y) * slice_words + (w - p * slice_words));
}

// ------------------------------------------------------------------------------------------------ unpack, device-side row count
__device__ __forceinline__ uint32_t spread4_dc(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

__global__ void __launch_bounds__(256)
unpack_bits_dc_kernel(const uint32_t *__restrict__ packed, const int32_t *__restrict__ n_rows_dev, int cap_rows, long long n, long long words,
                      uint8_t *__restrict__ dense, int wide) {
    const int rows = min(*n_rows_dev, cap_rows);
    const long long total = (long long)rows * n;
    if (wide) {
        // the (rows, n) byte matrix as one flat array, 16 bytes per thread (dense 16-byte aligned)
        const long long j0 = ((long long)blockIdx.x * 256 + threadIdx.x) * 16;
        if (j0 >= total) return;
        const long long row = j0 / n, i0 = j0 - row * n;
        if (i0 + 16 <= n) {
            const uint32_t *src = packed + row * words + (i0 >> 5);
            const int sh = (int)(i0 & 31);
            const uint32_t lo = __ldg(src);
            const uint32_t hi = sh > 16 ? __ldg(src + 1) : 0u;
            const uint32_t v = __funnelshift_r(lo, hi, sh) & 0xffffu;
            reinterpret_cast<uint4 *>(dense)[j0 >> 4] = make_uint4(spread4_dc(v & 15u), spread4_dc((v >> 4) & 15u), spread4_dc((v >> 8) & 15u), spread4_dc(v >> 12));
            return;
        }
        for (long long j = j0; j < j0 + 16 && j < total; ++j) {
            const long long r = j / n, i = j - r * n;
            dense[j] = (uint8_t)((__ldg(packed + r * words + (i >> 5)) >> (i & 31)) & 1u);
        }
    } else {
        const long long stride = (long long)gridDim.x * 256;
        for (long long j = (long long)blockIdx.x * 256 + threadIdx.x; j < total; j += stride) {
            const long long r = j / n, i = j - r * n;
            dense[j] = (uint8_t)((__ldg(packed + r * words + (i >> 5)) >> (i & 31)) & 1u);
        }
    }
}

}  // namespace bff

using namespace bff;

// ------------------------------------------------------------------------------------------------ host side
namespace {

struct Take {                                   // two-pass carving of one workspace block
    unsigned char *base;
    size_t off = 0;
    explicit Take(unsigned char *b) : base(b) {}
    template <typename T>
    T *get(size_t count) {
        off = (off + 255) & ~(size_t)255;
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += (count ? count : 1) * sizeof(T);
        return p;
    }
};

unsigned loop_rows(int cap) { return (unsigned)std::max(1, std::min(cap, kLoopRows)); }

int launch_intersect_dc(const uint32_t *a, const int32_t *n_a_dev, int n_a_host, int cap_a, const uint32_t *b, const int32_t *n_b_dev,
                        int n_b_host, long long words, int ld, int32_t *inter, int32_t *pop_a, int32_t *pop_b, cudaStream_t s) {
    if (words <= 0 || cap_a <= 0) return BFF_OK;
    const long long chunks = ceil_div64(words, kDcChunk);
    long long slices = std::min<long long>(chunks, kNumSMs);
    const long long wps = ceil_div64(chunks, slices) * kDcChunk;
    slices = ceil_div64(words, wps);
    dim3 grid((unsigned)std::min<long long>(ceil_div64(cap_a, kDcRows), 64), (unsigned)slices);
    intersect_dc_kernel<<<grid, kDcRows * 32, 0, s>>>(a, n_a_dev, n_a_host, b, n_b_dev, n_b_host, words, wps, ld, inter, pop_a, pop_b);
    BFF_LAUNCH_CHECK("intersect_dc_kernel");
    return BFF_OK;
}

}  // namespace

extern "C" int bff_stage2_device(bff_ctx *c, const bff_stage2_dev_desc *d, void *stream_) {
    cudaStream_t s = (cudaStream_t)stream_;
    BFF_REQUIRE(c && d, "bff_stage2_device: null argument");
    BFF_REQUIRE(d->n_rows > 0 && d->n_points > 0, "bff_stage2_device: empty input (the caller handles 'nothing lifted')");
    BFF_REQUIRE(d->rows && d->conf && d->mc && d->out_rows && d->result, "bff_stage2_device: null pointer");
    BFF_REQUIRE(d->min_aggregated >= 1, "bff_stage2_device: min_aggregated must be >= 1");
    BFF_REQUIRE(d->filter_mode >= 0 && d->filter_mode <= 2 && (d->filter_mode != 0 || d->vc), "bff_stage2_device: bad filter mode");
    BFF_REQUIRE(d->n_rows <= 25000, "bff_stage2_device: too many rows (the component search keeps 8 bytes per row in shared memory)");
    BFF_REQUIRE(d->cap_groups >= d->n_rows / d->min_aggregated && d->cap_groups >= 1, "bff_stage2_device: cap_groups %d below n_rows / min_aggregated", d->cap_groups);
    BFF_REQUIRE(d->n_vote_points >= 0 && d->vote_point0 >= 0, "bff_stage2_device: negative vote offsets");
    BFF_REQUIRE(d->mc_cap >= 0 && d->vc_cap >= 0, "bff_stage2_device: negative count bounds");
    BFF_CUDA(cudaSetDevice(c->device));
    const int n = d->n_rows, cap = d->cap_groups;
    const long long words = ceil_div64(d->n_points, 32);
    const int stride_words = ((n + 63) / 64) * 2;
    const long long n_votes = d->n_vote_points > 0 ? d->n_vote_points : d->n_points;
    const long long hist_cap = d->filter_mode == 2 ? 1 : ((long long)d->mc_cap + 1) * (d->filter_mode == 0 ? (long long)d->vc_cap + 1 : 1);
    long long table_cap = 1024;
    while (table_cap < 2 * hist_cap) table_cap <<= 1;
    BFF_REQUIRE(hist_cap < (1LL << 30), "bff_stage2_device: vote histogram bound too large");
    const size_t res_words = (size_t)bff_stage2_dev_result_words(n, cap);
    const long long prog_cap = std::max<long long>(1, (long long)cap * (cap - 1) / 2);

    int32_t *d_inter = nullptr, *d_pop = nullptr, *d_scratch = nullptr, *d_ginter = nullptr, *d_prog = nullptr, *d_max = nullptr;
    uint32_t *d_adj = nullptr, *d_agg = nullptr, *d_hist = nullptr, *d_table = nullptr, *d_keep = nullptr, *d_vals = nullptr;
    int32_t *d_nvals = nullptr;
    float *d_thr = nullptr;
    for (int pass = 0; pass < 2; ++pass) {
        Take t(pass ? c->dev.base() : nullptr);
        d_inter = t.get<int32_t>((size_t)n * n + n);          // intersections, then the row popcounts (one all-reduce)
        d_adj = t.get<uint32_t>((size_t)n * stride_words);
        d_scratch = t.get<int32_t>(2 * (size_t)n + (size_t)n * stride_words);
        d_agg = t.get<uint32_t>((size_t)cap * words);
        d_ginter = t.get<int32_t>((size_t)cap * cap);
        d_prog = t.get<int32_t>(2 * (size_t)prog_cap);
        d_max = t.get<int32_t>(2);
        d_hist = t.get<uint32_t>((size_t)hist_cap);
        d_table = t.get<uint32_t>((size_t)table_cap);
        d_vals = t.get<uint32_t>((size_t)hist_cap);
        d_nvals = t.get<int32_t>(1);
        d_thr = t.get<float>(1);
        d_keep = t.get<uint32_t>((size_t)words);
        if (pass == 0) {
            const int rc = c->dev.reserve(t.off + 256);
            if (rc != BFF_OK) return rc;
        }
    }
    d_pop = d_inter + (size_t)n * n;
    int32_t *res = d->result;
    int32_t *r_before = res + BFF_S2R_HEADER + 3 * (size_t)cap, *r_after = r_before + cap;
    int32_t *r_selected = res + BFF_S2R_HEADER + 5 * (size_t)cap;
    int32_t *r_seg = res + BFF_S2R_HEADER + 7 * (size_t)cap, *r_members = r_seg + cap + 1;
    int rc;
    BFF_CUDA(cudaMemsetAsync(res, 0, res_words * 4, s));

    // ---- point filter (:512-583) on the side stream: maxima -> histogram -> distinct values -> threshold -> keep bits.
    // Nothing of it depends on the grouping, so it runs next to the intersection chain (BFF_DEVPIPE_SIDE=0: in line).
    const int32_t *vc = d->filter_mode == 0 ? d->vc : nullptr;
    static const bool use_side = [] { const char *e = getenv("BFF_DEVPIPE_SIDE"); return !(e && e[0] == '0'); }();
    cudaStream_t vs = s;
    if (use_side) {
        if (!c->has_side) {
            BFF_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
            BFF_CUDA(cudaEventCreateWithFlags(&c->side_fork, cudaEventDisableTiming));
            BFF_CUDA(cudaEventCreateWithFlags(&c->side_join, cudaEventDisableTiming));
            c->has_side = 1;
        }
        vs = c->side;
        BFF_CUDA(cudaEventRecord(c->side_fork, s));
        BFF_CUDA(cudaStreamWaitEvent(vs, c->side_fork, 0));
    }
    if (d->filter_mode != 2) {
        if ((rc = bff_counts_max(d->mc, vc, n_votes, d_max, vs)) != BFF_OK) return rc;
        vote_prepare_kernel<<<kNumSMs, 256, 0, vs>>>(d_max, d->filter_mode, hist_cap, table_cap, d_hist, d_table, d_nvals, res);
        long long blocks = std::min<long long>(ceil_div64(n_votes, 256), kNumSMs * 4);
        vote_hist_dc_kernel<<<(unsigned)std::max<long long>(blocks, 1), 256, 0, vs>>>(d->mc, vc, n_votes, d_max, d->filter_mode, hist_cap, table_cap, d_hist);
        vote_unique_insert_dc_kernel<<<kNumSMs, 256, 0, vs>>>(d_hist, d_max, d->filter_mode, hist_cap, table_cap, d_table);
        vote_compact_dc_kernel<<<kNumSMs, 256, 0, vs>>>(d_table, d_max, d->filter_mode, hist_cap, table_cap, d_vals, d_nvals);
        vote_select_kernel<<<1, kDecT, 0, vs>>>(d_vals, d_nvals, d->filter_fraction, d_thr, res);
        BFF_LAUNCH_CHECK("vote threshold kernels");
    } else {
        BFF_CUDA(cudaMemsetAsync(d_thr, 0, 4, vs));
    }
    point_keep_dc_kernel<<<(unsigned)ceil_div64(words * 32, 256), 256, 0, vs>>>(d->mc + d->vote_point0, vc ? vc + d->vote_point0 : nullptr, d->n_points,
                                                                                words, d->filter_mode, d_thr, d_keep);
    BFF_LAUNCH_CHECK("point_keep_dc_kernel");
    if (use_side) BFF_CUDA(cudaEventRecord(c->side_join, vs));

    // ---- all-pairs intersections -> thresholded merge graph (calculate_iou :149-166, aggregate :113-122)
    if ((rc = bff_mask_iou(d->rows, n, d->rows, n, words, d_inter, d_pop, d_pop, nullptr, s)) != BFF_OK) return rc;
    if ((rc = comm_allreduce_sum_i32(d->comm, d_inter, (size_t)n * n + n, s)) != BFF_OK) return rc;
    if ((rc = launch_merge_adjacency(d_inter, d_pop, d->label_id, n, stride_words, d->iou_thres, d_adj, s)) != BFF_OK) return rc;
    // ---- connected components, groups, confidences (:190-274)
    {
        size_t smem = (size_t)n * 8;                         // labels + confidences
        BFF_REQUIRE(smem <= 200 * 1024, "bff_stage2_device: %d rows need %zu bytes of shared memory for the component search", n, smem);
        const size_t graph_bytes = 2 * (size_t)n * stride_words * 4;      // adjacency + member rows
        const int stage_graph = smem + graph_bytes <= 160 * 1024;
        if (stage_graph) smem += graph_bytes;
        if (smem > 48 * 1024) BFF_CUDA(cudaFuncSetAttribute((const void *)components_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        components_kernel<<<1, kDecT, smem, s>>>(d_adj, n, stride_words, d->min_aggregated, d->conf, cap, res, d_scratch, stage_graph);
        BFF_LAUNCH_CHECK("components_kernel");
    }
    // ---- OR-merge the members (:214-223)
    {
        dim3 grid((unsigned)ceil_div64(words, 256), loop_rows(cap));
        rows_or_reduce_dc_kernel<<<grid, 256, 0, s>>>(d->rows, words, r_members, r_seg, res + BFF_S2R_N_GROUPS, d_agg);
        BFF_LAUNCH_CHECK("rows_or_reduce_dc_kernel");
    }
    // ---- pair overlaps of the merged rows + row sums 'before' (:285-291, :592): the sums land in the result block
    BFF_CUDA(cudaMemsetAsync(d_ginter, 0, (size_t)cap * cap * 4, s));
    if ((rc = launch_intersect_dc(d_agg, res + BFF_S2R_N_GROUPS, 0, cap, d_agg, res + BFF_S2R_N_GROUPS, 0, words, cap, d_ginter, r_before, nullptr, s)) != BFF_OK)
        return rc;
    if (d->comm) {
        if ((rc = comm_allreduce_sum_i32(d->comm, d_ginter, (size_t)cap * cap, s)) != BFF_OK) return rc;
        if ((rc = comm_allreduce_sum_i32(d->comm, r_before, (size_t)cap, s)) != BFF_OK) return rc;
    }
    overlap_program_kernel<<<1, kDecT, 0, s>>>(d_ginter, cap, cap, res, d_prog);
    BFF_LAUNCH_CHECK("overlap_program_kernel");
    row_program_dc_kernel<false><<<(unsigned)ceil_div64(words, 128), 128, 0, s>>>(d_agg, words, d_prog, res + BFF_S2R_N_OVERLAP, res + BFF_S2R_N_GROUPS);
    BFF_LAUNCH_CHECK("row_program_dc_kernel");
    // ---- ins &= masked_points, row sums 'after' (:594-596), row filter (:601-623), selected rows
    if (use_side) BFF_CUDA(cudaStreamWaitEvent(s, c->side_join, 0));
    {
        dim3 grid((unsigned)ceil_div64(words, kAndPopWords), loop_rows(cap));
        rows_and_popcount_dc_kernel<<<grid, 256, 0, s>>>(d_agg, words, d_keep, res + BFF_S2R_N_GROUPS, r_after);
        BFF_LAUNCH_CHECK("rows_and_popcount_dc_kernel");
    }
    if ((rc = comm_allreduce_sum_i32(d->comm, r_after, (size_t)cap, s)) != BFF_OK) return rc;
    RowFilter f;
    f.small_is_float = d->small_is_float; f.ratio_is_float = d->ratio_is_float;
    f.small_i = d->small_i; f.ratio_i = d->ratio_i; f.small_f = d->small_f; f.ratio_f = d->ratio_f;
    row_filter_kernel<<<1, kDecT, 0, s>>>(cap, f, res);
    BFF_LAUNCH_CHECK("row_filter_kernel");
    {
        dim3 grid((unsigned)ceil_div64(words, 256), loop_rows(cap));
        rows_gather_dc_kernel<<<grid, 256, 0, s>>>(d_agg, words, r_selected, res + BFF_S2R_N_SELECTED, 0, d->out_rows);
        BFF_LAUNCH_CHECK("rows_gather_dc_kernel");
    }
    return BFF_OK;
}

namespace {
int refine_check(const bff_refine_dev_desc *d, const char *who) {
    BFF_REQUIRE(d, "%s: null descriptor", who);
    BFF_REQUIRE(d->n1 > 0 && d->cap2 > 0 && d->n_points > 0, "%s: empty input", who);
    BFF_REQUIRE(d->s1_rows && d->s2_rows && d->s2_conf && d->is_query && d->sim && d->s2_out && d->result, "%s: null pointer", who);
    BFF_REQUIRE(d->n2_dev || (d->n2_host >= 0 && d->n2_host <= d->cap2), "%s: n2_host outside [0, cap2]", who);
    BFF_REQUIRE(d->select_mode >= 0 && d->select_mode <= 2, "%s: bad select_mode", who);
    BFF_REQUIRE(d->select_mode == 0 || d->final_rows, "%s: final_rows missing", who);
    BFF_REQUIRE(((uintptr_t)d->result & 7) == 0, "%s: result block must be 8-byte aligned", who);
    return BFF_OK;
}

int refine_select(const bff_refine_dev_desc *d, int mode, double thres, cudaStream_t s) {
    const RefineLayout L{d->n1, d->cap2};
    const long long words = ceil_div64(d->n_points, 32);
    BFF_REQUIRE(d->cap2 <= 40960, "bff_refine_device: cap2 too large");
    refine_select_kernel<<<1, kDecT, (size_t)d->cap2, s>>>(L, mode, d->sim_percentile, thres, d->refine_iou_thres, d->result);
    BFF_LAUNCH_CHECK("refine_select_kernel");
    dim3 grid((unsigned)ceil_div64(words, 256), loop_rows(d->n1 + d->cap2));
    rows_gather2_dc_kernel<<<grid, 256, 0, s>>>(d->s1_rows, d->n1, d->s2_out, words, d->result + L.final_src(), d->result + BFF_RFR_N_FINAL, d->final_rows);
    BFF_LAUNCH_CHECK("rows_gather2_dc_kernel");
    return BFF_OK;
}
}  // namespace

extern "C" int bff_refine_device(bff_ctx *c, const bff_refine_dev_desc *d, void *stream_) {
    cudaStream_t s = (cudaStream_t)stream_;
    BFF_REQUIRE(c, "bff_refine_device: null context");
    int rc = refine_check(d, "bff_refine_device");
    if (rc != BFF_OK) return rc;
    BFF_CUDA(cudaSetDevice(c->device));
    const int n1 = d->n1, cap2 = d->cap2;
    const long long words = ceil_div64(d->n_points, 32);
    const RefineLayout L{n1, cap2};
    const size_t res_words = (size_t)bff_refine_dev_result_words(n1, cap2);
    // [inter (n1 x cap2) | pop1 (n1) | pop2 (cap2)] and [ginter (cap2 x cap2) | gpop (cap2)]: each one all-reduce
    const size_t iou_words = (size_t)n1 * cap2 + n1 + cap2, giou_words = (size_t)cap2 * cap2 + cap2;
    int32_t *d_iou = nullptr, *d_giou = nullptr, *d_n2 = nullptr, *d_prog = nullptr, *d_scratch = nullptr;
    uint32_t *d_sel = nullptr;
    for (int pass = 0; pass < 2; ++pass) {
        Take t(pass ? c->dev.base() : nullptr);
        d_iou = t.get<int32_t>(iou_words);
        d_giou = t.get<int32_t>(giou_words);
        d_n2 = t.get<int32_t>(1);
        d_prog = t.get<int32_t>(2 * (size_t)cap2 * cap2);
        d_scratch = t.get<int32_t>((size_t)n1);
        d_sel = t.get<uint32_t>((size_t)cap2 * words);
        if (pass == 0 && (rc = c->dev.reserve(t.off + 256)) != BFF_OK) return rc;
    }
    int32_t *d_inter = d_iou, *d_p1 = d_iou + (size_t)n1 * cap2, *d_p2 = d_p1 + n1;
    int32_t *d_ginter = d_giou, *d_gpop = d_giou + (size_t)cap2 * cap2;
    int32_t *res = d->result;
    BFF_CUDA(cudaMemsetAsync(res, 0, res_words * 4, s));
    const int32_t *n2 = d->n2_dev;
    if (!n2) {
        // the count travels through the result block so that every kernel reads it from the device
        if ((rc = launch_set_i32(d_n2, d->n2_host, s)) != BFF_OK) return rc;
        n2 = d_n2;
    }
    // ---- iou(stage 1, stage 2) and max_match = argmax over stage 1 (:208-213)
    BFF_CUDA(cudaMemsetAsync(d_iou, 0, iou_words * 4, s));
    if ((rc = launch_intersect_dc(d->s1_rows, nullptr, n1, n1, d->s2_rows, n2, 0, words, cap2, d_inter, d_p1, d_p2, s)) != BFF_OK) return rc;
    if ((rc = comm_allreduce_sum_i32(d->comm, d_iou, iou_words, s)) != BFF_OK) return rc;
    const unsigned arg_blocks = (unsigned)std::max<long long>(1, std::min<long long>(ceil_div64(cap2, 8), kNumSMs));
    argmax_cols_kernel<<<arg_blocks, 256, 0, s>>>(d_inter, cap2, d_p1, d_p2, n1, n2, res + L.first_match(), nullptr);
    BFF_LAUNCH_CHECK("argmax_cols_kernel");
    // ---- IoU among the matched stage-1 masks (:217-224)
    {
        dim3 grid((unsigned)ceil_div64(words, 256), loop_rows(cap2));
        rows_gather_dc_kernel<<<grid, 256, 0, s>>>(d->s1_rows, words, res + L.first_match(), n2, 0, d_sel);
        BFF_LAUNCH_CHECK("rows_gather_dc_kernel");
    }
    BFF_CUDA(cudaMemsetAsync(d_giou, 0, giou_words * 4, s));
    if ((rc = launch_intersect_dc(d_sel, n2, 0, cap2, d_sel, n2, 0, words, cap2, d_ginter, d_gpop, nullptr, s)) != BFF_OK) return rc;
    if ((rc = comm_allreduce_sum_i32(d->comm, d_giou, giou_words, s)) != BFF_OK) return rc;
    // ---- the sequential merge loop (:230-249) and the stage-2 merges (:258-281)
    {
        const size_t smem = ((size_t)cap2 + kMergeStage) * 4;
        BFF_REQUIRE(smem <= 200 * 1024, "bff_refine_device: cap2 too large");
        if (smem > 40 * 1024) BFF_CUDA(cudaFuncSetAttribute((const void *)refine_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        refine_merge_kernel<<<1, kDecT, smem, s>>>(L, n2, d_ginter, cap2, d_gpop, d->stage1_iou_thres, d->s2_conf, res, d_prog, d_scratch);
        BFF_LAUNCH_CHECK("refine_merge_kernel");
    }
    row_program_dc_kernel<true><<<(unsigned)ceil_div64(words, 128), 128, 0, s>>>(d->s1_rows, words, d_prog, res + BFF_RFR_N_OR, nullptr);
    BFF_LAUNCH_CHECK("row_program_dc_kernel");
    {
        dim3 grid((unsigned)ceil_div64(words, 256), loop_rows(cap2));
        rows_or_reduce_dc_kernel<<<grid, 256, 0, s>>>(d->s2_rows, words, res + L.idx2(), res + L.seg2(), res + BFF_RFR_N2_OUT, d->s2_out);
        BFF_LAUNCH_CHECK("rows_or_reduce_dc_kernel");
    }
    // ---- IoU of the merged stage-1 masks with the merged stage-2 masks, final argmax (:285-290, :308)
    BFF_CUDA(cudaMemsetAsync(d_iou, 0, iou_words * 4, s));
    if ((rc = launch_intersect_dc(d->s1_rows, nullptr, n1, n1, d->s2_out, res + BFF_RFR_N2_OUT, 0, words, cap2, d_inter, d_p1, d_p2, s)) != BFF_OK) return rc;
    if ((rc = comm_allreduce_sum_i32(d->comm, d_iou, iou_words, s)) != BFF_OK) return rc;
    argmax_cols_kernel<<<arg_blocks, 256, 0, s>>>(d_inter, cap2, d_p1, d_p2, n1, res + BFF_RFR_N2_OUT, res + L.match(),
                                                  reinterpret_cast<float *>(res + L.match_iou()));
    BFF_LAUNCH_CHECK("argmax_cols_kernel");
    refine_match_info_kernel<<<1, kDecT, 0, s>>>(L, d->is_query, d->sim, res, d_scratch);
    BFF_LAUNCH_CHECK("refine_match_info_kernel");
    if (d->select_mode != 0) return refine_select(d, d->select_mode, d->sim_thres, s);
    return BFF_OK;
}

extern "C" int bff_refine_select_device(bff_ctx *c, const bff_refine_dev_desc *d, double sim_thres, void *stream_) {
    BFF_REQUIRE(c, "bff_refine_select_device: null context");
    int rc = refine_check(d, "bff_refine_select_device");
    if (rc != BFF_OK) return rc;
    BFF_REQUIRE(d->final_rows, "bff_refine_select_device: final_rows missing");
    BFF_CUDA(cudaSetDevice(c->device));
    return refine_select(d, 2, sim_thres, (cudaStream_t)stream_);
}

extern "C" int bff_rows_gather_dev(const uint32_t *rows, int64_t words, const int32_t *idx_dev, int32_t n_idx, uint32_t *out, void *stream_) {
    BFF_REQUIRE(n_idx >= 0 && words >= 0, "bff_rows_gather_dev: bad arguments");
    if (n_idx == 0 || words == 0) return BFF_OK;
    BFF_REQUIRE(rows && idx_dev && out, "bff_rows_gather_dev: null pointer");
    dim3 grid((unsigned)ceil_div64(words, 256), (unsigned)std::min(n_idx, 4096));
    rows_gather_dc_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(rows, words, idx_dev, nullptr, n_idx, out);
    BFF_LAUNCH_CHECK("rows_gather_dc_kernel");
    return BFF_OK;
}

extern "C" int bff_unpack_bits_dc(const uint32_t *packed, const int32_t *n_rows_dev, int32_t cap_rows, int64_t n, uint8_t *dense, void *stream_) {
    BFF_REQUIRE(cap_rows >= 0 && n >= 0, "bff_unpack_bits_dc: negative size");
    if (cap_rows == 0 || n == 0) return BFF_OK;
    BFF_REQUIRE(packed && n_rows_dev && dense, "bff_unpack_bits_dc: null pointer");
    const long long words = ceil_div64(n, 32);
    const long long total = (long long)cap_rows * n;
    const int wide = (reinterpret_cast<uintptr_t>(dense) & 15) == 0 && n >= 16 && ceil_div64(total, 16 * 256) <= 0x7fffffffLL;
    const long long blocks = wide ? ceil_div64(total, 16 * 256) : std::min<long long>(ceil_div64(total, 256), (long long)kNumSMs * 32);
    unpack_bits_dc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream_>>>(packed, n_rows_dev, cap_rows, n, words, dense, wide);
    BFF_LAUNCH_CHECK("unpack_bits_dc_kernel");
    return BFF_OK;
}

extern "C" int bff_rows_slice(const uint32_t *rows, int32_t n_rows, int64_t words, int64_t word0, int64_t slice_words, uint32_t *out, void *stream_) {
    BFF_REQUIRE(n_rows >= 0 && words >= 0 && word0 >= 0 && slice_words >= 0, "bff_rows_slice: negative size");
    if (n_rows == 0 || slice_words == 0) return BFF_OK;
    BFF_REQUIRE(rows && out, "bff_rows_slice: null pointer");
    for (int32_t r0 = 0; r0 < n_rows; r0 += kMaxGridY) {
        dim3 grid((unsigned)ceil_div64(slice_words, 256), (unsigned)std::min(n_rows - r0, kMaxGridY));
        rows_slice_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(rows + (size_t)r0 * words, words, word0, slice_words, out + (size_t)r0 * slice_words);
        BFF_LAUNCH_CHECK("rows_slice_kernel");
    }
    return BFF_OK;
}

extern "C" int bff_rows_unslice(const uint32_t *slices, int32_t world, int32_t n_rows, int64_t slice_words, int64_t words, uint32_t *out, void *stream_) {
    BFF_REQUIRE(world >= 1 && n_rows >= 0 && slice_words > 0 && words >= 0 && (int64_t)world * slice_words >= words, "bff_rows_unslice: bad sizes");
    if (n_rows == 0 || words == 0) return BFF_OK;
    BFF_REQUIRE(slices && out, "bff_rows_unslice: null pointer");
    for (int32_t r0 = 0; r0 < n_rows; r0 += kMaxGridY) {       // n_rows stays the rank stride of `slices`; only the first row moves
        dim3 grid((unsigned)ceil_div64(words, 256), (unsigned)std::min(n_rows - r0, kMaxGridY));
        rows_unslice_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(slices + (size_t)r0 * slice_words, n_rows, slice_words, words, out + (size_t)r0 * words);
        BFF_LAUNCH_CHECK("rows_unslice_kernel");
    }
    return BFF_OK;
}
