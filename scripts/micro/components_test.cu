// Experiment for DESIGN.md 7b item 1 (device-side decisions): connected components of the merge graph on the GPU,
// in the reference's emission order, so that bff_stage2_aggregate needs no host round trip for them.
//
// Input: the bit adjacency merge_adjacency_kernel writes (row r = stride_words 32-bit words, bit c = edge (r, c)).
// Output: root[i] = smallest member of i's component (what the host flood fill of pipeline.cu computes), and from it
// the kept groups (root has an edge, size >= min_aggregated) in order of their smallest member, members ascending,
// with the sequential float32 confidence mean of the members (projection_2d_to_3d.py:203-226, :250-274).
//
// One CTA: min-label propagation over the bit rows (a warp per node, lanes over the row's words) with pointer jumping,
// until nothing changes; then a scan over the roots. NOT wired into the product; stand-alone check against the host
// algorithm + timing. First measurement (B200, random sparse graphs with one giant component, i.e. a far larger diameter
// than the near-cliques of real merge graphs): bit-exact; 0.17 ms at n = 525, 0.45 ms at n = 1351, 2.1 ms at n = 4299 --
// slower than the host flood fill + round trip (~35 us at n = 525) until the propagation and the per-group member
// walk are spread over more than one CTA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o components_test components_test.cu
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int kThreads = 1024;

__global__ void __launch_bounds__(kThreads)
components_kernel(const uint32_t *__restrict__ adj, int n, int stride_words, int min_aggregated, const float *__restrict__ conf,
                  int32_t *__restrict__ label /* n, scratch + output root[] */, int32_t *__restrict__ n_groups,
                  int32_t *__restrict__ group_first, int32_t *__restrict__ group_size, float *__restrict__ group_conf,
                  int32_t *__restrict__ members, int32_t *__restrict__ seg) {
    __shared__ int s_changed;
    __shared__ int s_scan[kThreads];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, n_warps = kThreads / 32;
    for (int i = tid; i < n; i += kThreads) label[i] = i;
    __syncthreads();
    // ---- min-label propagation: label[i] <- min(label[i], label[j] for every edge (i, j)); in-place updates are fine
    // (labels only decrease and stay member indices of the same component)
    while (true) {
        if (tid == 0) s_changed = 0;
        __syncthreads();
        for (int i = warp; i < n; i += n_warps) {
            const uint32_t *row = adj + (size_t)i * stride_words;
            int best = 0x7fffffff;
            for (int w = lane; w < stride_words; w += 32) {
                uint32_t m = row[w];
                while (m) {
                    const int j = w * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    if (j < n) best = min(best, ((volatile int32_t *)label)[j]);
                }
            }
            for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
            if (lane == 0) {
                const int cur = ((volatile int32_t *)label)[i];
                if (best < cur) { label[i] = best; s_changed = 1; }
            }
        }
        __syncthreads();
        // pointer jumping: label[label[i]] is a member of the same component and never larger
        for (int i = tid; i < n; i += kThreads) {
            int l = label[i];
            int ll = ((volatile int32_t *)label)[l];
            while (ll < l) { l = ll; ll = ((volatile int32_t *)label)[l]; }
            label[i] = l;
        }
        __syncthreads();
        if (!s_changed) break;
        __syncthreads();
    }
    // ---- group sizes: count[root] (reuse `seg` as scratch of n+1 ints is not possible, it holds g+1; use group_size as n scratch)
    // group_size / group_first have capacity n here (the harness allocates n), the product would use cap_groups after compaction
    for (int i = tid; i < n; i += kThreads) group_size[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kThreads) atomicAdd(&group_size[label[i]], 1);
    __syncthreads();
    // ---- kept roots in ascending order: exclusive scan of the keep flags over node index (chunks of kThreads)
    __shared__ int s_base;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int c0 = 0; c0 < n; c0 += kThreads) {
        const int i = c0 + tid;
        int keep = 0;
        if (i < n && label[i] == i && group_size[i] >= min_aggregated) {
            const uint32_t *row = adj + (size_t)i * stride_words;     // a node without any edge (empty mask) forms no group
            uint32_t any = 0;
            for (int w = 0; w < stride_words; ++w) any |= row[w];
            keep = any != 0;
        }
        s_scan[tid] = keep;
        __syncthreads();
        for (int o = 1; o < kThreads; o <<= 1) {                      // Hillis-Steele inclusive scan
            const int v = tid >= o ? s_scan[tid - o] : 0;
            __syncthreads();
            s_scan[tid] += v;
            __syncthreads();
        }
        const int pos = s_base + s_scan[tid] - keep;
        if (keep) group_first[pos] = i;                                // smallest member = the root itself
        __syncthreads();
        if (tid == kThreads - 1) s_base += s_scan[tid];
        __syncthreads();
    }
    const int g = s_base;
    if (tid == 0) *n_groups = g;
    // ---- members ascending + sequential float32 mean, one thread per group (g is small: tens)
    __syncthreads();
    if (tid == 0) {
        int off = 0;
        for (int k = 0; k < g; ++k) { seg[k] = off; off += group_size[group_first[k]]; }
        seg[g] = off;
    }
    __syncthreads();
    for (int k = tid; k < g; k += kThreads) {
        const int r = group_first[k];
        int off = seg[k];
        float acc = 0.0f;
        for (int i = 0; i < n; ++i)
            if (label[i] == r) { members[off++] = i; acc = __fadd_rn(acc, conf[i]); }
        group_conf[k] = __fdiv_rn(acc, (float)(off - seg[k]));
    }
}

// the host algorithm of pipeline.cu (agg_phase2_groups), on 32-bit words
static void host_groups(const std::vector<uint32_t> &adj, int n, int sw, int min_agg, const std::vector<float> &conf,
                        std::vector<int> &root, std::vector<int> &first, std::vector<int> &members, std::vector<float> &gconf) {
    root.assign(n, -1);
    std::vector<char> vis(n, 0);
    for (int i = 0; i < n; ++i) {
        if (vis[i]) continue;
        std::vector<int> stack{i};
        vis[i] = 1;
        while (!stack.empty()) {
            const int u = stack.back();
            stack.pop_back();
            root[u] = i;
            for (int j = 0; j < n; ++j)
                if (!vis[j] && ((adj[(size_t)u * sw + (j >> 5)] >> (j & 31)) & 1)) { vis[j] = 1; stack.push_back(j); }
        }
    }
    std::vector<int> count(n, 0);
    for (int i = 0; i < n; ++i) count[root[i]]++;
    first.clear(); members.clear(); gconf.clear();
    for (int r = 0; r < n; ++r) {
        bool any = false;
        for (int w = 0; w < sw; ++w) any |= adj[(size_t)r * sw + w] != 0;
        if (root[r] != r || !any || count[r] < min_agg) continue;
        first.push_back(r);
        float acc = 0.0f;
        int c = 0;
        for (int i = 0; i < n; ++i)
            if (root[i] == r) { members.push_back(i); acc = acc + conf[i]; ++c; }
        gconf.push_back(acc / (float)c);
    }
}

int main(int argc, char **argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 525;
    const double p_edge = argc > 2 ? atof(argv[2]) : 0.004;
    const int min_agg = argc > 3 ? atoi(argv[3]) : 3;
    const int sw = ((n + 63) / 64) * 2;
    long long bad_total = 0;
    for (int seed = 1; seed <= 5; ++seed) {
        srand(seed);
        std::vector<uint32_t> adj((size_t)n * sw, 0);
        std::vector<float> conf(n);
        for (int i = 0; i < n; ++i) {
            conf[i] = 0.2f + 0.25f * (float)rand() / (float)RAND_MAX;
            const bool empty = rand() % 17 == 0;                       // an empty lifted row: no self edge, no edges at all
            if (!empty) adj[(size_t)i * sw + (i >> 5)] |= 1u << (i & 31);
            for (int j = 0; j < i && !empty; ++j)
                if ((double)rand() / RAND_MAX < p_edge && (adj[(size_t)j * sw + (j >> 5)] >> (j & 31) & 1)) {
                    adj[(size_t)i * sw + (j >> 5)] |= 1u << (j & 31);
                    adj[(size_t)j * sw + (i >> 5)] |= 1u << (i & 31);
                }
        }
        std::vector<int> root, first, members;
        std::vector<float> gconf;
        host_groups(adj, n, sw, min_agg, conf, root, first, members, gconf);
        uint32_t *d_adj; float *d_conf, *d_gconf; int32_t *d_label, *d_ng, *d_first, *d_size, *d_members, *d_seg;
        cudaMalloc(&d_adj, adj.size() * 4); cudaMalloc(&d_conf, n * 4); cudaMalloc(&d_gconf, n * 4);
        cudaMalloc(&d_label, n * 4); cudaMalloc(&d_ng, 4); cudaMalloc(&d_first, n * 4); cudaMalloc(&d_size, n * 4);
        cudaMalloc(&d_members, n * 4); cudaMalloc(&d_seg, (n + 1) * 4);
        cudaMemcpy(d_adj, adj.data(), adj.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d_conf, conf.data(), n * 4, cudaMemcpyHostToDevice);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        float ms = 0;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            components_kernel<<<1, kThreads>>>(d_adj, n, sw, min_agg, d_conf, d_label, d_ng, d_first, d_size, d_gconf, d_members, d_seg);
            cudaEventRecord(e1);
            if (cudaEventSynchronize(e1) != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 2; }
            cudaEventElapsedTime(&ms, e0, e1);
        }
        int g = 0;
        cudaMemcpy(&g, d_ng, 4, cudaMemcpyDeviceToHost);
        std::vector<int32_t> h_label(n), h_first(n), h_members(n), h_seg(n + 1);
        std::vector<float> h_gconf(n);
        cudaMemcpy(h_label.data(), d_label, n * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(h_first.data(), d_first, n * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(h_members.data(), d_members, n * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(h_seg.data(), d_seg, (n + 1) * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(h_gconf.data(), d_gconf, n * 4, cudaMemcpyDeviceToHost);
        long long bad = 0;
        for (int i = 0; i < n; ++i) bad += h_label[i] != root[i];
        bad += g != (int)first.size();
        for (int k = 0; k < g && k < (int)first.size(); ++k) bad += h_first[k] != first[k] || h_gconf[k] != gconf[k];
        if (g == (int)first.size() && h_seg[g] == (int)members.size())
            for (size_t k = 0; k < members.size(); ++k) bad += h_members[k] != members[k];
        else
            ++bad;
        printf("seed %d: n=%d groups=%d (host %zu) members=%zu  %.1f us  mismatches=%lld\n", seed, n, g, first.size(), members.size(), ms * 1e3, bad);
        bad_total += bad;
        cudaFree(d_adj); cudaFree(d_conf); cudaFree(d_gconf); cudaFree(d_label); cudaFree(d_ng); cudaFree(d_first); cudaFree(d_size);
        cudaFree(d_members); cudaFree(d_seg);
    }
    return bad_total != 0;
}
