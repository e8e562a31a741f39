// Fused host+device pipelines (include/bff.h: bff_stage2_aggregate, bff_refine_match, bff_rows_gather).
//
// The reference takes its small decisions (which masks merge, which threshold, which pairs overlap, which
// stage-1 mask matches) with torch ops on tiny tensors between heavy tensor ops. Here the heavy ops are the
// kernels of masks.cu / votes.cu and the decisions are plain C++ on pinned host copies of the tiny matrices,
// so one call replaces dozens of Python-level launches. Every decision cites the reference lines it mirrors;
// float32 comparisons / divisions are done in float32 exactly like torch does them.
#include <math.h>
#include <string.h>

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <chrono>
#include <vector>

#include "common.cuh"

namespace bff {

// adjacency of the merge graph: same label AND iou > thr, IoU in float32 as calculate_iou computes it
// (projection_2d_to_3d.py:120-122,158-165). NaN (0/0) compares false.
// One BIT per pair: row r is `stride_words` 32-bit words (an even number, so the host walks 64-bit words), bit c of row r
// = edge (r, c); a warp's ballot is one word.
__global__ void merge_adjacency_kernel(const int32_t *__restrict__ inter, const int32_t *__restrict__ pop,
                                       const int32_t *__restrict__ label, int n, int stride_words, float thr, uint32_t *__restrict__ adj) {
    const int r = blockIdx.y;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    bool e = false;
    if (c < n) {
        const float fi = __int2float_rn(inter[(long long)r * n + c]);
        const float un = __fsub_rn(__fadd_rn(__int2float_rn(pop[r]), __int2float_rn(pop[c])), fi);
        e = (label == nullptr || label[r] == label[c]) && (__fdiv_rn(fi, un) > thr);
    }
    const unsigned w = __ballot_sync(0xffffffffu, e);
    if ((threadIdx.x & 31) == 0 && (c >> 5) < stride_words) adj[(long long)r * stride_words + (c >> 5)] = w;
}

__global__ void rows_gather_kernel(const uint32_t *__restrict__ rows, long long words, const int32_t *__restrict__ idx,
                                   uint32_t *__restrict__ out) {
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= words) return;
    out[(long long)blockIdx.y * words + w] = __ldg(rows + (long long)idx[blockIdx.y] * words + w);
}

__global__ void set_i32_kernel(int32_t *dst, int32_t value) { *dst = value; }

int launch_merge_adjacency(const int32_t *inter, const int32_t *pop, const int32_t *label, int n, int stride_words, float thr,
                           uint32_t *adj, cudaStream_t s) {
    dim3 grid((unsigned)(stride_words * 32 / 256 + 1), (unsigned)n);
    merge_adjacency_kernel<<<grid, 256, 0, s>>>(inter, pop, label, n, stride_words, thr, adj);
    BFF_LAUNCH_CHECK("merge_adjacency_kernel");
    return BFF_OK;
}

int launch_set_i32(int32_t *dst, int32_t value, cudaStream_t s) {
    set_i32_kernel<<<1, 1, 0, s>>>(dst, value);
    BFF_LAUNCH_CHECK("set_i32_kernel");
    return BFF_OK;
}

}  // namespace bff

using namespace bff;

// BFF_PIPE_TRACE=1: wall-clock marks inside the fused pipelines on stderr (dev aid; one getenv per call otherwise)
struct PipeTrace {
    bool on;
    std::chrono::steady_clock::time_point t0;
    explicit PipeTrace(const char *what) : on(getenv("BFF_PIPE_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {
        if (on) fprintf(stderr, "[bff trace] %s:", what);
    }
    void mark(const char *name) {
        if (!on) return;
        const auto t = std::chrono::steady_clock::now();
        fprintf(stderr, " %s %.0fus", name, std::chrono::duration<double, std::micro>(t - t0).count());
        t0 = t;
    }
    ~PipeTrace() { if (on) fprintf(stderr, "\n"); }
};

#define PIPE_TAKE(var, arena, type, count)                         \
    type *var = (arena).get<type>(count);                          \
    if (!var) return BFF_ENOMEM

static int upload_i32(bff_ctx *c, const std::vector<int32_t> &v, int32_t **dev_out, cudaStream_t s) {
    const size_t n = v.empty() ? 1 : v.size();
    int32_t *h = c->pin.get<int32_t>(n);
    int32_t *d = c->dev.get<int32_t>(n);
    if (!h || !d) return BFF_ENOMEM;
    if (!v.empty()) {
        // staged through the pinned arena and pulled over by the SM copy kernel, NOT a copy-engine transfer: while the
        // next scene's depth stack occupies the H2D engine (3.4 ms per 184 MB) a cudaMemcpyAsync here would wait for it
        memcpy(h, v.data(), v.size() * 4);
        const int rc = bff_upload(d, h, (int64_t)v.size() * 4, s);
        if (rc != BFF_OK) return rc;
    }
    *dev_out = d;
    return BFF_OK;
}

extern "C" int bff_rows_gather(bff_ctx *c, const uint32_t *rows, int64_t words, const int32_t *idx_host, int32_t n_idx,
                               uint32_t *out, void *stream_) {
    cudaStream_t s = (cudaStream_t)stream_;
    BFF_REQUIRE(c && n_idx >= 0 && words >= 0, "bff_rows_gather: bad arguments");
    if (n_idx == 0 || words == 0) return BFF_OK;
    BFF_REQUIRE(rows && idx_host && out, "bff_rows_gather: null pointer");
    c->dev.reset();
    c->pin.reset();
    int32_t *d_idx = nullptr;
    int rc = upload_i32(c, std::vector<int32_t>(idx_host, idx_host + n_idx), &d_idx, s);
    if (rc != BFF_OK) return rc;
    for (int32_t r0 = 0; r0 < n_idx; r0 += kMaxGridY) {
        dim3 grid((unsigned)ceil_div64(words, 256), (unsigned)std::min(n_idx - r0, kMaxGridY));
        rows_gather_kernel<<<grid, 256, 0, s>>>(rows, words, d_idx + r0, out + (size_t)r0 * words);
        BFF_LAUNCH_CHECK("rows_gather_kernel");
    }
    // the pinned staging buffer is reused by the next call: make sure the copy has been consumed
    BFF_CUDA(cudaStreamSynchronize(s));
    return BFF_OK;
}

// Side streams for the batched pipelines. With several jobs, job q is enqueued on lane q % n between fork() (the lanes
// wait for what is already on the caller's stream) and join() (the caller's stream waits for the lanes); one job keeps
// its lane through all phases. A single job, or BFF_BATCH_LANES=0, runs on the caller's stream.
struct Lanes {
    bff_ctx *c;
    cudaStream_t main;
    int n = 0;
    int init(bff_ctx *ctx, cudaStream_t s, int n_jobs) {
        c = ctx;
        main = s;
        static const int want = [] { const char *e = getenv("BFF_BATCH_LANES"); const int v = e ? atoi(e) : 4; return v < 0 ? 0 : (v > bff_ctx::kMaxLanes ? bff_ctx::kMaxLanes : v); }();
        if (n_jobs < 2 || want < 2) return BFF_OK;
        if (c->n_lanes == 0) {
            BFF_CUDA(cudaEventCreateWithFlags(&c->fork_ev, cudaEventDisableTiming));
            for (int i = 0; i < bff_ctx::kMaxLanes; ++i) {
                BFF_CUDA(cudaStreamCreateWithFlags(&c->lane[i], cudaStreamNonBlocking));
                BFF_CUDA(cudaEventCreateWithFlags(&c->lane_done[i], cudaEventDisableTiming));
                c->n_lanes = i + 1;
            }
        }
        n = want < n_jobs ? want : n_jobs;
        return BFF_OK;
    }
    cudaStream_t of(int q) const { return n ? c->lane[q % n] : main; }
    int fork() {
        if (!n) return BFF_OK;
        BFF_CUDA(cudaEventRecord(c->fork_ev, main));
        for (int i = 0; i < n; ++i) BFF_CUDA(cudaStreamWaitEvent(c->lane[i], c->fork_ev, 0));
        return BFF_OK;
    }
    int join() {
        if (!n) return BFF_OK;
        for (int i = 0; i < n; ++i) {
            BFF_CUDA(cudaEventRecord(c->lane_done[i], c->lane[i]));
            BFF_CUDA(cudaStreamWaitEvent(main, c->lane_done[i], 0));
        }
        return BFF_OK;
    }
    void drain() {                               // error path: nothing may still be writing into the pinned arena
        for (int i = 0; i < n; ++i) cudaStreamSynchronize(c->lane[i]);
        cudaStreamSynchronize(main);
    }
};

// ------------------------------------------------------------------------------------------------
// bff_stage2_aggregate / bff_stage2_aggregate_batch
//
// One query's aggregation is four phases with a stream synchronisation between them (each hands small device results
// to a host decision). The batch entry runs every phase for ALL jobs before it synchronises, so Q concurrent queries
// pay three synchronisations together instead of three each, and their small kernels queue back to back.
namespace {
struct AggJob {
    const bff_stage2_desc *d = nullptr;
    int n = 0, g = 0, stride_words = 0;
    int64_t words = 0;
    const int32_t *vc = nullptr;
    uint32_t *h_adj = nullptr;
    int32_t *h_max = nullptr;
    float *h_vals = nullptr, *d_vals_keep = nullptr;
    size_t first_vals = 0;
    int mb = 0, vb = 0;
    int32_t *h_ginter = nullptr, *h_before = nullptr, *h_after = nullptr;
    uint32_t *d_keep = nullptr;
};

int agg_check(const bff_stage2_desc *d) {
    BFF_REQUIRE(d, "bff_stage2_aggregate: null argument");
    BFF_REQUIRE(d->n_rows > 0 && d->n_points > 0, "bff_stage2_aggregate: empty input (the caller handles 'nothing lifted')");
    BFF_REQUIRE(d->rows && d->conf && d->label_id && d->mc && d->agg_rows && d->n_groups && d->group_first && d->group_size &&
                    d->group_conf && d->before && d->after,
                "bff_stage2_aggregate: null pointer");
    BFF_REQUIRE(d->min_aggregated >= 1, "bff_stage2_aggregate: min_aggregated must be >= 1");
    BFF_REQUIRE(d->filter_mode >= 0 && d->filter_mode <= 2 && (d->filter_mode != 0 || d->vc), "bff_stage2_aggregate: bad filter mode");
    BFF_REQUIRE(d->n_rows <= 46340, "bff_stage2_aggregate: too many rows");
    BFF_REQUIRE((d->inter_pre == nullptr) == (d->pop_pre == nullptr), "bff_stage2_aggregate: inter_pre and pop_pre come together");
    return BFF_OK;
}

// phase 1 (enqueue): all-pairs intersections -> thresholded merge graph (calculate_iou :149-166, aggregate :113-122),
// and the maxima that size the vote histogram
int agg_phase1(bff_ctx *c, AggJob &j, cudaStream_t s) {
    const bff_stage2_desc *d = j.d;
    const int n = j.n = d->n_rows;
    j.words = ceil_div64(d->n_points, 32);
    const size_t nn = (size_t)n * n;
    const bool have_pairs = d->inter_pre != nullptr;
    int32_t *d_inter = const_cast<int32_t *>(d->inter_pre), *d_pop = const_cast<int32_t *>(d->pop_pre);
    if (!have_pairs) {
        d_inter = c->dev.get<int32_t>(nn);
        d_pop = c->dev.get<int32_t>(n);
        if (!d_inter || !d_pop) return BFF_ENOMEM;
    }
    j.stride_words = ((n + 63) / 64) * 2;                 // 32-bit words per adjacency row, whole 64-bit words
    const size_t adj_bytes = (size_t)n * j.stride_words * 4;
    PIPE_TAKE(d_adj, c->dev, uint32_t, (size_t)n * j.stride_words);
    j.h_adj = c->pin.get<uint32_t>((size_t)n * j.stride_words);
    if (!j.h_adj) return BFF_ENOMEM;
    int32_t *d_label = nullptr;
    int rc = upload_i32(c, std::vector<int32_t>(d->label_id, d->label_id + n), &d_label, s);
    if (rc != BFF_OK) return rc;
    if (!have_pairs) {
        rc = bff_mask_iou(d->rows, n, d->rows, n, j.words, d_inter, d_pop, d_pop, nullptr, s);
        if (rc != BFF_OK) return rc;
    }
    if ((rc = launch_merge_adjacency(d_inter, d_pop, d_label, n, j.stride_words, d->iou_thres, d_adj, s)) != BFF_OK) return rc;
    BFF_CUDA(cudaMemcpyAsync(j.h_adj, d_adj, adj_bytes, cudaMemcpyDeviceToHost, s));
    // the point filter (:512-583) does not depend on the grouping: its first step (the maxima that size the vote
    // histogram) rides on this synchronisation instead of paying for one of its own
    j.vc = d->filter_mode == 0 ? d->vc : nullptr;
    if (d->filter_mode != 2) {
        PIPE_TAKE(d_max, c->dev, int32_t, 2);
        j.h_max = c->pin.get<int32_t>(2);
        if (!j.h_max) return BFF_ENOMEM;
        if ((rc = bff_counts_max(d->mc, j.vc, d->n_points, d_max, s)) != BFF_OK) return rc;
        BFF_CUDA(cudaMemcpyAsync(j.h_max, d_max, 8, cudaMemcpyDeviceToHost, s));
    }
    return BFF_OK;
}

// phase 2a (enqueue, needs the maxima): point filter (:512-583), first half -- the unique() over N values comes from the
// (masked, viewed) histogram. Enqueued BEFORE the host searches the components: the device builds the histogram and
// its distinct ratios while the host walks the graph, and the next synchronisation brings them along.
int agg_phase2_votes(bff_ctx *c, AggJob &j, cudaStream_t s) {
    const bff_stage2_desc *d = j.d;
    if (d->filter_mode == 2) return BFF_OK;
    int rc;
    j.mb = j.h_max[0] + 1;
    j.vb = j.vc ? j.h_max[1] + 1 : 1;
    const size_t bins = (size_t)j.mb * j.vb;
    PIPE_TAKE(d_hist, c->dev, uint32_t, bins);
    if ((rc = bff_vote_hist(d->mc, j.vc, d->n_points, j.mb, j.vb, d_hist, s)) != BFF_OK) return rc;
    // distinct ratios straight from the non-empty bins, on the device: only they travel to the host
    const int64_t tsize = bff_vote_unique_table_size(j.mb, j.vb);
    PIPE_TAKE(d_table, c->dev, uint32_t, (size_t)tsize);
    PIPE_TAKE(d_vals, c->dev, float, bins + 1);
    j.h_vals = c->pin.get<float>(bins + 1);
    if (!j.h_vals) return BFF_ENOMEM;
    int32_t *d_cnt = reinterpret_cast<int32_t *>(d_vals + bins);       // the count rides behind the values
    if ((rc = bff_vote_unique(d_hist, j.mb, j.vb, d->filter_mode, d_table, tsize, d_vals, d_cnt, s)) != BFF_OK) return rc;
    // the number of distinct values is not known yet: a first slice that nearly always suffices, the rest on demand
    j.first_vals = std::min<size_t>(bins, 16384);
    BFF_CUDA(cudaMemcpyAsync(j.h_vals, d_vals, j.first_vals * 4, cudaMemcpyDeviceToHost, s));
    BFF_CUDA(cudaMemcpyAsync(j.h_vals + bins, d_cnt, 4, cudaMemcpyDeviceToHost, s));
    j.d_vals_keep = d_vals;
    return BFF_OK;
}

// phase 2b (host + enqueue): connected components (find_unconnected_subgraphs_tensor :250-274; the graph is symmetric,
// so the reference's reachability lists are the components, emitted by smallest member, members ascending). A node
// without any edge (an empty row: IoU diagonal NaN) forms no group. Then the OR-merge and the pair overlaps.
int agg_phase2_groups(bff_ctx *c, AggJob &j, cudaStream_t s) {
    const bff_stage2_desc *d = j.d;
    const int n = j.n;
    // flood fill over the bit rows, 64 neighbours per step: root[i] = smallest member of i's component
    const int w64 = j.stride_words / 2;
    const uint64_t *adj64 = reinterpret_cast<const uint64_t *>(j.h_adj);
    std::vector<uint64_t> visited(w64, 0), frontier(w64), next(w64);
    std::vector<int> root(n, -1), count(n, 0);
    std::vector<char> has_edge(n, 0);
    for (int i = 0; i < n; ++i) {
        const uint64_t *row = adj64 + (size_t)i * w64;
        uint64_t any = 0;
        for (int w = 0; w < w64; ++w) any |= row[w];
        has_edge[i] = any != 0;
    }
    for (int i = 0; i < n; ++i) {
        if ((visited[i >> 6] >> (i & 63)) & 1) continue;
        std::fill(frontier.begin(), frontier.end(), 0);
        frontier[i >> 6] = 1ull << (i & 63);
        visited[i >> 6] |= 1ull << (i & 63);
        bool grew = true;
        while (grew) {
            grew = false;
            std::fill(next.begin(), next.end(), 0);
            for (int w = 0; w < w64; ++w) {
                for (uint64_t m = frontier[w]; m; m &= m - 1) {
                    const int u = w * 64 + __builtin_ctzll(m);
                    root[u] = i;
                    const uint64_t *row = adj64 + (size_t)u * w64;
                    for (int k = 0; k < w64; ++k) next[k] |= row[k];
                }
            }
            for (int k = 0; k < w64; ++k) {
                next[k] &= ~visited[k];
                visited[k] |= next[k];
                grew |= next[k] != 0;
            }
            frontier.swap(next);
        }
    }
    for (int i = 0; i < n; ++i) count[root[i]]++;
    // kept groups (:203): len >= min_aggregated; roots ascending == order of smallest member
    std::vector<int> group_of_root(n, -1);
    int g = 0;
    for (int r = 0; r < n; ++r)
        if (root[r] == r && has_edge[r] && count[r] >= d->min_aggregated) group_of_root[r] = g++;
    j.g = g;
    *d->n_groups = g;
    if (d->filter_value) *d->filter_value = 0.0f;
    if (g == 0) return BFF_OK;
    BFF_REQUIRE(g <= d->cap_groups, "bff_stage2_aggregate: %d groups exceed cap_groups %d", g, d->cap_groups);
    std::vector<int32_t> seg(g + 1, 0), idx;
    idx.reserve(n);
    {
        std::vector<std::vector<int32_t>> mem(g);
        for (int i = 0; i < n; ++i) { const int gi = group_of_root[root[i]]; if (gi >= 0) mem[gi].push_back(i); }
        for (int k = 0; k < g; ++k) {
            // conf = sum(conf_k) / len, summed sequentially in float32 in index order (:220-224)
            float acc = 0.0f;
            for (int32_t m : mem[k]) { acc = acc + d->conf[m]; idx.push_back(m); }
            d->group_conf[k] = acc / (float)mem[k].size();
            d->group_first[k] = mem[k][0];
            d->group_size[k] = (int32_t)mem[k].size();
            seg[k + 1] = (int32_t)idx.size();
        }
    }
    if (d->members) memcpy(d->members, idx.data(), idx.size() * 4);

    // ---- OR-merge the members (:214-223)
    int rc;
    int32_t *d_idx = nullptr, *d_seg = nullptr;
    {   // members and segment bounds travel together: one staging copy instead of two
        const size_t n_idx = idx.size();
        idx.insert(idx.end(), seg.begin(), seg.end());
        if ((rc = upload_i32(c, idx, &d_idx, s)) != BFF_OK) return rc;
        d_seg = d_idx + n_idx;
    }
    if ((rc = bff_rows_or_reduce(d->rows, j.words, d_idx, d_seg, g, d->agg_rows, s)) != BFF_OK) return rc;

    // ---- pair overlaps of the merged rows + row sums 'before' (:285-291, :592), issued now, read later
    PIPE_TAKE(d_ginter, c->dev, int32_t, (size_t)g * g);
    PIPE_TAKE(d_before, c->dev, int32_t, g);
    j.h_ginter = c->pin.get<int32_t>((size_t)g * g);
    j.h_before = c->pin.get<int32_t>(g);
    j.d_keep = c->dev.get<uint32_t>(j.words);
    if (!j.h_ginter || !j.h_before || !j.d_keep) return BFF_ENOMEM;
    if ((rc = bff_mask_iou(d->agg_rows, g, d->agg_rows, g, j.words, d_ginter, d_before, d_before, nullptr, s)) != BFF_OK) return rc;
    BFF_CUDA(cudaMemcpyAsync(j.h_ginter, d_ginter, (size_t)g * g * 4, cudaMemcpyDeviceToHost, s));
    BFF_CUDA(cudaMemcpyAsync(j.h_before, d_before, (size_t)g * 4, cudaMemcpyDeviceToHost, s));
    return BFF_OK;
}

// phase 3 (host + enqueue; pair overlaps, 'before' and the distinct vote ratios are on the host now)
int agg_phase3(bff_ctx *c, AggJob &j, cudaStream_t s, PipeTrace &tr) {
    const bff_stage2_desc *d = j.d;
    const int g = j.g;
    if (g == 0) return BFF_OK;
    int rc;
    // ---- solve_overlapping's sequential schedule (:285-299): pairs i<j that overlap in the incoming rows; the mask
    // merged from MORE members keeps the shared points, ties: i loses. Enqueued before the host picks the vote threshold.
    std::vector<int32_t> prog;
    for (int i = 0; i < g; ++i)
        for (int k = i + 1; k < g; ++k)
            if (j.h_ginter[(size_t)i * g + k] > 0) {
                if (d->group_size[i] > d->group_size[k]) { prog.push_back(k); prog.push_back(i); }
                else { prog.push_back(i); prog.push_back(k); }
            }
    if (!prog.empty()) {
        int32_t *d_prog = nullptr;
        if ((rc = upload_i32(c, prog, &d_prog, s)) != BFF_OK) return rc;
        if ((rc = bff_overlap_resolve(d->agg_rows, g, j.words, d_prog, (int32_t)(prog.size() / 2), s)) != BFF_OK) return rc;
    }
    float thr = 0.0f;
    if (d->filter_mode == 2) {
        if ((rc = bff_point_keep(d->mc, nullptr, d->n_points, 2, 0.0f, j.d_keep, s)) != BFF_OK) return rc;
    } else {
        const size_t bins = (size_t)j.mb * j.vb;
        int32_t n_unique;
        memcpy(&n_unique, j.h_vals + bins, 4);
        if ((size_t)n_unique > j.first_vals) {
            BFF_CUDA(cudaMemcpyAsync(j.h_vals + j.first_vals, j.d_vals_keep + j.first_vals, ((size_t)n_unique - j.first_vals) * 4,
                                     cudaMemcpyDeviceToHost, s));
            BFF_CUDA(cudaStreamSynchronize(s));
        }
        std::vector<float> u(j.h_vals, j.h_vals + n_unique);
        // u[floor(p * len(u))] of the SORTED unique values with p a Python float (:516-517, :575-576); out of range raises in the reference
        const double pos = floor(d->filter_fraction * (double)u.size());
        BFF_REQUIRE(pos >= 0.0 && pos < (double)u.size(), "bff_stage2_aggregate: filter fraction %g selects index %g of %zu unique values",
                    d->filter_fraction, pos, u.size());
        std::nth_element(u.begin(), u.begin() + (size_t)pos, u.end());
        thr = u[(size_t)pos];
        if (tr.on) fprintf(stderr, " [bins %dx%d unique %zu]", j.mb, j.vb, u.size());
        if ((rc = bff_point_keep(d->mc, j.vc, d->n_points, d->filter_mode, thr, j.d_keep, s)) != BFF_OK) return rc;
    }
    if (d->filter_value) *d->filter_value = thr;

    // ---- ins &= masked_points, row sums 'after' (:594-596)
    PIPE_TAKE(d_after, c->dev, int32_t, g);
    if ((rc = bff_rows_and_popcount(d->agg_rows, g, j.words, j.d_keep, d_after, s)) != BFF_OK) return rc;
    j.h_after = c->pin.get<int32_t>(g);
    if (!j.h_after) return BFF_ENOMEM;
    BFF_CUDA(cudaMemcpyAsync(j.h_after, d_after, (size_t)g * 4, cudaMemcpyDeviceToHost, s));
    return BFF_OK;
}

// All phases for `n_jobs` queries. On an error the stream is drained before returning: copies into the pinned arena may
// still be in flight and the next call reuses that arena.
int agg_run(bff_ctx *c, const bff_stage2_desc *descs, int n_jobs, cudaStream_t s) {
    int rc = BFF_OK;
    for (int q = 0; q < n_jobs; ++q)
        if ((rc = agg_check(descs + q)) != BFF_OK) return rc;
    BFF_CUDA(cudaSetDevice(c->device));
    c->dev.reset();
    c->pin.reset();
    PipeTrace tr(n_jobs == 1 ? "stage2_aggregate" : "stage2_aggregate_batch");
    std::vector<AggJob> jobs(n_jobs);
    for (int q = 0; q < n_jobs; ++q) jobs[q].d = descs + q;
    Lanes lanes;
    if ((rc = lanes.init(c, s, n_jobs)) != BFF_OK) return rc;
#define AGG_ALL(call)                                                      \
    for (int q = 0; q < n_jobs && rc == BFF_OK; ++q) rc = (call);          \
    if (rc != BFF_OK) { lanes.drain(); return rc; }
#define AGG_STEP(call)                                                     \
    if ((rc = (call)) != BFF_OK) { lanes.drain(); return rc; }
    AGG_STEP(lanes.fork());
    AGG_ALL(agg_phase1(c, jobs[q], lanes.of(q)));
    AGG_STEP(lanes.join());
    tr.mark("enqueue1");
    BFF_CUDA(cudaStreamSynchronize(s));
    tr.mark("sync1");
    AGG_ALL(agg_phase2_votes(c, jobs[q], lanes.of(q)));
    tr.mark("enqueue_votes");
    AGG_ALL(agg_phase2_groups(c, jobs[q], lanes.of(q)));
    AGG_STEP(lanes.join());
    tr.mark("components+enqueue2");
    BFF_CUDA(cudaStreamSynchronize(s));      // pair overlaps, 'before' and the histograms are on the host now
    tr.mark("sync2");
    AGG_ALL(agg_phase3(c, jobs[q], lanes.of(q), tr));
    AGG_STEP(lanes.join());
    tr.mark("threshold+enqueue3");
    BFF_CUDA(cudaStreamSynchronize(s));
    tr.mark("sync3");
#undef AGG_ALL
#undef AGG_STEP
    for (const AggJob &j : jobs)
        if (j.g > 0) {
            memcpy(j.d->before, j.h_before, (size_t)j.g * 4);
            memcpy(j.d->after, j.h_after, (size_t)j.g * 4);
        }
    return BFF_OK;
}
}  // namespace

extern "C" int bff_stage2_aggregate(bff_ctx *c, const bff_stage2_desc *d, void *stream_) {
    BFF_REQUIRE(c && d, "bff_stage2_aggregate: null argument");
    return agg_run(c, d, 1, (cudaStream_t)stream_);
}

extern "C" int bff_stage2_aggregate_batch(bff_ctx *c, const bff_stage2_desc *descs, int32_t n_jobs, void *stream_) {
    BFF_REQUIRE(c && n_jobs >= 0, "bff_stage2_aggregate_batch: bad arguments");
    if (n_jobs == 0) return BFF_OK;
    BFF_REQUIRE(descs, "bff_stage2_aggregate_batch: null descriptors");
    return agg_run(c, descs, n_jobs, (cudaStream_t)stream_);
}

// ------------------------------------------------------------------------------------------------
// bff_refine_match
// torch.argmax semantics (refinement.py:211,288): first maximum; NaN counts as the maximum (first NaN wins).
static int argmax_torch(const float *v, int n, int stride) {
    int best = 0;
    float cur = v[0];
    for (int i = 1; i < n; ++i) {
        const float x = v[(size_t)i * stride];
        if ((x > cur) || (isnan(x) && !isnan(cur))) { best = i; cur = x; }
    }
    return best;
}

// Like the aggregation, one job is phases separated by stream synchronisations; the batch entry runs each phase for all
// jobs (queries of one scene, each with its OWN copy of the stage-1 rows -- pass 1 modifies them in place) before it waits.
namespace {
struct RefineJob {
    const bff_refine_desc *d = nullptr;
    int n1 = 0, n2 = 0, m = 0;
    int64_t words = 0;
    int32_t *d_inter = nullptr, *d_p1 = nullptr, *d_p2 = nullptr;
    float *d_iou = nullptr, *h_iou = nullptr, *h_giou = nullptr;
    std::vector<int32_t> mm;
};

int refine_check(const bff_refine_desc *d) {
    BFF_REQUIRE(d, "bff_refine_match: null argument");
    BFF_REQUIRE(d->n1 > 0 && d->n2 > 0 && d->n_points > 0, "bff_refine_match: empty input (the caller handles empty stages)");
    BFF_REQUIRE(d->s1_rows && d->s2_rows && d->s2_conf && d->s2_out && d->n2_out && d->s2_conf_out && d->match && d->match_iou,
                "bff_refine_match: null pointer");
    BFF_REQUIRE(d->n2 <= 65535, "bff_refine_match: too many stage-2 masks");
    return BFF_OK;
}

// phase 1 (enqueue): iou(stage1, stage2) (:208-213)
int refine_phase1(bff_ctx *c, RefineJob &j, cudaStream_t s) {
    const bff_refine_desc *d = j.d;
    const int n1 = j.n1 = d->n1, n2 = j.n2 = d->n2;
    j.words = ceil_div64(d->n_points, 32);
    // scratch for an (n1 x k) IoU, k <= n2
    j.d_inter = c->dev.get<int32_t>((size_t)n1 * n2);
    j.d_p1 = c->dev.get<int32_t>(n1);
    j.d_p2 = c->dev.get<int32_t>(n2);
    j.d_iou = c->dev.get<float>((size_t)n1 * n2);
    j.h_iou = c->pin.get<float>((size_t)n1 * n2);
    if (!j.d_inter || !j.d_p1 || !j.d_p2 || !j.d_iou || !j.h_iou) return BFF_ENOMEM;
    int rc;
    if ((rc = bff_mask_iou(d->s1_rows, n1, d->s2_rows, n2, j.words, j.d_inter, j.d_p1, j.d_p2, j.d_iou, s)) != BFF_OK) return rc;
    BFF_CUDA(cudaMemcpyAsync(j.h_iou, j.d_iou, (size_t)n1 * n2 * 4, cudaMemcpyDeviceToHost, s));
    return BFF_OK;
}

// phase 2 (host + enqueue): max_match = argmax over stage-1; IoU among the matched stage-1 masks (:217-224)
int refine_phase2(bff_ctx *c, RefineJob &j, cudaStream_t s) {
    const bff_refine_desc *d = j.d;
    const int n1 = j.n1, n2 = j.n2;
    const int64_t words = j.words;
    int rc;
    j.mm.resize(n2);
    for (int k = 0; k < n2; ++k) j.mm[k] = argmax_torch(j.h_iou + k, n1, n2);
    PIPE_TAKE(d_sel, c->dev, uint32_t, (size_t)n2 * words);
    PIPE_TAKE(d_ginter, c->dev, int32_t, (size_t)n2 * n2);
    PIPE_TAKE(d_gpop, c->dev, int32_t, n2);
    PIPE_TAKE(d_giou, c->dev, float, (size_t)n2 * n2);
    j.h_giou = c->pin.get<float>((size_t)n2 * n2);
    if (!j.h_giou) return BFF_ENOMEM;
    int32_t *d_mm = nullptr;
    if ((rc = upload_i32(c, j.mm, &d_mm, s)) != BFF_OK) return rc;
    {
        dim3 grid((unsigned)ceil_div64(words, 256), (unsigned)n2);
        rows_gather_kernel<<<grid, 256, 0, s>>>(d->s1_rows, words, d_mm, d_sel);
        BFF_LAUNCH_CHECK("rows_gather_kernel");
    }
    if ((rc = bff_mask_iou(d_sel, n2, d_sel, n2, words, d_ginter, d_gpop, d_gpop, d_giou, s)) != BFF_OK) return rc;
    BFF_CUDA(cudaMemcpyAsync(j.h_giou, d_giou, (size_t)n2 * n2 * 4, cudaMemcpyDeviceToHost, s));
    return BFF_OK;
}

// phase 3 (host + enqueue): the sequential merge loop, the stage-2 merges and the final IoU
int refine_phase3(bff_ctx *c, RefineJob &j, cudaStream_t s) {
    const bff_refine_desc *d = j.d;
    const int n1 = j.n1, n2 = j.n2;
    const int64_t words = j.words;
    const std::vector<int32_t> &mm = j.mm;
    const float *h_giou = j.h_giou;
    int rc;
    // ---- sequential merge loop (:230-249): redirects + in-place ORs of stage-1 rows, in list order
    std::vector<int32_t> best(n2), redirect(n2, -1), or_prog;
    for (int i = 0; i < n2; ++i) {
        if (redirect[i] != -1) { best[i] = mm[redirect[i]]; continue; }
        best[i] = mm[i];
        for (int k = 0; k < n2; ++k) {
            if (k == i) continue;                                            // diagonal set to 0 (:219)
            if (h_giou[(size_t)i * n2 + k] > d->stage1_iou_thres) {
                redirect[k] = i;
                or_prog.push_back(mm[i]);
                or_prog.push_back(mm[k]);
            }
        }
    }
    if (d->best_match_log) memcpy(d->best_match_log, best.data(), (size_t)n2 * 4);
    if (!or_prog.empty()) {
        int32_t *d_prog = nullptr;
        if ((rc = upload_i32(c, or_prog, &d_prog, s)) != BFF_OK) return rc;
        if ((rc = bff_rows_or_into(d->s1_rows, n1, words, d_prog, (int32_t)(or_prog.size() / 2), s)) != BFF_OK) return rc;
    }

    // ---- stage-2 masks that share a stage-1 match are merged (:258-281): any() of the rows, mean of the
    // confidences, the merged mask moves to the END; matches are visited in ascending order (torch.unique)
    std::vector<std::vector<int32_t>> cur(n2);
    std::vector<float> cur_conf(d->s2_conf, d->s2_conf + n2);
    std::vector<int32_t> cur_best(best);
    for (int k = 0; k < n2; ++k) cur[k] = {k};
    std::vector<int32_t> uniq(best);
    std::sort(uniq.begin(), uniq.end());
    uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
    for (int32_t u : uniq) {
        if (std::count(best.begin(), best.end(), u) <= 1) continue;
        std::vector<std::vector<int32_t>> keep;
        std::vector<float> keep_conf;
        std::vector<int32_t> keep_best, merged;
        float sum = 0.0f;
        int cnt = 0;
        for (size_t k = 0; k < cur.size(); ++k) {
            if (cur_best[k] == u) {
                merged.insert(merged.end(), cur[k].begin(), cur[k].end());
                sum += cur_conf[k];
                ++cnt;
            } else {
                keep.push_back(cur[k]); keep_conf.push_back(cur_conf[k]); keep_best.push_back(cur_best[k]);
            }
        }
        keep.push_back(merged); keep_conf.push_back(sum / (float)cnt); keep_best.push_back(u);
        cur.swap(keep); cur_conf.swap(keep_conf); cur_best.swap(keep_best);
    }
    const int m = j.m = (int)cur.size();
    *d->n2_out = m;
    std::vector<int32_t> idx, seg(m + 1, 0);
    for (int k = 0; k < m; ++k) {
        idx.insert(idx.end(), cur[k].begin(), cur[k].end());
        seg[k + 1] = (int32_t)idx.size();
        d->s2_conf_out[k] = cur_conf[k];
    }
    int32_t *d_idx = nullptr, *d_seg = nullptr;
    {   // members and segment bounds travel together: one staging copy instead of two
        const size_t n_idx = idx.size();
        idx.insert(idx.end(), seg.begin(), seg.end());
        if ((rc = upload_i32(c, idx, &d_idx, s)) != BFF_OK) return rc;
        d_seg = d_idx + n_idx;
    }
    if ((rc = bff_rows_or_reduce(d->s2_rows, words, d_idx, d_seg, m, d->s2_out, s)) != BFF_OK) return rc;

    // ---- IoU of the merged stage-1 masks with the merged stage-2 masks (:285-290)
    if ((rc = bff_mask_iou(d->s1_rows, n1, d->s2_out, m, words, j.d_inter, j.d_p1, j.d_p2, j.d_iou, s)) != BFF_OK) return rc;
    BFF_CUDA(cudaMemcpyAsync(j.h_iou, j.d_iou, (size_t)n1 * m * 4, cudaMemcpyDeviceToHost, s));
    return BFF_OK;
}

int refine_run(bff_ctx *c, const bff_refine_desc *descs, int n_jobs, cudaStream_t s) {
    int rc = BFF_OK;
    for (int q = 0; q < n_jobs; ++q)
        if ((rc = refine_check(descs + q)) != BFF_OK) return rc;
    BFF_CUDA(cudaSetDevice(c->device));
    c->dev.reset();
    c->pin.reset();
    PipeTrace tr(n_jobs == 1 ? "refine_match" : "refine_match_batch");
    std::vector<RefineJob> jobs(n_jobs);
    for (int q = 0; q < n_jobs; ++q) jobs[q].d = descs + q;
    Lanes lanes;
    if ((rc = lanes.init(c, s, n_jobs)) != BFF_OK) return rc;
#define REFINE_ALL(call)                                                   \
    for (int q = 0; q < n_jobs && rc == BFF_OK; ++q) rc = (call);          \
    if (rc != BFF_OK) { lanes.drain(); return rc; }
#define REFINE_STEP(call)                                                  \
    if ((rc = (call)) != BFF_OK) { lanes.drain(); return rc; }
    REFINE_STEP(lanes.fork());
    REFINE_ALL(refine_phase1(c, jobs[q], lanes.of(q)));
    REFINE_STEP(lanes.join());
    tr.mark("enqueue1");
    BFF_CUDA(cudaStreamSynchronize(s));
    tr.mark("sync1");
    REFINE_ALL(refine_phase2(c, jobs[q], lanes.of(q)));
    REFINE_STEP(lanes.join());
    tr.mark("enqueue2");
    BFF_CUDA(cudaStreamSynchronize(s));
    tr.mark("sync2");
    REFINE_ALL(refine_phase3(c, jobs[q], lanes.of(q)));
    REFINE_STEP(lanes.join());
    tr.mark("enqueue3");
    BFF_CUDA(cudaStreamSynchronize(s));
    tr.mark("sync3");
#undef REFINE_ALL
#undef REFINE_STEP
    // final argmax (:288) and the matched IoU (:308)
    for (const RefineJob &j : jobs)
        for (int k = 0; k < j.m; ++k) {
            j.d->match[k] = argmax_torch(j.h_iou + k, j.n1, j.m);
            j.d->match_iou[k] = j.h_iou[(size_t)j.d->match[k] * j.m + k];
        }
    return BFF_OK;
}
}  // namespace

extern "C" int bff_refine_match(bff_ctx *c, const bff_refine_desc *d, void *stream_) {
    BFF_REQUIRE(c && d, "bff_refine_match: null argument");
    return refine_run(c, d, 1, (cudaStream_t)stream_);
}

extern "C" int bff_refine_match_batch(bff_ctx *c, const bff_refine_desc *descs, int32_t n_jobs, void *stream_) {
    BFF_REQUIRE(c && n_jobs >= 0, "bff_refine_match_batch: bad arguments");
    if (n_jobs == 0) return BFF_OK;
    BFF_REQUIRE(descs, "bff_refine_match_batch: null descriptors");
    return refine_run(c, descs, n_jobs, (cudaStream_t)stream_);
}
