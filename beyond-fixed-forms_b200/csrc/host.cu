// Error reporting, context and the host-buffer entry points of libbff.so.
#include <stdarg.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace bff {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

}  // namespace bff

using namespace bff;

extern "C" int bff_version(void) { return BFF_VERSION; }
extern "C" const char *bff_last_error(void) { return g_err; }

namespace bff {
void *Arena::take(size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    if (bytes == 0) bytes = 256;
    while (cur < n_blocks) {
        if (off + bytes <= blocks[cur].cap) {
            void *r = blocks[cur].p + off;
            off += bytes;
            return r;
        }
        ++cur;
        off = 0;
    }
    if (n_blocks == 32) {
        set_error("bff arena: too many blocks");
        return nullptr;
    }
    size_t cap = bytes > ((size_t)32 << 20) ? bytes : ((size_t)32 << 20);
    unsigned char *p = nullptr;
    cudaError_t e = pinned_host ? cudaMallocHost(&p, cap) : cudaMalloc(&p, cap);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("bff arena: cannot allocate %zu bytes of %s memory", cap, pinned_host ? "pinned host" : "device");
        return nullptr;
    }
    blocks[n_blocks] = {p, cap};
    cur = n_blocks++;
    off = bytes;
    return p;
}
void Arena::reset() {
    cur = 0;
    off = 0;
    if (n_blocks <= 1) return;
    // a call outgrew the first block: one block of twice the total instead of a growing list (the list is capped, and a
    // skipped block is wasted for the rest of a call). Nothing of the previous call may still be running on the old blocks.
    size_t total = 0;
    for (int i = 0; i < n_blocks; ++i) total += blocks[i].cap;
    cudaDeviceSynchronize();
    release();
    unsigned char *p = nullptr;
    const size_t cap = 2 * total;
    const cudaError_t e = pinned_host ? cudaMallocHost(&p, cap) : cudaMalloc(&p, cap);
    if (e != cudaSuccess) {              // fall back to on-demand blocks
        cudaGetLastError();
        return;
    }
    blocks[0] = {p, cap};
    n_blocks = 1;
}
int Arena::reserve(size_t bytes) {
    cur = 0;
    off = 0;
    if (n_blocks == 1 && blocks[0].cap >= bytes) return BFF_OK;
    cudaDeviceSynchronize();
    release();
    size_t cap = bytes + bytes / 4;
    if (cap < ((size_t)32 << 20)) cap = (size_t)32 << 20;
    unsigned char *p = nullptr;
    const cudaError_t e = pinned_host ? cudaMallocHost(&p, cap) : cudaMalloc(&p, cap);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("bff arena: cannot allocate %zu bytes of %s memory", cap, pinned_host ? "pinned host" : "device");
        return BFF_ENOMEM;
    }
    blocks[0] = {p, cap};
    n_blocks = 1;
    ++generation;
    return BFF_OK;
}
void Arena::release() {
    for (int i = 0; i < n_blocks; ++i) {
        if (pinned_host) cudaFreeHost(blocks[i].p);
        else cudaFree(blocks[i].p);
    }
    n_blocks = cur = 0;
    off = 0;
}
}  // namespace bff

static int ctx_reserve(bff_ctx *c, size_t bytes) {
    if (bytes <= c->ws_bytes) return BFF_OK;
    if (c->ws) cudaFree(c->ws);
    c->ws = nullptr;
    c->ws_bytes = 0;
    if (cudaMalloc(&c->ws, bytes) != cudaSuccess) {
        cudaGetLastError();
        set_error("bff_ctx: cannot allocate %zu bytes of device workspace", bytes);
        return BFF_ENOMEM;
    }
    c->ws_bytes = bytes;
    return BFF_OK;
}

extern "C" int bff_ctx_create(int device, bff_ctx **out) {
    BFF_REQUIRE(out != nullptr, "bff_ctx_create: null out");
    BFF_CUDA(cudaSetDevice(device));
    bff_ctx *c = new bff_ctx();
    c->device = device;
    c->ws = nullptr;
    c->ws_bytes = 0;
    c->pin.pinned_host = true;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete c;
        set_error("bff_ctx_create: %s", cudaGetErrorString(e));
        return BFF_ECUDA;
    }
    *out = c;
    return BFF_OK;
}

extern "C" void bff_ctx_destroy(bff_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->ws) cudaFree(c->ws);
    c->dev.release();
    c->pin.release();
    for (int i = 0; i < c->n_lanes; ++i) { cudaStreamDestroy(c->lane[i]); cudaEventDestroy(c->lane_done[i]); }
    if (c->n_lanes) cudaEventDestroy(c->fork_ev);
    if (c->has_side) { cudaStreamDestroy(c->side); cudaEventDestroy(c->side_fork); cudaEventDestroy(c->side_join); }
    cudaStreamDestroy(c->stream);
    delete c;
}

namespace {
struct Carver {
    unsigned char *base;
    size_t off = 0;
    explicit Carver(unsigned char *b) : base(b) {}
    template <typename T>
    T *take(size_t count) {
        off = (off + 255) & ~(size_t)255;
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += count * sizeof(T);
        return p;
    }
};
}  // namespace

extern "C" int bff_lift_scene_host(bff_ctx *c, const bff_lift_desc *h, int64_t *h2d_bytes, int64_t *d2h_bytes) {
    BFF_REQUIRE(c && h, "bff_lift_scene_host: null argument");
    BFF_REQUIRE(h->n_points >= 0 && h->n_frames >= 0 && h->n_masks >= 0 && h->n_queries >= 1 && h->height > 0 && h->width > 0,
                "bff_lift_scene_host: bad sizes");
    BFF_REQUIRE(h->debug_pix == nullptr && h->debug_vis == nullptr, "bff_lift_scene_host: debug outputs are device-API only");
    BFF_CUDA(cudaSetDevice(c->device));
    const size_t n = (size_t)h->n_points, f = (size_t)h->n_frames, hw = (size_t)h->height * h->width;
    const size_t words = (n + 31) / 32, mwords = (hw + 31) / 32;
    const size_t xyz_b = n * (h->xyz_kind == BFF_XYZ_F32X4 ? 16 : 24);
    const bool bilinear = h->depth_kind == BFF_DEPTH_U16_BILINEAR;
    BFF_REQUIRE(!bilinear || (h->src_height > 0 && h->src_width > 0), "bff_lift_scene_host: bad native depth size");
    const size_t depth_b = bilinear ? f * (size_t)h->src_height * h->src_width * 2
                                    : f * hw * (h->depth_kind == BFF_DEPTH_U16 ? 2 : 4);
    const size_t nm = (size_t)h->n_masks;
    // two passes over the same carve sequence: measure, then place
    unsigned char *base = nullptr;
    bff_lift_desc d = *h;
    size_t total = 0;
    for (int pass = 0; pass < 2; ++pass) {
        Carver cv(base);
        d.xyz = cv.take<unsigned char>(xyz_b);
        d.pose_inv = cv.take<double>(f * 16);
        d.depth = cv.take<unsigned char>(depth_b);
        d.resize_x = bilinear ? cv.take<bff_resize_tap>((size_t)h->width) : nullptr;
        d.resize_y = bilinear ? cv.take<bff_resize_tap>((size_t)h->height) : nullptr;
        d.masks2d = nm ? cv.take<uint32_t>(nm * mwords) : nullptr;
        d.mask_offset = (nm && h->mask_offset) ? cv.take<int32_t>(f + 1) : nullptr;
        d.mask_query = (nm && h->mask_query) ? cv.take<int32_t>(nm) : nullptr;
        d.frame_viewed = h->frame_viewed ? cv.take<uint8_t>(f) : nullptr;
        d.rows = (h->rows && nm) ? cv.take<uint32_t>(nm * words) : nullptr;
        d.masked_counts = h->masked_counts ? cv.take<int32_t>((size_t)h->n_queries * n) : nullptr;
        d.viewed_counts = h->viewed_counts ? cv.take<int32_t>(n) : nullptr;
        total = cv.off + 256;
        if (pass == 0) {
            int rc = ctx_reserve(c, total);
            if (rc != BFF_OK) return rc;
            base = c->ws;
        }
    }
    d.intrinsics = h->intrinsics;   // passed to the kernel by value from host memory
    cudaStream_t s = c->stream;
    int64_t up = 0, down = 0;
    auto h2d = [&](const void *dst, const void *src, size_t bytes) -> cudaError_t {
        if (!bytes) return cudaSuccess;
        up += (int64_t)bytes;
        return cudaMemcpyAsync(const_cast<void *>(dst), src, bytes, cudaMemcpyHostToDevice, s);
    };
    if (n && f) {
        BFF_REQUIRE(h->xyz && h->pose_inv && h->intrinsics && h->depth, "bff_lift_scene_host: null input pointer");
        BFF_CUDA(h2d(d.xyz, h->xyz, xyz_b));
        BFF_CUDA(h2d(d.pose_inv, h->pose_inv, f * 128));
        BFF_CUDA(h2d(d.depth, h->depth, depth_b));
        if (bilinear) {
            // host tables: the caller's, or built here (bff_resize_table) when it passes none
            std::vector<bff_resize_tap> tx, ty;
            const bff_resize_tap *hx = h->resize_x, *hy = h->resize_y;
            if (!hx) { tx.resize(h->width); bff_resize_table(h->width, h->src_width, tx.data()); hx = tx.data(); }
            if (!hy) { ty.resize(h->height); bff_resize_table(h->height, h->src_height, ty.data()); hy = ty.data(); }
            BFF_CUDA(cudaMemcpyAsync(const_cast<bff_resize_tap *>(d.resize_x), hx, (size_t)h->width * 8, cudaMemcpyHostToDevice, s));
            BFF_CUDA(cudaMemcpyAsync(const_cast<bff_resize_tap *>(d.resize_y), hy, (size_t)h->height * 8, cudaMemcpyHostToDevice, s));
            BFF_CUDA(cudaStreamSynchronize(s));      // the (pageable) table vectors die with this scope
            up += ((int64_t)h->width + h->height) * 8;
        }
        if (d.masks2d) BFF_CUDA(h2d(d.masks2d, h->masks2d, nm * mwords * 4));
        if (d.mask_offset) BFF_CUDA(h2d(d.mask_offset, h->mask_offset, (f + 1) * 4));
        if (d.mask_query) BFF_CUDA(h2d(d.mask_query, h->mask_query, nm * 4));
        if (d.frame_viewed) BFF_CUDA(h2d(d.frame_viewed, h->frame_viewed, f));
    }
    int rc = bff_lift_scene(&d, s);
    if (rc != BFF_OK) return rc;
    auto d2h = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
        if (!bytes) return cudaSuccess;
        down += (int64_t)bytes;
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s);
    };
    if (d.rows) BFF_CUDA(d2h(h->rows, d.rows, nm * words * 4));
    if (d.masked_counts) BFF_CUDA(d2h(h->masked_counts, d.masked_counts, (size_t)h->n_queries * n * 4));
    if (d.viewed_counts) BFF_CUDA(d2h(h->viewed_counts, d.viewed_counts, n * 4));
    BFF_CUDA(cudaStreamSynchronize(s));
    if (h2d_bytes) *h2d_bytes = up;
    if (d2h_bytes) *d2h_bytes = down;
    return BFF_OK;
}

extern "C" int bff_mask_iou_host(bff_ctx *c, const uint32_t *a, int32_t n_a, const uint32_t *b, int32_t n_b,
                                 int64_t words, int32_t *inter, int32_t *pop_a, int32_t *pop_b, float *iou,
                                 int64_t *h2d_bytes, int64_t *d2h_bytes) {
    BFF_REQUIRE(c, "bff_mask_iou_host: null context");
    BFF_REQUIRE(n_a >= 0 && n_b >= 0 && words >= 0, "bff_mask_iou_host: negative size");
    if (h2d_bytes) *h2d_bytes = 0;
    if (d2h_bytes) *d2h_bytes = 0;
    if (n_a == 0 || n_b == 0) return BFF_OK;
    BFF_REQUIRE(a && b, "bff_mask_iou_host: null rows");
    BFF_CUDA(cudaSetDevice(c->device));
    const bool same = (a == b && n_a == n_b);
    const size_t ab = (size_t)n_a * words * 4, bb = (size_t)n_b * words * 4, pairs = (size_t)n_a * n_b;
    unsigned char *base = nullptr;
    uint32_t *da = nullptr, *db = nullptr;
    int32_t *di = nullptr, *dpa = nullptr, *dpb = nullptr;
    float *dio = nullptr;
    for (int pass = 0; pass < 2; ++pass) {
        Carver cv(base);
        da = cv.take<uint32_t>((size_t)n_a * words);
        db = same ? da : cv.take<uint32_t>((size_t)n_b * words);
        di = cv.take<int32_t>(pairs);
        dpa = cv.take<int32_t>(n_a);
        dpb = cv.take<int32_t>(n_b);
        dio = iou ? cv.take<float>(pairs) : nullptr;
        if (pass == 0) {
            int rc = ctx_reserve(c, cv.off + 256);
            if (rc != BFF_OK) return rc;
            base = c->ws;
        }
    }
    cudaStream_t s = c->stream;
    int64_t up = 0, down = 0;
    if (ab) { BFF_CUDA(cudaMemcpyAsync(da, a, ab, cudaMemcpyHostToDevice, s)); up += ab; }
    if (!same && bb) { BFF_CUDA(cudaMemcpyAsync(db, b, bb, cudaMemcpyHostToDevice, s)); up += bb; }
    int rc = bff_mask_iou(da, n_a, db, n_b, words, di, dpa, dpb, dio, s);
    if (rc != BFF_OK) return rc;
    if (inter) { BFF_CUDA(cudaMemcpyAsync(inter, di, pairs * 4, cudaMemcpyDeviceToHost, s)); down += pairs * 4; }
    if (pop_a) { BFF_CUDA(cudaMemcpyAsync(pop_a, dpa, (size_t)n_a * 4, cudaMemcpyDeviceToHost, s)); down += (size_t)n_a * 4; }
    if (pop_b) { BFF_CUDA(cudaMemcpyAsync(pop_b, dpb, (size_t)n_b * 4, cudaMemcpyDeviceToHost, s)); down += (size_t)n_b * 4; }
    if (iou) { BFF_CUDA(cudaMemcpyAsync(iou, dio, pairs * 4, cudaMemcpyDeviceToHost, s)); down += pairs * 4; }
    BFF_CUDA(cudaStreamSynchronize(s));
    if (h2d_bytes) *h2d_bytes = up;
    if (d2h_bytes) *d2h_bytes = down;
    return BFF_OK;
}


// ---------------------------------------------------------------------------------------------------
// bff_upload: SM-driven host -> device copy (see include/bff.h). 8 CTAs x 512 threads x 4 x 16-byte loads keep 256 KB of
// reads in flight: enough for 50-51 GB/s over PCIe on the boxes measured (32 KB already reaches 51 GB/s on the fastest,
// 45-48 GB/s on the slowest), and no more than needed -- every read waiting in the GPU's outbound PCIe queue also delays
// the small copies, semaphores and completions of the pipeline that runs concurrently (measured at c2: a step that takes
// 1.4 ms alone takes 2.5 ms next to 8 x 4, 3.9 ms next to 16 x 2, 4.2 ms next to 32 x 4, and the scene loop becomes
// step-bound instead of upload-bound). BFF_UPLOAD_BLOCKS / BFF_UPLOAD_UNROLL override (scripts/upload_bw.py).
// Nothing here touches a copy engine (the tail bytes included): a copy-engine transfer queued behind a long kernel in
// the upload stream would sit at the head of the engine's queue and hold back every other stream's small copies.
namespace bff {
template <int UNROLL>
__global__ void __launch_bounds__(512) upload_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, long long n16, int tail) {
    const long long stride = (long long)gridDim.x * 512;
    long long i = (long long)blockIdx.x * 512 + threadIdx.x;
    if (blockIdx.x == 0 && (int)threadIdx.x < tail)       // the last (bytes % 16) bytes: no copy-engine call at all
        reinterpret_cast<unsigned char *>(dst + n16)[threadIdx.x] = reinterpret_cast<const unsigned char *>(src + n16)[threadIdx.x];
    for (; i + (UNROLL - 1) * stride < n16; i += UNROLL * stride) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = src[i + u * stride];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) dst[i + u * stride] = v[u];
    }
    for (; i < n16; i += stride) dst[i] = src[i];
}
}  // namespace bff

extern "C" int bff_upload(void *dst, const void *src_host, int64_t bytes, void *stream_) {
    cudaStream_t s = (cudaStream_t)stream_;
    BFF_REQUIRE(bytes >= 0, "bff_upload: negative size");
    if (bytes == 0) return BFF_OK;
    BFF_REQUIRE(dst && src_host, "bff_upload: null pointer");
    BFF_REQUIRE((((uintptr_t)dst | (uintptr_t)src_host) & 15) == 0, "bff_upload: pointers must be 16-byte aligned");
    void *src_dev = nullptr;
    cudaError_t e = cudaHostGetDevicePointer(&src_dev, const_cast<void *>(src_host), 0);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("bff_upload: source is not page-locked host memory (%s)", cudaGetErrorString(e));
        return BFF_EINVAL;
    }
    const long long n16 = bytes / 16;
    const int tail = (int)(bytes & 15);
    static const int max_blocks = [] { const char *e = getenv("BFF_UPLOAD_BLOCKS"); const int b = e ? atoi(e) : 8; return b > 0 ? b : 8; }();
    static const int unroll = [] { const char *e = getenv("BFF_UPLOAD_UNROLL"); return e ? atoi(e) : 4; }();
    const int blocks = (int)std::max<long long>(1, std::min<long long>(max_blocks, (n16 + 511) / 512));
    const uint4 *sp = reinterpret_cast<const uint4 *>(src_dev);
    uint4 *dp = reinterpret_cast<uint4 *>(dst);
    if (unroll == 1) upload_kernel<1><<<blocks, 512, 0, s>>>(sp, dp, n16, tail);
    else if (unroll == 2) upload_kernel<2><<<blocks, 512, 0, s>>>(sp, dp, n16, tail);
    else upload_kernel<4><<<blocks, 512, 0, s>>>(sp, dp, n16, tail);
    BFF_LAUNCH_CHECK("upload_kernel");
    return BFF_OK;
}
