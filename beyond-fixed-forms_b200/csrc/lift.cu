// bff_lift_scene: projection + depth visibility + 2D-mask gather + vote accumulation for ALL
// frames of a scene in one launch.
//
// Replaces the frame loops of tools/projection_2d_to_3d.py:413-461 and :538-567
// (compute_projected_pts_tensor :37-48, compute_visibility_mask_tensor :51-70,
// compute_visible_masked_pts_tensor :73-92, masked_counts / viewed_counts updates).
//
// Layout: one thread owns one point and keeps its xyz in registers while it walks a chunk of
// frames (grid.x = point tiles, grid.y = frame chunks). A warp therefore owns 32 consecutive
// points = exactly one word of every packed output row, so a __ballot_sync over the warp IS
// the row word and votes are accumulated in registers (Q = 1) or in thread-private shared
// memory columns (Q > 1) and flushed once per chunk. Inverse poses of the chunk are brought
// into shared memory with one bulk async copy (TMA, cp.async.bulk + mbarrier).
//
// Arithmetic: float64 with the reference's operation order -- the two BLAS matmuls are FMA
// chains from a zero accumulator, the divisor is camera z, rounding is
// half-to-even, the depth compare promotes float32 depth to double. Compiled with -fmad=false
// so only the explicit fma() calls fuse.
#include <cuda.h>
#include <math.h>

#include "common.cuh"

namespace bff {

constexpr int kLiftThreads = 256;
constexpr int kMaxFramesPerChunk = 256;

struct LiftParams {
    const void *xyz;
    const double *pose_inv;
    const void *depth;
    const uint32_t *masks2d;
    const int32_t *mask_offset;
    const int32_t *mask_query;
    const uint8_t *frame_viewed;
    uint32_t *rows;
    int32_t *mc;
    int32_t *vc;
    long long *dbg_pix;
    uint32_t *dbg_vis;
    double k[6];           // intrinsics rows 0 and 1
    double thresh;
    double z_lo, z_hi;     // camera-z window outside which no depth can match
    double w_hi, h_hi;     // (W + 1), (H + 1) for the conservative frustum pre-test
    long long n_points;
    long long words;       // Wd
    long long mask_words;  // ceil(H*W/32)
    int n_frames;
    int frames_per_chunk;
    int height, width;
    int n_queries;
    float depth_scale;
    int atomic_flush;      // >1 frame chunk: counts are combined with atomics
};

__device__ __forceinline__ long long f64_to_i64_numpy(double v) {
    // numpy astype(int64) on x86-64: NaN / inf / out of range -> INT64_MIN
    return (v >= -9223372036854775808.0 && v < 9223372036854775808.0) ? (long long)v : (long long)0x8000000000000000ULL;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int XYZ_KIND, int DEPTH_KIND, bool MULTI_Q, bool DEBUG>
__global__ void __launch_bounds__(kLiftThreads)
lift_kernel(const __grid_constant__ LiftParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *s_pose = reinterpret_cast<double *>(smem_raw);                       // fc x 16 doubles
    int32_t *s_acc = reinterpret_cast<int32_t *>(smem_raw + (size_t)p.frames_per_chunk * 128);  // Q x blockDim (MULTI_Q)
    __shared__ __align__(8) unsigned long long s_bar;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int f0 = blockIdx.y * p.frames_per_chunk;
    const int nf = min(p.frames_per_chunk, p.n_frames - f0);

    // ---- stage the chunk's inverse poses in shared memory: one bulk async copy (TMA)
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        const uint32_t bytes = (uint32_t)nf * 128u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_bar)), "r"(bytes) : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(s_pose)),
            "l"(p.pose_inv + (size_t)f0 * 16), "r"(bytes), "r"(smem_u32(&s_bar))
            : "memory");
    }

    // ---- this thread's point (overlaps the pose copy)
    const long long pt = (long long)blockIdx.x * kLiftThreads + tid;
    const bool valid = pt < p.n_points;
    double x = 0.0, y = 0.0, z = 0.0;
    if (valid) {
        if (XYZ_KIND == BFF_XYZ_F32X4) {
            const float4 v = __ldg(reinterpret_cast<const float4 *>(p.xyz) + pt);
            x = (double)v.x; y = (double)v.y; z = (double)v.z;
        } else {
            const double *v = reinterpret_cast<const double *>(p.xyz) + 3 * pt;
            x = __ldg(v); y = __ldg(v + 1); z = __ldg(v + 2);
        }
    }
    if (MULTI_Q) {
        for (int q = 0; q < p.n_queries; ++q) s_acc[q * kLiftThreads + tid] = 0;
    }
    int cnt_mask = 0, cnt_view = 0;
    const long long word_idx = pt >> 5;

    // wait for the poses
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred P1;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], 0;\n\t"
                "selp.u32 %0, 1, 0, P1;\n\t}"
                : "=r"(done)
                : "r"(smem_u32(&s_bar))
                : "memory");
        }
    }

    for (int fl = 0; fl < nf; ++fl) {
        const int f = f0 + fl;
        const double *P = s_pose + fl * 16;
        // camera z first: projection_2d_to_3d.py:424-425, row 2 of inv(pose) @ [x y z 1]
        double Z = fma(P[8], x, 0.0);
        Z = fma(P[9], y, Z);
        Z = fma(P[10], z, Z);
        Z = fma(P[11], 1.0, Z);
        bool vis = false;
        long long pidx = 0;
        bool cand = valid && (DEBUG || (Z >= p.z_lo && Z <= p.z_hi));
        if (cand) {
            double X = fma(P[0], x, 0.0);
            X = fma(P[1], y, X);
            X = fma(P[2], z, X);
            X = fma(P[3], 1.0, X);
            double Y = fma(P[4], x, 0.0);
            Y = fma(P[5], y, Y);
            Y = fma(P[6], z, Y);
            Y = fma(P[7], 1.0, Y);
            // K @ p, rows 0 and 1 (:43)
            double q0 = fma(p.k[0], X, 0.0);
            q0 = fma(p.k[1], Y, q0);
            q0 = fma(p.k[2], Z, q0);
            double q1 = fma(p.k[3], X, 0.0);
            q1 = fma(p.k[4], Y, q1);
            q1 = fma(p.k[5], Z, q1);
            bool inside = true;
            if (!DEBUG) {
                // conservative frustum pre-test: outside => rint(q/Z) is certainly not in [0, W)
                const double nz = -Z;
                const double bw = p.w_hi * Z, bh = p.h_hi * Z;
                inside = (q0 >= fmin(nz, bw)) && (q0 <= fmax(nz, bw)) && (q1 >= fmin(nz, bh)) && (q1 <= fmax(nz, bh));
            }
            if (inside) {
                const double u = __ddiv_rn(q0, Z);   // divisor is camera z (:43)
                const double v = __ddiv_rn(q1, Z);
                const double xr = rint(u), yr = rint(v);   // np.round: half to even (:47)
                if (DEBUG) {
                    long long *o = p.dbg_pix + ((long long)f * p.n_points + pt) * 2;
                    o[0] = f64_to_i64_numpy(xr);
                    o[1] = f64_to_i64_numpy(yr);
                }
                if (xr >= 0.0 && xr < (double)p.width && yr >= 0.0 && yr < (double)p.height) {
                    pidx = (long long)yr * p.width + (long long)xr;
                    float d;
                    const long long off = (long long)f * p.height * p.width + pidx;
                    if (DEPTH_KIND == BFF_DEPTH_U16)
                        d = __fdiv_rn((float)__ldg(reinterpret_cast<const unsigned short *>(p.depth) + off), p.depth_scale);
                    else
                        d = __ldg(reinterpret_cast<const float *>(p.depth) + off);
                    // :62-66: depth != 0 and |z - depth| < thresh in float64
                    vis = (d != 0.0f) && (fabs(Z - (double)d) < p.thresh);
                }
            }
        }
        if (p.dbg_vis != nullptr) {
            const unsigned vw = __ballot_sync(0xffffffffu, vis);
            if (lane == 0 && (pt < p.n_points)) p.dbg_vis[(long long)f * p.words + word_idx] = vw;
        }
        if (p.frame_viewed == nullptr || p.frame_viewed[f]) cnt_view += vis ? 1 : 0;
        if (p.mask_offset != nullptr) {
            const int m0 = p.mask_offset[f], m1 = p.mask_offset[f + 1];
            for (int m = m0; m < m1; ++m) {
                bool bit = false;
                if (vis) bit = (__ldg(p.masks2d + (long long)m * p.mask_words + (pidx >> 5)) >> (pidx & 31)) & 1u;
                const unsigned w = __ballot_sync(0xffffffffu, bit);
                if (w != 0u && lane == 0 && p.rows != nullptr) p.rows[(long long)m * p.words + word_idx] = w;
                if (MULTI_Q) {
                    if (bit) s_acc[(p.mask_query ? p.mask_query[m] : 0) * kLiftThreads + tid] += 1;
                } else {
                    cnt_mask += bit ? 1 : 0;
                }
            }
        }
    }

    // ---- flush votes once per (point, chunk)
    if (valid) {
        if (p.vc != nullptr) {
            if (!p.atomic_flush) p.vc[pt] = cnt_view;
            else if (cnt_view) atomicAdd(p.vc + pt, cnt_view);
        }
        if (p.mc != nullptr) {
            if (MULTI_Q) {
                for (int q = 0; q < p.n_queries; ++q) {
                    const int c = s_acc[q * kLiftThreads + tid];
                    if (!p.atomic_flush) p.mc[(long long)q * p.n_points + pt] = c;
                    else if (c) atomicAdd(p.mc + (long long)q * p.n_points + pt, c);
                }
            } else {
                if (!p.atomic_flush) p.mc[pt] = cnt_mask;
                else if (cnt_mask) atomicAdd(p.mc + pt, cnt_mask);
            }
        }
    }
}

using LiftKernel = void (*)(const LiftParams);

template <int X, int D>
static LiftKernel pick2(bool multi_q, bool debug) {
    if (debug) return multi_q ? lift_kernel<X, D, true, true> : lift_kernel<X, D, false, true>;
    return multi_q ? lift_kernel<X, D, true, false> : lift_kernel<X, D, false, false>;
}

static LiftKernel pick(int xyz_kind, int depth_kind, bool multi_q, bool debug) {
    if (xyz_kind == BFF_XYZ_F32X4)
        return depth_kind == BFF_DEPTH_U16 ? pick2<BFF_XYZ_F32X4, BFF_DEPTH_U16>(multi_q, debug)
                                           : pick2<BFF_XYZ_F32X4, BFF_DEPTH_F32>(multi_q, debug);
    return depth_kind == BFF_DEPTH_U16 ? pick2<BFF_XYZ_F64X3, BFF_DEPTH_U16>(multi_q, debug)
                                       : pick2<BFF_XYZ_F64X3, BFF_DEPTH_F32>(multi_q, debug);
}

int auto_frames_per_chunk(long long n_points, int n_frames) {
    const long long tiles = ceil_div64(n_points, kLiftThreads);
    // aim for >= ~16 CTAs per SM in total so the tail is short, keep chunks >= 8 frames
    long long chunks = ceil_div64((long long)kNumSMs * 16, tiles > 0 ? tiles : 1);
    if (chunks < 1) chunks = 1;
    long long fc = ceil_div64(n_frames, chunks);
    if (fc < 8) fc = 8;
    if (fc > kMaxFramesPerChunk) fc = kMaxFramesPerChunk;
    if (fc > n_frames) fc = n_frames;
    return (int)fc;
}

}  // namespace bff

using namespace bff;

static int validate(const bff_lift_desc *d) {
    BFF_REQUIRE(d != nullptr, "bff_lift_scene: null descriptor");
    BFF_REQUIRE(d->n_points >= 0 && d->n_frames >= 0, "bff_lift_scene: negative sizes");
    BFF_REQUIRE(d->n_points < (1LL << 31) * 32, "bff_lift_scene: too many points");
    BFF_REQUIRE(d->xyz_kind == BFF_XYZ_F32X4 || d->xyz_kind == BFF_XYZ_F64X3, "bff_lift_scene: bad xyz_kind %d", d->xyz_kind);
    BFF_REQUIRE(d->depth_kind == BFF_DEPTH_F32 || d->depth_kind == BFF_DEPTH_U16, "bff_lift_scene: bad depth_kind %d", d->depth_kind);
    BFF_REQUIRE(d->height > 0 && d->width > 0, "bff_lift_scene: bad image size %dx%d", d->height, d->width);
    BFF_REQUIRE((long long)d->height * d->width < (1LL << 31), "bff_lift_scene: image too large");
    BFF_REQUIRE(d->n_queries >= 1, "bff_lift_scene: n_queries must be >= 1");
    BFF_REQUIRE(d->n_masks >= 0, "bff_lift_scene: negative n_masks");
    if (d->n_points > 0 && d->n_frames > 0) {
        BFF_REQUIRE(d->xyz && d->pose_inv && d->intrinsics && d->depth, "bff_lift_scene: null input pointer");
        BFF_REQUIRE(d->n_masks == 0 || (d->masks2d && d->mask_offset), "bff_lift_scene: masks without masks2d/mask_offset");
        BFF_REQUIRE(((uintptr_t)d->pose_inv & 15) == 0, "bff_lift_scene: pose_inv must be 16-byte aligned");
        BFF_REQUIRE(d->xyz_kind != BFF_XYZ_F32X4 || ((uintptr_t)d->xyz & 15) == 0, "bff_lift_scene: float4 xyz must be 16-byte aligned");
    }
    BFF_REQUIRE(d->depth_kind != BFF_DEPTH_U16 || d->depth_scale != 0.0f, "bff_lift_scene: depth_scale is 0");
    BFF_REQUIRE(d->frames_per_chunk >= 0, "bff_lift_scene: negative frames_per_chunk");
    return BFF_OK;
}

extern "C" int bff_lift_scene_launches(const bff_lift_desc *d) {
    if (!d || d->n_points <= 0 || d->n_frames <= 0) return 0;
    return 1;
}

extern "C" int bff_lift_scene(const bff_lift_desc *d, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = validate(d);
    if (rc != BFF_OK) return rc;
    const long long words = ceil_div64(d->n_points, 32);
    const long long n = d->n_points;
    // outputs are fully overwritten: zero them first (rows get only their non-zero words stored)
    if (d->rows && d->n_masks > 0 && words > 0)
        BFF_CUDA(cudaMemsetAsync(d->rows, 0, (size_t)d->n_masks * words * 4, stream));
    if (d->masked_counts && n > 0) BFF_CUDA(cudaMemsetAsync(d->masked_counts, 0, (size_t)d->n_queries * n * 4, stream));
    if (d->viewed_counts && n > 0) BFF_CUDA(cudaMemsetAsync(d->viewed_counts, 0, (size_t)n * 4, stream));
    if (n == 0 || d->n_frames == 0) return BFF_OK;

    LiftParams p;
    p.xyz = d->xyz; p.pose_inv = d->pose_inv; p.depth = d->depth; p.masks2d = d->masks2d;
    p.mask_offset = d->n_masks > 0 ? d->mask_offset : nullptr;
    p.mask_query = d->mask_query; p.frame_viewed = d->frame_viewed;
    p.rows = d->rows; p.mc = d->masked_counts; p.vc = d->viewed_counts;
    p.dbg_pix = (long long *)d->debug_pix; p.dbg_vis = d->debug_vis;
    // intrinsics travel by value; they are host-readable only in the host entry point, so the
    // device-pointer entry point copies them D2H-free: the caller passes a HOST pointer here.
    for (int i = 0; i < 6; ++i) p.k[i] = d->intrinsics[i];
    p.thresh = d->depth_thresh;
    if (d->depth_min <= d->depth_max) {
        p.z_lo = d->depth_min - 2.0 * fabs(d->depth_thresh);
        p.z_hi = d->depth_max + 2.0 * fabs(d->depth_thresh);
    } else {
        p.z_lo = -INFINITY; p.z_hi = INFINITY;
    }
    p.w_hi = (double)d->width + 1.0; p.h_hi = (double)d->height + 1.0;
    p.n_points = n; p.words = words;
    p.mask_words = ceil_div64((long long)d->height * d->width, 32);
    p.n_frames = d->n_frames;
    int fc = d->frames_per_chunk > 0 ? d->frames_per_chunk : auto_frames_per_chunk(n, d->n_frames);
    if (fc > kMaxFramesPerChunk) fc = kMaxFramesPerChunk;
    if (fc > d->n_frames) fc = d->n_frames;
    p.frames_per_chunk = fc;
    p.height = d->height; p.width = d->width;
    p.n_queries = d->n_queries;
    p.depth_scale = d->depth_scale;
    const int chunks = (d->n_frames + fc - 1) / fc;
    p.atomic_flush = chunks > 1;

    const bool multi_q = d->n_queries > 1;
    const bool debug = d->debug_pix != nullptr;
    LiftKernel k = pick(d->xyz_kind, d->depth_kind, multi_q, debug);
    const size_t smem = (size_t)fc * 128 + (multi_q ? (size_t)d->n_queries * kLiftThreads * 4 : 0);
    BFF_REQUIRE(smem <= 200 * 1024, "bff_lift_scene: %d queries need %zu bytes of shared memory", d->n_queries, smem);
    BFF_CUDA(cudaFuncSetAttribute((const void *)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div64(n, kLiftThreads), (unsigned)chunks);
    k<<<grid, kLiftThreads, smem, stream>>>(p);
    BFF_LAUNCH_CHECK("lift_kernel");
    return BFF_OK;
}
