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
#include <string.h>

#include <atomic>
#include <cmath>

#include "common.cuh"

namespace bff {

#ifndef BFF_LIFT_THREADS
#define BFF_LIFT_THREADS 256
#endif
constexpr int kLiftThreads = BFF_LIFT_THREADS;
#ifndef BFF_LIFT_MIN_BLOCKS
#define BFF_LIFT_MIN_BLOCKS 6      // resident CTAs per SM the register allocation is tuned for. r02, c2, 30 frames per chunk:
                                   // 4: 0.276, 5: 0.237, 6: 0.224-0.230 (40 registers, 12 bytes spilled), 7/8: 0.241 ms (32 registers,
                                   // x/y/z and the counters live in local memory); 128-thread CTAs: 0.239 ms
#endif
#ifndef BFF_LIFT_PREFETCH
#define BFF_LIFT_PREFETCH 0        // software pipeline over the surviving frames (next frame's depth gather in flight): measured
                                   // 0.239 ms against 0.226 ms without at c2 -- the extra live registers cost more than the gather latency
#endif
#ifndef BFF_LIFT_AABB
#define BFF_LIFT_AABB 1            // cull with the warp's axis-aligned box (half extents against |n|) instead of its bounding sphere:
                                   // a census at c2 passes 30.2 % of the (warp, frame) pairs instead of 34.1 % (22.4 % hold a point in the frustum)
#endif
constexpr int kMaxFramesPerChunk = 256;   // <= kLiftThreads: one thread prepares one frame's culling planes
// Per-frame culling data in shared memory (structure of arrays, index [k * fc + frame]):
//   0..5   camera-z as a function of the world point: n(3), d, |n|(1+eps), eps|d|
//   6..25  four image-border planes (left, right, top, bottom), each n(3), d(+slack), |n|(1+eps)
constexpr int kPlaneDoubles = 26;
// A frame's inverse pose occupies kPoseStride doubles in shared memory, not 16: with 128-byte rows the lanes that build
// the culling tables (lane = frame) all hit the same banks (ncu r02: 8.4 M of the kernel's 25.5 M shared-memory wavefronts
// were conflict replays of those loads); 144-byte rows spread 32 consecutive frames over 8 bank groups and keep the rows
// 16-byte aligned for the main loop's LDS.128.
constexpr int kPoseStride = 18;
constexpr int kFrameSmemBytes = kPoseStride * 8 + kPlaneDoubles * 8 + 8 + 16;   // + depth offset (int64) + frame record (int4)
constexpr int kChunkSmemExtra = 16;                                  // alignment of the records

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
    double wd, hd;         // (double)W, (double)H
    long long hw;          // H * W (BFF_DEPTH_U16_BILINEAR: src_height * src_width, the stride of one depth frame)
    const bff_resize_tap *tap_x, *tap_y;   // BFF_DEPTH_U16_BILINEAR: per-column / per-row source taps
    int src_w, src_h;
    long long n_points;
    long long words;       // Wd
    long long mask_words;  // ceil(H*W/32)
    int n_frames;
    int frames_per_chunk;
    int height, width;
    int n_queries;
    float depth_scale;
    float depth_rcp;       // 1 / depth_scale in float32
    int fast_div;          // x / depth_scale == the 3-operation sequence of depth_value() for every uint16 x (checked on the host)
    int atomic_flush;      // >1 frame chunk: counts are combined with atomics
    // Row layout. One block: row m at rows + m * row_stride (row_stride = Wd). Sliced (n_slices > 1, the point-sharded
    // multi-GPU exchange): the words of every row are cut into n_slices column slices of row_stride words each and stored
    // slice-major, [slice][row][word in slice], so that slice p of ALL rows is one contiguous block to send to rank p.
    long long row_stride;
    long long slice_rows;  // rows per slice block (= n_masks)
    int n_slices;
    // fused exchange: slice p of the rows goes straight into rank p's receive buffer (peer memory), at row peer_row0 + m
    uint32_t *peer_rows[8];
    long long peer_row0;
    int use_peers;
};

__device__ __forceinline__ long long f64_to_i64_numpy(double v) {
    // numpy astype(int64) on x86-64: NaN / inf / out of range -> INT64_MIN
    return (v >= -9223372036854775808.0 && v < 9223372036854775808.0) ? (long long)v : (long long)0x8000000000000000ULL;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------------------------------------
// Conservative warp-level culling. It only ever SKIPS work whose result is known: a (point, frame)
// pair can be visible only if (a) its camera z lies in (z_lo, z_hi) -- the window around the non-zero
// depths of the frame set -- and (b) its pixel rounds into the image, i.e. for z > 0:
//   -0.5 <= q0/z <= W-0.5  and  -0.5 <= q1/z <= H-0.5   (q = K @ p_cam, projection_2d_to_3d.py:43-47).
// Both are linear in the world point, so a sphere (c, r) around the warp's 32 points is entirely
// outside when  n.c + d + |n| r < 0  for one of the planes. The planes are padded by half a pixel
// ([-1, W] instead of [-0.5, W-0.5]) and every bound carries a 1e-9 relative slack, 7 orders of
// magnitude above fp64 rounding; NaN/inf make every comparison false, i.e. "do not cull".
constexpr double kCullEps = 1e-9;

// |n| rounded UP (float32 square root with directed rounding; an upper bound is all the test needs), times (1 + eps)
__device__ __forceinline__ double norm_upper(double n0, double n1, double n2) {
    const float s = __double2float_ru(n0 * n0 + n1 * n1 + n2 * n2);
    return (double)__fsqrt_ru(s) * (1.0 + 2.0 * kCullEps) * (1.0 + 1e-6);
}

__device__ __forceinline__ void store_plane(double *dst, int fc, int k, double ax, double ay, double az, const double *P) {
    // camera-space functional a.p_cam  ->  world-space n.p + d with p_cam = R p + t (rows 0..2 of inv(pose))
    const double n0 = ax * P[0] + ay * P[4] + az * P[8];
    const double n1 = ax * P[1] + ay * P[5] + az * P[9];
    const double n2 = ax * P[2] + ay * P[6] + az * P[10];
    const double d = ax * P[3] + ay * P[7] + az * P[11];
    dst[(k + 0) * fc] = n0;
    dst[(k + 1) * fc] = n1;
    dst[(k + 2) * fc] = n2;
    dst[(k + 3) * fc] = d + kCullEps * fabs(d);
#if !BFF_LIFT_AABB
    dst[(k + 4) * fc] = norm_upper(n0, n1, n2);
#endif
}

// One of the five culling functionals of a frame (which = 0: camera z, 1..4: the image-border planes); the chunk's
// frames x 5 items are spread over all threads of the CTA so that no warp waits long for the tables.
__device__ __forceinline__ void build_frame_plane(const double *P, const LiftParams &p, double *dst, int fc, int which) {
    const double W = (double)p.width, H = (double)p.height;
    switch (which) {
    case 0:
        dst[0 * fc] = P[8];
        dst[1 * fc] = P[9];
        dst[2 * fc] = P[10];
        dst[3 * fc] = P[11];
#if !BFF_LIFT_AABB
        dst[4 * fc] = norm_upper(P[8], P[9], P[10]);
#endif
        dst[5 * fc] = kCullEps * fabs(P[11]);
        break;
    case 1: store_plane(dst, fc, 6, p.k[0], p.k[1], p.k[2] + 1.0, P); break;          // q0 + z >= 0      (u >= -1)
    case 2: store_plane(dst, fc, 11, -p.k[0], -p.k[1], W - p.k[2], P); break;         // W z - q0 >= 0    (u <= W)
    case 3: store_plane(dst, fc, 16, p.k[3], p.k[4], p.k[5] + 1.0, P); break;         // q1 + z >= 0      (v >= -1)
    default: store_plane(dst, fc, 21, -p.k[3], -p.k[4], H - p.k[5], P); break;        // H z - q1 >= 0    (v <= H)
    }
}

__device__ __forceinline__ double warp_min(double v) {
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Sphere (centre of the bounding box, radius to the farthest point) of the valid lanes' points.
// The radius is inflated so that n.c evaluated in floating point stays an upper bound (|n.c| error
// <= eps |n| (|cx|+|cy|+|cz|)). Returns false when the warp holds no point at all.
// Half extents of the box around the same centre, inflated like the radius (the box test adds |n_i| h_i instead of |n| r).
__device__ __forceinline__ double half_extent(double lo, double hi, double c) {
    return (0.5 * hi - 0.5 * lo) * (1.0 + kCullEps) + kCullEps * fabs(c);
}

__device__ __forceinline__ bool warp_sphere(bool valid, double x, double y, double z, double &cx, double &cy, double &cz, double &cr,
                                            double &ex, double &ey, double &ez) {
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    const double lx = warp_min(valid ? x : inf), hx = warp_max(valid ? x : -inf);
    const double ly = warp_min(valid ? y : inf), hy = warp_max(valid ? y : -inf);
    const double lz = warp_min(valid ? z : inf), hz = warp_max(valid ? z : -inf);
    cx = 0.5 * lx + 0.5 * hx;
    cy = 0.5 * ly + 0.5 * hy;
    cz = 0.5 * lz + 0.5 * hz;
    ex = half_extent(lx, hx, cx);
    ey = half_extent(ly, hy, cy);
    ez = half_extent(lz, hz, cz);
#if BFF_LIFT_AABB
    cr = 0.0;
#else
    const double dx = x - cx, dy = y - cy, dz = z - cz;
    const double d2 = warp_max(valid ? (dx * dx + dy * dy + dz * dz) : 0.0);
    cr = sqrt(d2) * (1.0 + kCullEps) + kCullEps * (fabs(cx) + fabs(cy) + fabs(cz));
#endif
    return __any_sync(0xffffffffu, valid);
}

// The same sphere for float32 clouds with one redux.sync per extreme instead of five shuffle rounds each:
// floats map monotonically to unsigned integers, the squared radius is rounded UP to float32 first, so
// the sphere only ever grows. Lanes with a NaN coordinate do not contribute (such points are never visible).
__device__ __forceinline__ unsigned f32_ordered(float v) {
    const unsigned u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ double f32_unordered(unsigned u) {
    return (double)__uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ bool warp_sphere_f32(bool valid, double x, double y, double z, double &cx, double &cy, double &cz, double &cr,
                                                double &ex, double &ey, double &ez) {
    const float fx = (float)x, fy = (float)y, fz = (float)z;      // exact: the cloud is float32
    const bool use = valid && fx == fx && fy == fy && fz == fz;
    const unsigned lo = 0xff800000u /* ordered(+inf) */, hi = 0x007fffffu /* ordered(-inf) */;
    const double lx = f32_unordered(__reduce_min_sync(0xffffffffu, use ? f32_ordered(fx) : lo));
    const double hx = f32_unordered(__reduce_max_sync(0xffffffffu, use ? f32_ordered(fx) : hi));
    const double ly = f32_unordered(__reduce_min_sync(0xffffffffu, use ? f32_ordered(fy) : lo));
    const double hy = f32_unordered(__reduce_max_sync(0xffffffffu, use ? f32_ordered(fy) : hi));
    const double lz = f32_unordered(__reduce_min_sync(0xffffffffu, use ? f32_ordered(fz) : lo));
    const double hz = f32_unordered(__reduce_max_sync(0xffffffffu, use ? f32_ordered(fz) : hi));
    cx = 0.5 * lx + 0.5 * hx;
    cy = 0.5 * ly + 0.5 * hy;
    cz = 0.5 * lz + 0.5 * hz;
    ex = half_extent(lx, hx, cx);
    ey = half_extent(ly, hy, cy);
    ez = half_extent(lz, hz, cz);
#if BFF_LIFT_AABB
    cr = 0.0;
#else
    const double dx = x - cx, dy = y - cy, dz = z - cz;
    const float d2 = use ? __double2float_ru(dx * dx + dy * dy + dz * dz) : 0.0f;   // >= 0: bit pattern order == value order
    const double r2 = (double)__uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(d2)));
    cr = sqrt(r2) * (1.0 + kCullEps) + kCullEps * (fabs(cx) + fabs(cy) + fabs(cz));
#endif
    return __any_sync(0xffffffffu, valid);
}

__device__ __forceinline__ bool warp_may_see(const double *pl, int fc, const double *sp, double z_lo, double z_hi) {
    // sp: centre (0..2), sphere radius (3), box half extents (4..6) of the warp's points
    const double cx = sp[0], cy = sp[1], cz = sp[2];
    const double zc = pl[0 * fc] * cx + pl[1 * fc] * cy + pl[2 * fc] * cz + pl[3 * fc];
#if BFF_LIFT_AABB
    // support of the box along n: sum |n_i| h_i  (<= |n| r, the sphere's); the 1e-9 factor covers the rounding of the sum itself
    const double ex = sp[4], ey = sp[5], ez = sp[6];
    const double zr = (fabs(pl[0 * fc]) * ex + fabs(pl[1 * fc]) * ey + fabs(pl[2 * fc]) * ez) * (1.0 + kCullEps) + pl[5 * fc];
#else
    const double cr = sp[3];
    const double zr = pl[4 * fc] * cr + pl[5 * fc];
#endif
    if (zc + zr < z_lo || zc - zr > z_hi) return false;      // no depth value can match anywhere in the box / sphere
    if (zc - zr > 0.0) {                                      // all of it in front of the camera: border planes apply
#pragma unroll
        for (int k = 6; k < kPlaneDoubles; k += 5) {
            const double n0 = pl[k * fc], n1 = pl[(k + 1) * fc], n2 = pl[(k + 2) * fc];
            const double v = n0 * cx + n1 * cy + n2 * cz + pl[(k + 3) * fc];
#if BFF_LIFT_AABB
            if (v + (fabs(n0) * ex + fabs(n1) * ey + fabs(n2) * ez) * (1.0 + kCullEps) < 0.0) return false;
#else
            if (v + pl[(k + 4) * fc] * cr < 0.0) return false;
#endif
        }
    }
    return true;
}

// rint(q0/z), rint(q1/z) without the two IEEE divisions whenever that is provably the same integer:
// one approximate reciprocal refined by two Newton steps (relative error ~2^-50), then u = q0 * r. The
// correctly rounded quotient and u differ by < 1e-10 for |u| < 65536, so both round to the same integer
// unless u lies within 1e-8 of a half-integer -- then (and for NaN/inf/huge values) the caller divides
// exactly. Pixel indices stay bit-exact; the divisions disappear from > 99.9999 % of the lanes.
__device__ __forceinline__ bool fast_rounded_quotients(double q0, double q1, double z, double &xr, double &yr) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(z));
    double e = fma(-z, r, 1.0);
    r = fma(r, e, r);
    e = fma(-z, r, 1.0);
    r = fma(r, e, r);
    const double u = q0 * r, v = q1 * r;
    xr = rint(u);
    yr = rint(v);
    const double lim = 0.5 - 1e-8;
    return (fabs(u) < 65536.0) && (fabs(v) < 65536.0) && (fabs(u - xr) < lim) && (fabs(v - yr) < lim);
}

// Projection of one point into one frame (projection_2d_to_3d.py:424-425, :37-48) up to the integer pixel.
// Returns true when the pixel is inside the image; Z = camera z, pix = y * W + x.
template <bool DEBUG>
__device__ __forceinline__ bool locate_pixel(const LiftParams &p, const double *P, int f, double x, double y, double z,
                                             bool valid, long long pt, double &Z, int &pix, int &px, int &py) {
    // camera z first: row 2 of inv(pose) @ [x y z 1]
    Z = fma(P[8], x, 0.0);
    Z = fma(P[9], y, Z);
    Z = fma(P[10], z, Z);
    Z = fma(P[11], 1.0, Z);
    pix = 0; px = 0; py = 0;
    if (!(valid && (DEBUG || (Z >= p.z_lo && Z <= p.z_hi)))) return false;
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
    if (!DEBUG && Z > 0.0) {
        // conservative frustum pre-test: outside => rint(q/Z) is certainly not in [0, W) x [0, H).
        // (Z <= 0 or NaN skips the pre-test and takes the exact path below.)
        if (!((q0 >= -Z) && (q0 <= p.w_hi * Z) && (q1 >= -Z) && (q1 <= p.h_hi * Z))) return false;
    }
    double xr, yr;
    if (DEBUG || !fast_rounded_quotients(q0, q1, Z, xr, yr)) {
        const double u = __ddiv_rn(q0, Z);   // divisor is camera z (:43)
        const double v = __ddiv_rn(q1, Z);
        xr = rint(u); yr = rint(v);          // np.round: half to even (:47)
    }
    if (DEBUG && p.dbg_pix != nullptr) {
        long long *o = p.dbg_pix + ((long long)f * p.n_points + pt) * 2;
        o[0] = f64_to_i64_numpy(xr);
        o[1] = f64_to_i64_numpy(yr);
    }
    if (!(xr >= 0.0 && xr < p.wd && yr >= 0.0 && yr < p.hd)) return false;
    px = (int)xr; py = (int)yr;
    pix = py * p.width + px;
    return true;
}

// uint16 sensor units -> float32 metres exactly as projection_2d_to_3d.py:431-435 (fp32 divide).
// fast_div: q = x * rcp, one residual correction -- the correctly rounded quotient for every uint16 x
// (bff_lift_scene checks all 65536 values on the host before it sets the flag); else the IEEE divide.
__device__ __forceinline__ float u16_to_metres(unsigned raw, float scale, float rcp, int fast_div) {
    const float x = (float)raw;
    if (fast_div) {
        const float q = __fmul_rn(x, rcp);
        return __fmaf_rn(__fmaf_rn(-q, scale, x), rcp, q);
    }
    return __fdiv_rn(x, scale);
}

// cv2.resize(depth, (W, H)) at ONE output pixel (projection_2d_to_3d.py:436): horizontal blends of the two source
// rows, then the vertical blend, each fmaf(s1 - s0, w, s0) in float32 (bff.h: bff_resize_table).
__device__ __forceinline__ float bilinear_sample(const unsigned short *__restrict__ img, int src_w, int src_h,
                                                 bff_resize_tap tx, bff_resize_tap ty, float scale, float rcp, int fast_div) {
    const int x0 = tx.index, x1 = min(x0 + 1, src_w - 1);
    const int y0 = ty.index, y1 = min(y0 + 1, src_h - 1);
    const unsigned short *r0 = img + (long long)y0 * src_w, *r1 = img + (long long)y1 * src_w;
    const float s00 = u16_to_metres(__ldg(r0 + x0), scale, rcp, fast_div), s01 = u16_to_metres(__ldg(r0 + x1), scale, rcp, fast_div);
    const float s10 = u16_to_metres(__ldg(r1 + x0), scale, rcp, fast_div), s11 = u16_to_metres(__ldg(r1 + x1), scale, rcp, fast_div);
    const float h0 = __fmaf_rn(__fsub_rn(s01, s00), tx.weight, s00);
    const float h1 = __fmaf_rn(__fsub_rn(s11, s10), tx.weight, s10);
    return __fmaf_rn(__fsub_rn(h1, h0), ty.weight, h0);
}

__device__ __forceinline__ bff_resize_tap load_tap(const bff_resize_tap *t) {
    const int2 v = __ldg(reinterpret_cast<const int2 *>(t));
    bff_resize_tap r;
    r.index = v.x;
    r.weight = __int_as_float(v.y);
    return r;
}

// raw depth word of an in-image pixel: the uint16 / float32 bits, or the float32 bits of the sampled value.
// frame_off = first element of the frame inside the depth stack (frame index x elements per frame).
template <int DEPTH_KIND>
__device__ __forceinline__ unsigned load_depth_raw(const LiftParams &p, long long frame_off, int pix, int px, int py) {
    if (DEPTH_KIND == BFF_DEPTH_U16_BILINEAR) {
        const unsigned short *img = reinterpret_cast<const unsigned short *>(p.depth) + frame_off;
        return __float_as_uint(bilinear_sample(img, p.src_w, p.src_h, load_tap(p.tap_x + px), load_tap(p.tap_y + py),
                                               p.depth_scale, p.depth_rcp, p.fast_div));
    }
    if (DEPTH_KIND == BFF_DEPTH_U16) return (unsigned)__ldg(reinterpret_cast<const unsigned short *>(p.depth) + frame_off + pix);
    return __ldg(reinterpret_cast<const unsigned *>(p.depth) + frame_off + pix);
}

template <int DEPTH_KIND>
__device__ __forceinline__ float depth_value(unsigned raw, const LiftParams &p) {
    if (DEPTH_KIND == BFF_DEPTH_U16) return u16_to_metres(raw, p.depth_scale, p.depth_rcp, p.fast_div);
    return __uint_as_float(raw);
}

// Shared memory of one CTA (dynamic): per frame of the chunk 16 doubles of inverse pose, kPlaneDoubles culling doubles
// (SoA), the frame's first element in the depth stack (int64) and one int4 {first mask, end mask, counts-as-viewed, -}.
template <int XYZ_KIND, int DEPTH_KIND, bool MULTI_Q, bool DEBUG>
__global__ void __launch_bounds__(kLiftThreads, BFF_LIFT_MIN_BLOCKS)
lift_kernel(const __grid_constant__ LiftParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int fc = p.frames_per_chunk;
    double *s_pose = reinterpret_cast<double *>(smem_raw);                       // fc x kPoseStride doubles (16 used)
    double *s_plane = s_pose + (size_t)fc * kPoseStride;                          // kPlaneDoubles x fc (SoA)
    long long *s_doff = reinterpret_cast<long long *>(s_plane + (size_t)fc * kPlaneDoubles);   // fc
    int4 *s_frame = reinterpret_cast<int4 *>(s_doff + fc + (fc & 1));             // fc (16-byte aligned)
    int32_t *s_acc = reinterpret_cast<int32_t *>(s_frame + fc);                   // Q x blockDim (MULTI_Q)
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ double s_sphere[kLiftThreads / 32][8];

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int f0 = blockIdx.y * fc;
    const int nf = min(fc, p.n_frames - f0);

    // ---- stage the chunk's inverse poses in shared memory: bulk async copies (TMA), one 128-byte row per frame into its
    // padded slot, all completing on one mbarrier
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        const uint32_t bytes = (uint32_t)nf * 128u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_bar)), "r"(bytes) : "memory");
        for (int i = 0; i < nf; ++i)
            asm volatile(
                "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(s_pose + i * kPoseStride)),
                "l"(p.pose_inv + (size_t)(f0 + i) * 16), "r"(128u), "r"(smem_u32(&s_bar))
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
    // where this warp's word of row 0 lives (a warp = 32 consecutive points = one word, hence one slice)
    long long row_base = word_idx;
    uint32_t *rows_out = p.rows;
    if (p.n_slices > 1) {
        const long long slice = word_idx / p.row_stride;
        if (p.use_peers) {
            rows_out = slice < p.n_slices ? p.peer_rows[slice] : nullptr;
            row_base = p.peer_row0 * p.row_stride + (word_idx - slice * p.row_stride);
        } else {
            row_base = slice * p.slice_rows * p.row_stride + (word_idx - slice * p.row_stride);
        }
    }
    // per-frame records (do not need the poses)
    for (int i = tid; i < nf; i += kLiftThreads) {
        int4 fr;
        fr.x = p.mask_offset != nullptr ? __ldg(p.mask_offset + f0 + i) : 0;
        fr.y = p.mask_offset != nullptr ? __ldg(p.mask_offset + f0 + i + 1) : 0;
        fr.z = (p.frame_viewed == nullptr || __ldg(p.frame_viewed + f0 + i)) ? 1 : 0;
        fr.w = 0;
        s_frame[i] = fr;
        s_doff[i] = (long long)(f0 + i) * p.hw;
    }

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

    // ---- per-frame culling planes (5 functionals per frame spread over the CTA), see warp_may_see()
    if (!DEBUG) {
        for (int item = tid; item < nf * 5; item += kLiftThreads) {
            const int fr = item % nf;
            build_frame_plane(s_pose + fr * kPoseStride, p, s_plane + fr, fc, item / nf);
        }
    }
    // ---- bounding sphere of the warp's 32 points (fp64, NaN coordinates ignored: such points are never visible);
    // parked in shared memory: it is read once per 32-frame batch and would otherwise hold 8 registers for the whole loop
    bool warp_has_points;
    {
        double cx, cy, cz, cr, ex, ey, ez;
        warp_has_points = XYZ_KIND == BFF_XYZ_F32X4 ? warp_sphere_f32(valid, x, y, z, cx, cy, cz, cr, ex, ey, ez)
                                                    : warp_sphere(valid, x, y, z, cx, cy, cz, cr, ex, ey, ez);
        // every lane stores (the values are warp-uniform). Not `if (lane == 0)`: with 64-byte rows nvcc 12.9 folds the row
        // address under that condition to s_sphere + 2 * tid and re-uses the folded form for the loads below, where the
        // lane bits are NOT zero -> misaligned LDS.64 (seen in the PTX; the 32-byte rows of the sphere-only table hid it)
        double *sp = s_sphere[tid >> 5];
        sp[0] = cx; sp[1] = cy; sp[2] = cz; sp[3] = cr;
        sp[4] = ex; sp[5] = ey; sp[6] = ez;
    }
    __syncthreads();
    const bool has_rows = rows_out != nullptr;

    for (int fb = 0; fb < nf; fb += 32) {
        // lane l decides for frame fb + l whether ANY point of this warp can be visible in it
        unsigned todo;
        {
            const int fl = fb + lane;
            bool see = fl < nf && warp_has_points;
            if (!DEBUG && see) {
                see = warp_may_see(s_plane + fl, fc, s_sphere[tid >> 5], p.z_lo, p.z_hi);
            }
            todo = __ballot_sync(0xffffffffu, see);
        }
        if (DEBUG && p.dbg_vis != nullptr && lane == 0 && valid) {
            // frames skipped below still owe their (all-zero) visibility word
            for (unsigned m = ~todo; m != 0u; m &= m - 1u) {
                const int fl = fb + __ffs(m) - 1;
                if (fl < nf) p.dbg_vis[(long long)(f0 + fl) * p.words + word_idx] = 0u;
            }
        }
#if BFF_LIFT_PREFETCH
        // Software pipeline over the surviving frames: the pixel of frame k+1 is located and its depth gather issued BEFORE
        // frame k's depth is consumed, so the gather's latency (L2 / DRAM: the depth stack does not fit the L2) overlaps the
        // compare -> ballot -> mask gather chain of the previous frame instead of stalling the warp.
        int fl_n = -1, pix_n = 0;
        double Z_n = 0.0;
        bool in_n = false;
        unsigned raw_n = 0u;
        if (todo != 0u) {
            fl_n = fb + __ffs(todo) - 1;
            todo &= todo - 1u;
            int px, py;
            in_n = locate_pixel<DEBUG>(p, s_pose + fl_n * kPoseStride, f0 + fl_n, x, y, z, valid, pt, Z_n, pix_n, px, py);
            raw_n = in_n ? load_depth_raw<DEPTH_KIND>(p, s_doff[fl_n], pix_n, px, py) : 0u;
        }
        while (fl_n >= 0) {
            const int fl = fl_n, pix = pix_n;
            const double Z = Z_n;
            const bool in = in_n;
            const unsigned raw = raw_n;
            fl_n = -1;
            if (todo != 0u) {
                fl_n = fb + __ffs(todo) - 1;
                todo &= todo - 1u;
                int px, py;
                in_n = locate_pixel<DEBUG>(p, s_pose + fl_n * kPoseStride, f0 + fl_n, x, y, z, valid, pt, Z_n, pix_n, px, py);
                raw_n = in_n ? load_depth_raw<DEPTH_KIND>(p, s_doff[fl_n], pix_n, px, py) : 0u;
            }
#else
        while (todo != 0u) {
            const int fl = fb + __ffs(todo) - 1;
            todo &= todo - 1u;
            double Z;
            int pix, px, py;
            const bool in = locate_pixel<DEBUG>(p, s_pose + fl * kPoseStride, f0 + fl, x, y, z, valid, pt, Z, pix, px, py);
            const unsigned raw = in ? load_depth_raw<DEPTH_KIND>(p, s_doff[fl], pix, px, py) : 0u;
#endif
            const float d = depth_value<DEPTH_KIND>(raw, p);
            // :62-66: in bounds, depth != 0 and |z - depth| < thresh in float64
            const bool vis = in && (d != 0.0f) && (fabs(Z - (double)d) < p.thresh);
            // the warp's visibility word; a warp without any visible point owes this frame nothing more
            const unsigned vw = __ballot_sync(0xffffffffu, vis);
            if (DEBUG && p.dbg_vis != nullptr && lane == 0 && valid) p.dbg_vis[(long long)(f0 + fl) * p.words + word_idx] = vw;
            if (vw == 0u) continue;
            const int4 fr = s_frame[fl];
            cnt_view += (vis ? 1 : 0) & fr.z;
            if (fr.x >= fr.y) continue;
            // masks of a frame are consecutive rows of both the packed 2D masks and the packed output rows
            const uint32_t *mp = p.masks2d + ((long long)fr.x * p.mask_words + (pix >> 5));
            long long ro = (long long)fr.x * p.row_stride + row_base;
            const int sh = pix & 31;
#pragma unroll 2
            for (int m = fr.x; m < fr.y; ++m) {
                const bool bit = vis && ((__ldg(mp) >> sh) & 1u);
                mp += p.mask_words;
                const unsigned w = __ballot_sync(0xffffffffu, bit);
                if (has_rows && w != 0u && lane == 0) rows_out[ro] = w;
                ro += p.row_stride;
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

template <int X>
static LiftKernel pick1(int depth_kind, bool multi_q, bool debug) {
    if (depth_kind == BFF_DEPTH_U16) return pick2<X, BFF_DEPTH_U16>(multi_q, debug);
    if (depth_kind == BFF_DEPTH_U16_BILINEAR) return pick2<X, BFF_DEPTH_U16_BILINEAR>(multi_q, debug);
    return pick2<X, BFF_DEPTH_F32>(multi_q, debug);
}

static LiftKernel pick(int xyz_kind, int depth_kind, bool multi_q, bool debug) {
    return xyz_kind == BFF_XYZ_F32X4 ? pick1<BFF_XYZ_F32X4>(depth_kind, multi_q, debug)
                                     : pick1<BFF_XYZ_F64X3>(depth_kind, multi_q, debug);
}

int auto_frames_per_chunk(long long n_points, int n_frames) {
    // One 32-frame ballot batch per CTA measured best on B200 (scripts/lift_variants.py: 0.33 ms at 30-38
    // frames per chunk vs 0.40 ms at 100 on the 250k-point / 300-frame scene): short CTAs balance the
    // uneven work that culling leaves and keep shared memory per CTA small. Larger chunks only when the
    // grid would become unreasonably big.
    const long long tiles = ceil_div64(n_points, kLiftThreads);
    long long fc = 32;
    while (fc < kMaxFramesPerChunk && tiles * ceil_div64(n_frames, fc) > (1LL << 21)) fc *= 2;
    if (fc > n_frames) fc = n_frames;
    // the same number of chunks, but of equal length (300 frames: 10 x 30 instead of 9 x 32 + 12; 0.2304 vs 0.2365 ms at c2)
    if (fc > 0) fc = ceil_div64(n_frames, ceil_div64(n_frames, fc));
    return (int)fc;
}

}  // namespace bff

using namespace bff;

// x / scale == fma(fma(-q, scale, x), rcp, q) with q = x * rcp for EVERY uint16 x? (exhaustive; the result for the
// last scale asked about is remembered -- scenes use one scale, 1000)
static bool fast_div_is_exact(float scale) {
    static std::atomic<unsigned long long> memo{0};   // scale bits << 2 | known << 1 | exact
    unsigned bits;
    memcpy(&bits, &scale, 4);
    const unsigned long long m = memo.load(std::memory_order_acquire);
    if ((m >> 2) == bits && (m & 2)) return m & 1;
    bool ok = scale > 0.0f && std::isfinite(scale);
    const float rc = 1.0f / scale;
    for (unsigned x = 0; ok && x < 65536u; ++x) {
        const float fx = (float)x, q = fx * rc;
        ok = fmaf(fmaf(-q, scale, fx), rc, q) == fx / scale;
    }
    memo.store(((unsigned long long)bits << 2) | 2ull | (ok ? 1ull : 0ull), std::memory_order_release);
    return ok;
}

static int validate(const bff_lift_desc *d) {
    BFF_REQUIRE(d != nullptr, "bff_lift_scene: null descriptor");
    BFF_REQUIRE(d->n_points >= 0 && d->n_frames >= 0, "bff_lift_scene: negative sizes");
    BFF_REQUIRE(d->n_points < (1LL << 31) * 32, "bff_lift_scene: too many points");
    BFF_REQUIRE(d->xyz_kind == BFF_XYZ_F32X4 || d->xyz_kind == BFF_XYZ_F64X3, "bff_lift_scene: bad xyz_kind %d", d->xyz_kind);
    BFF_REQUIRE(d->depth_kind == BFF_DEPTH_F32 || d->depth_kind == BFF_DEPTH_U16 || d->depth_kind == BFF_DEPTH_U16_BILINEAR,
                "bff_lift_scene: bad depth_kind %d", d->depth_kind);
    if (d->depth_kind == BFF_DEPTH_U16_BILINEAR) {
        BFF_REQUIRE(d->src_height > 0 && d->src_width > 0, "bff_lift_scene: bad native depth size %dx%d", d->src_height, d->src_width);
        BFF_REQUIRE(d->n_points == 0 || d->n_frames == 0 || (d->resize_x && d->resize_y), "bff_lift_scene: bilinear depth without resize tables");
        BFF_REQUIRE((((uintptr_t)d->resize_x | (uintptr_t)d->resize_y) & 7) == 0, "bff_lift_scene: resize tables must be 8-byte aligned");
    }
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
    BFF_REQUIRE(d->depth_kind == BFF_DEPTH_F32 || d->depth_scale != 0.0f, "bff_lift_scene: depth_scale is 0");
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
    if (d->ev_start) BFF_CUDA(cudaEventRecord((cudaEvent_t)d->ev_start, stream));
    // outputs are fully overwritten: zero them first (rows get only their non-zero words stored)
    const int n_slices = d->n_row_slices > 1 ? d->n_row_slices : 1;
    const long long row_stride = n_slices > 1 ? d->row_slice_words : words;
    if (n_slices > 1) BFF_REQUIRE(d->row_slice_words > 0 && (long long)n_slices * d->row_slice_words >= words, "bff_lift_scene: %d row slices of %lld words do not cover %lld words", n_slices, (long long)d->row_slice_words, words);
    bool use_peers = n_slices > 1 && d->peer_rows[0] != nullptr;
    if (use_peers) {
        BFF_REQUIRE(n_slices <= 8, "bff_lift_scene: at most 8 peers");
        for (int q = 0; q < n_slices; ++q) BFF_REQUIRE(d->peer_rows[q] != nullptr, "bff_lift_scene: peer_rows[%d] missing", q);
        BFF_REQUIRE(d->peer_row0 >= 0, "bff_lift_scene: negative peer_row0");
    }
    if (!use_peers && d->rows && d->n_masks > 0 && words > 0)
        BFF_CUDA(cudaMemsetAsync(d->rows, 0, (size_t)d->n_masks * (size_t)n_slices * row_stride * 4, stream));
    if (d->masked_counts && n > 0) BFF_CUDA(cudaMemsetAsync(d->masked_counts, 0, (size_t)d->n_queries * n * 4, stream));
    if (d->viewed_counts && n > 0) BFF_CUDA(cudaMemsetAsync(d->viewed_counts, 0, (size_t)n * 4, stream));
    if (n == 0 || d->n_frames == 0) {
        if (d->ev_stop) BFF_CUDA(cudaEventRecord((cudaEvent_t)d->ev_stop, stream));
        return BFF_OK;
    }

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
    p.wd = (double)d->width; p.hd = (double)d->height;
    p.hw = (long long)d->height * d->width;
    p.tap_x = p.tap_y = nullptr; p.src_w = p.src_h = 0;
    if (d->depth_kind == BFF_DEPTH_U16_BILINEAR) {
        p.hw = (long long)d->src_height * d->src_width;
        p.tap_x = d->resize_x; p.tap_y = d->resize_y;
        p.src_w = d->src_width; p.src_h = d->src_height;
    }
    p.n_points = n; p.words = words;
    p.row_stride = row_stride; p.slice_rows = d->n_masks; p.n_slices = n_slices;
    p.use_peers = use_peers ? 1 : 0;
    p.peer_row0 = d->peer_row0;
    for (int q = 0; q < 8; ++q) p.peer_rows[q] = use_peers && q < n_slices ? d->peer_rows[q] : nullptr;
    p.mask_words = ceil_div64((long long)d->height * d->width, 32);
    p.n_frames = d->n_frames;
    int fc = d->frames_per_chunk > 0 ? d->frames_per_chunk : auto_frames_per_chunk(n, d->n_frames);
    if (fc > kMaxFramesPerChunk) fc = kMaxFramesPerChunk;
    if (fc > d->n_frames) fc = d->n_frames;
    p.frames_per_chunk = fc;
    p.height = d->height; p.width = d->width;
    p.n_queries = d->n_queries;
    p.depth_scale = d->depth_scale;
    p.depth_rcp = d->depth_scale != 0.0f ? 1.0f / d->depth_scale : 0.0f;
    p.fast_div = d->depth_kind != BFF_DEPTH_F32 && fast_div_is_exact(d->depth_scale);
    const int chunks = (d->n_frames + fc - 1) / fc;
    p.atomic_flush = chunks > 1;

    const bool multi_q = d->n_queries > 1;
    const bool debug = d->debug_pix != nullptr || d->debug_vis != nullptr;
    LiftKernel k = pick(d->xyz_kind, d->depth_kind, multi_q, debug);
    const size_t smem = (size_t)fc * kFrameSmemBytes + kChunkSmemExtra + (multi_q ? (size_t)d->n_queries * kLiftThreads * 4 : 0);
    BFF_REQUIRE(smem <= 200 * 1024, "bff_lift_scene: %d queries need %zu bytes of shared memory", d->n_queries, smem);
    BFF_CUDA(cudaFuncSetAttribute((const void *)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div64(n, kLiftThreads), (unsigned)chunks);
    k<<<grid, kLiftThreads, smem, stream>>>(p);
    BFF_LAUNCH_CHECK("lift_kernel");
    if (d->ev_stop) BFF_CUDA(cudaEventRecord((cudaEvent_t)d->ev_stop, stream));
    return BFF_OK;
}


// ---------------------------------------------------------------------------------------------------
// Stand-alone forms of the reference's two per-frame helpers for callers that already hold pixel indices
// (bff_b200/reference_api.py): compute_visibility_mask_tensor (:51-70) and
// compute_visible_masked_pts_tensor (:73-92). The scene path uses the fused lift_kernel instead.
namespace bff {
__global__ void __launch_bounds__(256)
visibility_kernel(const long long *__restrict__ pix, const double *__restrict__ z, long long n, const float *__restrict__ depth,
                  int height, int width, double thresh, uint32_t *__restrict__ vis) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    bool v = false;
    if (i < n) {
        const long long px = pix[2 * i], py = pix[2 * i + 1];
        if (px >= 0 && px < width && py >= 0 && py < height) {
            const float d = __ldg(depth + py * width + px);
            v = (d != 0.0f) && (fabs(z[i] - (double)d) < thresh);
        }
    }
    const unsigned w = __ballot_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && i < n) vis[i >> 5] = w;
}

__global__ void __launch_bounds__(256)
mask_gather_kernel(const long long *__restrict__ pix, const uint32_t *__restrict__ vis, long long n, const uint32_t *__restrict__ masks2d,
                   long long mask_words, int height, int width, long long words, uint32_t *__restrict__ rows) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    const int m = blockIdx.y;
    bool bit = false;
    if (i < n && ((vis[i >> 5] >> (i & 31)) & 1u)) {
        // the caller's visibility flags are trusted for WHICH points to sample, not for the pixel being inside the image:
        // a flagged point with an out-of-range pixel reads nothing (the reference would raise or wrap there)
        const long long px = pix[2 * i], py = pix[2 * i + 1];
        if (px >= 0 && px < width && py >= 0 && py < height) {
            const long long p = py * width + px;
            bit = (__ldg(masks2d + (long long)m * mask_words + (p >> 5)) >> (p & 31)) & 1u;
        }
    }
    const unsigned w = __ballot_sync(0xffffffffu, bit);
    if ((threadIdx.x & 31) == 0 && i < n) rows[(long long)m * words + (i >> 5)] = w;
}
}  // namespace bff

extern "C" int bff_visibility(const int64_t *pix, const double *z, int64_t n, const float *depth, int32_t height, int32_t width,
                              double thresh, uint32_t *vis, void *stream_) {
    BFF_REQUIRE(n >= 0 && height > 0 && width > 0, "bff_visibility: bad sizes");
    if (n == 0) return BFF_OK;
    BFF_REQUIRE(pix && z && depth && vis, "bff_visibility: null pointer");
    visibility_kernel<<<(unsigned)ceil_div64(n, 256), 256, 0, (cudaStream_t)stream_>>>((const long long *)pix, z, n, depth, height, width, thresh, vis);
    BFF_LAUNCH_CHECK("visibility_kernel");
    return BFF_OK;
}

extern "C" int bff_mask_gather(const int64_t *pix, const uint32_t *vis, int64_t n, const uint32_t *masks2d, int32_t n_masks,
                               int32_t height, int32_t width, uint32_t *rows, void *stream_) {
    BFF_REQUIRE(n >= 0 && n_masks >= 0 && height > 0 && width > 0, "bff_mask_gather: bad sizes");
    if (n == 0 || n_masks == 0) return BFF_OK;
    BFF_REQUIRE(pix && vis && masks2d && rows, "bff_mask_gather: null pointer");
    BFF_REQUIRE(n_masks <= 65535, "bff_mask_gather: too many masks");
    dim3 grid((unsigned)ceil_div64(n, 256), (unsigned)n_masks);
    mask_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>((const long long *)pix, vis, n, masks2d,
                                                                ceil_div64((long long)height * width, 32), height, width, ceil_div64(n, 32), rows);
    BFF_LAUNCH_CHECK("mask_gather_kernel");
    return BFF_OK;
}


// ---------------------------------------------------------------------------------------------------
// Depth ingest of projection_2d_to_3d.py:431-436 as a stand-alone device op (bff.h: bff_depth_resize).
namespace bff {
__global__ void __launch_bounds__(256)
depth_resize_kernel(const unsigned short *__restrict__ src, int src_h, int src_w, float scale, float rcp, int fast_div,
                    const bff_resize_tap *__restrict__ tap_x, const bff_resize_tap *__restrict__ tap_y, int dst_h, int dst_w,
                    float *__restrict__ dst) {
    const int x = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y, f = blockIdx.z;
    if (x >= dst_w) return;
    const unsigned short *img = src + (long long)f * src_h * src_w;
    dst[((long long)f * dst_h + y) * dst_w + x] = bilinear_sample(img, src_w, src_h, load_tap(tap_x + x), load_tap(tap_y + y), scale, rcp, fast_div);
}
}  // namespace bff

extern "C" int bff_resize_table(int32_t n_dst, int32_t n_src, bff_resize_tap *taps) {
    BFF_REQUIRE(n_dst > 0 && n_src > 0 && taps, "bff_resize_table: bad arguments");
    const double scale = (double)n_src / (double)n_dst;
    for (int i = 0; i < n_dst; ++i) {
        const double f = ((double)i + 0.5) * scale - 0.5;
        double fl = floor(f);
        int idx = (int)fl;
        float w = (float)(f - fl);
        if (idx < 0) { idx = 0; w = 0.0f; }
        if (idx >= n_src - 1) { idx = n_src - 1; w = 0.0f; }
        taps[i].index = idx;
        taps[i].weight = w;
    }
    return BFF_OK;
}

extern "C" int bff_depth_resize(const uint16_t *src, int32_t n_frames, int32_t src_h, int32_t src_w, float depth_scale,
                                const bff_resize_tap *tap_x, const bff_resize_tap *tap_y, int32_t dst_h, int32_t dst_w,
                                float *dst, void *stream_) {
    BFF_REQUIRE(n_frames >= 0 && src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "bff_depth_resize: bad sizes");
    BFF_REQUIRE(depth_scale != 0.0f, "bff_depth_resize: depth_scale is 0");
    if (n_frames == 0) return BFF_OK;
    BFF_REQUIRE(src && tap_x && tap_y && dst, "bff_depth_resize: null pointer");
    BFF_REQUIRE(n_frames <= 65535 && dst_h <= 65535, "bff_depth_resize: too many frames / rows for one launch");
    dim3 grid((unsigned)ceil_div64(dst_w, 256), (unsigned)dst_h, (unsigned)n_frames);
    depth_resize_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(src, src_h, src_w, depth_scale, 1.0f / depth_scale,
                                                                  fast_div_is_exact(depth_scale) ? 1 : 0, tap_x, tap_y, dst_h, dst_w, dst);
    BFF_LAUNCH_CHECK("depth_resize_kernel");
    return BFF_OK;
}
