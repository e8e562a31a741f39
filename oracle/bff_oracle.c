/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the BeyondFF lifting / IoU arithmetic.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this library. The product path (beyond-fixed-forms_b200/) never links or calls it.
 *
 * Parity status: PINNED. tests/test_oracle_golden.py checks every function below against
 * outputs of the unmodified reference functions (imported from /root/reference by
 * oracle/make_golden.py in the build container, frozen under tests/golden/).
 *
 * Arithmetic follows the reference file:line cited at each function. The two small matmuls
 * of the reference run through OpenBLAS dgemm, whose result for K<=4 is the FMA chain
 * acc=0; acc=fma(a_k,b_k,acc) for k ascending -- restated here with
 * explicit fma(); build with -ffp-contract=off so nothing else is contracted.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* numpy .astype(np.int64) of a double on x86-64 (cvttsd2si): NaN, +-inf and anything
 * outside int64 become INT64_MIN. */
static inline int64_t to_i64(double v) {
    if (!(v >= -9223372036854775808.0 && v < 9223372036854775808.0)) return INT64_MIN;
    return (int64_t)v;
}

/* tools/projection_2d_to_3d.py:424-425 -- scene_pts = (inv(pose) @ pcd).T[:, :3]
 * tools/projection_2d_to_3d.py:37-48  -- (K @ pts / pts[2])[:2].T, np.round, astype(int64)
 * xyz    (N,3) f64   world points
 * pinv   (4,4) f64   row-major np.linalg.inv(cam_pose) (host LAPACK result, passed in)
 * kmat   (3,3) f64   row-major intrinsics
 * cam    (N,3) f64   out: camera-space points (rows 0..2 of the transform)
 * pix    (N,2) i64   out: [x, y] pixel indices */
ORACLE_API void oracle_project(const double *xyz, const double *pinv, const double *kmat,
                               int64_t n, double *cam, int64_t *pix) {
    for (int64_t i = 0; i < n; ++i) {
        const double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
        double c[3];
        for (int r = 0; r < 3; ++r) {
            double acc = 0.0;
            acc = fma(pinv[4 * r + 0], x, acc);
            acc = fma(pinv[4 * r + 1], y, acc);
            acc = fma(pinv[4 * r + 2], z, acc);
            acc = fma(pinv[4 * r + 3], 1.0, acc);
            c[r] = acc;
        }
        double q[2];
        for (int r = 0; r < 2; ++r) {
            double acc = 0.0;
            acc = fma(kmat[3 * r + 0], c[0], acc);
            acc = fma(kmat[3 * r + 1], c[1], acc);
            acc = fma(kmat[3 * r + 2], c[2], acc);
            q[r] = acc;
        }
        /* divisor is camera-space z (pts[2]), not the third projected row (:43) */
        const double u = q[0] / c[2], v = q[1] / c[2];
        cam[3 * i] = c[0]; cam[3 * i + 1] = c[1]; cam[3 * i + 2] = c[2];
        pix[2 * i] = to_i64(nearbyint(u));     /* np.round = half to even (:47) */
        pix[2 * i + 1] = to_i64(nearbyint(v));
    }
}

/* tools/projection_2d_to_3d.py:51-70 -- in-bounds, depth != 0, |z - depth| < thresh
 * (float32 depth promoted to double for the subtraction; strict <). */
ORACLE_API void oracle_visibility(const double *cam, const int64_t *pix, const float *depth,
                                  int64_t n, int h, int w, double thresh, uint8_t *vis) {
    for (int64_t i = 0; i < n; ++i) {
        const int64_t px = pix[2 * i], py = pix[2 * i + 1];
        uint8_t v = 0;
        if (px >= 0 && px < w && py >= 0 && py < h) {
            const float d = depth[py * (int64_t)w + px];
            v = (d != 0.0f) && (fabs(cam[3 * i + 2] - (double)d) < thresh);
        }
        vis[i] = v;
    }
}

/* tools/projection_2d_to_3d.py:73-92 -- rows[m, vis] = pred_masks[m, y, x] (nonzero -> True)
 * masks (M,H,W) u8 ; rows (M,N) u8 out */
ORACLE_API void oracle_mask_gather(const int64_t *pix, const uint8_t *vis, const uint8_t *masks,
                                   int64_t n, int m, int h, int w, uint8_t *rows) {
    for (int k = 0; k < m; ++k) {
        const uint8_t *mk = masks + (int64_t)k * h * w;
        uint8_t *row = rows + (int64_t)k * n;
        for (int64_t i = 0; i < n; ++i)
            row[i] = vis[i] ? (mk[pix[2 * i + 1] * (int64_t)w + pix[2 * i]] != 0) : 0;
    }
}

/* Whole-scene lifting: the frame loops of tools/projection_2d_to_3d.py:413-461 (mask frames:
 * rows + masked_counts, one increment per MASK, :459-461) and :538-567 (viewed_counts over
 * the frames flagged in `viewed`).
 * pinv (F,4,4); depth (F,H,W) f32; masks (sumM,H,W) u8; mask_off (F+1) i32; mask_query (sumM) i32
 * or NULL; viewed (F) u8 or NULL (= all frames).
 * rows (sumM,N) u8 out (may be NULL); mc (Q,N) i32 out; vc (N) i32 out. */
ORACLE_API void oracle_lift_scene(const double *xyz, int64_t n, const double *pinv, const double *kmat,
                                  int f, const float *depth, int h, int w, double thresh,
                                  const uint8_t *masks, const int32_t *mask_off,
                                  const int32_t *mask_query, const uint8_t *viewed, int q,
                                  uint8_t *rows, int32_t *mc, int32_t *vc) {
    double *cam = (double *)malloc(sizeof(double) * 3 * n);
    int64_t *pix = (int64_t *)malloc(sizeof(int64_t) * 2 * n);
    uint8_t *vis = (uint8_t *)malloc(n);
    memset(mc, 0, sizeof(int32_t) * n * (q > 0 ? q : 1));
    memset(vc, 0, sizeof(int32_t) * n);
    for (int fi = 0; fi < f; ++fi) {
        oracle_project(xyz, pinv + 16 * fi, kmat, n, cam, pix);
        oracle_visibility(cam, pix, depth + (int64_t)fi * h * w, n, h, w, thresh, vis);
        if (!viewed || viewed[fi]) {
            for (int64_t i = 0; i < n; ++i) vc[i] += vis[i];
        }
        for (int k = mask_off[fi]; k < mask_off[fi + 1]; ++k) {
            const uint8_t *mk = masks + (int64_t)k * h * w;
            int32_t *cnt = mc + (int64_t)(mask_query ? mask_query[k] : 0) * n;
            uint8_t *row = rows ? rows + (int64_t)k * n : NULL;
            for (int64_t i = 0; i < n; ++i) {
                const uint8_t b = vis[i] ? (mk[pix[2 * i + 1] * (int64_t)w + pix[2 * i]] != 0) : 0;
                if (row) row[i] = b;
                cnt[i] += b;
            }
        }
    }
    free(cam); free(pix); free(vis);
}

/* tools/projection_2d_to_3d.py:149-166 (calculate_iou) and tools/refinement.py:69-90
 * (calculate_iou_between_stages, before its final transpose): intersection by 0/1 product,
 * union = (sA_i + sB_j) - I in float32, IoU = I / union in float32 (0/0 -> NaN).
 * a (nA,N) u8, b (nB,N) u8 -> inter (nA,nB) i32, iou (nA,nB) f32 */
ORACLE_API void oracle_mask_iou(const uint8_t *a, const uint8_t *b, int na, int nb, int64_t n,
                                int32_t *inter, float *iou) {
    int32_t *sa = (int32_t *)calloc(na > 0 ? na : 1, sizeof(int32_t));
    int32_t *sb = (int32_t *)calloc(nb > 0 ? nb : 1, sizeof(int32_t));
    for (int i = 0; i < na; ++i) { int32_t s = 0; for (int64_t p = 0; p < n; ++p) s += a[i * n + p] != 0; sa[i] = s; }
    for (int j = 0; j < nb; ++j) { int32_t s = 0; for (int64_t p = 0; p < n; ++p) s += b[j * n + p] != 0; sb[j] = s; }
    for (int i = 0; i < na; ++i) {
        const uint8_t *ra = a + (int64_t)i * n;
        for (int j = 0; j < nb; ++j) {
            const uint8_t *rb = b + (int64_t)j * n;
            int32_t s = 0;
            for (int64_t p = 0; p < n; ++p) s += (ra[p] != 0) & (rb[p] != 0);
            inter[(int64_t)i * nb + j] = s;
            const float fi = (float)s;
            const float un = ((float)sa[i] + (float)sb[j]) - fi;
            iou[(int64_t)i * nb + j] = fi / un;
        }
    }
    free(sa); free(sb);
}

/* tools/utils/rle_encode_decode.py:35-61 and tools/refinement.py:26-39: int32 (start,len)
 * pairs, start 1-based, each run written with Python slice semantics mask[lo:hi] = 1. */
ORACLE_API void oracle_rle_decode(const int64_t *counts, int64_t ncounts, int64_t length, uint8_t *mask) {
    memset(mask, 0, (size_t)length);
    for (int64_t k = 0; k + 1 < ncounts; k += 2) {
        int32_t lo = (int32_t)counts[k] - 1;
        int32_t hi = (int32_t)((uint32_t)lo + (uint32_t)(int32_t)counts[k + 1]);
        int64_t l = lo, hh = hi;
        if (l < 0) { l += length; if (l < 0) l = 0; } else if (l > length) l = length;
        if (hh < 0) { hh += length; if (hh < 0) hh = 0; } else if (hh > length) hh = length;
        for (int64_t p = l; p < hh; ++p) mask[p] = 1;
    }
}

ORACLE_API int oracle_version(void) { return 1; }

/* tools/projection_2d_to_3d.py:431-436 (and :556-563) -- depth_im = imread(png).astype(float32) / 1000;
 * depth_im = cv2.resize(depth_im, (width_2d, height_2d)).
 * cv2.resize is OpenCV (INTER_LINEAR, float32), a dependency outside the reference tree; its IPP-backed pip
 * builds compute, per output pixel (x, y):
 *   f = (i + 0.5) * (n_src / n_dst) - 0.5 in float64 per axis; i0 = floor(f); w = float32(f - i0);
 *   i0 < 0 -> (0, w = 0);  i0 >= n_src - 1 -> (n_src - 1, w = 0);  i1 = min(i0 + 1, n_src - 1)
 *   h0 = fmaf(S[y0][x1] - S[y0][x0], wx, S[y0][x0]);  h1 = the same on row y1;  out = fmaf(h1 - h0, wy, h0)
 * PINNED against cv2.resize itself (tests/test_oracle_golden.py::test_resize_matches_cv2, 11 size pairs up
 * and down, bit for bit) and against the frozen cv2 output in tests/golden/resize.npz (oracle/make_golden_resize.py).
 * src (hs, ws) f32 -> dst (hd, wd) f32. */
static void resize_axis(int n_dst, int n_src, int *idx, float *wt) {
    const double scale = (double)n_src / (double)n_dst;
    for (int i = 0; i < n_dst; ++i) {
        const double f = ((double)i + 0.5) * scale - 0.5;
        const double fl = floor(f);
        int k = (int)fl;
        float w = (float)(f - fl);
        if (k < 0) { k = 0; w = 0.0f; }
        if (k >= n_src - 1) { k = n_src - 1; w = 0.0f; }
        idx[i] = k;
        wt[i] = w;
    }
}

ORACLE_API void oracle_resize_bilinear(const float *src, int hs, int ws, float *dst, int hd, int wd) {
    int *xi = (int *)malloc(sizeof(int) * (size_t)wd), *yi = (int *)malloc(sizeof(int) * (size_t)hd);
    float *xw = (float *)malloc(sizeof(float) * (size_t)wd), *yw = (float *)malloc(sizeof(float) * (size_t)hd);
    resize_axis(wd, ws, xi, xw);
    resize_axis(hd, hs, yi, yw);
    for (int y = 0; y < hd; ++y) {
        const int y0 = yi[y], y1 = y0 + 1 < hs ? y0 + 1 : hs - 1;
        const float *r0 = src + (size_t)y0 * ws, *r1 = src + (size_t)y1 * ws;
        for (int x = 0; x < wd; ++x) {
            const int x0 = xi[x], x1 = x0 + 1 < ws ? x0 + 1 : ws - 1;
            const float h0 = fmaf(r0[x1] - r0[x0], xw[x], r0[x0]);
            const float h1 = fmaf(r1[x1] - r1[x0], xw[x], r1[x0]);
            dst[(size_t)y * wd + x] = fmaf(h1 - h0, yw[y], h0);
        }
    }
    free(xi); free(yi); free(xw); free(yw);
}
