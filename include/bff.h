/*
 * bff.h -- C ABI of libbff.so: the sm_100a kernels behind BeyondFF's 2D->3D lifting,
 * aggregation/filtering and refinement hot path.
 *
 * The reference (zhenghao2519/Beyond-Fixed-Forms) has no FFI of its own: the path is plain
 * numpy/torch inside tools/projection_2d_to_3d.py and tools/refinement.py. Each entry point
 * below names the reference lines it replaces; INTEGRATION.md shows the ctypes stub a
 * maintainer of the reference would add at those lines.
 *
 * Conventions
 *  - every function returns 0 (BFF_OK) or a negative bff_status; bff_last_error() gives a
 *    thread-local message for the last failure on the calling thread;
 *  - *_dev / unsuffixed functions take DEVICE pointers owned by the caller and enqueue work
 *    on `stream` (a cudaStream_t passed as void*; NULL = legacy default stream); they never
 *    allocate and never synchronise the device;
 *  - bff_*_host functions take HOST pointers, stage through a bff_ctx-owned device
 *    workspace, and return after the results are back in host memory;
 *  - packed masks: row-major uint32 words, bit b of word w = element 32*w + b; a packed
 *    point row has Wd = ceil(N/32) words, a packed H x W image mask has ceil(H*W/32) words
 *    with element index y*W + x (the reference's flattening, rle_encode_decode.py:63-99);
 *  - bits beyond N in the last word of a row are always written as 0 and must be 0 on input.
 *
 * Size limits (checked, BFF_EINVAL beyond them; the host package routes larger jobs to its step-by-step path):
 *  - the row-indexed kernels address rows by blockIdx.y; their entry points loop over blocks of 65 535 rows, so
 *    bff_rows_or_reduce / bff_rows_and_popcount / bff_pack_bits / bff_unpack_bits / bff_rows_slice / bff_rows_unslice /
 *    bff_rows_gather / bff_rows_permute take any int32 row count. Still limited to 65 535: n_masks of bff_mask_gather,
 *    n_frames and the output height of bff_depth_resize;
 *  - bff_stage2_aggregate: n_rows <= 46 340 (n_rows^2 int32 intersections); bff_refine_match: n2 <= 65 535;
 *  - bff_stage2_device: n_rows <= 25 000 (8 bytes of shared memory per row in the component search);
 *    bff_refine_device: cap2 <= 40 960;
 *  - point counts < 2^31 x 32, image sizes H x W < 2^31, counts (votes, intersections) < 2^31; IoUs are exact while
 *    counts stay below 2^24 (float32), as in the reference.
 */
#ifndef BFF_H_
#define BFF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BFF_VERSION 100

#if defined(__GNUC__)
#define BFF_API __attribute__((visibility("default")))
#else
#define BFF_API
#endif

typedef enum bff_status {
    BFF_OK = 0,
    BFF_EINVAL = -1,     /* bad argument (null pointer, negative size, unsupported combination) */
    BFF_ECUDA = -2,      /* a CUDA runtime call or launch failed */
    BFF_ENOMEM = -3      /* workspace allocation failed (host-buffer entry points only) */
} bff_status;

typedef enum bff_depth_kind {
    BFF_DEPTH_F32 = 0,   /* float32 metres, as the reference holds it after /depth_scale and resize */
    BFF_DEPTH_U16 = 1,   /* raw uint16 sensor units; the kernel computes (float)u16 / depth_scale in fp32,
                            bit-identical to projection_2d_to_3d.py:431-435 when no resize is needed */
    BFF_DEPTH_U16_BILINEAR = 2
                         /* raw uint16 at the sensor's NATIVE size (src_height, src_width); the kernel samples it at
                            mask resolution with the arithmetic of cv2.resize(depth, (W, H)) (projection_2d_to_3d.py:436,
                            :563; default configs upsample 640x480 -> 1296x968), so the resized float32 image never
                            exists: see bff_resize_table / bff_depth_resize for the exact formula */
} bff_depth_kind;

typedef enum bff_xyz_kind {
    BFF_XYZ_F32X4 = 0,   /* float4 per point (x,y,z,unused); exact iff the float64 cloud is float32-representable */
    BFF_XYZ_F64X3 = 1    /* three doubles per point */
} bff_xyz_kind;

/* One output coordinate of a bilinear resize: the two source taps are index and min(index + 1, n_src - 1),
 * blended as fmaf(s1 - s0, weight, s0) in float32. */
typedef struct bff_resize_tap {
    int32_t index;
    float weight;
} bff_resize_tap;

/* Description of one scene's lifting job. All pointers are device pointers for
 * bff_lift_scene and host pointers for bff_lift_scene_host. */
typedef struct bff_lift_desc {
    /* cloud: tools/projection_2d_to_3d.py:385-390 */
    const void *xyz;            /* n_points elements of xyz_kind */
    int32_t xyz_kind;           /* bff_xyz_kind */
    int64_t n_points;           /* N */
    /* cameras: :370-376 (intrinsics), :422-425 (np.linalg.inv(cam_pose), computed by the caller) */
    const double *pose_inv;     /* (n_frames, 4, 4) row-major float64 inverse poses */
    const double *intrinsics;   /* (3, 3) row-major float64; ALWAYS a HOST pointer (passed to the kernel by value) */
    int32_t n_frames;           /* F */
    /* depth: :431-436 */
    const void *depth;          /* (n_frames, height, width) of depth_kind */
    int32_t depth_kind;         /* bff_depth_kind */
    float depth_scale;          /* divisor for BFF_DEPTH_U16 (reference: 1000) */
    int32_t height, width;      /* H, W = cfg.height_2d, cfg.width_2d */
    double depth_thresh;        /* reference: 0.08 (:438,:565) */
    /* Optional conservative depth range of NON-ZERO depth values (metres). Points whose camera z
     * lies outside [depth_min - 2*thresh, depth_max + 2*thresh] cannot pass |z - d| < thresh and
     * skip the divides. Set depth_min > depth_max (e.g. 1, 0) to disable. */
    double depth_min, depth_max;
    /* 2D masks: :395-421, packed bits, all frames concatenated in frame order */
    const uint32_t *masks2d;    /* (n_masks, ceil(H*W/32)) */
    const int32_t *mask_offset; /* (n_frames + 1) exclusive prefix of masks per frame; NULL = no masks. A DEVICE array, so its
                                 * contents are the caller's contract and are not checked: non-decreasing, mask_offset[0] = 0,
                                 * mask_offset[n_frames] = n_masks (bff_b200/lift.py:mask_layout builds it from the files) */
    const int32_t *mask_query;  /* (n_masks) query id in [0, n_queries); NULL = all 0 */
    int32_t n_masks;            /* sum of M_f */
    int32_t n_queries;          /* Q >= 1 */
    const uint8_t *frame_viewed;/* (n_frames) 1 = frame counts into viewed_counts (:538-567); NULL = all */
    /* outputs */
    uint32_t *rows;             /* (n_masks, Wd) packed lifted 3D rows (:441-457); may be NULL */
    int32_t *masked_counts;     /* (n_queries, N) votes, one per mask hit (:459-461); may be NULL */
    int32_t *viewed_counts;     /* (N) frames in which the point is visible (:567); may be NULL */
    int64_t *debug_pix;         /* optional (n_frames, N, 2) int64 [x, y] exactly as :37-48 returns */
    uint32_t *debug_vis;        /* optional (n_frames, Wd) packed visibility (:51-70) */
    /* launch shape: frames per CTA chunk (0 = automatic) */
    int32_t frames_per_chunk;
    int32_t reserved;
    /* BFF_DEPTH_U16_BILINEAR only: native depth size and the per-axis sampling tables of bff_resize_table
     * (resize_x: width entries over src_width, resize_y: height entries over src_height), device pointers
     * for bff_lift_scene */
    int32_t src_height, src_width;
    const bff_resize_tap *resize_x, *resize_y;
    /* optional cudaEvent_t handles recorded on `stream` right before the first and right after the last
     * device operation of the call (for callers that time the launch itself); NULL = none */
    void *ev_start, *ev_stop;
    /* Row layout for the point-sharded multi-GPU exchange. n_row_slices <= 1: rows is (n_masks, Wd). n_row_slices = G > 1:
     * the words of every row are cut into G column slices of row_slice_words words (G * row_slice_words >= Wd, the tail of
     * the last slice stays 0) and rows is (G, n_masks, row_slice_words): slice p of ALL rows is one contiguous block, which
     * is what rank p receives (bff_comm_exchange_rows). A warp owns one word of every row, so the layout costs nothing. */
    int32_t n_row_slices;
    int32_t reserved2;
    int64_t row_slice_words;
    /* Fused exchange (n_row_slices = G > 1, optional): peer_rows[p] != NULL for every p < G makes the kernel store slice p
     * of its rows DIRECTLY into rank p's receive buffer (peer memory over NVLink, opened with bff_ipc_open): row m of this
     * launch lands at peer_rows[p] + (peer_row0 + m) * row_slice_words. `rows` is then unused (may be NULL), nothing is
     * zeroed by the call (every owner zeroes its own buffer and the ranks synchronise before the launches), and no
     * separate row exchange follows. */
    uint32_t *peer_rows[8];
    int64_t peer_row0;
} bff_lift_desc;

BFF_API int bff_version(void);
BFF_API const char *bff_last_error(void);

/* Replaces the per-frame loops of tools/projection_2d_to_3d.py:413-461 and :538-567
 * (compute_projected_pts_tensor :37, compute_visibility_mask_tensor :51,
 * compute_visible_masked_pts_tensor :73, masked_counts/viewed_counts updates) with one launch
 * over all frames. Outputs are fully overwritten. */
BFF_API int bff_lift_scene(const bff_lift_desc *desc, void *stream);

/* The bilinear sampling of cv2.resize(src, (n_dst, .)) along one axis (INTER_LINEAR on float32, as OpenCV's
 * IPP-backed builds compute it; checked against cv2.resize in tests/ and at run time by the host package):
 *   f = (i + 0.5) * (n_src / n_dst) - 0.5 in float64; index = floor(f); weight = float32(f - index);
 *   index < 0 -> (0, 0);  index >= n_src - 1 -> (n_src - 1, 0).
 * Host function, `taps` is a HOST array of n_dst entries. */
BFF_API int bff_resize_table(int32_t n_dst, int32_t n_src, bff_resize_tap *taps);

/* Stand-alone form of the depth ingest of projection_2d_to_3d.py:431-436 (and :556-563) on the device:
 * src (n_frames, src_h, src_w) raw uint16 -> dst (n_frames, dst_h, dst_w) float32 metres =
 * cv2.resize((float)src / depth_scale, (dst_w, dst_h)): rows first blend horizontally, then vertically,
 * each blend fmaf(s1 - s0, weight, s0). tap_x (dst_w) / tap_y (dst_h) are DEVICE copies of bff_resize_table.
 * The scene path does not need it (BFF_DEPTH_U16_BILINEAR samples inside the lift kernel); callers that want
 * the resized image, and the run-time check against cv2.resize, do. */
BFF_API int bff_depth_resize(const uint16_t *src, int32_t n_frames, int32_t src_h, int32_t src_w, float depth_scale,
                             const bff_resize_tap *tap_x, const bff_resize_tap *tap_y, int32_t dst_h, int32_t dst_w,
                             float *dst, void *stream);

/* Number of kernels bff_lift_scene launches for this description (for launch accounting). */
BFF_API int bff_lift_scene_launches(const bff_lift_desc *desc);

/* Stand-alone per-frame helpers for callers that already hold pixel indices (the scene path is bff_lift_scene):
 * compute_visibility_mask_tensor (projection_2d_to_3d.py:51-70): pix (n,2) int64 [x,y], z (n) float64 camera z,
 * depth (H,W) float32 -> packed visibility bits (ceil(n/32)); and compute_visible_masked_pts_tensor (:73-92):
 * packed 2D masks (n_masks, ceil(H*W/32)) sampled at the visible points -> packed rows (n_masks, ceil(n/32)). */
BFF_API int bff_visibility(const int64_t *pix, const double *z, int64_t n, const float *depth, int32_t height, int32_t width,
                           double thresh, uint32_t *vis, void *stream);
BFF_API int bff_mask_gather(const int64_t *pix, const uint32_t *vis, int64_t n, const uint32_t *masks2d, int32_t n_masks,
                            int32_t height, int32_t width, uint32_t *rows, void *stream);

/* Replaces calculate_iou (projection_2d_to_3d.py:149-166) and calculate_iou_between_stages
 * (refinement.py:69-90, before its transpose): a (nA, Wd), b (nB, Wd) packed rows ->
 * inter (nA, nB) int32, pop_a (nA), pop_b (nB) int32, and optionally iou (nA, nB) float32 =
 * f32(I) / ((f32(sA) + f32(sB)) - f32(I)), 0/0 = NaN. b == a computes the square matrix.
 * inter / pop_a / pop_b / iou may each be NULL. */
BFF_API int bff_mask_iou(const uint32_t *a, int32_t n_a, const uint32_t *b, int32_t n_b, int64_t words,
                 int32_t *inter, int32_t *pop_a, int32_t *pop_b, float *iou, void *stream);
/* Which intersection kernel bff_mask_iou takes: 0 automatic (<= 32 columns: one-pass kernel; >= 96 rows on both sides:
 * tcgen05 int8; else POPC tiles), 1 / 2 / 3 force one of them (the tests cross-check the three; all are bit-exact).
 * Process-wide; the initial value comes from the environment variable BFF_IOU_PATH (skinny|popc|umma), read once. */
BFF_API int bff_set_iou_path(int32_t path);

/* Replaces the OR-merges of merge_masks (projection_2d_to_3d.py:214-223) and refinement.py:248,269:
 * out[g] = OR over k in [seg[g], seg[g+1]) of rows[idx[k]].  rows (n_rows, words), out (n_groups, words). */
BFF_API int bff_rows_or_reduce(const uint32_t *rows, int64_t words, const int32_t *idx, const int32_t *seg,
                       int32_t n_groups, uint32_t *out, void *stream);

/* Max of masked_counts (per query) and viewed_counts, so the histogram below can be sized:
 * out[0] = max(mc), out[1] = max(vc). vc may be NULL (out[1] = 0). */
BFF_API int bff_counts_max(const int32_t *mc, const int32_t *vc, int64_t n, int32_t *out2, void *stream);

/* 2-D histogram of (masked_count, viewed_count) pairs: hist[(mc * (vc_bins)) + vc] += 1.
 * The host derives the reference's detected_ratio.unique() / masked_counts.unique()
 * (projection_2d_to_3d.py:512-578) from the non-empty bins. vc may be NULL (vc_bins = 1). hist is overwritten. */
BFF_API int bff_vote_hist(const int32_t *mc, const int32_t *vc, int64_t n, int32_t mc_bins, int32_t vc_bins,
                  uint32_t *hist, void *stream);

/* The distinct values the reference's unique() sees (:512-517 masked_counts.unique(), :571-575 detected_ratio.unique()),
 * derived on the device from the non-empty bins of bff_vote_hist: mode 0 value = f32(mc) / (f32(vc) + 1), mode 1 value =
 * f32(mc). `table` is scratch of table_size uint32 (a power of two >= bff_vote_unique_table_size(mc_bins, vc_bins));
 * `values` receives *count distinct floats in unspecified order (capacity mc_bins * vc_bins). */
BFF_API int64_t bff_vote_unique_table_size(int32_t mc_bins, int32_t vc_bins);
BFF_API int bff_vote_unique(const uint32_t *hist, int32_t mc_bins, int32_t vc_bins, int32_t mode, uint32_t *table, int64_t table_size,
                            float *values, int32_t *count, void *stream);

/* bff_vote_unique without caller-provided scratch and without ANY allocation -- for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). No counterpart in the reference.
 * Same inputs, outputs and kernels as bff_vote_unique; the hash-set table is a
 * file-scope device array inside the library: BFF_VOTE_UNIQUE_STATIC_TABLE_WORDS uint32 = 1 MiB on every GPU that loads
 * the library, held until the process exits. Limit: mc_bins * vc_bins <= BFF_VOTE_UNIQUE_STATIC_MAX_BINS, beyond it the call
 * returns BFF_EINVAL (bff_last_error names the limit). The call uses the CURRENT device; the first call per device
 * resolves the array's device address (cudaGetSymbolAddress) -- make that call outside a capture. Every call on a device
 * shares the one table, which keeps the previous call's contents: each call re-initialises the slots it uses on its own
 * stream, and two calls that overlap on the GPU (different streams, no dependency) RACE -- serialise them. */
#define BFF_VOTE_UNIQUE_STATIC_MAX_BINS 131072
#define BFF_VOTE_UNIQUE_STATIC_TABLE_WORDS 262144
BFF_API int bff_vote_unique_static(const uint32_t *hist, int32_t mc_bins, int32_t vc_bins, int32_t mode, float *values, int32_t *count,
                                   void *stream);

/* Per-point keep predicate -> packed bits keep (Wd):
 *  mode 0 (detected ratio, :571-583): keep = mc > 0 && !(f32(mc) / (f32(vc) + 1) < thr)
 *  mode 1 (occurrence, :512-522,583):  keep = mc > 0 && !(f32(mc) < thr)
 *  mode 2 (no filter):                 keep = mc > 0 */
BFF_API int bff_point_keep(const int32_t *mc, const int32_t *vc, int64_t n, int32_t mode, float thr,
                   uint32_t *keep, void *stream);

/* Evaluator side (evaluation/eval/scannetv2_inst_eval.py:348-359): per-point encoded ground-truth labels (n) int64
 * -> rows (n_ids, ceil(n/32)) packed, bit p of row g = (labels[p] == ids[g]). ids is a DEVICE array. Together with
 * bff_pack_bits on the prediction masks and one bff_mask_iou call this yields every pred x gt intersection
 * (`np.count_nonzero(np.logical_and(gts == inst_id, pred_mask))`) at once. */
BFF_API int bff_labels_to_rows(const int64_t *labels, int64_t n, const int64_t *ids, int32_t n_ids, uint32_t *rows, void *stream);

/* Re-index packed rows by point: rows_out bit p = rows_in bit src_index[p] for every row (src_index: device, (n_points) int32,
 * a permutation or any gather map). The lifting host code uses it to return rows computed on a spatially sorted copy of the
 * cloud to the cloud's own order. rows_out must not alias rows_in. */
BFF_API int bff_rows_permute(const uint32_t *rows_in, int32_t n_rows, int64_t n_points, const int32_t *src_index,
                             uint32_t *rows_out, void *stream);

/* solve_overlapping's sequential pass (projection_2d_to_3d.py:295-299) on packed rows, in place.
 * prog (n_steps, 2) int32: step k clears from row prog[k][0] the bits of row prog[k][1]
 * (rows[dst] &= ~rows[src]) in list order; every word column runs the program independently. */
BFF_API int bff_overlap_resolve(uint32_t *rows, int32_t n_rows, int64_t words, const int32_t *prog,
                        int32_t n_steps, void *stream);

/* The in-place stage-1 merge of refinement.py:230-249 on packed rows: step k does
 * rows[prog[k][0]] |= rows[prog[k][1]] in list order (later steps see earlier results). */
BFF_API int bff_rows_or_into(uint32_t *rows, int32_t n_rows, int64_t words, const int32_t *prog,
                             int32_t n_steps, void *stream);

/* rows[r] &= keep (if keep != NULL, in place) and pop[r] = popcount(rows[r]) afterwards
 * (projection_2d_to_3d.py:592-596). */
BFF_API int bff_rows_and_popcount(uint32_t *rows, int32_t n_rows, int64_t words, const uint32_t *keep,
                          int32_t *pop, void *stream);

/* bool/uint8 (n_rows, n) <-> packed (n_rows, ceil(n/32)). Non-zero bytes are set bits. */
BFF_API int bff_pack_bits(const uint8_t *dense, int32_t n_rows, int64_t n, uint32_t *packed, void *stream);
BFF_API int bff_unpack_bits(const uint32_t *packed, int32_t n_rows, int64_t n, uint8_t *dense, void *stream);

/* RLE -> packed bits without a dense intermediate (rle_decode_batch, rle_encode_decode.py:35-61;
 * rle_decode, refinement.py:26-39). runs (n_runs, 3) int32 = [row, lo, hi) already clipped to
 * [0, length] by the host with Python-slice semantics. packed (n_rows, words) is overwritten. */
BFF_API int bff_rle_to_bits(const int32_t *runs, int64_t n_runs, int32_t n_rows, int64_t words,
                    uint32_t *packed, void *stream);

/* The same decode straight from the wire format: `counts` holds the masks' (1-based start, run length) int64 pairs back to
 * back (an unpaired trailing number of a mask dropped), pair_offset (n_rows + 1) the first pair of every row; both on the
 * DEVICE. The int32 casts, the int32 end = start + length and Python's slice clipping of the reference's decoder
 * (rle_encode_decode.py:53-57, refinement.py:33-37) happen in the kernel, so the host only concatenates arrays. */
BFF_API int bff_rle_pairs_to_bits(const int64_t *counts, const int64_t *pair_offset, int32_t n_rows, int64_t n_pairs, int64_t length,
                                  uint32_t *packed, void *stream);

/* ---- host-buffer entry points (what a ctypes/numpy binding in the reference would call) ---- */
typedef struct bff_ctx bff_ctx;
BFF_API int bff_ctx_create(int device, bff_ctx **out);
BFF_API void bff_ctx_destroy(bff_ctx *ctx);
/* Same contract as bff_lift_scene with every pointer in HOST memory (pinned or pageable).
 * Copies inputs to the device, runs the kernels, copies rows / counts back, and returns
 * after the outputs are valid. h2d_bytes / d2h_bytes (may be NULL) receive the bytes moved. */
BFF_API int bff_lift_scene_host(bff_ctx *ctx, const bff_lift_desc *host_desc, int64_t *h2d_bytes, int64_t *d2h_bytes);
/* Host-buffer form of bff_mask_iou. */
BFF_API int bff_mask_iou_host(bff_ctx *ctx, const uint32_t *a, int32_t n_a, const uint32_t *b, int32_t n_b,
                      int64_t words, int32_t *inter, int32_t *pop_a, int32_t *pop_b, float *iou,
                      int64_t *h2d_bytes, int64_t *d2h_bytes);

/* ---- fused host+device pipelines: the small decisions between kernels (component search, thresholds,
 * overlap schedule, argmax matching) run in C++ next to the launches instead of one Python/torch call each.
 * Device pointers are the caller's (torch) tensors; work is enqueued on `stream`, which the function
 * synchronises whenever a decision needs device results; scratch comes from the bff_ctx arenas. ---- */

/* aggregate + point filter + solve_overlapping of ONE query's lifted rows
 * (tools/projection_2d_to_3d.py:100-146,190-301,512-596). */
typedef struct bff_stage2_desc {
    const uint32_t *rows;       /* device (n_rows, Wd) lifted rows in file order (:481) */
    int32_t n_rows;
    int64_t n_points;
    const float *conf;          /* host (n_rows) float32 confidences (:484-486) */
    const int32_t *label_id;    /* host (n_rows); equal ids <=> equal label strings (:169-187) */
    const int32_t *mc;          /* device (N) masked_counts of this query */
    const int32_t *vc;          /* device (N) viewed_counts; required for filter_mode 0 */
    float iou_thres;            /* cfg.iou_thres as float32 (torch compares `f32 tensor > python float` in float32) */
    int32_t min_aggregated;     /* cfg.min_aggragated_masks, must be >= 1 here */
    int32_t filter_mode;        /* 0 detected ratio (:524-578), 1 occurrence (:512-522), 2 none */
    double filter_fraction;     /* cfg.detected_ratio_threshold / cfg.occurance_threshold */
    /* outputs */
    uint32_t *agg_rows;         /* device (cap_groups, Wd): merged rows after overlap resolution AND the point filter */
    int32_t cap_groups;
    int32_t *n_groups;          /* host: aggregated masks that passed min_aggregated (0 = 'nothing after aggregation') */
    int32_t *group_first;       /* host (cap_groups): smallest member = source of the group's label (:226) */
    int32_t *group_size;        /* host (cap_groups) */
    float *group_conf;          /* host (cap_groups): sequential float32 mean of the members' confidences (:220-224) */
    int32_t *before;            /* host (cap_groups): row sums before solve_overlapping (:592) */
    int32_t *after;             /* host (cap_groups): row sums after overlap resolution and `&= masked_points` (:596) */
    int32_t *members;           /* host (n_rows), optional: members of the kept groups, concatenated (for the log line) */
    float *filter_value;        /* host, optional: the threshold that was picked (:576) */
    /* optional, both or neither: the all-pairs intersections (n_rows, n_rows) and row popcounts (n_rows) of `rows`,
     * already on the device -- the frame-sharded multi-GPU path computes one row block per rank and all-gathers them
     * (bff_b200/dist.py), so the quadratic step is not repeated on every rank. NULL: computed here. */
    const int32_t *inter_pre;
    const int32_t *pop_pre;
} bff_stage2_desc;
BFF_API int bff_stage2_aggregate(bff_ctx *ctx, const bff_stage2_desc *desc, void *stream);
/* The same for `n_jobs` independent queries of one scene (descs[0..n_jobs)): every phase is enqueued for all jobs before
 * the stream is synchronised, so Q concurrent queries share three synchronisations instead of paying three each.
 * Results are identical to n_jobs separate calls. */
BFF_API int bff_stage2_aggregate_batch(bff_ctx *ctx, const bff_stage2_desc *descs, int32_t n_jobs, void *stream);

/* pass 1 of tools/refinement.py for one scene (:208-290): stage-1 / stage-2 matching, the in-place stage-1
 * OR-merges, merging of stage-2 masks that share a stage-1 match, and the final IoU + argmax. */
typedef struct bff_refine_desc {
    uint32_t *s1_rows;          /* device (n1, Wd) stage-1 masks, MODIFIED in place like the reference (:248) */
    int32_t n1;
    const uint32_t *s2_rows;    /* device (n2, Wd) stage-2 masks */
    int32_t n2;
    int64_t n_points;
    const float *s2_conf;       /* host (n2) */
    float stage1_iou_thres;     /* cfg.stage1_iou_thres as float32 */
    /* outputs */
    uint32_t *s2_out;           /* device (n2, Wd): stage-2 rows after merging (:258-281), n2_out rows used */
    int32_t *n2_out;            /* host */
    float *s2_conf_out;         /* host (n2) */
    int32_t *match;             /* host (n2): argmax stage-1 index per remaining stage-2 mask (:288) */
    float *match_iou;           /* host (n2): iou[k, match[k]] (:308) */
    int32_t *best_match_log;    /* host (n2), optional: best_match_after_iou_check before merging (:252) */
} bff_refine_desc;
BFF_API int bff_refine_match(bff_ctx *ctx, const bff_refine_desc *desc, void *stream);
/* The same for `n_jobs` independent jobs (queries of one scene; every job needs its own copy of the stage-1 rows, which
 * pass 1 modifies): phases are enqueued for all jobs before the stream is synchronised. Results equal n_jobs separate calls. */
BFF_API int bff_refine_match_batch(bff_ctx *ctx, const bff_refine_desc *descs, int32_t n_jobs, void *stream);

/* ---- device-resident pipelines: the same two stages with EVERY decision taken by small kernels on the device, so a
 * call only enqueues work on `stream`: no stream synchronisation, no host round trip, CUDA-graph capturable. Counts that
 * the host versions return (groups, selected rows, matches ...) live in a device RESULT BLOCK that later kernels read
 * through pointers; the caller copies the block to the host when (and if) it wants the numbers. Buffers whose row count
 * is only known on the device are sized by their upper bound (`cap_*`). The decisions are bit-identical to
 * bff_stage2_aggregate / bff_refine_match (tests/test_devpipe_gpu.py compares both and the oracle port).
 *
 * Point-sharded multi-GPU form (comm != NULL): every rank passes the SAME rows / masks restricted to its own contiguous
 * slice of the points (a multiple of 32 points; `n_points` = slice length), the additive integer results (intersections,
 * popcounts) are summed over the ranks with ncclAllReduce on `stream`, and every rank takes the same decisions. ---- */
typedef struct bff_comm bff_comm;
/* The library's own NCCL communicator (libnccl is resolved with dlopen when the first of these is called, so single-GPU
 * users never need it): rank 0 obtains 128 id bytes, the caller distributes them (torch.distributed broadcast), every
 * rank creates its communicator on its device. All-reduce op: 0 = sum, 1 = max. */
BFF_API int bff_comm_unique_id(void *id128);
BFF_API int bff_comm_create(int device, int rank, int world, const void *id128, bff_comm **out);
BFF_API void bff_comm_destroy(bff_comm *comm);
BFF_API int bff_comm_rank(const bff_comm *comm);
BFF_API int bff_comm_world(const bff_comm *comm);
BFF_API int bff_comm_allreduce_i32(bff_comm *comm, int32_t *buf, int64_t count, int32_t op, void *stream);
BFF_API int bff_comm_allgather_bytes(bff_comm *comm, const void *send, void *recv, int64_t bytes_per_rank, void *stream);
/* The one bulk exchange of the frame-sharded -> point-sharded hand-over: `send` is this rank's lifted rows in the sliced
 * layout of bff_lift_desc (world, n_local_rows, slice_words); slice p goes to rank p, which stores the blocks of all ranks
 * in rank order: `recv` is (sum of rows_per_rank, slice_words) = every lifted row of the scene in file order, restricted
 * to this rank's points. rows_per_rank: HOST array (world). One grouped ncclSend/ncclRecv; the own block is a local copy. */
BFF_API int bff_comm_exchange_rows(bff_comm *comm, const uint32_t *send, int32_t n_local_rows, const int32_t *rows_per_rank,
                                   int64_t slice_words, uint32_t *recv, void *stream);
/* Device buffers other processes of this node can map (CUDA IPC; the frame-sharded lifting stores row slices straight into
 * its peers' receive buffers). bff_ipc_alloc: cudaMalloc + the 64-byte handle to hand to the peers; bff_ipc_open: map a
 * peer's buffer into this process (peer access is enabled on demand); close / free undo them. */
BFF_API int bff_ipc_alloc(int64_t bytes, void **dev_ptr, void *handle64);
BFF_API int bff_ipc_open(const void *handle64, void **dev_ptr);
BFF_API int bff_ipc_close(void *dev_ptr);
BFF_API int bff_ipc_free(void *dev_ptr);

/* Column slices of packed rows: out (n_rows, slice_words) = words [word0, word0 + slice_words) of rows (n_rows, words),
 * zero beyond `words`; and the inverse for results: in (world, n_rows, slice_words) -> out (n_rows, words). */
BFF_API int bff_rows_slice(const uint32_t *rows, int32_t n_rows, int64_t words, int64_t word0, int64_t slice_words, uint32_t *out, void *stream);
BFF_API int bff_rows_unslice(const uint32_t *slices, int32_t world, int32_t n_rows, int64_t slice_words, int64_t words, uint32_t *out, void *stream);

/* Result block of bff_stage2_device: int32 words, R = n_rows, C = cap_groups. */
enum {
    BFF_S2R_N_GROUPS = 0,      /* aggregated masks that passed min_aggregated (:203) */
    BFF_S2R_N_SELECTED = 1,    /* rows that passed the row filter (:601-623) = rows of out_rows */
    BFF_S2R_N_OVERLAP = 2,     /* steps of solve_overlapping's schedule (:285-299) */
    BFF_S2R_N_UNIQUE = 3,      /* distinct vote values the threshold was picked from (:514,:573) */
    BFF_S2R_ERROR = 4,         /* bit 0: filter fraction selects an index outside the unique list (the reference raises
                                  IndexError); bit 1: a scratch bound (mc_cap / vc_cap / cap_groups) was exceeded */
    BFF_S2R_FILTER_VALUE = 5,  /* float32 bits of the picked threshold (:516,:576) */
    BFF_S2R_MC_MAX = 6,
    BFF_S2R_VC_MAX = 7,
    BFF_S2R_HEADER = 16
    /* then, in this order: group_first (C), group_size (C), group_conf (C, float32 bits), before (C), after (C),
       selected (C: group index of every selected row), selected_conf (C, float32 bits), seg (C + 1), members (R) */
};
static inline int64_t bff_stage2_dev_result_words(int32_t n_rows, int32_t cap_groups) {
    return BFF_S2R_HEADER + 7 * (int64_t)cap_groups + (cap_groups + 1) + n_rows;
}

typedef struct bff_stage2_dev_desc {
    const uint32_t *rows;       /* device (n_rows, Wd) lifted rows in file order (:481) */
    int32_t n_rows;
    int64_t n_points;           /* points of this rank's slice (single GPU: all of them); Wd = ceil(n_points / 32) */
    const float *conf;          /* DEVICE (n_rows) float32 confidences (:484-486) */
    const int32_t *label_id;    /* DEVICE (n_rows), equal ids <=> equal label strings (:169-187); NULL = one label for all */
    const int32_t *mc;          /* device (n_vote_points) masked_counts of this query, ALL points */
    const int32_t *vc;          /* device (n_vote_points) viewed_counts; required for filter_mode 0 */
    int64_t n_vote_points;      /* points the vote threshold is taken over (0 = n_points, the single-GPU case) */
    int64_t vote_point0;        /* first point of this rank's slice inside mc / vc (0 on a single GPU); mc / vc hold at
                                   least max(n_vote_points, vote_point0 + n_points) entries (padding points vote 0) */
    float iou_thres;            /* cfg.iou_thres as float32 */
    int32_t min_aggregated;     /* cfg.min_aggragated_masks, >= 1 */
    int32_t filter_mode;        /* 0 detected ratio, 1 occurrence, 2 none */
    double filter_fraction;
    /* row filter (:601-623) with torch's promotion rules: `after > small` is an int64 compare for a Python int and a
       float32 compare for a Python float; `after > ratio * before` multiplies in float32 for a Python float */
    int32_t small_is_float; int32_t ratio_is_float;
    int64_t small_i; int64_t ratio_i;
    float small_f; float ratio_f;
    int32_t mc_cap, vc_cap;     /* upper bounds of the counts (masks of the query, viewed frames): size the histogram scratch */
    /* outputs (device) */
    uint32_t *out_rows;         /* (cap_groups, Wd): the selected rows after overlap resolution and the point filter */
    int32_t cap_groups;         /* >= n_rows / min_aggregated */
    int32_t *result;            /* bff_stage2_dev_result_words(n_rows, cap_groups) int32 */
    bff_comm *comm;             /* NULL = single GPU */
} bff_stage2_dev_desc;
BFF_API int bff_stage2_device(bff_ctx *ctx, const bff_stage2_dev_desc *desc, void *stream);

/* Result block of bff_refine_device: int32 words, n1 stage-1 masks, C = cap2. */
enum {
    BFF_RFR_N2_IN = 0,         /* stage-2 masks that came in */
    BFF_RFR_N2_OUT = 1,        /* after merging the ones that share a stage-1 match (:258-281) */
    BFF_RFR_N_OR = 2,          /* in-place stage-1 ORs of the merge loop (:248) */
    BFF_RFR_N_OTHER = 3,       /* unmatched stage-1 masks of the query's class (:293) */
    BFF_RFR_N_FINAL = 4,       /* rows of final_rows (pass 2, :330-412) */
    BFF_RFR_ERROR = 5,         /* bit 0: similarity percentile selects an index outside the unique list (IndexError) */
    BFF_RFR_SIM_THRES = 6,     /* float64 bits (2 words) of the similarity threshold that was applied */
    BFF_RFR_HEADER = 16
    /* then, in this order: sims (2 C, float64 bits: :297-302; the block must be 8-byte aligned), first_match (C: argmax of
       :211), best (C: best_match_after_iou_check :252), match (C: :288), match_iou (C, float32 bits: :308), conf_out (C,
       float32 bits), idx2 (C), seg2 (C + 1), other (n1), final_src (n1 + C: < n1 = stage-1 row, else n1 + row of s2_out),
       final_conf (n1 + C, float32 bits) */
};
static inline int64_t bff_refine_dev_result_words(int32_t n1, int32_t cap2) {
    return BFF_RFR_HEADER + 8 * (int64_t)cap2 + (cap2 + 1) + n1 + 2 * ((int64_t)n1 + cap2);
}

typedef struct bff_refine_dev_desc {
    uint32_t *s1_rows;          /* device (n1, Wd) stage-1 masks, MODIFIED in place like the reference (:248) */
    int32_t n1;
    const uint32_t *s2_rows;    /* device (cap2, Wd) stage-2 masks, the first n2 rows are used */
    const int32_t *n2_dev;      /* device: number of stage-2 masks (e.g. &result[BFF_S2R_N_SELECTED] of bff_stage2_device);
                                   NULL = n2_host */
    int32_t n2_host;
    int32_t cap2;               /* upper bound of n2 */
    int64_t n_points;           /* of this rank's slice */
    const float *s2_conf;       /* DEVICE (cap2) */
    float stage1_iou_thres;     /* cfg.stage1_iou_thres as float32 */
    const uint8_t *is_query;    /* device (n1): stage-1 label == query with underscores (:293) */
    const double *sim;          /* device (n1): text similarity of the query and stage-1 mask i's label (:297-302) */
    int32_t select_mode;        /* 0: pass 1 only (the caller applies a global threshold later with bff_refine_select_device);
                                   1: threshold = sorted(set(sims of THIS scene))[int(len * percentile)] (:317-324 for one scene),
                                      0 when nothing matched; 2: threshold = sim_thres */
    double sim_percentile;      /* cfg.refinment_sim_percentile */
    double sim_thres;
    float refine_iou_thres;     /* cfg.refiment_iou_thres as float32 */
    /* outputs (device) */
    uint32_t *s2_out;           /* (cap2, Wd) stage-2 rows after merging */
    uint32_t *final_rows;       /* (n1 + cap2, Wd): rows of the saved masks, N_FINAL used (select_mode != 0); may be NULL with mode 0 */
    int32_t *result;            /* bff_refine_dev_result_words(n1, cap2) int32 */
    bff_comm *comm;
} bff_refine_dev_desc;
BFF_API int bff_refine_device(bff_ctx *ctx, const bff_refine_dev_desc *desc, void *stream);
/* Pass 2 alone (refinement.py:330-412) on the state bff_refine_device(select_mode 0) left in `result`, with the global
 * threshold of :317-324 that the caller computed over all scenes of the class. */
BFF_API int bff_refine_select_device(bff_ctx *ctx, const bff_refine_dev_desc *desc, double sim_thres, void *stream);

/* out[k] = rows[idx[k]], k < n_idx, with the index vector on the DEVICE: no staging copy, no synchronisation, any
 * number of rows (bff_rows_gather takes a host vector and waits for its staging copy). */
BFF_API int bff_rows_gather_dev(const uint32_t *rows, int64_t words, const int32_t *idx_dev, int32_t n_idx, uint32_t *out, void *stream);

/* bff_unpack_bits for a row count that lives on the device: rows [0, min(*n_rows_dev, cap_rows)) are expanded,
 * `dense` has room for cap_rows rows. */
BFF_API int bff_unpack_bits_dc(const uint32_t *packed, const int32_t *n_rows_dev, int32_t cap_rows, int64_t n, uint8_t *dense, void *stream);

/* Host -> device copy done by a small kernel that reads PINNED host memory directly (zero-copy loads over PCIe) instead
 * of a copy engine; the last bytes % 16 are moved by the same kernel, so the call never touches a copy engine. The fused
 * pipelines stage their small host vectors through it: while a scene-sized cudaMemcpyAsync of the NEXT scene occupies the
 * H2D engine (3.4 ms per 184 MB), a few-byte cudaMemcpyAsync of the running pipeline would wait for all of it. It also
 * serves as a bulk upload (50 GB/s) where the engines must stay free. `src_host` must be page-locked (cudaHostAlloc /
 * cudaHostRegister / torch pin_memory); dst and src 16-byte aligned, any size. */
BFF_API int bff_upload(void *dst_dev, const void *src_host, int64_t bytes, void *stream);

/* out[k] = rows[idx[k]] (device gather of packed rows; idx on the HOST). */
BFF_API int bff_rows_gather(bff_ctx *ctx, const uint32_t *rows, int64_t words, const int32_t *idx_host, int32_t n_idx,
                            uint32_t *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* BFF_H_ */
