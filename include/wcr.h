/* wcr.h -- C ABI of libwcr_b200.so: the B200 (sm_100a) hot path of Weak Cube R-CNN.
 *
 * The reference has no FFI/plugin interface for this path: it is
 * plain Python/torch. This header is therefore the boundary a maintainer would
 * bind with ctypes from the reference's own modules (INTEGRATION.md shows the
 * stub). Every entry point cites the reference code it replaces, paths relative
 * to the reference root.
 *
 * Conventions
 *  - plain C: pointers + sizes, no torch / C++ types;
 *  - every pointer is a DEVICE pointer of the current CUDA device unless the
 *    parameter name starts with h_ ; float4-typed buffers must be 16-byte aligned;
 *  - the caller owns every buffer, including workspaces; the library never
 *    allocates, frees or synchronises; all work is enqueued on `stream`
 *    (a cudaStream_t passed as void*);
 *  - return 0 on success, a negative WCR_E* code otherwise; wcr_last_error()
 *    gives a thread-local message. No exceptions cross the boundary;
 *  - no CPU fallback: on a device that is not sm_100 every call returns WCR_EARCH.
 */
#ifndef WCR_H_
#define WCR_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WCR_ABI_VERSION 2   /* 2: WcrLossIO.depth_images */

#define WCR_OK        0
#define WCR_EINVAL   -1   /* bad argument */
#define WCR_ELAUNCH  -2   /* CUDA launch / runtime error */
#define WCR_EARCH    -3   /* device is not sm_100 (B200) */

/* One image of a shard: un-padded depth (+ optional ground mask) and the
 * intrinsics THIS image is processed with. 48 bytes, lives in device memory. */
typedef struct WcrImage {
    const float   *depth;        /* H x W fp32 metres, row pitch depth_pitch elements */
    const uint8_t *ground;       /* H x W, >0 == ground; NULL == missing ((1,1) dummy) */
    int32_t H, W;
    int32_t depth_pitch, ground_pitch;
    float fx, fy, cx, cy;        /* K[0,0], K[1,1], K[0,2], K[1,2] */
} WcrImage;

int         wcr_abi_version(void);
const char *wcr_last_error(void);
/* 0 if the current device is sm_100, else WCR_EARCH. */
int         wcr_device_check(void);
/* number of kernel launches enqueued by this library (all threads of the process)
 * since the last wcr_reset_launch_count() (bench.py's gpu_launches). */
int64_t     wcr_launch_count(void);
void        wcr_reset_launch_count(void);

/* FP32-pipe probe (16 independent FFMA chains per thread, 8 CTAs x 256 threads
 * per SM). Returns the number of FLOPs enqueued (> 0) or a negative error; the
 * caller times it with CUDA events: the FP32 roofline denominator of this device
 * under this run's clocks. out must hold sm_count*8*256 floats. Not counted by
 * wcr_launch_count(). */
int64_t     wcr_probe_ffma(float *out, int64_t out_elems, int iters, void *stream);

/* ---------------------------------------------------------------- step 1 -- */

/* Strided unprojection in the "ground frame" + ordered compaction of the
 * ground-mask pixels. Replaces cubercnn/modeling/roi_heads/roi_heads.py:474-502
 * (normal_vector_from_maps, per-image body up to `points = stack(...)`).
 *   out_pts   [n_images][cap] float4 (x, y, z, 0), row-major scan order
 *   out_count [n_images]      number of points kept per image
 * cap must be >= ceil(H/use_nth)*ceil(W/use_nth) for every image. */
int wcr_unproject_ground(const WcrImage *images, int n_images, int use_nth,
                         float *out_pts /* float4 */, int cap, int *out_count,
                         void *stream);

/* Dense unprojection through K's principal point, the frame cubes
 * live in: X = z*(u-cx)/fx, Y = z*(v-cy)/fy as lifted at roi_heads.py:927-928.
 * Invalid depth (<=0 / non-finite) -> NaN point.
 *   max_H, max_W = largest image of the shard (sizes the grid);
 *   out_xyz [n_images][plane_elems] float4 with Hs=ceil(H/stride) rows of
 *   Ws=ceil(W/stride); image i starts at i*plane_elems float4s, row pitch Ws. */
int wcr_unproject_cloud(const WcrImage *images, int n_images, int stride,
                        int max_H, int max_W, float *out_xyz /* float4 */,
                        int64_t plane_elems, void *stream);

/* RANSAC plane fit + up-vector, all images of the shard in one call. Replaces
 * ProposalNetwork/utils/plane.py:79-134 (Plane_torch.fit_parallel) and the
 * fix-ups at roi_heads.py:508-525.
 *   pts, count      as written by wcr_unproject_ground
 *   sample_idx      [n_images][iters][3] int32 triples (replay of the
 *                   reference's random.sample draws) or NULL => in-kernel
 *                   Philox4x32-10 keyed by (seed, image, iteration)
 *   out_eq          [n_images][4]  NEGATED best plane (plane.py:134)
 *   out_normal      [n_images][3]  after the wall/side/flip fix-ups
 *   out_best        [n_images]     arg-max hypothesis (first max)
 *   out_counts      [n_images][iters] inlier counts (workspace AND output)
 *   out_inlier_mask [n_images][cap] u8 or NULL */
int wcr_ransac_ground(const float *pts /* float4 */, const int *count, int cap,
                      int n_images, const int32_t *sample_idx, int iters,
                      float thresh, uint64_t seed,
                      float *out_eq, float *out_normal, int *out_best,
                      int *out_counts, uint8_t *out_inlier_mask,
                      float *ws_eqs /* float4 [n_images][iters] workspace */,
                      void *stream);

/* Allocation-free variant of wcr_ransac_ground. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). Same arguments and results as
 * wcr_ransac_ground minus ws_eqs: the hypothesis workspace is a file-scope `static __device__`
 * array of WCR_RANSAC_STATIC_MAX_HYP float4 (1 MiB) that every GPU loading the library holds
 * until exit. The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   n_images * iters > WCR_RANSAC_STATIC_MAX_HYP -> WCR_EINVAL with a message (split the shard
 *   or use wcr_ransac_ground).
 * Every call on a device shares the one array and it keeps the last call's contents (nothing in
 * it needs to start at zero: every hypothesis slot is rewritten before it is read; out_counts is
 * cleared by the call): two calls that OVERLAP on the same GPU (different streams, or threads
 * without ordering) race on it -- serialise them on one stream or with events. */
#define WCR_RANSAC_STATIC_MAX_HYP 65536
int wcr_ransac_ground_static(const float *pts /* float4 */, const int *count, int cap,
                             int n_images, const int32_t *sample_idx, int iters,
                             float thresh, uint64_t seed,
                             float *out_eq, float *out_normal, int *out_best,
                             int *out_counts, uint8_t *out_inlier_mask, void *stream);

/* ------------------------------------------------------------ steps 2+3 -- */

/* One GT 2D box of the shard. 32 bytes, device memory. */
typedef struct WcrBox {
    float   x1, y1, x2, y2;   /* GT 2D box in pixels of its image */
    int32_t image;            /* index into the WcrImage table */
    int32_t cls;              /* category id -> priors[cls] */
    float   clamp0, clamp1;   /* the `clamp` tuple handed to Cubes.get_bube_corners
                                 (spaces.py:224): x is clamped with clamp0, y with
                                 clamp1. The ROI head passes [H, W] (roi_heads.py:
                                 801-806, 950) -- the caller decides, as in the
                                 reference. */
} WcrBox;

/* Free parameters of the proposal composite (project spec, the
 * reference ships only the building blocks). 48 bytes, passed BY VALUE fields
 * through a host pointer (h_params). */
typedef struct WcrSampleParams {
    float sigma_xy;       /* std of the 2D centre as a fraction of box w / h      */
    float sigma_z;        /* std of z as a fraction of the pseudo depth z0        */
    float z_lo, z_hi;     /* z truncated to [z0*z_lo, z0*z_hi]                    */
    float dim_lo;         /* dims truncated to [dim_lo, mean + dim_hi_sigmas*std] */
    float dim_hi_sigmas;
    float yaw_lo, yaw_hi; /* yaw ~ U[yaw_lo, yaw_hi)                              */
    int32_t max_tries;    /* bounded rejection loop of the in-kernel sampler      */
    int32_t reserved0, reserved1, reserved2;
} WcrSampleParams;

/* Score weights: score = s_iou^e_iou * s_pts^e_pts * s_seg^e_seg with small
 * non-negative integer exponents (0 == term skipped), SURVEY A3.12. */
typedef struct WcrScoreParams {
    int32_t e_iou, e_pts, e_seg;
    int32_t point_stride;   /* depth ROI sub-sampling (1 == every pixel)          */
    int32_t seg_mode;       /* 0 = mask_iou (utils.py:225), 1 = mod_mask_iou (:236) */
    float   map_beta;       /* > 0: every term goes through custom_mapping(x, map_beta)
                               (utils.py:207-223, the S-curve 1 / (1 + (x / (1 - x)) ** -beta),
                               evaluated in double) before its exponent; 0 = off (default) */
    int32_t seg_eval;       /* mask term: 0 = the library decides, 1 = fill every hull, 2 = on demand
                               (only where on demand is valid, see wcr_score_topk); results are
                               identical, only the time differs */
    int32_t pts_eval;       /* point term: 0 = the library decides, 1 = test every proposal against the
                               depth points, 2 = on demand: proposals whose score BOUND (IoU term x
                               share of the ROI's points their cull box can hold) is below the k-th
                               best score already known are never tested (only where valid, see
                               wcr_score_topk); results are identical, only the time differs.
                               (Was `reserved2`, which had to be 0.) */
} WcrScoreParams;

#define WCR_CUBE_FLOATS   15   /* [cx cy cz w h l R(9 row-major)], spaces.py:130-150 */
#define WCR_RECORD_FLOATS 20   /* idx(i32) score cube[15] s_iou s_pts s_seg           */
#define WCR_SAMPLE_ATTRS   7   /* u v z w h l yaw                                     */
#define WCR_MAX_TOPK      32

/* One block of the in-kernel generator (Philox4x32-10, counter (c0..c3), key =
 * the 64-bit seed, low word first): lets a test pin the device generator to the
 * published known-answer vectors.  out4 [4] uint32, device. */
int wcr_philox_block(uint64_t key, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                     uint32_t *out4, void *stream);

/* Pseudo depth per box: depth at the box centre with the reference's 10-px clamp
 * (roi_heads.py:645-652, pseudo_gt_z_point_loss applied to the GT box centre).
 *   out_z0 [n_boxes] */
int wcr_box_center_depth(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                         float *out_z0, void *stream);

/* Per-box cube sampling (north_star step 2; composite A2.7 built from
 * ProposalNetwork/utils/utils.py:41-59 sample_normal_in_range, :111-145
 * vec_perp_t / orthobasis_from_normal_t / rotate_vector_t and the lifting of
 * roi_heads.py:927-928).
 *   priors   [n_classes][2][3]  (mean w,h,l ; std w,h,l), math_util.py:419
 *   normals  [n_images][3]      up vectors from wcr_ransac_ground
 *   z0       [n_boxes]          from wcr_box_center_depth
 *   replay   [n_boxes][7][N] final samples (u v z w h l yaw) drawn by the
 *            reference's own sampler (parity mode), or NULL => in-kernel
 *            Philox4x32-10 keyed by (seed; proposal, box, attribute, try)
 *   out_cubes [n_boxes][N][15] */
int wcr_sample_cubes(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                     const float *priors, const float *normals, const float *z0,
                     const WcrSampleParams *h_params, int N, const float *replay,
                     uint64_t seed, float *out_cubes, void *stream);

/* Fused scoring + per-box top-k (north_star step 3; composite A3.12). Per
 * (box, proposal): 8 corners (math_util.py:198-217), projection + clamp
 * (spaces.py:224-245), projected box (conversions.py:25-48), 2D IoU vs the GT
 * box (utils.py:181-187), point-in-oriented-cube count over the depth pixels of
 * the GT box ROI unprojected through K (A3.11 / A1.4), segmentation-mask
 * coverage (utils.py:225-245, 490-493), combined score, top-k.
 *   R need not be orthonormal: the projected-box / depth-range cull (exact only for a rotation)
 *   is switched off for a proposal whose R^T R differs from I by more than 1e-4.
 *   cubes     [n_boxes][N][15] or NULL => sampled in-kernel (same generator and
 *             arguments as wcr_sample_cubes; priors/normals/z0/h_sample/seed are
 *             only read in that case)
 *   seg_bits  bit-packed per-box masks or NULL (term skipped); seg_off[n_boxes+1]
 *             word offsets; box b's mask covers its integer ROI
 *             [int(clip(x1)), int(clip(x2))) x [int(clip(y1)), int(clip(y2)))
 *             (Boxes.clip + .long(), roi_heads.py:590-597), row-major, bit (c & 31)
 *             of word r*ceil(ncols/32) + (c >> 5); padding bits must be zero.
 *             Mask term of a proposal (A3.12): projected corners clamped a second
 *             time to [0, clamp-1] (roi_heads.py:950-952) -> jarvis_march
 *             (utils.py:410-454) -> pixels strictly left of every hull edge
 *             (fill_polygon > 0, utils.py:456-488) -> mask_iou / mod_mask_iou
 *             (utils.py:225-245) against the box mask
 *   k <= WCR_MAX_TOPK; image sides <= 65535 pixels (the cull records hold 16-bit ROI indices)
 *   out_records [n_boxes][k][20]; slots beyond N hold idx = -1
 *   optional per-proposal outputs (NULL to skip), all [n_boxes][N]:
 *     out_scores, out_iou, out_seg, out_count (int32), out_proj [n_boxes][N][4]
 *   out_npts  [n_boxes] valid depth points per ROI (P_b), optional
 *   ws        workspace, wcr_score_workspace_bytes(n_boxes, N, k) bytes, 256-byte aligned.
 *             After the call the uint64 at byte 8 holds the number of point tests the
 *             kernel issued (pairs the projected-box cull could not rule out; the
 *             counts are those of the unculled evaluation) -- bench.py's frac_executed -- and the
 *             uint64 at byte 16 the number of hulls filled for the mask term.
 *   Mask term on demand: with seg_mode 0 and out_scores == out_seg == NULL, the hull of a
 *   proposal is filled only if its bound iou^e_iou * pts^e_pts can still reach the box's k-th
 *   best score (the ratio is <= 1, so the bound holds in floating point too); the records are
 *   those of the full evaluation, bit for bit. With plain terms (map_beta == 0) the bound also
 *   carries a bound of the mask ratio itself: |mask & hull| <= the pixels of the hull's bounding
 *   box inside the ROI, |hull| >= the largest projected face minus a perimeter margin.
 *   Point term on demand (pts_eval): with no per-proposal output and map_beta == 0 the depth
 *   pixels of a box are walked twice. The first pass tests the 64 proposals of each 1024-proposal
 *   chunk whose score BOUND is largest -- IoU term x (valid points of the 16 x 16 ROI cells the
 *   cube can reach / P_b) x the mask-ratio bound -- and takes T, the k-th best actual score among
 *   them; the second pass tests only proposals whose bound is not below T. The others cannot enter
 *   the top-k (every factor of the score is rounded like the bound's and is <= it), their counts
 *   are never formed; the records are those of the full evaluation, bit for bit. The uint64 at
 *   byte 40 of the workspace holds the number of proposals skipped.
 * wcr_score_bounds_offsets: byte offsets, inside the workspace, of the per-proposal bounds of the
 *   last call that used them: int32 [n_boxes][N] upper bounds of the inside-counts (0x7fffffff =
 *   none), int32 pairs [n_boxes][N][2] (max |mask & hull|, min |hull|), uint32 [n_boxes][N] cell
 *   ranges of the hulls' scan windows in the 32 x 32 mask grid (gx0 | gx1 << 6 | gy0 << 12 |
 *   gy1 << 18) and the grid's integral tables int32 [n_boxes][33][33] (max |mask & hull| is also
 *   bounded by the mask pixels of those cells). No reference
 *   counterpart; tests/test_score_bounds_gpu.py checks the bounds against the full evaluation. */
int64_t wcr_score_workspace_bytes(int n_boxes, int N, int k);
int wcr_score_bounds_offsets(int n_boxes, int N, int k, int64_t *off_ubcnt, int64_t *off_segb,
                             int64_t *off_segwin, int64_t *off_mint);
int wcr_score_topk(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                   const float *cubes, int N,
                   const float *priors, const float *normals, const float *z0,
                   const WcrSampleParams *h_sample, uint64_t seed,
                   const uint32_t *seg_bits, const int64_t *seg_off,
                   const WcrScoreParams *h_score, int k,
                   float *out_records, float *out_scores, float *out_iou, float *out_seg,
                   int32_t *out_count, float *out_proj, int32_t *out_npts,
                   void *ws, void *stream);
/* The persistent scoring kernel launches 3 CTAs per SM; wcr_score_set_reserved_ctas(n) makes the
 * following wcr_score_topk calls of this process launch n fewer (at least one per SM stays), which
 * leaves one CTA's worth of registers and shared memory free on n SMs for a kernel of another stream
 * -- wcr.pipeline.ShardPipeline runs the next shard's mask gather there. 0 = default. Results do not
 * depend on it. No reference counterpart. */
int wcr_score_set_reserved_ctas(int n);

/* wcr_score_topk whose scoring kernel -- the first launch of the chain that reads seg_bits -- waits
 * for the CUDA event `seg_ready` (cudaEvent_t or NULL) on `stream`: the kernels that prepare the
 * proposals do not need the masks and overlap a pack / upload still running on another stream
 * (wcr.pipeline.ShardPipeline). No reference counterpart. */
int wcr_score_topk_after(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                   const float *cubes, int N,
                   const float *priors, const float *normals, const float *z0,
                   const WcrSampleParams *h_sample, uint64_t seed,
                   const uint32_t *seg_bits, const int64_t *seg_off,
                   const WcrScoreParams *h_score, int k,
                   float *out_records, float *out_scores, float *out_iou, float *out_seg,
                   int32_t *out_count, float *out_proj, int32_t *out_npts,
                   void *ws, void *seg_ready, void *stream);

/* Dense per-box object masks (the reference's masks_all_images, roi_heads.py:273, 383-401: one
 * H x W map per object, non-zero == object) -> the seg_bits layout of wcr_score_topk, on the
 * device. Box b reads mask (mask_idx ? mask_idx[b] : b), mask_stride bytes apart, rows
 * mask_pitch bytes apart, over its integer ROI; words land at out_bits + seg_off[b] (the
 * offsets the caller derives from the ROIs: nrows * ceil(ncols / 32) words per box).
 * `masks` may also be PINNED host memory (mapped into the device's address space under unified
 * addressing): the kernel then reads the ROI bytes over PCIe and no copy is needed. */
int wcr_pack_box_masks(const uint8_t *masks, int64_t mask_stride, int mask_pitch,
                       const int32_t *mask_idx, const WcrImage *images, const WcrBox *boxes,
                       int n_boxes, const int64_t *seg_off, uint32_t *out_bits, void *stream);

/* Host -> device copy of just the ROI rectangle of every box's dense mask (one cudaMemcpy2DAsync
 * per box on `stream`): what wcr_pack_box_masks reads, and nothing else, crosses PCIe. h_masks:
 * PINNED host memory, same layout as d_masks; h_rois [n_boxes][4] = u0 v0 u1 v1, HOST array;
 * h_mask_idx HOST array or NULL. No device check: enqueues copies only. */
int wcr_h2d_box_mask_rois(const uint8_t *h_masks, uint8_t *d_masks, int64_t mask_stride,
                          int mask_pitch, const int32_t *h_rois, const int32_t *h_mask_idx,
                          int n_boxes, void *stream);
/* The same refresh as ONE contiguous copy per box, from the ROI's first pixel to its last with the
 * whole rows in between ((nrows - 1) * pitch + ncols bytes; *out_bytes = the total, may be NULL):
 * more bytes than the rectangles, but contiguous blocks move at the link's full rate on a copy
 * engine where 2-D copies of short rows do not. No reference counterpart. */
int wcr_h2d_box_mask_rows(const uint8_t *h_masks, uint8_t *d_masks, int64_t mask_stride,
                          int mask_pitch, const int32_t *h_rois, const int32_t *h_mask_idx,
                          int n_boxes, int64_t *out_bytes, void *stream);

/* ------------------------------------------------- block-level drop-ins -- */
/* One launch behind each helper the reference exposes by name (SURVEY A2.1-A2.5,
 * A3.2-A3.8); the fused entry points above use the same device functions. */

/* cubercnn/util/math_util.py:142-217 get_cuboid_verts_faces, as wrapped by
 * Cubes.get_all_corners (ProposalNetwork/utils/spaces.py:192-204).
 *   cubes [n][15] -> out_corners [n][8][3] */
int wcr_cube_corners(const float *cubes, int64_t n, float *out_corners, void *stream);

/* Cubes.get_bube_corners(K, clamp) (spaces.py:224-245) and cubes_to_box
 * (ProposalNetwork/utils/conversions.py:25-48): K @ corners, divide by z, clamp to
 * [int(-c/2+1), int(c-1+c)], min/max. Either output may be NULL.
 *   out_corners [n][8][2], out_boxes [n][4] (x1 y1 x2 y2) */
int wcr_bube_corners(const float *cubes, int64_t n, float fx, float fy, float cx, float cy,
                     float clamp0, float clamp1, float *out_corners, float *out_boxes,
                     void *stream);

/* ProposalNetwork/utils/utils.py:181-187 iou_2d (detectron2 pairwise_iou of ONE
 * GT box against P boxes).  gt_box [4] device, boxes [P][4] 16-byte aligned. */
int wcr_iou_2d(const float *gt_box, const float *boxes, int64_t P, float *out_iou, void *stream);

/* utils.py:490-493 convex_hull(mask, coords) = fill_polygon(jarvis_march(coords)):
 * soft mask in [0,1], product of the clamped half-plane values of every hull edge.
 *   coords [n_points][2] device (1..32 points), out_mask H x W, row pitch `pitch`.
 * An unclosed march (the reference would not terminate) gives zeros. */
int wcr_convex_hull(const float *coords, int n_points, int H, int W, float *out_mask,
                    int64_t pitch, void *stream);

/* utils.py:225-234 mask_iou (mode 0) / :236-245 mod_mask_iou (mode 1).
 *   seg: numel u8 (seg_is_float == 0) or f32 values; bube: numel f32;
 *   out [1]; ws: 32 bytes of device scratch, 8-byte aligned. */
int wcr_mask_iou(const void *seg, int seg_is_float, const float *bube, int64_t numel, int mode,
                 float *out, void *ws, void *stream);

/* utils.py:41-59 sample_normal_in_range with the in-kernel Philox4x32-10 generator
 * (counter = (column, row, tag, block), key = seed): out [M][count]; with
 * has_thresholds every element is redrawn until it lies in
 * [threshold_low, threshold_high[m]] (at most max_tries draws, then clamped). */
int wcr_sample_normal_in_range(const float *means, const float *stds, int M, int count,
                               int has_thresholds, float threshold_low,
                               const float *threshold_high, uint64_t seed, int max_tries,
                               float *out, void *stream);

/* utils.py:119-131 orthobasis_from_normal_t (with :111-117 vec_perp_t and :133-145
 * rotate_vector_t): normal [3], yaw [T] -> out_R [T][3][3], columns [x, n, y]. */
int wcr_orthobasis_from_normal(const float *normal, const float *yaw, int T, float *out_R,
                               void *stream);

/* utils.py:169-176 vectorized_linspace: out [M][steps] =
 * arange(steps) * ((end - start) / (steps - 1)) + start. */
int wcr_vectorized_linspace(const float *start, const float *end, int M, int steps, float *out,
                            void *stream);

/* ---------------------------------------------------------------- step 4 -- */

/* Weak-loss block of ROIHeads3DScore._forward_cube (roi_heads.py:786-905 decode,
 * :927-960 cubes + projected boxes, :973-1031 terms, :1075-1144 uncertainty scaling
 * and safe reductions; helpers :445-464 pose_loss, :612-656 dim_loss /
 * pseudo_gt_z_point_loss, :1307-1315 safely_reduce_losses; math_util.py:802-830
 * R_from_allocentric, :1132-1237 so3_relative_angle_batched; torchvision
 * generalized_box_iou_loss) as a fused forward / backward kernel pair. */
#define WCR_LOSS_IOU          1   /* 'iou'                 */
#define WCR_LOSS_POSE_ALIGN   2   /* 'pose_alignment'      */
#define WCR_LOSS_POSE_GROUND  4   /* 'pose_ground'         */
#define WCR_LOSS_Z_CENTER     8   /* 'z_pseudo_gt_center'  */
#define WCR_LOSS_DIMS        16   /* 'dims'                */
#define WCR_LOSS_Z_PATCH     32   /* 'z_pseudo_gt_patch' (takes precedence over the centre lookup,
                                     roi_heads.py:1010-1013; same output slot WCR_T_PSEUDO_Z) */
/* tier-2 terms */
#define WCR_LOSS_Z_SEARCH    64   /* 'z': z_loss, roi_heads.py:530-573 (50 offsets of 0.1 m, arg-min of
                                     the projected-area mismatch; value only, zero gradient) */
#define WCR_LOSS_DEPTH      128   /* 'depth': depth_range_loss, :658-684 (q90 - q10 of the depth under
                                     the object's mask vs the depth extent of the cube corners) */
#define WCR_LOSS_POSE_GROUND2 256 /* 'pose_ground2': normal_to_rotation + so3_relative_angle, :686-696,
                                     :997-1007; overrides 'pose_ground' in slot WCR_T_NORMAL */
#define WCR_LOSS_SEG        512   /* 'segmentation': segment_loss, :420-443 (soft convex-hull raster of the
                                     clamped projected corners vs the object mask, focal loss; the raster
                                     is differentiable in the hull vertices, utils.py:479-483) */

/* order of the scalar outputs (keys of the reference's loss dict, prefix 'Cube/') */
enum { WCR_T_UNCERT = 0, WCR_T_IOU, WCR_T_POSE, WCR_T_NORMAL, WCR_T_PSEUDO_Z,
       WCR_T_DIMS_W, WCR_T_DIMS_H, WCR_T_DIMS_L, WCR_T_Z, WCR_T_DEPTH, WCR_T_SEG, WCR_LOSS_TERMS };
/* slots of WcrLossIO.stats beyond the per-term valid counts */
enum { WCR_STAT_POSE = 16, WCR_STAT_FAIL = 17,
       WCR_STAT_TRACE_ERR = 18,  /* != 0: a relative rotation of 'pose_ground2' has its trace outside
                                    [-1-1e-4, 3+1e-4] (pytorch3d raises ValueError there)          */
       WCR_STAT_SHAPE_ERR = 19,  /* != 0: 'depth' with a mask whose shape differs from its depth map */
       WCR_LOSS_STATS = 32 };
#define WCR_Z_LOSS_STEPS 50      /* max_count of z_loss (roi_heads.py:531) */

/* torch.linspace(0, 4.9, 50) as z_loss builds it (roi_heads.py:547): host function, no GPU
 * needed; the fused kernel uses exactly this table. */
void wcr_z_loss_steps(float out[WCR_Z_LOSS_STEPS]);

typedef struct WcrLossParams {
    int32_t loss_flags;       /* WCR_LOSS_* bits (cfg.loss_functions, config.py:186)   */
    int32_t z_type;           /* 0 direct, 1 sigmoid, 2 log  (roi_heads.py:874-880)    */
    int32_t dims_func;        /* 0 exp, 1 sigmoid            (roi_heads.py:856-861)    */
    float   use_confidence;   /* cfg USE_CONFIDENCE; 0 disables uncertainty scaling    */
    float   w_3d, w_iou, w_pose, w_normal, w_z, w_dims;   /* LOSS_W_* (config.py:119-127) */
    float   w_depth;          /* LOSS_W_DEPTH (config.py:127)                          */
    float   w_seg;            /* LOSS_W_SEG (config.py:124)                            */
} WcrLossParams;

/* Every per-box array is indexed by foreground box, boxes of one image contiguous
 * (the reference's torch.cat order); img_off[n_images+1] are the box ranges. */
typedef struct WcrLossIO {
    int32_t n, n_images;
    /* head outputs, gathered per class (require grad in the reference) */
    const float *deltas;      /* [n][2]  cube_2d_deltas                       */
    const float *z_raw;       /* [n]     cube_z before z_type / virtual depth */
    const float *dims_raw;    /* [n][3]  cube_dims (log-scale for 'exp')      */
    const float *pose;        /* [n][9]  allocentric R_view, row-major        */
    const float *uncert;      /* [n]     cube_uncert                          */
    /* constants */
    const float *src_boxes;   /* [n][4]  proposal boxes the deltas apply to   */
    const float *K4;          /* [n][4]  fx fy cx cy of Ks_scaled_per_box     */
    const float *v2r;         /* [n]     virtual_to_real (math_util.py:732)   */
    const float *prior_mean;  /* [n][3]  gathered dims prior (w h l)          */
    const float *prior_std;   /* [n][3]                                       */
    const float *gt_boxes;    /* [n][4]                                       */
    const float *clamp;       /* [n][2]  im_sizes[i] handed to get_bube_corners */
    const int32_t *img_off;   /* [n_images+1]                                 */
    const WcrImage *images;   /* [n_images] depth maps (z_pseudo_gt_center)   */
    const float *normals;     /* [n_images][3] from wcr_ransac_ground         */
    const float *ground_conf; /* [n_images] 0.1 for the (1,1) dummy else 1 (roi_heads.py:988) */
    /* forward outputs / backward inputs */
    float   *losses;          /* [WCR_LOSS_TERMS] scalars, already times w_term * w_3d */
    int32_t *present;         /* [WCR_LOSS_TERMS + 2] 0 where the reference would not emit the key (None);
                                 then copies of the WCR_STAT_TRACE_ERR / WCR_STAT_SHAPE_ERR flags so one
                                 small D2H read tells the caller everything */
    float   *terms;           /* [WCR_LOSS_TERMS-1][n] raw per-box terms before uncertainty scaling,
                                 row t-1 for term t (row WCR_T_POSE-1 holds the broadcast scalar) */
    float   *R_ego;           /* [n][9] egocentric pose after R_from_allocentric */
    float   *proj;            /* [n][4] projected boxes                        */
    float   *stats;           /* [WCR_LOSS_STATS] n_valid per term, then the WCR_STAT_* slots */
    float   *img_pose;        /* [n_images] per-image pose term (workspace)    */
    /* backward only */
    const float *grad_losses; /* [WCR_LOSS_TERMS] upstream gradient of each scalar */
    float *d_deltas, *d_z_raw, *d_dims_raw, *d_pose, *d_uncert;  /* same shapes as inputs */
    float *d_R_pose;          /* [n][9] workspace                              */
    /* 'z_pseudo_gt_patch' only (NULL otherwise) */
    float   *z_dec;           /* [n] decoded cube_z                            */
    float   *z_gt;            /* [n] per-box pseudo depth (median / fallback)  */
    int32_t *z_inside;        /* [n] 1 = box has area inside the image         */
    float   *z_sign;          /* [n] sign(cube_z - pseudo depth), by position  */
    /* 'depth' and 'segmentation' (NULL otherwise): per-object masks as the reference's masks_all_images
     * (roi_heads.py:273), dense, non-zero == object. Every mask must have the shape of the
     * depth map of the image its boxes belong to (the reference's align_corners bilinear
     * resize, :672, is the identity then) unless depth_images (below) is given; a mismatch sets
     * stats[WCR_STAT_SHAPE_ERR]. */
    const uint8_t *masks;     /* [n_masks][mask_h][mask_w]                     */
    int32_t  mask_h, mask_w;
    const int32_t *mask_idx;  /* [n] at_which_mask_idx (:723-726)              */
    float   *depth_gt;        /* [n] out: q90 - q10 of the masked depth (GT box patch when the mask is empty) */
    /* 'pose_ground2' only (NULL otherwise) */
    float   *gt_rot;          /* [n_images][9] workspace: normal_to_rotation per image */
    /* 'z_pseudo_gt_patch' and 'depth' (NULL otherwise): wcr_select_ws_bytes(n) bytes, 256-byte aligned */
    void    *select_ws;
    /* 'segmentation', backward only: [n][16] d raw_seg / d (clamped projected corners) */
    float   *seg_dcorner;
    /* 'depth' with masks of ANOTHER shape than the depth maps (the reference's masks have the padded
     * batch-tensor size, the depth maps are un-padded): [n_images] depth maps already resized to
     * (mask_h, mask_w) by wcr_resize_bilinear_ac -- what F.interpolate(bilinear, align_corners=True)
     * at roi_heads.py:672 yields. NULL: `images` is used and a shape mismatch is an error. */
    const WcrImage *depth_images;
} WcrLossIO;

/* Workspace of the per-box order-statistic kernels (median of a depth patch, quantiles of
 * the depth under a mask): per box a 128-byte state and a 4 x 256-bin histogram, plus the
 * work-item prefix. Host function, no GPU needed. */
int64_t wcr_select_ws_bytes(int n);

/* ROIHeads3DScore.segment_loss (roi_heads.py:420-443), loss='focal': per box the mean over the
 * mask's pixels of sigmoid_focal_loss(inputs = object mask as 0/1, targets = convex_hull raster of
 * the 8 given corners (utils.py:410-493)). masks [n_masks][mask_h][mask_w] u8, mask_idx [n],
 * corners [n][8][2] (already clamped as roi_heads.py:950-953 does) -> out_loss [n].
 * Backward: d_corners [n][8][2] = d out_loss[i] / d corners[i] (the soft raster is differentiable
 * in the hull vertices; ties at val == 0 / val == 1 split in half like torch.max / torch.min). */
int wcr_segment_loss_fwd(const uint8_t *masks, int mask_h, int mask_w, const int32_t *mask_idx,
                         const float *corners, int n, float *out_loss, void *stream);
int wcr_segment_loss_bwd(const uint8_t *masks, int mask_h, int mask_w, const int32_t *mask_idx,
                         const float *corners, int n, float *d_corners, void *stream);

/* ---- callers and data formats either side of the path ------------ */

/* pytorch3d.transforms.rotation_6d_to_matrix as cube_head.py:174-176 calls it: Gram-Schmidt of
 * two 3-vectors, rows b1, b2, b1 x b2. d6 [n][6] -> out_R [n][9]; backward: grad_R [n][9] -> d_d6. */
int wcr_rotation_6d_fwd(const float *d6, int n, float *out_R, void *stream);
int wcr_rotation_6d_bwd(const float *d6, const float *grad_R, int n, float *d_d6, void *stream);

/* x[fg_inds, box_classes, ...] of roi_heads.py:823-840: src [n][C][k], classes [n] int64 ->
 * out [n][k]; backward scatters grad_out into the zero-filled d_src [n][C][k]. */
int wcr_class_gather_fwd(const float *src, int n, int C, int k, const int64_t *classes, float *out,
                         void *stream);
int wcr_class_gather_bwd(const float *grad_out, int n, int C, int k, const int64_t *classes,
                         float *d_src, void *stream);

/* Per-box constants of roi_heads.py:748-781, 810-814: Ks [B][9], im_scales_ratio [B],
 * im_current_dims [B][2] -> K4 [n][4] = (fx fy cx cy) / ratio, virtual_to_real [n]
 * (math_util.py:732-743; 1 when virtual_depth == 0), clamp [n][2] = im_current_dims. */
int wcr_box_constants(const float *Ks, const float *im_scales_ratio, const float *im_current_dims,
                      const int32_t *img_off, int n_images, int n, float virtual_focal,
                      int virtual_depth, float *out_K4, float *out_v2r, float *out_clamp, void *stream);

/* dataset_mapper.py:126-145: PIL Image.resize(size, Image.NEAREST) of the depth map (f32) and
 * the ground mask (bool). wcr_pil_nearest_table reproduces PIL's source-index table for one axis
 * (host function, no GPU needed: ImagingScaleAffine accumulates the coordinate in double);
 * wcr_resize_nearest gathers through the two tables on the device. elem_size 4 or 1;
 * pitches in elements. */
void wcr_pil_nearest_table(int in_size, int out_size, int32_t *table);
int wcr_resize_nearest(const void *src, int elem_size, int in_h, int in_w, int64_t src_pitch,
                       const int32_t *xtab, const int32_t *ytab, void *dst, int out_h, int out_w,
                       int64_t dst_pitch, void *stream);

/* F.interpolate(x, size=(out_h, out_w), mode='bilinear', align_corners=True) of one f32 plane, as
 * depth_range_loss applies it to a depth map (roi_heads.py:672). Pitches in elements. */
int wcr_resize_bilinear_ac(const float *src, int in_h, int in_w, int64_t src_pitch, float *dst,
                           int out_h, int out_w, int64_t dst_pitch, void *stream);

/* ---- 3D IoU ------------------------------------------------------------ */

/* pytorch3d.ops.box3d_overlap as iou_3d (ProposalNetwork/utils/utils.py:189-205), roi_heads.py:942
 * and omni3d_evaluation.py:106-166 call it: exact intersection volume and IoU of every pair of
 * oriented boxes. boxes1 [N][8][3], boxes2 [M][8][3] corner lists in pytorch3d's corner order
 * (what Cubes.get_all_corners returns); out_vol, out_iou [N][M]. */
int wcr_box3d_overlap(const float *boxes1, int N, const float *boxes2, int M, float *out_vol,
                      float *out_iou, void *stream);
/* volume of each box [n][8][3] -> out_vol [n] (pytorch3d raises on boxes of ~zero volume; the
 * Python shim uses this for the same check). */
int wcr_box3d_volume(const float *boxes, int n, float *out_vol, void *stream);

/* sizeof() of the ABI structs as the library was compiled, so a binding can verify its
 * mirror: 0 WcrImage, 1 WcrBox, 2 WcrSampleParams, 3 WcrScoreParams, 4 WcrLossParams,
 * 5 WcrLossIO; anything else -> 0. Host function, no GPU needed. */
int64_t wcr_sizeof(int which);

int wcr_weak_loss_fwd(const WcrLossIO *h_io, const WcrLossParams *h_params, void *stream);
int wcr_weak_loss_bwd(const WcrLossIO *h_io, const WcrLossParams *h_params, void *stream);

/* ---- stand-alone loss helpers (one forward + one backward launch each) ---------- */

/* ROIHeads3DScore.pose_loss (roi_heads.py:445-464) with so3_relative_angle_batched
 * (math_util.py:1132-1181, cos_angle=True): per image with n_i > 1 boxes the mean over
 * pairs i > j of 1 - |(tr(R_i R_j^T) - 1)/2|; summed and divided by (fail + 1), fail =
 * images with exactly one box.
 *   R [n][9], img_off [n_images+1]; img_pose [n_images] scratch;
 *   out3 = {loss, fail, n_images} (the caller returns None when fail == n_images). */
int wcr_pose_loss_fwd(const float *R, const int32_t *img_off, int n_images, float *img_pose,
                      float *out3, void *stream);
/* grad1 [1]: upstream gradient of the scalar; d_R [n][9]. */
int wcr_pose_loss_bwd(const float *R, const int32_t *img_off, int n_images, const float *fwd_out3,
                      const float *grad1, float *d_R, void *stream);

/* ROIHeads3DScore.pseudo_gt_z_point_loss (roi_heads.py:635-656): |pred_z - depth[y, x]|
 * at the 10-px-clamped, truncated predicted centre.  pred_xy [n][2], pred_z [n];
 * out_loss [n]; out_sign [n] = sign(pred_z - gt) for wcr_l1_loss_bwd. */
int wcr_pseudo_z_point_loss(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                            const float *pred_xy, const float *pred_z, float *out_loss,
                            float *out_sign, void *stream);

/* ROIHeads3DScore.pseudo_gt_z_box_loss (roi_heads.py:575-610): pseudo depth = LOWER
 * median (torch.median) of the depth patch under each box after Boxes.clip and .long();
 * boxes with zero area after clipping fall back to the centre lookup AND are appended
 * after the in-image boxes of their image, so the pairing with pred_z is by position in
 * that re-ordered list (reference behaviour, kept).
 *   boxes [n][4]; ws_gt [n], ws_inside [n] scratch (per-box pseudo depth / in-image flag);
 *   select_ws: wcr_select_ws_bytes(n) bytes. */
int wcr_pseudo_z_box_loss(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                          const float *boxes, const float *pred_z, float *out_loss, float *out_sign,
                          float *ws_gt, int32_t *ws_inside, void *select_ws, void *stream);

/* backward of smooth_l1(beta = 0) (roi_heads.py:352-353): d_pred[i] = grad[i] * sign[i]. */
int wcr_l1_loss_bwd(const float *sign, const float *grad, int n, float *d_pred, void *stream);

/* ROIHeads3DScore.dim_loss (roi_heads.py:612-633): max(|d - mean| / std - 1, 0) per axis.
 *   prior_mean, prior_std, dims [n][3]; out3n [3][n]; nan_flag [1] = number of NaN stds
 *   (the reference returns (None, None, None) when any is NaN). */
int wcr_dim_loss_fwd(const float *prior_mean, const float *prior_std, const float *dims, int n,
                     float *out3n, int32_t *nan_flag, void *stream);
int wcr_dim_loss_bwd(const float *prior_mean, const float *prior_std, const float *dims,
                     const float *grad3n, int n, float *d_dims, void *stream);

/* ---- second set of block-level drop-ins ------------------------ */

/* Backward of wcr_cube_corners: grad_corners [n][8][3] -> d_cubes [n][15] (centre, w h l, R).
 * The reference differentiates through Cubes.get_all_corners (depth_range_loss, roi_heads.py:665). */
int wcr_cube_corners_bwd(const float *cubes, const float *grad_corners, int64_t n, float *d_cubes,
                         void *stream);

/* Backward of wcr_bube_corners (get_bube_corners, spaces.py:224-245, and the min / max of
 * cubes_to_box, conversions.py:36-42; call site inside autograd: roi_heads.py:947-960).
 * grad_corners [n][8][2] and / or grad_boxes [n][4] (either may be NULL) -> d_cubes [n][15].
 * torch conventions: clamp passes the gradient on the closed interval, min / max send it to the
 * first extremum. */
int wcr_bube_corners_bwd(const float *cubes, int64_t n, float fx, float fy, float cx, float cy,
                         float clamp0, float clamp1, const float *grad_corners,
                         const float *grad_boxes, float *d_cubes, void *stream);

/* ProposalNetwork/utils/utils.py:207-223 custom_mapping: x <= 0 -> 0, else
 * 1 / (1 + (x / (1 - x)) ** -beta), evaluated in double like the reference's Python floats. */
int wcr_custom_mapping(const float *x, int64_t n, double beta, float *out, void *stream);

/* utils.py:61-68 randn_orthobasis_torch: z [n][3][3] standard-normal draws (replay of the
 * reference's torch.randn) or NULL => in-kernel Philox4x32-10 (counter (basis, "ORTH", block),
 * key = seed) -> out [n][3][3]: rows normalised, row 0 = cross(row 1, row 2), row 1 = cross(row 2,
 * row 0), each re-normalised. */
int wcr_randn_orthobasis(const float *z, int64_t n, uint64_t seed, float *out, void *stream);

/* utils.py:148-152 gt_in_norm_range, element-wise over n (range, gt) pairs:
 * out = (gt - lo) / |hi - lo|. */
int wcr_gt_in_norm_range(const float *lo, const float *hi, const float *gt, int64_t n, float *out,
                         void *stream);

/* utils.py:257-298 is_gt_included for n GT cubes at once. gt_cubes [n][15]; ranges [n][6] =
 * x_lo x_hi y_lo y_hi z_lo z_hi; priors [n][6] = (mean, std) of w, h, l. out_bits [n]: 0 ==
 * included, else OR of 1 x, 2 y, 4 z, 8 w-, 16 w+, 32 h-, 64 h+, 128 l-, 256 l+ (the reference's
 * `because_of` list); out_by [n][3]: the x / y / z distances it prints. */
int wcr_is_gt_included(const float *gt_cubes, const float *ranges, const float *priors, int64_t n,
                       int32_t *out_bits, float *out_by, void *stream);

/* ROIHeads3DScore.z_loss (roi_heads.py:530-573) on given cubes [n][15]: K4 [n][4], clamp [n][2]
 * (im_sizes[i]), proj_boxes [n][4] -> out [n]. Value only (the reference's graph has no path to
 * the cube either: the arg-min offset is a constant). */
int wcr_z_loss(const float *gt_boxes, const float *cubes, const float *K4, const float *clamp,
               const float *proj_boxes, int n, float *out, void *stream);

/* ROIHeads3DScore.normal_to_rotation (roi_heads.py:686-696): normal [n][3] -> out [n][3][3] with
 * rows [t0, t1, normal]; t0 and t1 are divided by the Frobenius norm of the WHOLE (n, 3) matrix
 * and the y-axis fallback is taken only when every t0 is zero (reference behaviour, kept). */
int wcr_normal_to_rotation(const float *normal, int n, float *out, void *stream);

/* q90 - q10 of the depth under each object's mask (GT box patch when the mask is empty): the GT
 * side of ROIHeads3DScore.depth_range_loss (roi_heads.py:658-684) as a stand-alone call; same
 * argument meaning as the WcrLossIO fields of the same names. stats: WCR_LOSS_STATS floats
 * (stats[WCR_STAT_SHAPE_ERR] != 0: a mask whose shape differs from its depth map). */
int wcr_masked_depth_range(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                           const uint8_t *masks, int mask_h, int mask_w, const int32_t *mask_idx,
                           const float *gt_boxes, float *out_range, float *stats, void *select_ws,
                           void *stream);

#ifdef __cplusplus
}
#endif
#endif /* WCR_H_ */
