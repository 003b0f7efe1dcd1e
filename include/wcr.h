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

#define WCR_ABI_VERSION 1

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
/* number of kernel launches enqueued by this library on the calling thread
 * since the last wcr_reset_launch_count() (bench.py's gpu_launches). */
int64_t     wcr_launch_count(void);
void        wcr_reset_launch_count(void);

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

#ifdef __cplusplus
}
#endif
#endif /* WCR_H_ */
