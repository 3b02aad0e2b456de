/*
 * svb.h — C ABI of the B200-native stereo depth engine (libsvb.so).
 *
 * The reference (jacksonroah/stereo-vision) has no FFI of its own: its dense hot path is the set of
 * cv2 calls its helpers make (`import cv2` at the top of every script).  Each entry point below names
 * the cv2 call it replaces and the reference call site that drives it:
 *
 *   svb_init_undistort_rectify_map  <- cv2.initUndistortRectifyMap(K,d,R,P,size,CV_32FC1)
 *        stereo_calibration/scripts/checkerboard_distance.py:54-57
 *        calibration_tests/calibrate_v4/distance.py:148-151, calibrate_v4/extrinsic.py:371-374
 *   svb_remap_u8                    <- cv2.remap(img,map1,map2,cv2.INTER_LINEAR)
 *        checkerboard_distance.py:71-72, calibrate_v4/distance.py:154-155, calibrate_v4/extrinsic.py:377-378
 *   svb_sgbm_*                      <- cv2.StereoSGBM_create(**kw).compute(left,right)   (north_star; the
 *        reference computes Q for it at checkerboard_distance.py:49-51 / calibrate_v4/extrinsic.py:312-324)
 *   svb_median3_s16 / svb_filter_speckles_s16  <- the medianBlur(3)/filterSpeckles stages inside compute()
 *   svb_reproject_to_3d             <- cv2.reprojectImageTo3D(disp,Q,handleMissingValues)
 *   svb_box_median_depth            <- per-YOLO-box depth (ball_drop.py:89-160 box -> :225-259 3-D point),
 *        re-defined on the dense map
 *
 * Conventions: every function returns 0 on success, <0 on error (message: svb_last_error(), thread-local).
 * All data pointers are DEVICE pointers owned by the caller (e.g. torch tensors' data_ptr()); `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  Nothing here allocates memory the caller
 * must free, except the context itself.  One context per device is enough: StereoSGBM jobs on one context
 * are serialised (host threads by a mutex while they enqueue; a job on another stream than the previous one
 * first waits on the device for it), because they share the context's scheduler counters and cached workspace.
 * There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef SVB_H
#define SVB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVB_OK 0
#define SVB_ERR_INVALID (-1)   /* bad argument (cv2.error -215 / -2 cases)           */
#define SVB_ERR_CUDA (-2)      /* CUDA runtime error                                  */
#define SVB_ERR_UNSUPPORTED (-3) /* valid for cv2 but outside this engine's contract  */
#define SVB_ERR_WORKSPACE (-4) /* caller workspace too small                          */

typedef struct svb_context svb_context;

/* The 11 StereoSGBM_create parameters, RAW (as the caller passed them to cv2); resolution to effective
 * values happens inside the library — see svb_sgbm_resolve_params. */
typedef struct svb_sgbm_params {
    int32_t minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff, preFilterCap, uniquenessRatio,
        speckleWindowSize, speckleRange, mode; /* mode: 0 = MODE_SGBM (5 paths), 1 = MODE_HH (8), 2 = MODE_SGBM_3WAY, 3 = MODE_HH4 (4) */
} svb_sgbm_params;

/* Geometry derived from params + image width (SURVEY.md §3.2.1 "Constants"). */
typedef struct svb_sgbm_geometry {
    int32_t minX1, maxX1, width1; /* matched columns [minX1,maxX1)                       */
    int32_t D, Dp;                /* numDisparities and its padded in-HBM extent          */
    int32_t invalid16;            /* (minDisparity-1)*16                                   */
    int32_t max_cost;             /* bound on C(p,d): bs^2 * channels * (2*ftzero+63)      */
    int32_t safe_packed;          /* 1 if S accumulates without per-add saturation        */
} svb_sgbm_geometry;

const char *svb_last_error(void);
const char *svb_version(void);

int svb_create(int device_id, svb_context **ctx);
int svb_destroy(svb_context *ctx);

/* ---- parameter resolution (host only; no device needed) ---- */
int svb_sgbm_resolve_params(const svb_sgbm_params *raw, svb_sgbm_params *effective);
int svb_sgbm_get_geometry(const svb_sgbm_params *raw, int W, int channels, svb_sgbm_geometry *g);

/* ---- StereoSGBM.compute ---- */
/* Bytes of device workspace needed for a batch of B pairs processed concurrently. */
int svb_sgbm_workspace_bytes(const svb_sgbm_params *raw, int H, int W, int channels, int B, size_t *bytes);
/* left/right: u8 [B][H][pitch] (channels interleaved, 1 or 3), disp: int16 [B][H][disp_pitch_elems].
 * ws may be NULL: the context then keeps a grow-only cached workspace of its own. */
int svb_sgbm_compute(svb_context *ctx, const svb_sgbm_params *raw, const uint8_t *left, const uint8_t *right,
                     int H, int W, int channels, size_t pitch_bytes, size_t image_stride_bytes, int B,
                     int16_t *disp, size_t disp_pitch_elems, size_t disp_stride_elems, void *ws, size_t ws_bytes,
                     void *stream);
/* Small jobs (B * ceil(H/15 or 16) <= 74 bands, modes 0 and 1) run in the LATENCY FORM by default: the forward and the
 * reverse sweep run side by side on two streams (forked from and joined to `stream` by events inside the call), each
 * into its own sum volume, and a third kernel adds them and selects the disparity — about half the wavefront time of a
 * single pair (the drop-in cv2.StereoSGBM_create(...).compute(L, R) call), at the price of one more volume in the
 * workspace (svb_sgbm_workspace_bytes accounts for it).  Results are identical.  svb_set_latency_form(ctx, 0) makes
 * every job run the two sweeps back to back (the throughput form larger batches always use); 1 restores the default. */
int svb_set_latency_form(svb_context *ctx, int enable);
/* Stage taps for parity tests (single pair): cost volume C and the pass-1 aggregate S (sum of the four
 * forward paths; the final S only ever exists in registers), both int16 [H][width1][Dp]; disp_raw = selection
 * + LR-check output before median/speckle.  Any output may be NULL. */
int svb_sgbm_debug_stages(svb_context *ctx, const svb_sgbm_params *raw, const uint8_t *left, const uint8_t *right,
                          int H, int W, int channels, size_t pitch_bytes, int16_t *Cvol, int16_t *Svol,
                          int16_t *disp_raw, void *stream);

/* Per-stage device times (ms) of the last svb_sgbm_compute on this context (CUDA events on the caller's stream):
 * ms[0..5] = planes, cost volume, aggregation pass 1, aggregation pass 2 + selection, LR check + median, speckle.
 * Enable with svb_set_timing(ctx,1); svb_sgbm_last_stage_ms blocks until that call has finished. */
int svb_set_timing(svb_context *ctx, int enable);
int svb_sgbm_last_stage_ms(svb_context *ctx, float *ms, int n);

/* ---- post filters (also usable stand-alone, like cv2.medianBlur(.,3) / cv2.filterSpeckles) ---- */
int svb_median3_s16(svb_context *ctx, const int16_t *src, int16_t *dst, int H, int W, size_t pitch_elems, int B,
                    size_t stride_elems, void *stream);
/* labels_ws: int32 workspace of at least 2*H*W elements per image (B images). In place on img. */
int svb_filter_speckles_s16(svb_context *ctx, int16_t *img, int H, int W, size_t pitch_elems, int B,
                            size_t stride_elems, int newVal, int maxSpeckleSize, int maxDiff, int32_t *labels_ws,
                            size_t labels_ws_elems, void *stream);

/* No-allocation variant of svb_filter_speckles_s16.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The label workspace is a static device array inside libsvb.so
 * (SVB_NOALLOC_SPECKLE_MAX_PIXELS * 8 bytes = 16 MiB, held by every GPU that loads the library until exit); the call
 * allocates nothing.  B*H*W must be <= SVB_NOALLOC_SPECKLE_MAX_PIXELS, otherwise SVB_ERR_WORKSPACE.  All calls on a
 * device share the array: calls that overlap on the GPU (different streams without ordering) race.  The first call
 * per device resolves the array's address (cudaGetSymbolAddress): make it once before starting a graph capture. */
#define SVB_NOALLOC_SPECKLE_MAX_PIXELS (1 << 21)   /* e.g. one 1920x1080 map, or 2 x 1280x720 */
int svb_filter_speckles_s16_noalloc(svb_context *ctx, int16_t *img, int H, int W, size_t pitch_elems, int B,
                                    size_t stride_elems, int newVal, int maxSpeckleSize, int maxDiff, void *stream);

/* ---- rectification ---- */
/* K[9], dist[ndist] (ndist in 0,4,5,8,12,14), R[9] (NULL = I), P33[9] = P[:3,:3] (NULL = K): HOST pointers
 * (calibration-time scalars).  mapx/mapy: device f32 [H][W]. */
int svb_init_undistort_rectify_map(svb_context *ctx, const double *K, const double *dist, int ndist,
                                   const double *R, const double *P33, int W, int H, float *mapx, float *mapy,
                                   void *stream);
/* src: u8 [B][sH][spitch] with C in {1,3}; mapx/mapy f32 [dH][dW] shared by the batch;
 * dst: u8 [B][dH][dpitch].  INTER_LINEAR, BORDER_CONSTANT(0): cv2's fixed-point tables. */
int svb_remap_u8(svb_context *ctx, const uint8_t *src, int sH, int sW, int C, size_t spitch, size_t sstride,
                 const float *mapx, const float *mapy, uint8_t *dst, int dH, int dW, size_t dpitch, size_t dstride,
                 int B, void *stream);
/* The CV_16SC2 form of the two calls above: map1 s16 [H][W][2] = floor of the coordinate, map2 u16 [H][W] =
 * (fy<<5)|fx, rounded from the f64 coordinate (cvRound(u*32)).  cv2.undistort(img, K, dist) is exactly these two
 * calls with R = I, P = K (reference: calibrate_v4/distance.py:125-126, calibrate/test_calibration.py undistort). */
int svb_init_undistort_rectify_map_fixed(svb_context *ctx, const double *K, const double *dist, int ndist,
                                         const double *R, const double *P33, int W, int H, int16_t *map1,
                                         uint16_t *map2, void *stream);
int svb_remap_u8_fixed(svb_context *ctx, const uint8_t *src, int sH, int sW, int C, size_t spitch, size_t sstride,
                       const int16_t *map1, const uint16_t *map2, uint8_t *dst, int dH, int dW, size_t dpitch,
                       size_t dstride, int B, void *stream);

/* ---- 3-D ---- */
/* dtype: 0=u8 1=s16 2=s32 3=f32 (raw value used, no /16 — as cv2).  Q: HOST double[16]. xyz: f32 [B][H][W][3].
 * min_ws: device workspace of B doubles (only read/written when handle_missing). */
int svb_reproject_to_3d(svb_context *ctx, const void *disp, int dtype, int H, int W, const double *Q,
                        int handle_missing, float *xyz, int B, double *min_ws, void *stream);
/* disp16: int16 [B][H][W] (x16, as compute returns); boxes: f32 [nboxes][4] xyxy; box_image: int32 [nboxes]
 * (frame index of each box, NULL = all 0); out: f32 [nboxes]: median Z over valid pixels, NaN if none.
 */
int svb_box_median_depth(svb_context *ctx, const int16_t *disp16, int H, int W, size_t stride_elems,
                         const double *Q, int invalid16, const float *boxes, const int32_t *box_image, int nboxes,
                         float *out, void *stream);

/* ---- sparse 3-D path + ingest ---- */
/* cv2.undistortPoints(pts, K, dist, R, P) — ball_drop.py:239-240, 3dpose.py:538-549, checkerboard_distance.py:83-86.
 * src/dst: device [n][2], dtype 3 = f32, 4 = f64; K[9], dist[ndist], R[9] (NULL = I), P33[9] = P[:, :3] (NULL = I):
 * HOST doubles.  Five fixed-point iterations in f64, as cv2's default criteria. */
int svb_undistort_points(svb_context *ctx, const void *src, void *dst, int dtype, int n, const double *K,
                         const double *dist, int ndist, const double *R, const double *P33, void *stream);
/* cv2.triangulatePoints(P1, P2, pts1, pts2) — ball_drop.py:254, 3dpose.py:553, checkerboard_distance.py:95.
 * P1/P2: HOST double[12] (3x4 row-major); pts: device, layout 0 = [2][n] (cv2's 2xN), 1 = [n][2];
 * out: device [4][n] homogeneous points (unit norm, sign arbitrary — callers divide by w). */
int svb_triangulate_points(svb_context *ctx, const double *P1, const double *P2, const void *pts1, const void *pts2,
                           int dtype, int n, int layout, void *out4xn, void *stream);
/* cv2.cvtColor(img, COLOR_BGR2GRAY), u8: (B*3735 + G*19235 + R*9798 + 16384) >> 15.  src [B][H][spitch] BGR
 * interleaved, dst [B][H][dpitch]. */
int svb_bgr2gray_u8(svb_context *ctx, const uint8_t *src, uint8_t *dst, int H, int W, size_t spitch, size_t sstride,
                    size_t dpitch, size_t dstride, int B, void *stream);

/* ---- measurement helper (bench.py): device peak of packed s16x2 ALU instructions (VIADDMNMX / VIMNMX chains on
 * all SMs), in thread-level packed instructions per second; x2 = 16-bit lane operations per second.  This is the
 * roofline denominator of the aggregation stage. */
int svb_measure_packed_alu_peak(svb_context *ctx, double *packed_instr_per_s, double *sm_clock_mhz, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SVB_H */
