/*
 * svb.h -- C ABI of libsvb.so, the B200-native (sm_100a) stereo depth engine.
 *
 * This is the drop-in boundary for the per-frame hot path of the reference
 * (AlexGr5/StereoVision: depth_map.py, fused_depth_map.py).  The reference has no
 * FFI of its own: it calls the OpenCV Python bindings (cv2.*).  Each entry point
 * below therefore names the cv2 call / reference call site it replaces; the
 * ctypes binding a maintainer would add is shown in INTEGRATION.md and lives in
 * stereovision_b200/_lib.py.
 *
 * Conventions
 *   - plain C types only; every function returns 0 on success or a negative
 *     svb_status; svb_last_error() gives the thread-local message.
 *   - "*_dev" pointers are DEVICE pointers (cudaMalloc'ed / torch CUDA tensors);
 *     work is enqueued on `stream` (a cudaStream_t passed as void*, NULL = legacy
 *     default stream) and the call does not synchronise.
 *   - "*_host" entry points take HOST buffers, copy in, compute, copy out and
 *     synchronise: this is what a cv2-shaped caller (NumPy in / NumPy out) binds.
 *   - images are dense row-major; a leading batch dimension B is always present
 *     (B = 1 for a single pair).
 *   - there is NO CPU fallback: without a CUDA device every call fails with
 *     SVB_ERR_CUDA.
 */
#ifndef SVB_H
#define SVB_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    SVB_OK = 0,
    SVB_ERR_PARAM = -1,       /* what cv2 would reject with cv2.error (bad sizes, D<=0 ...) */
    SVB_ERR_UNSUPPORTED = -2, /* valid for cv2 but outside this engine (stated in the message) */
    SVB_ERR_CUDA = -3,        /* CUDA runtime error, no device */
    SVB_ERR_NOMEM = -4
} svb_status;

/* cv2.STEREO_SGBM_MODE_* */
enum { SVB_MODE_SGBM = 0, SVB_MODE_HH = 1, SVB_MODE_SGBM_3WAY = 2, SVB_MODE_HH4 = 3 };

const char* svb_last_error(void);
int svb_version(void);
/* Select + initialise the CUDA device for the calling thread (cudaSetDevice). */
int svb_init(int device);
int svb_device_count(void);
/* Copies between PAGEABLE host arrays and device memory for the cv2-shaped calls that take NumPy arrays the way the
 * reference's call sites pass them (depth_map.py:815-826, 909; fused_depth_map.py:480-491).  Replaces the implicit
 * host<->device traffic of handing such an array to a device library: a few host threads move the data through two
 * pinned staging chunks while the DMA engine moves the previous chunk (a plain cudaMemcpy from pageable memory is a
 * single-threaded staged copy in the driver).  h2d returns once the host array has been read (the DMA of the last
 * chunk may still be in flight on `stream`); d2h returns when dst_host is complete.  Pinned or registered host
 * memory is copied with one cudaMemcpyAsync.  One staging area per process: calls serialise. */
int svb_copy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream);
int svb_copy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream);

/* ---- cv2.StereoSGBM_create(...)  (depth_map.py:894-906, fused_depth_map.py:988-1000) */
typedef struct {
    int minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff;
    int preFilterCap, uniquenessRatio, speckleWindowSize, speckleRange, mode;
} svb_sgbm_params;
typedef struct svb_sgbm svb_sgbm_t;

int svb_sgbm_create(const svb_sgbm_params* p, svb_sgbm_t** out);
void svb_sgbm_destroy(svb_sgbm_t* h);
/* Bytes of device workspace the handle will hold for this shape (grow-only, owned by the handle). */
size_t svb_sgbm_workspace_bytes(const svb_sgbm_t* h, int B, int H, int W);
/* Cap on the workspace; batches are processed in chunks that fit (0 = 60 % of the free device memory). */
int svb_sgbm_set_workspace_limit(svb_sgbm_t* h, size_t bytes);
/* stereo.compute(left, right) -> int16 disparity x16 (depth_map.py:909, fused_depth_map.py:1004).
 * left/right: u8 [B,H,W]; disp: s16 [B,H,W]. */
int svb_sgbm_compute_dev(svb_sgbm_t* h, const uint8_t* left_dev, const uint8_t* right_dev,
                         int B, int H, int W, int16_t* disp_dev, void* stream);
int svb_sgbm_compute_host(svb_sgbm_t* h, const uint8_t* left_host, const uint8_t* right_host,
                          int B, int H, int W, int16_t* disp_host);
/* The same with cn = 1 (u8 [B,H,W]) or cn = 3 (interleaved u8 [B,H,W,3]: cv2.StereoSGBM accepts colour
 * images and sums the Birchfield-Tomasi cost over the channels; the reference converts to gray first,
 * depth_map.py:871-879, so this is SURVEY 8(f) N2 surface completeness). */
int svb_sgbm_compute_cn_dev(svb_sgbm_t* h, const uint8_t* left_dev, const uint8_t* right_dev,
                            int B, int H, int W, int cn, int16_t* disp_dev, void* stream);
int svb_sgbm_compute_cn_host(svb_sgbm_t* h, const uint8_t* left_host, const uint8_t* right_host,
                             int B, int H, int W, int cn, int16_t* disp_host);
/* The device call on a PITCHED input batch: row_pitch / img_pitch are the bytes between consecutive rows / images of
 * both inputs (a crop of a larger frame, a padded capture buffer; img_pitch 0 = H * row_pitch, both 0 = dense).
 * What cv2 does with a non-contiguous Mat view; the output stays dense. */
int svb_sgbm_compute_pitched_dev(svb_sgbm_t* h, const uint8_t* left_dev, const uint8_t* right_dev, int B, int H, int W,
                                 int cn, ptrdiff_t row_pitch, ptrdiff_t img_pitch, int16_t* disp_dev, void* stream);
/* Test hook: copy internal stages of the LAST compute (chunk 0, image 0) to host buffers
 * (any may be NULL).  C,S: u16 [rows,width1,D] (3WAY: kept rows only are meaningful),
 * raw: s16 [H,W] disparity before median/speckle.  The aggregated volume S only exists in
 * HBM after svb_sgbm_set_debug(h, 1) (normally the last sweep consumes it in registers). */
int svb_sgbm_set_debug(svb_sgbm_t* h, int enable);
int svb_sgbm_debug_stages(svb_sgbm_t* h, uint16_t* C_host, uint16_t* S_host, int16_t* raw_host);
/* Number of kernel launches issued by the last compute call (bench.py's gpu_launches). */
int svb_sgbm_last_launches(const svb_sgbm_t* h);
/* The aggregation kernels hand rows from CTA to CTA through device-side progress counters; a wait that runs into
 * its spin limit (only a device reset or a foreign writer can cause that) raises a sticky flag instead of hanging.
 * svb_sgbm_check waits for the handle's last compute and returns SVB_ERR_CUDA if the flag was raised since the
 * previous check (those disparities are invalid).  The *_host entry points call it before they return; a *_dev
 * caller calls it at its own synchronisation points -- and the NEXT *_dev compute on the handle reports a flag
 * left by a finished predecessor without synchronising.  Successive computes of one handle are ordered on the
 * device even when they arrive on different streams (they share one workspace). */
int svb_sgbm_check(svb_sgbm_t* h);
/* Test hook: raise that flag by hand. */
int svb_sgbm_debug_trip_error(svb_sgbm_t* h);
/* Test hook: the first sweep at numDisparities <= 128 has two kernels (rows-per-warp k_wave for launches that fill
 * the GPU, warp-per-row k_sweep for a live frame); 1 / 0 forces one of them for every batch size, -1 = automatic. */
int svb_sgbm_debug_force_wave(svb_sgbm_t* h, int mode);

/* Per-stage device timing (CUDA events on the launching stream) for bench.py's roofline.
 * enable=1 starts/reset recording, 0 stops.  get_profile synchronises and returns the number
 * of stages filled: names (max_stages x 32 chars), accumulated ms and launch counts. */
int svb_sgbm_set_profiling(svb_sgbm_t* h, int enable);
int svb_sgbm_get_profile(svb_sgbm_t* h, int max_stages, char* names, float* ms, int* launches);

/* ---- cv2.StereoBM_create(numDisparities, blockSize) + setters (BASELINE config 2) */
typedef struct {
    int minDisparity, numDisparities, blockSize, preFilterCap, textureThreshold;
    int uniquenessRatio, speckleWindowSize, speckleRange, disp12MaxDiff;
    int preFilterType;   /* cv2.STEREO_BM_PREFILTER_NORMALIZED_RESPONSE = 0, _XSOBEL = 1 (cv2's default) */
    int preFilterSize;   /* window of the normalized-response prefilter: odd, 5..255 (cv2's default 9) */
} svb_bm_params;
typedef struct svb_bm svb_bm_t;
int svb_bm_create(const svb_bm_params* p, svb_bm_t** out);
void svb_bm_destroy(svb_bm_t* h);
int svb_bm_compute_dev(svb_bm_t* h, const uint8_t* left_dev, const uint8_t* right_dev,
                       int B, int H, int W, int16_t* disp_dev, void* stream);
int svb_bm_compute_host(svb_bm_t* h, const uint8_t* left_host, const uint8_t* right_host,
                        int B, int H, int W, int16_t* disp_host);
int svb_bm_last_launches(const svb_bm_t* h);
/* cv2.validateDisparity(disparity, cost, minDisparity, numberOfDisparities, disp12MaxDiff) (calib3d; what
 * StereoBM runs when disp12MaxDiff >= 0).  disp s16 [B,H,W] x16 fixed point, modified in place; cost s16
 * (cost_is_s32 == 0) or s32 [B,H,W]; scratch: B*H*W*10 bytes of device memory. */
int svb_validate_disparity_dev(int16_t* disp_dev, const void* cost_dev, int cost_is_s32, int B, int H, int W,
                               int minDisparity, int numberOfDisparities, int disp12MaxDiff,
                               void* scratch_dev, void* stream);

/* ---- cv2.remap(src, map1 CV_16SC2, map2 CV_16UC1, INTER_LINEAR), BORDER_CONSTANT 0
 *      (depth_map.py:815-826, fused_depth_map.py:480-491); to_gray!=0 fuses
 *      cv2.cvtColor(BGR2GRAY) (depth_map.py:871-879) into the epilogue (C must be 3).
 *      src u8 [B,Hs,Ws,C]; map1 s16 [Ho,Wo,2]; map2 u16 [Ho,Wo]; dst u8 [B,Ho,Wo,C] or [B,Ho,Wo]. */
int svb_remap_u8_dev(const uint8_t* src_dev, int B, int Hs, int Ws, int C,
                     const int16_t* map1_dev, const uint16_t* map2_dev, int Ho, int Wo,
                     uint8_t* dst_dev, int to_gray, void* stream);
/* ---- cv2.resize(src, (Wd, Hd), interpolation=cv2.INTER_LINEAR) (fused_depth_map.py:473-476: both camera frames on
 *      every processed frame, PROCESSING_SCALE 0.33; depth_map.py:776, 809-811; fused_depth_map.py:1214-1217: a
 *      float32 mono map to the stereo shape).  u8: C = 1 or 3 interleaved channels, src [B,Hs,Ws,C] -> dst
 *      [B,Hd,Wd,C], bit-exact with cv2 4.13 (11-bit fixed-point weights).  f32: one channel, 1e-6 relative. */
int svb_resize_u8_dev(const uint8_t* src_dev, int B, int Hs, int Ws, int C, uint8_t* dst_dev, int Hd, int Wd,
                      void* stream);
int svb_resize_f32_dev(const float* src_dev, int B, int Hs, int Ws, float* dst_dev, int Hd, int Wd, void* stream);
/* cv2.cvtColor(img, COLOR_BGR2GRAY): u8 [n,3] -> u8 [n] */
int svb_bgr2gray_dev(const uint8_t* bgr_dev, size_t npix, uint8_t* gray_dev, void* stream);

/* ---- cv2.medianBlur: 3x3 on s16 (inside StereoSGBM), 5x5 on f32 (depth_map.py:912) */
int svb_median3_s16_dev(const int16_t* src_dev, int B, int H, int W, int16_t* dst_dev, void* stream);
int svb_median5_f32_dev(const float* src_dev, int B, int H, int W, float* dst_dev, void* stream);

/* ---- cv2.filterSpeckles(img, newVal, maxSpeckleSize, maxDiff) in place on s16 [B,H,W].
 *      labels_dev: caller scratch of B*H*W int32 (2 arrays: 2*B*H*W*4 bytes). */
int svb_filter_speckles_s16_dev(int16_t* img_dev, int B, int H, int W, int newVal,
                                int maxSpeckleSize, int maxDiff, int32_t* scratch_dev, void* stream);
/* The same filter for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments minus the scratch: labels and counts live in a static device array inside libsvb.so
 * sized for B*H*W <= SVB_SPECKLE_STATIC_MAX_PIXELS (16 MiB, held by every GPU that loads the library until the
 * process exits); larger inputs are rejected with SVB_ERR_UNSUPPORTED.  The call performs no cudaMalloc / cudaFree
 * and calls no allocating routine; the first call on each device looks the array's device address up once
 * (cudaGetSymbolAddress; make that first call outside a graph capture).  Every call shares the one array:
 * calls that overlap on the GPU (different streams or threads) race -- serialise them. */
#define SVB_SPECKLE_STATIC_MAX_PIXELS 2097152   /* 2 Mi pixels: one 1920x1080 frame, or e.g. six 640x480 frames */
int svb_filter_speckles_s16_static(int16_t* img_dev, int B, int H, int W, int newVal,
                                   int maxSpeckleSize, int maxDiff, void* stream);

/* ---- depth_map.py:909-933 fused: disparity = s16/16 -> medianBlur 5 -> depth = fx*B/(d+1e-6),
 *      clip [min_depth,max_depth], zero where invalid.  disp16 s16 [B,H,W] -> disparity f32, depth f32. */
int svb_disp_to_depth_dev(const int16_t* disp16_dev, int B, int H, int W, double fx, double baseline,
                          double min_disp, double min_depth, double max_depth,
                          float* disparity_dev, float* depth_dev, void* stream);
/* ---- cv2.reprojectImageTo3D(disp, Q, handleMissingValues): disp f32 or s16 [B,H,W] -> f32 [B,H,W,3].
 *      Q_host16: 4x4 row-major doubles on the HOST; minscratch_dev: B floats (only read when
 *      handleMissingValues != 0, may be NULL otherwise). */
int svb_reproject3d_dev(const void* disp_dev, int is_s16, int B, int H, int W, const double* Q_host16,
                        int handleMissingValues, float* minscratch_dev, float* out_dev, void* stream);

/* ---- cv2.initUndistortRectifyMap(K, dist, R, P, size, CV_16SC2) (depth_map.py:633-638,
 *      fused_depth_map.py:404-409).  Host pointers to row-major doubles: K9 3x3, dist ndist
 *      (0, 4, 5, 8, 12 or 14 with zero tilt) coefficients, R9 3x3 or NULL (identity), P 3 x pcols
 *      (pcols 3 or 4: cv2 uses its left 3x3 block).  Device outputs: map1 s16 [H,W,2], map2 u16 [H,W]. */
int svb_init_undistort_rectify_map_dev(const double* K9, const double* dist, int ndist, const double* R9,
                                       const double* P, int pcols, int W, int H, int16_t* map1_dev,
                                       uint16_t* map2_dev, void* stream);

/* ==== Fusion chain of fused_depth_map.py.  All maps are f32 [B,H,W] on the
 *      device; no call synchronises the stream.  Scratch is caller-owned (sizes from the *_scratch_bytes
 *      functions). */

/* fused_depth_map.py:1010-1029: norm = uint8((clip(d, min, min+num-1) - min) / num * 255) as f32,
 * conf = 1 where min+1 < d < min+num-1 else 0. */
int svb_normalize_disparity_dev(const float* disp_dev, int B, int H, int W, int min_disp, int num_disp,
                                float* norm_dev, float* conf_dev, void* stream);

/* Scratch for svb_calibrate_midas_dev / svb_normalize_to_stereo_range_dev / svb_percentiles_f32_dev. */
size_t svb_range_scratch_bytes(int B);
/* calibrate_midas_to_stereo (fused_depth_map.py:1169-1257): np.percentile 10/90 over conf > 0.7
 * (5/95 over everything when fewer than 100 reliable pixels), then the affine / min-max mapping. */
int svb_calibrate_midas_dev(const float* midas_dev, const float* stereo_disp_dev, const float* stereo_conf_dev,
                            int B, int H, int W, float* out_dev, void* scratch_dev, size_t scratch_bytes, void* stream);
/* normalize_to_stereo_range (fused_depth_map.py:1503-1554). */
int svb_normalize_to_stereo_range_dev(const float* depth_dev, const float* stereo_disp_dev, int B, int H, int W,
                                      float* out_dev, void* scratch_dev, size_t scratch_bytes, void* stream);
/* np.percentile(vals[mask], 5 / 95) (exact order statistics + NumPy's linear interpolation).
 * mask_mode 0: all; 1: mask > mask_thr; 2: vals > 0; 3: mask > mask_thr when at least 100 such
 * pixels (then q = 10 / 90 in float64, as np.percentile(v, [10, 90])), else all.
 * out_dev: double [B][2]; count_dev: unsigned [B] selected elements (may be NULL). */
int svb_percentiles_f32_dev(const float* vals_dev, const float* mask_dev, float mask_thr, int mask_mode, int B,
                            size_t n, double* out_dev, unsigned* count_dev, void* scratch_dev, size_t scratch_bytes,
                            void* stream);

/* fuse_depth_maps (fused_depth_map.py:1560-1698); the module-global fusion_params / FUSION_SMOOTHING /
 * FUSION_METHODS dictionaries (fused_depth_map.py:80-116) become this struct. */
typedef struct svb_fuse_params {
    double stereo_weight, midas_fill_weight, flow_fill_weight, stereo_conf_threshold, flow_hole_threshold;
    double bilateral_sigma_color, bilateral_sigma_space;
    int blend_ksize, bilateral_d, use_midas_fill, use_flow_fill;
} svb_fuse_params;
size_t svb_fuse_scratch_bytes(int B, int H, int W);
/* stereo_dev / midas_dev / flow_dev may be NULL (method absent or disabled; flow_dev NULL when the
 * camera is not moving); conf_dev NULL = confidence 1.  out_u8_dev: the fused map; out_f32_dev
 * (may be NULL): the value before the uint8 cast. */
int svb_fuse_depth_maps_dev(const float* stereo_dev, const float* conf_dev, const float* midas_dev,
                            const float* flow_dev, int B, int H, int W, const svb_fuse_params* p,
                            uint8_t* out_u8_dev, float* out_f32_dev, void* scratch_dev, size_t scratch_bytes,
                            void* stream);

/* detect_camera_occlusion statistics (fused_depth_map.py:170-240) of u8 gray [B,H,W]:
 * stats_dev double [B][5] = {mean block std, low-variance block ratio, global std, entropy, mean}. */
size_t svb_occlusion_scratch_bytes(int B, int H, int W, int block);
int svb_occlusion_stats_u8_dev(const uint8_t* gray_dev, int B, int H, int W, int block, double* stats_dev,
                               void* scratch_dev, size_t scratch_bytes, void* stream);

/* ==== SURVEY 8(f) N3: the data-parallel code around the mono estimators of fused_depth_map.py.
 *      Maps are [B,H,W] on the device; scratch from the *_scratch_bytes functions; nothing synchronises. */

/* cv2.bilateralFilter(src f32, d, sigmaColor, sigmaSpace), BORDER_REFLECT_101 (fused_depth_map.py:1400,
 * 1409, 1690); d odd, <= 15. */
size_t svb_bilateral_scratch_bytes(int B);
int svb_bilateral_f32_dev(const float* src_dev, int B, int H, int W, int d, double sigma_color,
                          double sigma_space, float* dst_dev, void* scratch_dev, size_t scratch_bytes,
                          void* stream);

/* MiDaS post-processing, fused_depth_map.py:1134-1145: torch.nn.functional.interpolate(pred [B,Hi,Wi],
 * size=(Ho,Wo), mode="bicubic", align_corners=False), then (x - min) / (max - min + 1e-8) * 255 per image
 * (zeros when max - min < 1e-6). */
size_t svb_midas_scratch_bytes(int B);
int svb_midas_upsample_normalize_dev(const float* pred_dev, int B, int Hi, int Wi, float* out_dev, int Ho,
                                     int Wo, void* scratch_dev, size_t scratch_bytes, void* stream);

/* Contrast confidence, fused_depth_map.py:1150-1156: cv2.normalize(cv2.absdiff(gray, cv2.GaussianBlur(gray,
 * (5,5), 0)), None, 0, 1, cv2.NORM_MINMAX) on u8 gray [B,H,W] -> u8.  Scratch: svb_midas_scratch_bytes. */
int svb_contrast_confidence_u8_dev(const uint8_t* gray_dev, int B, int H, int W, uint8_t* conf_dev,
                                   void* scratch_dev, size_t scratch_bytes, void* stream);

/* Flow depth, fused_depth_map.py:1386-1398 + 1491-1497: expected flow = cv2.perspectiveTransform(pixel grid,
 * H) - grid, residual = flow - expected, current = 1 / (hypot(residual) + 0.5); when stable_dev is given it
 * becomes alpha * stable + (1 - alpha) * current (or a copy of current when stable_valid == 0).
 * flow_dev f32 [B,H,W,2]; homography9 HOST doubles [B][9]. */
int svb_flow_residual_depth_dev(const float* flow_dev, int B, int H, int W, const double* homography9,
                                float* current_dev, float* stable_dev, int stable_valid, double alpha,
                                void* stream);

/* ==== SURVEY 8(f) N4: display-side reductions, written to a headless device buffer (no window system). */

/* create_depth_map's depth in float64 -- what NumPy computes when fx is read from a float64 camera matrix
 * (depth_map.py:915-933).  Same as svb_disp_to_depth_dev otherwise. */
int svb_disp_to_depth_f64_dev(const int16_t* disp16_dev, int B, int H, int W, double fx, double baseline,
                              double min_disp, double min_depth, double max_depth, float* disparity_dev,
                              double* depth_dev, void* stream);

/* Normalise to an 8-bit index and look it up in cv2.COLORMAP_JET (2) / cv2.COLORMAP_TURBO (20)
 * (cv2.applyColorMap).  src_kind selects the reference expression that produces the index:
 *   0  u8 source as is                                    fused_depth_map.py:1016, 1700
 *   1  np.clip(f32, 0, 255).astype(u8)                    fused_depth_map.py:3026, 3052
 *   2  f32 disparity -> depth_clipped -> ((z - min) / (max - min) * 255).astype(u8)   depth_map.py:926-937
 *   3  the same in float64 (float64 fx)
 *   4  f32 depth, zeros where depth <= 0, clipped, same scaling   depth_map.py:1187-1192
 *   5  the same on f64 depth
 *   6  cv2.normalize(f32, None, lo, hi, NORM_MINMAX).astype(u8) per image   depth_map.py:1207-1208
 * params5 (host) = {fx * baseline, min_depth, max_depth, lo, hi}; kinds 0/1 accept NULL.  Outputs (either may be
 * NULL): index u8 [B,npix], bgr u8 [B,npix,3].  Scratch (kind 6 only): svb_display_scratch_bytes. */
size_t svb_display_scratch_bytes(int B);
int svb_colormap_dev(const void* src_dev, int src_kind, int B, size_t npix, int colormap, const double* params5,
                     uint8_t* index_dev, uint8_t* bgr_dev, void* scratch_dev, size_t scratch_bytes, void* stream);

/* Legend range of the fused map (fused_depth_map.py:1997-2004): per image {min, max} over the non-zero pixels of
 * u8 [B,npix] as int32 [B][2]; {0, 0} when every pixel is zero. */
int svb_range_nonzero_u8_dev(const uint8_t* src_dev, int B, size_t npix, int* minmax_dev, void* scratch_dev,
                             size_t scratch_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SVB_H */
