/* tmat_b200.h -- C ABI of libtmat_b200.so: the B200 (sm_100a) implementation of the
 * microvessel-branching segmentation hot path of fogg-lab/tissue-model-analysis-tools.
 *
 * The reference has no FFI on this path (plain Python functions on numpy arrays); these
 * entry points are what a ctypes binding inside the reference package would call.  Each
 * one cites the reference function it replaces (paths relative to the reference repo,
 * package dir fl_tissue_model_tools/ abbreviated to its file name).
 *
 * Conventions
 *   - every function returns 0 (TMAT_OK) or a negative TMAT_ERR_* code; the message for the
 *     calling thread is available from tmat_last_error().  No exception crosses the boundary.
 *   - pointers named *_dev are device pointers on the context's device; pointers named
 *     *_host are host pointers (pageable or pinned).  The caller owns every buffer.
 *   - images are row-major (C order), stacks are (Z, H, W) row-major, like the numpy
 *     arrays the reference passes around (helper.py:90).
 *   - device-pointer entry points are asynchronous on `stream` (a cudaStream_t passed as
 *     void*; NULL = the legacy default stream); *_host entry points copy in, run, copy out
 *     and synchronise before returning.
 *   - a context owns scratch device memory and may be used by one thread at a time; create
 *     one context per thread (compute_cell_area.py:108-113 fans out over dask threads).
 */
#ifndef TMAT_B200_H
#define TMAT_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TMAT_OK                  0
#define TMAT_ERR_INVALID        -1   /* bad argument */
#define TMAT_ERR_CUDA           -2   /* a CUDA call failed; see tmat_last_error() */
#define TMAT_ERR_UNSUPPORTED    -3   /* valid in the reference, not implemented here */
#define TMAT_ERR_NOMEM          -4
#define TMAT_ERR_NO_DEVICE      -5   /* no sm_100 device: there is no CPU fallback */

/* element types of images / stacks */
#define TMAT_U8   0
#define TMAT_U16  1
#define TMAT_F32  2
#define TMAT_F64  3

/* Z-projection methods (scripts/compute_zproj.py:26-32) */
#define TMAT_ZPROJ_MIN 0   /* zstacks.py:237 proj_min  -> same dtype  */
#define TMAT_ZPROJ_MAX 1   /* zstacks.py:221 proj_max  -> same dtype  */
#define TMAT_ZPROJ_AVG 2   /* zstacks.py:192 proj_avg  -> float64     */
#define TMAT_ZPROJ_MED 3   /* zstacks.py:206 proj_med  -> float64     */
#define TMAT_ZPROJ_FS  4   /* zstacks.py:153 proj_focus_stacking -> same dtype */

typedef struct tmat_ctx  tmat_ctx;
typedef struct tmat_unet tmat_unet;

/* ---- library / context ------------------------------------------------------------ */
int         tmat_version(void);
const char* tmat_last_error(void);
int  tmat_ctx_create(int device, tmat_ctx** out);
int  tmat_ctx_destroy(tmat_ctx* ctx);
int  tmat_ctx_sync(tmat_ctx* ctx);
/* number of kernels of this library launched through `ctx` since creation (bench.py's
 * gpu_launches claim is read from here). */
int  tmat_ctx_launch_count(tmat_ctx* ctx, int64_t* out);
/* Page-locked host buffer for *_host arguments (cudaHostAlloc): a copy to or from it runs at full PCIe
 * speed and is truly asynchronous.  The reference allocates its result with numpy (pageable,
 * smooth_tiled_predictions.py:244-262); the Python mirror hands out results backed by these buffers. */
int  tmat_host_alloc(int64_t bytes, void** out);
int  tmat_host_free(void* p);

/* Per-launch CUDA-event timing of the tensor-core convolution kernel (bench.py's roofline leg).
 * collect: out3[0] = summed kernel time in ms since the last collect, out3[1] = launches,
 * out3[2] = FLOPs executed by them; synchronises the device. */
int  tmat_ctx_profile_enable(tmat_ctx* ctx, int on);
int  tmat_ctx_profile_collect(tmat_ctx* ctx, double* out3);

/* ---- stage 1: Z projection  (zstacks.py:134-249) ------------------------------------ */
/* stack: (Z,H,W) of dtype TMAT_U8|TMAT_U16, or TMAT_F32 for MIN/MAX/AVG/MED (numpy semantics:
 * np.mean/np.median of a float32 stack stay float32; NaNs are not handled).  out: (H,W) of the
 * stack dtype for MIN/MAX/FS; for AVG/MED float64 (integer stacks) or float32 (float32 stacks).
 * ksize must be 5 for FS (the reference's default kernel_size, zstacks.py:153). */
int tmat_zproj_dev(tmat_ctx* ctx, const void* stack_dev, int dtype, int Z, int H, int W,
                   int method, int ksize, void* out_dev, void* stream);
int tmat_zproj_host(tmat_ctx* ctx, const void* stack_host, int dtype, int Z, int H, int W,
                    int method, int ksize, void* out_host);

/* ---- stage 2: smooth tiled prediction  (smooth_tiled_predictions.py:220-267) --------- */
/* Geometry helper: padded shape, tiles per orientation (rotated frame), aug.
 * geom[0]=aug geom[1]=PM geom[2]=PN geom[3]=step geom[4..5]=tiles (rows,cols) of even
 * orientations, geom[6..7]= of odd orientations. */
int tmat_stp_geometry(int H, int W, int ws, int subdiv, int* geom8);

/* Generic pred_func path, step 1: tiles of orientation g (0..7, the order of
 * _rotate_mirror_do, smooth_tiled_predictions.py:103-112) of the image padded by
 * aug with its global minimum (:68-79), in raster order of the rotated frame (:157-163).
 * tiles_dev: (n_tiles, ws, ws) float32. */
int tmat_stp_gather_dev(tmat_ctx* ctx, const float* img_dev, int H, int W, int ws, int subdiv,
                        int orient, float* tiles_dev, void* stream);
/* Generic path, step 2: window multiply (:185), overlap-add (:206-217), /subdiv^2, D4 undo
 * (:125-132), mean of 8 (:133), unpad + crop (:263-265).  preds_dev: the predictions of all 8
 * orientations concatenated in orientation order, (n_total, ws, ws) float32.  window1d_host:
 * ws float64 values of _spline_window (:26-41); NULL = computed by the library.
 * out_dev: (H,W) float64.  Bit-identical to the reference given identical predictions. */
int tmat_stp_blend_dev(tmat_ctx* ctx, const float* preds_dev, int H, int W, int ws, int subdiv,
                       const double* window1d_host, double* out_dev, void* stream);
int tmat_stp_gather_host(tmat_ctx* ctx, const float* img_host, int H, int W, int ws, int subdiv,
                         int orient, float* tiles_host);
int tmat_stp_blend_host(tmat_ctx* ctx, const float* preds_host, int H, int W, int ws, int subdiv,
                        const double* window1d_host, double* out_host);
/* _spline_window(ws, power=2) (:26-41) */
int tmat_spline_window(int ws, double* out_host);

/* UNet-Xception (models.py:85-171), forward only.  Weights are passed in Keras array layouts,
 * float32, as parallel arrays of names/pointers (names as in oracle/unet.py: "stem.kernel",
 * "down0.sep1.dw", "up2.ct1.kernel", "head.bias", ...). */
typedef struct {
  int patch;            /* window size, multiple of 32                         */
  int channels;         /* input channels: 1 (models.py:600 default)            */
  int filter_counts[4]; /* ascending, e.g. 64,128,256,512 (unet_patch_segmentor_1.json) */
  int impl;             /* 0 = auto (tcgen05 when filter counts are multiples of 64),
                           1 = force CUDA-core kernels (cross-check path), 2 = force tcgen05 */
} tmat_unet_desc;
int tmat_unet_load(tmat_ctx* ctx, const tmat_unet_desc* desc, int n_arrays,
                   const char* const* names, const float* const* arrays_host,
                   const int64_t* sizes, tmat_unet** out);
int tmat_unet_free(tmat_unet* net);
/* model.predict on a batch of patches: (n, ws, ws) float32 -> (n, ws, ws) float32 sigmoid
 * probabilities (models.py:166).  logits!=0 returns pre-sigmoid values. */
int tmat_unet_forward_dev(tmat_ctx* ctx, tmat_unet* net, const float* patches_dev, int n,
                          float* probs_dev, int logits, void* stream);
int tmat_unet_forward_host(tmat_ctx* ctx, tmat_unet* net, const float* patches_host, int n,
                           float* probs_host, int logits);
/* Fused fast path of predict_img_with_smooth_windowing with pred_func = this UNet
 * (models.py:640): D4 folded into the stem's tile read, all orientations in one canvas. */
int tmat_stp_predict_dev(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W,
                         int subdiv, double* out_dev, void* stream);
int tmat_stp_predict_host(tmat_ctx* ctx, tmat_unet* net, const float* img_host, int H, int W,
                          int subdiv, double* out_host);
/* Row-strip form of the fused path: output rows [row0, row1) only (out: (row1-row0) x W float64).  The UNet
 * runs on the patches whose window touches those rows, so N GPUs that each take H/N rows of ONE large image
 * (BASELINE config "large-well tiles", SURVEY 8(e) owner-computes row strips) need no cross-GPU reduction;
 * the strips are bit-identical to the rows of tmat_stp_predict_*.  n_patches (optional): patches evaluated. */
int tmat_stp_predict_rows_dev(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W, int subdiv,
                              int row0, int row1, double* out_rows_dev, int* n_patches, void* stream);
int tmat_stp_predict_rows_host(tmat_ctx* ctx, tmat_unet* net, const float* img_host, int H, int W, int subdiv,
                               int row0, int row1, double* out_rows_host, int* n_patches);
/* Host geometry of the row-strip form: ids (orientation * tiles_per_orientation + tile index in the rotated
 * frame) of the patches whose window touches output rows [row0, row1).  list may be NULL to query the count. */
int tmat_stp_rows_patches(int H, int W, int ws, int subdiv, int row0, int row1, int* list, int capacity,
                          int* n_out);
/* FLOPs (2*MAC of every convolution as written in models.py:110-171) of one patch */
int tmat_unet_flops_per_patch(const tmat_unet_desc* desc, double* flops);

/* ---- stage 3: thresholds and the Gaussian pre-filter -------------------------------- */
/* scripts/compute_branches.py:334  seg_mask = pred > 0.5   (mask: uint8 0/1) */
int tmat_threshold_gt_f64_dev(tmat_ctx* ctx, const double* x_dev, int64_t n, double thr,
                              uint8_t* mask_dev, void* stream);
/* preprocessing.py:13-23 apply_mask (out = img where mask!=0 else 0), float32 */
int tmat_apply_mask_f32_dev(tmat_ctx* ctx, const float* img_dev, const uint8_t* mask_dev,
                            int64_t n, float* out_dev, void* stream);
/* skimage.filters.gaussian(image, sigma) as called at well_mask_generation.py:245:
 * separable, mode='nearest', truncate=4; integer input is scaled to [0,1] float64 first and
 * the result is float64; float32 input gives float32.  out dtype: F64 for U8/U16, F32 for F32. */
int tmat_gaussian_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W, double sigma,
                      void* out_dev, void* stream);
int tmat_gaussian_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W, double sigma,
                       void* out_host);
/* preprocessing.py:44-93 exec_threshold: the statistics of GaussianMixture(n_components=2).fit on the
 * selected pixels (float32 vector x, n >= 2).  rand3: the three uniform [0,1) draws scikit-learn's k-means++
 * seeding takes from the caller's RandomState (random_sample() for the first centre, uniform(size=2) for the
 * candidate trials), made by the caller so its stream advances as in the reference.
 * result10: [0] threshold = min(255, mu_fg + sd_coef * sd_fg), [1,2] weights, [3,4] means, [5,6] variances,
 * [7] EM iterations (negative: not converged in 100), [8] k-means iterations, [9] mean log-likelihood. */
int tmat_gmm2_threshold_dev(tmat_ctx* ctx, const float* x_dev, int64_t n, double sd_coef,
                            const double* rand3_host, double* result10_host, void* stream);
int tmat_gmm2_threshold_host(tmat_ctx* ctx, const float* x_host, int64_t n, double sd_coef,
                             const double* rand3_host, double* result10_host);
/* preprocessing.py:91 np.where(masked <= thr, 0, masked), float32 */
int tmat_threshold_le_zero_f32_host(tmat_ctx* ctx, const float* img_host, int64_t n, double thr,
                                    float* out_host);
/* preprocessing.py:165-183 blur(img, blur_itr, k_size, gs): blur_itr x cv2.GaussianBlur((k,k), 0)
 * (fixed kernels for k = 3, 5, 7, BORDER_REFLECT_101), then the reference's final rounding.
 * dtype U8: exact fixed point (8-bit weights, (S + 32768) >> 16), out uint8, final_op ignored.
 * dtype F32: final_op 0 = raw float32, 1 = np.round -> float32 (gs=False), 2 = np.round -> uint8 (gs=True). */
int tmat_blur_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W, int ksize, int iters,
                  int final_op, void* out_dev, void* stream);
int tmat_blur_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W, int ksize, int iters,
                   int final_op, void* out_host);
/* well_mask_generation.py:236-277 auto_threshold_well: gaussian(sigma=1) -> rescale to uint8
 * -> corner medians -> optional invert -> Otsu -> '>=' -> binary_erosion(disk(5)).
 * mask: uint8 0/1.  info (optional, 4 ints): [0]=inverted [1]=otsu threshold [2]=u8 min [3]=u8 max */
int tmat_auto_threshold_well_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W,
                                 uint8_t* mask_dev, int* info, void* stream);
int tmat_auto_threshold_well_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W,
                                  uint8_t* mask_host, int* info);
/* The numpy-only pieces of generate_well_mask (well_mask_generation.py:16-140).
 * get_superellipse_hull(x, y, n, num_iters) :16-91 -- params (num_iters x 6: theta, d, s_a, s_b, c_x, c_y) are the
 * parameter sets as the reference draws them (np.random.rand mapped into the bounds, done by the caller with numpy so
 * that the global seed decides the same draw), cos/sin = np.cos / np.sin of theta (may be NULL for n == 2, where the
 * reference ignores theta), gamma1_sq = gamma(1 + 1/n)**2, gamma2 = gamma(1 + 2/n).  best_index = the first parameter
 * set of smallest area among those with max_j f(x_j, y_j) < 1, or -1 when none encloses the points (the reference's
 * np.argmin of an empty array raises ValueError there). */
int tmat_superellipse_hull_host(tmat_ctx* ctx, const double* x_host, const double* y_host, int n_points, int n,
                                const double* params_host, const double* cos_host, const double* sin_host,
                                int num_iters, double gamma1_sq, double gamma2, int* best_index);
/* gen_superellipse_mask :94-118: out[a][b] = |((xs[a]-c_x) cos - (ys[b]-c_y) sin) / d_sa|**n +
 * |((xs[a]-c_x) sin + (ys[b]-c_y) cos) / d_sb|**n < 1, xs = linspace(-1, 1, n0), ys = linspace(-1, 1, n1),
 * d_sa = d * s_a, d_sb = d * s_b; out uint8 0/1, n0 x n1 (the reference's swapaxes is folded in). */
int tmat_superellipse_mask_host(tmat_ctx* ctx, const double* xs_host, int n0, const double* ys_host, int n1,
                                double c_x, double c_y, double cos_t, double sin_t, double d_sa, double d_sb, int n,
                                uint8_t* out_host);
/* create_convex_hull_mask :121-140 (scipy Delaunay(hull).find_simplex(pixel) >= 0): pixels inside or on the convex
 * polygon verts (n_verts x 2 int32 (row, col), in polygon order, either orientation); exact integer arithmetic. */
int tmat_convex_hull_mask_host(tmat_ctx* ctx, int H, int W, const int32_t* verts_host, int n_verts, uint8_t* out_host);

/* ---- stage 4: binary morphology / component cleanup --------------------------------- */
/* skimage.filters.median(bool mask, disk(2)) (transforms.py:323) */
int tmat_median_disk2_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, uint8_t* out_dev,
                          void* stream);
/* skimage.morphology.binary_erosion(mask, disk(r)) (well_mask_generation.py:275,
 * scripts/compute_branches.py:130); outside the image counts as `border_value`. */
int tmat_erode_disk_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int r,
                        int border_value, uint8_t* out_dev, void* stream);
/* skimage.measure.label(mask, connectivity) (transforms.py:188,196,301,325): connectivity 1
 * (4-neighbour) or 2 (8-neighbour); labels 1..n numbered in raster order of each component's
 * first pixel.  Synchronises (n is returned to the host). */
int tmat_label_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int connectivity,
                   int32_t* labels_dev, int* n_labels, void* stream);
/* regionprops area + the 50-bin border-code histogram behind skimage's perimeter
 * (transforms.py:326-329).  area_host: n int64; perim_hist_host: n*50 int64. */
int tmat_region_stats_dev(tmat_ctx* ctx, const int32_t* labels_dev, int H, int W, int n_labels,
                          int64_t* area_host, int64_t* perim_hist_host, void* stream);
/* skimage.morphology.skeletonize, 2-D (transforms.py:331).  iterations (optional) receives
 * the number of {sub-pass 1, sub-pass 2} rounds executed. */
int tmat_skeletonize_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, uint8_t* out_dev,
                         int* iterations, void* stream);
/* transforms.py:306-361 filter_branch_seg_mask(mask, footprint=disk(2)|None, remove_isolated) */
int tmat_filter_branch_seg_mask_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W,
                                    int use_median, int remove_isolated, uint8_t* out_dev,
                                    void* stream);
int tmat_filter_branch_seg_mask_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W,
                                     int use_median, int remove_isolated, uint8_t* out_host);
/* transforms.py:170-206 remove_small_islands; returns TMAT_ERR_INVALID for non-{0,1} input
 * (the reference raises ValueError, :183-184). */
int tmat_remove_small_islands_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W,
                                   int min_area0, int min_area1, int connectivity0,
                                   int connectivity1, uint8_t* out_host);
/* transforms.py:291-303 regionprops_image(mask, 'area'): float64 image of component areas */
int tmat_area_image_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, double* out_host);
/* the same for the properties the reference's callers use (scripts/compute_branches.py:276-277):
 * prop 0 = 'area', 1 = 'equivalent_diameter_area' (= sqrt(4*area/pi), as skimage.measure.regionprops),
 * 2 = 'eccentricity' (sqrt(1 - l2/l1) of the inertia-tensor eigenvalues, 0 when l1 == 0; moments accumulated exactly
 * in integers -- skimage's float64 dot products differ by rounding, parity is a tolerance and unpinned). */
int tmat_region_prop_image_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int prop,
                                double* out_host);
/* Rank filter of a boolean image under an arbitrary footprint (fh x fw bytes, non-zero = member, both odd;
 * centre = (fh/2, fw/2)), for the footprints the specialised kernels above do not cover:
 * op 0 = skimage.filters.median on a bool image (-> scipy.ndimage.median_filter, mode 'nearest';
 *        transforms.py:323 with a caller-supplied footprint), border_value ignored;
 * op 1 = skimage.morphology.binary_erosion (-> scipy.ndimage.binary_erosion(border_value)). */
int tmat_footprint_filter_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W,
                              const uint8_t* footprint_host, int fh, int fw, int op, int border_value,
                              uint8_t* out_dev, void* stream);
int tmat_footprint_filter_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W,
                               const uint8_t* footprint_host, int fh, int fw, int op, int border_value,
                               uint8_t* out_host);
/* The 256-entry deletability table of the 2-D thinning (csrc/skeleton_lut.inc: derived from the
 * Zhang-Suen conditions; parity with scikit-image 0.22's hard-coded table is UNPINNED).  set
 * replaces it process-wide (entries 0..3; index = neighbour code, bit0 NW,1 N,2 NE,3 E,4 SE,5 S,
 * 6 SW,7 W; 1 = deleted in sub-pass 1, 2 = in sub-pass 2, 3 = either); get reads it back. */
int tmat_skeleton_lut_set(const uint8_t* lut256);
int tmat_skeleton_lut_get(uint8_t* lut256);
/* host variants of the primitives (used by the Python mirror of transforms.py) */
int tmat_median_disk2_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, uint8_t* out_host);
int tmat_erode_disk_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int r,
                         int border_value, uint8_t* out_host);
int tmat_label_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int connectivity,
                    int32_t* labels_host, int* n_labels);
int tmat_skeletonize_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, uint8_t* out_host,
                          int* iterations);

/* ---- centerline weighting handed to the Morse graph ------------------ */
/* scipy.ndimage.distance_transform_edt(mask) (scripts/compute_branches.py:342; also the `dist` that
 * skimage.morphology.medial_axis(return_distance=True) returns, :341): float64 distance of every non-zero pixel to
 * the nearest zero pixel, exact (sqrt of the integer squared distance).  invert != 0: transform of logical_not(mask).
 * An image without a zero pixel gives +inf everywhere (scipy's result is undefined there). */
int tmat_distance_transform_edt_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int invert,
                                    double* out_dev, void* stream);
int tmat_distance_transform_edt_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int invert,
                                     double* out_host);
/* Allocation-free variant of tmat_distance_transform_edt_dev.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same result, bit for bit.  Takes no context: it runs on the calling thread's CURRENT
 * device, on `stream`, and calls no cudaMalloc / cudaFree / allocating routine -- its one scratch plane is a
 * file-scope __device__ array of TMAT_EDT_STATIC_MAX_PIXELS uint32 (16 MiB per GPU that loads the library, held until
 * exit).  H * W beyond that limit is rejected with TMAT_ERR_INVALID (tmat_last_error names the limit).  Every call
 * shares the array: calls that overlap on the GPU (two streams, two threads on one device) race -- serialise them. */
#define TMAT_EDT_STATIC_MAX_PIXELS (1 << 22)
int tmat_distance_transform_edt_static_dev(const uint8_t* mask_dev, int H, int W, int invert,
                                           double* out_dev, void* stream);
/* scripts/compute_branches.py:341-344: out = pred * dist / (dist + centerline_dt), dist = edt(seg_mask),
 * centerline_dt = edt(~skel).  The skeleton is an input: medial_axis breaks ties with an unseeded RNG in the
 * reference and is not reproducible even against itself. */
int tmat_centerline_weight_dev(tmat_ctx* ctx, const double* pred_dev, const uint8_t* mask_dev,
                               const uint8_t* skel_dev, int H, int W, double* out_dev, void* stream);
int tmat_centerline_weight_host(tmat_ctx* ctx, const double* pred_host, const uint8_t* mask_host,
                                const uint8_t* skel_host, int H, int W, double* out_host);

/* ---- preprocessing.py:96-163: gen_circ_mask, dt_blur, sdt_blur (notebook / training helpers) ---- */
/* cv2.distanceTransform(mask, cv2.DIST_L2, 5) (preprocessing.py:132-134, :155-160): float32 5x5 chamfer distance of
 * every non-zero pixel to the nearest zero pixel (invert != 0: of logical_not(mask)).  Closed-form chamfer metric;
 * matches OpenCV's raster-scan float32 accumulation to ~1e-6 relative (a tolerance, not bit-exact). */
int tmat_distance_transform_cv5_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int invert,
                                    float* out_dev, void* stream);
int tmat_distance_transform_cv5_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int invert,
                                     float* out_host);
/* cv2.circle(zeros((H, W), uint8), (cx, cy), r, val, -1) (preprocessing.py:113-114): OpenCV's integer midpoint
 * circle, bit-exact. */
int tmat_circle_mask_host(tmat_ctx* ctx, int H, int W, int cx, int cy, int r, int val, uint8_t* out_host);

/* ---- the Morse graph built from the weighted prediction -------------- */
/* dmtgraph.py:38-99 compute_dmt_graph(img, delta1, delta2) (native twin: pydmtgraph DMTGraph.cpp:232-615).
 * Sequential union-find sweeps in filtration order: native HOST code in this library (no kernel; needs no GPU).
 * img: float32 (H, W).  vertices: int32 (n_vertices, 2) = (row, column); edges: int32 (n_edges, 2) indices into
 * vertices -- same content AND order as the reference.  Pass NULL for both arrays to query the counts. */
int tmat_dmt_graph_host(const float* img_host, int H, int W, float delta1, float delta2,
                        int32_t* vertices_host, int64_t cap_vertices, int64_t* n_vertices,
                        int32_t* edges_host, int64_t cap_edges, int64_t* n_edges);

/* ---- the resizes either side of the path ---------------------------- */
/* mode: which library call is reproduced.
 *   PIL_LANCZOS  Image.fromarray(float32).resize((Wout, Hout), Image.Resampling.LANCZOS)  models.py:630-634
 *   PIL_NEAREST  Image.fromarray(float32).resize((Wout, Hout), Image.Resampling.NEAREST)  models.py:647-650
 *                both float32 only and BIT-EXACT against Pillow (scalar C: double taps, float32 intermediate);
 *   CV_LINEAR    cv2.resize(img, (Wout, Hout)) -- what scripts/compute_cell_area.py:53-56 really calls: its
 *                cv2.INTER_AREA lands in the positional `dst` slot;
 *   CV_LANCZOS4  cv2.resize(img, (Wout, Hout), interpolation=cv2.INTER_LANCZOS4)  scripts/compute_branches.py:309-312
 *                uint8 / uint16 / float32; OpenCV routes these through vendor (IPP) code, so parity is a tolerance:
 *                <= 1 grey level for integer images, 1e-5 of the range for float32. */
#define TMAT_RESIZE_PIL_LANCZOS 0
#define TMAT_RESIZE_PIL_NEAREST 1
#define TMAT_RESIZE_CV_LINEAR   2
#define TMAT_RESIZE_CV_LANCZOS4 3
int tmat_resize_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W, void* out_dev, int Hout,
                    int Wout, int mode, void* stream);
int tmat_resize_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W, void* out_host, int Hout,
                     int Wout, int mode);

/* ---- whole path for one image (scripts/compute_branches.py:316-337) ------------------ */
/* stack (Z,H,W) uint16 -> max projection (:228) -> rescale_intensity to [0,1] float32 (:316)
 * -> predict_img_with_smooth_windowing (models.py:640) -> pred > 0.5 (:334)
 * -> filter_branch_seg_mask (:337).  pred_dev (optional) receives the float64 probabilities. */
int tmat_branch_mask_dev(tmat_ctx* ctx, tmat_unet* net, const uint16_t* stack_dev, int Z, int H,
                         int W, int subdiv, double* pred_dev, uint8_t* mask_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TMAT_B200_H */
