// Entry points declared in include/tmat_b200.h whose kernels are not written yet.  Each returns
// TMAT_ERR_UNSUPPORTED loudly; the file shrinks as stages land and is deleted when empty.
#include "common.cuh"
#define STUB(name, ...) extern "C" int name(__VA_ARGS__) { tmat_set_error(#name ": not implemented yet"); return TMAT_ERR_UNSUPPORTED; }
STUB(tmat_threshold_gt_f64_dev, tmat_ctx*, const double*, int64_t, double, uint8_t*, void*)
STUB(tmat_apply_mask_f32_dev, tmat_ctx*, const float*, const uint8_t*, int64_t, float*, void*)
STUB(tmat_gaussian_dev, tmat_ctx*, const void*, int, int, int, double, void*, void*)
STUB(tmat_gaussian_host, tmat_ctx*, const void*, int, int, int, double, void*)
STUB(tmat_auto_threshold_well_dev, tmat_ctx*, const void*, int, int, int, uint8_t*, int*, void*)
STUB(tmat_auto_threshold_well_host, tmat_ctx*, const void*, int, int, int, uint8_t*, int*)
STUB(tmat_median_disk2_dev, tmat_ctx*, const uint8_t*, int, int, uint8_t*, void*)
STUB(tmat_erode_disk_dev, tmat_ctx*, const uint8_t*, int, int, int, int, uint8_t*, void*)
STUB(tmat_label_dev, tmat_ctx*, const uint8_t*, int, int, int, int32_t*, int*, void*)
STUB(tmat_region_stats_dev, tmat_ctx*, const int32_t*, int, int, int, int64_t*, int64_t*, void*)
STUB(tmat_skeletonize_dev, tmat_ctx*, const uint8_t*, int, int, uint8_t*, int*, void*)
STUB(tmat_filter_branch_seg_mask_dev, tmat_ctx*, const uint8_t*, int, int, int, int, uint8_t*, void*)
STUB(tmat_filter_branch_seg_mask_host, tmat_ctx*, const uint8_t*, int, int, int, int, uint8_t*)
STUB(tmat_remove_small_islands_host, tmat_ctx*, const uint8_t*, int, int, int, int, int, int, uint8_t*)
STUB(tmat_area_image_host, tmat_ctx*, const uint8_t*, int, int, double*)
STUB(tmat_median_disk2_host, tmat_ctx*, const uint8_t*, int, int, uint8_t*)
STUB(tmat_erode_disk_host, tmat_ctx*, const uint8_t*, int, int, int, int, uint8_t*)
STUB(tmat_label_host, tmat_ctx*, const uint8_t*, int, int, int, int32_t*, int*)
STUB(tmat_skeletonize_host, tmat_ctx*, const uint8_t*, int, int, uint8_t*, int*)
STUB(tmat_branch_mask_dev, tmat_ctx*, tmat_unet*, const uint16_t*, int, int, int, int, double*, uint8_t*, void*)
