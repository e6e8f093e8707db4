// Whole branching-mask path for one image (reference: scripts/compute_branches.py:228, :316-337):
//   img.max(0) -> rescale_intensity(.., (0,1)).astype(float32) -> model.predict (tiled UNet + D4 + blend)
//   -> pred > 0.5 -> filter_branch_seg_mask.
// Every stage runs on the caller's stream on device-resident buffers; nothing returns to the host (the label count
// and the thinning convergence are read on the device).
#include "unet.cuh"
#include <algorithm>

int zproj_max_u16(tmat_ctx* ctx, const uint16_t* stack, int Z, int H, int W, uint16_t* out, cudaStream_t st);
int stp_predict_impl(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W, int subdiv, double* out_dev,
                     cudaStream_t st);
int filter_branch_impl(tmat_ctx* ctx, const uint8_t* mask, int H, int W, int use_median, int remove_isolated,
                       uint8_t* out, cudaStream_t st);
size_t stp_predict_scratch_bytes(const tmat_unet* net, int H, int W, int subdiv);

namespace {

__global__ void u16_minmax_init(unsigned int* mm) { mm[0] = 0xffffffffu; mm[1] = 0u; }
__global__ void __launch_bounds__(256) u16_minmax(const uint16_t* __restrict__ x, int64_t n, unsigned int* mm) {
  unsigned int lo = 0xffffffffu, hi = 0u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned int v = x[i];
    lo = min(lo, v);
    hi = max(hi, v);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(mm, lo); atomicMax(mm + 1, hi); }
}
// skimage rescale_intensity(img, out_range=(0, 1)) on an integer image: float64 (x - lo) / (hi - lo) * 1 + 0,
// then .astype(float32)  (compute_branches.py:316)
__global__ void __launch_bounds__(256) rescale_unit_f32(const uint16_t* __restrict__ x, int64_t n,
                                                        const unsigned int* __restrict__ mm, float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double lo = (double)mm[0], hi = (double)mm[1];
  double v = (double)x[i];
  if (hi != lo) v = __ddiv_rn(__dsub_rn(v, lo), __dsub_rn(hi, lo));
  else v = fmin(fmax(v, 0.0), 1.0);      // a constant image: skimage returns np.clip(image, omin, omax)
  out[i] = (float)v;
}
__global__ void __launch_bounds__(256) gt_half(const double* __restrict__ p, int64_t n, uint8_t* __restrict__ m) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) m[i] = p[i] > 0.5;
}

}  // namespace

extern "C" int tmat_branch_mask_dev(tmat_ctx* ctx, tmat_unet* net, const uint16_t* stack_dev, int Z, int H, int W,
                                    int subdiv, double* pred_dev, uint8_t* mask_dev, void* stream) {
  TMAT_REQUIRE(ctx && net && stack_dev && mask_dev, "tmat_branch_mask: NULL argument");
  TMAT_REQUIRE(Z >= 1 && H >= 1 && W >= 1, "tmat_branch_mask: bad shape (%d,%d,%d)", Z, H, W);
  cudaStream_t st = as_stream(stream);
  TMAT_TRY(arena_reset(ctx));
  int64_t n = (int64_t)H * W;
  TMAT_TRY(arena_reserve(ctx, stp_predict_scratch_bytes(net, H, W, subdiv) + (size_t)n * 64 + (8 << 20)));
  uint16_t* proj;
  unsigned int* mm;
  float* img;
  double* pred = pred_dev;
  uint8_t* raw_mask;
  TMAT_TRY(arena_get(ctx, (size_t)n, &proj));
  TMAT_TRY(arena_get(ctx, 2, &mm));
  TMAT_TRY(arena_get(ctx, (size_t)n, &img));
  TMAT_TRY(arena_get(ctx, (size_t)n, &raw_mask));
  if (!pred) TMAT_TRY(arena_get(ctx, (size_t)n, &pred));
  TMAT_TRY(zproj_max_u16(ctx, stack_dev, Z, H, W, proj, st));
  u16_minmax_init<<<1, 1, 0, st>>>(mm);
  KERNEL_CHECK(ctx);
  int blocks = (int)std::min<int64_t>(cdiv64(n, 256), (int64_t)ctx->sm_count * 8);
  u16_minmax<<<blocks, 256, 0, st>>>(proj, n, mm);
  KERNEL_CHECK(ctx);
  rescale_unit_f32<<<(unsigned)cdiv64(n, 256), 256, 0, st>>>(proj, n, mm, img);
  KERNEL_CHECK(ctx);
  TMAT_TRY(stp_predict_impl(ctx, net, img, H, W, subdiv, pred, st));
  gt_half<<<(unsigned)cdiv64(n, 256), 256, 0, st>>>(pred, n, raw_mask);
  KERNEL_CHECK(ctx);
  return filter_branch_impl(ctx, raw_mask, H, W, 1, 1, mask_dev, st);
}
