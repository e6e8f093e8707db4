// Stage 2 glue: smooth tiled prediction (reference: fl_tissue_model_tools/smooth_tiled_predictions.py)
//   _spline_window:26-41  _pad_img:68-79  _rotate_mirror_do:95-113  _windowed_subdivs:136-192
//   _recreate_from_subdivs:195-217  _rotate_mirror_undo:116-133  predict_img_with_smooth_windowing:220-267
// The blend is an owner-computes gather: one thread per output pixel walks the (orientation,
// tile) pairs that cover it in exactly the order the reference adds them, in float64 with
// separate multiply and add roundings, so the result is bit-identical to numpy for identical
// predictions and needs no atomics.
#include "stp.cuh"
#include <math.h>
#include <algorithm>

int stp_make_geom(int H, int W, int ws, int subdiv, StpGeom* g) {
  TMAT_REQUIRE(H >= 1 && W >= 1, "stp: bad image shape (%d,%d)", H, W);
  TMAT_REQUIRE(ws >= 4 && subdiv >= 1 && subdiv <= ws, "stp: bad window_size %d / subdivisions %d", ws, subdiv);
  g->H = H; g->W = W; g->ws = ws; g->subdiv = subdiv;
  g->aug = (int)nearbyint(ws * (1.0 - 1.0 / subdiv));   // Python round(): half-to-even, as nearbyint
  g->step = ws / subdiv;
  g->PM = H + 2 * g->aug;
  g->PN = W + 2 * g->aug;
  TMAT_REQUIRE(g->step >= 1, "stp: window_size / subdivisions is 0");
  TMAT_REQUIRE(g->PM >= ws && g->PN >= ws, "stp: padded image (%d,%d) smaller than the window %d", g->PM, g->PN, ws);
  g->nR = (g->PM - ws) / g->step + 1;
  g->nC = (g->PN - ws) / g->step + 1;
  return TMAT_OK;
}

extern "C" int tmat_stp_geometry(int H, int W, int ws, int subdiv, int* geom8) {
  TMAT_REQUIRE(geom8, "tmat_stp_geometry: NULL output");
  StpGeom g;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &g));
  geom8[0] = g.aug; geom8[1] = g.PM; geom8[2] = g.PN; geom8[3] = g.step;
  geom8[4] = g.tile_rows(0); geom8[5] = g.tile_cols(0);
  geom8[6] = g.tile_rows(1); geom8[7] = g.tile_cols(1);
  return TMAT_OK;
}

// ---- spline window on the host, numpy-compatible ------------------------------------------
// numpy's pairwise summation (np.add.reduce on a contiguous float64 vector), needed because
// the window is normalised by np.average (:40).
double np_pairwise_sum(const double* a, int64_t n) {
  if (n < 8) {
    double res = 0.;
    for (int64_t i = 0; i < n; i++) res += a[i];
    return res;
  } else if (n <= 128) {
    double r[8];
    for (int k = 0; k < 8; k++) r[k] = a[k];
    int64_t i;
    for (i = 8; i < n - (n % 8); i += 8)
      for (int k = 0; k < 8; k++) r[k] += a[i + k];
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; i++) res += a[i];
    return res;
  }
  int64_t n2 = n / 2;
  n2 -= n2 % 8;
  return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
}

void spline_window_host(int ws, double* out) {
  std::vector<double> t(ws), w(ws);
  // scipy.signal.windows.triang(ws), sym=True
  int half = (ws + 1) / 2;
  for (int n = 1; n <= half; ++n) {
    double v = (ws % 2 == 0) ? (2 * n - 1.0) / ws : 2 * n / (ws + 1.0);
    t[n - 1] = v;
    t[ws - n] = v;
  }
  int q = ws / 4;
  for (int i = 0; i < ws; ++i) {
    double o = fabs(2 * t[i]);
    o = (o * o) / 2;
    if (i >= q && i < ws - q) o = 0;
    double in = fabs(2 * (t[i] - 1));
    in = 1 - (in * in) / 2;
    if (i < q || i >= ws - q) in = 0;
    w[i] = in + o;
  }
  double avg = np_pairwise_sum(w.data(), ws) / (double)ws;
  for (int i = 0; i < ws; ++i) out[i] = w[i] / avg;
}

extern "C" int tmat_spline_window(int ws, double* out_host) {
  TMAT_REQUIRE(ws >= 4 && out_host, "tmat_spline_window: bad argument");
  spline_window_host(ws, out_host);
  return TMAT_OK;
}

// ---- device kernels --------------------------------------------------------------------------
namespace {

__device__ __forceinline__ uint32_t f32_order(float f) {
  uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float f32_unorder(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void minmax_init(uint32_t* enc) { enc[0] = 0xffffffffu; enc[1] = 0u; }

__global__ void __launch_bounds__(256) minmax_f32(const float* __restrict__ x, int64_t n, uint32_t* enc) {
  uint32_t lo = 0xffffffffu, hi = 0u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t e = f32_order(x[i]);
    lo = min(lo, e);
    hi = max(hi, e);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(enc + 0, lo);
    atomicMax(enc + 1, hi);
  }
}

__global__ void minmax_finish(const uint32_t* enc, float* out) {
  out[0] = f32_unorder(enc[0]);
  out[1] = f32_unorder(enc[1]);
}

__global__ void __launch_bounds__(256) stp_gather_kernel(const float* __restrict__ img, StpGeom G, int g,
                                                         const float* __restrict__ minmax,
                                                         float* __restrict__ tiles) {
  const int ws = G.ws;
  const int tile = blockIdx.y;
  const int ti = tile / G.tile_cols(g), tj = tile - ti * G.tile_cols(g);
  const float minval = minmax[0];
  float* dst = tiles + (int64_t)tile * ws * ws;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < ws * ws; p += gridDim.x * blockDim.x) {
    int i = p / ws, j = p - i * ws;
    dst[p] = stp_tile_pixel(img, G, g, ti, tj, i, j, minval);
  }
}

// One thread per output pixel.  For each orientation (in list order) the covering tiles are
// added in raster order of the rotated frame (:206-217), the sum is divided by subdiv^2 (:217),
// the 8 results are added left to right and divided by 8 (np.mean over the list, :133).
__global__ void __launch_bounds__(256) stp_blend_kernel(const float* __restrict__ preds, StpGeom G,
                                                        const double* __restrict__ w1d,
                                                        const int* __restrict__ slot_of_patch, int row0, int row1,
                                                        double* __restrict__ out) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  int y = row0 + blockIdx.y;
  if (x >= G.W || y >= row1) return;
  const int r = y + G.aug, c = x + G.aug;
  const int ws = G.ws, step = G.step;
  const int per_orient = G.tiles_per_orient();
  const double inv_div = (double)(G.subdiv * G.subdiv);
  double total = 0.0;
#pragma unroll 1
  for (int g = 0; g < 8; ++g) {
    int ri, rj;
    d4_orig_to_rot(g, r, c, G.PM, G.PN, &ri, &rj);
    const int a = G.tile_rows(g), b = G.tile_cols(g);
    int ti0 = ri - ws + 1;
    ti0 = ti0 <= 0 ? 0 : (ti0 + step - 1) / step;
    int ti1 = min(ri / step, a - 1);
    int tj0 = rj - ws + 1;
    tj0 = tj0 <= 0 ? 0 : (tj0 + step - 1) / step;
    int tj1 = min(rj / step, b - 1);
    double s = 0.0;
    for (int ti = ti0; ti <= ti1; ++ti) {
      int i = ri - ti * step;
      double wi = w1d[i];
      for (int tj = tj0; tj <= tj1; ++tj) {
        int j = rj - tj * step;
        double w2 = __dmul_rn(wi, w1d[j]);
        int pid = g * per_orient + ti * b + tj;
        if (slot_of_patch) pid = __ldg(slot_of_patch + pid);
        double p = (double)__ldg(preds + ((int64_t)pid * ws + i) * ws + j);
        s = __dadd_rn(s, __dmul_rn(p, w2));
      }
    }
    s = __ddiv_rn(s, inv_div);
    total = (g == 0) ? s : __dadd_rn(total, s);
  }
  out[(int64_t)(y - row0) * G.W + x] = __ddiv_rn(total, 8.0);
}

}  // namespace

int stp_image_min(tmat_ctx* ctx, const float* img_dev, int64_t n, float* minmax_dev, cudaStream_t st) {
  uint32_t* enc;
  TMAT_TRY(arena_get(ctx, 2, &enc));
  minmax_init<<<1, 1, 0, st>>>(enc);
  KERNEL_CHECK(ctx);
  int blocks = (int)std::min<int64_t>(cdiv64(n, 256), (int64_t)ctx->sm_count * 8);
  minmax_f32<<<blocks, 256, 0, st>>>(img_dev, n, enc);
  KERNEL_CHECK(ctx);
  minmax_finish<<<1, 1, 0, st>>>(enc, minmax_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

int stp_upload_window(tmat_ctx* ctx, int ws, const double* window1d_host, double** w1d_dev, cudaStream_t st) {
  TMAT_TRY(arena_get(ctx, ws, w1d_dev));
  // staged through the context's pinned buffer so the async copy never reads freed host memory
  TMAT_REQUIRE((size_t)ws * sizeof(double) <= ctx->pinned_cap, "stp: window_size %d too large", ws);
  double* stage = (double*)ctx->pinned;
  CUDA_TRY(cudaStreamSynchronize(st));   // previous user of the staging buffer
  if (window1d_host) memcpy(stage, window1d_host, ws * sizeof(double));
  else spline_window_host(ws, stage);
  CUDA_TRY(cudaMemcpyAsync(*w1d_dev, stage, ws * sizeof(double), cudaMemcpyHostToDevice, st));
  return TMAT_OK;
}

int stp_blend(tmat_ctx* ctx, const StpGeom& G, const float* preds_dev, const double* w1d_dev, double* out_dev,
              cudaStream_t st) {
  dim3 grid(cdiv(G.W, 256), G.H);
  stp_blend_kernel<<<grid, 256, 0, st>>>(preds_dev, G, w1d_dev, nullptr, 0, G.H, out_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

int stp_blend_rows(tmat_ctx* ctx, const StpGeom& G, const float* preds_dev, const int* slot_of_patch_dev,
                   const double* w1d_dev, int row0, int row1, double* out_rows_dev, cudaStream_t st) {
  TMAT_REQUIRE(row0 >= 0 && row0 < row1 && row1 <= G.H, "stp: bad row range [%d, %d) for H=%d", row0, row1, G.H);
  dim3 grid(cdiv(G.W, 256), row1 - row0);
  stp_blend_kernel<<<grid, 256, 0, st>>>(preds_dev, G, w1d_dev, slot_of_patch_dev, row0, row1, out_rows_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

void stp_patches_for_rows(const StpGeom& G, int row0, int row1, std::vector<int>* list) {
  list->clear();
  const int lo = row0 + G.aug, hi = row1 + G.aug;     // padded rows [lo, hi) of the ORIGINAL frame
  const int per = G.tiles_per_orient();
  for (int g = 0; g < 8; ++g) {
    const int a = G.tile_rows(g), b = G.tile_cols(g);
    for (int ti = 0; ti < a; ++ti)
      for (int tj = 0; tj < b; ++tj) {
        // the window is an axis-aligned square in both frames: map two opposite corners
        int r0, c0, r1, c1;
        d4_rot_to_orig(g, ti * G.step, tj * G.step, G.PM, G.PN, &r0, &c0);
        d4_rot_to_orig(g, ti * G.step + G.ws - 1, tj * G.step + G.ws - 1, G.PM, G.PN, &r1, &c1);
        const int rmin = r0 < r1 ? r0 : r1, rmax = r0 < r1 ? r1 : r0;
        if (rmax >= lo && rmin < hi) list->push_back(g * per + ti * b + tj);
      }
  }
}

// host geometry: which patches a row strip needs (for sharding one large image across GPUs by row strips)
extern "C" int tmat_stp_rows_patches(int H, int W, int ws, int subdiv, int row0, int row1, int* list, int capacity,
                                     int* n_out) {
  TMAT_REQUIRE(n_out, "tmat_stp_rows_patches: NULL output");
  StpGeom G;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  TMAT_REQUIRE(row0 >= 0 && row0 < row1 && row1 <= H, "tmat_stp_rows_patches: bad row range [%d, %d)", row0, row1);
  std::vector<int> v;
  stp_patches_for_rows(G, row0, row1, &v);
  *n_out = (int)v.size();
  if (list) {
    TMAT_REQUIRE(capacity >= (int)v.size(), "tmat_stp_rows_patches: capacity %d < %d", capacity, (int)v.size());
    memcpy(list, v.data(), v.size() * sizeof(int));
  }
  return TMAT_OK;
}

static int gather_impl(tmat_ctx* ctx, const float* img_dev, const StpGeom& G, int orient, float* tiles_dev,
                       cudaStream_t st) {
  float* minmax;
  TMAT_TRY(arena_get(ctx, 2, &minmax));
  TMAT_TRY(stp_image_min(ctx, img_dev, (int64_t)G.H * G.W, minmax, st));
  dim3 grid(std::min(cdiv(G.ws * G.ws, 256), 64), G.tiles_per_orient());
  stp_gather_kernel<<<grid, 256, 0, st>>>(img_dev, G, orient, minmax, tiles_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

extern "C" int tmat_stp_gather_dev(tmat_ctx* ctx, const float* img_dev, int H, int W, int ws, int subdiv,
                                   int orient, float* tiles_dev, void* stream) {
  TMAT_REQUIRE(ctx && img_dev && tiles_dev, "tmat_stp_gather: NULL argument");
  TMAT_REQUIRE(orient >= 0 && orient < 8, "tmat_stp_gather: orientation %d not in 0..7", orient);
  StpGeom G;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  TMAT_TRY(arena_reset(ctx));
  return gather_impl(ctx, img_dev, G, orient, tiles_dev, as_stream(stream));
}

extern "C" int tmat_stp_blend_dev(tmat_ctx* ctx, const float* preds_dev, int H, int W, int ws, int subdiv,
                                  const double* window1d_host, double* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && preds_dev && out_dev, "tmat_stp_blend: NULL argument");
  StpGeom G;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  TMAT_TRY(arena_reset(ctx));
  double* w1d;
  TMAT_TRY(stp_upload_window(ctx, ws, window1d_host, &w1d, as_stream(stream)));
  return stp_blend(ctx, G, preds_dev, w1d, out_dev, as_stream(stream));
}

extern "C" int tmat_stp_gather_host(tmat_ctx* ctx, const float* img_host, int H, int W, int ws, int subdiv,
                                    int orient, float* tiles_host) {
  TMAT_REQUIRE(ctx && img_host && tiles_host, "tmat_stp_gather_host: NULL argument");
  TMAT_REQUIRE(orient >= 0 && orient < 8, "tmat_stp_gather_host: orientation %d not in 0..7", orient);
  StpGeom G;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  TMAT_TRY(arena_reset(ctx));
  float *d_img, *d_tiles;
  size_t n_tile_px = (size_t)G.tiles_per_orient() * ws * ws;
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_img));
  TMAT_TRY(arena_get(ctx, n_tile_px, &d_tiles));
  CUDA_TRY(cudaMemcpyAsync(d_img, img_host, (size_t)H * W * 4, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(gather_impl(ctx, d_img, G, orient, d_tiles, 0));
  CUDA_TRY(cudaMemcpyAsync(tiles_host, d_tiles, n_tile_px * 4, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

extern "C" int tmat_stp_blend_host(tmat_ctx* ctx, const float* preds_host, int H, int W, int ws, int subdiv,
                                   const double* window1d_host, double* out_host) {
  TMAT_REQUIRE(ctx && preds_host && out_host, "tmat_stp_blend_host: NULL argument");
  StpGeom G;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  TMAT_TRY(arena_reset(ctx));
  size_t n_px = (size_t)8 * G.tiles_per_orient() * ws * ws;
  float* d_preds;
  double *d_out, *w1d;
  TMAT_TRY(arena_get(ctx, n_px, &d_preds));
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_out));
  TMAT_TRY(stp_upload_window(ctx, ws, window1d_host, &w1d, 0));
  CUDA_TRY(cudaMemcpyAsync(d_preds, preds_host, n_px * 4, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(stp_blend(ctx, G, d_preds, w1d, d_out, 0));
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, (size_t)H * W * 8, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}
