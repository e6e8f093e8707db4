// Exact Euclidean distance transform and the centerline weighting handed to the Morse graph.
// Reference: scripts/compute_branches.py:340-344
//     skel, dist = medial_axis(seg_mask, return_distance=True)      # dist = ndi.distance_transform_edt(seg_mask)
//     centerline_dt = distance_transform_edt(np.logical_not(skel))
//     relative_dt = dist / (dist + centerline_dt);   pred = pred * relative_dt
// scipy.ndimage.distance_transform_edt returns sqrt(exact integer squared distance to the nearest zero pixel) in
// float64, so an exact integer EDT followed by one IEEE square root is bit-identical to it.  Two separable passes:
// nearest zero along each column (two scans), then per pixel the minimum over the row of dx^2 + g^2, searched outwards
// and cut off as soon as dx^2 alone reaches the best value found (vessel masks: a handful of steps per pixel).
// medial_axis itself (skimage's priority-ordered thinning with RANDOM tie-breaking, rng=None in the reference's
// call) is not reproducible even against itself; the skeleton is an input here.
#include "common.cuh"

namespace {

constexpr unsigned int EDT_INF = 0xffffffffu;

// g[y][x] = distance along the column to the nearest zero pixel (EDT_INF if the column has none)
__global__ void __launch_bounds__(128) edt_columns(const uint8_t* __restrict__ in, int H, int W, int invert,
                                                   unsigned int* __restrict__ g) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= W) return;
  unsigned int d = EDT_INF;
  for (int y = 0; y < H; ++y) {
    const bool zero = ((in[(int64_t)y * W + x] != 0) == (invert != 0));
    d = zero ? 0u : (d == EDT_INF ? EDT_INF : d + 1u);
    g[(int64_t)y * W + x] = d;
  }
  d = EDT_INF;
  for (int y = H - 1; y >= 0; --y) {
    const bool zero = ((in[(int64_t)y * W + x] != 0) == (invert != 0));
    d = zero ? 0u : (d == EDT_INF ? EDT_INF : d + 1u);
    const int64_t i = (int64_t)y * W + x;
    if (d < g[i]) g[i] = d;
  }
}

__device__ __forceinline__ unsigned long long edt_row_min(const unsigned int* __restrict__ grow, int W, int x) {
  const unsigned int g0 = grow[x];
  unsigned long long best = g0 == EDT_INF ? ~0ull : (unsigned long long)g0 * g0;
  for (int d = 1; d < W; ++d) {
    const unsigned long long dd = (unsigned long long)d * d;
    if (dd >= best) break;
    const int xl = x - d, xr = x + d;
    if (xl < 0 && xr >= W) break;
    if (xl >= 0) {
      const unsigned int gv = grow[xl];
      if (gv != EDT_INF) { const unsigned long long c = dd + (unsigned long long)gv * gv; best = c < best ? c : best; }
    }
    if (xr < W) {
      const unsigned int gv = grow[xr];
      if (gv != EDT_INF) { const unsigned long long c = dd + (unsigned long long)gv * gv; best = c < best ? c : best; }
    }
  }
  return best;
}

__global__ void __launch_bounds__(256) edt_rows(const unsigned int* __restrict__ g, int H, int W, double* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const unsigned long long best = edt_row_min(g + (int64_t)y * W, W, x);
  // an image without any zero pixel has no defined distance: +inf
  out[(int64_t)y * W + x] = best == ~0ull ? __longlong_as_double(0x7ff0000000000000ll) : __dsqrt_rn((double)best);
}

// pred * dist / (dist + centerline_dt) with separately rounded IEEE operations, as numpy evaluates it
__global__ void __launch_bounds__(256) centerline_weight(const double* __restrict__ pred, const unsigned int* __restrict__ g_mask,
                                                         const unsigned int* __restrict__ g_skel, int H, int W,
                                                         double* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const int64_t i = (int64_t)y * W + x;
  const unsigned long long a = edt_row_min(g_mask + (int64_t)y * W, W, x);
  const unsigned long long b = edt_row_min(g_skel + (int64_t)y * W, W, x);
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  const double dist = a == ~0ull ? inf : __dsqrt_rn((double)a);
  const double cdt = b == ~0ull ? inf : __dsqrt_rn((double)b);
  const double rel = __ddiv_rn(dist, __dadd_rn(dist, cdt));
  out[i] = __dmul_rn(pred[i], rel);
}

}  // namespace

// scipy.ndimage.distance_transform_edt(mask): distance of every non-zero pixel to the nearest zero pixel.
// invert != 0 computes the transform of logical_not(mask) without materialising it.
extern "C" int tmat_distance_transform_edt_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int invert,
                                               double* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && mask_dev && out_dev && H >= 1 && W >= 1, "tmat_distance_transform_edt: bad argument");
  TMAT_REQUIRE(H <= 65535 && W <= 65535, "tmat_distance_transform_edt: image too large");
  TMAT_TRY(arena_reset(ctx));
  cudaStream_t st = as_stream(stream);
  unsigned int* g;
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &g));
  edt_columns<<<cdiv(W, 128), 128, 0, st>>>(mask_dev, H, W, invert, g);
  KERNEL_CHECK(ctx);
  edt_rows<<<dim3(cdiv(W, 256), H), 256, 0, st>>>(g, H, W, out_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

extern "C" int tmat_distance_transform_edt_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int invert,
                                                double* out_host) {
  TMAT_REQUIRE(ctx && mask_host && out_host && H >= 1 && W >= 1, "tmat_distance_transform_edt_host: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t n = (size_t)H * W;
  uint8_t* d_in = nullptr;
  double* d_out = nullptr;
  CUDA_TRY(cudaMalloc((void**)&d_in, n));
  if (cudaMalloc((void**)&d_out, n * 8) != cudaSuccess) { cudaFree(d_in); tmat_set_error("edt: out of memory"); return TMAT_ERR_NOMEM; }
  int rc = cudaMemcpyAsync(d_in, mask_host, n, cudaMemcpyHostToDevice, 0) == cudaSuccess ? TMAT_OK : TMAT_ERR_CUDA;
  if (rc == TMAT_OK) rc = tmat_distance_transform_edt_dev(ctx, d_in, H, W, invert, d_out, nullptr);
  if (rc == TMAT_OK && cudaMemcpyAsync(out_host, d_out, n * 8, cudaMemcpyDeviceToHost, 0) != cudaSuccess) rc = TMAT_ERR_CUDA;
  if (cudaStreamSynchronize(0) != cudaSuccess && rc == TMAT_OK) rc = TMAT_ERR_CUDA;
  cudaFree(d_in);
  cudaFree(d_out);
  if (rc == TMAT_ERR_CUDA) tmat_set_error("tmat_distance_transform_edt_host: %s", cudaGetErrorString(cudaGetLastError()));
  return rc;
}

// compute_branches.py:341-344 in one call: out = pred * edt(mask) / (edt(mask) + edt(~skel))
extern "C" int tmat_centerline_weight_dev(tmat_ctx* ctx, const double* pred_dev, const uint8_t* mask_dev,
                                          const uint8_t* skel_dev, int H, int W, double* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && pred_dev && mask_dev && skel_dev && out_dev && H >= 1 && W >= 1, "tmat_centerline_weight: bad argument");
  TMAT_REQUIRE(H <= 65535 && W <= 65535, "tmat_centerline_weight: image too large");
  TMAT_TRY(arena_reset(ctx));
  cudaStream_t st = as_stream(stream);
  unsigned int *gm, *gs;
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &gm));
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &gs));
  edt_columns<<<cdiv(W, 128), 128, 0, st>>>(mask_dev, H, W, 0, gm);
  KERNEL_CHECK(ctx);
  edt_columns<<<cdiv(W, 128), 128, 0, st>>>(skel_dev, H, W, 1, gs);
  KERNEL_CHECK(ctx);
  centerline_weight<<<dim3(cdiv(W, 256), H), 256, 0, st>>>(pred_dev, gm, gs, H, W, out_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

extern "C" int tmat_centerline_weight_host(tmat_ctx* ctx, const double* pred_host, const uint8_t* mask_host,
                                           const uint8_t* skel_host, int H, int W, double* out_host) {
  TMAT_REQUIRE(ctx && pred_host && mask_host && skel_host && out_host && H >= 1 && W >= 1, "tmat_centerline_weight_host: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t n = (size_t)H * W;
  char* blk = nullptr;                       // pred | out | mask | skel
  CUDA_TRY(cudaMalloc((void**)&blk, n * 18));
  double *d_pred = (double*)blk, *d_out = (double*)(blk + n * 8);
  uint8_t *d_mask = (uint8_t*)(blk + n * 16), *d_skel = d_mask + n;
  int rc = TMAT_OK;
  if (cudaMemcpyAsync(d_pred, pred_host, n * 8, cudaMemcpyHostToDevice, 0) != cudaSuccess ||
      cudaMemcpyAsync(d_mask, mask_host, n, cudaMemcpyHostToDevice, 0) != cudaSuccess ||
      cudaMemcpyAsync(d_skel, skel_host, n, cudaMemcpyHostToDevice, 0) != cudaSuccess) rc = TMAT_ERR_CUDA;
  if (rc == TMAT_OK) rc = tmat_centerline_weight_dev(ctx, d_pred, d_mask, d_skel, H, W, d_out, nullptr);
  if (rc == TMAT_OK && cudaMemcpyAsync(out_host, d_out, n * 8, cudaMemcpyDeviceToHost, 0) != cudaSuccess) rc = TMAT_ERR_CUDA;
  if (cudaStreamSynchronize(0) != cudaSuccess && rc == TMAT_OK) rc = TMAT_ERR_CUDA;
  cudaFree(blk);
  if (rc == TMAT_ERR_CUDA) tmat_set_error("tmat_centerline_weight_host: %s", cudaGetErrorString(cudaGetLastError()));
  return rc;
}

// ---- cv2.distanceTransform(mask, cv2.DIST_L2, 5) and cv2.circle(..., thickness=-1) -----------------------------------
// preprocessing.py:96-163 (gen_circ_mask, dt_blur, sdt_blur; notebook / training helpers).
// OpenCV's 5x5 chamfer transform walks the image in two raster scans adding 1.0f / 1.4f / 2.1969f per step in
// float32.  In free space the cheapest chamfer path between two pixels has a closed form (dx >= dy: dy knight moves +
// (dx - 2 dy) straight ones if 2 dy <= dx, else (dx - dy) knight moves + (2 dy - dx) diagonal ones), and the cost is
// monotone in |dy|, so the same two passes as the exact EDT apply with that metric.  The closed form adds the same
// steps in a different ORDER than the raster scan: parity with OpenCV is ~1e-6 relative (a tolerance, stated in
// tests/test_gpu_preproc.py), exact after the reference's .round().astype(uint8) except at ties.
namespace {

__device__ __forceinline__ float chamfer5(int dx, int dy) {
  const int mx = max(dx, dy), mn = min(dx, dy);
  if (2 * mn <= mx) return __fadd_rn(__fmul_rn(2.1969f, (float)mn), (float)(mx - 2 * mn));
  return __fadd_rn(__fmul_rn(2.1969f, (float)(mx - mn)), __fmul_rn(1.4f, (float)(2 * mn - mx)));
}
__global__ void __launch_bounds__(256) chamfer_rows(const unsigned int* __restrict__ g, int H, int W, float* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const unsigned int* grow = g + (int64_t)y * W;
  const unsigned int g0 = grow[x];
  float best = g0 == EDT_INF ? 3.0e38f : (float)g0;
  for (int d = 1; d < W; ++d) {
    if ((float)d >= best) break;                         // a path that moves d columns costs at least d
    const int xl = x - d, xr = x + d;
    if (xl < 0 && xr >= W) break;
    if (xl >= 0) { const unsigned int gv = grow[xl]; if (gv != EDT_INF) best = fminf(best, chamfer5(d, (int)gv)); }
    if (xr < W) { const unsigned int gv = grow[xr]; if (gv != EDT_INF) best = fminf(best, chamfer5(d, (int)gv)); }
  }
  out[(int64_t)y * W + x] = best;
}
__global__ void __launch_bounds__(256) circle_fill(int H, int W, int cx, int cy, int r, const int* __restrict__ half_width,
                                                   uint8_t val, uint8_t* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const int dy = abs(y - cy);
  out[(int64_t)y * W + x] = (dy <= r && abs(x - cx) <= half_width[dy]) ? val : (uint8_t)0;
}

}  // namespace

extern "C" int tmat_distance_transform_cv5_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int invert,
                                               float* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && mask_dev && out_dev && H >= 1 && W >= 1, "tmat_distance_transform_cv5: bad argument");
  TMAT_TRY(arena_reset(ctx));
  cudaStream_t st = as_stream(stream);
  unsigned int* g;
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &g));
  edt_columns<<<cdiv(W, 128), 128, 0, st>>>(mask_dev, H, W, invert, g);
  KERNEL_CHECK(ctx);
  chamfer_rows<<<dim3(cdiv(W, 256), H), 256, 0, st>>>(g, H, W, out_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

extern "C" int tmat_distance_transform_cv5_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int invert,
                                                float* out_host) {
  TMAT_REQUIRE(ctx && mask_host && out_host && H >= 1 && W >= 1, "tmat_distance_transform_cv5_host: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t n = (size_t)H * W;
  char* blk = nullptr;
  CUDA_TRY(cudaMalloc((void**)&blk, n * 5));
  float* d_out = (float*)blk;
  uint8_t* d_in = (uint8_t*)(blk + n * 4);
  int rc = cudaMemcpyAsync(d_in, mask_host, n, cudaMemcpyHostToDevice, 0) == cudaSuccess ? TMAT_OK : TMAT_ERR_CUDA;
  if (rc == TMAT_OK) rc = tmat_distance_transform_cv5_dev(ctx, d_in, H, W, invert, d_out, nullptr);
  if (rc == TMAT_OK && cudaMemcpyAsync(out_host, d_out, n * 4, cudaMemcpyDeviceToHost, 0) != cudaSuccess) rc = TMAT_ERR_CUDA;
  if (cudaStreamSynchronize(0) != cudaSuccess && rc == TMAT_OK) rc = TMAT_ERR_CUDA;
  cudaFree(blk);
  if (rc == TMAT_ERR_CUDA) tmat_set_error("tmat_distance_transform_cv5_host: %s", cudaGetErrorString(cudaGetLastError()));
  return rc;
}

// cv2.circle(img, (cx, cy), r, val, -1) on a zero image: OpenCV's integer midpoint circle (drawing.cpp Circle()),
// reproduced exactly -- the per-row half widths come from the same error recurrence.
extern "C" int tmat_circle_mask_host(tmat_ctx* ctx, int H, int W, int cx, int cy, int r, int val, uint8_t* out_host) {
  TMAT_REQUIRE(ctx && out_host && H >= 1 && W >= 1 && r >= 0, "tmat_circle_mask: bad argument");
  TMAT_REQUIRE(val >= 0 && val <= 255, "tmat_circle_mask: mask_val %d outside [0, 255]", val);
  CUDA_TRY(cudaSetDevice(ctx->device));
  std::vector<int> hw((size_t)r + 1, -1);
  int err = 0, dx = r, dy = 0, plus = 1, minus = (r << 1) - 1;
  while (dx >= dy) {
    hw[dy] = std::max(hw[dy], dx);
    hw[dx] = std::max(hw[dx], dy);
    dy++;
    err += plus;
    plus += 2;
    const int mask = (err <= 0) - 1;
    err -= minus & mask;
    dx += mask;
    minus -= mask & 2;
  }
  TMAT_TRY(arena_reset(ctx));
  int* d_hw;
  uint8_t* d_out;
  TMAT_TRY(arena_get(ctx, hw.size(), &d_hw));
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_hw, hw.data(), hw.size() * sizeof(int), cudaMemcpyHostToDevice, 0));
  circle_fill<<<dim3(cdiv(W, 256), H), 256>>>(H, W, cx, cy, r, d_hw, (uint8_t)val, d_out);
  KERNEL_CHECK(ctx);
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, (size_t)H * W, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

// ---- allocation-free variant of the exact EDT -------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  The column-distance plane `g` that
// tmat_distance_transform_edt_dev takes from the context arena lives here in a file-scope device array of
// TMAT_EDT_STATIC_MAX_PIXELS uint32 (16 MiB, held by every GPU that loads the library until the process exits).
// In host code the array's NAME is only a stand-in for the symbol, not a device address: the address is looked up once
// per device, with that device current, by cudaGetSymbolAddress and that pointer is what the kernels receive.
// Nothing here needs to start at zero: edt_columns overwrites every element of the H*W prefix that edt_rows reads, so
// the previous call's contents never show through.  All calls share the one array: calls whose kernels overlap on the
// GPU (different streams, different host threads on the same device) race -- serialise them.
#include <mutex>

static __device__ __align__(256) unsigned int g_edt_static_scratch[TMAT_EDT_STATIC_MAX_PIXELS];

namespace {
constexpr int EDT_STATIC_MAX_DEVICES = 64;
std::mutex g_edt_static_mu;
unsigned int* g_edt_static_ptr[EDT_STATIC_MAX_DEVICES];          // per device; nullptr until first use there
}  // namespace

extern "C" int tmat_distance_transform_edt_static_dev(const uint8_t* mask_dev, int H, int W, int invert,
                                                      double* out_dev, void* stream) {
  TMAT_REQUIRE(mask_dev && out_dev && H >= 1 && W >= 1, "tmat_distance_transform_edt_static: bad argument");
  TMAT_REQUIRE(H <= 65535 && W <= 65535 && (int64_t)H * W <= (int64_t)TMAT_EDT_STATIC_MAX_PIXELS,
               "tmat_distance_transform_edt_static: %d x %d = %lld pixels exceed the static scratch of %d pixels "
               "(TMAT_EDT_STATIC_MAX_PIXELS); use tmat_distance_transform_edt_dev for larger images",
               H, W, (long long)H * W, (int)TMAT_EDT_STATIC_MAX_PIXELS);
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  TMAT_REQUIRE(dev >= 0 && dev < EDT_STATIC_MAX_DEVICES, "tmat_distance_transform_edt_static: device %d out of range", dev);
  unsigned int* g = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_edt_static_mu);
    if (!g_edt_static_ptr[dev]) {                                // set-up, once per device, that device current
      void* p = nullptr;
      CUDA_TRY(cudaGetSymbolAddress(&p, g_edt_static_scratch));
      g_edt_static_ptr[dev] = (unsigned int*)p;
    }
    g = g_edt_static_ptr[dev];
  }
  cudaStream_t st = as_stream(stream);
  edt_columns<<<cdiv(W, 128), 128, 0, st>>>(mask_dev, H, W, invert, g);
  CUDA_TRY(cudaGetLastError());
  edt_rows<<<dim3(cdiv(W, 256), H), 256, 0, st>>>(g, H, W, out_dev);
  CUDA_TRY(cudaGetLastError());
  return TMAT_OK;
}
