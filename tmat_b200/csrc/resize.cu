// Resizes either side of the segmentation path.
//   * Pillow  Image.resize(size, LANCZOS) on a mode-'F' image and Image.resize(size, NEAREST)
//     -- UNetXceptionPatchSegmentor.predict(auto_resample=True), models.py:630-651.  Bit-exact: Pillow's
//     ImagingResample is scalar C (double coefficients, double accumulation, float32 intermediate image).
//   * cv2.resize(img, dsize, interpolation=INTER_LANCZOS4) -- scripts/compute_branches.py:309-312 -- and
//     cv2.resize(img, dsize, cv2.INTER_AREA) of scripts/compute_cell_area.py:53-56, where the flag lands in the
//     positional `dst` slot and the call is therefore INTER_LINEAR.  OpenCV's wheel routes these
//     through vendor code (IPP) whose rounding is not reproducible: parity is a TOLERANCE (<= 1 grey level for integer
//     images, 1e-5 of the range for float32; tests/test_gpu_resize.py), not bit-exactness.
// Every mode is a separable resampling: out[i] = sum_t in[idx[i][t]] * coef[i][t], taps in ascending order.  The tap
// tables are built on the host exactly as the libraries build them (same libm) and the two passes run on the device.
#include "common.cuh"
#include <math.h>

namespace {

struct Taps {
  int n_out = 0, ksize = 0;
  std::vector<int> idx;         // n_out * ksize source indices (always valid)
  std::vector<double> coef;     // n_out * ksize coefficients (0 for unused taps)
};

// ---- Pillow: precompute_coeffs (src/libImaging/Resample.c) with the Lanczos filter (support 3) -----------------
double pil_sinc(double x) {
  if (x == 0.0) return 1.0;
  x = x * M_PI;
  return sin(x) / x;
}
double pil_lanczos(double x) {
  if (-3.0 <= x && x < 3.0) return pil_sinc(x) * pil_sinc(x / 3);
  return 0.0;
}
Taps pil_lanczos_taps(int in_size, int out_size) {
  Taps t;
  const double support0 = 3.0;
  double scale = (double)((float)in_size - 0.0f) / out_size, filterscale = scale;
  if (filterscale < 1.0) filterscale = 1.0;
  const double support = support0 * filterscale;
  t.ksize = (int)ceil(support) * 2 + 1;
  t.n_out = out_size;
  t.idx.assign((size_t)out_size * t.ksize, 0);
  t.coef.assign((size_t)out_size * t.ksize, 0.0);
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = 0.0 + (xx + 0.5) * scale;
    double ww = 0.0;
    const double ss = 1.0 / filterscale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double* k = &t.coef[(size_t)xx * t.ksize];
    int* ix = &t.idx[(size_t)xx * t.ksize];
    for (int x = 0; x < xmax; ++x) {
      const double w = pil_lanczos((x + xmin - center + 0.5) * ss);
      k[x] = w;
      ww += w;
    }
    for (int x = 0; x < xmax; ++x)
      if (ww != 0.0) k[x] /= ww;
    for (int x = 0; x < t.ksize; ++x) ix[x] = x < xmax ? xmin + x : xmin;     // unused taps: coefficient 0
  }
  return t;
}
// Pillow NEAREST: ImagingScaleAffine (Geometry.c) -- positions accumulated in double, COORD() truncation
std::vector<int> pil_nearest_index(int in_size, int out_size) {
  std::vector<int> tab(out_size, -1);
  const double a = (double)in_size / out_size;
  double xo = 0.0 + a * 0.5;
  for (int x = 0; x < out_size; ++x) {
    const int xin = xo < 0.0 ? -1 : (int)xo;
    if (xin >= 0 && xin < in_size) tab[x] = xin;
    xo += a;
  }
  return tab;
}

// ---- OpenCV: imgproc/src/resize.cpp ------------------------------------------------------------------------------
// float_fx: OpenCV's own loops take the fraction in float32 ((float)(...) then fx -= sx); the path its wheel uses for
// 16-bit and float images (>= 2 rows) behaves like a fraction taken in double -- measured, see the test's docstring
Taps cv_linear_taps(int in_size, int out_size, bool float_fx = false) {
  Taps t;
  t.ksize = 2;
  t.n_out = out_size;
  t.idx.resize((size_t)out_size * 2);
  t.coef.resize((size_t)out_size * 2);
  const double scale = 1.0 / ((double)out_size / in_size);
  for (int d = 0; d < out_size; ++d) {
    double fx = (d + 0.5) * scale - 0.5;
    int sx;
    if (float_fx) {
      float f = (float)fx;
      sx = (int)floorf(f);
      f -= sx;
      fx = f;
    } else {
      sx = (int)floor(fx);
      fx -= sx;
    }
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= in_size - 1) { fx = 0; sx = in_size - 1; }
    t.idx[2 * d] = sx;
    t.idx[2 * d + 1] = sx + 1 < in_size ? sx + 1 : in_size - 1;
    t.coef[2 * d] = (double)(1.f - (float)fx);
    t.coef[2 * d + 1] = (double)(float)fx;
  }
  return t;
}
void cv_lanczos4_coeffs(float x, float* coeffs) {      // interpolateLanczos4
  static const double s45 = 0.70710678118654752440084436210485;
  static const double cs[][2] = {{1, 0}, {-s45, -s45}, {0, 1}, {s45, -s45}, {-1, 0}, {s45, s45}, {0, -1}, {-s45, s45}};
  float sum = 0;
  const double y0 = -(x + 3) * M_PI * 0.25, s0 = sin(y0), c0 = cos(y0);
  for (int i = 0; i < 8; i++) {
    const float y0_ = (x + 3 - i);
    if (fabs(y0_) >= 1e-6f) {
      const double y = -y0_ * M_PI * 0.25;
      coeffs[i] = (float)((cs[i][0] * s0 + cs[i][1] * c0) / (y * y));
    } else {
      coeffs[i] = 1e30f;
    }
    sum += coeffs[i];
  }
  sum = 1.f / sum;
  for (int i = 0; i < 8; i++) coeffs[i] *= sum;
}
Taps cv_lanczos4_taps(int in_size, int out_size) {
  Taps t;
  t.ksize = 8;
  t.n_out = out_size;
  t.idx.resize((size_t)out_size * 8);
  t.coef.resize((size_t)out_size * 8);
  const double scale = 1.0 / ((double)out_size / in_size);
  for (int d = 0; d < out_size; ++d) {
    float fx = (float)((d + 0.5) * scale - 0.5);
    const int sx = (int)floorf(fx);
    fx -= sx;
    float c[8];
    cv_lanczos4_coeffs(fx, c);
    for (int k = 0; k < 8; ++k) {
      int s = sx - 3 + k;                                // BORDER_REPLICATE outside the image
      s = s < 0 ? 0 : (s >= in_size ? in_size - 1 : s);
      t.idx[8 * d + k] = s;
      t.coef[8 * d + k] = (double)c[k];
    }
  }
  return t;
}

// ---- device: one separable pass ----------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ double px_f64(T v) { return (double)v; }
template <typename T> __device__ __forceinline__ float px_f32(T v) { return (float)v; }

// horizontal: out[y][xx] = sum_t in[y][idx[xx][t]] * coef[xx][t]; PIL = double accumulate, float32 result
template <typename TIn, bool PIL>
__global__ void __launch_bounds__(256) resample_h(const TIn* __restrict__ in, int H, int W, const int* __restrict__ idx,
                                                  const double* __restrict__ coef, int ksize, int Wout,
                                                  float* __restrict__ out) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (xx >= Wout) return;
  const TIn* row = in + (int64_t)y * W;
  const int* ix = idx + (int64_t)xx * ksize;
  const double* k = coef + (int64_t)xx * ksize;
  if (PIL) {
    double ss = 0.0;
    for (int t = 0; t < ksize; ++t) {
      const double c = k[t];
      if (c != 0.0 || t == 0) ss = __dadd_rn(ss, __dmul_rn(px_f64(row[ix[t]]), c));
    }
    out[(int64_t)y * Wout + xx] = __double2float_rn(ss);
  } else {
    float ss = __fmul_rn(px_f32(row[ix[0]]), (float)k[0]);
    for (int t = 1; t < ksize; ++t) ss = __fadd_rn(ss, __fmul_rn(px_f32(row[ix[t]]), (float)k[t]));
    out[(int64_t)y * Wout + xx] = ss;
  }
}
template <typename TOut> __device__ __forceinline__ TOut cv_cast(float v);
template <> __device__ __forceinline__ float cv_cast<float>(float v) { return v; }
template <> __device__ __forceinline__ uint8_t cv_cast<uint8_t>(float v) {
  int i = __float2int_rn(v);
  return (uint8_t)(i < 0 ? 0 : (i > 255 ? 255 : i));
}
template <> __device__ __forceinline__ uint16_t cv_cast<uint16_t>(float v) {
  int i = __float2int_rn(v);
  return (uint16_t)(i < 0 ? 0 : (i > 65535 ? 65535 : i));
}
// vertical: out[yy][x] = sum_t mid[idx[yy][t]][x] * coef[yy][t]
template <typename TOut, bool PIL>
__global__ void __launch_bounds__(256) resample_v(const float* __restrict__ mid, int H, int Wout, const int* __restrict__ idx,
                                                  const double* __restrict__ coef, int ksize, int Hout,
                                                  TOut* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, yy = blockIdx.y;
  if (x >= Wout) return;
  const int* ix = idx + (int64_t)yy * ksize;
  const double* k = coef + (int64_t)yy * ksize;
  if (PIL) {
    double ss = 0.0;
    for (int t = 0; t < ksize; ++t) {
      const double c = k[t];
      if (c != 0.0 || t == 0) ss = __dadd_rn(ss, __dmul_rn((double)mid[(int64_t)ix[t] * Wout + x], c));
    }
    out[(int64_t)yy * Wout + x] = (TOut)__double2float_rn(ss);
  } else {
    float ss = __fmul_rn(mid[(int64_t)ix[0] * Wout + x], (float)k[0]);
    for (int t = 1; t < ksize; ++t) ss = __fadd_rn(ss, __fmul_rn(mid[(int64_t)ix[t] * Wout + x], (float)k[t]));
    out[(int64_t)yy * Wout + x] = cv_cast<TOut>(ss);
  }
}
// uint8 images: OpenCV's own fixed-point loops (INTER_RESIZE_COEF_BITS = 11): integer taps cvRound(c * 2048), int32 rows
__global__ void __launch_bounds__(256) resample_h_u8fix(const uint8_t* __restrict__ in, int H, int W, const int* __restrict__ idx,
                                                        const double* __restrict__ coef, int ksize, int Wout,
                                                        int* __restrict__ out) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (xx >= Wout) return;
  const uint8_t* row = in + (int64_t)y * W;
  int ss = 0;
  for (int t = 0; t < ksize; ++t) ss += (int)row[idx[(int64_t)xx * ksize + t]] * (int)coef[(int64_t)xx * ksize + t];
  out[(int64_t)y * Wout + xx] = ss;
}
__global__ void __launch_bounds__(256) resample_v_u8fix(const int* __restrict__ mid, int Wout, const int* __restrict__ idx,
                                                        const double* __restrict__ coef, int ksize, int Hout,
                                                        uint8_t* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, yy = blockIdx.y;
  if (x >= Wout) return;
  const int* ix = idx + (int64_t)yy * ksize;
  const double* k = coef + (int64_t)yy * ksize;
  int v;
  if (ksize == 2) {      // VResizeLinear<uchar, int, short>: the form OpenCV's SIMD code fixes for scalar code too
    const int s0 = mid[(int64_t)ix[0] * Wout + x], s1 = mid[(int64_t)ix[1] * Wout + x];
    v = ((((int)k[0] * (s0 >> 4)) >> 16) + (((int)k[1] * (s1 >> 4)) >> 16) + 2) >> 2;
  } else {               // FixedPtCast<int, uchar, 22>
    int ss = 0;
    for (int t = 0; t < ksize; ++t) ss += mid[(int64_t)ix[t] * Wout + x] * (int)k[t];
    v = (ss + (1 << 21)) >> 22;
  }
  out[(int64_t)yy * Wout + x] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

template <typename T>
__global__ void __launch_bounds__(256) nearest_gather(const T* __restrict__ in, int W, const int* __restrict__ xtab,
                                                      const int* __restrict__ ytab, int Hout, int Wout, T* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= Wout) return;
  const int xi = xtab[x], yi = ytab[y];
  out[(int64_t)y * Wout + x] = (xi >= 0 && yi >= 0) ? in[(int64_t)yi * W + xi] : (T)0;     // Pillow fills with 0
}

template <typename V>
int upload(tmat_ctx* ctx, const std::vector<V>& h, V** d, cudaStream_t st) {
  TMAT_TRY(arena_get(ctx, h.size(), d));
  // pageable source: the runtime stages the bytes before returning, so `h` may go out of scope afterwards
  CUDA_TRY(cudaMemcpyAsync(*d, h.data(), h.size() * sizeof(V), cudaMemcpyHostToDevice, st));
  return TMAT_OK;
}

template <typename T, bool PIL>
int resample_typed(tmat_ctx* ctx, const T* in, int H, int W, T* out, int Hout, int Wout, const Taps& tx, const Taps& ty,
                   cudaStream_t st) {
  int *ixx, *ixy;
  double *cx, *cy;
  float* mid;
  TMAT_TRY(upload(ctx, tx.idx, &ixx, st));
  TMAT_TRY(upload(ctx, tx.coef, &cx, st));
  TMAT_TRY(upload(ctx, ty.idx, &ixy, st));
  TMAT_TRY(upload(ctx, ty.coef, &cy, st));
  TMAT_TRY(arena_get(ctx, (size_t)H * Wout, &mid));
  resample_h<T, PIL><<<dim3(cdiv(Wout, 256), H), 256, 0, st>>>(in, H, W, ixx, cx, tx.ksize, Wout, mid);
  KERNEL_CHECK(ctx);
  resample_v<T, PIL><<<dim3(cdiv(Wout, 256), Hout), 256, 0, st>>>(mid, H, Wout, ixy, cy, ty.ksize, Hout, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

}  // namespace

extern "C" int tmat_resize_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W, void* out_dev, int Hout,
                               int Wout, int mode, void* stream) {
  TMAT_REQUIRE(ctx && img_dev && out_dev, "tmat_resize: NULL argument");
  TMAT_REQUIRE(H >= 1 && W >= 1 && Hout >= 1 && Wout >= 1, "tmat_resize: bad shape %dx%d -> %dx%d", H, W, Hout, Wout);
  TMAT_REQUIRE(mode >= TMAT_RESIZE_PIL_LANCZOS && mode <= TMAT_RESIZE_CV_LANCZOS4, "tmat_resize: unknown mode %d", mode);
  TMAT_TRY(arena_reset(ctx));
  cudaStream_t st = as_stream(stream);
  if (mode == TMAT_RESIZE_PIL_NEAREST) {
    TMAT_REQUIRE(dtype == TMAT_F32, "tmat_resize: the Pillow modes work on float32 (mode 'F') images");
    int *xt, *yt;
    TMAT_TRY(upload(ctx, pil_nearest_index(W, Wout), &xt, st));
    TMAT_TRY(upload(ctx, pil_nearest_index(H, Hout), &yt, st));
    nearest_gather<float><<<dim3(cdiv(Wout, 256), Hout), 256, 0, st>>>((const float*)img_dev, W, xt, yt, Hout, Wout,
                                                                      (float*)out_dev);
    KERNEL_CHECK(ctx);
    return TMAT_OK;
  }
  if (mode == TMAT_RESIZE_PIL_LANCZOS) {
    TMAT_REQUIRE(dtype == TMAT_F32, "tmat_resize: the Pillow modes work on float32 (mode 'F') images");
    // Pillow skips a pass whose size does not change; the taps of an unchanged axis are the identity anyway, but
    // the intermediate float32 rounding must not be applied twice -- identical here because float32 -> float32
    return resample_typed<float, true>(ctx, (const float*)img_dev, H, W, (float*)out_dev, Hout, Wout,
                                       pil_lanczos_taps(W, Wout), pil_lanczos_taps(H, Hout), st);
  }
  const bool lin = (mode == TMAT_RESIZE_CV_LINEAR);
  Taps tx = lin ? cv_linear_taps(W, Wout, dtype == TMAT_U8) : cv_lanczos4_taps(W, Wout);
  Taps ty = lin ? cv_linear_taps(H, Hout, dtype == TMAT_U8) : cv_lanczos4_taps(H, Hout);
  if (dtype == TMAT_U8) {
    for (Taps* t : {&tx, &ty})
      for (double& c : t->coef) {
        const float f = (float)c * 2048.f;                // saturate_cast<short>(c * INTER_RESIZE_COEF_SCALE)
        c = (double)lrintf(f);
      }
    int *ixx, *ixy, *mid;
    double *cx, *cy;
    TMAT_TRY(upload(ctx, tx.idx, &ixx, st));
    TMAT_TRY(upload(ctx, tx.coef, &cx, st));
    TMAT_TRY(upload(ctx, ty.idx, &ixy, st));
    TMAT_TRY(upload(ctx, ty.coef, &cy, st));
    TMAT_TRY(arena_get(ctx, (size_t)H * Wout, &mid));
    resample_h_u8fix<<<dim3(cdiv(Wout, 256), H), 256, 0, st>>>((const uint8_t*)img_dev, H, W, ixx, cx, tx.ksize, Wout, mid);
    KERNEL_CHECK(ctx);
    resample_v_u8fix<<<dim3(cdiv(Wout, 256), Hout), 256, 0, st>>>(mid, Wout, ixy, cy, ty.ksize, Hout, (uint8_t*)out_dev);
    KERNEL_CHECK(ctx);
    return TMAT_OK;
  }
  switch (dtype) {
    case TMAT_U16: return resample_typed<uint16_t, false>(ctx, (const uint16_t*)img_dev, H, W, (uint16_t*)out_dev, Hout, Wout, tx, ty, st);
    case TMAT_F32: return resample_typed<float, false>(ctx, (const float*)img_dev, H, W, (float*)out_dev, Hout, Wout, tx, ty, st);
  }
  tmat_set_error("tmat_resize: dtype %d not supported (uint8 / uint16 / float32)", dtype);
  return TMAT_ERR_UNSUPPORTED;
}

extern "C" int tmat_resize_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W, void* out_host, int Hout,
                                int Wout, int mode) {
  TMAT_REQUIRE(ctx && img_host && out_host, "tmat_resize_host: NULL argument");
  const size_t es = dtype_size(dtype);
  TMAT_REQUIRE(es > 0 && dtype != TMAT_F64, "tmat_resize_host: dtype %d not supported", dtype);
  TMAT_REQUIRE(H >= 1 && W >= 1 && Hout >= 1 && Wout >= 1, "tmat_resize_host: bad shape");
  CUDA_TRY(cudaSetDevice(ctx->device));
  void *d_in = nullptr, *d_out = nullptr;      // outside the arena: tmat_resize_dev resets it
  CUDA_TRY(cudaMalloc(&d_in, (size_t)H * W * es));
  cudaError_t e = cudaMalloc(&d_out, (size_t)Hout * Wout * es);
  if (e != cudaSuccess) { cudaFree(d_in); tmat_set_error("tmat_resize_host: %s", cudaGetErrorString(e)); return TMAT_ERR_NOMEM; }
  int rc = TMAT_OK;
  if (cudaMemcpyAsync(d_in, img_host, (size_t)H * W * es, cudaMemcpyHostToDevice, 0) != cudaSuccess) rc = TMAT_ERR_CUDA;
  if (rc == TMAT_OK) rc = tmat_resize_dev(ctx, d_in, dtype, H, W, d_out, Hout, Wout, mode, nullptr);
  if (rc == TMAT_OK && cudaMemcpyAsync(out_host, d_out, (size_t)Hout * Wout * es, cudaMemcpyDeviceToHost, 0) != cudaSuccess) rc = TMAT_ERR_CUDA;
  if (cudaStreamSynchronize(0) != cudaSuccess && rc == TMAT_OK) rc = TMAT_ERR_CUDA;
  cudaFree(d_in);
  cudaFree(d_out);
  if (rc == TMAT_ERR_CUDA) tmat_set_error("tmat_resize_host: CUDA error %s", cudaGetErrorString(cudaGetLastError()));
  return rc;
}
