// Stage 3: thresholds and the Gaussian pre-filter.
// Reference: scripts/compute_branches.py:334 (pred > 0.5), fl_tissue_model_tools/preprocessing.py:13-23
// (apply_mask), fl_tissue_model_tools/well_mask_generation.py:236-277 (auto_threshold_well) and the
// scikit-image 0.22 calls it makes: filters.gaussian (= scipy.ndimage.gaussian_filter, mode 'nearest',
// truncate 4), exposure.rescale_intensity, filters.threshold_otsu, morphology.binary_erosion(disk(5)).
// Floating-point steps that feed an integer decision are evaluated in the reference's precision
// with the reference's operation order and no FMA contraction, so the uint8 image is bit-identical.
#include "common.cuh"
#include <stdlib.h>
#include <math.h>
#include <algorithm>
#include <type_traits>

int erode_disk_thr(tmat_ctx* ctx, const uint8_t* in, int H, int W, int r, int border_value, const int* thr_params_dev,
                   uint8_t* out, cudaStream_t st);

namespace {

__global__ void __launch_bounds__(256) threshold_gt_f64(const double* __restrict__ x, int64_t n, double thr,
                                                        uint8_t* __restrict__ mask) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) mask[i] = x[i] > thr ? 1 : 0;
}

__global__ void __launch_bounds__(256) apply_mask_f32(const float* __restrict__ img, const uint8_t* __restrict__ m,
                                                      int64_t n, float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = m[i] ? img[i] : 0.f;
}

// ---- separable Gaussian, scipy.ndimage.correlate1d semantics -----------------------------------
// symmetric kernel:  tmp = x[l]*w[0];  for j = -R..-1: tmp += (x[l+j] + x[l-j]) * w[j]   (double, no FMA)
// mode 'nearest' = index clamp.  Pass 1 filters along axis 0 (rows), pass 2 along axis 1, as
// gaussian_filter does; a float32 image is rounded to float32 between the passes (scipy writes
// the intermediate into the float32 output array).
constexpr int GAUSS_MAX_R = 32;
struct GaussW { double w[GAUSS_MAX_R + 1]; int R; };   // w[k] = weight at distance k

template <typename TIn>
__device__ __forceinline__ double load_as_f64(const TIn* p, int64_t i, double int_scale);
template <> __device__ __forceinline__ double load_as_f64<uint8_t>(const uint8_t* p, int64_t i, double s) { return __dmul_rn((double)p[i], s); }
template <> __device__ __forceinline__ double load_as_f64<uint16_t>(const uint16_t* p, int64_t i, double s) { return __dmul_rn((double)p[i], s); }
template <> __device__ __forceinline__ double load_as_f64<float>(const float* p, int64_t i, double) { return (double)p[i]; }
template <> __device__ __forceinline__ double load_as_f64<double>(const double* p, int64_t i, double) { return p[i]; }

// vertical pass (axis 0): thread per pixel, coalesced along x
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) gauss_axis0(const TIn* __restrict__ in, int H, int W, GaussW g, double int_scale,
                                                   TOut* __restrict__ out) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  int y = blockIdx.y;
  if (x >= W) return;
  double tmp = __dmul_rn(load_as_f64<TIn>(in, (int64_t)y * W + x, int_scale), g.w[0]);
  for (int j = g.R; j >= 1; --j) {
    int ya = max(y - j, 0), yb = min(y + j, H - 1);
    double a = load_as_f64<TIn>(in, (int64_t)ya * W + x, int_scale);
    double b = load_as_f64<TIn>(in, (int64_t)yb * W + x, int_scale);
    tmp = __dadd_rn(tmp, __dmul_rn(__dadd_rn(a, b), g.w[j]));
  }
  out[(int64_t)y * W + x] = (TOut)tmp;
}

// ---- min / max of a float image (exact), via order-preserving integer keys -------------------------
__device__ __forceinline__ unsigned long long f64_key(double d) {
  unsigned long long b = (unsigned long long)__double_as_longlong(d);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_unkey(unsigned long long k) {
  return __longlong_as_double((long long)((k >> 63) ? (k & 0x7fffffffffffffffull) : ~k));
}
// horizontal pass (axis 1): one CTA per row segment, row staged in shared memory with its halo; mm != nullptr also
// reduces the exact min / max of the OUTPUT into mm[0], mm[1] (order-preserving keys): auto_threshold_well needs them
// for rescale_intensity and a separate pass over the blurred image cost 20 us per 2048^2
template <typename T>
__global__ void __launch_bounds__(256) gauss_axis1(const T* __restrict__ in, int H, int W, GaussW g, T* __restrict__ out,
                                                   unsigned long long* __restrict__ mm) {
  __shared__ double s[256 + 2 * GAUSS_MAX_R];
  int y = blockIdx.y;
  int x0 = blockIdx.x * 256;
  const T* row = in + (int64_t)y * W;
  for (int i = threadIdx.x; i < 256 + 2 * g.R; i += 256) {
    int xx = min(max(x0 - g.R + i, 0), W - 1);
    s[i] = (double)row[xx];
  }
  __syncthreads();
  int x = x0 + threadIdx.x;
  unsigned long long lo = ~0ull, hi = 0ull;
  if (x < W) {
    int c = threadIdx.x + g.R;
    double tmp = __dmul_rn(s[c], g.w[0]);
    for (int j = g.R; j >= 1; --j) tmp = __dadd_rn(tmp, __dmul_rn(__dadd_rn(s[c - j], s[c + j]), g.w[j]));
    const T v = (T)tmp;
    out[(int64_t)y * W + x] = v;
    lo = hi = f64_key((double)v);
  }
  if (mm == nullptr) return;
  __shared__ unsigned long long s_lo[8], s_hi[8];
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
    atomicMin(mm, lo);
    atomicMax(mm + 1, hi);
  }
}

// ---- the same two passes for a compile-time radius (sigma = 1 -> R = 4, what auto_threshold_well uses) -----------
// ncu (round 2, 2048^2 uint16): the per-pixel kernels above spend their time on 2R+1 integer->float64 conversions per
// output (axis 0) and on 16 384 one-row CTAs of load -> barrier -> compute with two same-address atomics each (axis 1:
// 27 us, 40 us with the min/max).  Here a thread slides down GROWS rows (each input converted once: (GROWS + 2R) / GROWS
// conversions per output) resp. along 4 columns x 4 rows with every load issued before the first use, and a CTA
// issues one pair of atomics.  Per-output arithmetic and its order are unchanged (bit-identical results).
constexpr int GROWS = 8;
template <typename TIn, typename TOut, int R>
__global__ void __launch_bounds__(256) gauss_axis0_fixed(const TIn* __restrict__ in, int H, int W, GaussW g, double int_scale,
                                                         TOut* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y0 = blockIdx.y * GROWS;
  if (x >= W) return;
  double v[GROWS + 2 * R];
#pragma unroll
  for (int i = 0; i < GROWS + 2 * R; ++i) {
    const int yy = min(max(y0 - R + i, 0), H - 1);
    v[i] = load_as_f64<TIn>(in, (int64_t)yy * W + x, int_scale);
  }
#pragma unroll
  for (int k = 0; k < GROWS; ++k) {
    if (y0 + k >= H) break;
    double tmp = __dmul_rn(v[k + R], g.w[0]);
#pragma unroll
    for (int j = R; j >= 1; --j) tmp = __dadd_rn(tmp, __dmul_rn(__dadd_rn(v[k + R - j], v[k + R + j]), g.w[j]));
    out[(int64_t)(y0 + k) * W + x] = (TOut)tmp;
  }
}
constexpr int GX = 4, GY = 4;                  // outputs per thread of the horizontal pass: 4 columns x 4 rows
template <typename T, int R>
__global__ void __launch_bounds__(128) gauss_axis1_fixed(const T* __restrict__ in, int H, int W, GaussW g, T* __restrict__ out,
                                                         unsigned long long* __restrict__ mm) {
  const int x0 = (blockIdx.x * 128 + threadIdx.x) * GX;
  const int y0 = blockIdx.y * GY;
  unsigned long long lo = ~0ull, hi = 0ull;
  if (x0 < W) {
    double v[GY][GX + 2 * R];
#pragma unroll
    for (int r = 0; r < GY; ++r) {
      const T* row = in + (int64_t)min(y0 + r, H - 1) * W;
#pragma unroll
      for (int i = 0; i < GX + 2 * R; ++i) v[r][i] = (double)row[min(max(x0 - R + i, 0), W - 1)];
    }
#pragma unroll
    for (int r = 0; r < GY; ++r) {
      if (y0 + r >= H) break;
#pragma unroll
      for (int k = 0; k < GX; ++k) {
        if (x0 + k >= W) break;
        double tmp = __dmul_rn(v[r][k + R], g.w[0]);
#pragma unroll
        for (int j = R; j >= 1; --j) tmp = __dadd_rn(tmp, __dmul_rn(__dadd_rn(v[r][k + R - j], v[r][k + R + j]), g.w[j]));
        const T o = (T)tmp;
        out[(int64_t)(y0 + r) * W + x0 + k] = o;
        const unsigned long long key = f64_key((double)o);
        lo = min(lo, key); hi = max(hi, key);
      }
    }
  }
  if (mm == nullptr) return;
  __shared__ unsigned long long s_lo[4], s_hi[4];
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 4; ++w) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
    if (lo != ~0ull) { atomicMin(mm, lo); atomicMax(mm + 1, hi); }
  }
}

// Both passes in ONE kernel for sigma = 1 on integer images: a CTA of 256 threads takes a 56 x 32-pixel tile, 64 columns
// (tile + 4 halo columns each side, indices clamped = mode 'nearest') x 4 groups of 8 rows run the vertical pass from
// registers into a 32 x 64 float64 tile in shared memory, then every thread runs the horizontal pass for 7 consecutive
// outputs of one row.  The float64 intermediate (33.6 MB at 2048^2, written and re-read by the two-pass version) never
// leaves the SM.  Same operations in the same order per output as the two passes: bit-identical.
// MODE 0: write the blurred image (+ min / max keys when mm != nullptr); MODE 1: min / max only (nothing written);
// MODE 2: rescale_intensity -> uint8 + the five histograms straight from the recomputed blur (needs the keys).
constexpr int G2_TW = 56, G2_TH = 32, G2_PITCH = 66;
struct G2Corners { int xs, xe, ys, ye; };
template <typename TIn, int MODE>
__global__ void __launch_bounds__(256) gauss2d_sigma1(const TIn* __restrict__ in, int H, int W, GaussW g, double int_scale,
                                                      double* __restrict__ out, unsigned long long* __restrict__ mm,
                                                      G2Corners cn, uint8_t* __restrict__ u8, unsigned int* __restrict__ hist) {
  constexpr int R = 4;
  __shared__ double V[G2_TH * G2_PITCH];
  __shared__ unsigned int sh[MODE == 2 ? 5 * 256 : 1];
  __shared__ unsigned long long s_lo[8], s_hi[8];
  const int x0 = blockIdx.x * G2_TW, y0 = blockIdx.y * G2_TH;
  if (MODE == 2) {
    for (int i = threadIdx.x; i < 5 * 256; i += 256) sh[i] = 0;
  }
  {
    // vertical pass: column c of the 64, rows 8 grp .. 8 grp + 7 of the tile
    const int c = threadIdx.x & 63, grp = threadIdx.x >> 6;
    const int xc = min(max(x0 - R + c, 0), W - 1);
    double v[8 + 2 * R];
#pragma unroll
    for (int i = 0; i < 8 + 2 * R; ++i) {
      const int yy = min(max(y0 + 8 * grp - R + i, 0), H - 1);
      v[i] = load_as_f64<TIn>(in, (int64_t)yy * W + xc, int_scale);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      double tmp = __dmul_rn(v[k + R], g.w[0]);
#pragma unroll
      for (int j = R; j >= 1; --j) tmp = __dadd_rn(tmp, __dmul_rn(__dadd_rn(v[k + R - j], v[k + R + j]), g.w[j]));
      V[(8 * grp + k) * G2_PITCH + c] = tmp;
    }
  }
  __syncthreads();
  unsigned long long lo = ~0ull, hi = 0ull;
  double klo = 0.0, khi = 0.0;
  bool flat = false;
  if (MODE == 2) { klo = f64_unkey(mm[0]); khi = f64_unkey(mm[1]); flat = (klo == khi); }
  {
    // horizontal pass: row r of the tile, 7 consecutive outputs
    const int r = threadIdx.x >> 3, q = threadIdx.x & 7;
    const int y = y0 + r;
    const double* vr = V + r * G2_PITCH + 7 * q;          // vr[i] = column x0 + 7 q - R + i
    double v[7 + 2 * R];
#pragma unroll
    for (int i = 0; i < 7 + 2 * R; ++i) v[i] = vr[i];
    const bool top = y < cn.xs, bot = y >= cn.xe;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const int x = x0 + 7 * q + k;
      int qv = -1;
      if (y < H && x < W) {
        double tmp = __dmul_rn(v[k + R], g.w[0]);
#pragma unroll
        for (int j = R; j >= 1; --j) tmp = __dadd_rn(tmp, __dmul_rn(__dadd_rn(v[k + R - j], v[k + R + j]), g.w[j]));
        if (MODE == 0) out[(int64_t)y * W + x] = tmp;
        if (MODE != 2) {
          const unsigned long long key = f64_key(tmp);
          lo = min(lo, key); hi = max(hi, key);
        } else {
          double d = tmp;
          if (!flat) d = __ddiv_rn(__dsub_rn(d, klo), __dsub_rn(khi, klo));
          const double yv = __dadd_rn(__dmul_rn(d, 255.0), 0.0);
          const uint8_t qq = (uint8_t)(long long)yv;
          u8[(int64_t)y * W + x] = qq;
          qv = qq;
          if (top && x < cn.ys) atomicAdd(&sh[0 * 256 + qv], 1u);
          if (top && x >= cn.ye) atomicAdd(&sh[1 * 256 + qv], 1u);
          if (bot && x < cn.ys) atomicAdd(&sh[2 * 256 + qv], 1u);
          if (bot && x >= cn.ye) atomicAdd(&sh[3 * 256 + qv], 1u);
        }
      }
      if (MODE == 2) {
        const unsigned int peers = __match_any_sync(0xffffffffu, qv);
        if (qv >= 0 && (threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&sh[4 * 256 + qv], (unsigned int)__popc(peers));
      }
    }
  }
  if (MODE == 2) {
    __syncthreads();
    for (int i = threadIdx.x; i < 5 * 256; i += 256)
      if (sh[i]) atomicAdd(&hist[i], sh[i]);
    return;
  }
  if (mm == nullptr) return;
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
    if (lo != ~0ull) { atomicMin(mm, lo); atomicMax(mm + 1, hi); }
  }
}

// min / max keys and the 5 x 256 histogram counters of auto_threshold_well, initialised in one launch
__global__ void __launch_bounds__(256) minmax64_hist_init(unsigned long long* k, unsigned int* hist) {
  if (threadIdx.x == 0) { k[0] = ~0ull; k[1] = 0ull; }
  for (int i = threadIdx.x; i < 5 * 256; i += 256) hist[i] = 0;
}
// rescale_intensity(x, out_range=(0,255)).astype(uint8)  +  the 4 corner histograms + global histogram
//   y = (x - lo) / (hi - lo);  y = y * 255 + 0;  uint8(trunc(y))      in T's precision (float64 / float32)
template <typename T>
__global__ void __launch_bounds__(256) rescale_u8_hist(const T* __restrict__ x, int H, int W,
                                                       const unsigned long long* __restrict__ keys, int xs, int xe,
                                                       int ys, int ye, uint8_t* __restrict__ u8,
                                                       unsigned int* __restrict__ hist /*[5][256]*/) {
  __shared__ unsigned int sh[5 * 256];
  for (int i = threadIdx.x; i < 5 * 256; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const T lo = (T)f64_unkey(keys[0]), hi = (T)f64_unkey(keys[1]);
  const bool flat = (lo == hi);
  const int lane = threadIdx.x & 31;
  // a block walks whole rows (no 64-bit index division per pixel); the global histogram takes one shared-memory atomic
  // per distinct value per warp (neighbouring pixels of a blurred image mostly share a handful of values)
  for (int r = blockIdx.x; r < H; r += gridDim.x) {
    const T* xr = x + (int64_t)r * W;
    uint8_t* ur = u8 + (int64_t)r * W;
    const bool top = r < xs, bot = r >= xe;
    for (int cb = 0; cb < W; cb += 4 * 256) {
      T raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {             // four independent loads in flight before the first division
        const int c = cb + u * 256 + threadIdx.x;
        raw[u] = c < W ? xr[c] : (T)0;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c0 = cb + u * 256;
        if (c0 >= W) break;                     // block-uniform
        const int c = c0 + threadIdx.x;
        int q = -1;
        if (c < W) {
          T y;
          if (sizeof(T) == 8) {
            double d = (double)raw[u];
            if (!flat) d = __ddiv_rn(__dsub_rn(d, (double)lo), __dsub_rn((double)hi, (double)lo));
            y = (T)__dadd_rn(__dmul_rn(d, 255.0), 0.0);
          } else {
            float f = (float)raw[u];
            if (!flat) f = __fdiv_rn(__fsub_rn(f, (float)lo), __fsub_rn((float)hi, (float)lo));
            y = (T)__fadd_rn(__fmul_rn(f, 255.0f), 0.0f);
          }
          const uint8_t qq = (uint8_t)(long long)y;
          ur[c] = qq;
          q = qq;
          if (top && c < ys) atomicAdd(&sh[0 * 256 + q], 1u);
          if (top && c >= ye) atomicAdd(&sh[1 * 256 + q], 1u);
          if (bot && c < ys) atomicAdd(&sh[2 * 256 + q], 1u);
          if (bot && c >= ye) atomicAdd(&sh[3 * 256 + q], 1u);
        }
        const unsigned int peers = __match_any_sync(0xffffffffu, q);
        if (q >= 0 && lane == __ffs(peers) - 1) atomicAdd(&sh[4 * 256 + q], (unsigned int)__popc(peers));
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 5 * 256; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// np.median of a uint8 block from its 256-bin histogram (shared memory), computed by one warp: every lane owns 8
// consecutive bins, a shuffle scan gives the counts below each lane, and the lane in which the cumulative count
// first exceeds k walks its 8 bins.  Returns (v0 + v1) / 2 for the two middle order statistics (all lanes).
__device__ __forceinline__ int warp_hist_select(const unsigned int* h, unsigned long long below, unsigned long long mine,
                                                unsigned long long k, int lane) {
  // lane owning the k-th element (0-based): below <= k < below + mine
  const bool own = mine > 0 && below <= k && k < below + mine;
  int v = -1;
  if (own) {
    unsigned long long acc = below;
    for (int i = 0; i < 8; ++i) {
      acc += h[lane * 8 + i];
      if (acc > k) { v = lane * 8 + i; break; }
    }
  }
  const unsigned int who = __ballot_sync(0xffffffffu, own);
  return __shfl_sync(0xffffffffu, v, who ? __ffs(who) - 1 : 0);
}
__device__ double warp_hist_median(const unsigned int* h, int lane) {
  unsigned long long mine = 0;
  for (int i = 0; i < 8; ++i) mine += h[lane * 8 + i];
  unsigned long long incl = mine;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  const unsigned long long n = __shfl_sync(0xffffffffu, incl, 31), below = incl - mine;
  if (n == 0) return __longlong_as_double(0x7ff8000000000000ll);
  const int v0 = warp_hist_select(h, below, mine, (n - 1) / 2, lane);
  const int v1 = warp_hist_select(h, below, mine, n / 2, lane);
  return __ddiv_rn(__dadd_rn((double)v0, (double)v1), 2.0);
}
// Corner medians, invert decision and skimage.filters.threshold_otsu (well_mask_generation.py:248-272) on the device:
// 5 x 256 counters in, four ints out -- the host no longer waits for the histogram (round 1 synchronised here).
// One block of 256 threads, one bin each.  numpy's dtype semantics (float32 counts and class weights, float64 class
// means) are kept, and the cumulative sums may be taken in any order because they are exact: every partial count is an
// integer below 2^24 (n_pixels <= 2^24, checked by the caller) and every partial count*value sum below 2^53.
__global__ void __launch_bounds__(256) otsu_decide_kernel(const unsigned int* __restrict__ hist /*[5][256]*/,
                                                          int* __restrict__ params) {
  __shared__ unsigned int s_cnt[256];
  __shared__ unsigned long long s_a[256], s_b[256];
  __shared__ double s_v[256];
  __shared__ int s_i[256];
  __shared__ double s_med[4];
  __shared__ int s_misc[4];
  __shared__ unsigned int s_h[5 * 256];
  const int t = threadIdx.x;
  for (int i = t; i < 5 * 256; i += 256) s_h[i] = hist[i];
  if (t == 0) { s_misc[0] = 255; s_misc[1] = 0; }
  __syncthreads();
  const unsigned int* g = s_h + 4 * 256;
  if (t < 128) {                                   // warps 0..3: one corner block each
    const double m = warp_hist_median(s_h + (t >> 5) * 256, t & 31);
    if ((t & 31) == 0) s_med[t >> 5] = m;
  }
  if (g[t] != 0) { atomicMin(&s_misc[0], t); atomicMax(&s_misc[1], t); }      // occupied range of the image (n >= 400 pixels)
  __syncthreads();
  const int vmin = s_misc[0], vmax = s_misc[1];
  const double mmin = fmin(fmin(s_med[0], s_med[1]), fmin(s_med[2], s_med[3]));
  const double mmax = fmax(fmax(s_med[0], s_med[1]), fmax(s_med[2], s_med[3]));
  const int invert = fabs(__dsub_rn((double)vmin, mmin)) > fabs(__dsub_rn((double)vmax, mmax)) ? 1 : 0;
  // histogram of the (possibly inverted) image; its occupied range is [lo, hi]
  const int lo = invert ? 255 - vmax : vmin, hi = invert ? 255 - vmin : vmax;
  const unsigned int cnt = g[invert ? 255 - t : t];
  s_cnt[t] = cnt;
  // inclusive prefix sums of count and count*value (exact integers)
  s_a[t] = cnt;
  s_b[t] = (unsigned long long)cnt * (unsigned long long)t;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {
    const unsigned long long a = t >= o ? s_a[t - o] : 0ull, b2 = t >= o ? s_b[t - o] : 0ull;
    __syncthreads();
    s_a[t] += a; s_b[t] += b2;
    __syncthreads();
  }
  const unsigned long long tot_a = s_a[255], tot_b = s_b[255];
  // candidate split after bin t (lo <= t < hi): class 1 = bins lo..t, class 2 = bins t+1..hi
  double v = -2.0;
  if (lo < hi && t >= lo && t < hi) {
    const float w1 = (float)s_a[t], w2 = (float)(tot_a - s_a[t]);
    const double m1 = __ddiv_rn((double)s_b[t], (double)w1);
    const double m2 = __ddiv_rn((double)(tot_b - s_b[t]), (double)w2);
    const float ww = __fmul_rn(w1, w2);
    const double d = __dsub_rn(m1, m2);
    v = __dmul_rn((double)ww, __dmul_rn(d, d));
  }
  s_v[t] = v; s_i[t] = t;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {           // first maximum wins (np.argmax)
    if (t < o) {
      const double v2 = s_v[t + o];
      const int i2 = s_i[t + o];
      if (v2 > s_v[t] || (v2 == s_v[t] && i2 < s_i[t])) { s_v[t] = v2; s_i[t] = i2; }
    }
    __syncthreads();
  }
  if (t == 0) {
    const int thr = lo >= hi ? (lo > 255 ? 0 : lo) : s_i[0];
    params[0] = invert;
    params[1] = thr;
    params[2] = lo;
    params[3] = hi;
  }
}

}  // namespace

// ---- host side --------------------------------------------------------------------------------
static int make_gauss(double sigma, GaussW* g) {
  TMAT_REQUIRE(sigma > 0, "gaussian: sigma must be positive");
  int R = (int)(4.0 * sigma + 0.5);   // truncate = 4.0
  TMAT_REQUIRE(R >= 1 && R <= GAUSS_MAX_R, "gaussian: radius %d outside [1,%d]", R, GAUSS_MAX_R);
  // scipy.ndimage._filters._gaussian_kernel1d: exp(-0.5 / sigma^2 * x^2) / sum  (numpy pairwise sum)
  std::vector<double> phi(2 * R + 1);
  double s2 = sigma * sigma;
  for (int x = -R; x <= R; ++x) phi[x + R] = exp(-0.5 / s2 * (double)(x * x));
  double sum = np_pairwise_sum(phi.data(), 2 * R + 1);
  g->R = R;
  for (int k = 0; k <= R; ++k) g->w[k] = phi[R + k] / sum;
  return TMAT_OK;
}

// which sigma = 1 implementation runs (measurement switches): 0 fused 2-D kernel, 1 the two fixed-radius passes, 2 generic
static int gauss_sigma1_mode() {
  static const int mode = [] {
    const char* a = getenv("TMAT_GAUSS_GENERIC");
    const char* b = getenv("TMAT_GAUSS_TWO_PASS");
    return (a && atoi(a) == 1) ? 2 : ((b && atoi(b) == 1) ? 1 : 0);
  }();
  return mode;
}
// blur: float64 (H,W) for integer input, float32 for float32 input
static int gaussian_impl(tmat_ctx* ctx, const void* img, int dtype, int H, int W, double sigma, void* out,
                         cudaStream_t st, unsigned long long* mm = nullptr) {
  GaussW g;
  TMAT_TRY(make_gauss(sigma, &g));
  dim3 grid0(cdiv(W, 256), H), grid1(cdiv(W, 256), H);
  size_t n = (size_t)H * W;
  if (dtype == TMAT_U8 || dtype == TMAT_U16) {
    double* tmp;
    TMAT_TRY(arena_get(ctx, n, &tmp));
    // skimage img_as_float: np.multiply(image, 1. / imax, dtype=float64)
    const bool generic = gauss_sigma1_mode() == 2, two_pass = gauss_sigma1_mode() == 1;
    if (g.R == 4 && !generic && !two_pass) {
      const dim3 g2(cdiv(W, G2_TW), cdiv(H, G2_TH));
      const G2Corners none{0, 0, 0, 0};
      if (dtype == TMAT_U8) gauss2d_sigma1<uint8_t, 0><<<g2, 256, 0, st>>>((const uint8_t*)img, H, W, g, 1.0 / 255.0, (double*)out, mm, none, nullptr, nullptr);
      else gauss2d_sigma1<uint16_t, 0><<<g2, 256, 0, st>>>((const uint16_t*)img, H, W, g, 1.0 / 65535.0, (double*)out, mm, none, nullptr, nullptr);
      KERNEL_CHECK(ctx);
      return TMAT_OK;
    }
    if (g.R == 4 && !generic) {
      const dim3 gf0(cdiv(W, 256), cdiv(H, GROWS)), gf1(cdiv(W, 128 * GX), cdiv(H, GY));
      if (dtype == TMAT_U8) gauss_axis0_fixed<uint8_t, double, 4><<<gf0, 256, 0, st>>>((const uint8_t*)img, H, W, g, 1.0 / 255.0, tmp);
      else gauss_axis0_fixed<uint16_t, double, 4><<<gf0, 256, 0, st>>>((const uint16_t*)img, H, W, g, 1.0 / 65535.0, tmp);
      KERNEL_CHECK(ctx);
      gauss_axis1_fixed<double, 4><<<gf1, 128, 0, st>>>(tmp, H, W, g, (double*)out, mm);
      KERNEL_CHECK(ctx);
      return TMAT_OK;
    }
    if (dtype == TMAT_U8) gauss_axis0<uint8_t, double><<<grid0, 256, 0, st>>>((const uint8_t*)img, H, W, g, 1.0 / 255.0, tmp);
    else gauss_axis0<uint16_t, double><<<grid0, 256, 0, st>>>((const uint16_t*)img, H, W, g, 1.0 / 65535.0, tmp);
    KERNEL_CHECK(ctx);
    gauss_axis1<double><<<grid1, 256, 0, st>>>(tmp, H, W, g, (double*)out, mm);
    KERNEL_CHECK(ctx);
  } else if (dtype == TMAT_F32) {
    float* tmp;
    TMAT_TRY(arena_get(ctx, n, &tmp));
    gauss_axis0<float, float><<<grid0, 256, 0, st>>>((const float*)img, H, W, g, 1.0, tmp);
    KERNEL_CHECK(ctx);
    gauss_axis1<float><<<grid1, 256, 0, st>>>(tmp, H, W, g, (float*)out, mm);
    KERNEL_CHECK(ctx);
  } else if (dtype == TMAT_F64) {
    double* tmp;
    TMAT_TRY(arena_get(ctx, n, &tmp));
    gauss_axis0<double, double><<<grid0, 256, 0, st>>>((const double*)img, H, W, g, 1.0, tmp);
    KERNEL_CHECK(ctx);
    gauss_axis1<double><<<grid1, 256, 0, st>>>(tmp, H, W, g, (double*)out, mm);
    KERNEL_CHECK(ctx);
  } else {
    tmat_set_error("gaussian: unsupported dtype %d", dtype);
    return TMAT_ERR_UNSUPPORTED;
  }
  return TMAT_OK;
}

extern "C" int tmat_threshold_gt_f64_dev(tmat_ctx* ctx, const double* x_dev, int64_t n, double thr, uint8_t* mask_dev,
                                         void* stream) {
  TMAT_REQUIRE(ctx && x_dev && mask_dev && n >= 0, "tmat_threshold_gt_f64: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (n == 0) return TMAT_OK;
  threshold_gt_f64<<<(unsigned)cdiv64(n, 256), 256, 0, as_stream(stream)>>>(x_dev, n, thr, mask_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

extern "C" int tmat_apply_mask_f32_dev(tmat_ctx* ctx, const float* img_dev, const uint8_t* mask_dev, int64_t n,
                                       float* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && img_dev && mask_dev && out_dev && n >= 0, "tmat_apply_mask_f32: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (n == 0) return TMAT_OK;
  apply_mask_f32<<<(unsigned)cdiv64(n, 256), 256, 0, as_stream(stream)>>>(img_dev, mask_dev, n, out_dev);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

extern "C" int tmat_gaussian_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W, double sigma,
                                 void* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && img_dev && out_dev && H >= 1 && W >= 1, "tmat_gaussian: bad argument");
  TMAT_TRY(arena_reset(ctx));
  return gaussian_impl(ctx, img_dev, dtype, H, W, sigma, out_dev, as_stream(stream));
}

extern "C" int tmat_gaussian_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W, double sigma,
                                  void* out_host) {
  TMAT_REQUIRE(ctx && img_host && out_host && H >= 1 && W >= 1, "tmat_gaussian_host: bad argument");
  TMAT_REQUIRE(dtype_size(dtype) > 0, "tmat_gaussian_host: bad dtype %d", dtype);
  TMAT_TRY(arena_reset(ctx));
  size_t n = (size_t)H * W, es = dtype_size(dtype), os = (dtype == TMAT_F32) ? 4 : 8;
  void *d_in, *d_out;
  TMAT_TRY(arena_alloc(ctx, n * es, &d_in));
  TMAT_TRY(arena_alloc(ctx, n * os, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_in, img_host, n * es, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(gaussian_impl(ctx, d_in, dtype, H, W, sigma, d_out, 0));
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, n * os, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

// ---- auto_threshold_well (well_mask_generation.py:236-277) ------------------------------------------
static int auto_threshold_impl(tmat_ctx* ctx, const void* img, int dtype, int H, int W, uint8_t* mask, int* info,
                               cudaStream_t st) {
  TMAT_REQUIRE(dtype == TMAT_U8 || dtype == TMAT_U16 || dtype == TMAT_F32 || dtype == TMAT_F64,
               "auto_threshold_well: unsupported dtype %d", dtype);
  TMAT_REQUIRE(H >= 20 && W >= 20, "auto_threshold_well: image %dx%d too small for the 5%% corner blocks", H, W);
  TMAT_REQUIRE((int64_t)H * W <= (1ll << 24), "auto_threshold_well: images above 2^24 pixels are not supported (the float32 "
               "class weights of threshold_otsu stop being exact integers)");
  size_t n = (size_t)H * W;
  const bool f32 = (dtype == TMAT_F32);
  unsigned long long* keys;
  unsigned int* hist;
  uint8_t* u8;
  TMAT_TRY(arena_get(ctx, 2, &keys));
  TMAT_TRY(arena_get(ctx, 5 * 256, &hist));
  TMAT_TRY(arena_get(ctx, n, &u8));
  minmax64_hist_init<<<1, 256, 0, st>>>(keys, hist);
  KERNEL_CHECK(ctx);
  int xs = (int)(H * 0.05), xe = (int)(H * 0.95), ys = (int)(W * 0.05), ye = (int)(W * 0.95);
  if ((dtype == TMAT_U8 || dtype == TMAT_U16) && gauss_sigma1_mode() == 0) {
    // integer images: the blurred float64 image is never materialised -- one fused-Gaussian pass reduces its exact
    // min / max, a second one recomputes it and quantises straight to uint8 + histograms (2 x 8.4 MB read and 4.2 MB
    // written per 2048^2 uint16 image instead of 8.4 + 3 x 33.6 MB read and 2 x 33.6 + 4.2 MB written)
    GaussW g;
    TMAT_TRY(make_gauss(1.0, &g));
    const dim3 g2(cdiv(W, G2_TW), cdiv(H, G2_TH));
    const G2Corners none{0, 0, 0, 0}, cn{xs, xe, ys, ye};
    if (dtype == TMAT_U8) {
      gauss2d_sigma1<uint8_t, 1><<<g2, 256, 0, st>>>((const uint8_t*)img, H, W, g, 1.0 / 255.0, nullptr, keys, none, nullptr, nullptr);
      KERNEL_CHECK(ctx);
      gauss2d_sigma1<uint8_t, 2><<<g2, 256, 0, st>>>((const uint8_t*)img, H, W, g, 1.0 / 255.0, nullptr, keys, cn, u8, hist);
    } else {
      gauss2d_sigma1<uint16_t, 1><<<g2, 256, 0, st>>>((const uint16_t*)img, H, W, g, 1.0 / 65535.0, nullptr, keys, none, nullptr, nullptr);
      KERNEL_CHECK(ctx);
      gauss2d_sigma1<uint16_t, 2><<<g2, 256, 0, st>>>((const uint16_t*)img, H, W, g, 1.0 / 65535.0, nullptr, keys, cn, u8, hist);
    }
    KERNEL_CHECK(ctx);
  } else {
    void* blur;
    TMAT_TRY(arena_alloc(ctx, n * (f32 ? 4 : 8), &blur));
    TMAT_TRY(gaussian_impl(ctx, img, dtype, H, W, 1.0, blur, st, keys));     // min / max of the result reduced in its last pass
    const int rows_per_block = cdiv(H, ctx->sm_count * 8);        // rescale_u8_hist walks whole rows: the same number per block
    int blocks = cdiv(H, rows_per_block);
    if (f32) rescale_u8_hist<float><<<blocks, 256, 0, st>>>((const float*)blur, H, W, keys, xs, xe, ys, ye, u8, hist);
    else rescale_u8_hist<double><<<blocks, 256, 0, st>>>((const double*)blur, H, W, keys, xs, xe, ys, ye, u8, hist);
    KERNEL_CHECK(ctx);
  }
  int* params;
  TMAT_TRY(arena_get(ctx, 4, &params));
  otsu_decide_kernel<<<1, 256, 0, st>>>(hist, params);
  KERNEL_CHECK(ctx);
  // '>=' (well_mask_generation.py:274) is folded into the erosion's load
  TMAT_TRY(erode_disk_thr(ctx, u8, H, W, 5, 1, params, mask, st));
  if (info) {      // optional read-back: the only host synchronisation of the call
    int* h = (int*)ctx->pinned;
    CUDA_TRY(cudaMemcpyAsync(h, params, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < 4; ++k) info[k] = h[k];
  }
  return TMAT_OK;
}

extern "C" int tmat_auto_threshold_well_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W,
                                            uint8_t* mask_dev, int* info, void* stream) {
  TMAT_REQUIRE(ctx && img_dev && mask_dev, "tmat_auto_threshold_well: NULL argument");
  TMAT_TRY(arena_reset(ctx));
  return auto_threshold_impl(ctx, img_dev, dtype, H, W, mask_dev, info, as_stream(stream));
}

extern "C" int tmat_auto_threshold_well_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W,
                                             uint8_t* mask_host, int* info) {
  TMAT_REQUIRE(ctx && img_host && mask_host, "tmat_auto_threshold_well_host: NULL argument");
  TMAT_REQUIRE(dtype_size(dtype) > 0, "tmat_auto_threshold_well_host: bad dtype");
  TMAT_TRY(arena_reset(ctx));
  size_t n = (size_t)H * W, es = dtype_size(dtype);
  void* d_in;
  uint8_t* d_mask;
  TMAT_TRY(arena_alloc(ctx, n * es, &d_in));
  TMAT_TRY(arena_get(ctx, n, &d_mask));
  CUDA_TRY(cudaMemcpyAsync(d_in, img_host, n * es, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(auto_threshold_impl(ctx, d_in, dtype, H, W, d_mask, info, 0));
  CUDA_TRY(cudaMemcpyAsync(mask_host, d_mask, n, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

// ---- preprocessing.blur (preprocessing.py:165-183): blur_itr x cv2.GaussianBlur((k,k), 0), then round ----------
// cv2.GaussianBlur with sigma = 0 and k <= 7 uses the fixed kernels [1,2,1]/4, [1,4,6,4,1]/16,
// [0.03125, 0.109375, 0.21875, 0.28125, ...], BORDER_REFLECT_101.  Verified against cv2 4.13 in the build
// container (oracle/blur.py):  uint8  -> exact fixed point, 8-bit weights w = round(256 g), separable,
//                                        one rounding: (S + 32768) >> 16;
//                              float32 -> separable, per pass  acc = g[c] x[c]; acc = fma(g[c+j], x[c-j] + x[c+j], acc)
//                                        (bit-identical for k = 3, within 2 ulp of the image range for k = 5, 7).
namespace {

constexpr int BL_TW = 32, BL_TH = 16, BL_MAXR = 3;

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = (i < 0) ? -i : 2 * (n - 1) - i;
  return i;
}

// final: 0 keep, 1 rint (float32 only), 2 rint -> uint8 (float32 only; written to out_u8)
template <typename T>
__global__ void __launch_bounds__(BL_TW * BL_TH) blur_kernel(const T* __restrict__ in, int H, int W, int r,
                                                              const float* __restrict__ gk, T* __restrict__ out,
                                                              uint8_t* __restrict__ out_u8, int final_op) {
  using A = typename std::conditional<std::is_same<T, uint8_t>::value, int, float>::type;
  __shared__ A tile[BL_TH + 2 * BL_MAXR][BL_TW + 2 * BL_MAXR];
  __shared__ A tmp[BL_TH + 2 * BL_MAXR][BL_TW];
  __shared__ A wk[2 * BL_MAXR + 1];
  const int x0 = blockIdx.x * BL_TW, y0 = blockIdx.y * BL_TH;
  const int tid = threadIdx.y * BL_TW + threadIdx.x;
  const int k = 2 * r + 1;
  if (tid < k) {
    if (std::is_same<T, uint8_t>::value) wk[tid] = (A)__float2int_rn(gk[tid] * 256.f);
    else wk[tid] = (A)gk[tid];
  }
  const int th = BL_TH + 2 * r, tw = BL_TW + 2 * r;
  for (int i = tid; i < th * tw; i += BL_TW * BL_TH) {
    int ty = i / tw, tx = i - ty * tw;
    int yy = reflect101(y0 + ty - r, H), xx = reflect101(x0 + tx - r, W);
    tile[ty][tx] = (A)in[(int64_t)yy * W + xx];
  }
  __syncthreads();
  // horizontal pass over th rows
  for (int i = tid; i < th * BL_TW; i += BL_TW * BL_TH) {
    int ty = i / BL_TW, tx = i - ty * BL_TW;
    A acc;
    if (std::is_same<T, uint8_t>::value) {
      acc = wk[r] * tile[ty][tx + r];
      for (int j = 1; j <= r; ++j) acc += wk[r + j] * (tile[ty][tx + r - j] + tile[ty][tx + r + j]);
    } else {
      acc = (A)__fmul_rn((float)wk[r], (float)tile[ty][tx + r]);
      for (int j = 1; j <= r; ++j)
        acc = (A)__fmaf_rn((float)wk[r + j], __fadd_rn((float)tile[ty][tx + r - j], (float)tile[ty][tx + r + j]), (float)acc);
    }
    tmp[ty][tx] = acc;
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  if (x >= W || y >= H) return;
  const int ty = threadIdx.y, tx = threadIdx.x;
  if (std::is_same<T, uint8_t>::value) {
    int acc = (int)wk[r] * (int)tmp[ty + r][tx];
    for (int j = 1; j <= r; ++j) acc += (int)wk[r + j] * ((int)tmp[ty + r - j][tx] + (int)tmp[ty + r + j][tx]);
    out[(int64_t)y * W + x] = (T)((acc + 32768) >> 16);
  } else {
    float acc = __fmul_rn((float)wk[r], (float)tmp[ty + r][tx]);
    for (int j = 1; j <= r; ++j)
      acc = __fmaf_rn((float)wk[r + j], __fadd_rn((float)tmp[ty + r - j][tx], (float)tmp[ty + r + j][tx]), acc);
    if (final_op >= 1) acc = rintf(acc);          // np.round: half to even
    if (final_op == 2) out_u8[(int64_t)y * W + x] = (uint8_t)(int)acc;
    else out[(int64_t)y * W + x] = (T)acc;
  }
}

__global__ void round_f32_kernel(const float* in, int64_t n, float* out, uint8_t* out_u8) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v = rintf(in[i]);
  if (out_u8) out_u8[i] = (uint8_t)(int)v; else out[i] = v;
}

// cv2.getGaussianKernel(k, 0) for k = 3, 5, 7 (fixed small kernels)
const float BLUR_K3[3] = {0.25f, 0.5f, 0.25f};
const float BLUR_K5[5] = {0.0625f, 0.25f, 0.375f, 0.25f, 0.0625f};
const float BLUR_K7[7] = {0.03125f, 0.109375f, 0.21875f, 0.28125f, 0.21875f, 0.109375f, 0.03125f};

int blur_impl(tmat_ctx* ctx, const void* in, int dtype, int H, int W, int ksize, int iters, int final_op, void* out,
              cudaStream_t st) {
  TMAT_REQUIRE(dtype == TMAT_U8 || dtype == TMAT_F32, "tmat_blur: dtype must be uint8 or float32");
  TMAT_REQUIRE(ksize == 3 || ksize == 5 || ksize == 7, "tmat_blur: k_size %d not supported (3, 5, 7)", ksize);
  TMAT_REQUIRE(iters >= 0 && final_op >= 0 && final_op <= 2, "tmat_blur: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  const int64_t n = (int64_t)H * W;
  const size_t es = dtype_size(dtype);
  const float* hk = ksize == 3 ? BLUR_K3 : (ksize == 5 ? BLUR_K5 : BLUR_K7);
  float* gk;
  TMAT_TRY(arena_get(ctx, 8, &gk));
  CUDA_TRY(cudaMemcpyAsync(gk, hk, ksize * sizeof(float), cudaMemcpyHostToDevice, st));
  void *b0, *b1;
  TMAT_TRY(arena_alloc(ctx, n * es, &b0));
  TMAT_TRY(arena_alloc(ctx, n * es, &b1));
  dim3 grid(cdiv(W, BL_TW), cdiv(H, BL_TH)), block(BL_TW, BL_TH);
  const void* src = in;
  for (int it = 0; it < iters; ++it) {
    const bool last = (it == iters - 1);
    void* dst;
    uint8_t* dst_u8 = nullptr;
    if (!last) dst = (src == b0) ? b1 : b0;
    else if (dtype == TMAT_F32 && final_op == 2) { dst = (src == b0) ? b1 : b0; dst_u8 = (uint8_t*)out; }
    else dst = out;
    if (dtype == TMAT_U8)
      blur_kernel<uint8_t><<<grid, block, 0, st>>>((const uint8_t*)src, H, W, ksize / 2, gk, (uint8_t*)dst, nullptr, 0);
    else
      blur_kernel<float><<<grid, block, 0, st>>>((const float*)src, H, W, ksize / 2, gk, (float*)dst, dst_u8,
                                                 last ? final_op : 0);
    KERNEL_CHECK(ctx);
    src = dst;
  }
  if (iters == 0) {   // no blur: copy (+ round for float32)
    if (dtype == TMAT_U8 || final_op == 0) CUDA_TRY(cudaMemcpyAsync(out, in, n * es, cudaMemcpyDeviceToDevice, st));
    else {
      round_f32_kernel<<<(unsigned)cdiv64(n, 256), 256, 0, st>>>((const float*)in, n, (float*)out,
                                                                  final_op == 2 ? (uint8_t*)out : nullptr);
      KERNEL_CHECK(ctx);
    }
  }
  return TMAT_OK;
}

}  // namespace

extern "C" int tmat_blur_dev(tmat_ctx* ctx, const void* img_dev, int dtype, int H, int W, int ksize, int iters,
                             int final_op, void* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && img_dev && out_dev && H >= 1 && W >= 1, "tmat_blur: bad argument");
  TMAT_TRY(arena_reset(ctx));
  return blur_impl(ctx, img_dev, dtype, H, W, ksize, iters, final_op, out_dev, as_stream(stream));
}

extern "C" int tmat_blur_host(tmat_ctx* ctx, const void* img_host, int dtype, int H, int W, int ksize, int iters,
                              int final_op, void* out_host) {
  TMAT_REQUIRE(ctx && img_host && out_host && H >= 1 && W >= 1, "tmat_blur_host: bad argument");
  TMAT_REQUIRE(dtype == TMAT_U8 || dtype == TMAT_F32, "tmat_blur_host: dtype must be uint8 or float32");
  TMAT_TRY(arena_reset(ctx));
  size_t n = (size_t)H * W, es = dtype_size(dtype), os = (dtype == TMAT_F32 && final_op == 2) ? 1 : es;
  void *d_in, *d_out;
  TMAT_TRY(arena_alloc(ctx, n * es, &d_in));
  TMAT_TRY(arena_alloc(ctx, n * es, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_in, img_host, n * es, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(blur_impl(ctx, d_in, dtype, H, W, ksize, iters, final_op, d_out, 0));
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, n * os, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}
