// Stage 1: Z-stack projections (reference: fl_tissue_model_tools/zstacks.py:134-249).
//   proj_max:221  proj_min:237  proj_avg:192  proj_med:206  proj_focus_stacking:153 (+ _blur_and_lap:134)
// All kernels are HBM-bandwidth bound streaming reductions over Z: every voxel is read once
// (focus stacking re-reads a 4-pixel halo per tile), one output pixel is written.
#include "common.cuh"
#include <algorithm>
#include <stdlib.h>

namespace {

// ------------------------------------------------------------------------------------------
// min / max: 16-byte vector loads, packed SIMD-in-word compare (2 x u16 or 4 x u8 per op).
// ------------------------------------------------------------------------------------------
template <typename T, bool IS_MAX>
__device__ __forceinline__ uint32_t packed_sel(uint32_t a, uint32_t b);
template <> __device__ __forceinline__ uint32_t packed_sel<uint16_t, true>(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
template <> __device__ __forceinline__ uint32_t packed_sel<uint16_t, false>(uint32_t a, uint32_t b) { return __vminu2(a, b); }
template <> __device__ __forceinline__ uint32_t packed_sel<uint8_t, true>(uint32_t a, uint32_t b) { return __vmaxu4(a, b); }
template <> __device__ __forceinline__ uint32_t packed_sel<uint8_t, false>(uint32_t a, uint32_t b) { return __vminu4(a, b); }

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <typename T, bool IS_MAX>
__global__ void __launch_bounds__(256) zproj_minmax_vec(const uint4* __restrict__ stack, int Z,
                                                        int64_t n_vec, uint4* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_vec) return;
  uint4 acc = ld_stream(stack + i);
  int z = 1;
  for (; z + 4 <= Z; z += 4) {
    uint4 v0 = ld_stream(stack + (int64_t)(z + 0) * n_vec + i);
    uint4 v1 = ld_stream(stack + (int64_t)(z + 1) * n_vec + i);
    uint4 v2 = ld_stream(stack + (int64_t)(z + 2) * n_vec + i);
    uint4 v3 = ld_stream(stack + (int64_t)(z + 3) * n_vec + i);
#define SEL4(f) acc.f = packed_sel<T, IS_MAX>(packed_sel<T, IS_MAX>(acc.f, v0.f), packed_sel<T, IS_MAX>(v1.f, packed_sel<T, IS_MAX>(v2.f, v3.f)))
    SEL4(x); SEL4(y); SEL4(z); SEL4(w);
#undef SEL4
  }
  for (; z < Z; ++z) {
    uint4 v = ld_stream(stack + (int64_t)z * n_vec + i);
    acc.x = packed_sel<T, IS_MAX>(acc.x, v.x);
    acc.y = packed_sel<T, IS_MAX>(acc.y, v.y);
    acc.z = packed_sel<T, IS_MAX>(acc.z, v.z);
    acc.w = packed_sel<T, IS_MAX>(acc.w, v.w);
  }
  out[i] = acc;
}

// scalar path: any type, any alignment
template <typename T, bool IS_MAX>
__global__ void __launch_bounds__(256) zproj_minmax_scalar(const T* __restrict__ stack, int Z,
                                                           int64_t n, T* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  T acc = stack[i];
  for (int z = 1; z < Z; ++z) {
    T v = stack[(int64_t)z * n + i];
    acc = IS_MAX ? (v > acc ? v : acc) : (v < acc ? v : acc);
  }
  out[i] = acc;
}

// ------------------------------------------------------------------------------------------
// avg: exact integer sum (Z*65535 < 2^32 for Z <= 65537), one float64 divide
// (np.mean of an integer array accumulates in float64, exact here, then true_divide by Z).
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) zproj_avg(const T* __restrict__ stack, int Z, int64_t n,
                                                 double* __restrict__ out) {
  // each thread: 4 consecutive pixels
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  uint32_t s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  bool full = (i + 4 <= n);
  bool vec_ok = full && ((n * sizeof(T)) % (4 * sizeof(T)) == 0) &&
                ((reinterpret_cast<uintptr_t>(stack) % (4 * sizeof(T))) == 0) && (i % 4 == 0);
  if (vec_ok) {
    for (int z = 0; z < Z; ++z) {
      const T* p = stack + (int64_t)z * n + i;
      if (sizeof(T) == 2) {
        uint2 v = *reinterpret_cast<const uint2*>(p);
        s0 += v.x & 0xffffu; s1 += v.x >> 16; s2 += v.y & 0xffffu; s3 += v.y >> 16;
      } else {
        uint32_t v = *reinterpret_cast<const uint32_t*>(p);
        s0 += v & 0xffu; s1 += (v >> 8) & 0xffu; s2 += (v >> 16) & 0xffu; s3 += v >> 24;
      }
    }
  } else {
    for (int z = 0; z < Z; ++z) {
      const T* p = stack + (int64_t)z * n + i;
      s0 += p[0];
      if (i + 1 < n) s1 += p[1];
      if (i + 2 < n) s2 += p[2];
      if (i + 3 < n) s3 += p[3];
    }
  }
  double dz = (double)Z;
  out[i] = (double)s0 / dz;
  if (i + 1 < n) out[i + 1] = (double)s1 / dz;
  if (i + 2 < n) out[i + 2] = (double)s2 / dz;
  if (i + 3 < n) out[i + 3] = (double)s3 / dz;
}

// ------------------------------------------------------------------------------------------
// median: in-register Batcher odd-even merge network over Z.  The stack is padded to P = next power of two with
// 0x0000 / 0xFFFF sentinels placed so that the median pair always lands on positions P/2-1 and P/2 (for odd Z the
// median is P/2-1 alone): only those two outputs are used, so the compiler removes every min/max that cannot
// reach them (284 of the 382 operations of the full 32-sorter survive; the bitonic network of round 1 kept all
// 480 and was bound by the ALU pipe at 0.44 of HBM).  Two pixels per 32-bit register (VIMNMX.U16x2), two
// independent networks per thread.  np.median: middle value (odd Z) or the float64 mean of the two middle
// values (even Z) -> k or k + 0.5 exactly.
// ------------------------------------------------------------------------------------------
// Batcher's network in its recursive form; every index is a template parameter, so the arrays stay in registers.
struct MinMaxU16x2 {
  __device__ __forceinline__ static void cs(uint32_t& x, uint32_t& y) {
    const uint32_t lo = __vminu2(x, y), hi = __vmaxu2(x, y);
    x = lo; y = hi;
  }
};
struct MinMaxF32 {
  __device__ __forceinline__ static void cs(float& x, float& y) {
    const float lo = fminf(x, y), hi = fmaxf(x, y);
    x = lo; y = hi;
  }
};
template <typename Op, int LO, int N, int R, typename V, int P>
__device__ __forceinline__ void oddeven_merge(V (&a)[P]) {
  constexpr int M = 2 * R;
  if constexpr (M < N) {
    oddeven_merge<Op, LO, N, M>(a);
    oddeven_merge<Op, LO + R, N, M>(a);
#pragma unroll
    for (int i = LO + R; i + R < LO + N; i += M) Op::cs(a[i], a[i + R]);
  } else {
    Op::cs(a[LO], a[LO + R]);
  }
}
template <typename Op, int LO, int N, typename V, int P>
__device__ __forceinline__ void oddeven_sort(V (&a)[P]) {
  if constexpr (N > 1) {
    oddeven_sort<Op, LO, N / 2>(a);
    oddeven_sort<Op, LO + N / 2, N / 2>(a);
    oddeven_merge<Op, LO, N, 1>(a);
  }
}
template <int P>
__device__ __forceinline__ void oddeven_sort_u16x2(uint32_t (&a)[P]) { oddeven_sort<MinMaxU16x2, 0, P>(a); }

template <typename T, int P>
__global__ void __launch_bounds__(128) zproj_med(const T* __restrict__ stack, int Z, int64_t n,
                                                 double* __restrict__ out) {
  // each thread: 4 consecutive pixels = two packed pairs
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  uint32_t a[P], b[P];
  bool vec_ok = (i + 4 <= n) && (n % 4 == 0) &&
                ((reinterpret_cast<uintptr_t>(stack) % (4 * sizeof(T))) == 0);
  // P - Z sentinels: floor((P - Z) / 2) zeros in front, the rest 0xFFFF behind -> after sorting the Z samples
  // occupy positions n_lo .. n_lo + Z - 1 and their middle sits at P/2 - 1 (and P/2 for even Z)
  const int n_lo = (P - Z) >> 1;
#pragma unroll
  for (int s = 0; s < P; ++s) {
    const int z = s - n_lo;
    if (z >= 0 && z < Z) {
      const T* p = stack + (int64_t)z * n + i;
      uint32_t v0, v1, v2, v3;
      if (vec_ok) {
        if (sizeof(T) == 2) {
          uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
          a[s] = v.x; b[s] = v.y;
          continue;
        } else {
          uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(p));
          v0 = v & 0xffu; v1 = (v >> 8) & 0xffu; v2 = (v >> 16) & 0xffu; v3 = v >> 24;
        }
      } else {
        v0 = p[0];
        v1 = (i + 1 < n) ? p[1] : 0;
        v2 = (i + 2 < n) ? p[2] : 0;
        v3 = (i + 3 < n) ? p[3] : 0;
      }
      a[s] = v0 | (v1 << 16);
      b[s] = v2 | (v3 << 16);
    } else {
      a[s] = z < 0 ? 0u : 0xffffffffu;
      b[s] = z < 0 ? 0u : 0xffffffffu;
    }
  }
  oddeven_sort_u16x2<P>(a);
  oddeven_sort_u16x2<P>(b);
  const uint32_t alo = a[P / 2 - 1], blo = b[P / 2 - 1];
  const uint32_t ahi = (Z & 1) ? alo : a[P / 2], bhi = (Z & 1) ? blo : b[P / 2];
  double r0 = ((double)(alo & 0xffffu) + (double)(ahi & 0xffffu)) * 0.5;
  double r1 = ((double)(alo >> 16) + (double)(ahi >> 16)) * 0.5;
  double r2 = ((double)(blo & 0xffffu) + (double)(bhi & 0xffffu)) * 0.5;
  double r3 = ((double)(blo >> 16) + (double)(bhi >> 16)) * 0.5;
  if (vec_ok) {
    reinterpret_cast<double2*>(out + i)[0] = make_double2(r0, r1);
    reinterpret_cast<double2*>(out + i)[1] = make_double2(r2, r3);
    return;
  }
  out[i] = r0;
  if (i + 1 < n) out[i + 1] = r1;
  if (i + 2 < n) out[i + 2] = r2;
  if (i + 3 < n) out[i + 3] = r3;
}

// float32 stacks (np.median of float32 stays float32; the mean of the two middle values is one float32 add and an
// exact halving).  Same network on fminf/fmaxf with -inf / +inf sentinels, one pixel per register.  NaNs are not
// handled (numpy propagates them).
template <int P>
__global__ void __launch_bounds__(128) zproj_med_f32(const float* __restrict__ stack, int Z, int64_t n,
                                                     float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float a[P];
  const int n_lo = (P - Z) >> 1;
#pragma unroll
  for (int s = 0; s < P; ++s) {
    const int z = s - n_lo;
    a[s] = (z >= 0 && z < Z) ? __ldg(stack + (int64_t)z * n + i) : (z < 0 ? -INFINITY : INFINITY);
  }
  oddeven_sort<MinMaxF32, 0, P>(a);
  const float lo = a[P / 2 - 1];
  out[i] = (Z & 1) ? lo : __fmul_rn(__fadd_rn(lo, a[P / 2]), 0.5f);
}

// Any Z: per-pixel rank selection in O(Z^2) compares on values re-read from L2.
template <typename T>
__global__ void __launch_bounds__(256) zproj_med_generic(const T* __restrict__ stack, int Z, int64_t n,
                                                         double* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int hi = Z >> 1, lo = (Z - 1) >> 1;
  uint32_t vlo = 0, vhi = 0;
  for (int z = 0; z < Z; ++z) {
    uint32_t v = stack[(int64_t)z * n + i];
    int less = 0, eq = 0;
    for (int y = 0; y < Z; ++y) {
      uint32_t u = stack[(int64_t)y * n + i];
      less += (u < v);
      eq += (u == v);
    }
    if (less <= lo && lo < less + eq) vlo = v;
    if (less <= hi && hi < less + eq) vhi = v;
  }
  out[i] = ((double)vlo + (double)vhi) * 0.5;
}

// ------------------------------------------------------------------------------------------
// focus stacking (zstacks.py:134-189).  Per slice, in exact integers:
//   B = (S (x) S * raw + 128) >> 8            cv2.GaussianBlur(5x5, sigma 0), S = [1 4 6 4 1]
//   L = |(S_y (x) D2_x + D2_y (x) S_x) * B|   cv2.Laplacian(CV_64F, ksize 5), D2 = [1 0 -2 0 1]
// both with BORDER_REFLECT_101.  Running arg-max over Z with strict '>' (first slice wins ties);
// the output is the raw pixel of the winning slice.  One CTA owns a TX x TY tile for all Z; the
// slice tile (+4 halo) is staged in shared memory, the running maximum lives in registers.
// ------------------------------------------------------------------------------------------
constexpr int FS_TX = 64, FS_TY = 32, FS_THREADS = 256;
constexpr int FS_RW = FS_TX + 8, FS_RH = FS_TY + 8;   // raw tile with halo 4
constexpr int FS_BW = FS_TX + 4, FS_BH = FS_TY + 4;   // blurred tile with halo 2

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  int period = 2 * (n - 1);
  i %= period;
  if (i < 0) i += period;
  return i < n ? i : period - i;
}

template <typename T>
__global__ void __launch_bounds__(FS_THREADS) zproj_focus(const T* __restrict__ stack, int Z, int H, int W,
                                                          T* __restrict__ out) {
  __shared__ uint16_t s_raw[FS_RH][FS_RW];
  __shared__ int s_hb[FS_RH][FS_BW];
  __shared__ uint16_t s_blur[FS_BH][FS_BW];
  __shared__ int s_hd2[FS_BH][FS_TX];
  __shared__ int s_hs[FS_BH][FS_TX];
  __shared__ int s_row[FS_RH];
  __shared__ int s_col[FS_RW];

  const int x0 = blockIdx.x * FS_TX, y0 = blockIdx.y * FS_TY;
  const int tid = threadIdx.x;
  for (int i = tid; i < FS_RH; i += FS_THREADS) s_row[i] = reflect101(y0 - 4 + i, H);
  for (int i = tid; i < FS_RW; i += FS_THREADS) s_col[i] = reflect101(x0 - 4 + i, W);
  __syncthreads();

  constexpr int PER_THREAD = FS_TX * FS_TY / FS_THREADS;   // 8 pixels per thread
  int best[PER_THREAD];
  uint32_t val[PER_THREAD];
#pragma unroll
  for (int p = 0; p < PER_THREAD; ++p) { best[p] = -1; val[p] = 0; }

  const int64_t plane = (int64_t)H * W;
  for (int z = 0; z < Z; ++z) {
    const T* sl = stack + (int64_t)z * plane;
    for (int i = tid; i < FS_RH * FS_RW; i += FS_THREADS) {
      int r = i / FS_RW, c = i - r * FS_RW;
      s_raw[r][c] = (uint16_t)sl[(int64_t)s_row[r] * W + s_col[c]];
    }
    __syncthreads();
    // horizontal blur taps on every raw row
    for (int i = tid; i < FS_RH * FS_BW; i += FS_THREADS) {
      int r = i / FS_BW, c = i - r * FS_BW;
      const uint16_t* p = &s_raw[r][c];
      s_hb[r][c] = (int)p[0] + 4 * (int)p[1] + 6 * (int)p[2] + 4 * (int)p[3] + (int)p[4];
    }
    __syncthreads();
    // vertical blur taps, single rounding
    for (int i = tid; i < FS_BH * FS_BW; i += FS_THREADS) {
      int r = i / FS_BW, c = i - r * FS_BW;
      int s = s_hb[r][c] + 4 * s_hb[r + 1][c] + 6 * s_hb[r + 2][c] + 4 * s_hb[r + 3][c] + s_hb[r + 4][c];
      s_blur[r][c] = (uint16_t)((s + 128) >> 8);
    }
    __syncthreads();
    // horizontal passes of the two Laplacian terms
    for (int i = tid; i < FS_BH * FS_TX; i += FS_THREADS) {
      int r = i / FS_TX, c = i - r * FS_TX;
      const uint16_t* p = &s_blur[r][c];
      int a0 = p[0], a1 = p[1], a2 = p[2], a3 = p[3], a4 = p[4];
      s_hd2[r][c] = a0 - 2 * a2 + a4;
      s_hs[r][c] = a0 + 4 * a1 + 6 * a2 + 4 * a3 + a4;
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < PER_THREAD; ++p) {
      int i = tid + p * FS_THREADS;
      int r = i / FS_TX, c = i - r * FS_TX;
      int lap = s_hd2[r][c] + 4 * s_hd2[r + 1][c] + 6 * s_hd2[r + 2][c] + 4 * s_hd2[r + 3][c] + s_hd2[r + 4][c]
              + s_hs[r][c] - 2 * s_hs[r + 2][c] + s_hs[r + 4][c];
      lap = lap < 0 ? -lap : lap;
      if (lap > best[p]) { best[p] = lap; val[p] = s_raw[r + 4][c + 4]; }
    }
    // the next slice overwrites s_raw / s_hd2 / s_hs only after everyone passed the barrier below
    __syncthreads();
  }
#pragma unroll
  for (int p = 0; p < PER_THREAD; ++p) {
    int i = tid + p * FS_THREADS;
    int r = i / FS_TX, c = i - r * FS_TX;
    int y = y0 + r, x = x0 + c;
    if (y < H && x < W) out[(int64_t)y * W + x] = (T)val[p];
  }
}

// ------------------------------------------------------------------------------------------
// focus stacking, fast path (uint16 stacks with W % 4 == 0, Z <= 256, H, W >= 8).
//
// One WARP owns a tile of 124 x TY output pixels for all Z; warps of a CTA are independent (no block barrier
// anywhere).  Per slice the warp's raw tile (132 x (TY + 8) pixels: halo 4) arrives in its own shared-memory
// ring by 8-byte cp.async while the previous slice is computed.  A lane owns 4 adjacent columns and slides down the
// rows with the vertical filter windows in registers:
//   * the horizontal taps work on PACKED pairs of uint16 pixels with dp2a (two 16-bit x 8-bit products + add in
//     one instruction): no unpacking of the raw tile, three dp2a per blurred pixel;
//   * blurred pixels are exchanged between neighbouring lanes as packed words (two shuffles per row);
//   * the ksize-5 Laplacian S_y (x) D2_x + D2_y (x) S_x factors, in exact integers, as
//       L = 2 * ( [1 2 2 2 1]_y [1 2 2 2 1]_x B  -  4 * [1 2 1]_y [1 2 1]_x B )
//     (S = [1 2 1]*[1 2 1], D2 = [1 2 1]*[1 -2 1]); the arg-max compares |L| / 2;
//   * the running arg-max is ONE 32-bit key per pixel in shared memory: (|L|/2) << 8 | (255 - z), so that max()
//     keeps the largest response and, among equals, the earliest slice (the reference's strict '>').
// After the last slice the output pixel is fetched from the winning slice.  BORDER_REFLECT_101: raw halo pixels are
// loaded through the reflected index; blurring the reflected raw image gives the reflected blurred image because
// both kernels are symmetric.
// ------------------------------------------------------------------------------------------
constexpr int FSW_OUT = 124;                 // valid output columns per warp tile (32 lanes x 4 - 2 x 2 halo)
constexpr int FSW_ROWB = 264;                // bytes per raw row in shared memory: 132 pixels = 33 chunks of 8 bytes

__device__ __forceinline__ uint32_t fsw_dp2a(uint32_t a, uint32_t coef, uint32_t c) { return __dp2a_lo(a, coef, c); }
__device__ __forceinline__ constexpr uint32_t fsw_c(uint32_t lo, uint32_t hi) { return lo | (hi << 8); }
__device__ __forceinline__ int fsw_reflect(int i, int n) {       // i in [-4, n + 3] exact; beyond: clamped (unused)
  i = i < 0 ? -i : i;
  i = i >= n ? 2 * (n - 1) - i : i;
  return i < 0 ? 0 : i;
}

template <int TY, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) zproj_focus_u16_fast(const uint16_t* __restrict__ stack, int Z, int H,
                                                                      int W, uint16_t* __restrict__ out, int ntx,
                                                                      int nty) {
  extern __shared__ __align__(16) unsigned char fsw_smem[];
  constexpr int RAWH = TY + 8;
  static_assert(RAWH % 5 == 0, "the row loop is unrolled by the period of the 5-row filter windows");
  constexpr int STAGE_B = RAWH * FSW_ROWB;
  constexpr int RAW_B = 2 * STAGE_B;
  constexpr int WARP_B = RAW_B + TY * 128 * 4;
  constexpr int NCHUNK = RAWH * 33;            // 8-byte chunks of one raw tile
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wbase = fsw_smem + (size_t)warp * WARP_B;
  const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(wbase);
  uint4* keys = reinterpret_cast<uint4*>(wbase + RAW_B);
  const int64_t plane = (int64_t)H * W;

  for (int t = blockIdx.x * WARPS + warp; t < ntx * nty; t += gridDim.x * WARPS) {
    const int ty = t / ntx, tx = t - ty * ntx;
    const int ox = tx * FSW_OUT, oy = ty * TY;
    const int rx0 = ox - 4, ry0 = oy - 4;
    // W % 4 == 0 and rx0 % 4 == 0: a chunk lies entirely inside the image or entirely outside.  Chunks inside are
    // copied by cp.async; the one chunk left of column 0 and the one right of column W - 1 (BORDER_REFLECT_101) are
    // mirrored from shared memory once the slice has landed; chunks further out feed no valid output.
    const bool left_edge = (rx0 < 0);
    const int right_chunk = (W - rx0) >> 2;                  // chunk that starts at column W
    const bool right_edge = right_chunk < 33;

    // the chunks a lane copies are the same for every slice of the tile: their source offset (in 8-byte units from the
    // slice's first pixel) and shared-memory offset are computed once per tile and packed into one register each
    constexpr int NPL = (NCHUNK + 31) / 32;
    uint32_t plan[NPL];
#pragma unroll
    for (int k = 0; k < NPL; ++k) {
      const int c = lane + 32 * k;
      const int rr = c / 33, i = c - rr * 33;
      const int x = rx0 + 4 * i;
      plan[k] = 0xffffffffu;
      if (c < NCHUNK && x >= 0 && x < W) {
        const uint32_t goff8 = (uint32_t)(((int64_t)fsw_reflect(ry0 + rr, H) * W + x) >> 2);    // 4 pixels = 8 bytes
        plan[k] = (goff8 << 11) | (uint32_t)((rr * FSW_ROWB + 8 * i) >> 3);
      }
    }
    auto load_slice = [&](int z, int stage) {
      const uint16_t* sl = stack + (int64_t)z * plane;
      const uint32_t sbase = raw_s + stage * STAGE_B;
#pragma unroll
      for (int k = 0; k < NPL; ++k) {
        if (plan[k] != 0xffffffffu) {
          const uint16_t* src = sl + ((size_t)(plan[k] >> 11) << 2);
          asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sbase + ((plan[k] & 2047u) << 3)), "l"(src) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto mirror_edges = [&](int stage) {
      if (!(left_edge || right_edge)) return;
      unsigned char* sb = wbase + stage * STAGE_B;
      for (int rr = lane; rr < RAWH; rr += 32) {
        uint16_t* row = reinterpret_cast<uint16_t*>(sb + rr * FSW_ROWB);
        if (left_edge) {                                     // columns -4..-1 <- columns 4..1 (raw index = column + 4)
          row[0] = row[8]; row[1] = row[7]; row[2] = row[6]; row[3] = row[5];
        }
        if (right_edge) {                                    // columns W..W+3 <- columns W-2..W-5
          uint16_t* q = row + 4 * right_chunk;
          q[0] = q[-2]; q[1] = q[-3]; q[2] = q[-4]; q[3] = q[-5];
        }
      }
    };

    for (int i = lane; i < TY * 32; i += 32) keys[i] = make_uint4(0u, 0u, 0u, 0u);
    load_slice(0, 0);
    for (int z = 0; z < Z; ++z) {
      if (z + 1 < Z) {
        load_slice(z + 1, (z + 1) & 1);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      } else {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      __syncwarp();
      mirror_edges(z & 1);
      __syncwarp();
      const uint32_t sbase = raw_s + (z & 1) * STAGE_B + 8u * lane;
      const uint32_t zkey = 255u - (uint32_t)z;
      uint32_t hw[5][4];     // horizontal blur taps of the last 5 raw rows
      uint32_t pw[5][4];     // [1 2 2 2 1]_x B of the last 5 blurred rows
      uint32_t qw[5][4];     // [1 2 1]_x B of the last 5 blurred rows (4 are live)
#pragma unroll 1
      for (int r0 = 0; r0 < RAWH; r0 += 5) {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const int rr = r0 + k;                         // raw row; rr % 5 == k
          uint32_t w0, w1, w2, w3;
          asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(sbase + rr * FSW_ROWB));
          asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w2), "=r"(w3) : "r"(sbase + rr * FSW_ROWB + 8));
          hw[k][0] = fsw_dp2a(w0, fsw_c(1, 4), fsw_dp2a(w1, fsw_c(6, 4), fsw_dp2a(w2, fsw_c(1, 0), 0)));
          hw[k][1] = fsw_dp2a(w0, fsw_c(0, 1), fsw_dp2a(w1, fsw_c(4, 6), fsw_dp2a(w2, fsw_c(4, 1), 0)));
          hw[k][2] = fsw_dp2a(w1, fsw_c(1, 4), fsw_dp2a(w2, fsw_c(6, 4), fsw_dp2a(w3, fsw_c(1, 0), 0)));
          hw[k][3] = fsw_dp2a(w1, fsw_c(0, 1), fsw_dp2a(w2, fsw_c(4, 6), fsw_dp2a(w3, fsw_c(4, 1), 0)));
          if (rr >= 4) {
            // blurred row br = rr - 4 <-> image row oy - 2 + br; br % 5 == (k + 1) % 5
            constexpr int KB = 0;  (void)KB;
            uint32_t bl[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t sum = hw[(k + 1) % 5][j] + hw[k][j] + 4u * (hw[(k + 2) % 5][j] + hw[(k + 4) % 5][j]) +
                                   6u * hw[(k + 3) % 5][j];
              bl[j] = (sum + 128u) >> 8;                 // cv2.GaussianBlur: one rounding, fits 16 bits
            }
            const uint32_t b01 = __byte_perm(bl[0], bl[1], 0x5410), b23 = __byte_perm(bl[2], bl[3], 0x5410);
            const uint32_t lft = __shfl_up_sync(0xffffffffu, b23, 1);      // (B[-2], B[-1]) from the lane to the left
            const uint32_t rgt = __shfl_down_sync(0xffffffffu, b01, 1);    // (B[4], B[5]) from the lane to the right
            const uint32_t l01 = fsw_dp2a(lft, fsw_c(0, 1), 0);            // B[-1]
            const uint32_t m22 = fsw_dp2a(b01, fsw_c(2, 2), 0), n22 = fsw_dp2a(b23, fsw_c(2, 2), 0);
            const int kb = (k + 1) % 5;                  // slot of blurred row br in the 5-row windows
            pw[kb][0] = fsw_dp2a(lft, fsw_c(1, 2), fsw_dp2a(b23, fsw_c(1, 0), m22));
            pw[kb][1] = fsw_dp2a(b23, fsw_c(2, 1), m22 + l01);
            pw[kb][2] = fsw_dp2a(b01, fsw_c(1, 2), fsw_dp2a(rgt, fsw_c(1, 0), n22));
            pw[kb][3] = fsw_dp2a(b01, fsw_c(0, 1), fsw_dp2a(rgt, fsw_c(2, 1), n22));
            qw[kb][0] = fsw_dp2a(b01, fsw_c(2, 1), l01);
            qw[kb][1] = fsw_dp2a(b01, fsw_c(1, 2), fsw_dp2a(b23, fsw_c(1, 0), 0));
            qw[kb][2] = fsw_dp2a(b01, fsw_c(0, 1), fsw_dp2a(b23, fsw_c(2, 1), 0));
            qw[kb][3] = fsw_dp2a(b23, fsw_c(1, 2), fsw_dp2a(rgt, fsw_c(1, 0), 0));
            if (rr >= 8) {
              const int orow = rr - 8;                   // output row <-> image row oy + orow
              uint4* kp = keys + orow * 32 + lane;
              const uint4 old = *kp;
              uint32_t key[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                // window slots: blurred rows br-4 .. br sit at (kb+1)%5 .. kb
                const int v1 = (int)(pw[(kb + 1) % 5][j] + pw[kb][j] +
                                     2u * (pw[(kb + 2) % 5][j] + pw[(kb + 3) % 5][j] + pw[(kb + 4) % 5][j]));
                const int v2 = (int)(qw[(kb + 2) % 5][j] + 2u * qw[(kb + 3) % 5][j] + qw[(kb + 4) % 5][j]);
                const int half = v1 - 4 * v2;            // cv2.Laplacian(ksize=5) = 2 * half, |half| < 2^23
                key[j] = ((uint32_t)(half < 0 ? -half : half) << 8) | zkey;
              }
              *kp = make_uint4(max(key[0], old.x), max(key[1], old.y), max(key[2], old.z), max(key[3], old.w));
            }
          }
        }
      }
      __syncwarp();
    }
    // the output takes the ORIGINAL pixel of the winning slice (zstacks.py:181-187)
    for (int orow = 0; orow < TY; ++orow) {
      const int y = oy + orow;
      if (y >= H) break;
      const uint4 k = keys[orow * 32 + lane];
      const uint32_t kk[4] = {k.x, k.y, k.z, k.w};
      uint32_t v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int x = ox - 2 + 4 * lane + j;
        const bool valid = (lane > 0 || j >= 2) && (lane < 31 || j < 2) && x < W;
        v[j] = valid ? (uint32_t)__ldg(stack + (int64_t)(255 - (int)(kk[j] & 255u)) * plane + (int64_t)y * W + x) : 0u;
      }
      // x of j = 0 is even and W is even: (j0, j1) and (j2, j3) are aligned pairs
      const int x0 = ox - 2 + 4 * lane;
      uint32_t* o32 = reinterpret_cast<uint32_t*>(out + (int64_t)y * W);
      if (lane > 0 && x0 + 1 < W) o32[x0 >> 1] = v[0] | (v[1] << 16);
      if (lane < 31 && x0 + 3 < W) o32[(x0 + 2) >> 1] = v[2] | (v[3] << 16);
    }
    __syncwarp();
  }
}

template <int TY, int WARPS>
int launch_focus_fast(tmat_ctx* ctx, const uint16_t* stack, int Z, int H, int W, uint16_t* out, cudaStream_t st) {
  constexpr int SMEM = WARPS * (2 * (TY + 8) * FSW_ROWB + TY * 128 * 4);
  static bool attr_set = false;
  if (!attr_set) {
    CUDA_TRY(cudaFuncSetAttribute(zproj_focus_u16_fast<TY, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_set = true;
  }
  const int ntx = cdiv(W, FSW_OUT), nty = cdiv(H, TY);
  const int grid = std::min(cdiv(ntx * nty, WARPS), ctx->sm_count > 0 ? ctx->sm_count : 148);
  zproj_focus_u16_fast<TY, WARPS><<<grid, WARPS * 32, SMEM, st>>>(stack, Z, H, W, out, ntx, nty);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

// float32 stacks: np.mean keeps float32 -- np.add.reduce over the leading axis accumulates slice by slice in float32,
// then one float32 divide by Z.
__global__ void __launch_bounds__(256) zproj_avg_f32(const float* __restrict__ stack, int Z, int64_t n,
                                                     float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = stack[i];
  for (int z = 1; z < Z; ++z) s = __fadd_rn(s, stack[(int64_t)z * n + i]);
  out[i] = __fdiv_rn(s, (float)Z);
}
__global__ void __launch_bounds__(256) zproj_med_generic_f32(const float* __restrict__ stack, int Z, int64_t n,
                                                             float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int hi = Z >> 1, lo = (Z - 1) >> 1;
  float vlo = 0.f, vhi = 0.f;
  for (int z = 0; z < Z; ++z) {
    float v = stack[(int64_t)z * n + i];
    int less = 0, eq = 0;
    for (int y = 0; y < Z; ++y) {
      float u = stack[(int64_t)y * n + i];
      less += (u < v);
      eq += (u == v);
    }
    if (less <= lo && lo < less + eq) vlo = v;
    if (less <= hi && hi < less + eq) vhi = v;
  }
  out[i] = (Z & 1) ? vlo : __fmul_rn(__fadd_rn(vlo, vhi), 0.5f);
}

template <typename T>
int launch_minmax(tmat_ctx* ctx, const T* stack, int Z, int64_t n, bool is_max, T* out, cudaStream_t st) {
  const int vec = 16 / sizeof(T);
  bool vec_ok = (n % vec == 0) && (reinterpret_cast<uintptr_t>(stack) % 16 == 0) &&
                (reinterpret_cast<uintptr_t>(out) % 16 == 0);
  if (vec_ok) {
    int64_t nv = n / vec;
    dim3 grid((unsigned)cdiv64(nv, 256));
    if (is_max) zproj_minmax_vec<T, true><<<grid, 256, 0, st>>>((const uint4*)stack, Z, nv, (uint4*)out);
    else        zproj_minmax_vec<T, false><<<grid, 256, 0, st>>>((const uint4*)stack, Z, nv, (uint4*)out);
  } else {
    dim3 grid((unsigned)cdiv64(n, 256));
    if (is_max) zproj_minmax_scalar<T, true><<<grid, 256, 0, st>>>(stack, Z, n, out);
    else        zproj_minmax_scalar<T, false><<<grid, 256, 0, st>>>(stack, Z, n, out);
  }
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

template <typename T>
int launch_med(tmat_ctx* ctx, const T* stack, int Z, int64_t n, double* out, cudaStream_t st) {
  dim3 grid((unsigned)cdiv64(cdiv64(n, 4), 128));
  if (Z <= 2)       zproj_med<T, 2><<<grid, 128, 0, st>>>(stack, Z, n, out);
  else if (Z <= 4)  zproj_med<T, 4><<<grid, 128, 0, st>>>(stack, Z, n, out);
  else if (Z <= 8)  zproj_med<T, 8><<<grid, 128, 0, st>>>(stack, Z, n, out);
  else if (Z <= 16) zproj_med<T, 16><<<grid, 128, 0, st>>>(stack, Z, n, out);
  else if (Z <= 32) zproj_med<T, 32><<<grid, 128, 0, st>>>(stack, Z, n, out);
  else if (Z <= 64) zproj_med<T, 64><<<grid, 128, 0, st>>>(stack, Z, n, out);
  else zproj_med_generic<T><<<(unsigned)cdiv64(n, 256), 256, 0, st>>>(stack, Z, n, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

// fast focus-stacking path: uint16, 8-byte aligned rows (cp.async), keys hold the slice in 8 bits
template <typename T>
bool focus_fast_ok(const T*, int, int, int, const void*) { return false; }
template <>
bool focus_fast_ok<uint16_t>(const uint16_t* stack, int Z, int H, int W, const void* out) {
  static const bool off = [] { const char* e = getenv("TMAT_FS_GENERIC"); return e && atoi(e) != 0; }();
  return !off && Z <= 256 && H >= 8 && W >= 8 && (W % 4) == 0 && (int64_t)H * W <= (1ll << 23) &&       // 21-bit source offsets in the per-tile copy plan
         (reinterpret_cast<uintptr_t>(stack) % 8) == 0 &&
         (reinterpret_cast<uintptr_t>(out) % 4) == 0;
}
template <typename T>
int launch_focus_fast_any(tmat_ctx*, const T*, int, int, int, void*, cudaStream_t) { return TMAT_ERR_UNSUPPORTED; }
template <>
int launch_focus_fast_any<uint16_t>(tmat_ctx* ctx, const uint16_t* stack, int Z, int H, int W, void* out, cudaStream_t st) {
  static const int ty = [] { const char* e = getenv("TMAT_FS_TY"); return e ? atoi(e) : 12; }();
  if (ty == 12) return launch_focus_fast<12, 13>(ctx, stack, Z, H, W, (uint16_t*)out, st);
  if (ty == 22) return launch_focus_fast<22, 8>(ctx, stack, Z, H, W, (uint16_t*)out, st);
  return launch_focus_fast<32, 6>(ctx, stack, Z, H, W, (uint16_t*)out, st);
}

template <typename T>
int zproj_typed(tmat_ctx* ctx, const T* stack, int Z, int H, int W, int method, void* out, cudaStream_t st) {
  int64_t n = (int64_t)H * W;
  switch (method) {
    case TMAT_ZPROJ_MIN: return launch_minmax<T>(ctx, stack, Z, n, false, (T*)out, st);
    case TMAT_ZPROJ_MAX: return launch_minmax<T>(ctx, stack, Z, n, true, (T*)out, st);
    case TMAT_ZPROJ_AVG:
      zproj_avg<T><<<(unsigned)cdiv64(cdiv64(n, 4), 256), 256, 0, st>>>(stack, Z, n, (double*)out);
      KERNEL_CHECK(ctx);
      return TMAT_OK;
    case TMAT_ZPROJ_MED: return launch_med<T>(ctx, stack, Z, n, (double*)out, st);
    case TMAT_ZPROJ_FS: {
      if (focus_fast_ok<T>(stack, Z, H, W, out)) return launch_focus_fast_any(ctx, stack, Z, H, W, out, st);
      dim3 grid(cdiv(W, FS_TX), cdiv(H, FS_TY));
      zproj_focus<T><<<grid, FS_THREADS, 0, st>>>(stack, Z, H, W, (T*)out);
      KERNEL_CHECK(ctx);
      return TMAT_OK;
    }
  }
  tmat_set_error("tmat_zproj: unknown method %d", method);
  return TMAT_ERR_INVALID;
}

}  // namespace

// max projection of a uint16 stack, used by the whole-path driver too (compute_branches.py:228)
int zproj_max_u16(tmat_ctx* ctx, const uint16_t* stack, int Z, int H, int W, uint16_t* out, cudaStream_t st) {
  return launch_minmax<uint16_t>(ctx, stack, Z, (int64_t)H * W, true, out, st);
}

extern "C" int tmat_zproj_dev(tmat_ctx* ctx, const void* stack_dev, int dtype, int Z, int H, int W,
                              int method, int ksize, void* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && stack_dev && out_dev, "tmat_zproj: NULL argument");
  TMAT_REQUIRE(Z >= 1 && H >= 1 && W >= 1, "tmat_zproj: bad shape (%d,%d,%d)", Z, H, W);
  TMAT_REQUIRE(method >= TMAT_ZPROJ_MIN && method <= TMAT_ZPROJ_FS, "tmat_zproj: unknown method %d", method);
  if (method == TMAT_ZPROJ_FS && ksize != 5) {
    tmat_set_error("tmat_zproj: focus stacking implements kernel_size=5 (the reference default) only, got %d", ksize);
    return TMAT_ERR_UNSUPPORTED;
  }
  if (method == TMAT_ZPROJ_AVG && Z > 65536) {
    tmat_set_error("tmat_zproj: avg supports Z <= 65536");
    return TMAT_ERR_UNSUPPORTED;
  }
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  if (dtype == TMAT_U16) return zproj_typed<uint16_t>(ctx, (const uint16_t*)stack_dev, Z, H, W, method, out_dev, st);
  if (dtype == TMAT_U8) return zproj_typed<uint8_t>(ctx, (const uint8_t*)stack_dev, Z, H, W, method, out_dev, st);
  if (dtype == TMAT_F32 && method != TMAT_ZPROJ_FS) {
    // numpy semantics for a float32 stack: every projection stays float32 (zstacks.py:192-249 call np.mean/median/max/min)
    const float* stack = (const float*)stack_dev;
    const int64_t n = (int64_t)H * W;
    const unsigned g256 = (unsigned)cdiv64(n, 256);
    switch (method) {
      case TMAT_ZPROJ_MIN: zproj_minmax_scalar<float, false><<<g256, 256, 0, st>>>(stack, Z, n, (float*)out_dev); break;
      case TMAT_ZPROJ_MAX: zproj_minmax_scalar<float, true><<<g256, 256, 0, st>>>(stack, Z, n, (float*)out_dev); break;
      case TMAT_ZPROJ_AVG: zproj_avg_f32<<<g256, 256, 0, st>>>(stack, Z, n, (float*)out_dev); break;
      default: {
        const unsigned g128 = (unsigned)cdiv64(n, 128);
        if (Z <= 2)       zproj_med_f32<2><<<g128, 128, 0, st>>>(stack, Z, n, (float*)out_dev);
        else if (Z <= 4)  zproj_med_f32<4><<<g128, 128, 0, st>>>(stack, Z, n, (float*)out_dev);
        else if (Z <= 8)  zproj_med_f32<8><<<g128, 128, 0, st>>>(stack, Z, n, (float*)out_dev);
        else if (Z <= 16) zproj_med_f32<16><<<g128, 128, 0, st>>>(stack, Z, n, (float*)out_dev);
        else if (Z <= 32) zproj_med_f32<32><<<g128, 128, 0, st>>>(stack, Z, n, (float*)out_dev);
        else if (Z <= 64) zproj_med_f32<64><<<g128, 128, 0, st>>>(stack, Z, n, (float*)out_dev);
        else zproj_med_generic_f32<<<g256, 256, 0, st>>>(stack, Z, n, (float*)out_dev);
      }
    }
    KERNEL_CHECK(ctx);
    return TMAT_OK;
  }
  tmat_set_error("tmat_zproj: dtype %d not supported for method %d (uint8 / uint16 stacks; float32 for min/max/avg/med -- "
                 "cv2.GaussianBlur/Laplacian of a float stack is not restated)", dtype, method);
  return TMAT_ERR_UNSUPPORTED;
}

extern "C" int tmat_zproj_host(tmat_ctx* ctx, const void* stack_host, int dtype, int Z, int H, int W,
                               int method, int ksize, void* out_host) {
  TMAT_REQUIRE(ctx && stack_host && out_host, "tmat_zproj_host: NULL argument");
  TMAT_REQUIRE(dtype == TMAT_U8 || dtype == TMAT_U16 || dtype == TMAT_F32, "tmat_zproj_host: dtype %d not supported", dtype);
  TMAT_REQUIRE(Z >= 1 && H >= 1 && W >= 1, "tmat_zproj_host: bad shape (%d,%d,%d)", Z, H, W);
  TMAT_TRY(arena_reset(ctx));
  size_t es = dtype_size(dtype);
  size_t in_bytes = (size_t)Z * H * W * es;
  size_t out_es = (dtype != TMAT_F32 && (method == TMAT_ZPROJ_AVG || method == TMAT_ZPROJ_MED)) ? 8 : es;
  size_t out_bytes = (size_t)H * W * out_es;
  void *d_in, *d_out;
  TMAT_TRY(arena_alloc(ctx, in_bytes, &d_in));
  TMAT_TRY(arena_alloc(ctx, out_bytes, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_in, stack_host, in_bytes, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(tmat_zproj_dev(ctx, d_in, dtype, Z, H, W, method, ksize, d_out, nullptr));
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, out_bytes, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}
