// Stage 1: Z-stack projections (reference: fl_tissue_model_tools/zstacks.py:134-249).
//   proj_max:221  proj_min:237  proj_avg:192  proj_med:206  proj_focus_stacking:153 (+ _blur_and_lap:134)
// All kernels are HBM-bandwidth bound streaming reductions over Z: every voxel is read once
// (focus stacking re-reads a 4-pixel halo per tile), one output pixel is written.
#include "common.cuh"

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
// median: in-register bitonic sorting network over Z (padded to P = next power of two with
// 0xFFFF sentinels), two pixels packed per 32-bit register (u16x2 min/max = one instruction
// each), two independent networks per thread for ILP.  np.median: middle value (odd Z) or the
// float64 mean of the two middle values (even Z) -> k or k + 0.5 exactly.
// ------------------------------------------------------------------------------------------
template <int P>
__device__ __forceinline__ void bitonic_sort_u16x2(uint32_t (&a)[P]) {
#pragma unroll
  for (int k = 2; k <= P; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
      for (int i = 0; i < P; ++i) {
        int l = i ^ j;
        if (l > i) {
          uint32_t lo = __vminu2(a[i], a[l]);
          uint32_t hi = __vmaxu2(a[i], a[l]);
          bool asc = ((i & k) == 0);
          a[i] = asc ? lo : hi;
          a[l] = asc ? hi : lo;
        }
      }
    }
  }
}

template <typename T, int P>
__global__ void __launch_bounds__(128) zproj_med(const T* __restrict__ stack, int Z, int64_t n,
                                                 double* __restrict__ out) {
  // each thread: 4 consecutive pixels = two packed pairs
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  uint32_t a[P], b[P];
  bool vec_ok = (i + 4 <= n) && (n % 4 == 0) &&
                ((reinterpret_cast<uintptr_t>(stack) % (4 * sizeof(T))) == 0);
#pragma unroll
  for (int z = 0; z < P; ++z) {
    if (z < Z) {
      const T* p = stack + (int64_t)z * n + i;
      uint32_t v0, v1, v2, v3;
      if (vec_ok) {
        if (sizeof(T) == 2) {
          uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
          a[z] = v.x; b[z] = v.y;
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
      a[z] = v0 | (v1 << 16);
      b[z] = v2 | (v3 << 16);
    } else {
      a[z] = 0xffffffffu;
      b[z] = 0xffffffffu;
    }
  }
  bitonic_sort_u16x2<P>(a);
  bitonic_sort_u16x2<P>(b);
  // ascending; the Z real samples occupy positions 0..Z-1 (sentinels are >= every sample)
  int hi = Z >> 1, lo = (Z - 1) >> 1;
  uint32_t alo = 0, ahi = 0, blo = 0, bhi = 0;
#pragma unroll
  for (int z = 0; z < P; ++z) {
    if (z == lo) { alo = a[z]; blo = b[z]; }
    if (z == hi) { ahi = a[z]; bhi = b[z]; }
  }
  double r0 = ((double)(alo & 0xffffu) + (double)(ahi & 0xffffu)) * 0.5;
  double r1 = ((double)(alo >> 16) + (double)(ahi >> 16)) * 0.5;
  double r2 = ((double)(blo & 0xffffu) + (double)(bhi & 0xffffu)) * 0.5;
  double r3 = ((double)(blo >> 16) + (double)(bhi >> 16)) * 0.5;
  out[i] = r0;
  if (i + 1 < n) out[i + 1] = r1;
  if (i + 2 < n) out[i + 2] = r2;
  if (i + 3 < n) out[i + 3] = r3;
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
  tmat_set_error("tmat_zproj: dtype %d not supported (uint8 / uint16 stacks only)", dtype);
  return TMAT_ERR_UNSUPPORTED;
}

extern "C" int tmat_zproj_host(tmat_ctx* ctx, const void* stack_host, int dtype, int Z, int H, int W,
                               int method, int ksize, void* out_host) {
  TMAT_REQUIRE(ctx && stack_host && out_host, "tmat_zproj_host: NULL argument");
  TMAT_REQUIRE(dtype == TMAT_U8 || dtype == TMAT_U16, "tmat_zproj_host: dtype %d not supported", dtype);
  TMAT_REQUIRE(Z >= 1 && H >= 1 && W >= 1, "tmat_zproj_host: bad shape (%d,%d,%d)", Z, H, W);
  TMAT_TRY(arena_reset(ctx));
  size_t es = dtype_size(dtype);
  size_t in_bytes = (size_t)Z * H * W * es;
  size_t out_es = (method == TMAT_ZPROJ_AVG || method == TMAT_ZPROJ_MED) ? 8 : es;
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
