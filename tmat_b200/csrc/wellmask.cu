// The numpy-only pieces of generate_well_mask (well_mask_generation.py:16-140): the random search for
// the smallest enclosing superellipse, the superellipse mask and the convex-hull mask.
//   * get_superellipse_hull (:16-91): num_iters (25000) random parameter sets x P hull vertices -- a candidate is kept
//     when max_j f(vertex_j) < 1, the winner is the FIRST candidate of smallest area.  The parameter sets themselves
//     (np.random.rand, the affine map into the bounds, np.cos / np.sin of theta) are drawn by the Python mirror with
//     numpy, exactly as the reference draws them, so the same global seed gives the same winner; the device does the
//     O(num_iters x P) evaluation with numpy's operation order and separately rounded IEEE operations, and the arg-min.
//   * gen_superellipse_mask (:94-118): the same expression over the linspace(-1, 1) grid.
//   * create_convex_hull_mask (:121-140): scipy's Delaunay(hull_vertices).find_simplex(pixel) >= 0, i.e. inside or on
//     the hull.  Hull vertices are integer pixel coordinates, so the test is done in exact integer arithmetic (cross
//     products of int64): a lattice point is either exactly on an edge or at least 1 / |edge| away from it, far outside
//     qhull's 100 * DBL_EPSILON tolerance -- the two agree on every pixel.
// x ** n: numpy squares for n == 2 (exact); for any other exponent it calls libm's pow.  The device uses repeated
// squaring for n == 8 and pow() otherwise: the value only matters through '< 1', so a last-bit difference could flip a
// candidate / pixel only if the sum lay within an ulp of 1.0.
#include "common.cuh"
#include <math.h>

namespace {

__device__ __forceinline__ double se_pow(double v, int n) {
  const double a = fabs(v);
  if (n == 2) return __dmul_rn(a, a);
  if (n == 8) { double s = __dmul_rn(a, a); s = __dmul_rn(s, s); return __dmul_rn(s, s); }
  return pow(a, (double)n);
}
// ((x - c_x) * cos - (y - c_y) * sin) / (d * s_a)) ** n + (((x - c_x) * sin + (y - c_y) * cos) / (d * s_b)) ** n
__device__ __forceinline__ double se_value(double x, double y, double cx, double cy, double ct, double st, double da, double db,
                                           int n, bool rotate) {
  const double dx = __dsub_rn(x, cx), dy = __dsub_rn(y, cy);
  double u, v;
  if (rotate) {
    u = __dsub_rn(__dmul_rn(dx, ct), __dmul_rn(dy, st));
    v = __dadd_rn(__dmul_rn(dx, st), __dmul_rn(dy, ct));
  } else {
    u = dx; v = dy;
  }
  return __dadd_rn(se_pow(__ddiv_rn(u, da), n), se_pow(__ddiv_rn(v, db), n));
}

// one thread per parameter set: key[i] = area if every vertex is strictly inside, else +inf
__global__ void __launch_bounds__(128) superellipse_candidates(const double* __restrict__ params, const double* __restrict__ cos_t,
                                                               const double* __restrict__ sin_t, int iters,
                                                               const double* __restrict__ px, const double* __restrict__ py, int P,
                                                               int n, double g1sq, double g2, double* __restrict__ key) {
  extern __shared__ double pts[];              // x[P] | y[P]
  for (int j = threadIdx.x; j < P; j += blockDim.x) { pts[j] = px[j]; pts[P + j] = py[j]; }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= iters) return;
  const double* q = params + (int64_t)i * 6;
  const double d = q[1], sa = q[2], sb = q[3], cx = q[4], cy = q[5];
  const bool rotate = n != 2;                  // :53-57: for n == 2 theta is ignored
  const double ct = rotate ? cos_t[i] : 1.0, st = rotate ? sin_t[i] : 0.0;
  const double da = __dmul_rn(d, sa), db = __dmul_rn(d, sb);
  double m = -1.0;                             // the values are >= 0
  bool nan = false;
  for (int j = 0; j < P; ++j) {
    const double v = se_value(pts[j], pts[P + j], cx, cy, ct, st, da, db, n, rotate);
    nan |= (v != v);
    m = v > m ? v : m;
  }
  // :81-84: 4 * d**2 * s_a * s_b * gamma(1 + 1/n)**2 / gamma(1 + 2/n), left to right
  double area = __dmul_rn(4.0, __dmul_rn(d, d));
  area = __dmul_rn(area, sa);
  area = __dmul_rn(area, sb);
  area = __dmul_rn(area, g1sq);
  area = __ddiv_rn(area, g2);
  key[i] = (!nan && m < 1.0) ? area : __longlong_as_double(0x7ff0000000000000ll);
}
// np.argmin over the candidates in index order: smallest area, first index on ties; -1 when there is no candidate
__global__ void __launch_bounds__(1024) superellipse_argmin(const double* __restrict__ key, int iters, int* __restrict__ best) {
  __shared__ double s_val[1024];
  __shared__ int s_idx[1024];
  double v = __longlong_as_double(0x7ff0000000000000ll);
  int ix = -1;
  for (int i = threadIdx.x; i < iters; i += 1024) {
    const double k = key[i];
    if (k < v) { v = k; ix = i; }             // ascending i per thread: the first minimum stays
  }
  s_val[threadIdx.x] = v; s_idx[threadIdx.x] = ix;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) {
      const double ov = s_val[threadIdx.x + s];
      const int oi = s_idx[threadIdx.x + s];
      const double mv = s_val[threadIdx.x];
      const int mi = s_idx[threadIdx.x];
      if (oi >= 0 && (mi < 0 || ov < mv || (ov == mv && oi < mi))) { s_val[threadIdx.x] = ov; s_idx[threadIdx.x] = oi; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *best = s_idx[0];
}

// out[a][b] = f(xs[a], ys[b]) < 1 (the reference's swapaxes of the meshgrid, :108-116)
__global__ void __launch_bounds__(256) superellipse_mask(const double* __restrict__ xs, int n0, const double* __restrict__ ys, int n1,
                                                         double cx, double cy, double ct, double st, double da, double db, int n,
                                                         uint8_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y;
  if (b >= n1) return;
  const double v = se_value(xs[a], ys[b], cx, cy, ct, st, da, db, n, true);
  out[(int64_t)a * n1 + b] = v < 1.0;
}

// inside or on the convex polygon verts[0..nv) (row, col), either orientation
__global__ void __launch_bounds__(256) hull_mask(const int* __restrict__ verts, int nv, int H, int W, uint8_t* __restrict__ out) {
  extern __shared__ int sv[];
  for (int j = threadIdx.x; j < 2 * nv; j += blockDim.x) sv[j] = verts[j];
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (c >= W) return;
  bool pos = false, neg = false;
  for (int j = 0; j < nv; ++j) {
    const int k = j + 1 == nv ? 0 : j + 1;
    const long long er = sv[2 * k] - sv[2 * j], ec = sv[2 * k + 1] - sv[2 * j + 1];
    const long long cr = (long long)r - sv[2 * j], cc = (long long)c - sv[2 * j + 1];
    const long long cross = er * cc - ec * cr;
    pos |= cross > 0;
    neg |= cross < 0;
  }
  out[(int64_t)r * W + c] = !(pos && neg);
}

}  // namespace

extern "C" int tmat_superellipse_hull_host(tmat_ctx* ctx, const double* x_host, const double* y_host, int n_points, int n,
                                           const double* params_host, const double* cos_host, const double* sin_host,
                                           int num_iters, double gamma1_sq, double gamma2, int* best_index) {
  TMAT_REQUIRE(ctx && x_host && y_host && params_host && best_index, "tmat_superellipse_hull: NULL argument");
  TMAT_REQUIRE(n_points >= 1 && n_points <= 2048, "tmat_superellipse_hull: %d points outside [1, 2048]", n_points);
  TMAT_REQUIRE(num_iters >= 1 && n >= 1, "tmat_superellipse_hull: num_iters=%d, n=%d", num_iters, n);
  TMAT_REQUIRE(n == 2 || (cos_host && sin_host), "tmat_superellipse_hull: cos/sin of theta are needed for n != 2");
  CUDA_TRY(cudaSetDevice(ctx->device));
  TMAT_TRY(arena_reset(ctx));
  double *d_par, *d_cos = nullptr, *d_sin = nullptr, *d_x, *d_y, *d_key;
  int* d_best;
  TMAT_TRY(arena_get(ctx, (size_t)num_iters * 6, &d_par));
  TMAT_TRY(arena_get(ctx, (size_t)num_iters, &d_key));
  TMAT_TRY(arena_get(ctx, (size_t)n_points, &d_x));
  TMAT_TRY(arena_get(ctx, (size_t)n_points, &d_y));
  TMAT_TRY(arena_get(ctx, 1, &d_best));
  CUDA_TRY(cudaMemcpyAsync(d_par, params_host, (size_t)num_iters * 48, cudaMemcpyHostToDevice, 0));
  CUDA_TRY(cudaMemcpyAsync(d_x, x_host, (size_t)n_points * 8, cudaMemcpyHostToDevice, 0));
  CUDA_TRY(cudaMemcpyAsync(d_y, y_host, (size_t)n_points * 8, cudaMemcpyHostToDevice, 0));
  if (n != 2) {
    TMAT_TRY(arena_get(ctx, (size_t)num_iters, &d_cos));
    TMAT_TRY(arena_get(ctx, (size_t)num_iters, &d_sin));
    CUDA_TRY(cudaMemcpyAsync(d_cos, cos_host, (size_t)num_iters * 8, cudaMemcpyHostToDevice, 0));
    CUDA_TRY(cudaMemcpyAsync(d_sin, sin_host, (size_t)num_iters * 8, cudaMemcpyHostToDevice, 0));
  }
  superellipse_candidates<<<cdiv(num_iters, 128), 128, (size_t)n_points * 16>>>(d_par, d_cos, d_sin, num_iters, d_x, d_y, n_points, n,
                                                                               gamma1_sq, gamma2, d_key);
  KERNEL_CHECK(ctx);
  superellipse_argmin<<<1, 1024>>>(d_key, num_iters, d_best);
  KERNEL_CHECK(ctx);
  CUDA_TRY(cudaMemcpyAsync(best_index, d_best, sizeof(int), cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

extern "C" int tmat_superellipse_mask_host(tmat_ctx* ctx, const double* xs_host, int n0, const double* ys_host, int n1,
                                           double c_x, double c_y, double cos_t, double sin_t, double d_sa, double d_sb, int n,
                                           uint8_t* out_host) {
  TMAT_REQUIRE(ctx && xs_host && ys_host && out_host && n0 >= 1 && n1 >= 1 && n >= 1, "tmat_superellipse_mask: bad argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  TMAT_TRY(arena_reset(ctx));
  double *d_xs, *d_ys;
  uint8_t* d_out;
  TMAT_TRY(arena_get(ctx, (size_t)n0, &d_xs));
  TMAT_TRY(arena_get(ctx, (size_t)n1, &d_ys));
  TMAT_TRY(arena_get(ctx, (size_t)n0 * n1, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_xs, xs_host, (size_t)n0 * 8, cudaMemcpyHostToDevice, 0));
  CUDA_TRY(cudaMemcpyAsync(d_ys, ys_host, (size_t)n1 * 8, cudaMemcpyHostToDevice, 0));
  superellipse_mask<<<dim3(cdiv(n1, 256), n0), 256>>>(d_xs, n0, d_ys, n1, c_x, c_y, cos_t, sin_t, d_sa, d_sb, n, d_out);
  KERNEL_CHECK(ctx);
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, (size_t)n0 * n1, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

extern "C" int tmat_convex_hull_mask_host(tmat_ctx* ctx, int H, int W, const int32_t* verts_host, int n_verts, uint8_t* out_host) {
  TMAT_REQUIRE(ctx && verts_host && out_host && H >= 1 && W >= 1, "tmat_convex_hull_mask: bad argument");
  TMAT_REQUIRE(n_verts >= 3 && n_verts <= 4096, "tmat_convex_hull_mask: %d hull vertices outside [3, 4096]", n_verts);
  CUDA_TRY(cudaSetDevice(ctx->device));
  TMAT_TRY(arena_reset(ctx));
  int* d_v;
  uint8_t* d_out;
  TMAT_TRY(arena_get(ctx, (size_t)n_verts * 2, &d_v));
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_v, verts_host, (size_t)n_verts * 8, cudaMemcpyHostToDevice, 0));
  hull_mask<<<dim3(cdiv(W, 256), H), 256, (size_t)n_verts * 8>>>(d_v, n_verts, H, W, d_out);
  KERNEL_CHECK(ctx);
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, (size_t)H * W, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}
