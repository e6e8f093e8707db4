// exec_threshold's statistics on the GPU (reference: fl_tissue_model_tools/preprocessing.py:44-93):
//   GaussianMixture(n_components=2, random_state=rs).fit(pixels[:, None]) -> threshold = min(255, mu_fg + sd_coef * sd_fg)
// restated from scikit-learn's published algorithm (sklearn/mixture/_base.py, _gaussian_mixture.py,
// sklearn/cluster/_kmeans.py; defaults: init_params='kmeans' with one k-means++-seeded Lloyd run, tol 1e-3,
// max_iter 100, reg_covar 1e-6, full covariance), for 1-D data and two components:
//   1. k-means++ seeding from three uniform draws of the caller's RandomState (made on the host in the same
//      order scikit-learn makes them, so the caller's stream advances identically);
//   2. Lloyd iterations (tol = 1e-4 * var(x), max 300, stop on unchanged assignment or small centre shift);
//   3. one-hot responsibilities -> weights, means, variances (+ reg_covar);
//   4. EM until |delta mean log-likelihood| < 1e-3.
// Every pass over the samples is one deterministic two-stage float64 reduction (fixed chunk per block, fixed
// summation order), so results do not depend on scheduling.  scikit-learn itself computes in the input's
// float32; this code accumulates in float64 -- parity is a tolerance on the threshold, not bit-exactness.
#include "common.cuh"
#include <math.h>
#include <float.h>

namespace {

constexpr int GM_THREADS = 256;
constexpr int GM_CHUNK = 4096;      // samples per block: fixed, so partial sums are reproducible
constexpr int GM_MAXK = 6;

struct GmmArgs {
  double c0, c1;          // centres (k-means) / candidates
  double cand0, cand1;
  double w0, w1, m0, m1, p0, p1;   // EM: weights, means, precisions (1 / variance)
};

// functors: accumulate K sums for one sample
struct FMoments {   // sum x, sum x^2
  static constexpr int K = 2;
  __device__ void operator()(double x, const GmmArgs&, double* a) const { a[0] += x; a[1] += x * x; }
};
struct FPot {       // k-means++ potentials: sum d0^2, sum min(d0^2, dc0^2), sum min(d0^2, dc1^2)
  static constexpr int K = 3;
  __device__ void operator()(double x, const GmmArgs& g, double* a) const {
    double d0 = (x - g.c0) * (x - g.c0);
    double e0 = (x - g.cand0) * (x - g.cand0), e1 = (x - g.cand1) * (x - g.cand1);
    a[0] += d0; a[1] += fmin(d0, e0); a[2] += fmin(d0, e1);
  }
};
struct FLloyd {     // cluster 0 (nearest centre c0, ties to 0): count, sum x, sum x^2
  static constexpr int K = 3;
  __device__ void operator()(double x, const GmmArgs& g, double* a) const {
    double d0 = (x - g.c0) * (x - g.c0), d1 = (x - g.c1) * (x - g.c1);
    if (d0 <= d1) { a[0] += 1.0; a[1] += x; a[2] += x * x; }
  }
};
struct FEm {        // component 0 responsibilities: sum r, sum r x, sum r x^2; sum of log-likelihoods
  static constexpr int K = 4;
  __device__ void operator()(double x, const GmmArgs& g, double* a) const {
    const double LOG2PI = 1.8378770664093453;
    double l0 = -0.5 * (LOG2PI + (x - g.m0) * (x - g.m0) * g.p0) + 0.5 * log(g.p0) + log(g.w0);
    double l1 = -0.5 * (LOG2PI + (x - g.m1) * (x - g.m1) * g.p1) + 0.5 * log(g.p1) + log(g.w1);
    double mx = fmax(l0, l1);
    double lse = mx + log(exp(l0 - mx) + exp(l1 - mx));
    double r0 = exp(l0 - lse);
    a[0] += r0; a[1] += r0 * x; a[2] += r0 * x * x; a[3] += lse;
  }
};

template <typename F>
__global__ void __launch_bounds__(GM_THREADS) gmm_partial(const float* __restrict__ x, int64_t n, GmmArgs g,
                                                          double* __restrict__ partials) {
  constexpr int K = F::K;
  F f;
  double a[K];
#pragma unroll
  for (int k = 0; k < K; ++k) a[k] = 0.0;
  const int64_t beg = (int64_t)blockIdx.x * GM_CHUNK;
  const int64_t end = min(n, beg + GM_CHUNK);
  for (int64_t i = beg + threadIdx.x; i < end; i += GM_THREADS) f((double)x[i], g, a);
  __shared__ double sh[GM_THREADS / 32][K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double v = a[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double v = 0.0;
    for (int w = 0; w < GM_THREADS / 32; ++w) v += sh[w][threadIdx.x];
    partials[(int64_t)blockIdx.x * GM_MAXK + threadIdx.x] = v;
  }
}

// out[k] = sum over blocks, in block order (one thread per quantity)
__global__ void gmm_final(const double* __restrict__ partials, int G, int K, double* __restrict__ out) {
  int k = threadIdx.x;
  if (k >= K) return;
  double v = 0.0;
  for (int b = 0; b < G; ++b) v += partials[(int64_t)b * GM_MAXK + k];
  out[k] = v;
}

// np.searchsorted(np.cumsum(d0^2), target): first index whose running sum reaches the target.  The block is
// found on the host from the per-block sums (partials[.][0] of FPot); this walks that block in order.
__global__ void gmm_select(const float* __restrict__ x, int64_t n, double c0, const int* __restrict__ blk,
                           const double* __restrict__ before, const double* __restrict__ target, int ncand,
                           int64_t* __restrict__ idx_out, float* __restrict__ val_out) {
  int k = threadIdx.x;
  if (k >= ncand) return;
  int64_t beg = (int64_t)blk[k] * GM_CHUNK, end = min(n, beg + GM_CHUNK);
  double run = before[k];
  int64_t pick = end - 1;
  for (int64_t i = beg; i < end; ++i) {
    double d = (double)x[i] - c0;
    run += d * d;
    if (run >= target[k]) { pick = i; break; }
  }
  if (pick > n - 1) pick = n - 1;
  idx_out[k] = pick;
  val_out[k] = x[pick];
}

// the comparison is made in float64, as numpy makes it for a float32 array against a float64 scalar
__global__ void threshold_le_zero_f32(const float* __restrict__ in, int64_t n, double thr, float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { float v = in[i]; out[i] = ((double)v <= thr) ? 0.f : v; }
}

struct Reducer {
  tmat_ctx* ctx; const float* x; int64_t n; int G; double* partials; double* dsum; double* hsum; cudaStream_t st;
  template <typename F>
  int run(const GmmArgs& g, double* out) {
    gmm_partial<F><<<G, GM_THREADS, 0, st>>>(x, n, g, partials);
    KERNEL_CHECK(ctx);
    gmm_final<<<1, 32, 0, st>>>(partials, G, F::K, dsum);
    KERNEL_CHECK(ctx);
    CUDA_TRY(cudaMemcpyAsync(hsum, dsum, F::K * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < F::K; ++k) out[k] = hsum[k];
    return TMAT_OK;
  }
};

int gmm2_impl(tmat_ctx* ctx, const float* x, int64_t n, double sd_coef, const double* rnd, double* result,
              cudaStream_t st) {
  TMAT_REQUIRE(n >= 2, "tmat_gmm2_threshold: at least 2 samples are required (n=%lld)", (long long)n);
  CUDA_TRY(cudaSetDevice(ctx->device));
  const int G = (int)cdiv64(n, GM_CHUNK);
  Reducer R;
  R.ctx = ctx; R.x = x; R.n = n; R.G = G; R.st = st;
  TMAT_TRY(arena_get(ctx, (size_t)G * GM_MAXK, &R.partials));
  TMAT_TRY(arena_get(ctx, (size_t)GM_MAXK, &R.dsum));
  std::vector<double> hbuf(GM_MAXK), hpart;
  R.hsum = hbuf.data();
  GmmArgs g;
  memset(&g, 0, sizeof(g));
  double s[GM_MAXK];

  // ---- totals: sum x, sum x^2 -> variance (k-means tolerance) ----
  TMAT_TRY(R.run<FMoments>(g, s));
  const double N = (double)n, SX = s[0], SXX = s[1];
  const double mean = SX / N, var = fmax(SXX / N - mean * mean, 0.0);

  // ---- k-means++ seeding (sklearn/cluster/_kmeans.py::_kmeans_plusplus, n_local_trials = 2) ----
  // first centre: random_state.choice(n, p=uniform) = searchsorted(cdf, u, 'right') with cdf[i] = (i + 1) / n
  int64_t id0 = (int64_t)floor(rnd[0] * N);
  if (id0 > n - 1) id0 = n - 1;
  float c0f;
  CUDA_TRY(cudaMemcpyAsync(&c0f, x + id0, sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  g.c0 = (double)c0f;
  g.cand0 = g.cand1 = g.c0;
  TMAT_TRY(R.run<FPot>(g, s));
  const double pot = s[0];
  double c1 = g.c0;
  if (pot > 0.0) {
    // candidates: searchsorted(cumsum(d^2), rand * pot)
    hpart.resize((size_t)G * GM_MAXK);
    CUDA_TRY(cudaMemcpyAsync(hpart.data(), R.partials, hpart.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    int hblk[2]; double hbefore[2], htarget[2];
    for (int k = 0; k < 2; ++k) {
      htarget[k] = rnd[1 + k] * pot;
      double run = 0.0;
      int b = 0;
      for (; b < G - 1; ++b) {
        double nxt = run + hpart[(size_t)b * GM_MAXK];
        if (nxt >= htarget[k]) break;
        run = nxt;
      }
      hblk[k] = b; hbefore[k] = run;
    }
    int* dblk; double *dbefore, *dtarget; int64_t* didx; float* dval;
    TMAT_TRY(arena_get(ctx, 2, &dblk));
    TMAT_TRY(arena_get(ctx, 2, &dbefore));
    TMAT_TRY(arena_get(ctx, 2, &dtarget));
    TMAT_TRY(arena_get(ctx, 2, &didx));
    TMAT_TRY(arena_get(ctx, 2, &dval));
    CUDA_TRY(cudaMemcpyAsync(dblk, hblk, sizeof(hblk), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(dbefore, hbefore, sizeof(hbefore), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(dtarget, htarget, sizeof(htarget), cudaMemcpyHostToDevice, st));
    gmm_select<<<1, 32, 0, st>>>(x, n, g.c0, dblk, dbefore, dtarget, 2, didx, dval);
    KERNEL_CHECK(ctx);
    float hval[2];
    CUDA_TRY(cudaMemcpyAsync(hval, dval, sizeof(hval), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    g.cand0 = (double)hval[0]; g.cand1 = (double)hval[1];
    TMAT_TRY(R.run<FPot>(g, s));
    c1 = (s[1] <= s[2]) ? g.cand0 : g.cand1;     // np.argmin: first minimum
  }
  g.c1 = c1;

  // ---- Lloyd (max_iter 300, tol 1e-4 * var; stop on unchanged assignment or centre shift <= tol) ----
  const double ktol = 1e-4 * var;
  double cnt0 = -1.0, sum0 = 0.0, sq0 = 0.0;
  int kiter = 0;
  for (; kiter < 300; ++kiter) {
    TMAT_TRY(R.run<FLloyd>(g, s));
    const bool same = (s[0] == cnt0 && s[1] == sum0);
    cnt0 = s[0]; sum0 = s[1]; sq0 = s[2];
    if (same) break;                                            // strict convergence
    const double n0 = cnt0, n1 = N - cnt0;
    // an empty cluster keeps its centre (scikit-learn relocates it; with two 1-D clusters this needs identical data)
    const double nc0 = n0 > 0 ? sum0 / n0 : g.c0, nc1 = n1 > 0 ? (SX - sum0) / n1 : g.c1;
    const double shift = (nc0 - g.c0) * (nc0 - g.c0) + (nc1 - g.c1) * (nc1 - g.c1);
    g.c0 = nc0; g.c1 = nc1;
    if (shift <= ktol) {                                        // labels are re-derived from the final centres
      TMAT_TRY(R.run<FLloyd>(g, s));
      cnt0 = s[0]; sum0 = s[1]; sq0 = s[2];
      ++kiter;
      break;
    }
  }

  // ---- initial parameters from the one-hot responsibilities (_estimate_gaussian_parameters) ----
  const double EPS10 = 10.0 * DBL_EPSILON, REG = 1e-6;
  auto params_from = [&](double S0, double S1, double S2) {
    const double nk0 = S0 + EPS10, nk1 = (N - S0) + EPS10;
    g.m0 = S1 / nk0; g.m1 = (SX - S1) / nk1;
    double v0 = (S2 - 2.0 * g.m0 * S1 + g.m0 * g.m0 * S0) / nk0 + REG;
    double v1 = ((SXX - S2) - 2.0 * g.m1 * (SX - S1) + g.m1 * g.m1 * (N - S0)) / nk1 + REG;
    if (v0 < REG) v0 = REG;
    if (v1 < REG) v1 = REG;
    g.p0 = 1.0 / v0; g.p1 = 1.0 / v1;
    g.w0 = nk0 / N; g.w1 = nk1 / N;
    const double ws = g.w0 + g.w1;
    g.w0 /= ws; g.w1 /= ws;
  };
  params_from(cnt0, sum0, sq0);

  // ---- EM (tol 1e-3 on the mean log-likelihood, max_iter 100) ----
  double lower = -INFINITY;
  int it = 1, converged = 0;
  for (; it <= 100; ++it) {
    const double prev = lower;
    TMAT_TRY(R.run<FEm>(g, s));          // E-step with the current parameters
    params_from(s[0], s[1], s[2]);       // M-step
    lower = s[3] / N;
    if (fabs(lower - prev) < 1e-3) { converged = 1; break; }
  }

  const int fg = (g.m1 > g.m0) ? 1 : 0;       // np.argmax(means): first maximum
  const double mu = fg ? g.m1 : g.m0, sd = sqrt(1.0 / (fg ? g.p1 : g.p0));
  result[0] = fmin(255.0, mu + sd * sd_coef);
  result[1] = g.w0; result[2] = g.w1; result[3] = g.m0; result[4] = g.m1;
  result[5] = 1.0 / g.p0; result[6] = 1.0 / g.p1;
  result[7] = (double)(converged ? it : -it);
  result[8] = (double)kiter;
  result[9] = lower;
  return TMAT_OK;
}

}  // namespace

extern "C" int tmat_gmm2_threshold_dev(tmat_ctx* ctx, const float* x_dev, int64_t n, double sd_coef,
                                       const double* rand3_host, double* result10_host, void* stream) {
  TMAT_REQUIRE(ctx && x_dev && rand3_host && result10_host, "tmat_gmm2_threshold: NULL argument");
  TMAT_TRY(arena_reset(ctx));
  return gmm2_impl(ctx, x_dev, n, sd_coef, rand3_host, result10_host, as_stream(stream));
}

extern "C" int tmat_gmm2_threshold_host(tmat_ctx* ctx, const float* x_host, int64_t n, double sd_coef,
                                        const double* rand3_host, double* result10_host) {
  TMAT_REQUIRE(ctx && x_host && rand3_host && result10_host, "tmat_gmm2_threshold_host: NULL argument");
  TMAT_REQUIRE(n >= 2, "tmat_gmm2_threshold_host: at least 2 samples are required");
  TMAT_TRY(arena_reset(ctx));
  float* d;
  TMAT_TRY(arena_get(ctx, (size_t)n, &d));
  CUDA_TRY(cudaMemcpyAsync(d, x_host, (size_t)n * 4, cudaMemcpyHostToDevice, 0));
  return gmm2_impl(ctx, d, n, sd_coef, rand3_host, result10_host, 0);
}

// exec_threshold's last line: np.where(masked <= thr, 0, masked)
extern "C" int tmat_threshold_le_zero_f32_host(tmat_ctx* ctx, const float* img_host, int64_t n, double thr,
                                               float* out_host) {
  TMAT_REQUIRE(ctx && img_host && out_host && n >= 0, "tmat_threshold_le_zero_f32_host: bad argument");
  if (n == 0) return TMAT_OK;
  TMAT_TRY(arena_reset(ctx));
  float *d_in, *d_out;
  TMAT_TRY(arena_get(ctx, (size_t)n, &d_in));
  TMAT_TRY(arena_get(ctx, (size_t)n, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_in, img_host, (size_t)n * 4, cudaMemcpyHostToDevice, 0));
  threshold_le_zero_f32<<<(unsigned)cdiv64(n, 256), 256>>>(d_in, n, thr, d_out);
  KERNEL_CHECK(ctx);
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}
