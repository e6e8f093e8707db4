// Plain CUDA-core executor of a ConvOp (see unet.cuh).  It computes exactly what the tcgen05
// kernel (conv_tc.cu) computes -- same packed bf16 weights, same k-ordering of taps, same fused
// epilogue -- with one thread per (pixel, output channel).  It is the in-library cross-check
// for the tensor-core path and serves filter counts that are not multiples of 64.
#include "unet.cuh"

namespace {

struct SimtParams {
  const bf16* a0; int a0_C;
  const bf16* a1; int a1_C;
  int ntaps;
  int8_t dy[9], dx[9];
  const bf16* wmat; int K;
  EpiParams e;
};

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + __expf(-v)); }

// element-wise epilogue: value v (accumulator, no shift yet) for pixel (b,y,x), channel co
__device__ __forceinline__ void epilogue_elem(const EpiParams& e, int b, int y, int x, int co, float v) {
  v += e.shift[co];
  if (e.mode == EPI_HEAD) {
    int py = co >> 1, px = co & 1;
    float r = e.logits ? v : sigmoidf_(v);
    e.out_f32[((int64_t)b * 2 * e.H + 2 * y + py) * (2 * e.W) + 2 * x + px] = r;
    return;
  }
  if (e.mode == EPI_POOLADD) {
    int H2 = 2 * e.H, W2 = 2 * e.W;
    float m = -INFINITY;
    for (int ky = 0; ky < 3; ++ky) {
      int yy = 2 * y + ky;
      if (yy >= H2) continue;
      for (int kx = 0; kx < 3; ++kx) {
        int xx = 2 * x + kx;
        if (xx >= W2) continue;
        m = fmaxf(m, __bfloat162float(e.pool_src[(((int64_t)b * H2 + yy) * W2 + xx) * e.Cout + co]));
      }
    }
    v += m;
  }
  float primary = e.relu ? fmaxf(v, 0.f) : v;
  float second = fmaxf(v, 0.f);
  if (e.mode == EPI_UP2) {
    int H2 = 2 * e.H, W2 = 2 * e.W;
    for (int d = 0; d < 4; ++d) {
      int64_t o = (((int64_t)b * H2 + 2 * y + (d >> 1)) * W2 + 2 * x + (d & 1)) * e.Cout + co;
      e.out[o] = __float2bfloat16_rn(primary);
      if (e.out2) e.out2[o] = __float2bfloat16_rn(second);
    }
  } else {
    int64_t o = (((int64_t)b * e.H + y) * e.W + x) * e.Cout + co;
    e.out[o] = __float2bfloat16_rn(primary);
    if (e.out2) e.out2[o] = __float2bfloat16_rn(second);
  }
}

__global__ void __launch_bounds__(128) conv_simt_kernel(SimtParams p) {
  const EpiParams& e = p.e;
  int co = blockIdx.y * blockDim.x + threadIdx.x;
  int64_t pix = blockIdx.x;
  if (co >= e.Cout) return;
  int x = (int)(pix % e.W);
  int y = (int)((pix / e.W) % e.H);
  int b = (int)(pix / ((int64_t)e.W * e.H));
  const bf16* wrow = p.wmat + (int64_t)co * p.K;
  float acc = 0.f;
  for (int t = 0; t < p.ntaps; ++t) {
    int yy = y + p.dy[t], xx = x + p.dx[t];
    if (yy < 0 || yy >= e.H || xx < 0 || xx >= e.W) continue;
    const bf16* a = p.a0 + (((int64_t)b * e.H + yy) * e.W + xx) * p.a0_C;
    const bf16* wt = wrow + t * p.a0_C;
    for (int c = 0; c < p.a0_C; c += 2) {
      float2 av = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(a + c));
      float2 wv = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(wt + c));
      acc = fmaf(av.x, wv.x, acc);
      acc = fmaf(av.y, wv.y, acc);
    }
  }
  if (p.a1) {
    const bf16* a = p.a1 + (((int64_t)b * e.H + y) * e.W + x) * p.a1_C;
    const bf16* wt = wrow + p.ntaps * p.a0_C;
    for (int c = 0; c < p.a1_C; c += 2) {
      float2 av = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(a + c));
      float2 wv = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(wt + c));
      acc = fmaf(av.x, wv.x, acc);
      acc = fmaf(av.y, wv.y, acc);
    }
  }
  epilogue_elem(e, b, y, x, co, acc);
}

}  // namespace

int conv_run_simt(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st) {
  TMAT_REQUIRE(op.a0_C % 2 == 0 && op.a1_C % 2 == 0 && op.K % 2 == 0, "conv_simt(%s): odd channel count", op.name);
  SimtParams p;
  p.a0 = op.a0; p.a0_C = op.a0_C; p.a1 = op.a1; p.a1_C = op.a1_C;
  p.ntaps = op.ntaps;
  for (int t = 0; t < 9; ++t) { p.dy[t] = op.dy[t]; p.dx[t] = op.dx[t]; }
  p.wmat = op.wmat; p.K = op.K;
  p.e = epi_of(op);
  int64_t npix = (int64_t)op.B * op.H * op.W;
  TMAT_REQUIRE(npix <= 0x7fffffff, "conv_simt(%s): too many pixels", op.name);
  int threads = op.Cout >= 128 ? 128 : (op.Cout >= 64 ? 64 : 32);
  dim3 grid((unsigned)npix, cdiv(op.Cout, threads), 1);
  conv_simt_kernel<<<grid, threads, 0, st>>>(p);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}
