// Plain CUDA-core executor of a ConvOp (see unet.cuh).  It computes exactly what the tcgen05
// kernel (conv_tc.cu) computes -- same packed bf16 weights, same k-ordering of taps, same column
// groups, same fused epilogue -- with one thread per (pixel, GEMM column).  It is the in-library
// cross-check for the tensor-core path and serves filter counts that are not multiples of 64.
#include "unet.cuh"

namespace {

struct SimtParams {
  const bf16* a0; int a0_C; int a0_stride;
  const bf16* a1; int a1_C;
  int ngroups, ntaps, npad;
  int8_t dy[CONV_MAX_GROUPS][9], dx[CONV_MAX_GROUPS][9];
  const bf16* wmat; int K;
  EpiParams e;
};

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + __expf(-v)); }

// element-wise epilogue: accumulator v (no shift yet) of GEMM row (b,y,x), column n of group g
__device__ __forceinline__ void epilogue_elem(const EpiParams& e, int b, int y, int x, int g, int n, float v) {
  if (e.mode == EPI_HEAD) {
    v += e.shift[n];
    int py = n >> 1, px = n & 1;
    float r = e.logits ? v : sigmoidf_(v);
    e.out_f32[((int64_t)b * 2 * e.H + 2 * y + py) * (2 * e.W) + 2 * x + px] = r;
    return;
  }
  if (e.mode == EPI_PAR) {
    int q = g * e.ppg + n / e.Cout, co = n % e.Cout;
    v += e.shift[co];
    if (e.relu) v = fmaxf(v, 0.f);
    int64_t o = (((int64_t)b * 2 * e.H + 2 * y + (q >> 1)) * (2 * e.W) + 2 * x + (q & 1)) * e.Cout + co;
    e.out[o] = __float2bfloat16_rn(v);
    return;
  }
  v += e.shift[n];
  if (e.add_src) {
    int s = e.add_shift;
    int Ha = (e.H + (1 << s) - 1) >> s, Wa = (e.W + (1 << s) - 1) >> s;
    v += __bfloat162float(e.add_src[(((int64_t)b * Ha + (y >> s)) * Wa + (x >> s)) * e.Cout + n]);
  }
  int64_t o = (((int64_t)b * e.H + y) * e.W + x) * e.Cout + n;
  e.out[o] = __float2bfloat16_rn(e.relu ? fmaxf(v, 0.f) : v);
  if (e.out2) e.out2[o] = __float2bfloat16_rn(fmaxf(v, 0.f));
}

__global__ void __launch_bounds__(128) conv_simt_kernel(SimtParams p) {
  const EpiParams& e = p.e;
  int col = blockIdx.y * blockDim.x + threadIdx.x;   // column over all groups
  int64_t pix = blockIdx.x;
  int g = col / p.npad, n = col - g * p.npad;
  if (g >= p.ngroups) return;
  if (e.mode == EPI_HEAD && n >= 4) return;
  int x = (int)(pix % e.W);
  int y = (int)((pix / e.W) % e.H);
  int b = (int)(pix / ((int64_t)e.W * e.H));
  const bf16* wrow = p.wmat + (int64_t)col * p.K;
  float acc = 0.f;
  for (int t = 0; t < p.ntaps; ++t) {
    int yy = y + p.dy[g][t], xx = x + p.dx[g][t];
    if (yy < 0 || yy >= e.H || xx < 0 || xx >= e.W) continue;
    const int sa = p.a0_stride;
    const bf16* a = p.a0 + (((int64_t)b * e.H * sa + yy * sa) * e.W * sa + xx * sa) * p.a0_C;
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
  epilogue_elem(e, b, y, x, g, n, acc);
}

}  // namespace

int conv_run_simt(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st) {
  TMAT_REQUIRE(op.a0_C % 2 == 0 && op.a1_C % 2 == 0 && op.K % 2 == 0, "conv_simt(%s): odd channel count", op.name);
  TMAT_REQUIRE(op.ngroups >= 1 && op.ngroups <= CONV_MAX_GROUPS && (op.ngroups == 1 || !op.a1),
               "conv_simt(%s): bad group configuration", op.name);
  SimtParams p;
  p.a0 = op.a0; p.a0_C = op.a0_C; p.a0_stride = op.a0_stride; p.a1 = op.a1; p.a1_C = op.a1_C;
  TMAT_REQUIRE(op.a0_stride == 1 || (op.ntaps == 1 && op.ngroups == 1), "conv_simt(%s): strided a0 needs a single tap", op.name);
  p.ngroups = op.ngroups; p.ntaps = op.ntaps; p.npad = op.Npad;
  memcpy(p.dy, op.dy, sizeof(p.dy));
  memcpy(p.dx, op.dx, sizeof(p.dx));
  p.wmat = op.wmat; p.K = op.K;
  p.e = epi_of(op);
  int64_t npix = (int64_t)op.B * op.H * op.W;
  TMAT_REQUIRE(npix <= 0x7fffffff, "conv_simt(%s): too many pixels", op.name);
  int cols = op.ngroups * op.Npad;
  int threads = cols >= 128 ? 128 : (cols >= 64 ? 64 : 32);
  dim3 grid((unsigned)npix, cdiv(cols, threads), 1);
  conv_simt_kernel<<<grid, threads, 0, st>>>(p);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}
