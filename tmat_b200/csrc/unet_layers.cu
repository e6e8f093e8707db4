// CUDA-core layers of the UNet-Xception forward pass (reference: fl_tissue_model_tools/models.py:110-171):
//   stem Conv2D(3x3, stride 2) + BN + ReLU (:119-121), depthwise 3x3 of SeparableConv2D (:129,:134),
//   and a plain implicit-GEMM executor for ConvOp used to cross-check the tcgen05 kernel.
#include "unet.cuh"
#include <algorithm>

namespace {

// ---- stem ------------------------------------------------------------------------------------
// TF 'same' with stride 2 on an even size pads 0 before / 1 after: out[y,x] = sum in[2y+ky, 2x+kx].
// One thread computes 8 output channels of one output pixel (16-byte bf16 store).
__global__ void __launch_bounds__(256) stem_kernel(StemSrc src, int B, int ws, const float* __restrict__ w,
                                                   const float* __restrict__ shift, int C0,
                                                   bf16* __restrict__ out) {
  const int S = ws >> 1;
  const int groups = C0 >> 3;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)B * S * S * groups;
  if (idx >= total) return;
  int cg = (int)(idx % groups);
  int64_t pix = idx / groups;
  int x = (int)(pix % S);
  int y = (int)((pix / S) % S);
  int b = (int)(pix / ((int64_t)S * S));
  float in[9];
  if (src.use_geom) {
    const StpGeom& G = src.geom;
    int p = src.first_patch + b;
    int per = G.tiles_per_orient();
    int g = p / per, t = p - g * per;
    int tc = G.tile_cols(g);
    int ti = t / tc, tj = t - ti * tc;
    float minval = src.minmax[0];
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        int i = 2 * y + ky, j = 2 * x + kx;
        in[ky * 3 + kx] = (i < ws && j < ws) ? stp_tile_pixel(src.img, G, g, ti, tj, i, j, minval) : 0.f;
      }
  } else {
    const float* pp = src.patches + (int64_t)b * ws * ws;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        int i = 2 * y + ky, j = 2 * x + kx;
        in[ky * 3 + kx] = (i < ws && j < ws) ? __ldg(pp + (int64_t)i * ws + j) : 0.f;
      }
  }
  float acc[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) acc[c] = __ldg(shift + cg * 8 + c);
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const float4* wp = reinterpret_cast<const float4*>(w + t * C0 + cg * 8);
    float4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
    acc[0] = fmaf(in[t], w0.x, acc[0]); acc[1] = fmaf(in[t], w0.y, acc[1]);
    acc[2] = fmaf(in[t], w0.z, acc[2]); acc[3] = fmaf(in[t], w0.w, acc[3]);
    acc[4] = fmaf(in[t], w1.x, acc[4]); acc[5] = fmaf(in[t], w1.y, acc[5]);
    acc[6] = fmaf(in[t], w1.z, acc[6]); acc[7] = fmaf(in[t], w1.w, acc[7]);
  }
  uint4 o;
  __nv_bfloat162 h;
  h = __floats2bfloat162_rn(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f)); o.x = *reinterpret_cast<uint32_t*>(&h);
  h = __floats2bfloat162_rn(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f)); o.y = *reinterpret_cast<uint32_t*>(&h);
  h = __floats2bfloat162_rn(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f)); o.z = *reinterpret_cast<uint32_t*>(&h);
  h = __floats2bfloat162_rn(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f)); o.w = *reinterpret_cast<uint32_t*>(&h);
  *reinterpret_cast<uint4*>(out + pix * C0 + cg * 8) = o;
}

// ---- depthwise 3x3 ---------------------------------------------------------------------------
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x; f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 o;
  __nv_bfloat162 h;
  h = __floats2bfloat162_rn(f[0], f[1]); o.x = *reinterpret_cast<uint32_t*>(&h);
  h = __floats2bfloat162_rn(f[2], f[3]); o.y = *reinterpret_cast<uint32_t*>(&h);
  h = __floats2bfloat162_rn(f[4], f[5]); o.z = *reinterpret_cast<uint32_t*>(&h);
  h = __floats2bfloat162_rn(f[6], f[7]); o.w = *reinterpret_cast<uint32_t*>(&h);
  return o;
}

// one thread: 8 channels of one pixel
__global__ void __launch_bounds__(256) dwconv_kernel(const bf16* __restrict__ in, int B, int S, int C,
                                                     const float* __restrict__ w, int relu_in,
                                                     bf16* __restrict__ out, bf16* __restrict__ sub) {
  const int groups = C >> 3;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)B * S * S * groups;
  if (idx >= total) return;
  int cg = (int)(idx % groups);
  int64_t pix = idx / groups;
  int x = (int)(pix % S);
  int y = (int)((pix / S) % S);
  int b = (int)(pix / ((int64_t)S * S));
  float acc[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) acc[c] = 0.f;
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    int yy = y + ky - 1;
    if (yy < 0 || yy >= S) continue;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      int xx = x + kx - 1;
      if (xx < 0 || xx >= S) continue;
      uint4 v = __ldg(reinterpret_cast<const uint4*>(in + (((int64_t)b * S + yy) * S + xx) * C + cg * 8));
      if (ky == 1 && kx == 1 && sub != nullptr && ((x | y) & 1) == 0) {
        int S2 = S >> 1;
        *reinterpret_cast<uint4*>(sub + (((int64_t)b * S2 + (y >> 1)) * S2 + (x >> 1)) * C + cg * 8) = v;
      }
      float f[8];
      unpack8(v, f);
      const float4* wp = reinterpret_cast<const float4*>(w + (ky * 3 + kx) * C + cg * 8);
      float4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
      float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float a = relu_in ? fmaxf(f[c], 0.f) : f[c];
        acc[c] = fmaf(a, wv[c], acc[c]);
      }
    }
  }
  *reinterpret_cast<uint4*>(out + pix * C + cg * 8) = pack8(acc);
}

}  // namespace

int stem_run(tmat_ctx* ctx, const StemSrc& src, int B, int ws, const float* w, const float* shift, int C0,
             bf16* out, cudaStream_t st) {
  int S = ws / 2;
  int64_t total = (int64_t)B * S * S * (C0 / 8);
  stem_kernel<<<(unsigned)cdiv64(total, 256), 256, 0, st>>>(src, B, ws, w, shift, C0, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

int dwconv_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, const float* w, int relu_in, bf16* out,
               bf16* sub, cudaStream_t st) {
  int64_t total = (int64_t)B * S * S * (C / 8);
  dwconv_kernel<<<(unsigned)cdiv64(total, 256), 256, 0, st>>>(in, B, S, C, w, relu_in, out, sub);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}
