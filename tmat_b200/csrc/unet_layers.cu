// CUDA-core layers of the UNet-Xception forward pass (reference: fl_tissue_model_tools/models.py:110-171):
//   stem Conv2D(3x3, stride 2) + BN + ReLU (:119-121), depthwise 3x3 of SeparableConv2D (:129,:134),
//   and a plain implicit-GEMM executor for ConvOp used to cross-check the tcgen05 kernel.
#include "unet.cuh"
#include <algorithm>

namespace {

// ---- stem ------------------------------------------------------------------------------------
// TF 'same' with stride 2 on an even size pads 0 before / 1 after: out[y,x] = sum in[2y+ky, 2x+kx].
// A block produces a strip of STEM_ROWS output rows x 32 output columns of one patch: the input
// window (2*STEM_ROWS+1 rows x 65 columns) is fetched once into shared memory -- through the D4 tile
// map when the patches come straight from the image -- and every thread computes 8 output channels
// of one column for all rows of the strip with its 72 weights in registers.  A warp stores
// 4 pixels x 128 bytes = 512 contiguous bytes.
constexpr int STEM_ROWS = 16;
constexpr int STEM_COLS = 32;

__global__ void __launch_bounds__(256, 2) stem_kernel(StemSrc src, int B, int ws, const float* __restrict__ w,
                                                   const float* __restrict__ shift, int C0,
                                                   bf16* __restrict__ out, int tiles_x, int tiles_y) {
  // Persistent blocks walk the (patch, strip) tiles; the input window of the NEXT tile is fetched into
  // registers before the current tile is computed and lands in the other shared-memory buffer afterwards,
  // so the global-load latency of the fill hides behind the FFMA2 work.  Weights are loaded once per block.
  constexpr int IR = 2 * STEM_ROWS + 1, IC = 2 * STEM_COLS + 1;
  constexpr int FILL = (IR * IC + 255) / 256;
  // every window element is stored twice, as the (v, v) pair FFMA2 wants for a pixel broadcast over two channels:
  // one 8-byte ld.shared instead of a 4-byte one plus two register moves per tap (ptxas: 144 IMAD.MOV per 8 rows)
  __shared__ __align__(8) float2 tile[2][IR][IC + 1];
  const int S = ws >> 1;
  const int total = tiles_x * tiles_y * B;
  // window of tile t -> tile[buf], asynchronously (cp.async, 4 bytes per element; padding / out-of-range
  // elements are plain stores of the constant), so no registers are held while the loads are in flight
  auto fill_async = [&](int t, int buf) {
    const int b = t / (tiles_x * tiles_y), rem = t - b * (tiles_x * tiles_y);
    const int y0 = (rem / tiles_x) * STEM_ROWS, x0 = (rem % tiles_x) * STEM_COLS;
    float minval = 0.f;
    const float* pp = nullptr;
    // the D4 tile map is affine in (i, j): three probes give its coefficients for the whole block
    int r00 = 0, c00 = 0, dri = 0, dci = 0, drj = 0, dcj = 0;
    if (src.use_geom) {
      const StpGeom& G = src.geom;
      int p = src.first_patch + b;
      if (src.patch_list) p = __ldg(src.patch_list + p);
      const int per = G.tiles_per_orient();
      const int g = p / per;
      const int tt = p - g * per;
      const int tc = G.tile_cols(g);
      const int ti = tt / tc, tj = tt - ti * tc;
      minval = src.minmax[0];
      int r10, c10, r01, c01;
      d4_rot_to_orig(g, ti * G.step, tj * G.step, G.PM, G.PN, &r00, &c00);
      d4_rot_to_orig(g, ti * G.step + 1, tj * G.step, G.PM, G.PN, &r10, &c10);
      d4_rot_to_orig(g, ti * G.step, tj * G.step + 1, G.PM, G.PN, &r01, &c01);
      dri = r10 - r00; dci = c10 - c00; drj = r01 - r00; dcj = c01 - c00;
    } else {
      pp = src.patches + (int64_t)b * ws * ws;
    }
#pragma unroll
    for (int k = 0; k < FILL; ++k) {
      const int idx = threadIdx.x + k * 256;
      if (idx >= IR * IC) break;
      const int r = idx / IC, c = idx - r * IC;
      const int i = 2 * y0 + r, j = 2 * x0 + c;
      float2* dst = &tile[buf][r][c];
      const float* gp = nullptr;
      float cv = 0.f;
      if (i < ws && j < ws) {
        if (src.use_geom) {
          const StpGeom& G = src.geom;
          const int y = r00 + i * dri + j * drj - G.aug, x = c00 + i * dci + j * dcj - G.aug;
          if (y >= 0 && y < G.H && x >= 0 && x < G.W) gp = src.img + (int64_t)y * G.W + x;
          else cv = minval;
        } else {
          gp = pp + (int64_t)i * ws + j;
        }
      }
      if (gp != nullptr) {
        const uint32_t da = (uint32_t)__cvta_generic_to_shared(dst);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(da), "l"(gp) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(da + 4u), "l"(gp) : "memory");
      } else {
        *dst = make_float2(cv, cv);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int groups = C0 >> 3;                  // 8 when C0 == 64
  const int px_per_pass = blockDim.x / groups;
  const int cg = threadIdx.x % groups;
  f32x2 wr[9][4], sh[4];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const float4* wp = reinterpret_cast<const float4*>(w + t * C0 + cg * 8);
    float4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
    wr[t][0] = pack2(w0.x, w0.y); wr[t][1] = pack2(w0.z, w0.w);
    wr[t][2] = pack2(w1.x, w1.y); wr[t][3] = pack2(w1.z, w1.w);
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) sh[c] = pack2(__ldg(shift + cg * 8 + 2 * c), __ldg(shift + cg * 8 + 2 * c + 1));
  int t = blockIdx.x;
  if (t < total) fill_async(t, 0);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  for (int buf = 0; t < total; t += gridDim.x, buf ^= 1) {
    const int tn = t + gridDim.x;
    if (tn < total) fill_async(tn, buf ^ 1);
    const int b = t / (tiles_x * tiles_y), rem = t - b * (tiles_x * tiles_y);
    const int y0 = (rem / tiles_x) * STEM_ROWS, x0 = (rem % tiles_x) * STEM_COLS;
    for (int xl = threadIdx.x / groups; xl < STEM_COLS; xl += px_per_pass) {
      const int x = x0 + xl;
      if (x >= S) break;
      // one 64-bit address per column, advanced by a row per iteration
      bf16* op = out + (((int64_t)b * S + y0) * S + x) * C0 + cg * 8;
      const int64_t row_stride = (int64_t)S * C0;
#pragma unroll
      for (int r = 0; r < STEM_ROWS; ++r, op += row_stride) {
        const int y = y0 + r;
        if (y >= S) break;
        f32x2 acc[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[c] = sh[c];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const f32x2 vv = *reinterpret_cast<const f32x2*>(&tile[buf][2 * r + ky][2 * xl + kx]);
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[c] = ffma2(vv, wr[ky * 3 + kx][c], acc[c]);
          }
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float lo, hi;
          unpack2(acc[c], lo, hi);
          asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(ow[c]) : "f"(hi), "f"(lo));   // ReLU inside the convert
        }
        *reinterpret_cast<uint4*>(op) = o;
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
  }
}

// ---- depthwise 3x3 ---------------------------------------------------------------------------
__device__ __forceinline__ void unpack4(const uint2& v, bool relu, float (&f)[4]) {
  __nv_bfloat162 h0 = *reinterpret_cast<const __nv_bfloat162*>(&v.x);
  __nv_bfloat162 h1 = *reinterpret_cast<const __nv_bfloat162*>(&v.y);
  if (relu) {
    const __nv_bfloat162 z = __floats2bfloat162_rn(0.f, 0.f);
    h0 = __hmax2(h0, z); h1 = __hmax2(h1, z);
  }
  float2 a = __bfloat1622float2(h0), b = __bfloat1622float2(h1);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}

// Column-sliding depthwise 3x3: a thread owns 4 channels of one image column x and walks down a strip
// of rows.  Each input row (3 x-neighbours, 8 bytes each) is loaded and converted once and scattered
// into the three output rows it contributes to; the 36 weights stay in registers.  Consecutive
// threads are consecutive (x, channel-quad) pairs, so every load / store of a warp is 256 contiguous
// bytes and the x-neighbour loads hit L1.
__global__ void __launch_bounds__(256) dwconv_kernel(const bf16* __restrict__ in, int B, int S, int C, int rows,
                                                     const float* __restrict__ w, int relu_in,
                                                     bf16* __restrict__ out, bf16* __restrict__ sub) {
  const int quads = C >> 2;
  const int col = blockIdx.x * blockDim.x + threadIdx.x;   // (x, channel quad)
  if (col >= S * quads) return;
  const int x = col / quads, cq = col - x * quads;
  const int y0 = blockIdx.y * rows;
  const int y1 = min(S, y0 + rows);
  const int b = blockIdx.z;
  float wr[9][4];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    float4 wv = __ldg(reinterpret_cast<const float4*>(w + t * C + cq * 4));
    wr[t][0] = wv.x; wr[t][1] = wv.y; wr[t][2] = wv.z; wr[t][3] = wv.w;
  }
  const bool relu = relu_in != 0;
  const bool has_l = x > 0, has_r = x + 1 < S;
  const bf16* base = in + ((int64_t)b * S * S + x) * C + cq * 4;        // row 0, this column
  bf16* obase = out + ((int64_t)b * S * S + x) * C + cq * 4;
  const int64_t rstride = (int64_t)S * C;
  // acc0: output row r-1 (complete after input row r), acc1: output row r, acc2: output row r+1
  float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f}, acc2[4];
  const int rbeg = max(y0 - 1, 0), rend = min(y1 + 1, S);   // input rows needed
  uint2 nl = make_uint2(0, 0), nc = nl, nr = nl;
  {
    const bf16* p = base + rbeg * rstride;
    nc = __ldg(reinterpret_cast<const uint2*>(p));
    if (has_l) nl = __ldg(reinterpret_cast<const uint2*>(p - C));
    if (has_r) nr = __ldg(reinterpret_cast<const uint2*>(p + C));
  }
  for (int r = rbeg; r < rend; ++r) {
    const uint2 cl = nl, cc = nc, cr = nr;
    if (r + 1 < rend) {                               // prefetch the next input row
      const bf16* p = base + (int64_t)(r + 1) * rstride;
      nc = __ldg(reinterpret_cast<const uint2*>(p));
      if (has_l) nl = __ldg(reinterpret_cast<const uint2*>(p - C));
      if (has_r) nr = __ldg(reinterpret_cast<const uint2*>(p + C));
    }
    if (sub != nullptr && ((r | x) & 1) == 0 && r >= y0 && r < y1) {
      const int S2 = S >> 1;
      *reinterpret_cast<uint2*>(sub + (((int64_t)b * S2 + (r >> 1)) * S2 + (x >> 1)) * C + cq * 4) = cc;
    }
    float fl[4], fc[4], fr[4];
    unpack4(cl, relu, fl); unpack4(cc, relu, fc); unpack4(cr, relu, fr);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      // input row r is tap row ky=2 of output r-1, ky=1 of output r, ky=0 of output r+1
      acc0[c] = fmaf(fl[c], wr[6][c], acc0[c]); acc0[c] = fmaf(fc[c], wr[7][c], acc0[c]); acc0[c] = fmaf(fr[c], wr[8][c], acc0[c]);
      acc1[c] = fmaf(fl[c], wr[3][c], acc1[c]); acc1[c] = fmaf(fc[c], wr[4][c], acc1[c]); acc1[c] = fmaf(fr[c], wr[5][c], acc1[c]);
      acc2[c] = fl[c] * wr[0][c];               acc2[c] = fmaf(fc[c], wr[1][c], acc2[c]); acc2[c] = fmaf(fr[c], wr[2][c], acc2[c]);
    }
    const int yo = r - 1;                             // output row completed by this input row
    if (yo >= y0 && yo < y1) {
      __nv_bfloat162 h0 = __floats2bfloat162_rn(acc0[0], acc0[1]), h1 = __floats2bfloat162_rn(acc0[2], acc0[3]);
      uint2 o = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
      *reinterpret_cast<uint2*>(obase + (int64_t)yo * rstride) = o;
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) { acc0[c] = acc1[c]; acc1[c] = acc2[c]; }
  }
  // last output row of the strip when it is the last image row (no input row below it)
  if (rend == S && y1 == S) {
    const int yo = S - 1;
    if (yo >= y0) {
      __nv_bfloat162 h0 = __floats2bfloat162_rn(acc0[0], acc0[1]), h1 = __floats2bfloat162_rn(acc0[2], acc0[3]);
      uint2 o = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
      *reinterpret_cast<uint2*>(obase + (int64_t)yo * rstride) = o;
    }
  }
}

// ---- MaxPool(3, stride 2, same) ----------------------------------------------------------------
// one thread: 8 channels of one output pixel; window rows 2y..2y+2, cols 2x..2x+2 clipped to the image
__global__ void __launch_bounds__(256) maxpool_kernel(const bf16* __restrict__ in, int B, int S, int C,
                                                      bf16* __restrict__ out) {
  const int groups = C >> 3;
  const int S2 = S >> 1;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)B * S2 * S2 * groups;
  if (idx >= total) return;
  int cg = (int)(idx % groups);
  int64_t pix = idx / groups;
  int x = (int)(pix % S2);
  int y = (int)((pix / S2) % S2);
  int b = (int)(pix / ((int64_t)S2 * S2));
  const __nv_bfloat162 ninf = __floats2bfloat162_rn(-INFINITY, -INFINITY);
  __nv_bfloat162 m[4] = {ninf, ninf, ninf, ninf};
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    int yy = 2 * y + ky;
    if (yy >= S) continue;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      int xx = 2 * x + kx;
      if (xx >= S) continue;
      uint4 v = __ldg(reinterpret_cast<const uint4*>(in + (((int64_t)b * S + yy) * S + xx) * C + cg * 8));
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int i = 0; i < 4; ++i) m[i] = __hmax2(m[i], h[i]);
    }
  }
  uint4 o;
  o.x = *reinterpret_cast<uint32_t*>(&m[0]); o.y = *reinterpret_cast<uint32_t*>(&m[1]);
  o.z = *reinterpret_cast<uint32_t*>(&m[2]); o.w = *reinterpret_cast<uint32_t*>(&m[3]);
  *reinterpret_cast<uint4*>(out + pix * C + cg * 8) = o;
}

}  // namespace

int stem_run(tmat_ctx* ctx, const StemSrc& src, int B, int ws, const float* w, const float* shift, int C0,
             bf16* out, cudaStream_t st) {
  int S = ws / 2;
  TMAT_REQUIRE(C0 % 8 == 0 && C0 / 8 <= 256 && 256 % (C0 / 8) == 0, "stem: unsupported channel count %d", C0);
  TMAT_REQUIRE(B <= 65535, "stem: batch %d too large", B);
  const int tiles_x = cdiv(S, STEM_COLS), tiles_y = cdiv(S, STEM_ROWS);
  const int64_t total = (int64_t)tiles_x * tiles_y * B;
  TMAT_REQUIRE(total < (1ll << 31), "stem: too many tiles");
  const int grid = (int)std::min<int64_t>(total, 2 * (int64_t)ctx->sm_count);   // persistent, two blocks per SM
  stem_kernel<<<grid, 256, 0, st>>>(src, B, ws, w, shift, C0, out, tiles_x, tiles_y);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

int dwconv_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, const float* w, int relu_in, bf16* out,
               bf16* sub, cudaStream_t st) {
  TMAT_REQUIRE(C % 4 == 0 && B <= 65535, "dwconv: unsupported shape (C=%d, B=%d)", C, B);
  // strips of ~20 rows: 2 halo rows per strip are re-read (10 %), and the grid stays >= 2 waves
  int nstrips = cdiv(S, 20);
  int rows = cdiv(S, nstrips);
  if (sub && (rows & 1)) rows++;                // keep strip boundaries even (stride-2 sample rows)
  nstrips = cdiv(S, rows);
  dim3 grid(cdiv(S * (C / 4), 256), nstrips, B);
  dwconv_kernel<<<grid, 256, 0, st>>>(in, B, S, C, rows, w, relu_in, out, sub);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

int maxpool_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, bf16* out, cudaStream_t st) {
  TMAT_REQUIRE(C % 8 == 0 && S % 2 == 0, "maxpool: unsupported shape (S=%d, C=%d)", S, C);
  int64_t total = (int64_t)B * (S / 2) * (S / 2) * (C / 8);
  maxpool_kernel<<<(unsigned)cdiv64(total, 256), 256, 0, st>>>(in, B, S, C, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}
