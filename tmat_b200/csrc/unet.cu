// UNet-Xception patch segmentor: weight packing, forward pass, fused tiled prediction.
// Reference: fl_tissue_model_tools/models.py:85-171 (build_UNetXception), :624-653 (predict),
// fl_tissue_model_tools/smooth_tiled_predictions.py:220-267.
//
// Algebra used (exact in real arithmetic):
//   * BatchNorm (inference, eps = 1e-3) and the conv bias fold into a per-channel scale (multiplied
//     into the bf16 weights) and a per-channel fp32 shift applied in the epilogue.
//   * Conv2DTranspose(3, stride 1, same) is a 3x3 correlation with the spatially flipped kernel.
//   * Up block: up2(BN(ct2)) + conv1x1(up2(prev)) = up2(Y), Y = BN(ct2) + conv1x1(prev).  Only the
//     low-resolution Y is ever stored; the UpSampling2D tensors never exist:
//       - the next block's ct1, a 3x3 conv on up2(relu(Y)), is evaluated per output-pixel parity on
//         relu(Y) itself (taps that hit the same low-res pixel are summed: 4 taps per parity instead
//         of 9), written to the fine grid by parity-strided TMA stores (EPI_PAR);
//       - the next block's residual conv1x1(up2(up2(Y)..)) commutes with the upsampling: Z = conv1x1(Y)
//         is a small GEMM at Y's resolution, added as Z[y/2, x/2] in the ct2 epilogue.
//     In block 0 the residual input has the GEMM's own resolution and rides as extra k-steps of ct2.
//   * Head: the same parity evaluation (N = 4 parities) on the last block's Y, so the 320x320x64
//     tensor never exists.
#include "unet.cuh"
#include <math.h>
#include <stdlib.h>
#include <algorithm>

struct Tap {
  const void* ptr; int B, H, W, C; int is_f32; int up;   // up: report the 2x nearest-upsampled tensor
};

struct tmat_unet {
  tmat_unet_desc desc;
  int fc[4];
  int device = 0;
  bool use_tc = false;
  bool fuse_dw = true;     // SeparableConv2D as one tcgen05 kernel (TMAT_NO_FUSE_DW=1 runs depthwise separately)
  std::vector<void*> allocs;
  float* stem_w = nullptr; float* stem_shift = nullptr;
  struct Down { int cin, f; float *dw1, *dw2; bf16 *pw1, *pw2, *res; float *sh1, *sh2, *shres; } down[3];
  // up[0]: ct1 plain [f][9 cin], ct2 = [f][9 f + cin] (residual as extra k-steps);
  // up[i>0]: ct1 parity-packed [4 f][gtaps cin], ct2 = [f][9 f], z = [f][cin] (residual at low resolution)
  struct Up { int cin, f; bf16 *ct1, *ct2, *z; float *sh1, *sh2, *shz; int ngroups, gtaps; } up[4];
  bf16* head_w = nullptr; float* head_shift = nullptr;
  bf16* head_w2 = nullptr;        // [16 = (parity, 2x2 low-res tap)][C] for the taps-in-N head (conv_head.cu)
  std::map<std::string, Tap> taps;
  bool keep_taps = false;
};

namespace {

constexpr float BN_EPS = 1e-3f;

struct WeightSet {
  std::map<std::string, std::pair<const float*, int64_t>> m;
  int get(const std::string& name, int64_t expect, const float** out) const {
    auto it = m.find(name);
    if (it == m.end()) { tmat_set_error("tmat_unet_load: missing array '%s'", name.c_str()); return TMAT_ERR_INVALID; }
    if (it->second.second != expect) {
      tmat_set_error("tmat_unet_load: array '%s' has %lld elements, expected %lld", name.c_str(),
                     (long long)it->second.second, (long long)expect);
      return TMAT_ERR_INVALID;
    }
    *out = it->second.first;
    return TMAT_OK;
  }
};

// scale / shift of  BN(conv + bias)  (or identity BN when prefix is empty)
int bn_fold(const WeightSet& ws, const std::string& bn, const float* bias, int C, std::vector<float>& scale,
            std::vector<float>& shift) {
  scale.assign(C, 1.f);
  shift.assign(C, 0.f);
  if (bn.empty()) {
    for (int c = 0; c < C; ++c) shift[c] = bias[c];
    return TMAT_OK;
  }
  const float *g, *b, *m, *v;
  TMAT_TRY(ws.get(bn + ".gamma", C, &g));
  TMAT_TRY(ws.get(bn + ".beta", C, &b));
  TMAT_TRY(ws.get(bn + ".mean", C, &m));
  TMAT_TRY(ws.get(bn + ".var", C, &v));
  for (int c = 0; c < C; ++c) {
    float s = g[c] / sqrtf(v[c] + BN_EPS);
    scale[c] = s;
    shift[c] = s * (bias[c] - m[c]) + b[c];
  }
  return TMAT_OK;
}

template <typename T>
int upload(tmat_unet* net, const std::vector<T>& h, T** d) {
  CUDA_TRY(cudaMalloc((void**)d, h.size() * sizeof(T)));
  net->allocs.push_back(*d);
  CUDA_TRY(cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return TMAT_OK;
}

int upload_bf16(tmat_unet* net, const std::vector<float>& h, bf16** d) {
  std::vector<bf16> hb(h.size());
  for (size_t i = 0; i < h.size(); ++i) hb[i] = __float2bfloat16_rn(h[i]);
  return upload(net, hb, d);
}

// Parity evaluation of a 3x3 'same' convolution on UpSampling2D(2)(x): output pixel (2y+py, 2x+px)
// reads fine pixel (2y+py+ky-1, ..) = low-res pixel y + lowtap(py, ky): lowtap(0,.) = {-1,0,0},
// lowtap(1,.) = {0,0,1}.  The fine zero padding coincides with the low-res zero padding.
int parity_lowtap(int par, int kk) { int v = par + kk - 1; return v < 0 ? -1 : v / 2; }

// Column groups of the parity GEMM: enough parities are merged into one group (sharing the union of
// their taps, with zero weight blocks) to give the MMA an N of 128 .. 256.
struct ParityPlan { int ngroups, gtaps, ppg; int8_t dy[CONV_MAX_GROUPS][9], dx[CONV_MAX_GROUPS][9]; };
ParityPlan parity_plan(int f) {
  ParityPlan pl;
  memset(&pl, 0, sizeof(pl));
  // f = 64: two groups of N = 128 over 6 taps beat one group of N = 256 over 9 (a third fewer MMA cycles,
  // measured -0.8 ms per 2048^2 image); f = 128: two groups of N = 256 and four of N = 128 measure the same
  pl.ngroups = f >= 256 ? 4 : (f >= 64 ? 2 : 1);
  {
    // experiment hook: TMAT_PAR_GROUPS="<groups for f < 128>,<groups for 128 <= f < 256>"
    static const char* env = getenv("TMAT_PAR_GROUPS");
    int g64 = 0, g128 = 0;
    if (env && sscanf(env, "%d,%d", &g64, &g128) == 2) {
      if (f >= 64 && f < 128 && (g64 == 1 || g64 == 2 || g64 == 4)) pl.ngroups = g64;
      if (f >= 128 && f < 256 && (g128 == 1 || g128 == 2 || g128 == 4)) pl.ngroups = g128;
    }
  }
  pl.ppg = 4 / pl.ngroups;
  for (int g = 0; g < pl.ngroups; ++g) {
    int n = 0;
    for (int ly = -1; ly <= 1; ++ly)
      for (int lx = -1; lx <= 1; ++lx) {
        bool used = false;
        for (int ql = 0; ql < pl.ppg; ++ql) {
          int q = g * pl.ppg + ql, py = q >> 1, px = q & 1;
          bool uy = (ly == py - 1 || ly == py), ux = (lx == px - 1 || lx == px);
          used = used || (uy && ux);
        }
        if (used) { pl.dy[g][n] = (int8_t)ly; pl.dx[g][n] = (int8_t)lx; ++n; }
      }
    pl.gtaps = n;   // 4, 6 or 9 -- the same for every group of a plan
  }
  return pl;
}

int pack_weights(tmat_unet* net, const WeightSet& ws) {
  const int ch = net->desc.channels;
  const int* fc = net->fc;
  std::vector<float> scale, shift;
  const float *k, *b;
  // stem: Conv2D(fc0, 3, stride 2) + BN + ReLU   (models.py:119-121)
  {
    int C0 = fc[0];
    TMAT_TRY(ws.get("stem.kernel", 9ll * ch * C0, &k));
    TMAT_TRY(ws.get("stem.bias", C0, &b));
    TMAT_TRY(bn_fold(ws, "stem.bn", b, C0, scale, shift));
    std::vector<float> w(9 * C0);
    for (int t = 0; t < 9; ++t)
      for (int c = 0; c < C0; ++c) w[t * C0 + c] = k[(t * ch + 0) * C0 + c] * scale[c];
    TMAT_TRY(upload(net, w, &net->stem_w));
    TMAT_TRY(upload(net, shift, &net->stem_shift));
  }
  int cin = fc[0];
  for (int i = 0; i < 3; ++i) {   // models.py:126-146
    int f = fc[i + 1];
    auto& D = net->down[i];
    D.cin = cin; D.f = f;
    std::string p = "down" + std::to_string(i);
    for (int s = 1; s <= 2; ++s) {
      int ci = (s == 1) ? cin : f;
      std::string sp = p + ".sep" + std::to_string(s);
      const float *dw, *pw;
      TMAT_TRY(ws.get(sp + ".dw", 9ll * ci, &dw));
      TMAT_TRY(ws.get(sp + ".pw", (int64_t)ci * f, &pw));
      TMAT_TRY(ws.get(sp + ".bias", f, &b));
      TMAT_TRY(bn_fold(ws, p + ".bn" + std::to_string(s), b, f, scale, shift));
      std::vector<float> wd(dw, dw + 9 * ci);        // (3,3,ci,1) == [tap][c]
      std::vector<float> wp((size_t)f * ci);
      for (int co = 0; co < f; ++co)
        for (int c = 0; c < ci; ++c) wp[(size_t)co * ci + c] = pw[(size_t)c * f + co] * scale[co];
      TMAT_TRY(upload(net, wd, s == 1 ? &D.dw1 : &D.dw2));
      TMAT_TRY(upload_bf16(net, wp, s == 1 ? &D.pw1 : &D.pw2));
      TMAT_TRY(upload(net, shift, s == 1 ? &D.sh1 : &D.sh2));
    }
    TMAT_TRY(ws.get(p + ".res.kernel", (int64_t)cin * f, &k));
    TMAT_TRY(ws.get(p + ".res.bias", f, &b));
    std::vector<float> wr((size_t)f * cin), sr(b, b + f);
    for (int co = 0; co < f; ++co)
      for (int c = 0; c < cin; ++c) wr[(size_t)co * cin + c] = k[(size_t)c * f + co];
    TMAT_TRY(upload_bf16(net, wr, &D.res));
    TMAT_TRY(upload(net, sr, &D.shres));
    cin = f;
  }
  for (int i = 0; i < 4; ++i) {   // models.py:149-163
    int f = fc[3 - i];
    auto& U = net->up[i];
    U.cin = cin; U.f = f;
    std::string p = "up" + std::to_string(i);
    // ct1: kernel (3,3,f,cin); tap (ky,kx) of the equivalent correlation uses K[2-ky][2-kx][co][c]
    TMAT_TRY(ws.get(p + ".ct1.kernel", 9ll * f * cin, &k));
    TMAT_TRY(ws.get(p + ".ct1.bias", f, &b));
    TMAT_TRY(bn_fold(ws, p + ".bn1", b, f, scale, shift));
    auto wc1 = [&](int ky, int kx, int co, int c) {
      return k[((size_t)((2 - ky) * 3 + (2 - kx)) * f + co) * cin + c] * scale[co];
    };
    if (i == 0) {
      int K = 9 * cin;
      std::vector<float> w((size_t)f * K);
      for (int co = 0; co < f; ++co)
        for (int ky = 0; ky < 3; ++ky)
          for (int kx = 0; kx < 3; ++kx)
            for (int c = 0; c < cin; ++c) w[(size_t)co * K + (ky * 3 + kx) * cin + c] = wc1(ky, kx, co, c);
      U.ngroups = 1; U.gtaps = 9;
      TMAT_TRY(upload_bf16(net, w, &U.ct1));
    } else {
      // input is UpSampling2D(2) of the previous block's low-res output: evaluate per output parity
      ParityPlan pl = parity_plan(f);
      U.ngroups = pl.ngroups; U.gtaps = pl.gtaps;
      int K = pl.gtaps * cin;
      std::vector<float> w((size_t)4 * f * K, 0.f);
      for (int g = 0; g < pl.ngroups; ++g)
        for (int ql = 0; ql < pl.ppg; ++ql) {
          int q = g * pl.ppg + ql, py = q >> 1, px = q & 1;
          for (int t = 0; t < pl.gtaps; ++t)
            for (int ky = 0; ky < 3; ++ky) {
              if (parity_lowtap(py, ky) != pl.dy[g][t]) continue;
              for (int kx = 0; kx < 3; ++kx) {
                if (parity_lowtap(px, kx) != pl.dx[g][t]) continue;
                for (int co = 0; co < f; ++co) {
                  float* dst = &w[((size_t)(g * pl.ppg + ql) * f + co) * K + (size_t)t * cin];
                  for (int c = 0; c < cin; ++c) dst[c] += wc1(ky, kx, co, c);
                }
              }
            }
        }
      TMAT_TRY(upload_bf16(net, w, &U.ct1));
    }
    TMAT_TRY(upload(net, shift, &U.sh1));
    const float *kr, *br;
    TMAT_TRY(ws.get(p + ".ct2.kernel", 9ll * f * f, &k));
    TMAT_TRY(ws.get(p + ".ct2.bias", f, &b));
    TMAT_TRY(bn_fold(ws, p + ".bn2", b, f, scale, shift));
    TMAT_TRY(ws.get(p + ".res.kernel", (int64_t)cin * f, &kr));
    TMAT_TRY(ws.get(p + ".res.bias", f, &br));
    {
      int K = 9 * f + (i == 0 ? cin : 0);
      std::vector<float> w((size_t)f * K);
      for (int co = 0; co < f; ++co) {
        for (int ky = 0; ky < 3; ++ky)
          for (int kx = 0; kx < 3; ++kx)
            for (int c = 0; c < f; ++c)
              w[(size_t)co * K + (ky * 3 + kx) * f + c] =
                  k[((size_t)((2 - ky) * 3 + (2 - kx)) * f + co) * f + c] * scale[co];
        if (i == 0) {
          for (int c = 0; c < cin; ++c) w[(size_t)co * K + 9 * f + c] = kr[(size_t)c * f + co];
          shift[co] += br[co];
        }
      }
      TMAT_TRY(upload_bf16(net, w, &U.ct2));
      TMAT_TRY(upload(net, shift, &U.sh2));
    }
    U.z = nullptr; U.shz = nullptr;
    if (i > 0) {
      std::vector<float> wz((size_t)f * cin), sz(br, br + f);
      for (int co = 0; co < f; ++co)
        for (int c = 0; c < cin; ++c) wz[(size_t)co * cin + c] = kr[(size_t)c * f + co];
      TMAT_TRY(upload_bf16(net, wz, &U.z));
      TMAT_TRY(upload(net, sz, &U.shz));
    }
    cin = f;
  }
  // head: Conv2D(1, 3, same) on the x2-upsampled tensor -> 4 parity kernels on the low-res tensor
  {
    TMAT_TRY(ws.get("head.kernel", 9ll * cin, &k));
    TMAT_TRY(ws.get("head.bias", 1, &b));
    int K = 9 * cin, NP = 16;
    std::vector<float> w((size_t)NP * K, 0.f), sh(NP, 0.f);
    // low-res offset of fine tap kk for output parity par:  floor((par + kk - 1) / 2)
    auto lowtap = [](int par, int kk) { int v = par + kk - 1; return v < 0 ? -1 : v / 2; };
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        int n = py * 2 + px;
        sh[n] = b[0];
        for (int ky = 0; ky < 3; ++ky)
          for (int kx = 0; kx < 3; ++kx) {
            int t = (lowtap(py, ky) + 1) * 3 + (lowtap(px, kx) + 1);
            for (int c = 0; c < cin; ++c) w[(size_t)n * K + t * cin + c] += k[(size_t)(ky * 3 + kx) * cin + c];
          }
      }
    TMAT_TRY(upload_bf16(net, w, &net->head_w));
    TMAT_TRY(upload(net, sh, &net->head_shift));
    // the same sums arranged as 16 (parity, tap) rows of cin weights: tap (a, b) of parity (py, px) is the low-res pixel at
    // offset (a + py - 1, b + px - 1) and collects the fine taps that fall on it
    std::vector<float> w2((size_t)16 * cin, 0.f);
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px)
        for (int ky = 0; ky < 3; ++ky)
          for (int kx = 0; kx < 3; ++kx) {
            const int a = lowtap(py, ky) - (py - 1), bb = lowtap(px, kx) - (px - 1);       // 0 or 1
            const int n = (py * 2 + px) * 4 + a * 2 + bb;
            for (int c = 0; c < cin; ++c) w2[(size_t)n * cin + c] += k[(size_t)(ky * 3 + kx) * cin + c];
          }
    TMAT_TRY(upload_bf16(net, w2, &net->head_w2));
  }
  return TMAT_OK;
}

}  // namespace

// ---- forward -------------------------------------------------------------------------------------
namespace {

const int8_t TAP_DY[9] = {-1, -1, -1, 0, 0, 0, 1, 1, 1};
const int8_t TAP_DX[9] = {-1, 0, 1, -1, 0, 1, -1, 0, 1};

void set_taps3x3(ConvOp& op) {
  op.ngroups = 1; op.ntaps = 9;
  for (int t = 0; t < 9; ++t) { op.dy[0][t] = TAP_DY[t]; op.dx[0][t] = TAP_DX[t]; }
}
void set_tap1x1(ConvOp& op) { op.ngroups = 1; op.ntaps = 1; op.dy[0][0] = 0; op.dx[0][0] = 0; }
void set_taps_parity(ConvOp& op, int f) {
  ParityPlan pl = parity_plan(f);
  op.ngroups = pl.ngroups; op.ntaps = pl.gtaps;
  memcpy(op.dy, pl.dy, sizeof(op.dy));
  memcpy(op.dx, pl.dx, sizeof(op.dx));
  op.Cout = f; op.Npad = pl.ppg * f; op.epi = EPI_PAR;
}

int run_op(tmat_ctx* ctx, tmat_unet* net, const ConvOp& op, cudaStream_t st) {
  if (net->use_tc) return conv_run_tc(ctx, op, st);
  return conv_run_simt(ctx, op, st);
}

void record(tmat_unet* net, const char* name, const void* p, int B, int H, int W, int C, int f32 = 0, int up = 0) {
  if (net->keep_taps) net->taps[name] = Tap{p, B, H, W, C, f32, up};
}

// bytes of activation scratch one forward over B patches takes from the arena
size_t forward_scratch_bytes(const tmat_unet* net, int B) {
  const int* fc = net->fc;
  size_t S = net->desc.patch / 2;
  size_t px = (size_t)B * S * S;
  size_t total = px * fc[0] * 2 + 4096;            // stem out
  int cin = fc[0];
  for (int i = 0; i < 3; ++i) {
    int f = fc[i + 1];
    total += px * cin * 2 + px / 4 * cin * 2;        // dw1 out, subsampled input
    total += px * f * 2 * 3;                         // pw1 out, dw2 out, pw2 out
    total += px / 4 * f * 2 * 3;                     // pooled, block out (P and, for the last block, R)
    total += 8 * 512;
    px /= 4; cin = f;
  }
  for (int i = 0; i < 4; ++i) {
    int f = fc[3 - i];
    if (i > 0) { total += px * f * 2; px *= 4; }     // Z at the previous resolution
    total += px * f * 2 * 3;                         // ct1 out, Y (P and R)
    total += 4 * 512;
  }
  return total + (1 << 16);
}

int unet_forward(tmat_ctx* ctx, tmat_unet* net, const StemSrc& src, int B, float* probs, int logits,
                 cudaStream_t st) {
  const int* fc = net->fc;
  const int ws = net->desc.patch;
  int S = ws / 2;
  if (net->keep_taps) net->taps.clear();
  bf16* X;
  TMAT_TRY(arena_get(ctx, (size_t)B * S * S * fc[0], &X));
  TMAT_TRY(stem_run(ctx, src, B, ws, net->stem_w, net->stem_shift, fc[0], X, st));
  record(net, "stem", X, B, S, S, fc[0]);
  bf16* P = X;        // block input, pre-activation
  bf16* R = nullptr;  // ReLU'd copy (needed from the last down block on)
  for (int i = 0; i < 3; ++i) {
    auto& D = net->down[i];
    size_t px = (size_t)B * S * S;
    bf16 *d1 = nullptr, *a1, *d2 = nullptr, *b2, *pooled, *out, *outr = nullptr;
    const bool fused = net->use_tc && net->fuse_dw && dwpw_supported(B, S, D.cin, D.f) && dwpw_supported(B, S, D.f, D.f);
    if (!fused) {
      TMAT_TRY(arena_get(ctx, px * D.cin, &d1));
      TMAT_TRY(arena_get(ctx, px * D.f, &d2));
    }
    // MaxPool folded into the second SeparableConv2D's epilogue: its full-resolution output is never stored
    static const bool no_fuse_pool = getenv("TMAT_NO_FUSE_POOL") != nullptr;
    const bool fuse_pool = fused && !no_fuse_pool && !net->keep_taps && dwpw_pool_supported(B, S, D.f, D.f);
    TMAT_TRY(arena_get(ctx, px * D.f, &a1));
    if (!fuse_pool) TMAT_TRY(arena_get(ctx, px * D.f, &b2));
    else b2 = nullptr;
    TMAT_TRY(arena_get(ctx, px / 4 * D.f, &pooled));
    TMAT_TRY(arena_get(ctx, px / 4 * D.f, &out));
    if (i == 2) TMAT_TRY(arena_get(ctx, px / 4 * D.f, &outr));
    if (fused) {
      // SeparableConv2D as one kernel each: [ReLU] -> depthwise -> pointwise + BN [+ ReLU]
      TMAT_TRY(dwpw_run(ctx, P, B, S, D.cin, D.dw1, i != 0, D.pw1, D.sh1, D.f, 1, a1, st));
      TMAT_TRY(dwpw_run(ctx, a1, B, S, D.f, D.dw2, 0, D.pw2, D.sh2, D.f, 0, b2, st, fuse_pool ? pooled : nullptr));
    } else {
      TMAT_TRY(dwconv_run(ctx, P, B, S, D.cin, D.dw1, i != 0, d1, nullptr, st));
      ConvOp op;
      op.name = "down.pw1"; op.a0 = d1; op.a0_C = D.cin; op.B = B; op.H = S; op.W = S; set_tap1x1(op);
      op.wmat = D.pw1; op.K = D.cin; op.shift = D.sh1; op.Cout = op.Npad = D.f; op.epi = EPI_STORE; op.relu = 1;
      op.out = a1;
      TMAT_TRY(run_op(ctx, net, op, st));
      TMAT_TRY(dwconv_run(ctx, a1, B, S, D.f, D.dw2, 0, d2, nullptr, st));
      ConvOp op2;
      op2.name = "down.pw2"; op2.a0 = d2; op2.a0_C = D.f; op2.B = B; op2.H = S; op2.W = S; set_tap1x1(op2);
      op2.wmat = D.pw2; op2.K = D.f; op2.shift = D.sh2; op2.Cout = op2.Npad = D.f; op2.epi = EPI_STORE; op2.relu = 0;
      op2.out = b2;
      TMAT_TRY(run_op(ctx, net, op2, st));
    }
    record(net, (std::string("down") + std::to_string(i) + ".r1").c_str(), a1, B, S, S, D.f);
    if (b2) record(net, (std::string("down") + std::to_string(i) + ".b2").c_str(), b2, B, S, S, D.f);
    // MaxPool(3, 2, same)(b2) + residual 1x1 stride-2 conv of the block input
    if (!fuse_pool) TMAT_TRY(maxpool_run(ctx, b2, B, S, D.f, pooled, st));
    ConvOp op3;
    op3.name = "down.res"; op3.a0 = P; op3.a0_C = D.cin; op3.a0_stride = 2; op3.B = B; op3.H = S / 2; op3.W = S / 2;
    set_tap1x1(op3);
    op3.wmat = D.res; op3.K = D.cin; op3.shift = D.shres; op3.Cout = op3.Npad = D.f; op3.epi = EPI_STORE;
    op3.relu = 0; op3.add_src = pooled; op3.add_shift = 0; op3.out = out; op3.out2 = outr;
    TMAT_TRY(run_op(ctx, net, op3, st));
    S /= 2;
    record(net, (std::string("down") + std::to_string(i)).c_str(), out, B, S, S, D.f);
    P = out; R = outr;
  }
  // up blocks: block i runs at resolution S and leaves its LOW-RES output Y_i = BN(ct2) + residual
  // (pre-activation P and ReLU'd R); the reference's block output is UpSampling2D(2)(Y_i).
  for (int i = 0; i < 4; ++i) {
    auto& U = net->up[i];
    bf16* Z = nullptr;
    if (i > 0) {
      // residual of this block at the previous block's resolution: Z = conv1x1(Y_{i-1}) + bias
      TMAT_TRY(arena_get(ctx, (size_t)B * S * S * U.f, &Z));
      ConvOp oz;
      oz.name = "up.res"; oz.a0 = P; oz.a0_C = U.cin; oz.B = B; oz.H = S; oz.W = S; set_tap1x1(oz);
      oz.wmat = U.z; oz.K = U.cin; oz.shift = U.shz; oz.Cout = oz.Npad = U.f; oz.epi = EPI_STORE; oz.relu = 0;
      oz.out = Z;
      TMAT_TRY(run_op(ctx, net, oz, st));
    }
    const int Sin = S;              // resolution of the block's stored input (R, P)
    if (i > 0) S *= 2;
    size_t px = (size_t)B * S * S;
    bf16 *a1, *outp, *outr = nullptr;
    TMAT_TRY(arena_get(ctx, px * U.f, &a1));
    ConvOp op;
    op.name = "up.ct1"; op.a0 = R; op.a0_C = U.cin; op.B = B; op.H = Sin; op.W = Sin;
    op.wmat = U.ct1; op.shift = U.sh1; op.relu = 1; op.out = a1;
    if (i == 0) {
      set_taps3x3(op);
      op.K = 9 * U.cin; op.Cout = op.Npad = U.f; op.epi = EPI_STORE;
    } else {
      set_taps_parity(op, U.f);
      op.K = op.ntaps * U.cin;
    }
    TMAT_TRY(run_op(ctx, net, op, st));
    record(net, (std::string("up") + std::to_string(i) + ".r1").c_str(), a1, B, S, S, U.f);
    TMAT_TRY(arena_get(ctx, px * U.f, &outp));
    if (i < 3) TMAT_TRY(arena_get(ctx, px * U.f, &outr));
    ConvOp op2;
    op2.name = "up.ct2"; op2.a0 = a1; op2.a0_C = U.f; op2.B = B; op2.H = S; op2.W = S;
    set_taps3x3(op2);
    op2.wmat = U.ct2; op2.shift = U.sh2; op2.Cout = op2.Npad = U.f; op2.relu = 0; op2.epi = EPI_STORE;
    op2.out = outp; op2.out2 = outr;
    if (i == 0) { op2.a1 = P; op2.a1_C = U.cin; op2.K = 9 * U.f + U.cin; }
    else { op2.K = 9 * U.f; op2.add_src = Z; op2.add_shift = 1; }
    TMAT_TRY(run_op(ctx, net, op2, st));
    record(net, (std::string("up") + std::to_string(i)).c_str(), outp, B, S, S, U.f, 0, 1);
    P = outp; R = outr;
  }
  ConvOp oh;
  oh.name = "head"; oh.a0 = P; oh.a0_C = fc[0]; oh.B = B; oh.H = S; oh.W = S; set_taps3x3(oh);
  oh.wmat = net->head_w; oh.wmat2 = net->head_w2; oh.K = 9 * fc[0]; oh.shift = net->head_shift; oh.Cout = 4; oh.Npad = 16; oh.epi = EPI_HEAD;
  oh.out_f32 = probs; oh.logits = logits;
  TMAT_TRY(run_op(ctx, net, oh, st));
  record(net, "probs", probs, B, ws, ws, 1, 1);
  return TMAT_OK;
}

}  // namespace

// ---- C ABI -----------------------------------------------------------------------------------------
static double flops_per_patch(const tmat_unet_desc* d) {
  const int* fc = d->filter_counts;
  double s = d->patch / 2;
  double mac = s * s * 9 * d->channels * fc[0];
  int cin = fc[0];
  for (int i = 1; i < 4; ++i) {
    int f = fc[i];
    mac += s * s * (9.0 * cin + (double)cin * f) + s * s * (9.0 * f + (double)f * f);
    s /= 2;
    mac += s * s * cin * f;
    cin = f;
  }
  for (int i = 3; i >= 0; --i) {
    int f = fc[i];
    mac += s * s * 9 * ((double)cin * f + (double)f * f);
    s *= 2;
    mac += s * s * cin * f;
    cin = f;
  }
  mac += s * s * 9 * cin;
  return 2 * mac;
}

extern "C" int tmat_unet_flops_per_patch(const tmat_unet_desc* desc, double* flops) {
  TMAT_REQUIRE(desc && flops, "tmat_unet_flops_per_patch: NULL argument");
  *flops = flops_per_patch(desc);
  return TMAT_OK;
}

extern "C" int tmat_unet_load(tmat_ctx* ctx, const tmat_unet_desc* desc, int n_arrays, const char* const* names,
                              const float* const* arrays_host, const int64_t* sizes, tmat_unet** out) {
  TMAT_REQUIRE(ctx && desc && names && arrays_host && sizes && out, "tmat_unet_load: NULL argument");
  TMAT_REQUIRE(desc->patch >= 32 && desc->patch % 32 == 0, "tmat_unet_load: patch size %d must be a multiple of 32",
               desc->patch);
  if (desc->channels != 1) {
    tmat_set_error("tmat_unet_load: channels=%d; only the reference's single-channel configuration is implemented",
                   desc->channels);
    return TMAT_ERR_UNSUPPORTED;
  }
  bool mult64 = true;
  for (int i = 0; i < 4; ++i) {
    int f = desc->filter_counts[i];
    TMAT_REQUIRE(f >= 8 && f % 8 == 0, "tmat_unet_load: filter count %d must be a positive multiple of 8", f);
    if (i) TMAT_REQUIRE(f >= desc->filter_counts[i - 1], "tmat_unet_load: filter counts must be ascending");
    if (f % 64 != 0 || f > 512) mult64 = false;
  }
  if (desc->impl == 2 && !mult64) {
    tmat_set_error("tmat_unet_load: the tcgen05 path needs filter counts that are multiples of 64 (<= 512)");
    return TMAT_ERR_UNSUPPORTED;
  }
  CUDA_TRY(cudaSetDevice(ctx->device));
  tmat_unet* net = new tmat_unet();
  net->desc = *desc;
  for (int i = 0; i < 4; ++i) net->fc[i] = desc->filter_counts[i];
  net->device = ctx->device;
  net->use_tc = (desc->impl == 2) || (desc->impl == 0 && mult64);
  { const char* e = getenv("TMAT_NO_FUSE_DW"); net->fuse_dw = !(e && atoi(e) != 0); }
  WeightSet ws;
  for (int i = 0; i < n_arrays; ++i) ws.m[names[i]] = std::make_pair(arrays_host[i], sizes[i]);
  int r = pack_weights(net, ws);
  if (r != TMAT_OK) {
    tmat_unet_free(net);
    return r;
  }
  *out = net;
  return TMAT_OK;
}

extern "C" int tmat_unet_free(tmat_unet* net) {
  if (!net) return TMAT_OK;
  cudaSetDevice(net->device);
  cudaDeviceSynchronize();
  for (void* p : net->allocs) cudaFree(p);
  delete net;
  return TMAT_OK;
}

// Patches per chunk for n patches: at most ~704 patches of 320x320 (~40 MB of bf16 activations each, 27 GB per chunk),
// split evenly so that the last chunk is not a short one (1352 patches -> 2 x 676).  Large chunks keep the persistent
// GEMM kernels at >= 20 tiles per SM even at the 20x20 bottleneck and halve the launches and their tails (measured per
// 2048^2 image: 64 -> 65.8, 128 -> 61.9, 256 -> 60.4 ms on the first path; 256 -> 43.0, 352 -> 42.3, 512 -> 42.5 ms at
// the end of round 1; with the taps-in-N head, A/B pairs on one box: 4 x 338 -> 41.9 / 42.3, 3 x 451 -> 42.1,
// 2 x 676 -> 41.6 / 41.3, 1 x 1352 -> 42.0 / 42.2 ms, identical outputs).  All tensor indexing is 64-bit: a chunk's
// largest tensor (676 x 160 x 160 x 128) has 2.2e9 elements.
static int chunk_size(const tmat_unet* net, int n) {
  const char* e = getenv("TMAT_CHUNK");
  if (e && atoi(e) > 0) return std::min(n, atoi(e));
  int64_t px = (int64_t)net->desc.patch * net->desc.patch;
  int cap = (int)(704 * 102400 / px);
  cap = cap < 8 ? 8 : (cap > 2048 ? 2048 : cap);
  const int n_chunks = (n + cap - 1) / cap;
  return (n + n_chunks - 1) / n_chunks;
}

extern "C" int tmat_unet_forward_dev(tmat_ctx* ctx, tmat_unet* net, const float* patches_dev, int n,
                                     float* probs_dev, int logits, void* stream) {
  TMAT_REQUIRE(ctx && net && patches_dev && probs_dev, "tmat_unet_forward: NULL argument");
  TMAT_REQUIRE(n >= 1, "tmat_unet_forward: n=%d", n);
  TMAT_TRY(arena_reset(ctx));
  const int ws = net->desc.patch;
  const int CH = chunk_size(net, n);
  TMAT_TRY(arena_reserve(ctx, forward_scratch_bytes(net, CH)));
  net->keep_taps = (n <= CH);
  for (int p0 = 0; p0 < n; p0 += CH) {
    int B = std::min(CH, n - p0);
    size_t mark = arena_mark(ctx);
    StemSrc src;
    memset(&src, 0, sizeof(src));
    src.patches = patches_dev + (size_t)p0 * ws * ws;
    src.use_geom = 0;
    TMAT_TRY(unet_forward(ctx, net, src, B, probs_dev + (size_t)p0 * ws * ws, logits, as_stream(stream)));
    arena_release(ctx, mark);
  }
  return TMAT_OK;
}

extern "C" int tmat_unet_forward_host(tmat_ctx* ctx, tmat_unet* net, const float* patches_host, int n,
                                      float* probs_host, int logits) {
  TMAT_REQUIRE(ctx && net && patches_host && probs_host, "tmat_unet_forward_host: NULL argument");
  TMAT_REQUIRE(n >= 1, "tmat_unet_forward_host: n=%d", n);
  CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t px = (size_t)n * net->desc.patch * net->desc.patch;
  float *d_in, *d_out;
  CUDA_TRY(cudaMalloc((void**)&d_in, px * 4));
  CUDA_TRY(cudaMalloc((void**)&d_out, px * 4));
  int r = TMAT_OK;
  if (cudaMemcpyAsync(d_in, patches_host, px * 4, cudaMemcpyHostToDevice, 0) != cudaSuccess) r = TMAT_ERR_CUDA;
  if (r == TMAT_OK) r = tmat_unet_forward_dev(ctx, net, d_in, n, d_out, logits, nullptr);
  if (r == TMAT_OK && cudaMemcpyAsync(probs_host, d_out, px * 4, cudaMemcpyDeviceToHost, 0) != cudaSuccess) r = TMAT_ERR_CUDA;
  cudaError_t e = cudaStreamSynchronize(0);
  if (r == TMAT_OK && e != cudaSuccess) {
    tmat_set_error("tmat_unet_forward_host: %s", cudaGetErrorString(e));
    r = TMAT_ERR_CUDA;
  }
  cudaFree(d_in);
  cudaFree(d_out);
  return r;
}

// scratch bytes stp_predict_impl takes from the arena (patch probabilities + one chunk of activations)
size_t stp_predict_scratch_bytes(const tmat_unet* net, int H, int W, int subdiv) {
  StpGeom G;
  if (stp_make_geom(H, W, net->desc.patch, subdiv, &G) != TMAT_OK) return 0;
  const int n_total = 8 * G.tiles_per_orient();
  const int ws = net->desc.patch;
  return (size_t)n_total * ws * ws * 4 + forward_scratch_bytes(net, chunk_size(net, n_total)) + (4 << 20);
}

// predict_img_with_smooth_windowing with pred_func = this UNet; arena must have been reset by the caller
int stp_predict_impl(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W, int subdiv,
                     double* out_dev, cudaStream_t st) {
  StpGeom G;
  const int ws = net->desc.patch;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  const int n_total = 8 * G.tiles_per_orient();
  const int CH = chunk_size(net, n_total);
  size_t probs_bytes = (size_t)n_total * ws * ws * 4;
  TMAT_TRY(arena_reserve(ctx, probs_bytes + forward_scratch_bytes(net, CH) + (1 << 20)));
  float* minmax;
  float* probs;
  double* w1d;
  TMAT_TRY(arena_get(ctx, 2, &minmax));
  TMAT_TRY(arena_get(ctx, (size_t)n_total * ws * ws, &probs));
  TMAT_TRY(stp_upload_window(ctx, ws, nullptr, &w1d, st));
  TMAT_TRY(stp_image_min(ctx, img_dev, (int64_t)H * W, minmax, st));
  net->keep_taps = false;
  for (int p0 = 0; p0 < n_total; p0 += CH) {
    int B = std::min(CH, n_total - p0);
    size_t mark = arena_mark(ctx);
    StemSrc src;
    memset(&src, 0, sizeof(src));
    src.img = img_dev; src.minmax = minmax; src.geom = G; src.first_patch = p0; src.use_geom = 1;
    TMAT_TRY(unet_forward(ctx, net, src, B, probs + (size_t)p0 * ws * ws, 0, st));
    arena_release(ctx, mark);
  }
  return stp_blend(ctx, G, probs, w1d, out_dev, st);
}

// Output rows [row0, row1) of predict_img_with_smooth_windowing only: the UNet runs on the patches whose window
// touches those rows (owner-computes row strips: sharding ONE large image over GPUs with no reduction)
int stp_predict_rows_impl(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W, int subdiv, int row0,
                          int row1, double* out_rows_dev, int* n_patches, cudaStream_t st) {
  StpGeom G;
  const int ws = net->desc.patch;
  TMAT_TRY(stp_make_geom(H, W, ws, subdiv, &G));
  TMAT_REQUIRE(row0 >= 0 && row0 < row1 && row1 <= H, "tmat_stp_predict_rows: bad row range [%d, %d) for H=%d", row0,
               row1, H);
  std::vector<int> list;
  stp_patches_for_rows(G, row0, row1, &list);
  const int n_total = 8 * G.tiles_per_orient();
  const int n_sel = (int)list.size();
  if (n_patches) *n_patches = n_sel;
  std::vector<int> slot(n_total, -1);
  for (int i = 0; i < n_sel; ++i) slot[list[i]] = i;
  const int CH = chunk_size(net, n_sel);
  TMAT_TRY(arena_reserve(ctx, (size_t)n_sel * ws * ws * 4 + forward_scratch_bytes(net, CH) + (size_t)n_total * 8 + (4 << 20)));
  float *minmax, *probs;
  double* w1d;
  int *d_list, *d_slot;
  TMAT_TRY(arena_get(ctx, 2, &minmax));
  TMAT_TRY(arena_get(ctx, (size_t)n_sel * ws * ws, &probs));
  TMAT_TRY(arena_get(ctx, (size_t)n_sel, &d_list));
  TMAT_TRY(arena_get(ctx, (size_t)n_total, &d_slot));
  CUDA_TRY(cudaMemcpyAsync(d_list, list.data(), (size_t)n_sel * 4, cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemcpyAsync(d_slot, slot.data(), (size_t)n_total * 4, cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaStreamSynchronize(st));        // the host vectors die with this frame
  TMAT_TRY(stp_upload_window(ctx, ws, nullptr, &w1d, st));
  TMAT_TRY(stp_image_min(ctx, img_dev, (int64_t)H * W, minmax, st));
  net->keep_taps = false;
  for (int p0 = 0; p0 < n_sel; p0 += CH) {
    int B = std::min(CH, n_sel - p0);
    size_t mark = arena_mark(ctx);
    StemSrc src;
    memset(&src, 0, sizeof(src));
    src.img = img_dev; src.minmax = minmax; src.geom = G; src.first_patch = p0; src.use_geom = 1;
    src.patch_list = d_list;
    TMAT_TRY(unet_forward(ctx, net, src, B, probs + (size_t)p0 * ws * ws, 0, st));
    arena_release(ctx, mark);
  }
  return stp_blend_rows(ctx, G, probs, d_slot, w1d, row0, row1, out_rows_dev, st);
}

extern "C" int tmat_stp_predict_rows_dev(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W, int subdiv,
                                         int row0, int row1, double* out_rows_dev, int* n_patches, void* stream) {
  TMAT_REQUIRE(ctx && net && img_dev && out_rows_dev, "tmat_stp_predict_rows: NULL argument");
  TMAT_TRY(arena_reset(ctx));
  return stp_predict_rows_impl(ctx, net, img_dev, H, W, subdiv, row0, row1, out_rows_dev, n_patches, as_stream(stream));
}

extern "C" int tmat_stp_predict_rows_host(tmat_ctx* ctx, tmat_unet* net, const float* img_host, int H, int W,
                                          int subdiv, int row0, int row1, double* out_rows_host, int* n_patches) {
  TMAT_REQUIRE(ctx && net && img_host && out_rows_host, "tmat_stp_predict_rows_host: NULL argument");
  TMAT_REQUIRE(H >= 1 && W >= 1 && row0 >= 0 && row0 < row1 && row1 <= H, "tmat_stp_predict_rows_host: bad shape / rows");
  TMAT_TRY(arena_reset(ctx));
  float* d_img;
  double* d_out;
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_img));
  TMAT_TRY(arena_get(ctx, (size_t)(row1 - row0) * W, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_img, img_host, (size_t)H * W * 4, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(stp_predict_rows_impl(ctx, net, d_img, H, W, subdiv, row0, row1, d_out, n_patches, 0));
  CUDA_TRY(cudaMemcpyAsync(out_rows_host, d_out, (size_t)(row1 - row0) * W * 8, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

extern "C" int tmat_stp_predict_dev(tmat_ctx* ctx, tmat_unet* net, const float* img_dev, int H, int W, int subdiv,
                                    double* out_dev, void* stream) {
  TMAT_REQUIRE(ctx && net && img_dev && out_dev, "tmat_stp_predict: NULL argument");
  TMAT_TRY(arena_reset(ctx));
  return stp_predict_impl(ctx, net, img_dev, H, W, subdiv, out_dev, as_stream(stream));
}

extern "C" int tmat_stp_predict_host(tmat_ctx* ctx, tmat_unet* net, const float* img_host, int H, int W,
                                     int subdiv, double* out_host) {
  TMAT_REQUIRE(ctx && net && img_host && out_host, "tmat_stp_predict_host: NULL argument");
  TMAT_REQUIRE(H >= 1 && W >= 1, "tmat_stp_predict_host: bad shape");
  TMAT_TRY(arena_reset(ctx));
  StpGeom G;
  TMAT_TRY(stp_make_geom(H, W, net->desc.patch, subdiv, &G));
  const int n_total = 8 * G.tiles_per_orient();
  const int ws = net->desc.patch;
  TMAT_TRY(arena_reserve(ctx, (size_t)n_total * ws * ws * 4 + forward_scratch_bytes(net, chunk_size(net, n_total)) +
                                  (size_t)H * W * 12 + (4 << 20)));
  float* d_img;
  double* d_out;
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_img));
  TMAT_TRY(arena_get(ctx, (size_t)H * W, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_img, img_host, (size_t)H * W * 4, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(stp_predict_impl(ctx, net, d_img, H, W, subdiv, d_out, 0));
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, (size_t)H * W * 8, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

namespace {
// bf16 (B,H,W,C) -> float32, optionally 2x nearest-upsampled to (B,2H,2W,C)
__global__ void tap_to_f32(const bf16* in, int64_t n, int H, int W, int C, int up, float* out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!up) { out[i] = __bfloat162float(in[i]); return; }
  int c = (int)(i % C);
  int64_t pix = i / C;
  int x = (int)(pix % (2 * W));
  int y = (int)((pix / (2 * W)) % (2 * H));
  int64_t b = pix / ((int64_t)4 * W * H);
  out[i] = __bfloat162float(in[((b * H + (y >> 1)) * W + (x >> 1)) * C + c]);
}
}  // namespace

extern "C" int tmat_unet_debug_tap(tmat_ctx* ctx, tmat_unet* net, const char* name, float* out_host,
                                   int64_t capacity, int64_t* n_written) {
  TMAT_REQUIRE(ctx && net && name && n_written, "tmat_unet_debug_tap: NULL argument");
  auto it = net->taps.find(name);
  TMAT_REQUIRE(it != net->taps.end(), "tmat_unet_debug_tap: no tap '%s' (taps are kept for single-chunk forwards)", name);
  const Tap& t = it->second;
  int64_t n = (int64_t)t.B * t.H * t.W * t.C * (t.up ? 4 : 1);
  *n_written = n;
  if (!out_host) return TMAT_OK;
  TMAT_REQUIRE(capacity >= n, "tmat_unet_debug_tap: capacity %lld < %lld", (long long)capacity, (long long)n);
  CUDA_TRY(cudaSetDevice(ctx->device));
  CUDA_TRY(cudaDeviceSynchronize());
  if (t.is_f32) {
    CUDA_TRY(cudaMemcpy(out_host, t.ptr, n * 4, cudaMemcpyDeviceToHost));
    return TMAT_OK;
  }
  float* tmp;
  CUDA_TRY(cudaMalloc((void**)&tmp, n * 4));
  tap_to_f32<<<(unsigned)cdiv64(n, 256), 256>>>((const bf16*)t.ptr, n, t.H, t.W, t.C, t.up, tmp);
  cudaError_t e = cudaMemcpy(out_host, tmp, n * 4, cudaMemcpyDeviceToHost);
  cudaFree(tmp);
  CUDA_TRY(e);
  return TMAT_OK;
}

// ---- self-test hook: one ConvOp on random data through both executors ------------------------------
namespace {
__global__ void fill_random_bf16(bf16* p, int64_t n, uint64_t seed, float scale) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t z = (uint64_t)i * 0x9E3779B97F4A7C15ull + seed;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  float u = (float)(z >> 40) * (1.f / 16777216.f) * 2.f - 1.f;
  p[i] = __float2bfloat16_rn(u * scale);
}
__global__ void fill_random_f32(float* p, int64_t n, uint64_t seed, float scale) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t z = (uint64_t)i * 0x9E3779B97F4A7C15ull + seed;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  p[i] = ((float)(z >> 40) * (1.f / 16777216.f) * 2.f - 1.f) * scale;
}
template <typename T>
__global__ void max_abs_diff_kernel(const T* a, const T* b, int64_t n, float* out /*[2]: max|a-b|, max|a|*/) {
  float d = 0.f, m = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x = (float)a[i], y = (float)b[i];
    float dd = fabsf(x - y);
    if (!(dd <= 3.0e38f)) dd = 3.0e38f;   // NaN / inf -> huge
    d = fmaxf(d, dd);
    m = fmaxf(m, fabsf(x));
  }
  for (int o = 16; o > 0; o >>= 1) {
    d = fmaxf(d, __shfl_xor_sync(0xffffffffu, d, o));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(reinterpret_cast<int*>(out), __float_as_int(d));
    atomicMax(reinterpret_cast<int*>(out + 1), __float_as_int(m));
  }
}
}  // namespace

// epi: 0 = EPI_STORE, 1 = EPI_PAR (parity groups chosen from Cout as in the UNet), 3 = EPI_HEAD.
// addend: 0 none, 1 same-resolution addend, 2 half-resolution addend (EPI_STORE only).
extern "C" int tmat_debug_conv_selftest(tmat_ctx* ctx, int B, int H, int W, int C0, int C1, int Cout, int ntaps,
                                        int epi, int relu, int dual, int addend, uint64_t seed,
                                        double* max_abs_diff, double* max_abs_ref) {
  TMAT_REQUIRE(ctx && max_abs_diff && max_abs_ref, "tmat_debug_conv_selftest: NULL argument");
  TMAT_REQUIRE(ntaps == 1 || ntaps == 9, "tmat_debug_conv_selftest: ntaps must be 1 or 9");
  TMAT_REQUIRE(epi == EPI_STORE || epi == EPI_PAR || epi == EPI_HEAD, "tmat_debug_conv_selftest: bad epi %d", epi);
  TMAT_TRY(arena_reset(ctx));
  ConvOp op;
  op.name = "selftest";
  op.B = B; op.H = H; op.W = W; op.a0_C = C0; op.a1_C = C1;
  op.epi = epi; op.relu = relu;
  if (epi == EPI_PAR) {
    TMAT_REQUIRE(C1 == 0 && !dual && !addend, "tmat_debug_conv_selftest: EPI_PAR takes no a1 / out2 / addend");
    set_taps_parity(op, Cout);
    op.K = op.ntaps * C0;
  } else {
    if (ntaps == 9) set_taps3x3(op); else set_tap1x1(op);
    op.K = ntaps * C0 + C1;
    op.Cout = (epi == EPI_HEAD) ? 4 : Cout;
    op.Npad = (epi == EPI_HEAD) ? 16 : Cout;
  }
  const int rows = op.ngroups * op.Npad;
  size_t px = (size_t)B * H * W;
  bf16 *a0, *a1 = nullptr, *w, *add = nullptr;
  float* shift;
  TMAT_TRY(arena_get(ctx, px * C0, &a0));
  if (C1) TMAT_TRY(arena_get(ctx, px * C1, &a1));
  TMAT_TRY(arena_get(ctx, (size_t)rows * op.K, &w));
  TMAT_TRY(arena_get(ctx, (size_t)rows, &shift));
  auto fill = [&](bf16* p, size_t n, uint64_t s, float sc) {
    fill_random_bf16<<<(unsigned)cdiv64(n, 256), 256>>>(p, (int64_t)n, s, sc);
  };
  fill(a0, px * C0, seed + 1, 1.f);
  if (C1) fill(a1, px * C1, seed + 2, 1.f);
  fill(w, (size_t)rows * op.K, seed + 3, 1.f / sqrtf((float)op.K));
  fill_random_f32<<<cdiv(rows, 256), 256>>>(shift, rows, seed + 4, 0.5f);
  size_t out_elems = px * op.Cout * (epi == EPI_PAR ? 4 : 1);
  if (addend) {
    int s = addend - 1;
    size_t n = (size_t)B * ((H + s) >> s) * ((W + s) >> s) * Cout;
    TMAT_TRY(arena_get(ctx, n, &add));
    fill(add, n, seed + 5, 1.f);
    op.add_src = add; op.add_shift = s;
  }
  op.a0 = a0; op.a1 = a1; op.wmat = w; op.shift = shift;
  float* stats;
  TMAT_TRY(arena_get(ctx, 2, &stats));
  CUDA_TRY(cudaMemsetAsync(stats, 0, 8, 0));
  if (epi == EPI_HEAD) {
    float *o1, *o2;
    size_t n = px * 4;
    TMAT_TRY(arena_get(ctx, n, &o1));
    TMAT_TRY(arena_get(ctx, n, &o2));
    CUDA_TRY(cudaMemsetAsync(o1, 0xff, n * 4, 0));
    CUDA_TRY(cudaMemsetAsync(o2, 0x7f, n * 4, 0));
    op.out_f32 = o1; op.logits = relu;   // 'relu' flag doubles as the logits switch for the head
    TMAT_TRY(conv_run_simt(ctx, op, 0));
    op.out_f32 = o2;
    TMAT_TRY(conv_run_tc(ctx, op, 0));
    max_abs_diff_kernel<float><<<256, 256>>>(o1, o2, (int64_t)n, stats);
  } else {
    bf16 *o1, *o2, *r1 = nullptr, *r2 = nullptr;
    TMAT_TRY(arena_get(ctx, out_elems, &o1));
    TMAT_TRY(arena_get(ctx, out_elems, &o2));
    CUDA_TRY(cudaMemsetAsync(o1, 0xff, out_elems * 2, 0));
    CUDA_TRY(cudaMemsetAsync(o2, 0x7f, out_elems * 2, 0));
    if (dual) {
      TMAT_TRY(arena_get(ctx, out_elems, &r1));
      TMAT_TRY(arena_get(ctx, out_elems, &r2));
      CUDA_TRY(cudaMemsetAsync(r1, 0xff, out_elems * 2, 0));
      CUDA_TRY(cudaMemsetAsync(r2, 0x7f, out_elems * 2, 0));
    }
    op.out = o1; op.out2 = r1;
    TMAT_TRY(conv_run_simt(ctx, op, 0));
    op.out = o2; op.out2 = r2;
    TMAT_TRY(conv_run_tc(ctx, op, 0));
    max_abs_diff_kernel<bf16><<<256, 256>>>(o1, o2, (int64_t)out_elems, stats);
    if (dual) max_abs_diff_kernel<bf16><<<256, 256>>>(r1, r2, (int64_t)out_elems, stats);
  }
  float h[2];
  CUDA_TRY(cudaMemcpy(h, stats, 8, cudaMemcpyDeviceToHost));
  *max_abs_diff = h[0];
  *max_abs_ref = h[1];
  return TMAT_OK;
}

// Fused SeparableConv2D with the pooled epilogue vs the same kernel followed by maxpool_kernel: bit-identical.
extern "C" int tmat_debug_dwpw_pool_selftest(tmat_ctx* ctx, int B, int S, int C, int N, int relu_in, int relu_out,
                                             uint64_t seed, double* max_abs_diff, double* max_abs_ref) {
  TMAT_REQUIRE(ctx && max_abs_diff && max_abs_ref, "tmat_debug_dwpw_pool_selftest: NULL argument");
  TMAT_REQUIRE(dwpw_pool_supported(B, S, C, N), "tmat_debug_dwpw_pool_selftest: shape not supported by the pooled epilogue");
  TMAT_TRY(arena_reset(ctx));
  size_t px = (size_t)B * S * S;
  bf16 *in, *w, *full, *p1, *p2;
  float *dw, *shift, *stats;
  TMAT_TRY(arena_get(ctx, px * C, &in));
  TMAT_TRY(arena_get(ctx, (size_t)N * C, &w));
  TMAT_TRY(arena_get(ctx, px * N, &full));
  TMAT_TRY(arena_get(ctx, px / 4 * N, &p1));
  TMAT_TRY(arena_get(ctx, px / 4 * N, &p2));
  TMAT_TRY(arena_get(ctx, (size_t)9 * C, &dw));
  TMAT_TRY(arena_get(ctx, (size_t)N, &shift));
  TMAT_TRY(arena_get(ctx, 2, &stats));
  fill_random_bf16<<<(unsigned)cdiv64(px * C, 256), 256>>>(in, (int64_t)(px * C), seed + 1, 1.f);
  fill_random_bf16<<<(unsigned)cdiv64((int64_t)N * C, 256), 256>>>(w, (int64_t)N * C, seed + 2, 1.f / sqrtf((float)C));
  fill_random_f32<<<cdiv(9 * C, 256), 256>>>(dw, 9 * C, seed + 3, 0.5f);
  fill_random_f32<<<cdiv(N, 256), 256>>>(shift, N, seed + 4, 0.5f);
  CUDA_TRY(cudaMemsetAsync(stats, 0, 8, 0));
  CUDA_TRY(cudaMemsetAsync(p1, 0xff, px / 4 * N * 2, 0));
  CUDA_TRY(cudaMemsetAsync(p2, 0x7f, px / 4 * N * 2, 0));
  TMAT_TRY(dwpw_run(ctx, in, B, S, C, dw, relu_in, w, shift, N, relu_out, full, 0));
  TMAT_TRY(maxpool_run(ctx, full, B, S, N, p1, 0));
  TMAT_TRY(dwpw_run(ctx, in, B, S, C, dw, relu_in, w, shift, N, relu_out, nullptr, 0, p2));
  max_abs_diff_kernel<bf16><<<256, 256>>>(p1, p2, (int64_t)(px / 4 * N), stats);
  float h[2];
  CUDA_TRY(cudaMemcpy(h, stats, 8, cudaMemcpyDeviceToHost));
  *max_abs_diff = h[0];
  *max_abs_ref = h[1];
  return TMAT_OK;
}

// Fused depthwise+pointwise kernel vs the unfused pair (dwconv_kernel + CUDA-core GEMM) on random data.
extern "C" int tmat_debug_dwpw_selftest(tmat_ctx* ctx, int B, int S, int C, int N, int relu_in, int relu_out,
                                        uint64_t seed, double* max_abs_diff, double* max_abs_ref) {
  TMAT_REQUIRE(ctx && max_abs_diff && max_abs_ref, "tmat_debug_dwpw_selftest: NULL argument");
  TMAT_REQUIRE(dwpw_supported(B, S, C, N), "tmat_debug_dwpw_selftest: shape not supported by the fused kernel");
  TMAT_TRY(arena_reset(ctx));
  size_t px = (size_t)B * S * S;
  bf16 *in, *d1, *w, *o1, *o2;
  float *dw, *shift, *stats;
  TMAT_TRY(arena_get(ctx, px * C, &in));
  TMAT_TRY(arena_get(ctx, px * C, &d1));
  TMAT_TRY(arena_get(ctx, (size_t)N * C, &w));
  TMAT_TRY(arena_get(ctx, px * N, &o1));
  TMAT_TRY(arena_get(ctx, px * N, &o2));
  TMAT_TRY(arena_get(ctx, (size_t)9 * C, &dw));
  TMAT_TRY(arena_get(ctx, (size_t)N, &shift));
  TMAT_TRY(arena_get(ctx, 2, &stats));
  fill_random_bf16<<<(unsigned)cdiv64(px * C, 256), 256>>>(in, (int64_t)(px * C), seed + 1, 1.f);
  fill_random_bf16<<<(unsigned)cdiv64((int64_t)N * C, 256), 256>>>(w, (int64_t)N * C, seed + 2, 1.f / sqrtf((float)C));
  fill_random_f32<<<cdiv(9 * C, 256), 256>>>(dw, 9 * C, seed + 3, 0.5f);
  fill_random_f32<<<cdiv(N, 256), 256>>>(shift, N, seed + 4, 0.5f);
  CUDA_TRY(cudaMemsetAsync(stats, 0, 8, 0));
  CUDA_TRY(cudaMemsetAsync(o1, 0xff, px * N * 2, 0));
  CUDA_TRY(cudaMemsetAsync(o2, 0x7f, px * N * 2, 0));
  TMAT_TRY(dwconv_run(ctx, in, B, S, C, dw, relu_in, d1, nullptr, 0));
  ConvOp op;
  op.name = "dwpw.ref"; op.a0 = d1; op.a0_C = C; op.B = B; op.H = S; op.W = S; set_tap1x1(op);
  op.wmat = w; op.K = C; op.shift = shift; op.Cout = op.Npad = N; op.epi = EPI_STORE; op.relu = relu_out; op.out = o1;
  TMAT_TRY(conv_run_simt(ctx, op, 0));
  TMAT_TRY(dwpw_run(ctx, in, B, S, C, dw, relu_in, w, shift, N, relu_out, o2, 0));
  max_abs_diff_kernel<bf16><<<256, 256>>>(o1, o2, (int64_t)(px * N), stats);
  float h[2];
  CUDA_TRY(cudaMemcpy(h, stats, 8, cudaMemcpyDeviceToHost));
  *max_abs_diff = h[0];
  *max_abs_ref = h[1];
  return TMAT_OK;
}
