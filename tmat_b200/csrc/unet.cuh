// UNet-Xception forward engine (reference: fl_tissue_model_tools/models.py:85-171).
// Activations are NHWC bf16 in HBM; every convolution that carries real FLOPs is expressed as
// one "ConvOp" = implicit GEMM  D[pixels, Cout] = sum_taps A(shifted pixels, Cin) * W  with a
// fused epilogue, executed either by the tcgen05/TMEM/TMA kernel (conv_tc.cu) or by a plain
// CUDA-core kernel (unet_layers.cu) that exists to cross-check the tensor-core path and to
// serve filter counts that are not multiples of 64.
#pragma once
#include "common.cuh"
#include "stp.cuh"

typedef __nv_bfloat16 bf16;

#ifdef __CUDACC__
// packed fp32 pairs (Blackwell FFMA2 / FMUL2): two channels per instruction
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
  return d;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 fmul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
#endif

enum EpiMode {
  EPI_STORE = 0,    // out[b,y,x,:] (same resolution), optional addend and ReLU'd second copy
  EPI_PAR = 1,      // parity store: GEMM column block q = (py,px) -> out[b,2y+py,2x+px,:]; this is a
                    // 3x3 convolution applied to UpSampling2D(2)(a0) (models.py:150-157) evaluated on
                    // the low-resolution tensor with per-parity summed taps
  EPI_HEAD = 3      // 4 output-pixel parities -> sigmoid -> float32 (B, 2H, 2W)  (models.py:166)
};

constexpr int CONV_MAX_GROUPS = 4;

struct ConvOp {
  // implicit-GEMM A operand: per group g, gtaps shifted views of a0 (channels a0_C); then (only
  // when ngroups == 1) one 1x1 view of a1
  const bf16* a0 = nullptr; int a0_C = 0;
  const bf16* a1 = nullptr; int a1_C = 0;
  int a0_stride = 1;                // a0 is a dense (B, H*s, W*s, a0_C) tensor read at pixels (y*s, x*s): the
                                    // stride-2 1x1 residual conv of the down blocks (1 tap only)
  int B = 0, H = 0, W = 0;          // pixel grid of the GEMM rows (= resolution of a0/a1)
  int ngroups = 1;                  // independent column groups with their own tap lists
  int ntaps = 1;                    // taps per group
  int8_t dy[CONV_MAX_GROUPS][9] = {{0}}, dx[CONV_MAX_GROUPS][9] = {{0}};
  const bf16* wmat = nullptr;       // [ngroups*Npad][K] K-major, BN scale folded in; K = ntaps*a0_C + a1_C
  const bf16* wmat2 = nullptr;      // EPI_HEAD only: [16 = (parity, 2x2 low-res tap)][a0_C] for the taps-in-N head (conv_head.cu)
  int K = 0;
  const float* shift = nullptr;     // [Cout] bias / BN shift (EPI_HEAD: [Npad])
  int Cout = 0;                     // channels of the output tensor
  int Npad = 0;                     // GEMM columns per group (EPI_STORE: Cout; EPI_PAR: Cout * parities per group)
  int epi = EPI_STORE;
  int relu = 0;                     // ReLU on `out`
  bf16* out = nullptr;
  bf16* out2 = nullptr;             // EPI_STORE only: ReLU'd copy (when `out` must stay pre-activation)
  const bf16* add_src = nullptr;    // EPI_STORE only: v += add_src[b, y >> add_shift, x >> add_shift, :]
  int add_shift = 0;                //   add_src has shape (B, ceil(H / 2^s), ceil(W / 2^s), Cout)
  float* out_f32 = nullptr;         // EPI_HEAD
  int logits = 0;
  const char* name = "";
};

// parameters of the epilogue, shared by both executors (device-visible POD)
struct EpiParams {
  int mode, relu, logits;
  int B, H, W, Cout;
  int ppg;                          // EPI_PAR: parities per group (Npad / Cout)
  int add_shift;
  const float* shift;
  bf16* out;
  bf16* out2;
  const bf16* add_src;
  float* out_f32;
};

static inline EpiParams epi_of(const ConvOp& op) {
  EpiParams e;
  e.mode = op.epi; e.relu = op.relu; e.logits = op.logits;
  e.B = op.B; e.H = op.H; e.W = op.W; e.Cout = op.Cout;
  e.ppg = (op.epi == EPI_PAR) ? op.Npad / op.Cout : 1;
  e.add_shift = op.add_shift;
  e.shift = op.shift; e.out = op.out; e.out2 = op.out2; e.add_src = op.add_src; e.out_f32 = op.out_f32;
  return e;
}

// executors
int conv_run_simt(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st);
int conv_run_tc(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st);
bool conv_tc_supported(const ConvOp& op);

// CUDA-core layers (unet_layers.cu)
// stem: Conv2D(C0, 3, stride 2, same) + bias + BN + ReLU on float32 patches -> bf16 (B, S, S, C0).
// Patches come either from a dense (B, ws, ws) float32 buffer or straight from the image through the
// D4 tile map (geom != nullptr): patch p = first_patch + b.
struct StemSrc {
  const float* patches;   // dense source, or nullptr
  const float* img;       // image source (H, W) float32
  const float* minmax;    // device [2]: global min (pad value), max
  StpGeom geom;
  int first_patch;
  int use_geom;
  const int* patch_list;  // optional (device): patch id = patch_list[first_patch + b] (row-strip prediction)
};
int stem_run(tmat_ctx* ctx, const StemSrc& src, int B, int ws, const float* w /*[9][C0] scale folded*/,
             const float* shift, int C0, bf16* out, cudaStream_t st);
// depthwise 3x3 'same' (no bias): in (B,S,S,C) bf16 -> out bf16; optional ReLU on the input;
// optional sub = in[:, ::2, ::2, :] (raw, for the stride-2 1x1 residual conv).
int dwconv_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, const float* w /*[9][C]*/, int relu_in,
               bf16* out, bf16* sub, cudaStream_t st);
// MaxPool2D(3, strides=2, padding="same") on an even size (pads 0 before / 1 after, padding excluded
// from the max; models.py:139): in (B,S,S,C) bf16 -> out (B,S/2,S/2,C) bf16.
int maxpool_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, bf16* out, cudaStream_t st);

// Fused SeparableConv2D on the tcgen05 path (conv_dwpw.cu): out = [relu](BN(pointwise(depthwise([relu](in)))))
// pool_out != nullptr: the epilogue applies MaxPool2D(3, strides=2, 'same') (models.py:139) and leaves only the pooled
// tensor (B, S/2, S/2, N) in pool_out -- `out` is not written (bit-identical to dwpw_run + maxpool_run).
bool dwpw_supported(int B, int S, int C, int N);
bool dwpw_pool_supported(int B, int S, int C, int N);
int dwpw_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, const float* dw /*[9][C]*/, int relu_in,
             const bf16* wmat /*[N][C], BN scale folded*/, const float* shift, int N, int relu_out, bf16* out,
             cudaStream_t st, bf16* pool_out = nullptr);

// 3x3 convolution with the haloed A tile loaded once per tile (conv_halo.cu); conv_run_tc dispatches to it
// for single-group 3x3 ops with N <= 128 (and the head) when conv_halo_supported().
bool conv_halo_supported(const ConvOp& op);
int conv_halo_mode();
int conv_run_halo(tmat_ctx* ctx, const ConvOp& op, int mode, cudaStream_t st);
// Head with the (parity, tap) pairs as the 16 GEMM columns of a 1x1 GEMM over the haloed tile (conv_head.cu)
bool conv_head_taps_supported(const ConvOp& op);
int conv_run_head_taps(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st);
