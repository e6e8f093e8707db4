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

enum EpiMode {
  EPI_STORE = 0,    // out[b,y,x,:]  (same resolution)
  EPI_UP2 = 1,      // out[b,2y+{0,1},2x+{0,1},:]   UpSampling2D(2) fused (models.py:157)
  EPI_POOLADD = 2,  // v += MaxPool(3, s2, same)(pool_src)[b,y,x,:]   (models.py:139-146)
  EPI_HEAD = 3      // 4 output-pixel parities -> sigmoid -> float32 (B, 2H, 2W)  (models.py:166)
};

struct ConvOp {
  // implicit-GEMM A operand: ntaps shifted views of a0 (channels a0_C), then one 1x1 view of a1
  const bf16* a0 = nullptr; int a0_C = 0;
  const bf16* a1 = nullptr; int a1_C = 0;
  int B = 0, H = 0, W = 0;          // pixel grid of the GEMM rows (= resolution of a0/a1)
  int ntaps = 1;
  int8_t dy[9] = {0}, dx[9] = {0};
  const bf16* wmat = nullptr;       // [Npad][K] K-major, BN scale folded in; K = ntaps*a0_C + a1_C
  int K = 0;
  const float* shift = nullptr;     // [Npad] bias / BN shift
  int Cout = 0, Npad = 0;
  int epi = EPI_STORE;
  int relu = 0;                     // ReLU on `out`
  bf16* out = nullptr;
  bf16* out2 = nullptr;             // optional: ReLU'd copy (when `out` must stay pre-activation)
  const bf16* pool_src = nullptr;   // EPI_POOLADD: (B, 2H, 2W, Cout)
  float* out_f32 = nullptr;         // EPI_HEAD
  int logits = 0;
  const char* name = "";
};

// parameters of the epilogue, shared by both executors (device-visible POD)
struct EpiParams {
  int mode, relu, logits;
  int B, H, W, Cout;
  const float* shift;
  bf16* out;
  bf16* out2;
  const bf16* pool_src;
  float* out_f32;
};

static inline EpiParams epi_of(const ConvOp& op) {
  EpiParams e;
  e.mode = op.epi; e.relu = op.relu; e.logits = op.logits;
  e.B = op.B; e.H = op.H; e.W = op.W; e.Cout = op.Cout;
  e.shift = op.shift; e.out = op.out; e.out2 = op.out2; e.pool_src = op.pool_src; e.out_f32 = op.out_f32;
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
};
int stem_run(tmat_ctx* ctx, const StemSrc& src, int B, int ws, const float* w /*[9][C0] scale folded*/,
             const float* shift, int C0, bf16* out, cudaStream_t st);
// depthwise 3x3 'same' (no bias): in (B,S,S,C) bf16 -> out bf16; optional ReLU on the input;
// optional sub = in[:, ::2, ::2, :] (raw, for the stride-2 1x1 residual conv).
int dwconv_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, const float* w /*[9][C]*/, int relu_in,
               bf16* out, bf16* sub, cudaStream_t st);
