/* tmat_b200_debug.h -- test and self-check hooks of libtmat_b200.so.  NOT part of the drop-in boundary
 * (include/tmat_b200.h): nothing in the reference corresponds to these; tests/ uses them to compare the
 * tcgen05 kernels with the CUDA-core executors of the same operation. */
#ifndef TMAT_B200_DEBUG_H
#define TMAT_B200_DEBUG_H
#include "tmat_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* debug / test hook: copy a named intermediate activation of the last forward (float32 NHWC)
 * to the host; returns TMAT_ERR_INVALID when the tap does not exist. */
int tmat_unet_debug_tap(tmat_ctx* ctx, tmat_unet* net, const char* name, float* out_host,
                        int64_t capacity, int64_t* n_written);
/* debug / test hook: one implicit-GEMM convolution (the unit the UNet is built from) on seeded
 * random data through the CUDA-core executor and through the tcgen05 executor; returns the
 * max-abs difference of the outputs and the max-abs of the CUDA-core output.
 * epi: 0 store, 1 parity store (3x3 conv on the 2x nearest-upsampled input, evaluated at low
 * resolution), 3 head (sigmoid unless relu!=0).  dual: also write the ReLU'd copy (epi 0).
 * addend: 0 none, 1 same-resolution addend, 2 half-resolution addend (epi 0). */
int tmat_debug_conv_selftest(tmat_ctx* ctx, int B, int H, int W, int C0, int C1, int Cout, int ntaps,
                             int epi, int relu, int dual, int addend, uint64_t seed,
                             double* max_abs_diff, double* max_abs_ref);
/* Debug switch of the haloed-tile 3x3 kernel (conv_halo.cu): 0 off, 1 on.  Default: env TMAT_HALO or 1. */
int tmat_debug_set_halo_mode(int mode);
/* Self-test hook: the fused SeparableConv2D kernel (depthwise 3x3 computed into the tcgen05 A operand)
 * vs the unfused pair (depthwise kernel, CUDA-core GEMM) on random data. */
int tmat_debug_dwpw_selftest(tmat_ctx* ctx, int B, int S, int C, int N, int relu_in, int relu_out,
                             uint64_t seed, double* max_abs_diff, double* max_abs_ref);
/* the same kernel with MaxPool2D(3, strides=2, 'same') (models.py:139) folded into its epilogue, against the unfused
 * kernel followed by the stand-alone pooling kernel: max_abs_diff must be 0. */
int tmat_debug_dwpw_pool_selftest(tmat_ctx* ctx, int B, int S, int C, int N, int relu_in, int relu_out,
                                  uint64_t seed, double* max_abs_diff, double* max_abs_ref);
#ifdef __cplusplus
}
#endif
#endif /* TMAT_B200_DEBUG_H */
