// Geometry of the smooth tiled prediction (reference: fl_tissue_model_tools/smooth_tiled_predictions.py).
// One struct describes the padded image and the per-orientation tile grids; the D4 maps are
// device functions so the tile gather (stem of the UNet, or tmat_stp_gather) and the blend
// read the image / write the canvas in the ORIGINAL frame -- the 8 rotated copies of
// _rotate_mirror_do (:95-113) are never materialised.
#pragma once
#include "common.cuh"

struct StpGeom {
  int H, W;        // input image
  int ws, subdiv;  // window_size, subdivisions
  int aug;         // round(ws * (1 - 1/subdiv))          (:73)
  int step;        // int(ws / subdiv)                    (:150)
  int PM, PN;      // padded shape                        (:74-79)
  int nR, nC;      // tile positions along PM / PN: len(range(0, L - ws + 1, step))  (:157-163)
  // orientation g: rotated frame has shape (Lm, Ln) = (PM, PN) for even g, (PN, PM) for odd g;
  // tiles (a, b) = (nR, nC) for even g, (nC, nR) for odd g; n_tiles = nR * nC either way.
  __host__ __device__ int tiles_per_orient() const { return nR * nC; }
  __host__ __device__ int tile_rows(int g) const { return (g & 1) ? nC : nR; }
  __host__ __device__ int tile_cols(int g) const { return (g & 1) ? nR : nC; }
};

int stp_make_geom(int H, int W, int ws, int subdiv, StpGeom* g);

// Position (ri, rj) in the frame of orientation g  ->  position (r, c) in the padded original
// frame.  do_g(X) = rot90(X[:, ::-1] if g >= 4 else X, k = g % 4)   (:103-112)
//   rot90(A,1)[i,j] = A[j, N-1-i]   rot90(A,2)[i,j] = A[M-1-i, N-1-j]   rot90(A,3)[i,j] = A[M-1-j, i]
__host__ __device__ __forceinline__ void d4_rot_to_orig(int g, int ri, int rj, int PM, int PN, int* r, int* c) {
  int k = g & 3, ip, jp;
  if (k == 0)      { ip = ri;           jp = rj; }
  else if (k == 1) { ip = rj;           jp = PN - 1 - ri; }
  else if (k == 2) { ip = PM - 1 - ri;  jp = PN - 1 - rj; }
  else             { ip = PM - 1 - rj;  jp = ri; }
  *r = ip;
  *c = (g >= 4) ? PN - 1 - jp : jp;
}

// Inverse: padded original (r, c) -> rotated-frame (ri, rj).
__host__ __device__ __forceinline__ void d4_orig_to_rot(int g, int r, int c, int PM, int PN, int* ri, int* rj) {
  int k = g & 3;
  int ip = r, jp = (g >= 4) ? PN - 1 - c : c;
  if (k == 0)      { *ri = ip;           *rj = jp; }
  else if (k == 1) { *ri = PN - 1 - jp;  *rj = ip; }
  else if (k == 2) { *ri = PM - 1 - ip;  *rj = PN - 1 - jp; }
  else             { *ri = jp;           *rj = PM - 1 - ip; }
}

// Value of the padded image at padded coordinates (constant padding = global minimum, :76-78).
__device__ __forceinline__ float stp_padded_value(const float* __restrict__ img, const StpGeom& G, int r, int c,
                                                  float minval) {
  int y = r - G.aug, x = c - G.aug;
  return (y >= 0 && y < G.H && x >= 0 && x < G.W) ? __ldg(img + (int64_t)y * G.W + x) : minval;
}

// Pixel (i, j) of tile (ti, tj) of orientation g.
__device__ __forceinline__ float stp_tile_pixel(const float* __restrict__ img, const StpGeom& G, int g, int ti,
                                                int tj, int i, int j, float minval) {
  int r, c;
  d4_rot_to_orig(g, ti * G.step + i, tj * G.step + j, G.PM, G.PN, &r, &c);
  return stp_padded_value(img, G, r, c, minval);
}

// device-side helpers implemented in stp.cu
int stp_image_min(tmat_ctx* ctx, const float* img_dev, int64_t n, float* minmax_dev /*[2]*/, cudaStream_t st);
int stp_blend(tmat_ctx* ctx, const StpGeom& G, const float* preds_dev, const double* w1d_dev, double* out_dev,
              cudaStream_t st);
// Blend of output rows [row0, row1) only; preds_dev holds the patches of a row strip in list order and
// slot_of_patch (device, one int per patch of the image, -1 = not computed) maps patch id -> slot.
int stp_blend_rows(tmat_ctx* ctx, const StpGeom& G, const float* preds_dev, const int* slot_of_patch_dev,
                   const double* w1d_dev, int row0, int row1, double* out_rows_dev, cudaStream_t st);
// Patches (id = orientation * tiles_per_orient + tile index in the rotated frame) whose window touches
// output rows [row0, row1), in id order.  Host geometry only.
void stp_patches_for_rows(const StpGeom& G, int row0, int row1, std::vector<int>* list);
int stp_upload_window(tmat_ctx* ctx, int ws, const double* window1d_host, double** w1d_dev, cudaStream_t st);
void spline_window_host(int ws, double* out);
