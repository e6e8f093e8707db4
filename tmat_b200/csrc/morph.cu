// Stage 4: binary morphology and connected-component cleanup.
// Reference: fl_tissue_model_tools/transforms.py:170-206 (remove_small_islands), :291-303
// (regionprops_image), :306-361 (filter_branch_seg_mask), well_mask_generation.py:275 (binary_erosion) and
// the scikit-image 0.22 primitives they call (filters.median on a bool image, measure.label,
// regionprops area / perimeter, morphology.skeletonize, remove_small_objects, disk).
// All integer work, bit-exact against oracle/morphology.py; the scikit-image primitives that oracle restates --
// above all the thinning table, csrc/skeleton_lut.inc -- are PARITY UNPINNED (scikit-image is installable on no
// machine this project reaches).  Masks are uint8 images; any non-zero value is foreground.
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>

namespace {

// ---- median with the disk(2) footprint on a boolean image: >= 7 of the 13 pixels set -----------------
// (skimage.filters.median -> scipy.ndimage.median_filter, mode 'nearest' = index clamp)
__global__ void __launch_bounds__(256) median_disk2_kernel(const uint8_t* __restrict__ in, int H, int W,
                                                           uint8_t* __restrict__ out) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int cnt = 0;
#pragma unroll
  for (int dy = -2; dy <= 2; ++dy) {
    int hw = (dy == 0) ? 2 : ((dy == -1 || dy == 1) ? 1 : 0);
    int yy = min(max(y + dy, 0), H - 1);
    for (int dx = -hw; dx <= hw; ++dx) {
      int xx = min(max(x + dx, 0), W - 1);
      cnt += in[(int64_t)yy * W + xx] != 0;
    }
  }
  out[(int64_t)y * W + x] = cnt >= 7;
}

// Bit-sliced version (the default): a warp owns 32 columns and a segment of rows.  A row (32 + 4 pixels, indices clamped
// = mode 'nearest') becomes a 64-bit mask by two ballots; its horizontal counts over half-widths 0 / 1 / 2 are held as
// bit planes (S1: the pixel, S3 = 2 planes, S5 = 3 planes; full adders are one XOR3 and one MAJ3 logic operation on 32
// pixels at once), and output row y adds S5[y] + S3[y-1] + S3[y+1] + S1[y-2] + S1[y+2] with six more full adders:
// count >= 7 <=> bit3 | (bit2 & bit1 & bit0).  ~30 logic operations per 32 pixels instead of 13 clamped loads and
// compares per pixel (26 us -> per 2048^2, ALU pipe 71 %, round-2 ncu).
constexpr int MED_SEG = 32;
__device__ __forceinline__ unsigned int xor3(unsigned int a, unsigned int b, unsigned int c) { return a ^ b ^ c; }
__device__ __forceinline__ unsigned int maj3(unsigned int a, unsigned int b, unsigned int c) { return (a & b) | (c & (a | b)); }
__global__ void __launch_bounds__(128) median_disk2_bits(const uint8_t* __restrict__ in, int H, int W, uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x0 = blockIdx.x * 32;
  const int y0 = (blockIdx.y * 4 + warp) * MED_SEG;
  if (y0 >= H) return;
  const int y1 = min(y0 + MED_SEG, H);
  const int xa = min(max(x0 - 2 + lane, 0), W - 1), xb = min(x0 + 30 + lane, W - 1);     // clamped column indices
  // planes of the last five input rows, slot = (row - (y0 - 2)) mod 5, kept static by unrolling five rows
  unsigned int s1[5], a0[5], a1[5], t0[5], u1[5], u2[5];
  for (int base = y0 - 2; base < y1 + 2; base += 5) {
    uint8_t va[5], vb[5];
#pragma unroll
    for (int s = 0; s < 5; ++s) {
      const int yi = min(max(base + s, 0), H - 1);
      va[s] = in[(int64_t)yi * W + xa];
      vb[s] = in[(int64_t)yi * W + xb];
    }
#pragma unroll
    for (int s = 0; s < 5; ++s) {
      const int yi = base + s;                               // rows past the segment only fill slots that are never read
      const unsigned int lo = __ballot_sync(0xffffffffu, va[s] != 0);
      const unsigned int hi = __ballot_sync(0xffffffffu, vb[s] != 0);     // only bits 0..3 are used
      // w[d]: bit i = pixel x0 + i + d
      const unsigned int wm2 = lo, wm1 = __funnelshift_r(lo, hi, 1), w0 = __funnelshift_r(lo, hi, 2),
                         wp1 = __funnelshift_r(lo, hi, 3), wp2 = __funnelshift_r(lo, hi, 4);
      s1[s] = w0;
      a0[s] = xor3(wm1, w0, wp1);                            // S3 = a1 a0
      a1[s] = maj3(wm1, w0, wp1);
      t0[s] = xor3(a0[s], wm2, wp2);                         // S5 = u2 u1 t0
      const unsigned int c1 = maj3(a0[s], wm2, wp2);
      u1[s] = a1[s] ^ c1;
      u2[s] = a1[s] & c1;
      // output row yo = yi - 2 has now seen rows yo - 2 .. yo + 2: slots (s - 4 .. s) mod 5
      const int yo = yi - 2;
      if (yo >= y0 && yo < y1) {
        constexpr int M = 5;
        const int sm2 = (s + 1) % M, sm1 = (s + 2) % M, sc = (s + 3) % M, sp1 = (s + 4) % M, sp2 = s;   // rows yo-2 .. yo+2
        const unsigned int x0b = xor3(t0[sc], a0[sm1], a0[sp1]), cx = maj3(t0[sc], a0[sm1], a0[sp1]);
        const unsigned int r0 = xor3(x0b, s1[sm2], s1[sp2]), cy = maj3(x0b, s1[sm2], s1[sp2]);
        const unsigned int x1b = xor3(u1[sc], a1[sm1], a1[sp1]), dx = maj3(u1[sc], a1[sm1], a1[sp1]);
        const unsigned int r1 = xor3(x1b, cx, cy), dy = maj3(x1b, cx, cy);
        const unsigned int r2 = xor3(u2[sc], dx, dy), r3 = maj3(u2[sc], dx, dy);
        const unsigned int res = r3 | (r2 & r1 & r0);
        const int x = x0 + lane;
        if (x < W) out[(int64_t)yo * W + x] = (res >> lane) & 1u;
      }
    }
  }
}

// ---- erosion with disk(r): all footprint pixels set; outside the image counts as border_value --------
// Bit-parallel: a warp owns a strip of 32 columns and a segment of rows.  Each input row (32 + 2r pixels) becomes one
// 64-bit mask by two ballots; E_h = AND of the row shifted by -h..h (two operations per h) is the row eroded by a
// horizontal segment of half-width h; the row of the disk at vertical offset dy has half-width hw(dy), so the input
// row y contributes E_hw(dy) to output row y - dy: 2r + 1 running 32-bit accumulators, one stored per input row.
// ~50 instructions per 32 pixels instead of (2r+1)^2 shared-memory loads per pixel (round 1: 84 us per 2048^2 at r = 5).
// thr_params != nullptr fuses the '>= Otsu threshold' of auto_threshold_well into the load (in = the uint8 image,
// params[0] = invert, params[1] = threshold; well_mask_generation.py:265-274).
constexpr int ER_MAX_R = 8, ER_SEG = 64;
// half-width of the disk's row at vertical offset dy: largest h with h^2 + dy^2 <= R^2 (skimage.morphology.disk)
__host__ __device__ constexpr int disk_half_width(int R, int dy) {
  int h = 0;
  while ((h + 1) * (h + 1) + dy * dy <= R * R) ++h;
  return h;
}
template <int R, int ER_SEGT = ER_SEG>
__global__ void __launch_bounds__(128) erode_disk_bits(const uint8_t* __restrict__ in, int H, int W, int border_value,
                                                       const int* __restrict__ thr_params, uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x0 = blockIdx.x * 32;
  const int y0 = (blockIdx.y * 4 + warp) * ER_SEGT;
  if (y0 >= H) return;
  const int y1 = min(y0 + ER_SEGT, H);
  // one predicate for both uses: a plain mask is 'v >= 1', the fused Otsu threshold '(invert ? 255 - v : v) >= thr'
  const int thr = thr_params ? thr_params[1] : 1;
  const int vxor = (thr_params && thr_params[0]) ? 255 : 0;            // 255 - v == v ^ 255 for a byte
  const bool bord = border_value != 0;
  unsigned int acc[2 * R + 1];
#pragma unroll
  for (int k = 0; k <= 2 * R; ++k) acc[k] = 0xffffffffu;
  // input rows y0 - R .. y1 - 1 + R; slot of output row yo: (yo - y0 + R... ) kept static by unrolling 2R+1 rows.
  // Straight-line row body: loads go to clamped addresses and are overridden by selects, rows past the segment's end
  // are processed like any other (they only reach accumulator slots that are never stored) -- the first version spent
  // ~155 instructions per row, most of them compares, branches and reconvergence barriers around the two ballots.
  const int xa = x0 - R + lane, xb = xa + 32;
  const bool xa_in = xa >= 0 && xa < W, xb_in = lane < 2 * R && xb < W;     // xb >= 0 always (x0 >= 0, lane >= 0)
  const int xa_c = min(max(xa, 0), W - 1), xb_c = min(xb, W - 1);
  const bool hi_lane = lane < 2 * R;
  const int xo = x0 + lane;
  for (int base = y0 - R; base < y1 + R; base += 2 * R + 1) {
    // all loads of the batch first: 2 (2R + 1) independent loads in flight per lane
    int va[2 * R + 1], vb[2 * R + 1];
#pragma unroll
    for (int s = 0; s <= 2 * R; ++s) {
      const int yc = min(max(base + s, 0), H - 1);
      const uint8_t* p = in + (int64_t)yc * W;
      va[s] = p[xa_c];
      vb[s] = p[xb_c];
    }
#pragma unroll
    for (int s = 0; s <= 2 * R; ++s) {
      const int yi = base + s;
      {
        const bool row_in = yi >= 0 && yi < H;
        const bool ba = (row_in && xa_in) ? ((va[s] ^ vxor) >= thr) : bord;
        const bool bb = hi_lane && ((row_in && xb_in) ? ((vb[s] ^ vxor) >= thr) : bord);
        const unsigned int rlo = __ballot_sync(0xffffffffu, ba);
        const unsigned int rhi = __ballot_sync(0xffffffffu, bb);
        // E[h]: bit j set iff pixels j-h .. j+h of the row are all set.  Only the 32 bits R .. R+31 are ever used, so the
        // shifted copies are taken as 32-bit windows of the 64-bit row (one funnel shift each) and E[h] = E[h-1] & w[-h] &
        // w[+h] is one three-input logic operation: 2R + 1 shifts + R LOP3 per row instead of 64-bit shift/AND chains
        // (ncu, round 2: the ALU pipe was 78 % busy with those, 37 us per 2048^2 image).
        unsigned int eh[R + 1];
        eh[0] = __funnelshift_r(rlo, rhi, R);
#pragma unroll
        for (int h = 1; h <= R; ++h)
          eh[h] = eh[h - 1] & __funnelshift_r(rlo, rhi, R - h) & __funnelshift_r(rlo, rhi, R + h);
        // this input row is row (yi - yo) of the footprint for output row yo = yi - dy
#pragma unroll
        for (int k = 0; k <= 2 * R; ++k) {
          // dy = k - R; output row yo = yi - dy has slot (s - dy) mod (2R+1)  [slot of row y = (y - (y0 - R)) mod (2R+1)]
          acc[(s - (k - R) + 2 * R + 1) % (2 * R + 1)] &= eh[disk_half_width(R, k - R)];
        }
        // output row yi - R is complete (it has now seen input rows yi - 2R .. yi): its slot is (s - R) mod (2R+1)
        const int yo = yi - R;
        constexpr int M = 2 * R + 1;
        const int slot = (s - R + M) % M;
        if (yo >= y0 && yo < y1 && xo < W) out[(int64_t)yo * W + xo] = (acc[slot] >> lane) & 1u;
        acc[slot] = 0xffffffffu;
      }
    }
  }
}

// ---- eight pixels per lane: erosion and median for widths on the 8-pixel grid ---------------------------------------
// The ballot kernels above pack 32 pixels per WARP instruction; here a lane loads one 8-byte word, packs its pixels into
// 8 bits (SIMD byte compare + multiply), gets its neighbours' bits by two shuffles and runs the same shift / AND / adder
// logic on its own 24-bit window -- 256 pixels per warp instruction.  A lane outside the row supplies the border
// (erosion: border_value; median: the row's last pixel replicated = mode 'nearest'), lane 0 / lane 31 load the word
// beyond the warp's 256-column span themselves.  Output bits are spread back to 0/1 bytes and stored as one word.
__device__ __forceinline__ uint32_t pack4_ge(uint32_t w, uint32_t vx4, uint32_t thr4) {
  const uint32_t m = __vcmpgeu4(w ^ vx4, thr4) & 0x01010101u;              // byte k -> bit 8k
  return ((m * 0x01020408u) >> 24) & 0xfu;                                  // -> bits 0..3
}
__device__ __forceinline__ uint32_t pack8_ge(unsigned long long w, uint32_t vx4, uint32_t thr4) {
  return pack4_ge((uint32_t)w, vx4, thr4) | (pack4_ge((uint32_t)(w >> 32), vx4, thr4) << 4);
}
__device__ __forceinline__ unsigned long long unpack8(uint32_t bits) {
  unsigned long long x = ((unsigned long long)(bits & 0xffu) * 0x0101010101010101ull) & 0x8040201008040201ull;   // bit k of byte k
  return ((x + 0x7f7f7f7f7f7f7f7full) >> 7) & 0x0101010101010101ull;
}
constexpr int EW_SEG = 16;
template <int R>
__global__ void __launch_bounds__(128) erode_disk_words(const uint8_t* __restrict__ in, int H, int W, int border_value,
                                                        const int* __restrict__ thr_params, uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wpr = W >> 3;
  const int xw = blockIdx.x * 32 + lane;                      // this lane's word in the row
  const int y0 = (blockIdx.y * 4 + warp) * EW_SEG;
  if (y0 >= H) return;
  const int y1 = min(y0 + EW_SEG, H);
  const uint32_t thr4 = (uint32_t)(thr_params ? thr_params[1] : 1) * 0x01010101u;
  const uint32_t vx4 = (thr_params && thr_params[0]) ? 0xffffffffu : 0u;
  const uint32_t bord8 = border_value ? 0xffu : 0u;
  const bool in_row = xw < wpr;
  // the word beyond the warp's span that this lane fetches (lane 0: left, lane 31: right), -1: none / outside
  const int xe = lane == 0 ? xw - 1 : (lane == 31 ? xw + 1 : -1);
  const bool e_in = xe >= 0 && xe < wpr;
  const int xw_c = min(xw, wpr - 1), xe_c = min(max(xe, 0), wpr - 1);
  unsigned int acc[2 * R + 1];
#pragma unroll
  for (int k = 0; k <= 2 * R; ++k) acc[k] = 0xffu;
  for (int base = y0 - R; base < y1 + R; base += 2 * R + 1) {
    unsigned long long wa[2 * R + 1], we[2 * R + 1];
#pragma unroll
    for (int s = 0; s <= 2 * R; ++s) {
      const int yc = min(max(base + s, 0), H - 1);
      const unsigned long long* row = reinterpret_cast<const unsigned long long*>(in + (int64_t)yc * W);
      wa[s] = row[xw_c];
      we[s] = row[xe_c];
    }
#pragma unroll
    for (int s = 0; s <= 2 * R; ++s) {
      const int yi = base + s;
      const bool row_in = yi >= 0 && yi < H;
      const uint32_t own = (row_in && in_row) ? pack8_ge(wa[s], vx4, thr4) : bord8;
      const uint32_t ext = (row_in && e_in) ? pack8_ge(we[s], vx4, thr4) : bord8;
      uint32_t l = __shfl_up_sync(0xffffffffu, own, 1), r = __shfl_down_sync(0xffffffffu, own, 1);
      if (lane == 0) l = ext;
      if (lane == 31) r = ext;
      const uint32_t w = l | (own << 8) | (r << 16);          // bit j = pixel 8 xw - 8 + j
      unsigned int eh[R + 1];
      eh[0] = w >> 8;
#pragma unroll
      for (int h = 1; h <= R; ++h) eh[h] = eh[h - 1] & (w >> (8 - h)) & (w >> (8 + h));
#pragma unroll
      for (int k = 0; k <= 2 * R; ++k) acc[(s - (k - R) + 2 * R + 1) % (2 * R + 1)] &= eh[disk_half_width(R, k - R)];
      const int yo = yi - R;
      constexpr int M = 2 * R + 1;
      const int slot = (s - R + M) % M;
      if (yo >= y0 && yo < y1 && in_row)
        reinterpret_cast<unsigned long long*>(out + (int64_t)yo * W)[xw] = unpack8(acc[slot]);
      acc[slot] = 0xffu;
    }
  }
}
constexpr int MW_SEG = 16;
__global__ void __launch_bounds__(128) median_disk2_words(const uint8_t* __restrict__ in, int H, int W, uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wpr = W >> 3;
  const int xw = blockIdx.x * 32 + lane;
  const int y0 = (blockIdx.y * 4 + warp) * MW_SEG;
  if (y0 >= H) return;
  const int y1 = min(y0 + MW_SEG, H);
  const bool in_row = xw < wpr;
  const int xe = lane == 0 ? xw - 1 : (lane == 31 ? xw + 1 : -1);
  const int xw_c = min(xw, wpr - 1), xe_c = min(max(xe, 0), wpr - 1);
  unsigned int s1[5], a0[5], a1[5], t0[5], u1[5], u2[5];
  for (int base = y0 - 2; base < y1 + 2; base += 5) {
    unsigned long long wa[5], we[5];
#pragma unroll
    for (int s = 0; s < 5; ++s) {
      const int yc = min(max(base + s, 0), H - 1);
      const unsigned long long* row = reinterpret_cast<const unsigned long long*>(in + (int64_t)yc * W);
      wa[s] = row[xw_c];
      we[s] = row[xe_c];
    }
#pragma unroll
    for (int s = 0; s < 5; ++s) {
      const int yi = base + s;
      uint32_t own = pack8_ge(wa[s], 0u, 0x01010101u);
      if (!in_row) own = (own & 0x80u) ? 0xffu : 0u;           // beyond the row: its last pixel, replicated
      uint32_t ext = pack8_ge(we[s], 0u, 0x01010101u);
      if (xe < 0) ext = (ext & 1u) ? 0xffu : 0u;               // left of column 0: the first pixel, replicated
      if (xe >= wpr) ext = (ext & 0x80u) ? 0xffu : 0u;
      uint32_t l = __shfl_up_sync(0xffffffffu, own, 1), r = __shfl_down_sync(0xffffffffu, own, 1);
      if (lane == 0) l = ext;
      if (lane == 31) r = ext;
      const uint32_t w = l | (own << 8) | (r << 16);          // bit j = pixel 8 xw - 8 + j (clamped)
      const uint32_t wm2 = w >> 6, wm1 = w >> 7, w0 = w >> 8, wp1 = w >> 9, wp2 = w >> 10;   // bit i = pixel 8 xw + i + d
      s1[s] = w0;
      a0[s] = xor3(wm1, w0, wp1);
      a1[s] = maj3(wm1, w0, wp1);
      t0[s] = xor3(a0[s], wm2, wp2);
      const unsigned int c1 = maj3(a0[s], wm2, wp2);
      u1[s] = a1[s] ^ c1;
      u2[s] = a1[s] & c1;
      const int yo = yi - 2;
      if (yo >= y0 && yo < y1 && in_row) {
        constexpr int M = 5;
        const int sm2 = (s + 1) % M, sm1 = (s + 2) % M, sc = (s + 3) % M, sp1 = (s + 4) % M, sp2 = s;
        const unsigned int x0b = xor3(t0[sc], a0[sm1], a0[sp1]), cx = maj3(t0[sc], a0[sm1], a0[sp1]);
        const unsigned int r0 = xor3(x0b, s1[sm2], s1[sp2]), cy = maj3(x0b, s1[sm2], s1[sp2]);
        const unsigned int x1b = xor3(u1[sc], a1[sm1], a1[sp1]), dx = maj3(u1[sc], a1[sm1], a1[sp1]);
        const unsigned int r1 = xor3(x1b, cx, cy), dy = maj3(x1b, cx, cy);
        const unsigned int r2 = xor3(u2[sc], dx, dy), r3 = maj3(u2[sc], dx, dy);
        reinterpret_cast<unsigned long long*>(out + (int64_t)yo * W)[xw] = unpack8(r3 | (r2 & r1 & r0));
      }
    }
  }
}

// ---- rank filter of a boolean image under an arbitrary footprint ------------------------------------------
// offs: (dy, dx) of the nfp footprint members relative to the centre.  clamp != 0: indices clamped to the image
// (ndimage mode 'nearest', the median); else pixels outside count as border_value (binary_erosion).
// out = (number of set pixels under the footprint) >= need.
__global__ void __launch_bounds__(256) footprint_count_kernel(const uint8_t* __restrict__ in, int H, int W,
                                                              const short2* __restrict__ offs, int nfp, int clamp,
                                                              int border_value, int need, uint8_t* __restrict__ out) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int cnt = 0;
  for (int k = 0; k < nfp; ++k) {
    short2 o = offs[k];
    int yy = y + o.x, xx = x + o.y;
    if (clamp) {
      yy = min(max(yy, 0), H - 1);
      xx = min(max(xx, 0), W - 1);
      cnt += in[(int64_t)yy * W + xx] != 0;
    } else {
      bool inside = yy >= 0 && yy < H && xx >= 0 && xx < W;
      cnt += inside ? (in[(int64_t)yy * W + xx] != 0) : (border_value != 0);
    }
  }
  out[(int64_t)y * W + x] = cnt >= need;
}

// ---- connected components: union-find on pixel indices; the root of a component is its first
// pixel in raster order (links always point from the larger root to the smaller) ------------------------
__device__ __forceinline__ int uf_find(int* parent, int i) {
  int p = parent[i];
  while (p != i) { i = p; p = parent[i]; }
  return i;
}
__device__ __forceinline__ void uf_union(int* parent, int a, int b) {
  while (true) {
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a == b) return;
    if (a < b) { int t = a; a = b; b = t; }
    int old = atomicMin(&parent[a], b);
    if (old == a) return;
    a = old;
  }
}

// Step 1: every foreground pixel points at the start of its horizontal run inside its 32-pixel warp segment (one
// ballot per warp instead of one atomic per pixel pair; a tree is at most one link deep afterwards).
constexpr int CCL_RPB = 4;    // rows per CTA of ccl_init
__global__ void __launch_bounds__(256) ccl_init(const uint8_t* __restrict__ m, int H, int W, int fg_is_zero,
                                                int* __restrict__ parent) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, yb = blockIdx.y * CCL_RPB;
  const bool in = x < W;
  const int lane = threadIdx.x & 31;
  uint8_t v[CCL_RPB];
#pragma unroll
  for (int k = 0; k < CCL_RPB; ++k) v[k] = (in && yb + k < H) ? m[(int64_t)(yb + k) * W + x] : (uint8_t)(fg_is_zero ? 1 : 0);
#pragma unroll
  for (int k = 0; k < CCL_RPB; ++k) {
    const int y = yb + k;
    if (y >= H) break;                                   // CTA-uniform
    const bool fg = in && (fg_is_zero ? (v[k] == 0) : (v[k] != 0));
    const unsigned int bits = __ballot_sync(0xffffffffu, fg);
    if (!in) continue;
    // highest zero bit below this lane -> the run starts one past it
    const unsigned int below = ~bits & ((1u << lane) - 1u);
    const int start = below ? 32 - __clz(below) : 0;
    parent[y * W + x] = fg ? y * W + (x - lane + start) : -1;
  }
}
// Step 2: link the runs.  A pixel only issues the unions its left neighbour has not already made:
//   * a run that continues across a 32-pixel segment boundary is joined to the previous segment;
//   * upwards (8-connectivity): W set -> only NE, and only if N is clear (W already met N and NW);
//     W clear -> N if set (N's own run reaches NW and NE), else NW and NE separately.
//   * 4-connectivity: N only, and only for the first pixel of a run or where the left neighbour had no N.
// Foreground tests read the byte mask (not the 4-byte parent array) and a CTA walks MRG_RPB rows, each row's three
// neighbour bytes serving as the "row above" of the next.
constexpr int MRG_RPB = 4;
__global__ void __launch_bounds__(256) ccl_merge(const uint8_t* __restrict__ m, int fg_is_zero, int* __restrict__ parent, int H, int W,
                                                 int conn8) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, yb = blockIdx.y * MRG_RPB;
  if (x >= W) return;
  const uint8_t bg = fg_is_zero ? 1 : 0;                 // a byte that reads as background
  uint8_t L[MRG_RPB + 1], C[MRG_RPB + 1], Rt[MRG_RPB + 1];
#pragma unroll
  for (int k = 0; k <= MRG_RPB; ++k) {
    const int y = yb - 1 + k;
    const bool row_in = y >= 0 && y < H;
    const uint8_t* p = m + (int64_t)(row_in ? y : 0) * W;
    C[k] = row_in ? p[x] : bg;
    L[k] = (row_in && x > 0) ? p[x - 1] : bg;
    Rt[k] = (row_in && x + 1 < W) ? p[x + 1] : bg;
  }
  auto fg = [&](uint8_t v) -> bool { return fg_is_zero ? (v == 0) : (v != 0); };
#pragma unroll
  for (int k = 1; k <= MRG_RPB; ++k) {
    const int y = yb + k - 1;
    if (y >= H) break;
    if (!fg(C[k])) continue;
    const int i = y * W + x;
    const bool w = fg(L[k]);
    if (w && (x & 31) == 0) uf_union(parent, i, i - 1);
    if (y == 0) continue;
    const bool n = fg(C[k - 1]), ne = fg(Rt[k - 1]), nw = fg(L[k - 1]);
    if (conn8) {
      if (w) {
        if (ne && !n) uf_union(parent, i, i - W + 1);
      } else if (n) {
        uf_union(parent, i, i - W);
      } else {
        if (nw) uf_union(parent, i, i - W - 1);
        if (ne) uf_union(parent, i, i - W + 1);
      }
    } else if (n && !(w && nw)) {
      uf_union(parent, i, i - W);        // W and NW both set: the left neighbour already joined this run to the run above
    }
  }
}

// The streaming kernels of this file take EPT elements per thread, 256 apart (coalesced, no alignment requirement, all
// loads issued before the first use): one element per thread meant 16 384 one-shot CTAs per 2048^2 image and 13-16 us
// per pass whatever the pass did (ncu, round 2).
constexpr int EPT = 4;
__global__ void __launch_bounds__(256) ccl_flatten(int* __restrict__ parent, int64_t n) {
  const int64_t base = (int64_t)blockIdx.x * (256 * EPT) + threadIdx.x;
  int p[EPT];
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; p[k] = i < n ? parent[i] : -1; }
#pragma unroll
  for (int k = 0; k < EPT; ++k) {
    const int64_t i = base + k * 256;
    if (p[k] >= 0) parent[i] = uf_find(parent, (int)i);
  }
}
// ccl_flatten that also zeroes the per-root statistics of the cleanup driver (area, the 10 perimeter counters, the kill
// flag) at every root it meets: one pass less over the parent array than a separate initialisation kernel
__global__ void __launch_bounds__(256) ccl_flatten_init(int* __restrict__ parent, int64_t n, unsigned int* __restrict__ area,
                                                        unsigned int* __restrict__ hist10, uint8_t* __restrict__ kill) {
  const int64_t base = (int64_t)blockIdx.x * (256 * EPT) + threadIdx.x;
  int p[EPT];
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; p[k] = i < n ? parent[i] : -1; }
#pragma unroll
  for (int k = 0; k < EPT; ++k) {
    const int64_t i = base + k * 256;
    if (p[k] < 0) continue;
    const int r = uf_find(parent, (int)i);
    parent[i] = r;
    if (r == (int)i) {
      area[i] = 0;
#pragma unroll
      for (int c = 0; c < 10; ++c) hist10[i * 10 + c] = 0;
      kill[i] = 0;
    }
  }
}

// raster-order renumbering: rank of each root among the roots (3-kernel scan over 1024-pixel blocks)
constexpr int SCAN_BLOCK = 1024;
__global__ void __launch_bounds__(256) roots_count(const int* __restrict__ parent, int64_t n, int* __restrict__ bsum) {
  __shared__ int ws[8];
  int64_t base = (int64_t)blockIdx.x * SCAN_BLOCK;
  int c = 0;
  for (int k = threadIdx.x; k < SCAN_BLOCK; k += 256) {
    int64_t i = base + k;
    c += (i < n && parent[i] == (int)i);
  }
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += ws[w];
    bsum[blockIdx.x] = t;
  }
}
// single block: exclusive scan of bsum[nb] in place; total -> total_out
__global__ void __launch_bounds__(1024) scan_block_sums(int* __restrict__ bsum, int nb, int* __restrict__ total_out) {
  __shared__ int s[1024];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    int i = base + threadIdx.x;
    int v = (i < nb) ? bsum[i] : 0;
    s[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      int t = (threadIdx.x >= o) ? s[threadIdx.x - o] : 0;
      __syncthreads();
      s[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < nb) bsum[i] = carry + s[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 0) carry += s[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}
// rank[root] = 1-based raster-order label, written at the root's own position
__global__ void __launch_bounds__(256) roots_rank(const int* __restrict__ parent, int64_t n, const int* __restrict__ boff,
                                                  int* __restrict__ rank) {
  // one warp-serial pass per 1024 block: 256 threads x 4 consecutive pixels, block-level exclusive scan
  __shared__ int ws[8];
  int64_t base = (int64_t)blockIdx.x * SCAN_BLOCK + threadIdx.x * 4;
  int f[4], c = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) { int64_t i = base + k; f[k] = (i < n && parent[i] == (int)i); c += f[k]; }
  int incl = c;
  for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += t; }
  if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = incl;
  __syncthreads();
  int woff = 0;
  for (int w = 0; w < (threadIdx.x >> 5); ++w) woff += ws[w];
  int run = boff[blockIdx.x] + woff + incl - c;
#pragma unroll
  for (int k = 0; k < 4; ++k) { int64_t i = base + k; if (f[k]) rank[i] = ++run; }
}
__global__ void __launch_bounds__(256) labels_from_rank(const int* __restrict__ parent, const int* __restrict__ rank,
                                                        int64_t n, int32_t* __restrict__ labels) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int p = parent[i];
  labels[i] = p < 0 ? 0 : rank[p];
}

// ---- region statistics: area and skimage's perimeter border-code histogram ---------------------------
// border = fg with a 4-neighbour that is bg (outside = bg); code = sum over the 3x3 neighbourhood of
// border * [[10,2,10],[2,1,2],[10,2,10]]  (skimage.measure.perimeter, neighbourhood 4)
__global__ void __launch_bounds__(256) border_kernel(const int32_t* __restrict__ lab, int H, int W,
                                                     uint8_t* __restrict__ border) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int64_t i = (int64_t)y * W + x;
  bool fg = lab[i] > 0;
  bool interior = fg && y > 0 && y + 1 < H && x > 0 && x + 1 < W && lab[i - W] > 0 && lab[i + W] > 0 &&
                  lab[i - 1] > 0 && lab[i + 1] > 0;
  border[i] = fg && !interior;
}
__global__ void __launch_bounds__(256) region_stats_kernel(const int32_t* __restrict__ lab, const uint8_t* __restrict__ border,
                                                           int H, int W, unsigned int* __restrict__ area,
                                                           unsigned int* __restrict__ hist /*[n][50]*/) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int64_t i = (int64_t)y * W + x;
  int l = lab[i];
  if (l <= 0) return;
  atomicAdd(&area[l - 1], 1u);
  if (!border[i]) return;
  int code = 0;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      int yy = y + dy, xx = x + dx;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      int wgt = (dy == 0 && dx == 0) ? 1 : ((dy == 0 || dx == 0) ? 2 : 10);
      code += border[(int64_t)yy * W + xx] ? wgt : 0;
    }
  atomicAdd(&hist[(int64_t)(l - 1) * 50 + code], 1u);
}
// circularity 4*pi*area / (perimeter^2 + 1e-7) > 0.8  (transforms.py:326-329, :346), per label
__global__ void __launch_bounds__(256) circ_flag_kernel(const unsigned int* __restrict__ area,
                                                        const unsigned int* __restrict__ hist, int n,
                                                        uint8_t* __restrict__ round_flag /*[n+1]*/) {
  int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n) return;
  const double w1 = 1.0, w2 = 1.4142135623730951, w3 = (1.0 + 1.4142135623730951) / 2.0;
  const unsigned int* h = hist + (int64_t)l * 50;
  double perim = 0.0;
  for (int c = 0; c < 50; ++c) {
    double w = (c == 5 || c == 7 || c == 15 || c == 17 || c == 25 || c == 27) ? w1
             : ((c == 21 || c == 33) ? w2 : ((c == 13 || c == 23) ? w3 : 0.0));
    perim = __dadd_rn(perim, __dmul_rn((double)h[c], w));
  }
  double circ = __ddiv_rn(__dmul_rn(__dmul_rn(4.0, 3.141592653589793), (double)area[l]),
                          __dadd_rn(__dmul_rn(perim, perim), 1e-7));
  round_flag[l + 1] = circ > 0.8;
}

// the ten border codes that carry a perimeter weight, in code order (= the order skimage's 50-bin dot product adds them)
__device__ __forceinline__ int perim_slot(int code) {
  switch (code) {
    case 5: return 0; case 7: return 1; case 13: return 2; case 15: return 3; case 17: return 4;
    case 21: return 5; case 23: return 6; case 25: return 7; case 27: return 8; case 33: return 9;
  }
  return -1;
}

// ---- the cleanup driver's statistics, indexed by ROOT PIXEL instead of by label rank ------------------------------
// filter_branch_seg_mask never needs the raster-order label numbers, only "which component": the root pixel index does.
// Per-root arrays have one entry per pixel but only root entries are ever touched (zeroed by ccl_flatten_init), so the
// four renumbering kernels and every memset of a label-sized array disappear from the path.
// border pixels of the mask (skimage perimeter: foreground with a background or outside 4-neighbour), from the 0/1 mask
// itself (a byte per pixel instead of the 4-byte parent array), RPB rows per CTA
constexpr int RPB = 4;
static inline dim3 grid2d_rpb(int H, int W) { return dim3(cdiv(W, 256), cdiv(H, RPB)); }
__global__ void __launch_bounds__(256) border_from_mask(const uint8_t* __restrict__ m, int H, int W, uint8_t* __restrict__ border) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, yb = blockIdx.y * RPB;
  if (x >= W) return;
  // rows yb - 1 .. yb + RPB of column x and of its two neighbours, loaded once
  uint8_t c[RPB + 2], l[RPB], r[RPB];
#pragma unroll
  for (int k = 0; k < RPB + 2; ++k) {
    const int y = yb - 1 + k;
    c[k] = (y >= 0 && y < H) ? m[(int64_t)y * W + x] : (uint8_t)0;
  }
#pragma unroll
  for (int k = 0; k < RPB; ++k) {
    const int y = yb + k;
    const bool in = y < H;
    l[k] = (in && x > 0) ? m[(int64_t)y * W + x - 1] : (uint8_t)0;
    r[k] = (in && x + 1 < W) ? m[(int64_t)y * W + x + 1] : (uint8_t)0;
  }
#pragma unroll
  for (int k = 0; k < RPB; ++k) {
    const int y = yb + k;
    if (y >= H) break;
    const bool fg = c[k + 1] != 0;
    const bool interior = fg && c[k] && c[k + 2] && l[k] && r[k];       // outside the image counts as background
    border[(int64_t)y * W + x] = fg && !interior;
  }
}
// one atomic per distinct key per warp: the pixels of a warp mostly belong to one or two components
__device__ __forceinline__ void warp_count(unsigned int* __restrict__ counters, long long key) {
  const unsigned int peers = __match_any_sync(0xffffffffu, key);
  if (key >= 0 && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) atomicAdd(&counters[key], (unsigned int)__popc(peers));
}
__global__ void __launch_bounds__(256) root_stats_kernel(const int* __restrict__ parent, const uint8_t* __restrict__ border,
                                                         int H, int W, unsigned int* __restrict__ area,
                                                         unsigned int* __restrict__ hist10) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  const bool in = x < W;
  int64_t i = (int64_t)y * W + (in ? x : 0);
  const int r = in ? parent[i] : -1;
  warp_count(area, (long long)r);
  long long hkey = -1;
  if (r >= 0 && border[i]) {
    int code = 0;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        int yy = y + dy, xx = x + dx;
        if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
        int wgt = (dy == 0 && dx == 0) ? 1 : ((dy == 0 || dx == 0) ? 2 : 10);
        code += border[(int64_t)yy * W + xx] ? wgt : 0;
      }
    const int slot = perim_slot(code);
    if (slot >= 0) hkey = (long long)r * 10 + slot;
  }
  warp_count(hist10, hkey);
}
// skimage: circularity = 4 pi area / (perimeter^2 + 1e-7), perimeter = sum of the 10 code counters x their weights
__device__ __forceinline__ bool root_is_round(const unsigned int* __restrict__ area, const unsigned int* __restrict__ hist10, int r) {
  const double w2 = 1.4142135623730951, w3 = (1.0 + 1.4142135623730951) / 2.0;
  const double wt[10] = {1.0, 1.0, w3, 1.0, 1.0, w2, w3, 1.0, 1.0, w2};
  double perim = 0.0;
#pragma unroll
  for (int c = 0; c < 10; ++c) perim = __dadd_rn(perim, __dmul_rn((double)hist10[(int64_t)r * 10 + c], wt[c]));
  const double circ = __ddiv_rn(__dmul_rn(__dmul_rn(4.0, 3.141592653589793), (double)area[r]),
                                __dadd_rn(__dmul_rn(perim, perim), 1e-7));
  return circ > 0.8;
}
// one thread per skeleton component (its root pixel): decide whether the mask component under it dies.  Every mask
// component carries a skeleton component (thinning never removes a component's last pixel), so the circularity test
// (transforms.py:346-351) is evaluated here, once per skeleton root, instead of in a pass of its own over all pixels.
__global__ void __launch_bounds__(256) root_kill_kernel(const int* __restrict__ sk_root, const uint8_t* __restrict__ has_fork,
                                                        const int* __restrict__ parent, const unsigned int* __restrict__ area,
                                                        const unsigned int* __restrict__ hist10, int64_t n, int remove_isolated,
                                                        uint8_t* __restrict__ kill) {
  const int64_t base = (int64_t)blockIdx.x * (256 * EPT) + threadIdx.x;
  int sr[EPT];
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; sr[k] = i < n ? sk_root[i] : -1; }
#pragma unroll
  for (int k = 0; k < EPT; ++k) {
    const int64_t i = base + k * 256;
    if (i >= n || sr[k] != (int)i) continue;
    const int r = parent[i];                    // the skeleton is a subset of the mask: r >= 0
    if ((remove_isolated && !has_fork[i]) || root_is_round(area, hist10, r)) kill[r] = 1;
  }
}
__global__ void __launch_bounds__(256) root_apply_kill(const int* __restrict__ parent, const uint8_t* __restrict__ kill, int64_t n,
                                                       uint8_t* __restrict__ out) {
  const int64_t base = (int64_t)blockIdx.x * (256 * EPT) + threadIdx.x;
  int r[EPT];
  uint8_t dead[EPT];
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; r[k] = i < n ? parent[i] : -1; }
#pragma unroll
  for (int k = 0; k < EPT; ++k) dead[k] = r[k] >= 0 ? kill[r[k]] : (uint8_t)1;
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; if (i < n) out[i] = !dead[k]; }
}

// ---- Zhang-Suen thinning (skimage.morphology.skeletonize, 2-D) -----------------------------------------
__constant__ uint8_t c_zs_lut[256];
__device__ __forceinline__ int nb_code(const uint8_t* __restrict__ m, int H, int W, int y, int x) {
  auto at = [&](int yy, int xx) -> int { return (yy >= 0 && yy < H && xx >= 0 && xx < W) ? (m[(int64_t)yy * W + xx] != 0) : 0; };
  return at(y - 1, x - 1) | (at(y - 1, x) << 1) | (at(y - 1, x + 1) << 2) | (at(y, x + 1) << 3) | (at(y + 1, x + 1) << 4) |
         (at(y + 1, x) << 5) | (at(y + 1, x - 1) << 6) | (at(y, x - 1) << 7);
}
__global__ void __launch_bounds__(256) thin_subpass(const uint8_t* __restrict__ in, int H, int W, int bit,
                                                    uint8_t* __restrict__ out, unsigned int* __restrict__ removed) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int64_t i = (int64_t)y * W + x;
  uint8_t v = in[i] != 0;
  if (v && (c_zs_lut[nb_code(in, H, W, y, x)] & bit)) {
    v = 0;
    atomicAdd(removed, 1u);
  }
  out[i] = v;
}
// ---- the whole thinning in ONE cooperative launch (W % 8 == 0) ----------------------------------------------------
// Persistent grid, 8 pixels (one 64-bit word) per work item, both sub-passes of every round separated by a grid-wide
// barrier, convergence decided on the device: no host round trip, no per-pass launch (round 1: 2 launches per round
// plus a device->host copy every 8 rounds, 19 us per pass on byte-wide accesses).  The image (4 MB at 2048^2) and its
// ping-pong copy stay L2-resident after the first pass.  A word without a set pixel is copied without looking at its
// neighbours (vessel masks: ~90 % of the words).  Co-residency of the grid comes from cudaLaunchCooperativeKernel.
__device__ __forceinline__ void grid_barrier(unsigned int* bar, unsigned int& goal) {
  __syncthreads();
  if (threadIdx.x == 0) {
    goal += gridDim.x;
    __threadfence();
    atomicAdd(bar, 1u);
    // release/acquire instead of the fences and polling without back-off all measured the same (0.112-0.114 ms per 2048^2 skeleton)
    while (*(volatile unsigned int*)bar < goal) __nanosleep(40);
    __threadfence();
  }
  __syncthreads();
}
// bits 1..8 = the word's eight pixels, bit 0 / bit 9 = the pixel left / right of the word
__device__ __forceinline__ uint32_t row_bits(const uint8_t* __restrict__ img, int W, int y, int xw, int H) {
  if (y < 0 || y >= H) return 0u;
  const uint8_t* row = img + (int64_t)y * W;
  const unsigned long long w = *reinterpret_cast<const unsigned long long*>(row + 8 * xw);
  // bytes are 0/1: gather their low bits (byte k -> bit k)
  uint32_t b = (uint32_t)(((w & 0x0101010101010101ull) * 0x0102040810204080ull) >> 56);
  b <<= 1;
  if (xw > 0) b |= row[8 * xw - 1] & 1u;
  if (8 * xw + 8 < W) b |= (uint32_t)(row[8 * xw + 8] & 1u) << 9;
  return b;
}
__device__ __forceinline__ unsigned int thin_pass_words(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W,
                                                        const uint8_t* __restrict__ s_lut, int bit) {
  const int wpr = W >> 3;
  const int nw = H * wpr;                       // < 2^31: the image itself is limited to 2^31 pixels
  unsigned int removed = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nw; i += gridDim.x * blockDim.x) {
    const int y = i / wpr, xw = i - y * wpr;
    unsigned long long c = *reinterpret_cast<const unsigned long long*>(in + (int64_t)y * W + 8 * xw);
    if (c != 0ull) {
      const uint32_t U = row_bits(in, W, y - 1, xw, H), C = row_bits(in, W, y, xw, H), D = row_bits(in, W, y + 1, xw, H);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if ((C >> (k + 1)) & 1u) {
          const uint32_t code = ((U >> k) & 7u) | (((C >> (k + 2)) & 1u) << 3) | (((D >> (k + 2)) & 1u) << 4) |
                                (((D >> (k + 1)) & 1u) << 5) | (((D >> k) & 1u) << 6) | (((C >> k) & 1u) << 7);
          if (s_lut[code] & bit) { c &= ~(0xffull << (8 * k)); ++removed; }
        }
      }
    }
    *reinterpret_cast<unsigned long long*>(out + (int64_t)y * W + 8 * xw) = c;
  }
  return removed;
}
// Thinning only removes pixels: a word that is zero stays zero.  The words with a set pixel are therefore listed ONCE
// (warp-aggregated append; the same scan copies the image into the ping-pong buffer so that both agree everywhere) and
// every sub-pass walks the list instead of the image -- a vessel mask has a set pixel in 10-30 % of its words, and the
// full-image sub-passes (read 4 MB, write 4 MB, 14 times) were 130 of the 316 us of filter_branch_seg_mask.
// src != nullptr: the caller's mask (any non-zero byte = set) is binarised on the way in and written to BOTH buffers, which
// saves the separate binarize pass; src and img may be the same array (every thread reads and writes its own word).
__device__ __forceinline__ void thin_build_list(const uint8_t* src, uint8_t* img, uint8_t* __restrict__ tmp, int nw,
                                                int* __restrict__ list, unsigned int* __restrict__ n_active) {
  const int lane = threadIdx.x & 31;
  const int stride = gridDim.x * blockDim.x;
  for (int i0 = blockIdx.x * blockDim.x; i0 < nw; i0 += stride) {          // block-uniform trip count
    const int i = i0 + threadIdx.x;
    unsigned long long c = 0ull;
    if (i < nw) {
      if (src) {
        c = reinterpret_cast<const unsigned long long*>(src)[i];
        // byte != 0 -> 1: the low seven bits carry into bit 7, OR the byte's own bit 7, keep that bit
        c = ((((c & 0x7f7f7f7f7f7f7f7full) + 0x7f7f7f7f7f7f7f7full) | c) >> 7) & 0x0101010101010101ull;
        reinterpret_cast<unsigned long long*>(img)[i] = c;
      } else {
        c = reinterpret_cast<const unsigned long long*>(img)[i];
      }
      reinterpret_cast<unsigned long long*>(tmp)[i] = c;
    }
    const unsigned int nz = __ballot_sync(0xffffffffu, c != 0ull);
    if (nz) {
      unsigned int base = 0;
      if (lane == 0) base = atomicAdd(n_active, (unsigned int)__popc(nz));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (c != 0ull) list[base + __popc(nz & ((1u << lane) - 1u))] = i;
    }
  }
}
// TLB listed words per thread and iteration, their loads issued together: a word's chain list -> word -> 3 neighbour rows
// is three dependent L2 round trips, and with one word at a time that latency, not the barrier, was the sub-pass.
constexpr int TLB = 4;
__device__ __forceinline__ unsigned int thin_pass_list(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W,
                                                       const uint8_t* __restrict__ s_lut, int bit, const int* __restrict__ list,
                                                       int n_list, int k_begin) {
  const int wpr = W >> 3;
  const int stride = gridDim.x * blockDim.x;
  unsigned int removed = 0;
  for (int k0 = k_begin + blockIdx.x * blockDim.x + threadIdx.x; k0 < n_list; k0 += TLB * stride) {
    int idx[TLB], yy[TLB], xx[TLB];
    unsigned long long c[TLB];
    uint32_t U[TLB], C[TLB], D[TLB];
#pragma unroll
    for (int b = 0; b < TLB; ++b) { const int k = k0 + b * stride; idx[b] = k < n_list ? list[k] : -1; }
#pragma unroll
    for (int b = 0; b < TLB; ++b) {
      c[b] = 0ull;
      if (idx[b] >= 0) {
        yy[b] = idx[b] / wpr; xx[b] = idx[b] - yy[b] * wpr;
        c[b] = reinterpret_cast<const unsigned long long*>(in)[idx[b]];
      }
    }
#pragma unroll
    for (int b = 0; b < TLB; ++b) {
      if (c[b] != 0ull) {
        U[b] = row_bits(in, W, yy[b] - 1, xx[b], H); C[b] = row_bits(in, W, yy[b], xx[b], H); D[b] = row_bits(in, W, yy[b] + 1, xx[b], H);
      }
    }
#pragma unroll
    for (int b = 0; b < TLB; ++b) {
      if (idx[b] < 0) continue;
      unsigned long long cw = c[b];
      if (cw != 0ull) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if ((C[b] >> (q + 1)) & 1u) {
            const uint32_t code = ((U[b] >> q) & 7u) | (((C[b] >> (q + 2)) & 1u) << 3) | (((D[b] >> (q + 2)) & 1u) << 4) |
                                  (((D[b] >> (q + 1)) & 1u) << 5) | (((D[b] >> q) & 1u) << 6) | (((C[b] >> q) & 1u) << 7);
            if (s_lut[code] & bit) { cw &= ~(0xffull << (8 * q)); ++removed; }
          }
        }
      }
      reinterpret_cast<unsigned long long*>(out)[idx[b]] = cw;      // also when it became zero: out holds an older state
    }
  }
  return removed;
}
// The first TLB listed words of a thread are ITS words in every sub-pass (static assignment), so their indices and their
// current values stay in registers across the whole kernel: a sub-pass then starts with the neighbour-row loads instead
// of the dependent chain list -> word -> neighbours (three L2 round trips -> one).
struct ThinOwn { int idx[TLB]; unsigned long long cur[TLB]; };
__device__ __forceinline__ unsigned int thin_pass_own(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int H, int W,
                                                      const uint8_t* __restrict__ s_lut, int bit, ThinOwn& o) {
  const int wpr = W >> 3;
  unsigned int removed = 0;
  uint32_t U[TLB], C[TLB], D[TLB];
#pragma unroll
  for (int b = 0; b < TLB; ++b) {
    if (o.idx[b] >= 0 && o.cur[b] != 0ull) {
      const int y = o.idx[b] / wpr, xw = o.idx[b] - y * wpr;
      U[b] = row_bits(in, W, y - 1, xw, H);
      D[b] = row_bits(in, W, y + 1, xw, H);
      const uint8_t* row = in + (int64_t)y * W;
      uint32_t c = (uint32_t)(((o.cur[b] & 0x0101010101010101ull) * 0x0102040810204080ull) >> 56) << 1;
      if (xw > 0) c |= row[8 * xw - 1] & 1u;
      if (8 * xw + 8 < W) c |= (uint32_t)(row[8 * xw + 8] & 1u) << 9;
      C[b] = c;
    }
  }
#pragma unroll
  for (int b = 0; b < TLB; ++b) {
    if (o.idx[b] < 0) continue;
    unsigned long long cw = o.cur[b];
    if (cw != 0ull) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if ((C[b] >> (q + 1)) & 1u) {
          const uint32_t code = ((U[b] >> q) & 7u) | (((C[b] >> (q + 2)) & 1u) << 3) | (((D[b] >> (q + 2)) & 1u) << 4) |
                                (((D[b] >> (q + 1)) & 1u) << 5) | (((D[b] >> q) & 1u) << 6) | (((C[b] >> q) & 1u) << 7);
          if (s_lut[code] & bit) { cw &= ~(0xffull << (8 * q)); ++removed; }
        }
      }
      o.cur[b] = cw;
    }
    reinterpret_cast<unsigned long long*>(out)[o.idx[b]] = cw;
  }
  return removed;
}
// THIN_THREADS threads per CTA, one CTA per SM by default: with the list a sub-pass is a few microseconds of work and the
// grid barrier (one atomic and one spinning thread per CTA) is what a round costs -- 148 participants instead of 592.
constexpr int THIN_THREADS = 512;
__global__ void __launch_bounds__(THIN_THREADS) thin_persistent(uint8_t* img, uint8_t* __restrict__ tmp, int H, int W,
                                                       const uint8_t* __restrict__ lut, unsigned int* __restrict__ counters,
                                                       int max_rounds, unsigned int* __restrict__ bar, int* __restrict__ iters_out,
                                                       int* __restrict__ list, unsigned int* __restrict__ n_active,
                                                       const uint8_t* src) {
  __shared__ uint8_t s_lut[256];
  __shared__ unsigned int s_removed;
  if (threadIdx.x < 256) s_lut[threadIdx.x] = lut[threadIdx.x];
  unsigned int goal = 0;
  int round = 0;
  const bool use_list = list != nullptr;
  int n_list = 0;
  ThinOwn own;
  const int stride = gridDim.x * blockDim.x;
  if (use_list) {
    thin_build_list(src, img, tmp, H * (W >> 3), list, n_active);
    grid_barrier(bar, goal);
    n_list = (int)*(volatile unsigned int*)n_active;
#pragma unroll
    for (int b = 0; b < TLB; ++b) {
      const int k = blockIdx.x * blockDim.x + threadIdx.x + b * stride;
      own.idx[b] = k < n_list ? list[k] : -1;
    }
#pragma unroll
    for (int b = 0; b < TLB; ++b) own.cur[b] = own.idx[b] >= 0 ? reinterpret_cast<const unsigned long long*>(img)[own.idx[b]] : 0ull;
  }
  for (; round < max_rounds; ++round) {
    if (threadIdx.x == 0) s_removed = 0;
    __syncthreads();
    unsigned int r;
    if (use_list) {
      r = thin_pass_own(img, tmp, H, W, s_lut, 1, own);
      if (n_list > TLB * stride) r += thin_pass_list(img, tmp, H, W, s_lut, 1, list, n_list, TLB * stride);
    } else {
      r = thin_pass_words(img, tmp, H, W, s_lut, 1);
    }
    grid_barrier(bar, goal);
    if (use_list) {
      r += thin_pass_own(tmp, img, H, W, s_lut, 2, own);
      if (n_list > TLB * stride) r += thin_pass_list(tmp, img, H, W, s_lut, 2, list, n_list, TLB * stride);
    } else {
      r += thin_pass_words(tmp, img, H, W, s_lut, 2);
    }
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    if ((threadIdx.x & 31) == 0 && r) atomicAdd(&s_removed, r);
    __syncthreads();
    if (threadIdx.x == 0 && s_removed) atomicAdd(&counters[round], s_removed);
    grid_barrier(bar, goal);
    if (*(volatile unsigned int*)&counters[round] == 0u) { ++round; break; }    // this round removed nothing: done
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && iters_out) *iters_out = round;
}

// ---- tiled variant: several rounds per grid barrier -----------------------------------------------------------------
// A block takes a 64 x 64 tile with an 8-pixel halo into shared memory and runs FOUR rounds (eight sub-passes) on it
// before anything goes back to global memory: a sub-pass looks one pixel around each pixel, so after k sub-passes
// everything further than k pixels from the halo's edge is still exact, and the 64 x 64 core is exact after all eight.
// Removals are counted per round for core pixels only; the grid meets once per four rounds instead of twice per round
// (7 rounds at 2048^2: 2 barriers instead of 14).  Rounds after convergence remove nothing, so running up to three of
// them too many changes neither the image nor the reported round count (= first round without a removal).  Any width:
// this is the path for widths that are not multiples of 8 (no host round trip either).  Measured on a 2048^2 vessel
// mask it is SLOWER than the word kernel above (0.20-0.33 ms against 0.17 ms): every pixel of every tile is visited
// in all eight sub-passes, while the word kernel skips an empty 8-pixel word with one 64-bit compare.
constexpr int TH_T = 64, TH_RPL = 4, TH_HL = 2 * TH_RPL, TH_S = TH_T + 2 * TH_HL, TH_P = TH_S + 2;   // +1 zero border
__device__ __forceinline__ unsigned int thin_tile_subpass(const uint8_t (*__restrict__ a)[TH_P], uint8_t (*__restrict__ b)[TH_P],
                                                          const uint8_t* __restrict__ s_lut, int bit, int cy0, int cy1, int cx0,
                                                          int cx1) {
  unsigned int removed = 0;
  for (int idx = threadIdx.x; idx < TH_S * TH_S; idx += blockDim.x) {
    const int y = idx / TH_S + 1, x = idx - (y - 1) * TH_S + 1;
    uint8_t v = a[y][x];
    if (v) {
      const int code = a[y - 1][x - 1] | (a[y - 1][x] << 1) | (a[y - 1][x + 1] << 2) | (a[y][x + 1] << 3) |
                       (a[y + 1][x + 1] << 4) | (a[y + 1][x] << 5) | (a[y + 1][x - 1] << 6) | (a[y][x - 1] << 7);
      if (s_lut[code] & bit) {
        v = 0;
        removed += (y >= cy0 && y < cy1 && x >= cx0 && x < cx1);
      }
    }
    b[y][x] = v;
  }
  return removed;
}
__global__ void __launch_bounds__(256) thin_tiled_persistent(uint8_t* __restrict__ img, uint8_t* __restrict__ tmp, int H, int W,
                                                             const uint8_t* __restrict__ lut, unsigned int* __restrict__ counters,
                                                             int max_rounds, unsigned int* __restrict__ bar,
                                                             int* __restrict__ iters_out) {
  __shared__ uint8_t s_a[TH_P][TH_P], s_b[TH_P][TH_P];
  __shared__ uint8_t s_lut[256];
  __shared__ unsigned int s_cnt[TH_RPL];
  __shared__ int s_any;
  s_lut[threadIdx.x] = lut[threadIdx.x];
  for (int i = threadIdx.x; i < TH_P * TH_P; i += blockDim.x) { (&s_a[0][0])[i] = 0; (&s_b[0][0])[i] = 0; }
  const int ntx = (W + TH_T - 1) / TH_T, nty = (H + TH_T - 1) / TH_T;
  uint8_t *src = img, *dst = tmp;
  unsigned int goal = 0;
  int iters = max_rounds;
  for (int sr = 0; sr * TH_RPL < max_rounds; ++sr) {
    for (int t = blockIdx.x; t < ntx * nty; t += gridDim.x) {
      const int ty = t / ntx, tx = t - ty * ntx;
      const int gy0 = ty * TH_T - TH_HL, gx0 = tx * TH_T - TH_HL;
      __syncthreads();                                  // the previous tile's write-back has read s_a
      if (threadIdx.x == 0) s_any = 0;
      if (threadIdx.x < TH_RPL) s_cnt[threadIdx.x] = 0;
      __syncthreads();
      int any = 0;
      for (int idx = threadIdx.x; idx < TH_S * TH_S; idx += blockDim.x) {
        const int y = idx / TH_S, x = idx - y * TH_S;
        const int gy = gy0 + y, gx = gx0 + x;
        const uint8_t v = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? src[(int64_t)gy * W + gx] : (uint8_t)0;
        s_a[y + 1][x + 1] = v;
        any |= v;
      }
      if (any) s_any = 1;
      __syncthreads();
      // core of this tile in shared-memory coordinates, clipped to the image
      const int cy0 = TH_HL + 1, cx0 = TH_HL + 1;
      const int cy1 = cy0 + min(TH_T, H - ty * TH_T), cx1 = cx0 + min(TH_T, W - tx * TH_T);
      if (s_any) {
        for (int r = 0; r < TH_RPL; ++r) {
          unsigned int rem = thin_tile_subpass(s_a, s_b, s_lut, 1, cy0, cy1, cx0, cx1);
          __syncthreads();
          rem += thin_tile_subpass(s_b, s_a, s_lut, 2, cy0, cy1, cx0, cx1);
          for (int o = 16; o > 0; o >>= 1) rem += __shfl_xor_sync(0xffffffffu, rem, o);
          if ((threadIdx.x & 31) == 0 && rem) atomicAdd(&s_cnt[r], rem);
          __syncthreads();
        }
      }
      for (int idx = threadIdx.x; idx < TH_T * TH_T; idx += blockDim.x) {
        const int y = idx / TH_T, x = idx - y * TH_T;
        const int gy = ty * TH_T + y, gx = tx * TH_T + x;
        if (gy < H && gx < W) dst[(int64_t)gy * W + gx] = s_a[cy0 + y][cx0 + x];
      }
      if (threadIdx.x < TH_RPL && s_cnt[threadIdx.x]) atomicAdd(&counters[sr * TH_RPL + threadIdx.x], s_cnt[threadIdx.x]);
    }
    grid_barrier(bar, goal);
    uint8_t* t2 = src; src = dst; dst = t2;
    bool done = false;
    for (int r = 0; r < TH_RPL && !done; ++r) {
      if (*(volatile unsigned int*)&counters[sr * TH_RPL + r] == 0u) { iters = sr * TH_RPL + r + 1; done = true; }
    }
    if (done) break;
  }
  if (src != img) {                                     // the result must end up in `img`
    const int64_t n = (int64_t)H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) img[i] = src[i];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && iters_out) *iters_out = iters;
}

// degree > 2 in the 8-neighbour pixel graph (transforms.py:248, :334)
__global__ void __launch_bounds__(256) fork_kernel(const uint8_t* __restrict__ sk, int H, int W, const int* __restrict__ sk_root,
                                                   uint8_t* __restrict__ has_fork /*indexed by root pixel*/) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int64_t i = (int64_t)y * W + x;
  if (!sk[i]) return;
  if (__popc(nb_code(sk, H, W, y, x)) > 2) has_fork[sk_root[i]] = 1;
}
__global__ void __launch_bounds__(256) binarize(const uint8_t* __restrict__ m, int64_t n, uint8_t* __restrict__ out) {
  const int64_t base = (int64_t)blockIdx.x * (256 * EPT) + threadIdx.x;
  uint8_t v[EPT];
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; v[k] = i < n ? m[i] : (uint8_t)0; }
#pragma unroll
  for (int k = 0; k < EPT; ++k) { const int64_t i = base + k * 256; if (i < n) out[i] = v[k] != 0; }
}

// ---- the skeleton's components, forks and the kill decision over the thinning's word list ------------------------
// The skeleton has a set pixel in a few per cent of the image's words, all of them on the list the thinning kernel
// built.  Labelling it with the dense kernels (three passes over every pixel), marking forks and deciding the kills
// cost ~85 us per 2048^2 image; here one thread takes one listed word (8 pixels), sk_root is preset to -1 by a memset.
// Same union rules as ccl_merge, with runs that start at word boundaries.
__global__ void __launch_bounds__(256) sk_init_list(const uint8_t* __restrict__ sk, int H, int W, const int* __restrict__ list,
                                                    const unsigned int* __restrict__ n_list, int* __restrict__ sk_root,
                                                    uint8_t* __restrict__ has_fork) {
  const int wpr = W >> 3, n = (int)*n_list;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    const int i = list[k];
    const int y = i / wpr, xw = i - y * wpr;
    const unsigned long long c = *reinterpret_cast<const unsigned long long*>(sk + (int64_t)y * W + 8 * xw);
    if (c == 0ull) continue;
    const uint32_t bits = (uint32_t)(((c & 0x0101010101010101ull) * 0x0102040810204080ull) >> 56);
    const int base = y * W + 8 * xw;
    int start = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      if ((bits >> q) & 1u) { sk_root[base + q] = base + start; has_fork[base + q] = 0; }
      else start = q + 1;
    }
  }
}
// One thread per PIXEL of a listed word from here on (8 consecutive threads share a word): a union or a find is a chain
// of dependent global accesses, and a thread that walked its word's pixels one after the other made the merge pass 35 us
// for ~1 % of the image's pixels.
__global__ void __launch_bounds__(256) sk_merge_list(const uint8_t* __restrict__ sk, int H, int W, const int* __restrict__ list,
                                                     const unsigned int* __restrict__ n_list, int* __restrict__ sk_root) {
  const int wpr = W >> 3;
  const int64_t n = (int64_t)*n_list * 8;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
    const int wi = list[t >> 3], q = (int)(t & 7) + 1;
    const int y = wi / wpr, xw = wi - y * wpr;
    const uint32_t C = row_bits(sk, W, y, xw, H);
    if (!((C >> q) & 1u)) continue;
    const uint32_t U = row_bits(sk, W, y - 1, xw, H);
    const int i = y * W + 8 * xw + q - 1;
    const bool w = (C >> (q - 1)) & 1u;
    if (w && q == 1) uf_union(sk_root, i, i - 1);                 // the run continues from the previous word
    const bool n_ = (U >> q) & 1u, ne = (U >> (q + 1)) & 1u, nw = (U >> (q - 1)) & 1u;
    if (w) {
      if (ne && !n_) uf_union(sk_root, i, i - W + 1);
    } else if (n_) {
      uf_union(sk_root, i, i - W);
    } else {
      if (nw) uf_union(sk_root, i, i - W - 1);
      if (ne) uf_union(sk_root, i, i - W + 1);
    }
  }
}
// flatten + fork marking: degree > 2 in the 8-neighbour pixel graph (transforms.py:248, :334)
__global__ void __launch_bounds__(256) sk_flatten_fork_list(const uint8_t* __restrict__ sk, int H, int W, const int* __restrict__ list,
                                                            const unsigned int* __restrict__ n_list, int* __restrict__ sk_root,
                                                            uint8_t* __restrict__ has_fork) {
  const int wpr = W >> 3;
  const int64_t n = (int64_t)*n_list * 8;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
    const int wi = list[t >> 3], q = (int)(t & 7) + 1;
    const int y = wi / wpr, xw = wi - y * wpr;
    const uint32_t C = row_bits(sk, W, y, xw, H);
    if (!((C >> q) & 1u)) continue;
    const uint32_t U = row_bits(sk, W, y - 1, xw, H), D = row_bits(sk, W, y + 1, xw, H);
    const int i = y * W + 8 * xw + q - 1;
    const int r = uf_find(sk_root, i);
    sk_root[i] = r;
    const int deg = __popc((U >> (q - 1)) & 7u) + __popc((D >> (q - 1)) & 7u) + (int)((C >> (q - 1)) & 1u) + (int)((C >> (q + 1)) & 1u);
    if (deg > 2) has_fork[r] = 1;
  }
}
__global__ void __launch_bounds__(256) sk_kill_list(const uint8_t* __restrict__ sk, int W, const int* __restrict__ list,
                                                    const unsigned int* __restrict__ n_list, const int* __restrict__ sk_root,
                                                    const uint8_t* __restrict__ has_fork, const int* __restrict__ parent,
                                                    const unsigned int* __restrict__ area, const unsigned int* __restrict__ hist10,
                                                    int remove_isolated, uint8_t* __restrict__ kill) {
  const int wpr = W >> 3;
  const int64_t n = (int64_t)*n_list * 8;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
    const int wi = list[t >> 3];
    const int y = wi / wpr, xw = wi - y * wpr;
    const int i = y * W + 8 * xw + (int)(t & 7);
    if (!sk[i] || sk_root[i] != i) continue;
    const int r = parent[i];
    if ((remove_isolated && !has_fork[i]) || root_is_round(area, hist10, r)) kill[r] = 1;
  }
}

// ---- per-root areas (no renumbering needed) ------------------------------------------------------------
__global__ void __launch_bounds__(256) root_area(const int* __restrict__ parent, int64_t n, unsigned int* __restrict__ area) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && parent[i] >= 0) atomicAdd(&area[parent[i]], 1u);
}
// remove_small_islands step: pixels whose component (of the tested polarity) is smaller than min_size flip
__global__ void __launch_bounds__(256) flip_small(const int* __restrict__ parent, const unsigned int* __restrict__ area,
                                                  int64_t n, unsigned int min_size, uint8_t set_to, uint8_t* __restrict__ mask) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && parent[i] >= 0 && area[parent[i]] < min_size) mask[i] = set_to;
}
__global__ void __launch_bounds__(256) area_image_kernel(const int* __restrict__ parent, const unsigned int* __restrict__ area,
                                                         int64_t n, int prop, double* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double a = parent[i] >= 0 ? (double)area[parent[i]] : 0.0;
  // skimage regionprops: equivalent_diameter_area = sqrt(4 * area / pi), IEEE operations in that order
  if (prop == 1) a = __dsqrt_rn(__ddiv_rn(4.0 * a, 3.141592653589793));
  out[i] = a;
}

// ---- regionprops 'eccentricity' (compute_branches.py:276) ---------------------------------------------------------
// skimage: eigenvalues l1 >= l2 of the inertia tensor [[mu02, -mu11], [-mu11, mu20]] / mu00 of the region's pixels,
// eccentricity = 0 if l1 == 0 else sqrt(1 - l2 / l1).  The raw moments are integers: they are accumulated exactly
// (64-bit atomics, one per distinct component per warp) and the central moments scaled by mu00^2,
//   A = n Sxx - Sx^2,  C = n Syy - Sy^2,  B = n Sxy - Sx Sy,
// are formed exactly in 128-bit integers before anything is rounded -- skimage's float64 dot products carry ~1e-13
// relative noise, so parity with it is a tolerance (and unpinned here: scikit-image is not installable).
struct RootMoments { unsigned long long sx, sy, sxx, syy, sxy; };
__global__ void __launch_bounds__(256) root_moments_kernel(const int* __restrict__ parent, int H, int W, unsigned int* __restrict__ area,
                                                           RootMoments* __restrict__ mom) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  const bool in = x < W;
  const int r = in ? parent[(int64_t)y * W + x] : -1;
  const unsigned int peers = __match_any_sync(0xffffffffu, r);
  const int lane = threadIdx.x & 31;
  if (r < 0 || lane != __ffs(peers) - 1) return;
  // all lanes of a warp share y; lane l holds column x - lane + l
  unsigned long long cnt = 0, sx = 0, sxx = 0;
  for (unsigned int m = peers; m; m &= m - 1) {
    const unsigned long long xx = (unsigned long long)(x - lane + (__ffs(m) - 1));
    cnt += 1; sx += xx; sxx += xx * xx;
  }
  const unsigned long long yy = (unsigned long long)y;
  atomicAdd(&area[r], (unsigned int)cnt);
  atomicAdd(&mom[r].sx, sx);
  atomicAdd(&mom[r].sy, yy * cnt);
  atomicAdd(&mom[r].sxx, sxx);
  atomicAdd(&mom[r].syy, yy * yy * cnt);
  atomicAdd(&mom[r].sxy, yy * sx);
}
__device__ __forceinline__ double i128_to_double(__int128 v) {
  const bool neg = v < 0;
  const unsigned __int128 u = neg ? (unsigned __int128)(-v) : (unsigned __int128)v;
  const double d = __dadd_rn(__dmul_rn((double)(unsigned long long)(u >> 64), 18446744073709551616.0), (double)(unsigned long long)u);
  return neg ? -d : d;
}
__global__ void __launch_bounds__(256) root_eccentricity_kernel(const int* __restrict__ parent, int64_t n, const unsigned int* __restrict__ area,
                                                                const RootMoments* __restrict__ mom, double* __restrict__ val) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || parent[i] != (int)i) return;
  const RootMoments m = mom[i];
  const __int128 cnt = (__int128)area[i];
  const double A = i128_to_double(cnt * (__int128)m.sxx - (__int128)m.sx * (__int128)m.sx);     // columns
  const double C = i128_to_double(cnt * (__int128)m.syy - (__int128)m.sy * (__int128)m.sy);     // rows
  const double B = i128_to_double(cnt * (__int128)m.sxy - (__int128)m.sx * (__int128)m.sy);
  // l1 = (A + C) / 2 + sqrt(((A - C) / 2)^2 + B^2); l2 = det / l1 (no cancellation for elongated regions)
  const double half = 0.5 * (A + C), hd = 0.5 * (A - C);
  const double l1 = half + sqrt(hd * hd + B * B);
  double ecc = 0.0;
  if (l1 > 0.0) {
    const double p = A * C, pe = fma(A, C, -p), q = B * B, qe = fma(B, B, -q);
    double det = (p - q) + (pe - qe);
    det = det > 0.0 ? det : 0.0;
    double ratio = det / l1 / l1;
    ratio = ratio < 1.0 ? ratio : 1.0;
    ecc = sqrt(1.0 - ratio);
  }
  val[i] = ecc;
}
__global__ void __launch_bounds__(256) paint_root_value(const int* __restrict__ parent, const double* __restrict__ val, int64_t n,
                                                        double* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = parent[i] >= 0 ? val[parent[i]] : 0.0;
}

}  // namespace

// ---- host side --------------------------------------------------------------------------------------
// The deletability table is data, not code: csrc/skeleton_lut.inc (shared with oracle/morphology.py; derived from the
// Zhang-Suen conditions, parity with scikit-image's shipped table UNPINNED -- see the file).  tmat_skeleton_lut_set
// replaces it process-wide at run time.
static uint8_t g_zs_lut[256] = {
#include "skeleton_lut.inc"
};
static const uint8_t* zs_lut_host() { return g_zs_lut; }
extern "C" int tmat_skeleton_lut_set(const uint8_t* lut256) {
  TMAT_REQUIRE(lut256, "tmat_skeleton_lut_set: NULL table");
  for (int v = 0; v < 256; ++v) TMAT_REQUIRE(lut256[v] <= 3, "tmat_skeleton_lut_set: entry %d = %d not in 0..3", v, lut256[v]);
  memcpy(g_zs_lut, lut256, 256);
  return TMAT_OK;
}
extern "C" int tmat_skeleton_lut_get(uint8_t* lut256) {
  TMAT_REQUIRE(lut256, "tmat_skeleton_lut_get: NULL output");
  memcpy(lut256, g_zs_lut, 256);
  return TMAT_OK;
}

static inline dim3 grid2d(int H, int W) { return dim3(cdiv(W, 256), H); }
static inline unsigned grid1d(int64_t n) { return (unsigned)cdiv64(n, 256); }
static inline unsigned grid1d_ept(int64_t n) { return (unsigned)cdiv64(n, 256 * EPT); }

int erode_disk_thr(tmat_ctx* ctx, const uint8_t* in, int H, int W, int r, int border_value, const int* thr_params_dev,
                   uint8_t* out, cudaStream_t st) {
  TMAT_REQUIRE(r >= 0 && r <= ER_MAX_R, "erode_disk: radius %d outside [0,%d]", r, ER_MAX_R);
  // rows per warp: more, shorter segments hide the per-row dependency chain (ballot -> shifts -> accumulators) better
  // than the halo rows cost (measured at r = 5 on 2048^2, TMAT_ERODE_SEG)
  static const int seg = [] { const char* v = getenv("TMAT_ERODE_SEG"); int s = v ? atoi(v) : 32; return s == 16 || s == 32 ? s : 64; }();
  static const bool ballot_only = [] { const char* v = getenv("TMAT_MORPH_BALLOT"); return v && atoi(v) == 1; }();
  const bool words_ok = !ballot_only && (W % 8) == 0 && (reinterpret_cast<uintptr_t>(in) % 8) == 0 && (reinterpret_cast<uintptr_t>(out) % 8) == 0;
  if (r == 5 && words_ok) {
    erode_disk_words<5><<<dim3(cdiv(W / 8, 32), cdiv(cdiv(H, EW_SEG), 4)), 128, 0, st>>>(in, H, W, border_value, thr_params_dev, out);
    KERNEL_CHECK(ctx);
    return TMAT_OK;
  }
  if (r == 5 && seg != ER_SEG) {
    dim3 g2(cdiv(W, 32), cdiv(cdiv(H, seg), 4));
    if (seg == 32) erode_disk_bits<5, 32><<<g2, 128, 0, st>>>(in, H, W, border_value, thr_params_dev, out);
    else erode_disk_bits<5, 16><<<g2, 128, 0, st>>>(in, H, W, border_value, thr_params_dev, out);
    KERNEL_CHECK(ctx);
    return TMAT_OK;
  }
  dim3 grid(cdiv(W, 32), cdiv(cdiv(H, ER_SEG), 4));
  switch (r) {
#define ER_CASE(R) case R: erode_disk_bits<R><<<grid, 128, 0, st>>>(in, H, W, border_value, thr_params_dev, out); break;
    ER_CASE(0) ER_CASE(1) ER_CASE(2) ER_CASE(3) ER_CASE(4) ER_CASE(5) ER_CASE(6) ER_CASE(7) ER_CASE(8)
#undef ER_CASE
  }
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}
int erode_disk(tmat_ctx* ctx, const uint8_t* in, int H, int W, int r, int border_value, uint8_t* out, cudaStream_t st) {
  return erode_disk_thr(ctx, in, H, W, r, border_value, nullptr, out, st);
}

int median_disk2(tmat_ctx* ctx, const uint8_t* in, int H, int W, uint8_t* out, cudaStream_t st) {
  static const bool per_pixel = [] { const char* v = getenv("TMAT_MEDIAN_PER_PIXEL"); return v && atoi(v) == 1; }();
  static const bool ballot_only = [] { const char* v = getenv("TMAT_MORPH_BALLOT"); return v && atoi(v) == 1; }();
  const bool words_ok = !ballot_only && (W % 8) == 0 && (reinterpret_cast<uintptr_t>(in) % 8) == 0 && (reinterpret_cast<uintptr_t>(out) % 8) == 0;
  if (per_pixel) median_disk2_kernel<<<grid2d(H, W), 256, 0, st>>>(in, H, W, out);
  else if (words_ok) median_disk2_words<<<dim3(cdiv(W / 8, 32), cdiv(cdiv(H, MW_SEG), 4)), 128, 0, st>>>(in, H, W, out);
  else median_disk2_bits<<<dim3(cdiv(W, 32), cdiv(cdiv(H, MED_SEG), 4)), 128, 0, st>>>(in, H, W, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

// op 0: median of a bool image (element of rank nfp/2 of the sorted neighbourhood is 1 iff at least nfp - nfp/2
// members are set); op 1: erosion (all members set).
int footprint_filter(tmat_ctx* ctx, const uint8_t* in, int H, int W, const uint8_t* fp_host, int fh, int fw, int op,
                     int border_value, uint8_t* out, cudaStream_t st) {
  TMAT_REQUIRE(fp_host && fh >= 1 && fw >= 1 && (fh & 1) && (fw & 1) && fh <= 255 && fw <= 255,
               "footprint_filter: footprint must be odd-sized, got %dx%d", fh, fw);
  TMAT_REQUIRE(op == 0 || op == 1, "footprint_filter: op %d", op);
  std::vector<short2> offs;
  for (int r = 0; r < fh; ++r)
    for (int c = 0; c < fw; ++c)
      if (fp_host[r * fw + c]) offs.push_back(make_short2((short)(r - fh / 2), (short)(c - fw / 2)));
  const int nfp = (int)offs.size();
  TMAT_REQUIRE(nfp >= 1, "footprint_filter: empty footprint");
  short2* d_offs;
  TMAT_TRY(arena_get(ctx, (size_t)nfp, &d_offs));
  // the offsets are small: staged through the context's pinned buffer so the copy is stream-ordered
  TMAT_REQUIRE((size_t)nfp * sizeof(short2) <= 4096, "footprint_filter: footprint with %d members is too large", nfp);
  CUDA_TRY(cudaStreamSynchronize(st));
  memcpy(ctx->pinned, offs.data(), (size_t)nfp * sizeof(short2));
  CUDA_TRY(cudaMemcpyAsync(d_offs, ctx->pinned, (size_t)nfp * sizeof(short2), cudaMemcpyHostToDevice, st));
  const int need = op == 0 ? nfp - nfp / 2 : nfp;
  footprint_count_kernel<<<grid2d(H, W), 256, 0, st>>>(in, H, W, d_offs, nfp, op == 0, border_value, need, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

// parent[i] = index of the first pixel (raster order) of i's component, -1 for background
static int ccl_roots(tmat_ctx* ctx, const uint8_t* mask, int H, int W, int connectivity, int fg_is_zero, int* parent,
                     cudaStream_t st, unsigned int* init_area = nullptr, unsigned int* init_hist10 = nullptr,
                     uint8_t* init_kill = nullptr) {
  TMAT_REQUIRE(connectivity == 1 || connectivity == 2, "label: connectivity must be 1 or 2, got %d", connectivity);
  TMAT_REQUIRE((int64_t)H * W < (1ll << 31), "label: image too large");
  int64_t n = (int64_t)H * W;
  ccl_init<<<dim3(cdiv(W, 256), cdiv(H, CCL_RPB)), 256, 0, st>>>(mask, H, W, fg_is_zero, parent);
  KERNEL_CHECK(ctx);
  ccl_merge<<<dim3(cdiv(W, 256), cdiv(H, MRG_RPB)), 256, 0, st>>>(mask, fg_is_zero, parent, H, W, connectivity == 2);
  KERNEL_CHECK(ctx);
  if (init_area) ccl_flatten_init<<<grid1d_ept(n), 256, 0, st>>>(parent, n, init_area, init_hist10, init_kill);
  else ccl_flatten<<<grid1d_ept(n), 256, 0, st>>>(parent, n);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

// labels 1..n in raster order of each component's first pixel; synchronises to return n
static int label_impl(tmat_ctx* ctx, const uint8_t* mask, int H, int W, int connectivity, int32_t* labels, int* n_labels,
                      int** parent_out, cudaStream_t st) {
  int64_t n = (int64_t)H * W;
  int nb = (int)cdiv64(n, SCAN_BLOCK);
  int *parent, *rank, *bsum, *total;
  TMAT_TRY(arena_get(ctx, (size_t)n, &parent));
  TMAT_TRY(arena_get(ctx, (size_t)n, &rank));
  TMAT_TRY(arena_get(ctx, (size_t)nb + 1, &bsum));
  TMAT_TRY(arena_get(ctx, 1, &total));
  TMAT_TRY(ccl_roots(ctx, mask, H, W, connectivity, 0, parent, st));
  roots_count<<<nb, 256, 0, st>>>(parent, n, bsum);
  KERNEL_CHECK(ctx);
  scan_block_sums<<<1, 1024, 0, st>>>(bsum, nb, total);
  KERNEL_CHECK(ctx);
  roots_rank<<<nb, 256, 0, st>>>(parent, n, bsum, rank);
  KERNEL_CHECK(ctx);
  labels_from_rank<<<grid1d(n), 256, 0, st>>>(parent, rank, n, labels);
  KERNEL_CHECK(ctx);
  if (parent_out) *parent_out = parent;
  int* h = (int*)ctx->pinned;
  CUDA_TRY(cudaMemcpyAsync(h, total, sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  *n_labels = *h;
  return TMAT_OK;
}

// iterations (optional, host): rounds executed -- asking for it costs a stream synchronisation on the cooperative path.
static int skeletonize_impl(tmat_ctx* ctx, const uint8_t* mask, int H, int W, uint8_t* out, int* iterations,
                            cudaStream_t st, int** list_out = nullptr, unsigned int** n_list_out = nullptr) {
  if (list_out) *list_out = nullptr;
  int64_t n = (int64_t)H * W;
  uint8_t* tmp;
  TMAT_TRY(arena_get(ctx, (size_t)n, &tmp));
  static const bool no_coop = [] { const char* e = getenv("TMAT_THIN_MULTILAUNCH"); return e && atoi(e) != 0; }();
  static const int variant = [] { const char* e = getenv("TMAT_THIN_VARIANT"); return e ? atoi(e) : 1; }();   // 1: 8-pixel words (default where W % 8 == 0), 2: tiles
  static const bool full_scan = [] { const char* e = getenv("TMAT_THIN_FULLSCAN"); return e && atoi(e) == 1; }();
  const bool words_ok = (W % 8) == 0 && (reinterpret_cast<uintptr_t>(out) % 8) == 0 && (reinterpret_cast<uintptr_t>(tmp) % 8) == 0;
  // the list-building scan of the word kernel binarises the mask itself when it can read it in 8-byte words
  const bool fuse_bin = !no_coop && variant != 2 && words_ok && !full_scan && (reinterpret_cast<uintptr_t>(mask) % 8) == 0;
  if (!fuse_bin) {
    binarize<<<grid1d_ept(n), 256, 0, st>>>(mask, n, out);
    KERNEL_CHECK(ctx);
  }
  if (!no_coop && (variant == 2 || words_ok)) {
    const bool tiled = variant == 2 || !words_ok;
    int max_rounds = std::min(H, W) / 2 + 2;               // a round that removes something peels one layer
    if (tiled) max_rounds = (max_rounds + TH_RPL - 1) / TH_RPL * TH_RPL;
    uint8_t* d_lut;
    unsigned int* counters;
    int* d_iters;
    TMAT_TRY(arena_get(ctx, 256, &d_lut));
    TMAT_TRY(arena_get(ctx, (size_t)max_rounds + 3, &counters));      // [max_rounds] counters, the barrier word, the list length
    TMAT_TRY(arena_get(ctx, 1, &d_iters));
    // the table goes through the context's pinned page so that the copy is stream-ordered and the call stays asynchronous
    uint8_t* stage = (uint8_t*)ctx->pinned + 8192 + 256 * (ctx->lut_slot++ & 7);
    memcpy(stage, zs_lut_host(), 256);
    CUDA_TRY(cudaMemcpyAsync(d_lut, stage, 256, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemsetAsync(counters, 0, ((size_t)max_rounds + 3) * sizeof(unsigned int), st));
    if (tiled) {
      static int tiled_bps = 0;
      if (!tiled_bps) {
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tiled_bps, thin_tiled_persistent, 256, 0));
        const char* e = getenv("TMAT_THIN_BLOCKS");
        const int cap_b = e && atoi(e) > 0 ? atoi(e) : 4;
        tiled_bps = std::max(1, std::min(tiled_bps, cap_b));
      }
      const int ntiles = cdiv(W, TH_T) * cdiv(H, TH_T);
      int grid = std::max(1, std::min(ctx->sm_count * tiled_bps, ntiles));
      unsigned int* bar = counters + max_rounds + 1;
      int Hh = H, Ww = W, mr = max_rounds;
      const uint8_t* lut_c = d_lut;
      void* args[] = {(void*)&out, (void*)&tmp, (void*)&Hh, (void*)&Ww, (void*)&lut_c, (void*)&counters, (void*)&mr, (void*)&bar,
                      (void*)&d_iters};
      CUDA_TRY(cudaLaunchCooperativeKernel((const void*)thin_tiled_persistent, dim3(grid), dim3(256), args, 0, st));
      ctx->launches++;
      if (iterations) {
        int* h = (int*)ctx->pinned;
        CUDA_TRY(cudaMemcpyAsync(h, d_iters, sizeof(int), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        *iterations = *h;
      }
      return TMAT_OK;
    }
    static int blocks_per_sm = 0;
    if (!blocks_per_sm) {
      CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, thin_persistent, THIN_THREADS, 0));
      const char* e = getenv("TMAT_THIN_BLOCKS");
      const int cap_b = e && atoi(e) > 0 ? atoi(e) : 1;
      if (blocks_per_sm > cap_b) blocks_per_sm = cap_b;      // every block is a participant of the grid barrier
      if (blocks_per_sm < 1) blocks_per_sm = 1;
    }
    const int64_t words = n / 8;
    int grid = (int)std::min<int64_t>((int64_t)ctx->sm_count * blocks_per_sm, cdiv64(words, THIN_THREADS));
    if (grid < 1) grid = 1;
    unsigned int* bar = counters + max_rounds + 1;
    unsigned int* n_active = counters + max_rounds + 2;
    int* list = nullptr;
    if (!full_scan) TMAT_TRY(arena_get(ctx, (size_t)words, &list));
    if (list_out) { *list_out = list; *n_list_out = n_active; }
    int Hh = H, Ww = W, mr = max_rounds;
    const uint8_t* lut_c = d_lut;
    const uint8_t* src = fuse_bin ? mask : nullptr;
    void* args[] = {(void*)&out, (void*)&tmp, (void*)&Hh, (void*)&Ww, (void*)&lut_c, (void*)&counters, (void*)&mr, (void*)&bar,
                    (void*)&d_iters, (void*)&list, (void*)&n_active, (void*)&src};
    CUDA_TRY(cudaLaunchCooperativeKernel((const void*)thin_persistent, dim3(grid), dim3(THIN_THREADS), args, 0, st));
    ctx->launches++;
    if (iterations) {
      int* h = (int*)ctx->pinned;
      CUDA_TRY(cudaMemcpyAsync(h, d_iters, sizeof(int), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaStreamSynchronize(st));
      *iterations = *h;
    }
    return TMAT_OK;
  }
  // any width: one launch per sub-pass, convergence counters read back every 8 rounds
  CUDA_TRY(cudaMemcpyToSymbolAsync(c_zs_lut, zs_lut_host(), 256, 0, cudaMemcpyHostToDevice, st));
  unsigned int* counters;
  constexpr int BATCH = 8;
  TMAT_TRY(arena_get(ctx, BATCH, &counters));
  unsigned int* h = (unsigned int*)ctx->pinned;
  int iters = 0;
  for (;;) {
    CUDA_TRY(cudaMemsetAsync(counters, 0, BATCH * sizeof(unsigned int), st));
    for (int k = 0; k < BATCH; ++k) {
      thin_subpass<<<grid2d(H, W), 256, 0, st>>>(out, H, W, 1, tmp, counters + k);
      KERNEL_CHECK(ctx);
      thin_subpass<<<grid2d(H, W), 256, 0, st>>>(tmp, H, W, 2, out, counters + k);
      KERNEL_CHECK(ctx);
    }
    CUDA_TRY(cudaMemcpyAsync(h, counters, BATCH * sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    int k = 0;
    while (k < BATCH && h[k] != 0) ++k;
    if (k < BATCH) { iters += k + 1; break; }   // round k removed nothing: the reference stops there
    iters += BATCH;
  }
  if (iterations) *iterations = iters;
  return TMAT_OK;
}

// transforms.py:306-361.  `mask` and `out` may alias.
int filter_branch_impl(tmat_ctx* ctx, const uint8_t* mask, int H, int W, int use_median, int remove_isolated,
                       uint8_t* out, cudaStream_t st) {
  TMAT_REQUIRE((int64_t)H * W < (1ll << 31), "filter_branch_seg_mask: image too large");
  int64_t n = (int64_t)H * W;
  uint8_t *m, *sk, *border, *kill, *has_fork;
  int *parent, *sk_root;
  unsigned int *area, *hist;
  TMAT_TRY(arena_get(ctx, (size_t)n, &m));
  TMAT_TRY(arena_get(ctx, (size_t)n, &sk));
  TMAT_TRY(arena_get(ctx, (size_t)n, &border));
  TMAT_TRY(arena_get(ctx, (size_t)n, &parent));
  TMAT_TRY(arena_get(ctx, (size_t)n, &sk_root));
  TMAT_TRY(arena_get(ctx, (size_t)n, &area));
  TMAT_TRY(arena_get(ctx, (size_t)n * 10, &hist));
  TMAT_TRY(arena_get(ctx, (size_t)n, &kill));
  TMAT_TRY(arena_get(ctx, (size_t)n, &has_fork));
  if (use_median) TMAT_TRY(median_disk2(ctx, mask, H, W, m, st));
  else { binarize<<<grid1d_ept(n), 256, 0, st>>>(mask, n, m); KERNEL_CHECK(ctx); }
  // components by root pixel; no renumbering, no host round trip anywhere below
  TMAT_TRY(ccl_roots(ctx, m, H, W, 2, 0, parent, st, area, hist, kill));
  border_from_mask<<<grid2d_rpb(H, W), 256, 0, st>>>(m, H, W, border);
  KERNEL_CHECK(ctx);
  root_stats_kernel<<<grid2d(H, W), 256, 0, st>>>(parent, border, H, W, area, hist);
  KERNEL_CHECK(ctx);
  int* sk_list = nullptr;
  unsigned int* sk_n = nullptr;
  TMAT_TRY(skeletonize_impl(ctx, m, H, W, sk, nullptr, st, &sk_list, &sk_n));
  static const bool dense_sk = [] { const char* e = getenv("TMAT_SK_DENSE"); return e && atoi(e) == 1; }();
  if (sk_list && !dense_sk) {
    // the skeleton lives in the words the thinning kernel listed
    const int g = ctx->sm_count * 8;
    CUDA_TRY(cudaMemsetAsync(sk_root, 0xff, (size_t)n * sizeof(int), st));
    sk_init_list<<<g, 256, 0, st>>>(sk, H, W, sk_list, sk_n, sk_root, has_fork);
    KERNEL_CHECK(ctx);
    sk_merge_list<<<g, 256, 0, st>>>(sk, H, W, sk_list, sk_n, sk_root);
    KERNEL_CHECK(ctx);
    sk_flatten_fork_list<<<g, 256, 0, st>>>(sk, H, W, sk_list, sk_n, sk_root, has_fork);
    KERNEL_CHECK(ctx);
    sk_kill_list<<<g, 256, 0, st>>>(sk, W, sk_list, sk_n, sk_root, has_fork, parent, area, hist, remove_isolated, kill);
    KERNEL_CHECK(ctx);
  } else {
    TMAT_TRY(ccl_roots(ctx, sk, H, W, 2, 0, sk_root, st));
    CUDA_TRY(cudaMemsetAsync(has_fork, 0, (size_t)n, st));
    fork_kernel<<<grid2d(H, W), 256, 0, st>>>(sk, H, W, sk_root, has_fork);
    KERNEL_CHECK(ctx);
    root_kill_kernel<<<grid1d_ept(n), 256, 0, st>>>(sk_root, has_fork, parent, area, hist, n, remove_isolated, kill);
    KERNEL_CHECK(ctx);
  }
  root_apply_kill<<<grid1d_ept(n), 256, 0, st>>>(parent, kill, n, out);
  KERNEL_CHECK(ctx);
  return TMAT_OK;
}

// ---- C ABI: device-pointer entry points ----------------------------------------------------------------
#define MORPH_ARGS_OK(name) TMAT_REQUIRE(ctx && mask_dev && H >= 1 && W >= 1, name ": bad argument")

extern "C" int tmat_median_disk2_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, uint8_t* out_dev, void* stream) {
  MORPH_ARGS_OK("tmat_median_disk2");
  TMAT_REQUIRE(out_dev && out_dev != mask_dev, "tmat_median_disk2: out must be a distinct buffer");
  CUDA_TRY(cudaSetDevice(ctx->device));
  return median_disk2(ctx, mask_dev, H, W, out_dev, as_stream(stream));
}

extern "C" int tmat_erode_disk_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int r, int border_value,
                                   uint8_t* out_dev, void* stream) {
  MORPH_ARGS_OK("tmat_erode_disk");
  TMAT_REQUIRE(out_dev && out_dev != mask_dev, "tmat_erode_disk: out must be a distinct buffer");
  CUDA_TRY(cudaSetDevice(ctx->device));
  return erode_disk(ctx, mask_dev, H, W, r, border_value, out_dev, as_stream(stream));
}

extern "C" int tmat_footprint_filter_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W,
                                         const uint8_t* footprint_host, int fh, int fw, int op, int border_value,
                                         uint8_t* out_dev, void* stream) {
  MORPH_ARGS_OK("tmat_footprint_filter");
  TMAT_REQUIRE(out_dev && out_dev != mask_dev, "tmat_footprint_filter: out must be a distinct buffer");
  CUDA_TRY(cudaSetDevice(ctx->device));
  TMAT_TRY(arena_reset(ctx));
  return footprint_filter(ctx, mask_dev, H, W, footprint_host, fh, fw, op, border_value, out_dev, as_stream(stream));
}

extern "C" int tmat_label_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int connectivity,
                              int32_t* labels_dev, int* n_labels, void* stream) {
  MORPH_ARGS_OK("tmat_label");
  TMAT_REQUIRE(labels_dev && n_labels, "tmat_label: NULL output");
  TMAT_TRY(arena_reset(ctx));
  return label_impl(ctx, mask_dev, H, W, connectivity, labels_dev, n_labels, nullptr, as_stream(stream));
}

extern "C" int tmat_region_stats_dev(tmat_ctx* ctx, const int32_t* labels_dev, int H, int W, int n_labels,
                                     int64_t* area_host, int64_t* perim_hist_host, void* stream) {
  TMAT_REQUIRE(ctx && labels_dev && area_host && perim_hist_host && H >= 1 && W >= 1 && n_labels >= 0,
               "tmat_region_stats: bad argument");
  if (n_labels == 0) return TMAT_OK;
  TMAT_TRY(arena_reset(ctx));
  cudaStream_t st = as_stream(stream);
  int64_t n = (int64_t)H * W;
  uint8_t* border;
  unsigned int *area, *hist;
  TMAT_TRY(arena_get(ctx, (size_t)n, &border));
  TMAT_TRY(arena_get(ctx, (size_t)n_labels, &area));
  TMAT_TRY(arena_get(ctx, (size_t)n_labels * 50, &hist));
  CUDA_TRY(cudaMemsetAsync(area, 0, (size_t)n_labels * 4, st));
  CUDA_TRY(cudaMemsetAsync(hist, 0, (size_t)n_labels * 200, st));
  border_kernel<<<grid2d(H, W), 256, 0, st>>>(labels_dev, H, W, border);
  KERNEL_CHECK(ctx);
  region_stats_kernel<<<grid2d(H, W), 256, 0, st>>>(labels_dev, border, H, W, area, hist);
  KERNEL_CHECK(ctx);
  std::vector<unsigned int> ha(n_labels), hh((size_t)n_labels * 50);
  CUDA_TRY(cudaMemcpyAsync(ha.data(), area, (size_t)n_labels * 4, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(hh.data(), hist, (size_t)n_labels * 200, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  for (int i = 0; i < n_labels; ++i) area_host[i] = ha[i];
  for (size_t i = 0; i < hh.size(); ++i) perim_hist_host[i] = hh[i];
  return TMAT_OK;
}

extern "C" int tmat_skeletonize_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, uint8_t* out_dev,
                                    int* iterations, void* stream) {
  MORPH_ARGS_OK("tmat_skeletonize");
  TMAT_REQUIRE(out_dev && out_dev != mask_dev, "tmat_skeletonize: out must be a distinct buffer");
  TMAT_TRY(arena_reset(ctx));
  return skeletonize_impl(ctx, mask_dev, H, W, out_dev, iterations, as_stream(stream));
}

extern "C" int tmat_filter_branch_seg_mask_dev(tmat_ctx* ctx, const uint8_t* mask_dev, int H, int W, int use_median,
                                               int remove_isolated, uint8_t* out_dev, void* stream) {
  MORPH_ARGS_OK("tmat_filter_branch_seg_mask");
  TMAT_REQUIRE(out_dev, "tmat_filter_branch_seg_mask: NULL output");
  TMAT_TRY(arena_reset(ctx));
  return filter_branch_impl(ctx, mask_dev, H, W, use_median, remove_isolated, out_dev, as_stream(stream));
}

// ---- C ABI: host-pointer entry points --------------------------------------------------------------------
// copy a host mask in, run `body(d_in, d_out)`, copy the uint8 result out
template <typename F>
static int with_host_mask(tmat_ctx* ctx, const uint8_t* in_host, int H, int W, uint8_t* out_host, F body) {
  TMAT_REQUIRE(ctx && in_host && out_host && H >= 1 && W >= 1, "morphology(host): bad argument");
  TMAT_TRY(arena_reset(ctx));
  size_t n = (size_t)H * W;
  uint8_t *d_in, *d_out;
  TMAT_TRY(arena_get(ctx, n, &d_in));
  TMAT_TRY(arena_get(ctx, n, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_in, in_host, n, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(body(d_in, d_out));
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, n, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

extern "C" int tmat_median_disk2_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, uint8_t* out_host) {
  return with_host_mask(ctx, mask_host, H, W, out_host,
                        [&](uint8_t* a, uint8_t* b) { return median_disk2(ctx, a, H, W, b, 0); });
}
extern "C" int tmat_erode_disk_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int r, int border_value,
                                    uint8_t* out_host) {
  return with_host_mask(ctx, mask_host, H, W, out_host,
                        [&](uint8_t* a, uint8_t* b) { return erode_disk(ctx, a, H, W, r, border_value, b, 0); });
}
extern "C" int tmat_footprint_filter_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W,
                                          const uint8_t* footprint_host, int fh, int fw, int op, int border_value,
                                          uint8_t* out_host) {
  return with_host_mask(ctx, mask_host, H, W, out_host, [&](uint8_t* a, uint8_t* b) {
    return footprint_filter(ctx, a, H, W, footprint_host, fh, fw, op, border_value, b, 0);
  });
}
extern "C" int tmat_skeletonize_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, uint8_t* out_host,
                                     int* iterations) {
  return with_host_mask(ctx, mask_host, H, W, out_host,
                        [&](uint8_t* a, uint8_t* b) { return skeletonize_impl(ctx, a, H, W, b, iterations, 0); });
}
extern "C" int tmat_filter_branch_seg_mask_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int use_median,
                                                int remove_isolated, uint8_t* out_host) {
  return with_host_mask(ctx, mask_host, H, W, out_host, [&](uint8_t* a, uint8_t* b) {
    return filter_branch_impl(ctx, a, H, W, use_median, remove_isolated, b, 0);
  });
}

extern "C" int tmat_label_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int connectivity,
                               int32_t* labels_host, int* n_labels) {
  TMAT_REQUIRE(ctx && mask_host && labels_host && n_labels && H >= 1 && W >= 1, "tmat_label_host: bad argument");
  TMAT_TRY(arena_reset(ctx));
  size_t n = (size_t)H * W;
  uint8_t* d_in;
  int32_t* d_lab;
  TMAT_TRY(arena_get(ctx, n, &d_in));
  TMAT_TRY(arena_get(ctx, n, &d_lab));
  CUDA_TRY(cudaMemcpyAsync(d_in, mask_host, n, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(label_impl(ctx, d_in, H, W, connectivity, d_lab, n_labels, nullptr, 0));
  CUDA_TRY(cudaMemcpyAsync(labels_host, d_lab, n * 4, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}

// transforms.py:170-206: fill holes smaller than min_area0 (background components, connectivity0), then
// drop objects smaller than min_area1 (foreground components, connectivity1).
extern "C" int tmat_remove_small_islands_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int min_area0,
                                              int min_area1, int connectivity0, int connectivity1, uint8_t* out_host) {
  TMAT_REQUIRE(ctx && mask_host && out_host && H >= 1 && W >= 1, "tmat_remove_small_islands: bad argument");
  size_t n = (size_t)H * W;
  uint8_t lo = 255, hi = 0;
  for (size_t i = 0; i < n; ++i) { lo = std::min(lo, mask_host[i]); hi = std::max(hi, mask_host[i]); }
  // the reference raises ValueError (transforms.py:183-184)
  TMAT_REQUIRE(lo == 0 && hi <= 1, "this function expects a binary mask of values 0 and 1");
  return with_host_mask(ctx, mask_host, H, W, out_host, [&](uint8_t* a, uint8_t* b) {
    int* parent;
    unsigned int* area;
    TMAT_TRY(arena_get(ctx, n, &parent));
    TMAT_TRY(arena_get(ctx, n, &area));
    CUDA_TRY(cudaMemcpyAsync(b, a, n, cudaMemcpyDeviceToDevice, 0));
    // background components of the input
    TMAT_TRY(ccl_roots(ctx, b, H, W, connectivity0, 1, parent, 0));
    CUDA_TRY(cudaMemsetAsync(area, 0, n * 4, 0));
    root_area<<<grid1d((int64_t)n), 256>>>(parent, (int64_t)n, area);
    KERNEL_CHECK(ctx);
    flip_small<<<grid1d((int64_t)n), 256>>>(parent, area, (int64_t)n, (unsigned)std::max(min_area0, 0), 1, b);
    KERNEL_CHECK(ctx);
    // foreground components of the hole-filled mask
    TMAT_TRY(ccl_roots(ctx, b, H, W, connectivity1, 0, parent, 0));
    CUDA_TRY(cudaMemsetAsync(area, 0, n * 4, 0));
    root_area<<<grid1d((int64_t)n), 256>>>(parent, (int64_t)n, area);
    KERNEL_CHECK(ctx);
    flip_small<<<grid1d((int64_t)n), 256>>>(parent, area, (int64_t)n, (unsigned)std::max(min_area1, 0), 0, b);
    KERNEL_CHECK(ctx);
    return TMAT_OK;
  });
}

// transforms.py:291-303 with prop = 'area' (0), 'equivalent_diameter_area' (1) or 'eccentricity' (2): label (full
// connectivity) and paint each component's value
extern "C" int tmat_region_prop_image_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int prop,
                                           double* out_host);
extern "C" int tmat_area_image_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, double* out_host) {
  return tmat_region_prop_image_host(ctx, mask_host, H, W, 0, out_host);
}
extern "C" int tmat_region_prop_image_host(tmat_ctx* ctx, const uint8_t* mask_host, int H, int W, int prop,
                                           double* out_host) {
  TMAT_REQUIRE(ctx && mask_host && out_host && H >= 1 && W >= 1, "tmat_region_prop_image: bad argument");
  if (prop != 0 && prop != 1 && prop != 2) {
    tmat_set_error("tmat_region_prop_image: only area (0), equivalent_diameter_area (1) and eccentricity (2) are built");
    return TMAT_ERR_UNSUPPORTED;
  }
  TMAT_TRY(arena_reset(ctx));
  size_t n = (size_t)H * W;
  uint8_t* d_in;
  int* parent;
  unsigned int* area;
  double* d_out;
  TMAT_TRY(arena_get(ctx, n, &d_in));
  TMAT_TRY(arena_get(ctx, n, &parent));
  TMAT_TRY(arena_get(ctx, n, &area));
  TMAT_TRY(arena_get(ctx, n, &d_out));
  CUDA_TRY(cudaMemcpyAsync(d_in, mask_host, n, cudaMemcpyHostToDevice, 0));
  TMAT_TRY(ccl_roots(ctx, d_in, H, W, 2, 0, parent, 0));
  CUDA_TRY(cudaMemsetAsync(area, 0, n * 4, 0));
  if (prop == 2) {
    TMAT_REQUIRE(H <= 65535 && W <= 65535, "tmat_region_prop_image: eccentricity needs H, W <= 65535");
    RootMoments* mom;
    double* val;
    TMAT_TRY(arena_get(ctx, n, &mom));
    TMAT_TRY(arena_get(ctx, n, &val));
    CUDA_TRY(cudaMemsetAsync(mom, 0, n * sizeof(RootMoments), 0));
    root_moments_kernel<<<grid2d(H, W), 256>>>(parent, H, W, area, mom);
    KERNEL_CHECK(ctx);
    root_eccentricity_kernel<<<grid1d((int64_t)n), 256>>>(parent, (int64_t)n, area, mom, val);
    KERNEL_CHECK(ctx);
    paint_root_value<<<grid1d((int64_t)n), 256>>>(parent, val, (int64_t)n, d_out);
    KERNEL_CHECK(ctx);
  } else {
    root_area<<<grid1d((int64_t)n), 256>>>(parent, (int64_t)n, area);
    KERNEL_CHECK(ctx);
    area_image_kernel<<<grid1d((int64_t)n), 256>>>(parent, area, (int64_t)n, prop, d_out);
    KERNEL_CHECK(ctx);
  }
  CUDA_TRY(cudaMemcpyAsync(out_host, d_out, n * 8, cudaMemcpyDeviceToHost, 0));
  CUDA_TRY(cudaStreamSynchronize(0));
  return TMAT_OK;
}
