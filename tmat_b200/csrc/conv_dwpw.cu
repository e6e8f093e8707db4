// Fused SeparableConv2D (models.py:129,134): [ReLU] -> depthwise 3x3 -> pointwise 1x1 + bias + BN [+ ReLU]
// as ONE tcgen05 kernel.  The depthwise result never goes to HBM: it is produced straight into the
// 128-byte-swizzled A-operand stage of the pointwise GEMM.
//
// Five warpgroups (20 warps); roles are warpgroup-aligned so that setmaxnreg can move registers between them
// (launch: 96 per thread = 61440 per CTA, and a CTA can only re-distribute what it was launched with: warpgroup 0
// keeps 40, the epilogue warpgroups stay at 96, the depthwise ones grow to 120 -- 60416 in total):
//   warp 0        raw-tile producer: per (tile, 64-channel slab) one haloed raw input tile
//                 (64 ch x (TW+2) x (TH+2) x TN, zero-filled outside the image = 'same' padding) into the raw ring
//   warp 1        MMA issuer (tcgen05.mma M=128, N=BN, K=16 x4 per slab), accumulators in all 512 TMEM columns
//   warp 2        weight producer: all pointwise slabs once (resident, <= 64 KB) or a two-slab ring
//   warps 4..11   epilogue, two warpgroups taking alternate tiles (ncu: the single epilogue warpgroup was busy all
//                 the time with one dependent instruction stream per scheduler while the MMA warp waited for a free
//                 accumulator 60-70 % of its time and the depthwise warps for a free A stage 45 %): tcgen05.ld ->
//                 +shift (staged in smem), ReLU in the convert -> bf16 -> swizzled staging tile -> TMA store, or
//                 -> 3x3/2 max-pool of the staged tile -> TMA max-reduce.  Each warpgroup has its own staging
//                 buffers, named barrier and store leader.
//   warps 12..19  depthwise, two groups of four warps taking alternate items: each thread owns 4 channels of two
//                 adjacent tile columns and slides down the rows (8-byte ld.shared per raw pixel, 36 fp32 weights
//                 in registers, FFMA2 on fp32 pairs), rounds to bf16 and writes the A ring in UMMA layout
#include "tc_ptx.cuh"
#include <algorithm>

using namespace tcptx;

namespace {

constexpr int DW_THREADS = 640;     // 20 warps = 5 warpgroups
constexpr int DW_GROUP_WARPS = 4;    // depthwise warps 12..19: two groups of four, each owning every other (tile, slab) item
constexpr int WARP_RAW = 0, WARP_MMA = 1, WARP_WEIGHTS = 2, WARP_EPI0 = 4, WARP_DW0 = 12;
constexpr int REGS_LIGHT = 40, REGS_EPI = 96, REGS_DW = 120;
static_assert(128 * REGS_LIGHT + 256 * REGS_EPI + 256 * REGS_DW <= DW_THREADS * 96, "register budget of the CTA");
constexpr int MAX_A = 4;             // A (depthwise output) ring
constexpr int MAX_B = 8;             // pointwise-weight slabs: ring entries, or all nk slabs when resident
constexpr int MAX_RAW = 8;           // raw haloed input tiles
constexpr int BAR_BYTES = 1024;      // barrier block
constexpr int MAX_N = 1024;          // shift vector staged in shared memory
constexpr int TAIL_BYTES = BAR_BYTES + MAX_N * 4 + 1024;   // barriers + shift + alignment slack

struct DwpwParams {
  int tw_log, th_log, tn_log;
  int tiles_x, tiles_y, tiles_b, tiles_n, total_tiles;
  int nk;            // 64-channel slabs of the input
  int C;             // input channels
  int relu_in;
  int raw_bytes;     // one raw ring slot (1024-aligned)
  int raw_tx;        // bytes one raw TMA load delivers
  int n_raw, n_a, n_b;
  int b_resident;    // all nk weight slabs stay in smem for the whole kernel (n_b == nk, tiles_n == 1)
  int pool;          // epilogue max-pools (3, stride 2, same) the tile and max-reduces it into the pooled tensor
  int pw, ph;        // pooled box of a tile: (TW/2 + 1) x (TH/2 + 1), origin (x0/2 - 1, y0/2 - 1)
  int n_epi;         // epilogue warpgroups in use (2, or 1 when shared memory has no room for a second set of buffers)
  int stg_wg_bytes;  // staging bytes of one epilogue warpgroup
  int spw;           // store staging tiles per epilogue warpgroup (plain epilogue)
  int npb;           // pooled boxes per epilogue warpgroup (pooled epilogue)
  const float* dw;   // [9][C]
  EpiParams e;
};

__device__ __forceinline__ uint2 ld_shared_v2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void st_shared_v2(uint32_t addr, uint2 v) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ uint32_t f2_to_bf2(f32x2 v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  return pack_bf16x2(lo, hi);
}
// bf16x2 (channel c in the low half, c+1 in the high half) -> fp32 pair, with the block's input ReLU fused
template <bool RELU>
__device__ __forceinline__ f32x2 bf2_to_f2(uint32_t x) {
  if (RELU) {
    __nv_bfloat162 h = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&x), __floats2bfloat162_rn(0.f, 0.f));
    x = *reinterpret_cast<uint32_t*>(&h);
  }
  return pack2(__uint_as_float(x << 16), __uint_as_float(x & 0xffff0000u));
}

// Depthwise 3x3 of two adjacent tile columns (x even, x + 1) for a thread's 4 channels, RH = TH_C + 2 raw rows,
// fully unrolled: raw row yy is tap row ky of output row yy - ky; the three live accumulators rotate by
// compile-time index.  The four raw pixels x .. x+3 of a row are loaded and unpacked once and feed both columns,
// so the per-output cost of ld.shared + bf16 unpack is 2 instead of 3.
template <int TH_C, bool RELU>
__device__ __forceinline__ void dw_column2(uint32_t rbase_sub, uint32_t abase_sub, uint32_t rp0, uint32_t ar0,
                                           uint32_t RW, uint32_t TW, uint32_t jch, const f32x2 (&wr)[9][2]) {
  f32x2 acc[3][2][2];   // [live output row][column][channel pair]
  auto raw_addr = [&](uint32_t rp) { return rbase_sub + (rp << 7) + ((jch ^ (rp & 7u)) << 4); };
#pragma unroll
  for (int yy = 0; yy < TH_C + 2; ++yy) {
    const uint32_t rp = rp0 + (uint32_t)yy * RW;
    uint2 q[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = ld_shared_v2(raw_addr(rp + k));
    f32x2 f[4][2];
#pragma unroll
    for (int k = 0; k < 4; ++k) { f[k][0] = bf2_to_f2<RELU>(q[k].x); f[k][1] = bf2_to_f2<RELU>(q[k].y); }
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int o = yy - ky;
      if (o < 0 || o >= TH_C) continue;
#pragma unroll
      for (int col = 0; col < 2; ++col) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          f32x2 a = (ky == 0) ? fmul2(f[col][c], wr[0][c]) : ffma2(f[col][c], wr[3 * ky][c], acc[o % 3][col][c]);
          a = ffma2(f[col + 1][c], wr[3 * ky + 1][c], a);
          acc[o % 3][col][c] = ffma2(f[col + 2][c], wr[3 * ky + 2][c], a);
        }
      }
    }
    if (yy >= 2) {
      const int o = yy - 2;
#pragma unroll
      for (int col = 0; col < 2; ++col) {
        const uint32_t ar = ar0 + (uint32_t)o * TW + col;
        st_shared_v2(abase_sub + (ar << 7) + ((jch ^ (ar & 7u)) << 4),
                     make_uint2(f2_to_bf2(acc[o % 3][col][0]), f2_to_bf2(acc[o % 3][col][1])));
      }
    }
  }
}

// dw_column2 with the tile width known at compile time: pixel offsets become immediates and the 128-byte swizzle
// term (pixel index mod 8) reduces to eight per-thread base registers for the loads and two for the stores, so
// the row loop carries no address arithmetic at all.
template <int TH_C, int TW_C, bool RELU>
__device__ __forceinline__ void dw_column2_fixed(uint32_t rbase_sub, uint32_t abase_sub, uint32_t rp0, uint32_t ar0,
                                                 uint32_t jch, const f32x2 (&wr)[9][2]) {
  constexpr int RW = TW_C + 2;
  uint32_t R[8], S[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    R[c] = rbase_sub + (rp0 << 7) + ((jch ^ ((rp0 + c) & 7u)) << 4);
    S[c] = abase_sub + (ar0 << 7) + ((jch ^ ((ar0 + c) & 7u)) << 4);
  }
  f32x2 acc[3][2][2];
#pragma unroll
  for (int yy = 0; yy < TH_C + 2; ++yy) {
    uint2 q[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = ld_shared_v2(R[(yy * RW + k) & 7] + (uint32_t)((yy * RW + k) * 128));
    f32x2 f[4][2];
#pragma unroll
    for (int k = 0; k < 4; ++k) { f[k][0] = bf2_to_f2<RELU>(q[k].x); f[k][1] = bf2_to_f2<RELU>(q[k].y); }
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int o = yy - ky;
      if (o < 0 || o >= TH_C) continue;
#pragma unroll
      for (int col = 0; col < 2; ++col) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          f32x2 a = (ky == 0) ? fmul2(f[col][c], wr[0][c]) : ffma2(f[col][c], wr[3 * ky][c], acc[o % 3][col][c]);
          a = ffma2(f[col + 1][c], wr[3 * ky + 1][c], a);
          acc[o % 3][col][c] = ffma2(f[col + 2][c], wr[3 * ky + 2][c], a);
        }
      }
    }
    if (yy >= 2) {
      const int o = yy - 2;
#pragma unroll
      for (int col = 0; col < 2; ++col)
        st_shared_v2(S[(o * TW_C + col) & 7] + (uint32_t)((o * TW_C + col) * 128),
                     make_uint2(f2_to_bf2(acc[o % 3][col][0]), f2_to_bf2(acc[o % 3][col][1])));
    }
  }
}

// One 64-column chunk of a thread's accumulator row: TMEM -> + shift (from shared memory) -> [ReLU] -> bf16 ->
// swizzled staging row.  Packed fp32 pairs: one FADD2 per two columns, ReLU folded into the bf16 convert.
__device__ __forceinline__ void ld_shared_v4f(uint32_t addr, f32x2& a, f32x2& b) {
  asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}
template <bool RELU>
__device__ __forceinline__ void dwpw_epi_chunk(uint32_t taddr, uint32_t shift_addr, uint32_t buf_row, uint32_t sw) {
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    uint32_t rr[32];
    tmem_ld32(taddr + half * 32, rr);
    f32x2 sh[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) ld_shared_v4f(shift_addr + (uint32_t)(half * 32 + 4 * i) * 4u, sh[2 * i], sh[2 * i + 1]);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint4 o;
      o.x = cvt_bf16x2<RELU>(fadd2(pack2(__uint_as_float(rr[8 * j]), __uint_as_float(rr[8 * j + 1])), sh[4 * j]));
      o.y = cvt_bf16x2<RELU>(fadd2(pack2(__uint_as_float(rr[8 * j + 2]), __uint_as_float(rr[8 * j + 3])), sh[4 * j + 1]));
      o.z = cvt_bf16x2<RELU>(fadd2(pack2(__uint_as_float(rr[8 * j + 4]), __uint_as_float(rr[8 * j + 5])), sh[4 * j + 2]));
      o.w = cvt_bf16x2<RELU>(fadd2(pack2(__uint_as_float(rr[8 * j + 6]), __uint_as_float(rr[8 * j + 7])), sh[4 * j + 3]));
      st_shared_v4(buf_row + ((((uint32_t)(half * 4 + j)) ^ sw) << 4), o);
    }
  }
}

__device__ __forceinline__ uint32_t bf2_max(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

constexpr int POOL_BUF_BYTES = STG_BYTES / 2;      // two pooled boxes share the second staging tile
constexpr int POOL_MAX_ROWS = POOL_BUF_BYTES / 128;

// MaxPool2D(3, strides=2, 'same') of one staged 64-channel tile (models.py:139; pads 0 before / 1 after, padding
// excluded).  Pooled pixel (k, l) of the tile's box covers tile rows 2k-2 .. 2k and columns 2l-2 .. 2l clipped to the
// tile: rows/columns k = 0, l = 0 and the last ones are partial and are completed by the neighbouring tiles' boxes
// through the max-reducing store.  Eight consecutive threads take the eight 16-byte chunks of one pooled pixel, so
// every quarter-warp reads and writes one whole 128-byte swizzle row.
// (ox, oy) = 1 for tiles on the left / top image edge: their box starts one pooled pixel later (a TMA store must not
// start at coordinate -1) and its last column / row, which then belongs to the next tile, stays at -inf.
__device__ __forceinline__ void dwpw_pool_tile(uint32_t stg, uint32_t pbuf, int et, int TW, int TH, int TN, int PW,
                                               int PH, int ox, int oy) {
  const uint32_t c = (uint32_t)(et & 7);
  const int rows = TN * PH * PW;
  for (int pr = et >> 3; pr < rows; pr += 16) {
    const int n = pr / (PH * PW), r2 = pr - n * (PH * PW);
    const int kb = r2 / PW;
    const int k = kb + oy, l = r2 - kb * PW + ox;
    const int ylo = max(2 * k - 2, 0), yhi = min(2 * k, TH - 1);
    const int xlo = max(2 * l - 2, 0), xhi = min(2 * l, TW - 1);
    uint4 m = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);   // -inf
    for (int yy = ylo; yy <= yhi; ++yy) {
      for (int xx = xlo; xx <= xhi; ++xx) {
        const uint32_t srow = (uint32_t)((n * TH + yy) * TW + xx);
        const uint4 v = ld_shared_v4(stg + srow * 128u + ((c ^ (srow & 7u)) << 4));
        m.x = bf2_max(m.x, v.x); m.y = bf2_max(m.y, v.y); m.z = bf2_max(m.z, v.z); m.w = bf2_max(m.w, v.w);
      }
    }
    st_shared_v4(pbuf + (uint32_t)pr * 128u + ((c ^ ((uint32_t)pr & 7u)) << 4), m);
  }
}

// The same for the tile shapes the UNet uses (TW a multiple of 8, so a staged pixel's swizzle phase is its column & 7),
// split in two so that the index arithmetic leaves the per-chunk path (ncu: 230 of the 460 instructions an epilogue
// warp spent per 64-channel chunk were clamps, swizzles and multiplies of the pooling addresses):
//  * pool_offsets_fixed computes, once per (ox, oy) -- i.e. once for all interior tiles -- the byte addresses a thread
//    needs for its up to four pooled items: three tap rows and three tap columns each, clamped into the tile (a
//    duplicate tap does not change a maximum), and the item's offset in the pooled box;
//  * pool_apply is then nine independent 16-byte loads, eight 3-input maxima per 32-bit lane and one store per item.
struct PoolOfs {
  uint32_t yo[4][3];   // staging address of the tap row (includes the staging base)
  uint32_t xo[4][3];   // byte offset of the tap column inside a row, swizzled for the thread's 16-byte chunk
  uint32_t st[4];      // byte offset of the item in the pooled box
  uint32_t flags;      // bit 2j: item j exists; bit 2j+1: it has taps (else it stays at -inf)
};
template <int TWL, int THL>
__device__ __forceinline__ void pool_offsets_fixed(uint32_t stg, int et, int ox, int oy, PoolOfs& o) {
  constexpr int TW = 1 << TWL, TH = 1 << THL, TN = 128 / (TW * TH), PW = TW / 2 + 1, PH = TH / 2 + 1;
  constexpr int ROWS = TN * PH * PW;
  static_assert(TW % 8 == 0 && ROWS <= POOL_MAX_ROWS && ROWS <= 64, "tile shape");
  const uint32_t c = (uint32_t)(et & 7);
  o.flags = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int pr = (et >> 3) + 16 * j;
    const int n = pr / (PH * PW), r2 = pr - n * (PH * PW);
    const int kb = r2 / PW, lb = r2 - kb * PW;
    const int k = kb + oy, l = lb + ox;
    if (pr < ROWS) o.flags |= 1u << (2 * j);
    if (pr < ROWS && k <= TH / 2 && l <= TW / 2) o.flags |= 2u << (2 * j);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const int y = min(max(2 * k - 2 + d, 0), TH - 1), x = min(max(2 * l - 2 + d, 0), TW - 1);
      o.yo[j][d] = stg + (uint32_t)((min(n, TN - 1) * TH + y) * TW) * 128u;
      o.xo[j][d] = (uint32_t)x * 128u + ((c ^ ((uint32_t)x & 7u)) << 4);
    }
    o.st[j] = (uint32_t)pr * 128u + ((c ^ ((uint32_t)pr & 7u)) << 4);
  }
}
__device__ __forceinline__ void pool_apply(uint32_t pbuf, const PoolOfs& o) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (o.flags & (1u << (2 * j))) {
      uint4 m = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);   // -inf
      if (o.flags & (2u << (2 * j))) {
        uint4 v[9];
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) v[3 * dy + dx] = ld_shared_v4(o.yo[j][dy] + o.xo[j][dx]);
        m = v[0];
#pragma unroll
        for (int i = 1; i < 9; ++i) {
          m.x = bf2_max(m.x, v[i].x); m.y = bf2_max(m.y, v[i].y); m.z = bf2_max(m.z, v[i].z); m.w = bf2_max(m.w, v[i].w);
        }
      }
      st_shared_v4(pbuf + o.st[j], m);
    }
  }
}

__global__ void __launch_bounds__(256) fill_neg_inf_bf16(uint4* __restrict__ p, int64_t n16) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n16) p[i] = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);
}

template <int BN>
__global__ void __launch_bounds__(DW_THREADS, 1)
conv_dwpw_kernel(const __grid_constant__ CUtensorMap tmRaw, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmPool,
                 const __grid_constant__ DwpwParams p) {
  constexpr int B_BYTES = BN * KSLAB * 2;
  constexpr int N_ACC = (BN <= 128) ? 4 : 2;      // accumulator buffers: all 512 TMEM columns
  constexpr int TMEM_COLS = N_ACC * BN < 32 ? 32 : N_ACC * BN;
  extern __shared__ uint8_t smem_raw[];
  // [A ring | B ring or resident slabs | raw ring | store staging | barriers]
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_base = smem_base + p.n_a * A_STAGE_BYTES;
  const uint32_t raw_base = b_base + p.n_b * B_BYTES;
  const uint32_t stg_base = raw_base + p.n_raw * p.raw_bytes;
  const uint32_t bar_base = stg_base + p.n_epi * p.stg_wg_bytes;
  auto afull_bar = [&](int s) { return bar_base + 8u * s; };
  auto aempty_bar = [&](int s) { return bar_base + 8u * (MAX_A + s); };
  auto bfull_bar = [&](int s) { return bar_base + 8u * (2 * MAX_A + s); };
  auto bempty_bar = [&](int s) { return bar_base + 8u * (2 * MAX_A + MAX_B + s); };
  auto rfull_bar = [&](int s) { return bar_base + 8u * (2 * MAX_A + 2 * MAX_B + s); };
  auto rempty_bar = [&](int s) { return bar_base + 8u * (2 * MAX_A + 2 * MAX_B + MAX_RAW + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * MAX_A + 2 * MAX_B + 2 * MAX_RAW + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * MAX_A + 2 * MAX_B + 2 * MAX_RAW + 4 + s); };
  constexpr int TMEM_SLOT = 8 * (2 * MAX_A + 2 * MAX_B + 2 * MAX_RAW + 8);
  constexpr int ISSUED_SLOT = TMEM_SLOT + 8;
  const uint32_t shift_base = bar_base + BAR_BYTES;      // the layer's N shift values (fp32), staged once
  const uint32_t tmem_slot = bar_base + TMEM_SLOT;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (bar_base - smem_base) + TMEM_SLOT);
  // number of raw tiles the producer has issued so far: a depthwise group only starts to wait on a slot's
  // barrier once the load it wants has been issued (two groups share the ring, so a plain parity wait could
  // be satisfied by the slot's previous use)
  volatile int* issued = reinterpret_cast<volatile int*>(smem_gen + (bar_base - smem_base) + ISSUED_SLOT);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.n_a; ++s) { mbar_init(afull_bar(s), DW_GROUP_WARPS); mbar_init(aempty_bar(s), 1); }
    for (int s = 0; s < p.n_b; ++s) { mbar_init(bfull_bar(s), 1); mbar_init(bempty_bar(s), 1); }
    for (int s = 0; s < p.n_raw; ++s) { mbar_init(rfull_bar(s), 1); mbar_init(rempty_bar(s), DW_GROUP_WARPS); }
    *issued = 0;
    for (int s = 0; s < N_ACC; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 128); }
    fence_barrier_init();
    fence_proxy_async();
    tma_prefetch_desc(&tmRaw);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmOut);
    tma_prefetch_desc(&tmPool);
  }
  {
    float* shift_gen = reinterpret_cast<float*>(smem_gen + (bar_base - smem_base) + BAR_BYTES);
    for (int i = threadIdx.x; i < p.e.Cout; i += DW_THREADS) shift_gen[i] = __ldg(p.e.shift + i);
  }
  if (warp == WARP_MMA) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  const int TW = 1 << p.tw_log, TH = 1 << p.th_log, TN = 1 << p.tn_log;
  const int tiles_xy = p.tiles_x * p.tiles_y;
  auto decode = [&](int tile, int& nt, int& x0, int& y0, int& b0) {
    int mt = tile / p.tiles_n;
    nt = tile - mt * p.tiles_n;
    int tb = mt / tiles_xy, rr = mt - tb * tiles_xy;
    int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
    x0 = tx << p.tw_log; y0 = ty << p.th_log; b0 = tb << p.tn_log;
  };

  // registers follow the work: warpgroup 0 (three single-warp roles) gives its registers up, the epilogue and
  // depthwise warpgroups grow (setmaxnreg sits at the head of each warpgroup's own branch)
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS_LIGHT));
  if (warp == WARP_RAW) {
    // ================= raw-tile producer =================
    // Only the raw ring gates this warp (the weight slabs have their own producer), so it runs as far ahead of
    // the depthwise groups as the ring allows: n_raw - 2 haloed tiles in flight while two are being worked on.
    int rs = 0; uint32_t rphase = 0; int item = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      int nt, x0, y0, b0;
      decode(tile, nt, x0, y0, b0);
      for (int ks = 0; ks < p.nk; ++ks) {
        mbar_wait(rempty_bar(rs), rphase ^ 1);
        if (lane == 0) {
          mbar_expect_tx(rfull_bar(rs), p.raw_tx);
          tma_load_4d(raw_base + rs * p.raw_bytes, &tmRaw, rfull_bar(rs), ks * KSLAB, x0 - 1, y0 - 1, b0);
          *issued = ++item;
        }
        __syncwarp();
        if (++rs == p.n_raw) { rs = 0; rphase ^= 1; }
      }
    }
  } else if (warp == WARP_WEIGHTS) {
    // ================= pointwise-weight producer =================
    if (p.b_resident) {
      if (lane == 0) {
        for (int ks = 0; ks < p.nk; ++ks) {
          mbar_expect_tx(bfull_bar(ks), B_BYTES);
          tma_load_2d(b_base + ks * B_BYTES, &tmB, bfull_bar(ks), ks * KSLAB, 0);
        }
      }
    } else {
      int sb = 0; uint32_t bphase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int nt = tile % p.tiles_n;
        for (int ks = 0; ks < p.nk; ++ks) {
          mbar_wait(bempty_bar(sb), bphase ^ 1);
          if (lane == 0) {
            mbar_expect_tx(bfull_bar(sb), B_BYTES);
            tma_load_2d(b_base + sb * B_BYTES, &tmB, bfull_bar(sb), ks * KSLAB, nt * BN);
          }
          __syncwarp();
          if (++sb == p.n_b) { sb = 0; bphase ^= 1; }
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // ================= MMA issuer (warp-convergent loop, one elected lane issues) =================
    {
      constexpr uint32_t idesc = umma_idesc<BN>();
      int sa = 0, sb = 0; uint32_t pa = 0, pb = 0;
      int it = 0;
      if (p.b_resident) {
        for (int ks = 0; ks < p.nk; ++ks) mbar_wait(bfull_bar(ks), 0);
      }
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int as = it % N_ACC;
        const uint32_t aphase = (it / N_ACC) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int ks = 0; ks < p.nk; ++ks) {
          if (!p.b_resident) mbar_wait(bfull_bar(sb), pb);
          mbar_wait(afull_bar(sa), pa);
          tc_fence_after();
          const uint32_t a_addr = smem_base + sa * A_STAGE_BYTES;
          const uint32_t b_addr = b_base + (p.b_resident ? ks : sb) * B_BYTES;
          // the four K = 16 MMAs of the slab: +32 bytes along K inside the swizzle atom per MMA
          umma_k64_elect(tmem_d, umma_desc_sw128(a_addr), umma_desc_sw128(b_addr), idesc, (uint32_t)(ks != 0));
          umma_commit_elect(aempty_bar(sa));                       // A stage reusable once these MMAs retire
          if (!p.b_resident) {
            umma_commit_elect(bempty_bar(sb));
            if (++sb == p.n_b) { sb = 0; pb ^= 1; }
          }
          if (ks == p.nk - 1) umma_commit_elect(tfull_bar(as));    // accumulator complete
          if (++sa == p.n_a) { sa = 0; pa ^= 1; }
        }
      }
    }
    __syncwarp();
  }
  } else if (warp < WARP_DW0) {
    // ================= epilogue (two warpgroups on alternate tiles; TMEM lane quadrant = warp & 3) =================
    static_assert(REGS_EPI == 96, "the epilogue warpgroups keep the launch allocation");
    const int ewg = (warp - WARP_EPI0) >> 2;
    const uint32_t bar_id = 1u + (uint32_t)ewg;
    auto epi_bar_sync = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory"); };
    const uint32_t stg_wg = stg_base + (uint32_t)(ewg * p.stg_wg_bytes);
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const bool leader = (threadIdx.x == (WARP_EPI0 + 4 * ewg) * 32);
    const uint32_t stg_row = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const EpiParams& e = p.e;
    int sbuf = 0;
    int it = 0;
    // pooling addresses of this thread (fixed tile shapes), recomputed only when the tile's edge flags change
    const bool pool_fixed = p.pool && p.th_log == 3 && (p.tw_log == 4 || p.tw_log == 3);
    PoolOfs pofs;
    int pofs_ox = -1, pofs_oy = -1;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      if (ewg >= p.n_epi || (p.n_epi == 2 && (it & 1) != ewg)) continue;
      const int as = it % N_ACC;
      const uint32_t aphase = (it / N_ACC) & 1;
      int nt, x0, y0, b0;
      decode(tile, nt, x0, y0, b0);
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
      if (p.pool) {
        // fused MaxPool: the 64-channel tile is staged (generic proxy only), pooled into one of two small boxes and
        // max-reduced into the pooled tensor; the full-resolution tensor is never written
        const int ox = (x0 == 0), oy = (y0 == 0);
        if (pool_fixed && (ox != pofs_ox || oy != pofs_oy)) {
          if (p.tw_log == 4) pool_offsets_fixed<4, 3>(stg_wg, row, ox, oy, pofs);
          else pool_offsets_fixed<3, 3>(stg_wg, row, ox, oy, pofs);
          pofs_ox = ox; pofs_oy = oy;
        }
#pragma unroll 1
        for (int c64 = 0; c64 < BN; c64 += 64) {
          const int ch0 = nt * BN + c64;
          const uint32_t pbuf = stg_wg + STG_BYTES + (uint32_t)sbuf * POOL_BUF_BYTES;
          if (leader) { if (p.npb == 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>(); }   // the reduce that last read this pooled box is done
          if (e.relu) dwpw_epi_chunk<true>(taddr + c64, shift_base + (uint32_t)ch0 * 4u, stg_wg + stg_row, sw);
          else dwpw_epi_chunk<false>(taddr + c64, shift_base + (uint32_t)ch0 * 4u, stg_wg + stg_row, sw);
          if (c64 + 64 >= BN) {
            tc_fence_before();
            mbar_arrive(tempty_bar(as));
          }
          epi_bar_sync();                                  // tile staged; leader's wait is behind everyone
          if (pool_fixed) pool_apply(pbuf, pofs);
          else dwpw_pool_tile(stg_wg, pbuf, row, TW, TH, TN, p.pw, p.ph, ox, oy);
          fence_proxy_async();
          epi_bar_sync();                                  // box complete, staging tile free again
          if (leader) {
            tma_reduce_max_4d(&tmPool, pbuf, ch0, (x0 >> 1) - 1 + ox, (y0 >> 1) - 1 + oy, b0);
            tma_store_commit();
          }
          if (p.npb == 2) sbuf ^= 1;
        }
        continue;
      }
#pragma unroll 1
      for (int c64 = 0; c64 < BN; c64 += 64) {
        const int ch0 = nt * BN + c64;
        if (leader) { if (p.spw == 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>(); }
        epi_bar_sync();
        const uint32_t buf = stg_wg + sbuf * STG_BYTES;
        if (e.relu) dwpw_epi_chunk<true>(taddr + c64, shift_base + (uint32_t)ch0 * 4u, buf + stg_row, sw);
        else dwpw_epi_chunk<false>(taddr + c64, shift_base + (uint32_t)ch0 * 4u, buf + stg_row, sw);
        if (c64 + 64 >= BN) {
          tc_fence_before();
          mbar_arrive(tempty_bar(as));
        }
        fence_proxy_async();
        epi_bar_sync();
        if (leader) {
          tma_store_4d(&tmOut, buf, ch0, x0, y0, b0);
          tma_store_commit();
        }
        if (p.spw == 2) sbuf ^= 1;
      }
    }
    if (leader) tma_store_wait_all();
  } else {
    // ================= depthwise (warps 12..19) =================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_DW));
    const int t = threadIdx.x - WARP_DW0 * 32;   // 0..255
    const int grp = t >> 7;                 // group 0 takes even items, group 1 odd items
    const int tg = t & 127;
    const int h = tg & 15;                  // 4-channel group within the 64-channel slab
    const uint32_t jch = (uint32_t)(h >> 1);            // 16-byte chunk
    const uint32_t sub = (uint32_t)(h & 1) << 3;        // 8-byte half of the chunk
    const int cslot = tg >> 4;              // column(-pair) slot 0..7
    const int ncols = TW * TN;
    const int RW = TW + 2, RH = TH + 2;
    const bool relu = p.relu_in != 0;
    int item = 0;
    int w_ks = -1;
    f32x2 wr[9][2];
    // ring slot and phase of the current item, advanced per item (every item, also the other group's) instead of
    // item % n and item / n: the two integer divisions were ~100 instructions of each item's prologue
    int sa = 0, rs = 0;
    uint32_t pa = 0, rphase = 0;
    auto ring_advance = [&]() {
      if (++sa == p.n_a) { sa = 0; pa ^= 1u; }
      if (++rs == p.n_raw) { rs = 0; rphase ^= 1u; }
    };
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      for (int ks = 0; ks < p.nk; ++ks, ++item, ring_advance()) {
        if ((item & 1) != grp) continue;
        // a group sees the same slab every time when nk <= 2 (item parity = slab parity): its 36 depthwise weights
        // are then loaded once, not once per item (ncu: 16 % of the depthwise warps' samples sat behind these loads)
        if (ks != w_ks) {
          w_ks = ks;
#pragma unroll
          for (int tp = 0; tp < 9; ++tp) {
            float4 wv = __ldg(reinterpret_cast<const float4*>(p.dw + tp * p.C + ks * KSLAB + h * 4));
            wr[tp][0] = pack2(wv.x, wv.y); wr[tp][1] = pack2(wv.z, wv.w);
          }
        }
        while (*issued <= item) { }
        mbar_wait(rfull_bar(rs), rphase);
        mbar_wait(aempty_bar(sa), pa ^ 1);
        const uint32_t rbase = raw_base + rs * p.raw_bytes + sub;
        const uint32_t abase = smem_base + sa * A_STAGE_BYTES + sub;
        {
          for (int ci = 2 * cslot; ci < ncols; ci += 16) {
            const int n = ci >> p.tw_log, x = ci & (TW - 1);
            const uint32_t rp0 = (uint32_t)((n * RH) * RW + x);   // raw pixel (n, yy = 0, xx = x)
            const uint32_t ar0 = (uint32_t)((n * TH) * TW + x);   // A row of output (n, y = 0, x)
            if (TH == 8 && TW == 16) {
              if (!relu) dw_column2_fixed<8, 16, false>(rbase, abase, rp0, ar0, jch, wr);
              else dw_column2_fixed<8, 16, true>(rbase, abase, rp0, ar0, jch, wr);
            } else if (TH == 8 && TW == 8) {
              if (!relu) dw_column2_fixed<8, 8, false>(rbase, abase, rp0, ar0, jch, wr);
              else dw_column2_fixed<8, 8, true>(rbase, abase, rp0, ar0, jch, wr);
            } else if (!relu) {
              if (TH == 8) dw_column2<8, false>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else if (TH == 4) dw_column2<4, false>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else if (TH == 16) dw_column2<16, false>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else if (TH == 2) dw_column2<2, false>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else dw_column2<1, false>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
            } else {
              if (TH == 8) dw_column2<8, true>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else if (TH == 4) dw_column2<4, true>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else if (TH == 16) dw_column2<16, true>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else if (TH == 2) dw_column2<2, true>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
              else dw_column2<1, true>(rbase, abase, rp0, ar0, RW, TW, jch, wr);
            }
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(afull_bar(sa));
          mbar_arrive(rempty_bar(rs));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) tmem_dealloc<TMEM_COLS>(tmem_base);
}

struct DwpwPlan {
  int tw_log, th_log, tn_log, BN, n_raw, n_a, n_b, b_resident, raw_bytes, raw_tx, smem;
  int n_epi, stg_wg_bytes, spw, npb;
};

bool plan_dwpw(int B, int S, int C, int N, bool pool, DwpwPlan* pl) {
  if (C % KSLAB != 0 || N % 64 != 0 || C / KSLAB > 64 || N > MAX_N) return false;
  tc_choose_tile(B, S, S, &pl->tw_log, &pl->th_log, &pl->tn_log);
  {
    // the depthwise warps re-read 2 halo rows per tile column: prefer 8-row tiles (16x8x1, 8x8x2, 4x8x4 ...)
    // when they cover the image with at most 10 % more tiles than the best shape
    int64_t best = (int64_t)cdiv(S, 1 << pl->tw_log) * cdiv(S, 1 << pl->th_log) * cdiv(B, 1 << pl->tn_log);
    int64_t b8 = -1; int a8 = 0;
    for (int a = 4; a >= 0; --a) {
      int64_t n = (int64_t)cdiv(S, 1 << a) * cdiv(S, 8) * cdiv(B, 1 << (4 - a));
      if (b8 < 0 || n < b8) { b8 = n; a8 = a; }
    }
    if (b8 * 10 <= best * 11) { pl->tw_log = a8; pl->th_log = 3; pl->tn_log = 4 - a8; }
  }
  const int TW = 1 << pl->tw_log, TH = 1 << pl->th_log, TN = 1 << pl->tn_log;
  if (TH > 16 || TW < 2 || TW + 2 > 256) return false;   // the depthwise warps work on column pairs
  pl->BN = N % 256 == 0 ? 256 : (N % 128 == 0 ? 128 : 64);
  pl->raw_tx = (TW + 2) * (TH + 2) * TN * 128;
  pl->raw_bytes = (pl->raw_tx + 1023) / 1024 * 1024;
  // Shared memory goes to the raw ring first: the depthwise groups hold two slots and only the slots beyond
  // those keep loads in flight.  Weights stay resident when all slabs fit in 64 KB, else a two-slab ring;
  // the store staging drops to one buffer and the A ring to two stages when that buys another raw slot.
  const int b_bytes = pl->BN * KSLAB * 2;
  const int nk = C / KSLAB;
  static const int raw_max = [] { const char* v = getenv("TMAT_DW_RAW"); int r = v ? atoi(v) : MAX_RAW; return r < 2 ? 2 : (r > MAX_RAW ? MAX_RAW : r); }();
  // two slots are in work (one per depthwise group), the rest are loads in flight.  With one epilogue warpgroup the
  // depth did not matter (2..6 measured the same); with two, the depthwise groups wait for raw tiles (ncu), so the
  // ring takes what is left after the other buffers, up to raw_want.
  static const int raw_want = [] { const char* v = getenv("TMAT_DW_RAWWANT"); int r = v ? atoi(v) : 4; return r < 2 ? 2 : (r > MAX_RAW ? MAX_RAW : r); }();
  pl->b_resident = (N == pl->BN && nk * b_bytes <= 64 * 1024 && nk <= MAX_B) ? 1 : 0;
  pl->n_b = pl->b_resident ? nk : 2;
  const int budget = 227 * 1024 - TAIL_BYTES - pl->n_b * b_bytes;
  // Options in order of preference: two epilogue warpgroups before one, three raw slots before two, three A stages
  // before two, full staging (two store tiles, or the staged tile + two pooled boxes, per warpgroup) before minimal.
  pl->n_raw = 0;
  static const int epi_max = [] { const char* v = getenv("TMAT_DW_EPI"); return v && atoi(v) == 1 ? 1 : 2; }();
  for (int n_epi = epi_max; n_epi >= 1 && pl->n_raw == 0; --n_epi) {
    for (int want = std::min(raw_want, raw_max); want >= 2 && pl->n_raw == 0; --want) {
      for (int n_a = 3; n_a >= 2 && pl->n_raw == 0; --n_a) {
        for (int full = 1; full >= 0 && pl->n_raw == 0; --full) {
          const int wg_bytes = pool ? STG_BYTES + (full ? 2 : 1) * POOL_BUF_BYTES : (full ? 2 : 1) * STG_BYTES;
          const int n_raw = (budget - n_a * A_STAGE_BYTES - n_epi * wg_bytes) / pl->raw_bytes;
          if (n_raw >= want) {
            pl->n_raw = want; pl->n_a = n_a; pl->n_epi = n_epi; pl->stg_wg_bytes = wg_bytes;
            pl->spw = full ? 2 : 1; pl->npb = full ? 2 : 1;
          }
        }
      }
    }
  }
  if (pl->n_raw < 2) return false;
  pl->smem = pl->n_a * A_STAGE_BYTES + pl->n_b * b_bytes + pl->n_raw * pl->raw_bytes + pl->n_epi * pl->stg_wg_bytes + TAIL_BYTES;
  return true;
}

template <int BN>
int launch_dwpw(tmat_ctx* ctx, const CUtensorMap& raw, const CUtensorMap& b, const CUtensorMap& out,
                const CUtensorMap& pool, const DwpwParams& p, int smem, cudaStream_t st) {
  static int configured = 0;
  if (configured < smem) {
    CUDA_TRY(cudaFuncSetAttribute(conv_dwpw_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = 227 * 1024;
  }
  int grid = std::min(p.total_tiles, ctx->sm_count);
  tmat_ctx::ProfRec rec{nullptr, nullptr, 0.0};
  if (ctx->profiling) {
    for (cudaEvent_t* e : {&rec.a, &rec.b}) {
      if (!ctx->event_pool.empty()) { *e = ctx->event_pool.back(); ctx->event_pool.pop_back(); }
      else CUDA_TRY(cudaEventCreate(e));
    }
    rec.flops = 2.0 * p.e.B * p.e.H * p.e.W * ((double)(p.tiles_n * BN) * (double)p.C + 9.0 * p.C);
    CUDA_TRY(cudaEventRecord(rec.a, st));
  }
  conv_dwpw_kernel<BN><<<grid, DW_THREADS, smem, st>>>(raw, b, out, pool, p);
  KERNEL_CHECK(ctx);
  if (ctx->profiling) {
    CUDA_TRY(cudaEventRecord(rec.b, st));
    ctx->prof.push_back(rec);
  }
  return TMAT_OK;
}

}  // namespace

bool dwpw_supported(int B, int S, int C, int N) {
  DwpwPlan pl;
  return plan_dwpw(B, S, C, N, false, &pl);
}

// The pooled epilogue needs whole even tiles inside the image, both staging tiles and a pooled box that fits half of one.
static bool pool_plan_ok(const DwpwPlan& pl, int S) {
  const int TW = 1 << pl.tw_log, TH = 1 << pl.th_log, TN = 1 << pl.tn_log;
  return S % 2 == 0 && TW >= 2 && TH >= 2 && S % TW == 0 && S % TH == 0 &&
         TN * (TW / 2 + 1) * (TH / 2 + 1) <= POOL_MAX_ROWS;
}

bool dwpw_pool_supported(int B, int S, int C, int N) {
  DwpwPlan pl;
  return plan_dwpw(B, S, C, N, true, &pl) && pool_plan_ok(pl, S);
}

int dwpw_run(tmat_ctx* ctx, const bf16* in, int B, int S, int C, const float* dw, int relu_in, const bf16* wmat,
             const float* shift, int N, int relu_out, bf16* out, cudaStream_t st, bf16* pool_out) {
  DwpwPlan pl;
  TMAT_REQUIRE(plan_dwpw(B, S, C, N, pool_out != nullptr, &pl), "dwpw: shape (B=%d, S=%d, C=%d, N=%d) not supported", B, S, C, N);
  TMAT_REQUIRE(pool_out == nullptr || pool_plan_ok(pl, S), "dwpw: pooled epilogue not supported for (B=%d, S=%d, C=%d, N=%d)",
               B, S, C, N);
  TMAT_REQUIRE(out != nullptr || pool_out != nullptr, "dwpw: no output");
  const int TW = 1 << pl.tw_log, TH = 1 << pl.th_log, TN = 1 << pl.tn_log;
  DwpwParams p;
  memset(&p, 0, sizeof(p));
  p.tw_log = pl.tw_log; p.th_log = pl.th_log; p.tn_log = pl.tn_log;
  p.tiles_x = cdiv(S, TW); p.tiles_y = cdiv(S, TH); p.tiles_b = cdiv(B, TN); p.tiles_n = N / pl.BN;
  int64_t total = (int64_t)p.tiles_x * p.tiles_y * p.tiles_b * p.tiles_n;
  TMAT_REQUIRE(total < (1ll << 31), "dwpw: too many tiles");
  p.total_tiles = (int)total;
  p.nk = C / KSLAB; p.C = C; p.relu_in = relu_in;
  p.raw_bytes = pl.raw_bytes; p.raw_tx = pl.raw_tx;
  p.n_raw = pl.n_raw; p.n_a = pl.n_a; p.n_b = pl.n_b; p.b_resident = pl.b_resident;
  p.dw = dw;
  p.n_epi = pl.n_epi; p.stg_wg_bytes = pl.stg_wg_bytes; p.spw = pl.spw; p.npb = pl.npb;
  memset(&p.e, 0, sizeof(p.e));
  p.e.mode = EPI_STORE; p.e.relu = relu_out; p.e.B = B; p.e.H = S; p.e.W = S; p.e.Cout = N; p.e.ppg = 1;
  p.e.shift = shift; p.e.out = out;
  CUtensorMap mRaw, mB, mOut, mPool;
  TMAT_TRY(tc_make_act_map(&mRaw, in, B, S, S, C, TW + 2, TH + 2, TN));
  TMAT_TRY(tc_make_weight_map(&mB, wmat, N, C, pl.BN));
  if (pool_out != nullptr) {
    // the pooled tensor starts at -inf: every tile max-reduces its (partial) box into it
    p.pool = 1; p.pw = TW / 2 + 1; p.ph = TH / 2 + 1;
    const int64_t n16 = (int64_t)B * (S / 2) * (S / 2) * N / 8;
    fill_neg_inf_bf16<<<(unsigned)cdiv64(n16, 256), 256, 0, st>>>(reinterpret_cast<uint4*>(pool_out), n16);
    KERNEL_CHECK(ctx);
    TMAT_TRY(tc_make_act_map(&mPool, pool_out, B, S / 2, S / 2, N, p.pw, p.ph, TN));
    mOut = mPool;
  } else {
    TMAT_TRY(tc_make_act_map(&mOut, out, B, S, S, N, TW, TH, TN));
    mPool = mOut;
  }
  switch (pl.BN) {
    case 256: return launch_dwpw<256>(ctx, mRaw, mB, mOut, mPool, p, pl.smem, st);
    case 128: return launch_dwpw<128>(ctx, mRaw, mB, mOut, mPool, p, pl.smem, st);
    case 64: return launch_dwpw<64>(ctx, mRaw, mB, mOut, mPool, p, pl.smem, st);
  }
  return TMAT_ERR_INVALID;
}
