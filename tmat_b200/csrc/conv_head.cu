// Head of the UNet (models.py:165-166: Conv2D(1, 3, 'same', sigmoid) on the 2x-upsampled last block) with the TAPS IN THE
// GEMM'S N DIMENSION.
//
// Evaluated on the low-resolution tensor Y (B, S, S, 64), output parity q = (py, px) of fine pixel (2y + py, 2x + px) sees
// 2 x 2 low-resolution pixels (unet.cu: parity_lowtap), i.e. 4 parities x 4 taps = exactly 16 (parity, tap) weight vectors
// of 64 channels.  conv_halo_kernel<16> ran this as a 3x3 convolution with N = 4 (padded to 16): 9 taps x 4 K-steps = 36
// MMAs of M = 128 x N = 16 per 128-pixel tile, each re-reading the 16 KB A tile from shared memory (32 cycles per MMA at
// 128 B/clk -- that, not the tensor pipe, bounded the layer: 95 us per 64 patches, tensor pipe 15 % active).  Here
//     P[r][(q, t)] = sum_c Y[r][c] * W[(q, t)][c]
// is ONE 1x1 GEMM over the haloed 10 x 18-pixel tile (two M = 128 blocks x 4 K-steps = 8 MMAs, all 16 columns useful) and
// the output is a 4-term shifted sum per fine pixel in the epilogue:
//     out[2y + py][2x + px] = bias + sum_{a, b in {0,1}} P[(y + dy_a(py), x + dx_b(px))][(q, 2a + b)],
//     dy_0(0) = -1, dy_1(0) = 0, dy_0(1) = 0, dy_1(1) = +1 (same for dx).
// Same bf16 weights (the fine taps that fall on one low-resolution pixel are summed before rounding, as before), fp32
// accumulation; only the order of the fp32 additions differs from the 9-tap formulation.
//
//   warp 0      TMA producer: the (parity, tap) matrix once (2 KB), then one haloed raw tile per output tile (ring)
//   warp 1      MMA issuer; owns TMEM (one accumulator buffer per epilogue warpgroup: 2 M-blocks x 16 columns each)
//   warps 2..17  four epilogue warpgroups taking every fourth tile (the epilogue's latency chain, not its instruction
//               count, bounded the layer with two): TMEM -> P tile in shared memory (256 rows x 16 fp32)
//               -> 16 gathered values per low-resolution pixel -> bias, sigmoid -> 2 x 2 float32 outputs
#include "tc_ptx.cuh"
#include <algorithm>
#include <stdlib.h>

using namespace tcptx;

namespace {

constexpr int HD_WGS = 4;                             // epilogue warpgroups = accumulator buffers (tile it -> warpgroup it % HD_WGS)
constexpr int HD_THREADS = 64 + 128 * HD_WGS;
constexpr int HD_TW = 8, HD_TH = 16;                 // low-resolution pixels per tile
constexpr int HD_PITCH = HD_TW + 2;                  // raw row pitch in pixels
constexpr int HD_RAW_TX = HD_PITCH * (HD_TH + 2) * 128;          // 23040 bytes delivered per raw tile
constexpr int HD_RAW_BYTES = (HD_RAW_TX + 1023) / 1024 * 1024;   // ring slot
constexpr int HD_NRAW = 4;
constexpr int HD_B_BYTES = 16 * KSLAB * 2;           // 2 KB
constexpr int HD_P_ROW = 80;                         // bytes per P row (16 fp32 + 16 bytes of padding against bank conflicts)
constexpr int HD_P_BYTES = 256 * HD_P_ROW;           // per epilogue warpgroup
// the second M-block of the last ring slot reads 256 x 128 bytes from the slot's start, i.e. past the slot into the P area
static_assert(256 * 128 - HD_RAW_BYTES <= HD_WGS * HD_P_BYTES, "M-block 1 of the last raw slot must stay inside the allocation");
constexpr int HD_SMEM = 2048 + HD_NRAW * HD_RAW_BYTES + HD_WGS * HD_P_BYTES + 256 + 1024;
static_assert(HD_SMEM <= 227 * 1024 && (HD_WGS == 2 || HD_WGS == 4), "shared memory / TMEM columns");

struct HeadParams {
  int tiles_x, tiles_y, total_tiles;
  int B, H, W, logits;
  const float* shift;      // [4] bias per parity
  float* out;              // (B, 2H, 2W) float32
};

__global__ void __launch_bounds__(HD_THREADS, 1)
head_taps_kernel(const __grid_constant__ CUtensorMap tmRaw, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ HeadParams p) {
  constexpr int TMEM_COLS = 32 * HD_WGS;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_base = smem_base;
  const uint32_t raw_base = smem_base + 2048;
  const uint32_t p_base = raw_base + HD_NRAW * HD_RAW_BYTES;
  const uint32_t bar_base = p_base + HD_WGS * HD_P_BYTES;
  auto rfull_bar = [&](int s) { return bar_base + 8u * s; };
  auto rempty_bar = [&](int s) { return bar_base + 8u * (HD_NRAW + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * HD_NRAW + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * HD_NRAW + HD_WGS + s); };
  const uint32_t bfull_bar = bar_base + 8u * (2 * HD_NRAW + 2 * HD_WGS);
  constexpr int TMEM_SLOT = 8 * (2 * HD_NRAW + 2 * HD_WGS + 1);
  const uint32_t tmem_slot = bar_base + TMEM_SLOT;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + (bar_base - smem_base) + TMEM_SLOT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < HD_NRAW; ++s) { mbar_init(rfull_bar(s), 1); mbar_init(rempty_bar(s), 1); }
    for (int s = 0; s < HD_WGS; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 128); }
    mbar_init(bfull_bar, 1);
    fence_barrier_init();
    fence_proxy_async();
    tma_prefetch_desc(&tmRaw);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  const int tiles_xy = p.tiles_x * p.tiles_y;
  auto decode = [&](int tile, int& x0, int& y0, int& b0) {
    const int tb = tile / tiles_xy, rr = tile - tb * tiles_xy;
    const int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
    x0 = tx * HD_TW; y0 = ty * HD_TH; b0 = tb;
  };
  const int my_tiles = (p.total_tiles > (int)blockIdx.x) ? (p.total_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == 0) {
    // ================= TMA producer =================
    if (my_tiles > 0 && lane == 0) {
      mbar_expect_tx(bfull_bar, HD_B_BYTES);
      tma_load_2d(b_base, &tmB, bfull_bar, 0, 0);
    }
    __syncwarp();
    int rs = 0; uint32_t rphase = 0;
    for (int it = 0; it < my_tiles; ++it) {
      mbar_wait(rempty_bar(rs), rphase ^ 1);
      if (lane == 0) {
        int x0, y0, b0;
        decode((int)blockIdx.x + it * (int)gridDim.x, x0, y0, b0);
        mbar_expect_tx(rfull_bar(rs), HD_RAW_TX);
        tma_load_4d(raw_base + rs * HD_RAW_BYTES, &tmRaw, rfull_bar(rs), 0, x0 - 1, y0 - 1, b0);
      }
      __syncwarp();
      if (++rs == HD_NRAW) { rs = 0; rphase ^= 1; }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (my_tiles > 0) {
      constexpr uint32_t idesc = umma_idesc<16>();
      const uint64_t bdesc = umma_desc_sw128(b_base);
      int rs = 0; uint32_t rphase = 0;
      mbar_wait(bfull_bar, 0);
      for (int it = 0; it < my_tiles; ++it) {
        const int as = it % HD_WGS;
        const uint32_t aphase = (it / HD_WGS) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        mbar_wait(rfull_bar(rs), rphase);
        tc_fence_after();
        const uint32_t raw = raw_base + rs * HD_RAW_BYTES;
        const uint32_t tmem_d = tmem_base + as * 32;
        umma_k64_elect(tmem_d, umma_desc_sw128(raw), bdesc, idesc, 0u);                      // raw pixels 0..127
        umma_k64_elect(tmem_d + 16, umma_desc_sw128(raw + 128 * 128), bdesc, idesc, 0u);     // raw pixels 128..255 (180.. unused)
        umma_commit_elect(rempty_bar(rs));
        umma_commit_elect(tfull_bar(as));
        if (++rs == HD_NRAW) { rs = 0; rphase ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue =================
    const int ewg = (warp - 2) >> 2;
    const uint32_t bar_id = 1u + (uint32_t)ewg;
    auto epi_sync = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory"); };
    const int q = warp & 3;
    const int row = q * 32 + lane;                       // TMEM lane = raw pixel (mod 128); also this thread's low-res pixel
    const int tx_in = row & (HD_TW - 1), ty_in = row >> 3;
    const uint32_t P = p_base + (uint32_t)ewg * HD_P_BYTES;
    const float sh0 = __ldg(p.shift + 0), sh1 = __ldg(p.shift + 1), sh2 = __ldg(p.shift + 2), sh3 = __ldg(p.shift + 3);
    for (int it = ewg; it < my_tiles; it += HD_WGS) {
      const int as = ewg;                                // tile it -> accumulator buffer it % HD_WGS
      const uint32_t aphase = (it / HD_WGS) & 1;
      int x0, y0, b0;
      decode((int)blockIdx.x + it * (int)gridDim.x, x0, y0, b0);
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * 32;
      uint32_t r0[16], r1[16];
      tmem_ld16(taddr, r0);
      tmem_ld16(taddr + 16, r1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(tempty_bar(as));
      epi_sync();                                        // everyone has finished reading the previous tile's P
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        st_shared_v4(P + (uint32_t)row * HD_P_ROW + 16u * j, make_uint4(r0[4 * j], r0[4 * j + 1], r0[4 * j + 2], r0[4 * j + 3]));
        st_shared_v4(P + (uint32_t)(128 + row) * HD_P_ROW + 16u * j, make_uint4(r1[4 * j], r1[4 * j + 1], r1[4 * j + 2], r1[4 * j + 3]));
      }
      epi_sync();
      const int x = x0 + tx_in, y = y0 + ty_in;
      if (x < p.W && y < p.H) {
        // raw pixel of low-res (ty_in + dy, tx_in + dx) in the haloed tile: (ty_in + 1 + dy) * PITCH + tx_in + 1 + dx
        float o[4];
#pragma unroll
        for (int par = 0; par < 4; ++par) {
          const int py = par >> 1, px = par & 1;
          float acc[4];
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int dy = (t >> 1) + py - 1, dx = (t & 1) + px - 1;        // {-1, 0} for parity 0, {0, +1} for parity 1
            const uint32_t rp = (uint32_t)((ty_in + 1 + dy) * HD_PITCH + tx_in + 1 + dx);
            float v;
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(P + rp * HD_P_ROW + 4u * (uint32_t)(par * 4 + t)));
            acc[t] = v;
          }
          o[par] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
        }
        o[0] += sh0; o[1] += sh1; o[2] += sh2; o[3] += sh3;
        if (!p.logits) {
#pragma unroll
          for (int i = 0; i < 4; ++i) o[i] = sigmoid_f(o[i]);
        }
        float* dst = p.out + ((int64_t)b0 * 2 * p.H + 2 * y) * (2 * p.W) + 2 * x;
        *reinterpret_cast<float2*>(dst) = make_float2(o[0], o[1]);
        *reinterpret_cast<float2*>(dst + 2 * p.W) = make_float2(o[2], o[3]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

int g_head_taps = -1;
int head_taps_mode() {      // TMAT_HEAD_TAPS=0: the head stays on conv_halo_kernel<16> (3x3 formulation)
  if (g_head_taps < 0) {
    const char* e = getenv("TMAT_HEAD_TAPS");
    g_head_taps = e ? atoi(e) : 1;
  }
  return g_head_taps;
}

}  // namespace

bool conv_head_taps_supported(const ConvOp& op) {
  return head_taps_mode() == 1 && op.epi == EPI_HEAD && op.wmat2 != nullptr && op.a0_C == KSLAB && op.a1 == nullptr &&
         op.a0_stride == 1 && op.out_f32 != nullptr;
}

int conv_run_head_taps(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st) {
  TMAT_REQUIRE(conv_head_taps_supported(op), "conv_head(%s): not a head op of 64 channels", op.name);
  HeadParams p;
  memset(&p, 0, sizeof(p));
  p.tiles_x = cdiv(op.W, HD_TW); p.tiles_y = cdiv(op.H, HD_TH);
  const int64_t total = (int64_t)p.tiles_x * p.tiles_y * op.B;
  TMAT_REQUIRE(total < (1ll << 31), "conv_head(%s): too many tiles", op.name);
  p.total_tiles = (int)total;
  p.B = op.B; p.H = op.H; p.W = op.W; p.logits = op.logits;
  p.shift = op.shift; p.out = op.out_f32;
  CUtensorMap mRaw, mB;
  TMAT_TRY(tc_make_act_map(&mRaw, op.a0, op.B, op.H, op.W, op.a0_C, HD_PITCH, HD_TH + 2, 1));
  TMAT_TRY(tc_make_weight_map(&mB, op.wmat2, 16, KSLAB, 16));
  static bool configured = false;
  if (!configured) {
    CUDA_TRY(cudaFuncSetAttribute(head_taps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HD_SMEM));
    configured = true;
  }
  const int grid = std::min(p.total_tiles, ctx->sm_count);
  tmat_ctx::ProfRec rec{nullptr, nullptr, 0.0};
  if (ctx->profiling) {
    for (cudaEvent_t* e : {&rec.a, &rec.b}) {
      if (!ctx->event_pool.empty()) { *e = ctx->event_pool.back(); ctx->event_pool.pop_back(); }
      else CUDA_TRY(cudaEventCreate(e));
    }
    rec.flops = 2.0 * (double)p.total_tiles * 256.0 * 16.0 * (double)KSLAB;      // what the tensor pipe executes
    CUDA_TRY(cudaEventRecord(rec.a, st));
  }
  head_taps_kernel<<<grid, HD_THREADS, HD_SMEM, st>>>(mRaw, mB, p);
  KERNEL_CHECK(ctx);
  if (ctx->profiling) {
    CUDA_TRY(cudaEventRecord(rec.b, st));
    ctx->prof.push_back(rec);
  }
  return TMAT_OK;
}
