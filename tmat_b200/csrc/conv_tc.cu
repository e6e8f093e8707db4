// tcgen05 / TMEM / TMA implicit-GEMM executor of a ConvOp (see unet.cuh) for sm_100a.
//
//   D[128 pixels, BN channels] (fp32, TMEM) = sum over k-steps  A_k[128, 64] (bf16, smem) * B_k[BN, 64]^T
//
// One k-step = one filter tap x one 64-channel slab.  The A tile of a k-step is a 4-D TMA box
// (64 channels, TW x TH pixels, TN images) of the NHWC activation tensor whose origin is shifted by
// the tap offset; out-of-image pixels are zero-filled by the TMA unit, which implements the 'same'
// padding of the reference's Conv2D / Conv2DTranspose (models.py:119-166) without an im2col buffer.
// The B tile is a 2-D TMA box of the packed weight matrix [Npad][K].  Both land in shared memory in
// the 128-byte-swizzled K-major layout that the UMMA shared-memory descriptors describe.
//
// A ConvOp may carry up to 4 column groups (EPI_PAR: the output-pixel parities of a 3x3 convolution
// on a 2x nearest-upsampled tensor, evaluated on the low-resolution tensor): every group has its own
// tap list and its own rows of the weight matrix, and shares the A tensor.
//
// Warp roles (192 threads, persistent over output tiles):
//   warp 0      TMA producer       (ring of STAGES stages, full/empty mbarriers)
//   warp 1      MMA issuer         (one lane issues tcgen05.mma, tcgen05.commit frees stages);
//               also owns the TMEM allocation (2 accumulator buffers of BN columns)
//   warps 2..5  epilogue           (tcgen05.ld -> +shift / +addend / ReLU -> bf16 -> 128B-swizzled smem
//                                   staging tile -> one TMA tensor store per 64 channels), overlapped
//                                   with the next tile's MMAs.  The TMA store clips partial tiles and
//                                   writes full 128-byte lines, so the epilogue issues no per-thread
//                                   global stores (the head, 4 floats per pixel, stores directly).
#include "tc_ptx.cuh"
#include <algorithm>

using namespace tcptx;

namespace {

constexpr int MAX_KSTEPS = 160;                    // over all groups
constexpr int TC_THREADS = 320;                    // producer, MMA, two epilogue warpgroups (warps 2..5, 6..9)

struct TcParams {
  int tw_log, th_log, tn_log;   // tile = 2^tw x 2^th pixels x 2^tn images (product 128)
  int tiles_x, tiles_y, tiles_b, tiles_n, ngroups, total_tiles;
  int nk;                       // k-steps per group
  int npad;                     // GEMM columns per group
  int n_epi;                    // epilogue warpgroups in use: 2 (alternate tiles, one staging tile each) or 1
  uint16_t kstep[MAX_KSTEPS];   // [group][nk]: bit0 src | bits1-2 dy+1 | bits3-4 dx+1 | bits5.. c0/64
  EpiParams e;
};

// ---- the kernel -------------------------------------------------------------------------------
template <int BN, int STAGES>
struct TcSmem {
  static constexpr int B_STAGE_BYTES = BN * KSLAB * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STG_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int BAR_OFFSET = STG_OFFSET + N_STG * STG_BYTES;
  static constexpr int SHIFT_OFFSET = BAR_OFFSET + 256;     // BN fp32 shift values of the current tile
  static constexpr int SHIFT_BYTES = (BN < 64 ? 64 : BN) * 4;
  static constexpr int EPI_WGS = (BN >= 64) ? 2 : 1;
  // BN = 256 has no room for a second shift area: its two warpgroups share one, which is only used when every tile
  // has the same column tile (tiles_n == 1), i.e. both write the same values
  static constexpr int SHIFT_AREAS = (BN == 128 || BN == 64) ? 2 : 1;
  static constexpr int BYTES = SHIFT_OFFSET + SHIFT_AREAS * SHIFT_BYTES + 1024;  // barriers + shift + alignment slack
  static_assert(BYTES <= 227 * 1024, "shared memory budget");
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ TcOutMaps om,
               const __grid_constant__ TcParams p) {
  using L = TcSmem<BN, STAGES>;
  constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : 2 * BN;   // power of two >= 32
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + L::BAR_OFFSET;
  // barrier slots (8 bytes each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then tmem ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + L::BAR_OFFSET + 8 * (2 * STAGES + 4));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 128); }
    fence_barrier_init();
    fence_proxy_async();
    tma_prefetch_desc(&tmA0);
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  const int TW = 1 << p.tw_log, TH = 1 << p.th_log;
  const int tiles_xy = p.tiles_x * p.tiles_y;
  const int tiles_gn = p.ngroups * p.tiles_n;
  // tile index -> (pixel tile mt, group g, column tile nt); nt fastest so neighbouring CTAs share A in L2
  auto decode = [&](int tile, int& g, int& nt, int& x0, int& y0, int& b0) {
    int mt = tile / tiles_gn, r = tile - mt * tiles_gn;
    g = r / p.tiles_n; nt = r - g * p.tiles_n;
    int tb = mt / tiles_xy, rr = mt - tb * tiles_xy;
    int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
    x0 = tx << p.tw_log; y0 = ty << p.th_log; b0 = tb << p.tn_log;
  };

  if (warp == 0) {
    // ================= TMA producer =================
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      int g, nt, x0, y0, b0;
      decode(tile, g, nt, x0, y0, b0);
      const uint16_t* ktab = p.kstep + g * p.nk;
      const int brow = g * p.npad + nt * BN;
      for (int ks = 0; ks < p.nk; ++ks) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        if (lane == 0) {
          uint32_t kd = ktab[ks];
          int src = kd & 1, dy = (int)((kd >> 1) & 3) - 1, dx = (int)((kd >> 3) & 3) - 1, c0 = (int)(kd >> 5) * KSLAB;
          uint32_t sa = smem_base + stage * L::STAGE_BYTES;
          uint32_t sb = sa + A_STAGE_BYTES;
          mbar_expect_tx(full_bar(stage), L::STAGE_BYTES);
          tma_load_4d(sa, src ? &tmA1 : &tmA0, full_bar(stage), c0, x0 + dx, y0 + dy, b0);
          tma_load_2d(sb, &tmB, full_bar(stage), ks * KSLAB, brow);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (warp-convergent loop, one elected lane issues) =================
    {
      constexpr uint32_t idesc = umma_idesc<BN>();
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int ks = 0; ks < p.nk; ++ks) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * L::STAGE_BYTES;
          // the four K = 16 MMAs of the stage: +32 bytes along K inside the swizzle atom per MMA
          umma_k64_elect(tmem_d, umma_desc_sw128(sa), umma_desc_sw128(sa + A_STAGE_BYTES), idesc, (uint32_t)(ks != 0));
          umma_commit_elect(empty_bar(stage));                     // stage reusable once these MMAs retire
          if (ks == p.nk - 1) umma_commit_elect(tfull_bar(as));    // accumulator complete
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5, and 6..9 on alternate tiles when p.n_epi == 2: the layers with few
    // k-steps per tile are bound by the epilogue's single dependent instruction stream per scheduler) =================
    const int ewg = (warp - 2) >> 2;
    const bool two = (p.n_epi == 2);
    const int q = warp & 3;                   // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;            // GEMM row = pixel within the tile
    const int tx_in = row & (TW - 1);
    const int ty_in = (row >> p.tw_log) & (TH - 1);
    const int tb_in = row >> (p.tw_log + p.th_log);
    const bool leader = (threadIdx.x == 64 + 128 * ewg);  // issues, commits and waits on every TMA store
    const uint32_t stg_base = smem_base + L::STG_OFFSET + (two ? (uint32_t)(ewg * STG_BYTES) : 0u);
    const uint32_t shift_smem = smem_base + L::SHIFT_OFFSET + ((two && L::SHIFT_AREAS == 2) ? (uint32_t)(ewg * L::SHIFT_BYTES) : 0u);
    const uint32_t stg_row = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const EpiParams& e = p.e;
    int sbuf = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      if (ewg >= p.n_epi || (two && (it & 1) != ewg)) continue;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      int g, nt, x0, y0, b0;
      decode(tile, g, nt, x0, y0, b0);
      const int x = x0 + tx_in, y = y0 + ty_in, b = b0 + tb_in;
      const bool valid = (x < e.W) && (y < e.H) && (b < e.B);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
      if (BN >= 64 && e.add_src != nullptr && tile + 2 * (int)gridDim.x < p.total_tiles) {
        // addend rows of the tile after next -> L2
        int gn, ntn, x0n, y0n, b0n;
        decode(tile + 2 * (int)gridDim.x, gn, ntn, x0n, y0n, b0n);
        tc_prefetch_addend(e, ntn * BN, BN, x0n + tx_in, y0n + ty_in, b0n + tb_in);
      }
      tc_epilogue_tile<BN>(e, om, taddr, g, nt, x0, y0, b0, x, y, b, valid, stg_base, stg_row, sw, leader, sbuf,
                           tempty_bar(as), tfull_bar(as), aphase, shift_smem, 0, 0, 1u + (uint32_t)ewg, two);
    }
    if (leader) tma_store_wait_all();            // smem must outlive the last store's read
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

}  // namespace

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// NHWC activation tensor (B, H, W, C) bf16 -> 4-D map, box = (64 ch, TW, TH, TN), 128B swizzle, zero OOB fill
// (sx, sy = pixel strides in units of pixels of the underlying dense tensor of row length Wfull, height Hfull:
//  the parity views of EPI_PAR use sx = sy = 2)
int tc_make_act_map(CUtensorMap* m, const bf16* base, int B, int H, int W, int C, int TW, int TH, int TN,
                    int sx, int sy) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { tmat_set_error("conv_tc: cuTensorMapEncodeTiled unavailable"); return TMAT_ERR_CUDA; }
  const cuuint64_t Wfull = (cuuint64_t)W * sx, Hfull = (cuuint64_t)H * sy;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2 * sx, Wfull * C * 2 * sy, Hfull * Wfull * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)KSLAB, (cuuint32_t)TW, (cuuint32_t)TH, (cuuint32_t)TN};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    tmat_set_error("conv_tc: cuTensorMapEncodeTiled(activation %dx%dx%dx%d, box %d,%d,%d) failed: %d", B, H, W, C, TW,
                   TH, TN, (int)r);
    return TMAT_ERR_CUDA;
  }
  return TMAT_OK;
}

int tc_make_weight_map(CUtensorMap* m, const bf16* base, int Npad, int K, int BN) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { tmat_set_error("conv_tc: cuTensorMapEncodeTiled unavailable"); return TMAT_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Npad};
  cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {(cuuint32_t)KSLAB, (cuuint32_t)BN};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    tmat_set_error("conv_tc: cuTensorMapEncodeTiled(weights %dx%d) failed: %d", Npad, K, (int)r);
    return TMAT_ERR_CUDA;
  }
  return TMAT_OK;
}

// pick the 128-pixel tile (2^a x 2^b pixels x 2^c images) that covers (B, H, W) with the fewest tiles
void tc_choose_tile(int B, int H, int W, int* tw_log, int* th_log, int* tn_log) {
  int64_t best = -1;
  for (int a = 7; a >= 0; --a)
    for (int b = 7 - a; b >= 0; --b) {
      int c = 7 - a - b;
      int64_t n = (int64_t)cdiv(W, 1 << a) * cdiv(H, 1 << b) * cdiv(B, 1 << c);
      // tie-break: fewer images per tile, then wider tiles (longer contiguous runs for the TMA unit)
      if (best < 0 || n < best) { best = n; *tw_log = a; *th_log = b; *tn_log = c; }
    }
}

bool conv_tc_supported(const ConvOp& op) {
  if (op.a0_C % KSLAB != 0 || (op.a1 && op.a1_C % KSLAB != 0)) return false;
  if (op.ngroups < 1 || op.ngroups > CONV_MAX_GROUPS || (op.ngroups > 1 && op.a1)) return false;
  if (op.a0_stride != 1 && (op.ntaps != 1 || op.ngroups != 1 || op.dy[0][0] != 0 || op.dx[0][0] != 0)) return false;
  int nk = op.ntaps * (op.a0_C / KSLAB) + (op.a1 ? op.a1_C / KSLAB : 0);
  if (nk * op.ngroups > MAX_KSTEPS) return false;
  if (op.epi == EPI_HEAD) return op.Npad == 16 && op.ngroups == 1;
  if (op.Cout % 64 != 0) return false;
  if (op.epi == EPI_PAR) return op.Npad % op.Cout == 0 && op.ngroups * (op.Npad / op.Cout) == 4 && !op.out2 && !op.add_src;
  return op.Npad == op.Cout && op.ngroups == 1;
}

template <int BN, int STAGES>
static int launch_tc(tmat_ctx* ctx, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                     const TcOutMaps& om, const TcParams& p, cudaStream_t st) {
  using L = TcSmem<BN, STAGES>;
  static bool configured = false;
  if (!configured) {
    CUDA_TRY(cudaFuncSetAttribute(conv_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
    configured = true;
  }
  int grid = std::min(p.total_tiles, ctx->sm_count);
  // two epilogue warpgroups share the staging pair (one tile each): not with dual stores, which need both tiles; at
  // BN = 256 only for layers with few k-steps per tile (the MMA-bound ones measured 1.5 % slower with it)
  static const int epi_max = [] { const char* v = getenv("TMAT_TC_EPI"); return v ? atoi(v) : 3; }();   // 1: one warpgroup; 2: two for BN <= 128; 3: also BN = 256
  TcParams pp = p;
  pp.n_epi = (L::EPI_WGS == 2 && p.e.out2 == nullptr && epi_max >= 2 && (L::SHIFT_AREAS == 2 || (p.tiles_n == 1 && p.nk <= 8 && epi_max == 3))) ? 2 : 1;
  tmat_ctx::ProfRec rec{nullptr, nullptr, 0.0};
  if (ctx->profiling) {
    for (cudaEvent_t* e : {&rec.a, &rec.b}) {
      if (!ctx->event_pool.empty()) { *e = ctx->event_pool.back(); ctx->event_pool.pop_back(); }
      else CUDA_TRY(cudaEventCreate(e));
    }
    rec.flops = 2.0 * p.e.B * p.e.H * p.e.W * (double)(p.ngroups * p.tiles_n * BN) * (double)(p.nk * KSLAB);
    CUDA_TRY(cudaEventRecord(rec.a, st));
  }
  conv_tc_kernel<BN, STAGES><<<grid, TC_THREADS, L::BYTES, st>>>(a0, a1, b, om, pp);
  KERNEL_CHECK(ctx);
  if (ctx->profiling) {
    CUDA_TRY(cudaEventRecord(rec.b, st));
    ctx->prof.push_back(rec);
  }
  return TMAT_OK;
}

int conv_run_tc(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st) {
  if (conv_head_taps_supported(op)) return conv_run_head_taps(ctx, op, st);
  if (conv_halo_supported(op)) return conv_run_halo(ctx, op, conv_halo_mode(), st);
  TMAT_REQUIRE(conv_tc_supported(op), "conv_tc(%s): shape not supported by the tcgen05 path", op.name);
  TcParams p;
  memset(&p, 0, sizeof(p));
  tc_choose_tile(op.B, op.H, op.W, &p.tw_log, &p.th_log, &p.tn_log);
  const int TW = 1 << p.tw_log, TH = 1 << p.th_log, TN = 1 << p.tn_log;
  int BN = op.epi == EPI_HEAD ? 16 : (op.Npad % 256 == 0 ? 256 : (op.Npad % 128 == 0 ? 128 : 64));
  p.tiles_x = cdiv(op.W, TW);
  p.tiles_y = cdiv(op.H, TH);
  p.tiles_b = cdiv(op.B, TN);
  p.tiles_n = op.Npad / BN;
  p.ngroups = op.ngroups;
  p.npad = op.Npad;
  int64_t total = (int64_t)p.tiles_x * p.tiles_y * p.tiles_b * p.tiles_n * p.ngroups;
  TMAT_REQUIRE(total < (1ll << 31), "conv_tc(%s): too many tiles", op.name);
  p.total_tiles = (int)total;
  int nk = 0;
  for (int g = 0; g < op.ngroups; ++g) {
    nk = 0;
    uint16_t* tab = p.kstep + g * (op.K / KSLAB);
    for (int t = 0; t < op.ntaps; ++t)
      for (int c = 0; c < op.a0_C / KSLAB; ++c)
        tab[nk++] = (uint16_t)(0 | ((op.dy[g][t] + 1) << 1) | ((op.dx[g][t] + 1) << 3) | (c << 5));
    if (op.a1)
      for (int c = 0; c < op.a1_C / KSLAB; ++c) tab[nk++] = (uint16_t)(1 | (1 << 1) | (1 << 3) | (c << 5));
  }
  p.nk = nk;
  TMAT_REQUIRE(nk * KSLAB == op.K, "conv_tc(%s): K mismatch (%d k-steps, K=%d)", op.name, nk, op.K);
  p.e = epi_of(op);

  CUtensorMap mA0, mA1, mB;
  TcOutMaps om;
  TMAT_TRY(tc_make_act_map(&mA0, op.a0, op.B, op.H, op.W, op.a0_C, TW, TH, TN, op.a0_stride, op.a0_stride));
  if (op.a1) TMAT_TRY(tc_make_act_map(&mA1, op.a1, op.B, op.H, op.W, op.a1_C, TW, TH, TN));
  else mA1 = mA0;
  TMAT_TRY(tc_make_weight_map(&mB, op.wmat, op.ngroups * op.Npad, op.K, BN));
  if (op.epi == EPI_STORE) {
    TMAT_TRY(tc_make_act_map(&om.m[0], op.out, op.B, op.H, op.W, op.Cout, TW, TH, TN));
    om.m[1] = om.m[2] = om.m[3] = om.m[0];
    if (op.out2) TMAT_TRY(tc_make_act_map(&om.m2, op.out2, op.B, op.H, op.W, op.Cout, TW, TH, TN));
    else om.m2 = om.m[0];
  } else if (op.epi == EPI_PAR) {
    for (int q = 0; q < 4; ++q) {
      const bf16* base = op.out + ((int64_t)(q >> 1) * 2 * op.W + (q & 1)) * op.Cout;
      TMAT_TRY(tc_make_act_map(&om.m[q], base, op.B, op.H, op.W, op.Cout, TW, TH, TN, 2, 2));
    }
    om.m2 = om.m[0];
  } else {
    om.m[0] = om.m[1] = om.m[2] = om.m[3] = om.m2 = mA0;   // unused by the head
  }
  switch (BN) {
    case 256: return launch_tc<256, 4>(ctx, mA0, mA1, mB, om, p, st);
    case 128: return launch_tc<128, 6>(ctx, mA0, mA1, mB, om, p, st);
    case 64: return launch_tc<64, 8>(ctx, mA0, mA1, mB, om, p, st);
    case 16: return launch_tc<16, 8>(ctx, mA0, mA1, mB, om, p, st);
  }
  return TMAT_ERR_INVALID;
}
