// tcgen05 3x3 'same' convolution with the A operand loaded ONCE per tile: the haloed input tile
// (64 channels x RW x 18 pixels, RW >= 10) is brought into shared memory by one TMA load per 64-channel
// slab, and the nine taps are nine UMMA shared-memory descriptors into that same buffer -- tap (ky, kx)
// starts (ky * RW + kx) pixels = 128-byte rows further on, the 8-row groups of the M = 128 = 8 x 16 pixel
// tile are RW rows apart (descriptor stride byte offset = RW * 128).  Compared with conv_tc.cu, which
// re-loads the shifted 16 KB tile for every tap, the A traffic from L2 drops 9-fold (6.3-fold counting the
// halo), which is what bounds the layers with few output channels (N <= 128: arithmetic intensity per
// A byte is N / 2 flop per byte per tap).  The 128-byte swizzle is a function of the shared-memory address
// bits (TMA writes it that way and UMMA reads it that way), so descriptors that start at any 128-byte row
// of the 1024-byte-aligned buffer see consistent data.
//
//   warp 0      TMA producer: raw tiles (ring, issued one or two items ahead) and, unless the whole
//               weight matrix of the layer is resident in shared memory, one B tile per tap (ring)
//   warp 1      MMA issuer; owns TMEM (2 accumulators of BN columns)
//   warps 2..5  epilogue (tc_epilogue_tile: staging + TMA store, or the head's direct store)
#include "tc_ptx.cuh"
#include <algorithm>
#include <stdlib.h>

using namespace tcptx;

namespace {

constexpr int HALO_THREADS = 320;                // producer, MMA, two epilogue warpgroups (warps 2..5, 6..9) on alternate tiles
constexpr int HALO_EPI_WGS = 2;
constexpr int HTW = 8, HTH = 16;                 // tile: 8 x 16 pixels of one image
constexpr int MAX_STAGES = 8;
constexpr int MAX_RAW = 4;
constexpr int MAX_Z = 4;                          // addend (half-resolution Z) tiles staged by the producer
constexpr int PITCH = HTW + 2;                    // raw row pitch in pixels = box width of the raw load

struct HaloParams {
  int tiles_x, tiles_y, tiles_b, tiles_n, total_tiles;
  int nslab, Cin;
  int raw_bytes, raw_tx, n_raw, n_stage;
  int n_z;           // > 0: ring of addend tiles (add_shift == 1): 4 x 8 pixels x 64 channels per 64 output columns
  EpiParams e;
};

__device__ __forceinline__ uint64_t umma_desc_sw128_ex(uint32_t saddr, uint32_t sbo_bytes, uint32_t base_off) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(sbo_bytes >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(base_off & 7u) << 49;
  d |= (uint64_t)2 << 61;
  return d;
}

template <int BN, bool BRES>
__global__ void __launch_bounds__(HALO_THREADS, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tmRaw, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmZ, const __grid_constant__ TcOutMaps om, const __grid_constant__ HaloParams p) {
  constexpr int B_TILE_BYTES = BN * KSLAB * 2;
  constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : 2 * BN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int n_btiles = BRES ? 9 * p.nslab : p.n_stage;
  const uint32_t b_base = smem_base;
  const uint32_t raw_base = (b_base + n_btiles * B_TILE_BYTES + 1023u) & ~1023u;
  const uint32_t stg_base = raw_base + p.n_raw * p.raw_bytes;
  constexpr int Z_BYTES = (BN >= 64 ? BN / 64 : 1) * 4096;
  const uint32_t z_base = stg_base + HALO_EPI_WGS * N_STG * STG_BYTES;
  const uint32_t bar_base = z_base + p.n_z * Z_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (MAX_STAGES + s); };
  auto rfull_bar = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + s); };
  auto rempty_bar = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + MAX_RAW + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + 2 * MAX_RAW + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + 2 * MAX_RAW + 2 + s); };
  const uint32_t bres_bar = bar_base + 8u * (2 * MAX_STAGES + 2 * MAX_RAW + 4);
  auto zfull_bar = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + 2 * MAX_RAW + 5 + s); };
  auto zempty_bar = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + 2 * MAX_RAW + 5 + MAX_Z + s); };
  constexpr int TMEM_SLOT = 8 * (2 * MAX_STAGES + 2 * MAX_RAW + 5 + 2 * MAX_Z);
  const uint32_t tmem_slot = bar_base + TMEM_SLOT;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (bar_base - smem_base) + TMEM_SLOT);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < MAX_RAW; ++s) { mbar_init(rfull_bar(s), 1); mbar_init(rempty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 128); }
    mbar_init(bres_bar, 1);
    for (int s = 0; s < MAX_Z; ++s) { mbar_init(zfull_bar(s), 1); mbar_init(zempty_bar(s), 128); }
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
  auto decode = [&](int tile, int& nt, int& x0, int& y0, int& b0) {
    int mt = tile / p.tiles_n;
    nt = tile - mt * p.tiles_n;
    int tb = mt / tiles_xy, rr = mt - tb * tiles_xy;
    int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
    x0 = tx * HTW; y0 = ty * HTH; b0 = tb;
  };
  const int my_tiles = (p.total_tiles > (int)blockIdx.x) ? (p.total_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == 0) {
    // ================= TMA producer =================
    if (BRES && my_tiles > 0) {
      if (lane == 0) {
        mbar_expect_tx(bres_bar, 9 * p.nslab * B_TILE_BYTES);
        for (int c = 0; c < p.nslab; ++c)
          for (int t = 0; t < 9; ++t)
            tma_load_2d(b_base + (c * 9 + t) * B_TILE_BYTES, &tmB, bres_bar, t * p.Cin + c * KSLAB, 0);
      }
      __syncwarp();
    }
    const int items = my_tiles * p.nslab;
    int stage = 0; uint32_t phase = 0;
    int ri = 0, rs = 0; uint32_t rphase = 0;       // next raw item to issue
    int zs = 0; uint32_t zphase = 0;
    // Raw tiles run ahead of the B loads.  With streamed B the look-ahead is n_raw - 2 items, so the slot
    // a raw load waits for was released two items ago and the wait never delays this item's B loads;
    // with resident B the producer has nothing else to do and simply keeps the ring full.
    const int look = BRES ? items : (p.n_raw >= 2 ? p.n_raw - 2 : 0);
    for (int item = 0; item < items; ++item) {
      while (ri < items && ri <= item + look) {
        mbar_wait(rempty_bar(rs), rphase ^ 1);
        if (lane == 0) {
          int k = ri / p.nslab, c = ri - k * p.nslab;
          int nt, x0, y0, b0;
          decode((int)blockIdx.x + k * (int)gridDim.x, nt, x0, y0, b0);
          mbar_expect_tx(rfull_bar(rs), p.raw_tx);
          tma_load_4d(raw_base + rs * p.raw_bytes, &tmRaw, rfull_bar(rs), c * KSLAB, x0 - 1, y0 - 1, b0);
        }
        __syncwarp();
        if (p.n_z > 0 && ri % p.nslab == 0) {
          // the tile's addend: Z[b, y0/2 .. +8, x0/2 .. +4, nt*BN .. +BN), long before the epilogue needs it
          mbar_wait(zempty_bar(zs), zphase ^ 1);
          if (lane == 0) {
            int k = ri / p.nslab;
            int nt, x0, y0, b0;
            decode((int)blockIdx.x + k * (int)gridDim.x, nt, x0, y0, b0);
            mbar_expect_tx(zfull_bar(zs), Z_BYTES);
            for (int cc = 0; cc < Z_BYTES / 4096; ++cc)
              tma_load_4d(z_base + zs * Z_BYTES + cc * 4096, &tmZ, zfull_bar(zs), nt * BN + cc * 64, x0 >> 1, y0 >> 1, b0);
          }
          __syncwarp();
          if (++zs == p.n_z) { zs = 0; zphase ^= 1; }
        }
        ++ri;
        if (++rs == p.n_raw) { rs = 0; rphase ^= 1; }
      }
      if (!BRES) {
        int k = item / p.nslab, c = item - k * p.nslab;
        int nt, x0, y0, b0;
        decode((int)blockIdx.x + k * (int)gridDim.x, nt, x0, y0, b0);
        for (int t = 0; t < 9; ++t) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          if (lane == 0) {
            mbar_expect_tx(full_bar(stage), B_TILE_BYTES);
            tma_load_2d(b_base + stage * B_TILE_BYTES, &tmB, full_bar(stage), t * p.Cin + c * KSLAB, nt * BN);
          }
          __syncwarp();
          if (++stage == p.n_stage) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (warp-convergent loop, one elected lane issues; taps fully unrolled) =====
    if (my_tiles > 0) {
      constexpr uint32_t idesc = umma_idesc<BN>();
      // descriptor bits that do not change: SBO = PITCH rows of 128 bytes (A), 1024 bytes (B)
      const uint64_t a_hi = umma_desc_sw128_ex(0, PITCH * 128u, 0);
      const uint64_t b_hi = umma_desc_sw128(0);
      int stage = 0; uint32_t phase = 0;
      int rs = 0; uint32_t rphase = 0;
      if (BRES) mbar_wait(bres_bar, 0);
      for (int it = 0; it < my_tiles; ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int c = 0; c < p.nslab; ++c) {
          mbar_wait(rfull_bar(rs), rphase);
          tc_fence_after();
          const uint64_t a0 = a_hi | (uint64_t)((raw_base + rs * p.raw_bytes) >> 4);
          const uint64_t bres0 = b_hi | (uint64_t)((b_base + c * 9 * B_TILE_BYTES) >> 4);
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            uint64_t bdesc;
            if (BRES) bdesc = bres0 + (uint64_t)((t * B_TILE_BYTES) >> 4);
            else {
              mbar_wait(full_bar(stage), phase);
              tc_fence_after();
              bdesc = b_hi | (uint64_t)((b_base + stage * B_TILE_BYTES) >> 4);
            }
            // tap (ky, kx) = (t / 3, t % 3): the A tile starts (ky * PITCH + kx) 128-byte rows further on
            const uint64_t adesc = a0 + (uint64_t)((((t / 3) * PITCH + (t % 3)) * 128) >> 4);
            umma_k64_elect(tmem_d, adesc, bdesc, idesc, (t == 0) ? (uint32_t)(c != 0) : 1u);
            if (!BRES) {
              umma_commit_elect(empty_bar(stage));
              if (++stage == p.n_stage) { stage = 0; phase ^= 1; }
            }
          }
          umma_commit_elect(rempty_bar(rs));
          if (c == p.nslab - 1) umma_commit_elect(tfull_bar(as));
          if (++rs == p.n_raw) { rs = 0; rphase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5 and 6..9: one accumulator buffer, staging pair, shift area, named
    // barrier and store leader per warpgroup; tile it goes to warpgroup it & 1) =================
    const int ewg = (warp - 2) >> 2;
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int tx_in = row & (HTW - 1), ty_in = row >> 3;
    const bool leader = (threadIdx.x == 64 + 128 * ewg);
    const uint32_t stg_row = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const EpiParams& e = p.e;
    const uint32_t stg_wg = stg_base + (uint32_t)(ewg * N_STG * STG_BYTES);
    int sbuf = 0;
    int zs = 0; uint32_t zphase = 0;
    for (int it = 0; it < my_tiles; ++it) {
      if ((it & 1) != ewg) {          // the other warpgroup's tile (and addend slot)
        if (p.n_z > 0 && ++zs == p.n_z) { zs = 0; zphase ^= 1; }
        continue;
      }
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      int nt, x0, y0, b0;
      decode((int)blockIdx.x + it * (int)gridDim.x, nt, x0, y0, b0);
      const int x = x0 + tx_in, y = y0 + ty_in, b = b0;
      const bool valid = (x < e.W) && (y < e.H);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
      uint32_t z_row = 0, zsw = 0;
      if (p.n_z > 0) {
        const uint32_t pz = (uint32_t)((ty_in >> 1) * (HTW / 2) + (tx_in >> 1));
        mbar_wait(zfull_bar(zs), zphase);
        z_row = z_base + zs * Z_BYTES + pz * 128u;
        zsw = pz & 7u;
      }
      tc_epilogue_tile<BN>(e, om, taddr, 0, nt, x0, y0, b0, x, y, b, valid, stg_wg, stg_row, sw, leader, sbuf,
                           tempty_bar(as), tfull_bar(as), aphase, bar_base + 512u + 512u * (uint32_t)ewg, z_row, zsw,
                           1u + (uint32_t)ewg);
      if (p.n_z > 0) {
        mbar_arrive(zempty_bar(zs));
        if (++zs == p.n_z) { zs = 0; zphase ^= 1; }
      }
    }
    if (leader) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

struct HaloPlan { int BN, n_raw, n_stage, raw_bytes, raw_tx, b_resident, n_z, smem; };

int g_halo_mode = -1;
int halo_mode() {   // 0 off, 1 on (env TMAT_HALO)
  if (g_halo_mode < 0) {
    const char* e = getenv("TMAT_HALO");
    g_halo_mode = e ? atoi(e) : 1;
  }
  return g_halo_mode;
}

bool plan_halo(const ConvOp& op, int mode, HaloPlan* pl) {
  if (mode != 1) return false;
  if (op.ngroups != 1 || op.ntaps != 9 || op.a1 || op.a0_stride != 1 || op.a0_C % KSLAB != 0) return false;
  for (int t = 0; t < 9; ++t)
    if (op.dy[0][t] != t / 3 - 1 || op.dx[0][t] != t % 3 - 1) return false;
  if (op.epi == EPI_HEAD) { if (op.Npad != 16) return false; pl->BN = 16; }
  else if (op.epi == EPI_STORE) {
    if (op.Cout % 64 != 0 || op.Npad != op.Cout || op.Npad > 128) return false;
    pl->BN = op.Npad % 128 == 0 ? 128 : 64;
  } else return false;
  // tile efficiency: the fixed 8 x 16 tile must not waste much of the image
  double eff = (double)op.W * op.H / ((double)cdiv(op.W, HTW) * HTW * cdiv(op.H, HTH) * HTH);
  if (eff < 0.8) return false;
  pl->raw_tx = PITCH * (HTH + 2) * 128;
  pl->raw_bytes = (pl->raw_tx + 1023) / 1024 * 1024;
  const int nslab = op.a0_C / KSLAB;
  const int btile = pl->BN * KSLAB * 2;
  // the half-resolution addend of the up path comes through a producer-fed ring instead of a per-thread gather
  // (only next to resident weights: with streamed weights the ring would throttle the weight loads)
  const bool z_ok = (op.epi == EPI_STORE && op.add_src && op.add_shift == 1 && op.W % 2 == 0 && op.H % 2 == 0);
  const int tiles_n = op.Npad / pl->BN;
  for (int with_z = z_ok ? 1 : 0; with_z >= 0; --with_z) {
    pl->n_z = with_z ? MAX_Z : 0;
    const int z_bytes = pl->n_z * (pl->BN >= 64 ? pl->BN / 64 : 1) * 4096;
    const int budget = 227 * 1024 - 3072 - HALO_EPI_WGS * N_STG * STG_BYTES - z_bytes;
    for (int n_raw = 3; n_raw >= 2; --n_raw) {
      int left = budget - n_raw * pl->raw_bytes;
      if (left < 2 * btile) continue;
      pl->n_raw = n_raw;
      if (tiles_n == 1 && 9 * nslab * btile <= left) { pl->b_resident = 1; pl->n_stage = 1; }
      else { pl->b_resident = 0; pl->n_stage = std::min(MAX_STAGES, left / btile); }
      if (with_z && !pl->b_resident) continue;
      int nb = pl->b_resident ? 9 * nslab : pl->n_stage;
      pl->smem = nb * btile + n_raw * pl->raw_bytes + HALO_EPI_WGS * N_STG * STG_BYTES + z_bytes + 3072;
      return true;
    }
  }
  return false;
}

template <int BN, bool BRES>
int launch_halo(tmat_ctx* ctx, const CUtensorMap& raw, const CUtensorMap& b, const CUtensorMap& z, const TcOutMaps& om, const HaloParams& p,
                int smem, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    CUDA_TRY(cudaFuncSetAttribute(conv_halo_kernel<BN, BRES>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  int grid = std::min(p.total_tiles, ctx->sm_count);
  tmat_ctx::ProfRec rec{nullptr, nullptr, 0.0};
  if (ctx->profiling) {
    for (cudaEvent_t* e : {&rec.a, &rec.b}) {
      if (!ctx->event_pool.empty()) { *e = ctx->event_pool.back(); ctx->event_pool.pop_back(); }
      else CUDA_TRY(cudaEventCreate(e));
    }
    rec.flops = 2.0 * p.e.B * p.e.H * p.e.W * (double)(p.tiles_n * BN) * (double)(9 * p.Cin);
    CUDA_TRY(cudaEventRecord(rec.a, st));
  }
  conv_halo_kernel<BN, BRES><<<grid, HALO_THREADS, smem, st>>>(raw, b, z, om, p);
  KERNEL_CHECK(ctx);
  if (ctx->profiling) {
    CUDA_TRY(cudaEventRecord(rec.b, st));
    ctx->prof.push_back(rec);
  }
  return TMAT_OK;
}

}  // namespace

int conv_halo_mode() { return halo_mode(); }

extern "C" int tmat_debug_set_halo_mode(int mode) {
  g_halo_mode = mode;
  return TMAT_OK;
}

bool conv_halo_supported(const ConvOp& op) {
  HaloPlan pl;
  return plan_halo(op, halo_mode(), &pl);
}

int conv_run_halo(tmat_ctx* ctx, const ConvOp& op, int mode, cudaStream_t st) {
  HaloPlan pl;
  memset(&pl, 0, sizeof(pl));
  TMAT_REQUIRE(plan_halo(op, mode, &pl), "conv_halo(%s): shape not supported", op.name);
  HaloParams p;
  memset(&p, 0, sizeof(p));
  p.tiles_x = cdiv(op.W, HTW); p.tiles_y = cdiv(op.H, HTH); p.tiles_b = op.B; p.tiles_n = op.Npad / pl.BN;
  int64_t total = (int64_t)p.tiles_x * p.tiles_y * p.tiles_b * p.tiles_n;
  TMAT_REQUIRE(total < (1ll << 31), "conv_halo(%s): too many tiles", op.name);
  p.total_tiles = (int)total;
  p.nslab = op.a0_C / KSLAB; p.Cin = op.a0_C;
  p.raw_bytes = pl.raw_bytes; p.raw_tx = pl.raw_tx; p.n_raw = pl.n_raw; p.n_stage = pl.n_stage;
  p.n_z = pl.n_z;
  p.e = epi_of(op);
  CUtensorMap mRaw, mB, mZ;
  TcOutMaps om;
  TMAT_TRY(tc_make_act_map(&mRaw, op.a0, op.B, op.H, op.W, op.a0_C, PITCH, HTH + 2, 1));
  TMAT_TRY(tc_make_weight_map(&mB, op.wmat, op.Npad, op.K, pl.BN));
  if (op.epi == EPI_STORE) {
    TMAT_TRY(tc_make_act_map(&om.m[0], op.out, op.B, op.H, op.W, op.Cout, HTW, HTH, 1));
    om.m[1] = om.m[2] = om.m[3] = om.m[0];
    if (op.out2) TMAT_TRY(tc_make_act_map(&om.m2, op.out2, op.B, op.H, op.W, op.Cout, HTW, HTH, 1));
    else om.m2 = om.m[0];
  } else {
    om.m[0] = om.m[1] = om.m[2] = om.m[3] = om.m2 = mRaw;
  }
  if (pl.n_z > 0) TMAT_TRY(tc_make_act_map(&mZ, op.add_src, op.B, op.H / 2, op.W / 2, op.Cout, HTW / 2, HTH / 2, 1));
  else mZ = mRaw;
  const int key = pl.BN * 2 + (pl.b_resident ? 1 : 0);
  switch (key) {
    case 128 * 2 + 0: return launch_halo<128, false>(ctx, mRaw, mB, mZ, om, p, pl.smem, st);
    case 128 * 2 + 1: return launch_halo<128, true>(ctx, mRaw, mB, mZ, om, p, pl.smem, st);
    case 64 * 2 + 0: return launch_halo<64, false>(ctx, mRaw, mB, mZ, om, p, pl.smem, st);
    case 64 * 2 + 1: return launch_halo<64, true>(ctx, mRaw, mB, mZ, om, p, pl.smem, st);
    case 16 * 2 + 0: return launch_halo<16, false>(ctx, mRaw, mB, mZ, om, p, pl.smem, st);
    case 16 * 2 + 1: return launch_halo<16, true>(ctx, mRaw, mB, mZ, om, p, pl.smem, st);
  }
  return TMAT_ERR_INVALID;
}
