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
// Warp roles (192 threads, persistent over output tiles):
//   warp 0      TMA producer       (ring of STAGES stages, full/empty mbarriers)
//   warp 1      MMA issuer         (one lane issues tcgen05.mma, tcgen05.commit frees stages);
//               also owns the TMEM allocation (2 accumulator buffers of BN columns)
//   warps 2..5  epilogue           (tcgen05.ld -> +shift / ReLU / pool-add / upsample / sigmoid ->
//                                   global), overlapped with the next tile's MMAs
#include "unet.cuh"
#include <cuda.h>
#include <algorithm>

namespace {

constexpr int TILE_M = 128;
constexpr int KSLAB = 64;                          // bf16 elements per k-step = 128 bytes = swizzle span
constexpr int A_STAGE_BYTES = TILE_M * KSLAB * 2;  // 16 KiB
constexpr int MAX_KSTEPS = 96;
constexpr int TC_THREADS = 192;

struct TcParams {
  int tw_log, th_log, tn_log;   // tile = 2^tw x 2^th pixels x 2^tn images (product 128)
  int tiles_x, tiles_y, tiles_b, tiles_n, total_tiles;
  int nk;
  uint16_t kstep[MAX_KSTEPS];   // bit0 src | bits1-2 dy+1 | bits3-4 dx+1 | bits5.. c0/64
  EpiParams e;
};

// ---- PTX wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (reported as a launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "n"(COLS) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (bf16 inputs, fp32 accumulate)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// shared-memory matrix descriptor: K-major, 128-byte swizzle, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);   // start address  [0,14)
  d |= (uint64_t)1 << 16;                    // leading byte offset (ignored for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;          // stride byte offset [32,46)
  d |= (uint64_t)1 << 46;                    // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
  return d;
}

// instruction descriptor: c=f32, a=b=bf16, both K-major, M=128, N=BN
template <int BN>
__device__ __forceinline__ constexpr uint32_t umma_idesc() {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float sigmoid_f(float v) { return 1.f / (1.f + __expf(-v)); }

// ---- epilogue: 32 consecutive channels of one pixel -------------------------------------------
__device__ __forceinline__ void store32(bf16* dst, const float (&v)[32], bool relu) {
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = relu ? fmaxf(v[8 * q + i], 0.f) : v[8 * q + i];
    uint4 o;
    o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
    o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
    d[q] = o;
  }
}

__device__ __forceinline__ void epilogue32(const EpiParams& e, int b, int y, int x, int col, float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    float4 s = __ldg(reinterpret_cast<const float4*>(e.shift + col + i));
    v[i] += s.x; v[i + 1] += s.y; v[i + 2] += s.z; v[i + 3] += s.w;
  }
  if (e.mode == EPI_POOLADD) {
    const int H2 = 2 * e.H, W2 = 2 * e.W;
    float m[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) m[i] = -INFINITY;
    for (int ky = 0; ky < 3; ++ky) {
      int yy = 2 * y + ky;
      if (yy >= H2) continue;
      for (int kx = 0; kx < 3; ++kx) {
        int xx = 2 * x + kx;
        if (xx >= W2) continue;
        const uint4* src = reinterpret_cast<const uint4*>(e.pool_src + (((int64_t)b * H2 + yy) * W2 + xx) * e.Cout + col);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 u = __ldg(src + q);
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float2 t = __bfloat1622float2(h[i]);
            m[8 * q + 2 * i] = fmaxf(m[8 * q + 2 * i], t.x);
            m[8 * q + 2 * i + 1] = fmaxf(m[8 * q + 2 * i + 1], t.y);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += m[i];
  }
  if (e.mode == EPI_UP2) {
    const int H2 = 2 * e.H, W2 = 2 * e.W;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      int64_t o = (((int64_t)b * H2 + 2 * y + (d >> 1)) * W2 + 2 * x + (d & 1)) * e.Cout + col;
      store32(e.out + o, v, e.relu != 0);
      if (e.out2) store32(e.out2 + o, v, true);
    }
  } else {
    int64_t o = (((int64_t)b * e.H + y) * e.W + x) * e.Cout + col;
    store32(e.out + o, v, e.relu != 0);
    if (e.out2) store32(e.out2 + o, v, true);
  }
}

// ---- the kernel -------------------------------------------------------------------------------
template <int BN, int STAGES>
struct TcSmem {
  static constexpr int B_STAGE_BYTES = BN * KSLAB * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int BYTES = BAR_OFFSET + 256 + 1024;   // barriers + alignment slack
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ TcParams p) {
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

  if (warp == 0) {
    // ================= TMA producer =================
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      int nt = tile % p.tiles_n, mt = tile / p.tiles_n;
      int tb = mt / tiles_xy, r = mt - tb * tiles_xy;
      int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
      int x0 = tx << p.tw_log, y0 = ty << p.th_log, b0 = tb << p.tn_log;
      for (int ks = 0; ks < p.nk; ++ks) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        if (lane == 0) {
          uint32_t kd = p.kstep[ks];
          int src = kd & 1, dy = (int)((kd >> 1) & 3) - 1, dx = (int)((kd >> 3) & 3) - 1, c0 = (int)(kd >> 5) * KSLAB;
          uint32_t sa = smem_base + stage * L::STAGE_BYTES;
          uint32_t sb = sa + A_STAGE_BYTES;
          mbar_expect_tx(full_bar(stage), L::STAGE_BYTES);
          tma_load_4d(sa, src ? &tmA1 : &tmA0, full_bar(stage), c0, x0 + dx, y0 + dy, b0);
          tma_load_2d(sb, &tmB, full_bar(stage), ks * KSLAB, nt * BN);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
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
        if (lane == 0) {
          uint32_t sa = smem_base + stage * L::STAGE_BYTES;
          uint64_t adesc = umma_desc_sw128(sa);
          uint64_t bdesc = umma_desc_sw128(sa + A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < KSLAB / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in the (>>4) address field
            umma_bf16(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (ks | k) != 0);
          }
          umma_commit(empty_bar(stage));                     // stage reusable once these MMAs retire
          if (ks == p.nk - 1) umma_commit(tfull_bar(as));    // accumulator complete
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    const int q = warp & 3;                   // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;            // GEMM row = pixel within the tile
    const int tx_in = row & (TW - 1);
    const int ty_in = (row >> p.tw_log) & (TH - 1);
    const int tb_in = row >> (p.tw_log + p.th_log);
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      int nt = tile % p.tiles_n, mt = tile / p.tiles_n;
      int tb = mt / tiles_xy, r = mt - tb * tiles_xy;
      int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
      const int x = (tx << p.tw_log) + tx_in, y = (ty << p.th_log) + ty_in, b = (tb << p.tn_log) + tb_in;
      const bool valid = (x < p.e.W) && (y < p.e.H) && (b < p.e.B);
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
      if constexpr (BN >= 32) {
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t rr[32];
          tmem_ld32(taddr + c, rr);
          tmem_ld_wait();
          int col = nt * BN + c;
          if (valid && col < p.e.Cout) {
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(rr[i]);
            epilogue32(p.e, b, y, x, col, v);
          }
        }
      } else {
        // head: 4 output-pixel parities in columns 0..3
        uint32_t rr[16];
        tmem_ld16(taddr, rr);
        tmem_ld_wait();
        if (valid) {
          float o[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float v = __uint_as_float(rr[i]) + __ldg(p.e.shift + i);
            o[i] = p.e.logits ? v : sigmoid_f(v);
          }
          float* dst = p.e.out_f32 + ((int64_t)b * 2 * p.e.H + 2 * y) * (2 * p.e.W) + 2 * x;
          *reinterpret_cast<float2*>(dst) = make_float2(o[0], o[1]);
          *reinterpret_cast<float2*>(dst + 2 * p.e.W) = make_float2(o[2], o[3]);
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(as));
    }
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
static int make_act_map(CUtensorMap* m, const bf16* base, int B, int H, int W, int C, int TW, int TH, int TN) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { tmat_set_error("conv_tc: cuTensorMapEncodeTiled unavailable"); return TMAT_ERR_CUDA; }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
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

static int make_weight_map(CUtensorMap* m, const bf16* base, int Npad, int K, int BN) {
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
static void choose_tile(int B, int H, int W, int* tw_log, int* th_log, int* tn_log) {
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
  int nk = op.ntaps * (op.a0_C / KSLAB) + (op.a1 ? op.a1_C / KSLAB : 0);
  if (nk > MAX_KSTEPS) return false;
  if (op.epi == EPI_HEAD) return op.Npad == 16;
  return (op.Cout % 64 == 0) && op.Npad == op.Cout;
}

template <int BN, int STAGES>
static int launch_tc(tmat_ctx* ctx, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                     const TcParams& p, cudaStream_t st) {
  using L = TcSmem<BN, STAGES>;
  static bool configured = false;
  if (!configured) {
    CUDA_TRY(cudaFuncSetAttribute(conv_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
    configured = true;
  }
  int grid = std::min(p.total_tiles, ctx->sm_count);
  tmat_ctx::ProfRec rec{nullptr, nullptr, 0.0};
  if (ctx->profiling) {
    for (cudaEvent_t* e : {&rec.a, &rec.b}) {
      if (!ctx->event_pool.empty()) { *e = ctx->event_pool.back(); ctx->event_pool.pop_back(); }
      else CUDA_TRY(cudaEventCreate(e));
    }
    rec.flops = 2.0 * p.e.B * p.e.H * p.e.W * (double)(p.tiles_n * BN) * (double)(p.nk * KSLAB);
    CUDA_TRY(cudaEventRecord(rec.a, st));
  }
  conv_tc_kernel<BN, STAGES><<<grid, TC_THREADS, L::BYTES, st>>>(a0, a1, b, p);
  KERNEL_CHECK(ctx);
  if (ctx->profiling) {
    CUDA_TRY(cudaEventRecord(rec.b, st));
    ctx->prof.push_back(rec);
  }
  return TMAT_OK;
}

int conv_run_tc(tmat_ctx* ctx, const ConvOp& op, cudaStream_t st) {
  TMAT_REQUIRE(conv_tc_supported(op), "conv_tc(%s): shape not supported by the tcgen05 path", op.name);
  TcParams p;
  memset(&p, 0, sizeof(p));
  choose_tile(op.B, op.H, op.W, &p.tw_log, &p.th_log, &p.tn_log);
  const int TW = 1 << p.tw_log, TH = 1 << p.th_log, TN = 1 << p.tn_log;
  int BN = op.epi == EPI_HEAD ? 16 : (op.Cout % 256 == 0 ? 256 : (op.Cout % 128 == 0 ? 128 : 64));
  p.tiles_x = cdiv(op.W, TW);
  p.tiles_y = cdiv(op.H, TH);
  p.tiles_b = cdiv(op.B, TN);
  p.tiles_n = op.Npad / BN;
  int64_t total = (int64_t)p.tiles_x * p.tiles_y * p.tiles_b * p.tiles_n;
  TMAT_REQUIRE(total < (1ll << 31), "conv_tc(%s): too many tiles", op.name);
  p.total_tiles = (int)total;
  int nk = 0;
  for (int t = 0; t < op.ntaps; ++t)
    for (int c = 0; c < op.a0_C / KSLAB; ++c)
      p.kstep[nk++] = (uint16_t)(0 | ((op.dy[t] + 1) << 1) | ((op.dx[t] + 1) << 3) | (c << 5));
  if (op.a1)
    for (int c = 0; c < op.a1_C / KSLAB; ++c) p.kstep[nk++] = (uint16_t)(1 | (1 << 1) | (1 << 3) | (c << 5));
  p.nk = nk;
  TMAT_REQUIRE(nk * KSLAB == op.K, "conv_tc(%s): K mismatch (%d k-steps, K=%d)", op.name, nk, op.K);
  p.e = epi_of(op);

  CUtensorMap mA0, mA1, mB;
  TMAT_TRY(make_act_map(&mA0, op.a0, op.B, op.H, op.W, op.a0_C, TW, TH, TN));
  if (op.a1) TMAT_TRY(make_act_map(&mA1, op.a1, op.B, op.H, op.W, op.a1_C, TW, TH, TN));
  else mA1 = mA0;
  TMAT_TRY(make_weight_map(&mB, op.wmat, op.Npad, op.K, BN));
  switch (BN) {
    case 256: return launch_tc<256, 4>(ctx, mA0, mA1, mB, p, st);
    case 128: return launch_tc<128, 6>(ctx, mA0, mA1, mB, p, st);
    case 64: return launch_tc<64, 8>(ctx, mA0, mA1, mB, p, st);
    case 16: return launch_tc<16, 8>(ctx, mA0, mA1, mB, p, st);
  }
  return TMAT_ERR_INVALID;
}
