// PTX wrappers shared by the tcgen05 kernels (conv_tc.cu, conv_dwpw.cu): mbarrier, TMA load / store,
// TMEM allocation, UMMA descriptors and issue, tcgen05.ld.
#pragma once
#include <type_traits>
#include "unet.cuh"
#include <cuda.h>

namespace tcptx {

constexpr int TILE_M = 128;
constexpr int KSLAB = 64;                          // bf16 elements per k-step = 128 bytes = swizzle span
constexpr int A_STAGE_BYTES = TILE_M * KSLAB * 2;  // 16 KiB
constexpr int STG_BYTES = TILE_M * 128;            // one staging tile: 128 pixels x 64 bf16 channels

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
// smem tile -> global tensor (bulk async group of the issuing thread)
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
      ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
// element-wise max of the shared-memory box into the tensor (performed at L2; out-of-range parts of the box are
// clipped like a store's)
__device__ __forceinline__ void tma_reduce_max_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.4d.global.shared::cta.max.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
      ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// named barrier among the 128 epilogue threads (barrier 0 is __syncthreads)
__device__ __forceinline__ void epi_bar_sync(uint32_t id = 1) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
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
// Warp-convergent issue (the whole warp calls, one elected lane issues): the four K = 16 MMAs of one 64-wide
// k-step.  Keeping the loop around it warp-uniform lets the compiler hold descriptors, barrier addresses and
// ring state in uniform registers, which is what keeps the single issuing thread ahead of the tensor pipe.
__device__ __forceinline__ void umma_k64_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate_first) {
  asm volatile(
      "{\n\t.reg .pred pe, pa, pt;\n\t.reg .b64 a1, b1;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "setp.eq.b32 pt, %3, %3;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, pa;\n\t"
      "add.s64 a1, %1, 2;\n\tadd.s64 b1, %2, 2;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, pt;\n\t"
      "add.s64 a1, %1, 4;\n\tadd.s64 b1, %2, 4;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, pt;\n\t"
      "add.s64 a1, %1, 6;\n\tadd.s64 b1, %2, 6;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, pt;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate_first) : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      ::"r"(bar) : "memory");
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
__device__ __forceinline__ uint32_t umma_idesc_rt(int bn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
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
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void ld_shared_2xf32x2(uint32_t addr, f32x2& a, f32x2& b) {
  asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}
__device__ __forceinline__ f32x2 fadd2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// fp32 pair -> bf16x2 (lo in the low half), round to nearest even; RELU clamps negatives to +0 inside the convert
template <bool RELU>
__device__ __forceinline__ uint32_t cvt_bf16x2(f32x2 v) {
  float lo, hi;
  unpack2(v, lo, hi);
  uint32_t d;
  if (RELU) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ uint32_t cvt_bf16x2(f32x2 v, bool relu) {
  return relu ? cvt_bf16x2<true>(v) : cvt_bf16x2<false>(v);
}
// bf16x2 -> fp32 pair (exact)
__device__ __forceinline__ f32x2 bf16x2_to_f32x2(uint32_t x) {
  return pack2(__uint_as_float(x << 16), __uint_as_float(x & 0xffff0000u));
}
__device__ __forceinline__ float sigmoid_f(float v) { return 1.f / (1.f + __expf(-v)); }


}  // namespace tcptx

// tensor maps of the output: m[q] = view of parity q (EPI_STORE: m[0] = the tensor), m2 = ReLU'd copy
struct TcOutMaps {
  CUtensorMap m[4];
  CUtensorMap m2;
};

namespace tcptx {

// L2 prefetch of the addend row a thread will need for a later tile (columns [col0, col0 + ncols) of pixel
// (b, y, x)): the gather otherwise waits out a full DRAM latency per tile with nothing else in flight.
__device__ __forceinline__ void tc_prefetch_addend(const EpiParams& e, int col0, int ncols, int x, int y, int b) {
  if (e.add_src == nullptr || x >= e.W || y >= e.H || b >= e.B) return;
  const int s = e.add_shift;
  const int Ha = (e.H + (1 << s) - 1) >> s, Wa = (e.W + (1 << s) - 1) >> s;
  const bf16* a = e.add_src + (((int64_t)b * Ha + (y >> s)) * Wa + (x >> s)) * e.Cout + col0;
  for (int c = 0; c < ncols; c += 64) asm volatile("prefetch.global.L2 [%0];" ::"l"(a + c));
}

constexpr int N_STG = 2;   // staging tiles of STG_BYTES each at stg_base

// Epilogue of one output tile, executed by the 128 epilogue threads (thread = GEMM row = TMEM lane):
// accumulator columns [0, BN) at taddr -> +shift (+addend) (ReLU) -> bf16 -> swizzled staging tile -> one
// TMA tensor store per 64 channels (the head, BN = 16, stores its 4 floats per pixel directly).  `tempty`
// is arrived on as soon as the accumulator has been read; the wait for the accumulator (`tfull`, `tphase`)
// happens inside, after the first addend loads have been issued.
template <int BN>
__device__ __forceinline__ void tc_epilogue_tile(const EpiParams& e, const TcOutMaps& om, uint32_t taddr, int g, int nt,
                                                 int x0, int y0, int b0, int x, int y, int b, bool valid,
                                                 uint32_t stg_base, uint32_t stg_row, uint32_t sw, bool leader,
                                                 int& sbuf, uint32_t tempty, uint32_t tfull, uint32_t tphase,
                                                 uint32_t shift_smem, uint32_t z_row = 0, uint32_t zsw = 0,
                                                 uint32_t bar_id = 1, bool single = false) {
      // single: this epilogue warpgroup owns ONE staging tile (two warpgroups share the staging pair); not with dual
      // z_row != 0: the addend of this thread's pixel was staged in shared memory by the producer (one
      // 128-byte-swizzled row of 64 channels per 4096-byte block, block c64 / 64); else it is gathered from e.add_src
      if constexpr (BN >= 64) {
        // this tile's BN shift values -> shared memory (read back as broadcast 16-byte loads below); the
        // hand-over barrier of the first chunk orders the writes, the last barrier of the previous tile the reuse
        for (int c = (int)(stg_row >> 7); c < BN; c += 128) {
          const int col = nt * BN + c;
          st_shared_f32(shift_smem + 4u * (uint32_t)c, __ldg(e.shift + (e.mode == EPI_PAR ? col % e.Cout : col)));
        }
        const bool dual = (e.out2 != nullptr);
        const bf16* addp = nullptr;
        if (e.add_src != nullptr && valid && z_row == 0) {
          const int s = e.add_shift;
          const int Ha = (e.H + (1 << s) - 1) >> s, Wa = (e.W + (1 << s) - 1) >> s;
          addp = e.add_src + (((int64_t)b * Ha + (y >> s)) * Wa + (x >> s)) * e.Cout;
        }
#pragma unroll 1
        for (int c64 = 0; c64 < BN; c64 += 64) {
          const int col = nt * BN + c64;                        // column within the group
          int par = 0, ch0 = col;
          if (e.mode == EPI_PAR) { par = g * e.ppg + col / e.Cout; ch0 = col % e.Cout; }
          // the addend (8 x 16 bytes per thread) is fetched before anything that waits: its L2 latency hides
          // behind the accumulator wait / the staging hand-over
          uint4 av[8];
          if (z_row != 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) av[j] = ld_shared_v4(z_row + (uint32_t)(c64 >> 6) * 4096u + ((((uint32_t)j) ^ zsw) << 4));
          } else if (addp != nullptr) {
            const uint4* ap = reinterpret_cast<const uint4*>(addp + ch0);
#pragma unroll
            for (int j = 0; j < 8; ++j) av[j] = __ldg(ap + j);
          }
          const bool has_add = (addp != nullptr) || (z_row != 0);
          if (c64 == 0) {
            mbar_wait(tfull, tphase);
            tc_fence_after();
          }
          // the staging buffer(s) about to be written must have been read out by their last store
          if (leader) { if (dual || single) tma_store_wait_read<0>(); else tma_store_wait_read<1>(); }
          epi_bar_sync(bar_id);
          const uint32_t buf0 = stg_base + ((dual || single) ? 0 : sbuf * STG_BYTES) + stg_row;
          const uint32_t buf1 = stg_base + STG_BYTES + stg_row;
          auto chunk = [&](auto relu_c) {
            constexpr bool RELU = decltype(relu_c)::value;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              uint32_t rr[32];
              tmem_ld32(taddr + c64 + half * 32, rr);
              // packed fp32 pairs: one FADD2 per two columns, ReLU folded into the bf16 convert
              f32x2 v[16];
#pragma unroll
              for (int i = 0; i < 8; ++i)
                ld_shared_2xf32x2(shift_smem + 4u * (uint32_t)(c64 + half * 32 + 4 * i), v[2 * i], v[2 * i + 1]);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i)
                v[i] = fadd2(pack2(__uint_as_float(rr[2 * i]), __uint_as_float(rr[2 * i + 1])), v[i]);
              if (has_add) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint4 u = av[half * 4 + j];
                  v[4 * j] = fadd2(v[4 * j], bf16x2_to_f32x2(u.x));
                  v[4 * j + 1] = fadd2(v[4 * j + 1], bf16x2_to_f32x2(u.y));
                  v[4 * j + 2] = fadd2(v[4 * j + 2], bf16x2_to_f32x2(u.z));
                  v[4 * j + 3] = fadd2(v[4 * j + 3], bf16x2_to_f32x2(u.w));
                }
              }
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint32_t off = (((uint32_t)(half * 4 + j)) ^ sw) << 4;   // 128-byte swizzle: chunk ^= row % 8
                uint4 o;
                o.x = cvt_bf16x2<RELU>(v[4 * j]); o.y = cvt_bf16x2<RELU>(v[4 * j + 1]);
                o.z = cvt_bf16x2<RELU>(v[4 * j + 2]); o.w = cvt_bf16x2<RELU>(v[4 * j + 3]);
                st_shared_v4(buf0 + off, o);
                if (dual) {
                  o.x = cvt_bf16x2<true>(v[4 * j]); o.y = cvt_bf16x2<true>(v[4 * j + 1]);
                  o.z = cvt_bf16x2<true>(v[4 * j + 2]); o.w = cvt_bf16x2<true>(v[4 * j + 3]);
                  st_shared_v4(buf1 + off, o);
                }
              }
            }
          };
          if (e.relu) chunk(std::true_type{}); else chunk(std::false_type{});
          if (c64 + 64 >= BN) {                  // accumulator fully read: hand the TMEM buffer back
            tc_fence_before();
            mbar_arrive(tempty);
          }
          fence_proxy_async();                   // generic-proxy smem writes -> visible to the TMA unit
          epi_bar_sync(bar_id);
          if (leader) {
            tma_store_4d(&om.m[par], buf0 - stg_row, ch0, x0, y0, b0);
            if (dual) tma_store_4d(&om.m2, buf1 - stg_row, ch0, x0, y0, b0);
            tma_store_commit();
          }
          if (!dual) sbuf ^= 1;
        }
      } else {
        // head: 4 output-pixel parities in columns 0..3
        mbar_wait(tfull, tphase);
        tc_fence_after();
        uint32_t rr[16];
        tmem_ld16(taddr, rr);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(tempty);
        if (valid) {
          float o[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float v = __uint_as_float(rr[i]) + __ldg(e.shift + i);
            o[i] = e.logits ? v : sigmoid_f(v);
          }
          float* dst = e.out_f32 + ((int64_t)b * 2 * e.H + 2 * y) * (2 * e.W) + 2 * x;
          *reinterpret_cast<float2*>(dst) = make_float2(o[0], o[1]);
          *reinterpret_cast<float2*>(dst + 2 * e.W) = make_float2(o[2], o[3]);
        }
      }
}

}  // namespace tcptx

// host helpers (conv_tc.cu)
// NHWC bf16 tensor (B, H*sy, W*sx, C) viewed at pixel strides (sx, sy) as (B, H, W, C) -> 4-D map with
// box (64 channels, BW, BH, BI), 128-byte swizzle, zero fill outside the view
int tc_make_act_map(CUtensorMap* m, const bf16* base, int B, int H, int W, int C, int BW, int BH, int BI,
                    int sx = 1, int sy = 1);
// packed weights [Npad][K] bf16 -> 2-D map with box (64, BN)
int tc_make_weight_map(CUtensorMap* m, const bf16* base, int Npad, int K, int BN);
// 128-pixel tile (2^a x 2^b pixels x 2^c images) covering (B, H, W) with the fewest tiles
void tc_choose_tile(int B, int H, int W, int* tw_log, int* th_log, int* tn_log);
