// Shared host-side plumbing for libtmat_b200: context, scratch arena, error reporting.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <vector>
#include <string>
#include <map>
#include "../../include/tmat_b200.h"
#include "../../include/tmat_b200_debug.h"

#define TMAT_SM_TARGET 100

void tmat_set_error(const char* fmt, ...);

#define CUDA_TRY(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      tmat_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return TMAT_ERR_CUDA;                                                              \
    }                                                                                    \
  } while (0)

#define TMAT_TRY(expr)            \
  do {                            \
    int _r = (expr);              \
    if (_r != TMAT_OK) return _r; \
  } while (0)

#define TMAT_REQUIRE(cond, ...)      \
  do {                               \
    if (!(cond)) {                   \
      tmat_set_error(__VA_ARGS__);   \
      return TMAT_ERR_INVALID;       \
    }                                \
  } while (0)

// Check the launch that was just issued and count it.
#define KERNEL_CHECK(ctx)                 \
  do {                                    \
    CUDA_TRY(cudaGetLastError());         \
    (ctx)->launches++;                    \
  } while (0)

struct tmat_ctx {
  int device = 0;
  int sm_count = 0;
  int cc_major = 0, cc_minor = 0;
  int64_t launches = 0;
  // bump arena for scratch; reset at the start of every public entry point that uses it
  char* arena = nullptr;
  size_t arena_cap = 0;
  size_t arena_off = 0;
  std::vector<void*> overflow;   // blocks allocated when the arena was too small
  size_t overflow_bytes = 0;
  cudaStream_t own_stream = nullptr;
  // small pinned staging area for scalars coming back from the device
  void* pinned = nullptr;
  size_t pinned_cap = 0;
  unsigned int lut_slot = 0;     // rotating 256-byte slots (pinned + 8192 ...) for small stream-ordered uploads
  // optional per-launch timing of the tensor-core convolution kernel (bench.py's roofline leg)
  bool profiling = false;
  struct ProfRec { cudaEvent_t a, b; double flops; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> event_pool;
};

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Reset the arena.  If the previous call overflowed, grow the arena so the next one fits.
int arena_reset(tmat_ctx* ctx);
// Allocate scratch device memory valid until the next arena_reset.
int arena_alloc(tmat_ctx* ctx, size_t bytes, void** out);
int arena_reserve(tmat_ctx* ctx, size_t bytes);
static inline size_t arena_mark(tmat_ctx* ctx) { return ctx->arena_off; }
static inline void arena_release(tmat_ctx* ctx, size_t mark) { if (mark <= ctx->arena_off) ctx->arena_off = mark; }
template <typename T>
static inline int arena_get(tmat_ctx* ctx, size_t count, T** out) {
  void* p = nullptr;
  int r = arena_alloc(ctx, count * sizeof(T), &p);
  *out = (T*)p;
  return r;
}

// numpy's pairwise summation of a contiguous float64 vector (np.add.reduce), defined in stp.cu
double np_pairwise_sum(const double* a, int64_t n);

static inline cudaStream_t as_stream(void* s) { return (cudaStream_t)s; }

static inline size_t dtype_size(int dtype) {
  switch (dtype) {
    case TMAT_U8: return 1;
    case TMAT_U16: return 2;
    case TMAT_F32: return 4;
    case TMAT_F64: return 8;
  }
  return 0;
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline int64_t cdiv64(int64_t a, int64_t b) { return (a + b - 1) / b; }
