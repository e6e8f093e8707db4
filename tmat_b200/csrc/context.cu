// Context, scratch arena and error plumbing of libtmat_b200 (no reference counterpart: the
// reference is a Python library; this is the state a ctypes binding holds per thread).
#include "common.cuh"

static thread_local char g_err[1024] = "";

void tmat_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* tmat_last_error(void) { return g_err; }
extern "C" int tmat_version(void) { return 100; }

extern "C" int tmat_ctx_create(int device, tmat_ctx** out) {
  if (!out) {
    tmat_set_error("tmat_ctx_create: out is NULL");
    return TMAT_ERR_INVALID;
  }
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    tmat_set_error("tmat_ctx_create: no CUDA device (%s); libtmat_b200 has no CPU fallback",
                   e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    cudaGetLastError();
    return TMAT_ERR_NO_DEVICE;
  }
  TMAT_REQUIRE(device >= 0 && device < n, "tmat_ctx_create: device %d out of range [0,%d)", device, n);
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    tmat_set_error("tmat_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only",
                   device, prop.major, prop.minor);
    return TMAT_ERR_NO_DEVICE;
  }
  tmat_ctx* ctx = new tmat_ctx();
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->cc_major = prop.major;
  ctx->cc_minor = prop.minor;
  ctx->pinned_cap = 1 << 16;
  CUDA_TRY(cudaMallocHost(&ctx->pinned, ctx->pinned_cap));
  *out = ctx;
  return TMAT_OK;
}

static void free_overflow(tmat_ctx* ctx) {
  for (void* p : ctx->overflow) cudaFree(p);
  ctx->overflow.clear();
  ctx->overflow_bytes = 0;
}

extern "C" int tmat_ctx_destroy(tmat_ctx* ctx) {
  if (!ctx) return TMAT_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  free_overflow(ctx);
  if (ctx->arena) cudaFree(ctx->arena);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (auto& r : ctx->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (auto e : ctx->event_pool) cudaEventDestroy(e);
  delete ctx;
  return TMAT_OK;
}

extern "C" int tmat_host_alloc(int64_t bytes, void** out) {
  TMAT_REQUIRE(out && bytes > 0, "tmat_host_alloc: bad argument");
  *out = nullptr;
  CUDA_TRY(cudaHostAlloc(out, (size_t)bytes, cudaHostAllocDefault));
  return TMAT_OK;
}

extern "C" int tmat_host_free(void* p) {
  if (p) CUDA_TRY(cudaFreeHost(p));
  return TMAT_OK;
}

extern "C" int tmat_ctx_sync(tmat_ctx* ctx) {
  TMAT_REQUIRE(ctx, "tmat_ctx_sync: ctx is NULL");
  CUDA_TRY(cudaSetDevice(ctx->device));
  CUDA_TRY(cudaDeviceSynchronize());
  return TMAT_OK;
}

extern "C" int tmat_ctx_launch_count(tmat_ctx* ctx, int64_t* out) {
  TMAT_REQUIRE(ctx && out, "tmat_ctx_launch_count: NULL argument");
  *out = ctx->launches;
  return TMAT_OK;
}

int arena_reset(tmat_ctx* ctx) {
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->overflow.empty()) {
    // the previous call did not fit: wait for it, then grow so the steady state is malloc-free
    CUDA_TRY(cudaDeviceSynchronize());
    size_t want = align_up(ctx->arena_off + ctx->overflow_bytes + (ctx->overflow.size() + 1) * 512, 1 << 20);
    want += want / 4;
    free_overflow(ctx);
    if (ctx->arena) CUDA_TRY(cudaFree(ctx->arena));
    ctx->arena = nullptr;
    ctx->arena_cap = 0;
    CUDA_TRY(cudaMalloc((void**)&ctx->arena, want));
    ctx->arena_cap = want;
  }
  ctx->arena_off = 0;
  return TMAT_OK;
}

int arena_alloc(tmat_ctx* ctx, size_t bytes, void** out) {
  bytes = align_up(bytes ? bytes : 1, 512);
  if (ctx->arena && ctx->arena_off + bytes <= ctx->arena_cap) {
    *out = ctx->arena + ctx->arena_off;
    ctx->arena_off += bytes;
    return TMAT_OK;
  }
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    tmat_set_error("arena_alloc: cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    cudaGetLastError();
    return TMAT_ERR_NOMEM;
  }
  ctx->overflow.push_back(p);
  ctx->overflow_bytes += bytes;
  *out = p;
  return TMAT_OK;
}

// Make sure the arena can hold `bytes` (plus what is already allocated) without overflowing.
int arena_reserve(tmat_ctx* ctx, size_t bytes) {
  size_t want = align_up(ctx->arena_off + bytes + (1 << 20), 1 << 20);
  if (ctx->arena && want <= ctx->arena_cap) return TMAT_OK;
  if (ctx->arena_off != 0) return TMAT_OK;   // live allocations: fall back to overflow blocks
  CUDA_TRY(cudaDeviceSynchronize());
  if (ctx->arena) CUDA_TRY(cudaFree(ctx->arena));
  ctx->arena = nullptr;
  ctx->arena_cap = 0;
  cudaError_t e = cudaMalloc((void**)&ctx->arena, want);
  if (e != cudaSuccess) {
    tmat_set_error("arena_reserve: cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    cudaGetLastError();
    return TMAT_ERR_NOMEM;
  }
  ctx->arena_cap = want;
  return TMAT_OK;
}

extern "C" int tmat_ctx_profile_enable(tmat_ctx* ctx, int on) {
  TMAT_REQUIRE(ctx, "tmat_ctx_profile_enable: ctx is NULL");
  ctx->profiling = on != 0;
  return TMAT_OK;
}

// out[0] = summed duration (ms) of the profiled tcgen05 convolution launches since the last collect,
// out[1] = their count, out[2] = the FLOPs (2*M*N*K) they executed.  Synchronises the device.
extern "C" int tmat_ctx_profile_collect(tmat_ctx* ctx, double* out3) {
  TMAT_REQUIRE(ctx && out3, "tmat_ctx_profile_collect: NULL argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  CUDA_TRY(cudaDeviceSynchronize());
  double ms = 0, fl = 0;
  for (auto& r : ctx->prof) {
    float t = 0;
    CUDA_TRY(cudaEventElapsedTime(&t, r.a, r.b));
    ms += t;
    fl += r.flops;
    ctx->event_pool.push_back(r.a);
    ctx->event_pool.push_back(r.b);
  }
  out3[0] = ms;
  out3[1] = (double)ctx->prof.size();
  out3[2] = fl;
  ctx->prof.clear();
  return TMAT_OK;
}
