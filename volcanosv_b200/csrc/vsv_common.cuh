// Shared plumbing of libvolcano_b200.so: context, error handling, host/device IO
// staging.  Everything that matters for speed lives in the kernel files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/volcanosv_b200.h"

#define VSV_SM_COUNT_B200 148

struct vsv_ctx {
  int device = 0;
  int sm_count = VSV_SM_COUNT_B200;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t evt0 = nullptr, evt1 = nullptr;  // user stopwatch
  cudaEvent_t evk0 = nullptr, evk1 = nullptr;  // dominant kernel of the last scan call
  bool main_open = false;
  float last_main_ms = 0.f;
  bool device_io = false;
  bool timing_open = false;
  float last_kernel_ms = 0.f;
  int64_t launches = 0;
  std::string err;
  // bump scratch arena on the device, reset at the start of every API call
  char* arena = nullptr;
  size_t arena_cap = 0, arena_used = 0;
  std::vector<void*> overflow;  // allocations that did not fit the arena (freed at reset)
  vsv_params params;            // reference constants (vsv_params_default; replaced by vsv_ctx_set_params)
  bool scan_attr_set = false;   // cudaFuncSetAttribute of the streaming kernels done
};

#define VSV_CUDA(ctx, call)                                                            \
  do {                                                                                 \
    cudaError_t e__ = (call);                                                          \
    if (e__ != cudaSuccess) {                                                          \
      (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                \
      return VSV_E_CUDA;                                                               \
    }                                                                                  \
  } while (0)

#define VSV_FAIL(ctx, code, msg) \
  do {                           \
    (ctx)->err = (msg);          \
    return (code);               \
  } while (0)

#define VSV_TRY(expr)        \
  do {                       \
    int r__ = (expr);        \
    if (r__ != VSV_OK) return r__; \
  } while (0)

int vsv_arena_reset(vsv_ctx* ctx);
int vsv_arena_alloc(vsv_ctx* ctx, size_t bytes, void** out);

template <typename T>
static inline int vsv_tmp(vsv_ctx* ctx, size_t n, T** out) {
  void* p = nullptr;
  int r = vsv_arena_alloc(ctx, (n ? n : 1) * sizeof(T), &p);
  *out = (T*)p;
  return r;
}

// Input staging: host pointer -> device copy in the arena (async on ctx->stream), or
// the pointer itself when the ctx is in device-IO mode.  nullptr stays nullptr.
template <typename T>
static inline int vsv_in(vsv_ctx* ctx, const T* p, size_t n, const T** dev) {
  if (p == nullptr || ctx->device_io) {
    *dev = p;
    return VSV_OK;
  }
  T* d = nullptr;
  VSV_TRY(vsv_tmp(ctx, n, &d));
  if (n) VSV_CUDA(ctx, cudaMemcpyAsync(d, p, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  *dev = d;
  return VSV_OK;
}

// Output staging: device buffer the kernels write to.
template <typename T>
static inline int vsv_out(vsv_ctx* ctx, T* p, size_t n, T** dev) {
  if (p == nullptr || ctx->device_io) {
    *dev = p;
    return VSV_OK;
  }
  return vsv_tmp(ctx, n, dev);
}

template <typename T>
static inline int vsv_out_done(vsv_ctx* ctx, T* host, const T* dev, size_t n) {
  if (host == nullptr || ctx->device_io || n == 0) return VSV_OK;
  VSV_CUDA(ctx, cudaMemcpyAsync(host, dev, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  return VSV_OK;
}

static inline int vsv_begin(vsv_ctx* ctx) {
  if (!ctx) return VSV_E_ARG;
  ctx->err.clear();
  VSV_CUDA(ctx, cudaSetDevice(ctx->device));
  VSV_TRY(vsv_arena_reset(ctx));
  return VSV_OK;
}
// kernel-time bracket (events on the launching stream)
static inline int vsv_t0(vsv_ctx* ctx) {
  VSV_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  return VSV_OK;
}
static inline int vsv_t1(vsv_ctx* ctx) {
  VSV_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  ctx->timing_open = true;
  return VSV_OK;
}
// end of an API call: host-IO mode waits for the copies; device-IO mode stays async.
static inline int vsv_end(vsv_ctx* ctx, bool force_sync = false) {
  VSV_CUDA(ctx, cudaGetLastError());
  if (!ctx->device_io || force_sync) VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSV_OK;
}

#define VSV_LAUNCH(ctx, kernel, grid, block, smem, ...)                         \
  do {                                                                          \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);            \
    (ctx)->launches++;                                                          \
  } while (0)

static inline int64_t vsv_cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- small device helpers shared by the kernels --------------------------------
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_cg_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

// device-wide exclusive scan of int32 -> int64 (small helper arrays only)
int vsv_exclusive_scan_i32(vsv_ctx* ctx, const int32_t* in, int64_t n, int64_t* out_excl,
                           int64_t* total_dev);
