// Context, scratch arena and pinned-memory helpers of libvolcano_b200.so.
#include <cub/device/device_scan.cuh>

#include "vsv_common.cuh"

static const size_t kArenaInitial = size_t(256) << 20;
static const size_t kAlign = 256;

extern "C" const char* vsv_version(void) { return "volcanosv_b200 0.1 (sm_100a)"; }

extern "C" int vsv_ctx_create(int device, vsv_ctx** out) {
  if (!out) return VSV_E_ARG;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
    // no CPU fallback by design: the caller must fail loudly
    return VSV_E_CUDA;
  }
  vsv_ctx* c = new vsv_ctx();
  c->device = device;
  if (cudaSetDevice(device) != cudaSuccess) { delete c; return VSV_E_CUDA; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
      cudaEventCreate(&c->evt0) != cudaSuccess || cudaEventCreate(&c->evt1) != cudaSuccess ||
      cudaEventCreate(&c->evk0) != cudaSuccess || cudaEventCreate(&c->evk1) != cudaSuccess) {
    delete c;
    return VSV_E_CUDA;
  }
  if (cudaMalloc((void**)&c->arena, kArenaInitial) != cudaSuccess) {
    delete c;
    return VSV_E_NOMEM;
  }
  c->arena_cap = kArenaInitial;
  *out = c;
  return VSV_OK;
}

extern "C" void vsv_ctx_destroy(vsv_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (void* p : c->overflow) cudaFree(p);
  if (c->arena) cudaFree(c->arena);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->evt0) cudaEventDestroy(c->evt0);
  if (c->evt1) cudaEventDestroy(c->evt1);
  if (c->evk0) cudaEventDestroy(c->evk0);
  if (c->evk1) cudaEventDestroy(c->evk1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

extern "C" const char* vsv_last_error(vsv_ctx* c) { return c ? c->err.c_str() : "null ctx"; }

extern "C" int vsv_ctx_set_device_io(vsv_ctx* c, int on) {
  if (!c) return VSV_E_ARG;
  c->device_io = on != 0;
  return VSV_OK;
}

extern "C" int vsv_sync(vsv_ctx* c) {
  if (!c) return VSV_E_ARG;
  VSV_CUDA(c, cudaStreamSynchronize(c->stream));
  return VSV_OK;
}

extern "C" int vsv_alloc_pinned(void** p, size_t bytes) {
  if (!p) return VSV_E_ARG;
  return cudaHostAlloc(p, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? VSV_OK : VSV_E_NOMEM;
}
extern "C" int vsv_free_pinned(void* p) { return cudaFreeHost(p) == cudaSuccess ? VSV_OK : VSV_E_CUDA; }

extern "C" int vsv_get_timing(vsv_ctx* c, float* kernel_ms, int64_t* launches) {
  if (!c) return VSV_E_ARG;
  if (c->timing_open) {
    VSV_CUDA(c, cudaEventSynchronize(c->ev1));
    VSV_CUDA(c, cudaEventElapsedTime(&c->last_kernel_ms, c->ev0, c->ev1));
    c->timing_open = false;
  }
  if (kernel_ms) *kernel_ms = c->last_kernel_ms;
  if (launches) *launches = c->launches;
  return VSV_OK;
}

extern "C" int vsv_timer_start(vsv_ctx* c) {
  if (!c) return VSV_E_ARG;
  VSV_CUDA(c, cudaSetDevice(c->device));
  VSV_CUDA(c, cudaStreamSynchronize(c->stream));
  VSV_CUDA(c, cudaEventRecord(c->evt0, c->stream));
  return VSV_OK;
}
extern "C" int vsv_timer_stop(vsv_ctx* c, float* ms) {
  if (!c || !ms) return VSV_E_ARG;
  VSV_CUDA(c, cudaEventRecord(c->evt1, c->stream));
  VSV_CUDA(c, cudaEventSynchronize(c->evt1));
  VSV_CUDA(c, cudaEventElapsedTime(ms, c->evt0, c->evt1));
  return VSV_OK;
}
extern "C" int vsv_get_main_kernel_ms(vsv_ctx* c, float* ms) {
  if (!c || !ms) return VSV_E_ARG;
  if (c->main_open) {
    VSV_CUDA(c, cudaEventSynchronize(c->evk1));
    VSV_CUDA(c, cudaEventElapsedTime(&c->last_main_ms, c->evk0, c->evk1));
    c->main_open = false;
  }
  *ms = c->last_main_ms;
  return VSV_OK;
}

int vsv_arena_reset(vsv_ctx* c) {
  if (!c->overflow.empty()) {
    // the previous call outgrew the arena: wait for it, free the spill, grow once
    VSV_CUDA(c, cudaStreamSynchronize(c->stream));
    for (void* p : c->overflow) cudaFree(p);
    c->overflow.clear();
    size_t want = c->arena_used + (c->arena_used >> 2) + (size_t(64) << 20);
    if (want > c->arena_cap) {
      cudaFree(c->arena);
      c->arena = nullptr;
      c->arena_cap = 0;
      if (cudaMalloc((void**)&c->arena, want) != cudaSuccess) {
        cudaGetLastError();
        c->err = "scratch arena allocation failed";
        return VSV_E_NOMEM;
      }
      c->arena_cap = want;
    }
  }
  c->arena_used = 0;
  return VSV_OK;
}

int vsv_arena_alloc(vsv_ctx* c, size_t bytes, void** out) {
  size_t need = (bytes + kAlign - 1) / kAlign * kAlign;
  if (c->arena_used + need <= c->arena_cap && c->overflow.empty()) {
    *out = c->arena + c->arena_used;
    c->arena_used += need;
    return VSV_OK;
  }
  void* p = nullptr;
  if (cudaMalloc(&p, need) != cudaSuccess) {
    cudaGetLastError();
    c->err = "device allocation of " + std::to_string(need) + " bytes failed";
    return VSV_E_NOMEM;
  }
  c->overflow.push_back(p);
  c->arena_used += need;  // bookkeeping for the next reset's growth
  *out = p;
  return VSV_OK;
}

// exclusive scan int32 -> int64 with the grand total in total_dev (device, may be null)
__global__ void k_scan_total(const int32_t* in, const int64_t* excl, int64_t n, int64_t* total) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *total = n ? excl[n - 1] + in[n - 1] : 0;
}

struct SumI64 {
  __host__ __device__ int64_t operator()(int64_t a, int64_t b) const { return a + b; }
};

int vsv_exclusive_scan_i32(vsv_ctx* c, const int32_t* in, int64_t n, int64_t* out_excl,
                           int64_t* total_dev) {
  if (n > 0) {
    // int64 init value => int64 accumulator (counts can exceed 2^31 in total)
    size_t tb = 0;
    VSV_CUDA(c, cub::DeviceScan::ExclusiveScan(nullptr, tb, in, out_excl, SumI64(), (int64_t)0, n, c->stream));
    void* tmp = nullptr;
    VSV_TRY(vsv_arena_alloc(c, tb, &tmp));
    VSV_CUDA(c, cub::DeviceScan::ExclusiveScan(tmp, tb, in, out_excl, SumI64(), (int64_t)0, n, c->stream));
    c->launches += 2;
  }
  if (total_dev) VSV_LAUNCH(c, k_scan_total, 1, 32, 0, in, out_excl, n, total_dev);
  return VSV_OK;
}
