// Context, scratch arena and pinned-memory helpers of libvolcano_b200.so.
#include <cub/device/device_scan.cuh>

#include "vsv_common.cuh"

static const size_t kArenaInitial = size_t(256) << 20;
static const size_t kAlign = 256;

extern "C" const char* vsv_version(void) { return "volcanosv_b200 0.1 (sm_100a)"; }

// Reference defaults: argparse defaults and literals of the LI scripts, LI/para/
// GT_correction_para_<dtype>_<type>.txt, LI/filter_para.csv rows asm=volcano.
extern "C" int vsv_params_default(vsv_params* p, int dtype) {
  if (!p || dtype < 0 || dtype > 2) return VSV_E_ARG;
  memset(p, 0, sizeof(*p));
  p->dtype = dtype;
  p->reads_min_len = 30; p->reads_min_mapq = 50;
  p->cute_min_len = 10; p->cute_min_mapq = 20; p->cute_min_read_len = 500;
  p->sa_sv_size = 30; p->sa_max_size = 100000; p->sa_max_parts = 7; p->batch_size = 10000000;
  p->contig_min_len = 30; p->contig_min_mapq = 50; p->cluster_max_shift = 100; p->pair_max_shift = 200;
  p->clr_max_ins_pct = 0.13; p->clr_min_var_dist = 200.0;
  p->fp_max_comp = 250; p->fp_max_shift = 500;
  p->cov_flanking = 1000; p->cov_min_size = 30;
  static const double cov[3][4] = {{0.097, 2.754, 0.2, 2.605}, {0.206, 2.79, 0.242, 2.77}, {0.075, 2.383, 0.186, 3.018}};
  p->cov_lb_ins = cov[dtype][0]; p->cov_rb_ins = cov[dtype][1]; p->cov_lb_del = cov[dtype][2]; p->cov_rb_del = cov[dtype][3];
  p->gt_ins_min_len = 30; p->gt_min_size_sim = 0.6; p->gt_max_shift_ratio = 2.3; p->gt_min_shift = 500.0;
  p->gt_ins_flank = 100; p->gt_del_large = 1000; p->gt_del_off = 150; p->gt_del_win = 100;
  const double nan = __builtin_nan("");
  static const double td[3][4] = {{0.6, 0.69, 0.71, 0.91}, {0.61, 0.61, 0.68, 0.79}, {0.55, 0.59, 0.65, 0.75}};
  const double ti[3][4] = {{nan, 0.59, 0.63, 0.79}, {nan, nan, 0.67, 0.72}, {nan, nan, 0.64, 0.75}};
  for (int k = 0; k < 4; ++k) { p->thr_del[k] = td[dtype][k]; p->thr_ins[k] = ti[dtype][k]; }
  return VSV_OK;
}
extern "C" int vsv_ctx_set_params(vsv_ctx* c, const vsv_params* p) {
  if (!c || !p) return VSV_E_ARG;
  if (p->sa_max_parts < 2 || p->sa_max_parts > 16 || p->batch_size <= 0 || p->cov_flanking < 0 || p->reads_min_len < 0 ||
      p->cute_min_len < 0 || p->reads_min_len >= (1 << 27) || p->cute_min_len >= (1 << 27)) {
    c->err = "vsv_ctx_set_params: parameter out of range";
    return VSV_E_ARG;
  }
  c->params = *p;
  return VSV_OK;
}
extern "C" int vsv_ctx_get_params(vsv_ctx* c, vsv_params* p) {
  if (!c || !p) return VSV_E_ARG;
  *p = c->params;
  return VSV_OK;
}

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
  vsv_params_default(&c->params, VSV_DTYPE_HIFI);
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
