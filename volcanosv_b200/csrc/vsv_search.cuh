// Binary searches and small CUB scan wrappers shared by the stage-2/3 kernels.
#pragma once
#include <cub/device/device_scan.cuh>

#include "vsv_common.cuh"

// --------------------------------------------------------------------------------
// binary searches on sorted int32 arrays
// --------------------------------------------------------------------------------
__device__ __forceinline__ int64_t lower_bound_i32(const int32_t* __restrict__ a, int64_t lo, int64_t hi, int32_t x) {
  while (lo < hi) {  // first i with a[i] >= x
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int64_t upper_bound_i32(const int32_t* __restrict__ a, int64_t lo, int64_t hi, int32_t x) {
  while (lo < hi) {  // first i with a[i] > x
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(a + mid) <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
// comparisons against a double bound (the reference compares ints with floats)
__device__ __forceinline__ int64_t lower_bound_f64(const int32_t* __restrict__ a, int64_t lo, int64_t hi, double x) {
  while (lo < hi) {  // first i with a[i] >= x
    int64_t mid = (lo + hi) >> 1;
    if ((double)__ldg(a + mid) < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int64_t upper_bound_f64(const int32_t* __restrict__ a, int64_t lo, int64_t hi, double x) {
  while (lo < hi) {  // first i with a[i] > x
    int64_t mid = (lo + hi) >> 1;
    if ((double)__ldg(a + mid) <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

struct MaxI32 {
  __host__ __device__ int32_t operator()(int32_t a, int32_t b) const { return a > b ? a : b; }
};
struct LastValid {  // inclusive scan "last value >= 0 seen so far"
  __host__ __device__ int32_t operator()(int32_t a, int32_t b) const { return b >= 0 ? b : a; }
};

template <typename Op>
static int inclusive_scan_i32(vsv_ctx* c, const int32_t* in, int32_t* out, int64_t n, Op op) {
  if (n <= 0) return VSV_OK;
  size_t tb = 0;
  VSV_CUDA(c, cub::DeviceScan::InclusiveScan(nullptr, tb, in, out, op, n, c->stream));
  void* tmp = nullptr;
  VSV_TRY(vsv_arena_alloc(c, tb, &tmp));
  VSV_CUDA(c, cub::DeviceScan::InclusiveScan(tmp, tb, in, out, op, n, c->stream));
  c->launches += 2;
  return VSV_OK;
}

