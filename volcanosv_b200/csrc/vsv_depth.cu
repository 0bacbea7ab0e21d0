// K6 span depth + genotype decision.
//
// span depth = count_reads_span_region (LI/correct_gt_del_real_data.py:140-146,
// LI/correct_gt_ins_real_data.py:188-196): per window, the number of records with
// reference_start < S and reference_end > E (no flag / mapq filter).
//
// Closed form (no sweep over the reads a window spans): with the reads sorted by start (BAM order)
// and a second copy of their ends sorted ascending,
//     depth(S,E) = #{start < S} - #{end <= E} + #{start in [S,E] and end <= E}
//                = lower_bound(start, S) - upper_bound(sorted_end, E) + (short scan of the reads that
//                  START inside the window)
// because a record with start >= S and end <= E starts inside [S,E] (end >= start).  Two binary searches
// and a scan of (window length x read-start density) elements per window, whatever the read lengths are;
// round 1's kernel walked every read between the first one that can reach E and S, i.e. O(coverage x
// longest read) per window (tens of thousands of reads with ONT's 500 kb tail).
// Cost per shard: one radix sort of the ends (8 n_reads bytes, the B3 term of SURVEY.md 8d).
#include <cub/device/device_radix_sort.cuh>
#include <math_constants.h>

#include "vsv_depth.cuh"

__global__ void __launch_bounds__(256)
k_span_depth_closed(const int32_t* __restrict__ rstart, const int32_t* __restrict__ rend,
                    const int32_t* __restrict__ send, int64_t n_reads, const int32_t* __restrict__ S,
                    const int32_t* __restrict__ E, int64_t n_win, int32_t* __restrict__ depth) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_win) return;
  const int32_t s = S[w];
  depth[w] = (s == VSV_WIN_OFF) ? 0 : span_depth_closed(rstart, rend, send, n_reads, s, E[w]);
}

// ascending copy of the read ends (temporary of one call)
static int sort_ends(vsv_ctx* ctx, const int32_t* re, int64_t n_reads, int32_t** out) {
  int32_t* se = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)n_reads, &se));
  *out = se;
  if (n_reads == 0) return VSV_OK;
  if (n_reads >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, "more than 2^31-1 reads per call");
  size_t tb = 0;
  VSV_CUDA(ctx, cub::DeviceRadixSort::SortKeys(nullptr, tb, re, se, (int)n_reads, 0, 32, ctx->stream));
  void* tmp = nullptr;
  VSV_TRY(vsv_arena_alloc(ctx, tb, &tmp));
  VSV_CUDA(ctx, cub::DeviceRadixSort::SortKeys(tmp, tb, re, se, (int)n_reads, 0, 32, ctx->stream));
  ctx->launches += 6;
  return VSV_OK;
}

static int span_depth_dev(vsv_ctx* ctx, const int32_t* rs, const int32_t* re, const int32_t* se, int64_t n_reads,
                          const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth) {
  const int32_t* sorted = se;
  if (!sorted) {
    int32_t* tmp = nullptr;
    VSV_TRY(sort_ends(ctx, re, n_reads, &tmp));
    sorted = tmp;
  }
  VSV_LAUNCH(ctx, k_span_depth_closed, (unsigned)vsv_cdiv(n_win, 256), 256, 0, rs, re, sorted, n_reads, S, E, n_win, depth);
  return VSV_OK;
}

static int span_depth_api(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, const int32_t* sorted_end,
                          int64_t n_reads, const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth, const char* fn) {
  VSV_TRY(vsv_begin(ctx));
  if (n_reads < 0 || n_win < 0) VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": bad argument");
  if (n_win == 0) return VSV_OK;
  const int32_t *rs, *re, *se, *dS, *dE;
  VSV_TRY(vsv_in(ctx, read_start, (size_t)n_reads, &rs));
  VSV_TRY(vsv_in(ctx, read_end, (size_t)n_reads, &re));
  VSV_TRY(vsv_in(ctx, sorted_end, (size_t)n_reads, &se));
  VSV_TRY(vsv_in(ctx, S, (size_t)n_win, &dS));
  VSV_TRY(vsv_in(ctx, E, (size_t)n_win, &dE));
  int32_t* dd;
  VSV_TRY(vsv_out(ctx, depth, (size_t)n_win, &dd));
  VSV_TRY(vsv_t0(ctx));
  VSV_TRY(span_depth_dev(ctx, rs, re, se, n_reads, dS, dE, n_win, dd));
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, depth, dd, (size_t)n_win));
  return vsv_end(ctx);
}

extern "C" int vsv_span_depth(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, int64_t n_reads,
                              const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth) {
  return span_depth_api(ctx, read_start, read_end, nullptr, n_reads, S, E, n_win, depth, "vsv_span_depth");
}
extern "C" int vsv_span_depth_sorted(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end,
                                     const int32_t* sorted_end, int64_t n_reads, const int32_t* S, const int32_t* E,
                                     int64_t n_win, int32_t* depth) {
  if (ctx && n_reads > 0 && !sorted_end) VSV_FAIL(ctx, VSV_E_ARG, "vsv_span_depth_sorted: sorted_end is NULL");
  return span_depth_api(ctx, read_start, read_end, sorted_end, n_reads, S, E, n_win, depth, "vsv_span_depth_sorted");
}

// --------------------------------------------------------------------------------
// genotype: windows per candidate -> depth -> ratio -> thresholds
// --------------------------------------------------------------------------------
// One thread per candidate: windows (GTD:149-168, GTI:188-196), closed-form depth, ratio, thresholds.
__global__ void __launch_bounds__(256)
k_genotype(const int32_t* __restrict__ rstart, const int32_t* __restrict__ rend, const int32_t* __restrict__ send,
           int64_t n_reads, const int32_t* __restrict__ pos, const int32_t* __restrict__ svlen,
           const int32_t* __restrict__ type, const int32_t* __restrict__ call_gt, const int32_t* __restrict__ nsup,
           int64_t n, GtParams gp, int32_t* __restrict__ depth2, double* __restrict__ ratio,
           int32_t* __restrict__ new_gt) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t d2, out;
  double r;
  genotype_one(rstart, rend, send, n_reads, pos[i], svlen[i], type[i], call_gt[i], nsup[i], gp, d2, r, out);
  depth2[i] = d2; ratio[i] = r; new_gt[i] = out;
}

extern "C" int vsv_genotype(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, int64_t n_reads,
                            const int32_t* cand_pos, const int32_t* cand_svlen, const int32_t* cand_type,
                            const int32_t* call_gt, const int32_t* n_support, int64_t n_cand,
                            const double* thr_del, const double* thr_ins, int32_t* depth2, double* ratio,
                            int32_t* new_gt) {
  VSV_TRY(vsv_begin(ctx));
  if (n_reads < 0 || n_cand < 0 || !thr_del || !thr_ins) VSV_FAIL(ctx, VSV_E_ARG, "vsv_genotype: bad argument");
  if (n_cand == 0) return VSV_OK;
  const int32_t *rs, *re, *cp, *cl, *ct, *cg, *ns;
  VSV_TRY(vsv_in(ctx, read_start, (size_t)n_reads, &rs));
  VSV_TRY(vsv_in(ctx, read_end, (size_t)n_reads, &re));
  VSV_TRY(vsv_in(ctx, cand_pos, (size_t)n_cand, &cp));
  VSV_TRY(vsv_in(ctx, cand_svlen, (size_t)n_cand, &cl));
  VSV_TRY(vsv_in(ctx, cand_type, (size_t)n_cand, &ct));
  VSV_TRY(vsv_in(ctx, call_gt, (size_t)n_cand, &cg));
  VSV_TRY(vsv_in(ctx, n_support, (size_t)n_cand, &ns));
  int32_t *d2, *ng, *se;
  double* rt;
  VSV_TRY(vsv_out(ctx, depth2, (size_t)n_cand, &d2));
  VSV_TRY(vsv_out(ctx, ratio, (size_t)n_cand, &rt));
  VSV_TRY(vsv_out(ctx, new_gt, (size_t)n_cand, &ng));
  GtParams gp = gt_params_of(ctx->params);
  for (int k = 0; k < 4; ++k) { gp.thr_del[k] = thr_del[k]; gp.thr_ins[k] = thr_ins[k]; }
  const unsigned g = (unsigned)vsv_cdiv(n_cand, 256);
  VSV_TRY(vsv_t0(ctx));
  VSV_TRY(sort_ends(ctx, re, n_reads, &se));
  VSV_LAUNCH(ctx, k_genotype, g, 256, 0, rs, re, (const int32_t*)se, n_reads, cp, cl, ct, cg, ns, n_cand, gp, d2, rt, ng);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, depth2, d2, (size_t)n_cand));
  VSV_TRY(vsv_out_done(ctx, ratio, rt, (size_t)n_cand));
  VSV_TRY(vsv_out_done(ctx, new_gt, ng, (size_t)n_cand));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// vsv_span_depth_noalloc: the span-depth entry for callers that cannot allocate
// --------------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  vsv_span_depth keeps two temporaries in the
// context's arena (the ascending copy of the read ends and the radix sort's work space); here both live in ONE file-scope
// device array of fixed size, so the call allocates nothing: no cudaMalloc / cudaFree, no arena, no allocating library
// routine (cub::DeviceRadixSort is handed its work space).  Every GPU that loads the library holds the array until exit.
//   * limit: n_reads <= VSV_NOALLOC_MAX_READS (2^21); more is VSV_E_CAPACITY.  n_win is unlimited (the caller owns S/E/depth).
//   * the array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
//     zero: the sort overwrites the n_reads ends it is about to read and clears its own counters on the stream.  Calls that
//     overlap on the GPU (two streams or two contexts of one device) race on it: serialise them.
//   * in host code `g_noalloc_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, taken once per
//     device with that device current (first call on the device; make that call outside a graph capture).
#define VSV_NOALLOC_SORT_BYTES ((size_t)VSV_NOALLOC_MAX_READS * sizeof(int32_t))
#define VSV_NOALLOC_WORK_BYTES (VSV_NOALLOC_SORT_BYTES + (size_t)(1 << 20))   // alternate key buffer + histograms of the sort
#define VSV_NOALLOC_BYTES (VSV_NOALLOC_SORT_BYTES + VSV_NOALLOC_WORK_BYTES)   // 17 MiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[VSV_NOALLOC_BYTES];

#define VSV_NOALLOC_MAX_DEVICES 64
static unsigned char* g_noalloc_ptr[VSV_NOALLOC_MAX_DEVICES];   // device address of the array, per device (nullptr: not taken yet)

extern "C" int vsv_span_depth_noalloc(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, int64_t n_reads,
                                      const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth) {
  if (!ctx) return VSV_E_ARG;
  ctx->err.clear();
  if (!ctx->device_io)
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_span_depth_noalloc: device pointers only (vsv_ctx_set_device_io(ctx, 1)): staging host buffers would allocate");
  if (n_reads < 0 || n_win < 0 || (n_reads > 0 && (!read_start || !read_end)) || (n_win > 0 && (!S || !E || !depth)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_span_depth_noalloc: bad argument");
  if (n_reads > (int64_t)VSV_NOALLOC_MAX_READS)
    VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_span_depth_noalloc: " + std::to_string((long long)n_reads) + " reads exceed the fixed scratch of " +
                                      std::to_string((long long)VSV_NOALLOC_MAX_READS) + " reads (VSV_NOALLOC_MAX_READS); use vsv_span_depth");
  if (ctx->device < 0 || ctx->device >= VSV_NOALLOC_MAX_DEVICES) VSV_FAIL(ctx, VSV_E_ARG, "vsv_span_depth_noalloc: device index out of range");
  if (n_win == 0) return VSV_OK;
  VSV_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned char* base = g_noalloc_ptr[ctx->device];
  if (base == nullptr) {   // set-up, once per device, with that device current
    void* p = nullptr;
    VSV_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (p == nullptr) VSV_FAIL(ctx, VSV_E_CUDA, "vsv_span_depth_noalloc: cudaGetSymbolAddress returned a null address");
    base = g_noalloc_ptr[ctx->device] = (unsigned char*)p;
  }
  int32_t* sorted = reinterpret_cast<int32_t*>(base);
  void* work = base + VSV_NOALLOC_SORT_BYTES;
  if (n_reads > 0) {
    size_t need = 0;
    VSV_CUDA(ctx, cub::DeviceRadixSort::SortKeys(nullptr, need, read_end, sorted, (int)n_reads, 0, 32, ctx->stream));
    if (need > VSV_NOALLOC_WORK_BYTES) VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_span_depth_noalloc: the sort's work space does not fit the fixed scratch");
    size_t have = VSV_NOALLOC_WORK_BYTES;
    VSV_CUDA(ctx, cub::DeviceRadixSort::SortKeys(work, have, read_end, sorted, (int)n_reads, 0, 32, ctx->stream));
    ctx->launches += 6;
  }
  VSV_LAUNCH(ctx, k_span_depth_closed, (unsigned)vsv_cdiv(n_win, 256), 256, 0, read_start, read_end, (const int32_t*)sorted, n_reads, S, E,
             n_win, depth);
  VSV_CUDA(ctx, cudaGetLastError());
  return VSV_OK;   // asynchronous on the context's stream, like every device-IO call
}
