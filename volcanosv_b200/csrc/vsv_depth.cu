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
