// Closed-form spanning depth and the genotype decision as device functions (shared by vsv_depth.cu and the
// resident shard engine).  Reference: LI/correct_gt_del_real_data.py:140-170,206-217,291-318;
// LI/correct_gt_ins_real_data.py:188-196,228,258-283.
#pragma once
#include <math_constants.h>

#include "vsv_search.cuh"

#define VSV_WIN_OFF INT32_MIN

// number of records with start < S and end > E; rstart ascending (BAM order), send = ends ascending
__device__ __forceinline__ int32_t span_depth_closed(const int32_t* __restrict__ rstart, const int32_t* __restrict__ rend,
                                                     const int32_t* __restrict__ send, int64_t n, int32_t S, int32_t E) {
  const int64_t a = lower_bound_i32(rstart, 0, n, S);   // #{start < S}
  const int64_t ue = upper_bound_i32(send, 0, n, E);    // #{end <= E}
  int64_t c = 0;                                        // #{start in [S,E] and end <= E}
  if (E >= S) {
    const int64_t b = upper_bound_i32(rstart, a, n, E);
    for (int64_t i = a; i < b; ++i) c += (__ldg(rend + i) <= E) ? 1 : 0;
  }
  return (int32_t)(a - ue + c);
}

struct GtParams {
  int32_t ins_flank, del_large, del_off, del_win;
  double thr_del[4], thr_ins[4];
};
static inline GtParams gt_params_of(const vsv_params& p) {
  GtParams g;
  g.ins_flank = p.gt_ins_flank; g.del_large = p.gt_del_large; g.del_off = p.gt_del_off; g.del_win = p.gt_del_win;
  for (int k = 0; k < 4; ++k) { g.thr_del[k] = p.thr_del[k]; g.thr_ins[k] = p.thr_ins[k]; }
  return g;
}

// One candidate: windows -> depth (x2, exact: the mean of two windows may be x.5) -> ratio -> thresholds.
__device__ __forceinline__ void genotype_one(const int32_t* __restrict__ rstart, const int32_t* __restrict__ rend,
                                             const int32_t* __restrict__ send, int64_t n_reads, int32_t p, int32_t L,
                                             int32_t type, int32_t g, int32_t ns, const GtParams& gp, int32_t& d2,
                                             double& r, int32_t& out) {
  if (type == VSV_INS) {            // GTI:188-196, flanking 100 (GTI:327)
    d2 = 2 * span_depth_closed(rstart, rend, send, n_reads, p - gp.ins_flank, p + gp.ins_flank);
  } else if (L <= gp.del_large) {   // GTD:150-154
    d2 = 2 * span_depth_closed(rstart, rend, send, n_reads, p, p + L);
  } else {                          // GTD:157-168: two 100 bp windows 150 bp off the breakpoints, mean
    const int32_t s0 = p - gp.del_off, s1 = p + L + gp.del_off;
    d2 = span_depth_closed(rstart, rend, send, n_reads, s0, s0 + gp.del_win) +
         span_depth_closed(rstart, rend, send, n_reads, s1, s1 + gp.del_win);
  }
  const double depth = (double)d2 / 2.0;  // GTD:168
  if (type == VSV_DEL) {
    r = (d2 == 0) ? 1.0 : (double)ns / depth;  // GTD:213-217
  } else {                                      // GTI:228 (pandas division: inf / nan)
    r = (d2 == 0) ? (ns == 0 ? CUDART_NAN : CUDART_INF) : (double)ns / depth;
  }
  out = g;
  if (g == 0 || g == 1) {
    const double* t4 = type == VSV_DEL ? gp.thr_del : gp.thr_ins;
    const bool large = L > gp.del_large;
    const double t = g == 1 ? (large ? t4[0] : t4[1]) : (large ? t4[2] : t4[3]);
    if (!isnan(t)) {  // GTD:302-318
      if (r > t) out = 1;
      else if (r <= t) out = 0;
    }
  }
}
