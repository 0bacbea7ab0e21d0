// K6 span depth + genotype decision.
//
// span depth = count_reads_span_region (LI/correct_gt_del_real_data.py:140-146,
// LI/correct_gt_ins_real_data.py:188-196): per window, the number of records with
// reference_start < S and reference_end > E (no flag / mapq filter).
//
// Reads are sorted by start.  hi = first read with start >= S; lo = first read whose
// running max of `end` exceeds E (everything before it ends at or before E).  Only
// reads in [lo,hi) can span; among those the predicate is end > E.
// A CTA takes 64 consecutive windows.  When their ranges overlap (dense candidates:
// the stress sweep) the union range is staged through shared memory in 16 KB tiles and
// every warp counts its windows from the tile with ballot+popc; sparse windows read
// their own range straight from L2/HBM (each byte still read once per CTA).
#include <math_constants.h>

#include "vsv_search.cuh"

#define DEPTH_THREADS 256
#define DEPTH_WARPS (DEPTH_THREADS / 32)
#define DEPTH_WPB 64      // windows per block
#define DEPTH_WPW (DEPTH_WPB / DEPTH_WARPS)
#define DEPTH_CHUNK 4096  // read ends per shared-memory tile (16 KB)
#define VSV_WIN_OFF INT32_MIN

__global__ void __launch_bounds__(DEPTH_THREADS)
k_span_depth(const int32_t* __restrict__ rstart, const int32_t* __restrict__ rend,
             const int32_t* __restrict__ pmax_end, int64_t n_reads, const int32_t* __restrict__ S,
             const int32_t* __restrict__ E, int64_t n_win, int32_t* __restrict__ depth) {
  __shared__ int32_t s_end[DEPTH_CHUNK];
  __shared__ int64_t s_lo[DEPTH_WPB], s_hi[DEPTH_WPB];
  __shared__ int32_t s_E[DEPTH_WPB];
  __shared__ int64_t s_union[2];
  __shared__ unsigned long long s_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t w0 = (int64_t)blockIdx.x * DEPTH_WPB;
  if (tid < DEPTH_WPB) {
    const int64_t w = w0 + tid;
    int64_t lo = 0, hi = 0;
    int32_t e = 0;
    if (w < n_win) {
      const int32_t s = S[w];
      e = E[w];
      if (s != VSV_WIN_OFF) {
        hi = lower_bound_i32(rstart, 0, n_reads, s);
        lo = upper_bound_i32(pmax_end, 0, hi, e);
      }
    }
    s_lo[tid] = lo; s_hi[tid] = hi; s_E[tid] = e;
  }
  __syncthreads();
  if (warp == 0) {  // union of the 64 ranges and their summed length
    int64_t mn = INT64_MAX, mx = 0;
    unsigned long long tot = 0;
    for (int k = lane; k < DEPTH_WPB; k += 32) {
      if (s_hi[k] > s_lo[k]) {
        mn = min(mn, s_lo[k]); mx = max(mx, s_hi[k]);
        tot += (unsigned long long)(s_hi[k] - s_lo[k]);
      }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, d));
      mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
      tot += __shfl_xor_sync(0xffffffffu, tot, d);
    }
    if (lane == 0) { s_union[0] = mn; s_union[1] = mx; s_total = tot; }
  }
  __syncthreads();
  const int64_t u0 = s_union[0], u1 = s_union[1];
  // stage through shared memory when the windows re-read each element >= 2x on average
  const bool tiled = (u1 > u0) && ((unsigned long long)(u1 - u0) * 2ull <= s_total);
  int32_t cnt[DEPTH_WPW];
#pragma unroll
  for (int q = 0; q < DEPTH_WPW; ++q) cnt[q] = 0;
  if (tiled) {
    for (int64_t base = u0; base < u1; base += DEPTH_CHUNK) {
      const int64_t lim = min((int64_t)DEPTH_CHUNK, u1 - base);
      for (int i = tid; i < lim; i += DEPTH_THREADS) s_end[i] = __ldg(rend + base + i);
      __syncthreads();
#pragma unroll
      for (int q = 0; q < DEPTH_WPW; ++q) {
        const int k = warp * DEPTH_WPW + q;
        const int a = (int)(max(s_lo[k], base) - base);
        const int b = (int)(min(s_hi[k], base + lim) - base);
        const int32_t e = s_E[k];
        for (int i0 = a; i0 < b; i0 += 32) {  // warp-uniform trip count
          const int i = i0 + lane;
          const bool p = (i < b) && (s_end[i] > e);
          cnt[q] += __popc(__ballot_sync(0xffffffffu, p));
        }
      }
      __syncthreads();
    }
  } else {
#pragma unroll
    for (int q = 0; q < DEPTH_WPW; ++q) {
      const int k = warp * DEPTH_WPW + q;
      const int64_t a = s_lo[k], b = s_hi[k];
      const int32_t e = s_E[k];
      for (int64_t i0 = a; i0 < b; i0 += 32) {
        const int64_t i = i0 + lane;
        const bool p = (i < b) && (__ldg(rend + i) > e);
        cnt[q] += __popc(__ballot_sync(0xffffffffu, p));
      }
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int q = 0; q < DEPTH_WPW; ++q) {
      const int64_t w = w0 + warp * DEPTH_WPW + q;
      if (w < n_win) depth[w] = cnt[q];
    }
  }
}

static int span_depth_dev(vsv_ctx* ctx, const int32_t* rs, const int32_t* re, int64_t n_reads, const int32_t* S,
                          const int32_t* E, int64_t n_win, int32_t* depth) {
  int32_t* pm = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)n_reads, &pm));
  VSV_TRY(inclusive_scan_i32(ctx, re, pm, n_reads, MaxI32()));
  VSV_LAUNCH(ctx, k_span_depth, (unsigned)vsv_cdiv(n_win, DEPTH_WPB), DEPTH_THREADS, 0, rs, re, pm, n_reads, S, E,
             n_win, depth);
  return VSV_OK;
}

extern "C" int vsv_span_depth(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, int64_t n_reads,
                              const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth) {
  VSV_TRY(vsv_begin(ctx));
  if (n_reads < 0 || n_win < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_span_depth: bad argument");
  if (n_win == 0) return VSV_OK;
  const int32_t *rs, *re, *dS, *dE;
  VSV_TRY(vsv_in(ctx, read_start, (size_t)n_reads, &rs));
  VSV_TRY(vsv_in(ctx, read_end, (size_t)n_reads, &re));
  VSV_TRY(vsv_in(ctx, S, (size_t)n_win, &dS));
  VSV_TRY(vsv_in(ctx, E, (size_t)n_win, &dE));
  int32_t* dd;
  VSV_TRY(vsv_out(ctx, depth, (size_t)n_win, &dd));
  VSV_TRY(vsv_t0(ctx));
  VSV_TRY(span_depth_dev(ctx, rs, re, n_reads, dS, dE, n_win, dd));
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, depth, dd, (size_t)n_win));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// genotype: windows per candidate -> depth -> ratio -> thresholds
// --------------------------------------------------------------------------------
__global__ void k_gt_windows(const int32_t* __restrict__ pos, const int32_t* __restrict__ svlen,
                             const int32_t* __restrict__ type, int64_t n, int32_t* __restrict__ S,
                             int32_t* __restrict__ E) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t p = pos[i], L = svlen[i];
  int32_t s0, e0, s1 = VSV_WIN_OFF, e1 = 0;
  if (type[i] == VSV_INS) {  // GTI:188-196, flanking 100 (GTI:327)
    s0 = p - 100; e0 = p + 100;
  } else if (L <= 1000) {    // GTD:150-154
    s0 = p; e0 = p + L;
  } else {                   // GTD:157-168: two 100 bp windows 150 bp off the breakpoints
    s0 = p - 150; e0 = s0 + 100;
    s1 = p + L + 150; e1 = s1 + 100;
  }
  S[2 * i] = s0; E[2 * i] = e0; S[2 * i + 1] = s1; E[2 * i + 1] = e1;
}

__global__ void k_gt_decide(const int32_t* __restrict__ svlen, const int32_t* __restrict__ type,
                            const int32_t* __restrict__ call_gt, const int32_t* __restrict__ nsup,
                            const int32_t* __restrict__ S, const int32_t* __restrict__ wdepth, int64_t n,
                            double4 thr_del, double4 thr_ins, int32_t* __restrict__ depth2,
                            double* __restrict__ ratio, int32_t* __restrict__ new_gt) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool two = S[2 * i + 1] != VSV_WIN_OFF;
  const int32_t d2 = two ? wdepth[2 * i] + wdepth[2 * i + 1] : 2 * wdepth[2 * i];
  const double depth = (double)d2 / 2.0;  // GTD:168
  const int32_t ns = nsup[i];
  double r;
  if (type[i] == VSV_DEL) {
    r = (d2 == 0) ? 1.0 : (double)ns / depth;  // GTD:213-217
  } else {                                      // GTI:228 (pandas division: inf / nan)
    r = (d2 == 0) ? (ns == 0 ? CUDART_NAN : CUDART_INF) : (double)ns / depth;
  }
  const int32_t g = call_gt[i];
  int32_t out = g;
  if (g == 0 || g == 1) {
    const double4 t4 = type[i] == VSV_DEL ? thr_del : thr_ins;
    const bool large = svlen[i] > 1000;
    const double t = g == 1 ? (large ? t4.x : t4.y) : (large ? t4.z : t4.w);
    if (!isnan(t)) {  // GTD:302-318
      if (r > t) out = 1;
      else if (r <= t) out = 0;
    }
  }
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
  int32_t *d2, *ng, *S, *E, *wd;
  double* rt;
  VSV_TRY(vsv_out(ctx, depth2, (size_t)n_cand, &d2));
  VSV_TRY(vsv_out(ctx, ratio, (size_t)n_cand, &rt));
  VSV_TRY(vsv_out(ctx, new_gt, (size_t)n_cand, &ng));
  VSV_TRY(vsv_tmp(ctx, (size_t)2 * n_cand, &S));
  VSV_TRY(vsv_tmp(ctx, (size_t)2 * n_cand, &E));
  VSV_TRY(vsv_tmp(ctx, (size_t)2 * n_cand, &wd));
  const double4 td = make_double4(thr_del[0], thr_del[1], thr_del[2], thr_del[3]);
  const double4 ti = make_double4(thr_ins[0], thr_ins[1], thr_ins[2], thr_ins[3]);
  const unsigned g = (unsigned)vsv_cdiv(n_cand, 256);
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_gt_windows, g, 256, 0, cp, cl, ct, n_cand, S, E);
  VSV_TRY(span_depth_dev(ctx, rs, re, n_reads, S, E, 2 * n_cand, wd));
  VSV_LAUNCH(ctx, k_gt_decide, g, 256, 0, cl, ct, cg, ns, S, wd, n_cand, td, ti, d2, rt, ng);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, depth2, d2, (size_t)n_cand));
  VSV_TRY(vsv_out_done(ctx, ratio, rt, (size_t)n_cand));
  VSV_TRY(vsv_out_done(ctx, new_gt, ng, (size_t)n_cand));
  return vsv_end(ctx);
}
