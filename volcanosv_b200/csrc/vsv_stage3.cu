// Stage 3: genotyping.  Signature grouping (GTD:65-86), supporting-signature count
// with the reference's last_match_idx double count (GTD:92-137, GTI:107-166), spanning
// read depth as a shared-memory tiled interval join (GTD:140-170, GTI:188-196) and the
// ratio / threshold decision (GTD:206-217,291-318; GTI:228,258-283).
#include "vsv_search.cuh"

// --------------------------------------------------------------------------------
// group_sigs: run-length encode adjacent equal (chrom,pos,len)
// --------------------------------------------------------------------------------
__global__ void k_group_heads(const int32_t* __restrict__ c, const int32_t* __restrict__ p,
                              const int32_t* __restrict__ l, int64_t n, int min_len,
                              int32_t* __restrict__ head) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // rows shorter than min_len are dropped before grouping (GTI:81); equal keys share a len
  head[i] = (l[i] < min_len) ? 0 : (i == 0 || (c && c[i] != c[i - 1]) || p[i] != p[i - 1] || l[i] != l[i - 1]) ? 1 : 0;
}
__global__ void k_group_write(const int32_t* __restrict__ c, const int32_t* __restrict__ p,
                              const int32_t* __restrict__ l, int64_t n, const int32_t* __restrict__ head,
                              const int64_t* __restrict__ excl, int32_t* __restrict__ uc, int32_t* __restrict__ up,
                              int32_t* __restrict__ ul, int32_t* __restrict__ ucnt) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !head[i]) return;
  int64_t e = i + 1;
  while (e < n && !head[e] && p[e] == p[i] && l[e] == l[i] && (!c || c[e] == c[i])) ++e;
  int64_t o = excl[i];
  uc[o] = c ? c[i] : 0; up[o] = p[i]; ul[o] = l[i]; ucnt[o] = (int32_t)(e - i);
}

extern "C" int vsv_group_sigs(vsv_ctx* ctx, const int32_t* chrom, const int32_t* pos, const int32_t* len,
                              int64_t n, int min_len, int32_t* u_chrom, int32_t* u_pos, int32_t* u_len,
                              int32_t* u_cnt, int64_t* n_unique) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0 || !n_unique) VSV_FAIL(ctx, VSV_E_ARG, "vsv_group_sigs: bad argument");
  *n_unique = 0;
  if (n == 0) return VSV_OK;
  const int32_t *dc, *dp, *dl;
  VSV_TRY(vsv_in(ctx, chrom, (size_t)n, &dc));
  VSV_TRY(vsv_in(ctx, pos, (size_t)n, &dp));
  VSV_TRY(vsv_in(ctx, len, (size_t)n, &dl));
  int32_t *oc, *op, *ol, *on, *head;
  VSV_TRY(vsv_out(ctx, u_chrom, (size_t)n, &oc));
  VSV_TRY(vsv_out(ctx, u_pos, (size_t)n, &op));
  VSV_TRY(vsv_out(ctx, u_len, (size_t)n, &ol));
  VSV_TRY(vsv_out(ctx, u_cnt, (size_t)n, &on));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &head));
  int64_t *excl, *tot;
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &excl));
  VSV_TRY(vsv_tmp(ctx, 1, &tot));
  const unsigned g = (unsigned)vsv_cdiv(n, 256);
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_group_heads, g, 256, 0, dc, dp, dl, n, min_len, head);
  VSV_TRY(vsv_exclusive_scan_i32(ctx, head, n, excl, tot));
  VSV_LAUNCH(ctx, k_group_write, g, 256, 0, dc, dp, dl, n, head, excl, oc, op, ol, on);
  VSV_TRY(vsv_t1(ctx));
  VSV_CUDA(ctx, cudaMemcpyAsync(n_unique, tot, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  VSV_TRY(vsv_out_done(ctx, u_chrom, oc, (size_t)*n_unique));
  VSV_TRY(vsv_out_done(ctx, u_pos, op, (size_t)*n_unique));
  VSV_TRY(vsv_out_done(ctx, u_len, ol, (size_t)*n_unique));
  VSV_TRY(vsv_out_done(ctx, u_cnt, on, (size_t)*n_unique));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// sig_support
// --------------------------------------------------------------------------------
#define VSV_MAX_CHROM 4096

__global__ void k_chrom_runs(const int32_t* __restrict__ sc, int64_t n, int32_t* __restrict__ seg_lo,
                             int32_t* __restrict__ seg_hi, int32_t* __restrict__ bad) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t c = sc[i];
  if (c < 0 || c >= VSV_MAX_CHROM) { *bad = 1; return; }
  if (i == 0 || sc[i - 1] != c) {
    if (atomicCAS(seg_lo + c, -1, (int32_t)i) != -1) *bad = 2;  // chromosome appears in two runs
  }
  if (i == n - 1 || sc[i + 1] != c) seg_hi[c] = (int32_t)(i + 1);
}

__global__ void k_single_chrom(int64_t n, int32_t* __restrict__ seg_lo, int32_t* __restrict__ seg_hi) {
  if (threadIdx.x == 0 && blockIdx.x == 0) { seg_lo[0] = 0; seg_hi[0] = (int32_t)n; }
}

#define SUP_GROUP 8  // lanes cooperating on one candidate
__global__ void k_support_window(const int32_t* __restrict__ cc, const int32_t* __restrict__ cp,
                                 const int32_t* __restrict__ cl, int64_t n_cand,
                                 const int32_t* __restrict__ sp, const int32_t* __restrict__ sl,
                                 const int32_t* __restrict__ sn, const int32_t* __restrict__ seg_lo,
                                 const int32_t* __restrict__ seg_hi, double min_size_sim, double max_shift_ratio,
                                 int32_t* __restrict__ lb_out, int32_t* __restrict__ ub_out,
                                 int32_t* __restrict__ valid_lb, int32_t* __restrict__ support) {
  const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / SUP_GROUP;
  const int g = threadIdx.x % SUP_GROUP;
  const unsigned gmask = 0xFFu << ((threadIdx.x & 31) / SUP_GROUP * SUP_GROUP);
  if (k >= n_cand) return;  // whole groups exit together (blockDim multiple of SUP_GROUP)
  const int32_t c = cc ? cc[k] : 0;
  const double sv = (double)cl[k];
  const double max_size = sv / min_size_sim;      // GTD:103
  const double min_size = sv * min_size_sim;      // GTD:104
  const double shift = fmax(sv * max_shift_ratio, 500.0);  // GTD:105
  const double lo_p = (double)cp[k] - shift, hi_p = (double)cp[k] + shift;
  int64_t a = 0, b = 0;
  if (c >= 0 && c < VSV_MAX_CHROM && seg_lo[c] >= 0) {
    a = lower_bound_f64(sp, seg_lo[c], seg_hi[c], lo_p);
    b = upper_bound_f64(sp, a, seg_hi[c], hi_p);
  }
  int32_t s = 0;
  for (int64_t i = a + g; i < b; i += SUP_GROUP) {
    double L = (double)__ldg(sl + i);
    if (min_size <= L && L <= max_size) s += __ldg(sn + i);
  }
#pragma unroll
  for (int d = SUP_GROUP / 2; d >= 1; d >>= 1) s += __shfl_xor_sync(gmask, s, d);
  if (g == 0) {
    lb_out[k] = (int32_t)a;
    ub_out[k] = (int32_t)b;
    valid_lb[k] = (b > a) ? (int32_t)a : -1;
    support[k] = s;
  }
}

__global__ void k_support_fixup(const int32_t* __restrict__ cc, const int32_t* __restrict__ cl, int64_t n_cand,
                                const int32_t* __restrict__ sc, const int32_t* __restrict__ sl,
                                const int32_t* __restrict__ sn, int64_t n_sig, const int32_t* __restrict__ lb,
                                const int32_t* __restrict__ ub, const int32_t* __restrict__ last_scan,
                                int32_t last_in, double min_size_sim, int32_t* __restrict__ support,
                                int32_t* __restrict__ match_id, int32_t* __restrict__ last_out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_cand) return;
  int32_t idx0 = last_in;
  if (k > 0 && last_scan[k - 1] >= 0) idx0 = last_scan[k - 1];
  if (match_id) match_id[k] = idx0;
  // both of the reference's scans visit last_match_idx itself (GTD:112-130): double count
  if (idx0 >= lb[k] && idx0 < ub[k] && idx0 < n_sig && (sc == nullptr || cc == nullptr || sc[idx0] == cc[k])) {
    const double sv = (double)cl[k];
    const double L = (double)sl[idx0];
    if (sv * min_size_sim <= L && L <= sv / min_size_sim) support[k] += sn[idx0];
  }
  if (k == n_cand - 1) *last_out = last_scan[k] >= 0 ? last_scan[k] : last_in;
}

extern "C" int vsv_sig_support(vsv_ctx* ctx, const int32_t* cand_chrom, const int32_t* cand_pos,
                               const int32_t* cand_svlen, int64_t n_cand, const int32_t* sig_chrom,
                               const int32_t* sig_pos, const int32_t* sig_len, const int32_t* sig_cnt,
                               int64_t n_sig, int32_t last_in, int32_t* support, int32_t* match_id,
                               int32_t* last_out) {
  VSV_TRY(vsv_begin(ctx));
  // last_out may be NULL for a single-chromosome call (no chromosome arrays): nothing has to come
  // back to the host then, so the call stays asynchronous (no stream synchronisation)
  if (n_cand < 0 || n_sig < 0 || n_sig >= (int64_t(1) << 31) || (!last_out && (cand_chrom || sig_chrom)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_sig_support: bad argument");
  if (last_out) *last_out = last_in;
  if (n_cand == 0) return VSV_OK;
  const int32_t *cc, *cp, *cl, *sc, *sp, *sl, *sn;
  VSV_TRY(vsv_in(ctx, cand_chrom, (size_t)n_cand, &cc));
  VSV_TRY(vsv_in(ctx, cand_pos, (size_t)n_cand, &cp));
  VSV_TRY(vsv_in(ctx, cand_svlen, (size_t)n_cand, &cl));
  VSV_TRY(vsv_in(ctx, sig_chrom, (size_t)n_sig, &sc));
  VSV_TRY(vsv_in(ctx, sig_pos, (size_t)n_sig, &sp));
  VSV_TRY(vsv_in(ctx, sig_len, (size_t)n_sig, &sl));
  VSV_TRY(vsv_in(ctx, sig_cnt, (size_t)n_sig, &sn));
  int32_t *d_sup, *d_mid;
  VSV_TRY(vsv_out(ctx, support, (size_t)n_cand, &d_sup));
  VSV_TRY(vsv_out(ctx, match_id, (size_t)n_cand, &d_mid));
  int32_t *seg, *lb, *ub, *vlb, *lscan, *d_last;
  VSV_TRY(vsv_tmp(ctx, 2 * VSV_MAX_CHROM + 2, &seg));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_cand, &lb));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_cand, &ub));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_cand, &vlb));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_cand, &lscan));
  int32_t* seg_lo = seg;
  int32_t* seg_hi = seg + VSV_MAX_CHROM;
  int32_t* bad = seg + 2 * VSV_MAX_CHROM;
  d_last = bad + 1;
  VSV_TRY(vsv_t0(ctx));
  VSV_CUDA(ctx, cudaMemsetAsync(seg, 0xFF, sizeof(int32_t) * 2 * VSV_MAX_CHROM, ctx->stream));
  VSV_CUDA(ctx, cudaMemsetAsync(bad, 0, 2 * sizeof(int32_t), ctx->stream));
  if ((cand_chrom == nullptr) != (sig_chrom == nullptr))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_sig_support: pass both chromosome arrays or neither");
  if (n_sig > 0 && sig_chrom == nullptr) VSV_LAUNCH(ctx, k_single_chrom, 1, 32, 0, n_sig, seg_lo, seg_hi);
  else if (n_sig > 0) VSV_LAUNCH(ctx, k_chrom_runs, (unsigned)vsv_cdiv(n_sig, 256), 256, 0, sc, n_sig, seg_lo, seg_hi, bad);
  VSV_LAUNCH(ctx, k_support_window, (unsigned)vsv_cdiv(n_cand * SUP_GROUP, 256), 256, 0, cc, cp, cl, n_cand, sp, sl, sn,
             seg_lo, seg_hi, 0.6, 2.3, lb, ub, vlb, d_sup);
  VSV_TRY(inclusive_scan_i32(ctx, vlb, lscan, n_cand, LastValid()));
  VSV_LAUNCH(ctx, k_support_fixup, (unsigned)vsv_cdiv(n_cand, 256), 256, 0, cc, cl, n_cand, sc, sl, sn, n_sig, lb, ub,
             lscan, last_in, 0.6, d_sup, d_mid, d_last);
  VSV_TRY(vsv_t1(ctx));
  if (last_out) {
    int32_t h[2] = {0, 0};
    VSV_CUDA(ctx, cudaMemcpyAsync(h, bad, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h[0] == 1) VSV_FAIL(ctx, VSV_E_ARG, "vsv_sig_support: chromosome id out of range [0,4096)");
    if (h[0] == 2) VSV_FAIL(ctx, VSV_E_ARG, "vsv_sig_support: signature table is not grouped by chromosome");
    *last_out = h[1];
  }
  VSV_TRY(vsv_out_done(ctx, support, d_sup, (size_t)n_cand));
  VSV_TRY(vsv_out_done(ctx, match_id, d_mid, (size_t)n_cand));
  return vsv_end(ctx);
}
