// K3b: SA-tag split analysis (analysis_split_read, LI/sig_extract.py:193-326) on the
// segment lists organize_split_signal (SE:345-371) builds from a primary record and its SA
// entries.  One thread per read: the list (<= max_parts <= 16 segments) is sorted by query
// start in local memory (stable insertion sort == Python's sorted), then the reference's
// two-segment rule, sliding triples with the tail pair, and the INS-in-TRA rescue run in
// order.  Rows go to a per-read slot of a staging array (a read with k segments emits at
// most 2(k-1)+1 rows) and are compacted in (read, emission) order: count -> exclusive scan
// -> copy.  TRA rows are not produced (SE:637-638 drops them with `grep -w`), but the
// different-chromosome trigger of the rescue rule is kept.
#include "vsv_rules.cuh"

__global__ void __launch_bounds__(128) k_sa_analysis(SaIn in, int64_t n_read, int sv_size, int max_size, int max_parts,
                                                     vsv_sig* __restrict__ tmp, int32_t* __restrict__ cnt) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_read) return;
  const int64_t a = in.seg_off[r], b = in.seg_off[r + 1];
  const int k = (int)(b - a);
  SaEmit em;
  em.dst = tmp + 2 * a; em.rec = (int32_t)r; em.n = 0; em.rev = 0; em.sv_size = sv_size; em.max_size = max_size;
  if (k >= 2 && k <= max_parts) {
    SaSeg sp[SA_MAXSEG];
    for (int i = 0; i < k; ++i) {
      SaSeg s;
      s.qs = in.q_s[a + i]; s.qe = in.q_e[a + i]; s.rs = in.r_s[a + i]; s.re = in.r_e[a + i];
      s.chr = in.chrom[a + i]; s.st = in.strand[a + i] ? 1 : 0;
      sa_insert(sp, i, s);
    }
    sa_analyze(sp, k, in.read_len[r], em);
  }
  cnt[r] = em.n;
}

__global__ void k_sa_compact(const int64_t* __restrict__ seg_off, const int32_t* __restrict__ cnt,
                             const int64_t* __restrict__ excl, const vsv_sig* __restrict__ tmp, int64_t n_read,
                             vsv_sig* __restrict__ out, int64_t cap) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_read) return;
  const int32_t c = cnt[r];
  const int64_t o = excl[r], s = 2 * seg_off[r];
  for (int32_t i = 0; i < c; ++i)
    if (o + i < cap) out[o + i] = tmp[s + i];
}

extern "C" int vsv_sa_analysis(vsv_ctx* ctx, int sv_size, int max_size, int max_parts, const int64_t* seg_off,
                               const int32_t* q_s, const int32_t* q_e, const int32_t* r_s, const int32_t* r_e,
                               const int32_t* chrom, const uint8_t* strand, const int32_t* read_len, int64_t n_read,
                               int64_t n_seg, vsv_sig* out, int64_t cap, int64_t* n_out) {
  VSV_TRY(vsv_begin(ctx));
  if (n_read < 0 || n_seg < 0 || cap < 0 || !n_out || (n_read > 0 && (!seg_off || !read_len)) ||
      (n_seg > 0 && (!q_s || !q_e || !r_s || !r_e || !chrom || !strand)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_sa_analysis: bad argument");
  if (max_parts < 2 || max_parts > SA_MAXSEG)
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_sa_analysis: max_parts must be in [2,16] (reference: 7, SE:370)");
  *n_out = 0;
  if (n_read == 0) return VSV_OK;
  SaIn in;
  VSV_TRY(vsv_in(ctx, seg_off, (size_t)n_read + 1, &in.seg_off));
  VSV_TRY(vsv_in(ctx, q_s, (size_t)n_seg, &in.q_s));
  VSV_TRY(vsv_in(ctx, q_e, (size_t)n_seg, &in.q_e));
  VSV_TRY(vsv_in(ctx, r_s, (size_t)n_seg, &in.r_s));
  VSV_TRY(vsv_in(ctx, r_e, (size_t)n_seg, &in.r_e));
  VSV_TRY(vsv_in(ctx, chrom, (size_t)n_seg, &in.chrom));
  VSV_TRY(vsv_in(ctx, strand, (size_t)n_seg, &in.strand));
  VSV_TRY(vsv_in(ctx, read_len, (size_t)n_read, &in.read_len));
  vsv_sig *d_out, *tmp;
  int32_t* cnt;
  int64_t *excl, *tot;
  VSV_TRY(vsv_out(ctx, out, (size_t)cap, &d_out));
  VSV_TRY(vsv_tmp(ctx, 2 * (size_t)n_seg + 2, &tmp));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_read, &cnt));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_read, &excl));
  VSV_TRY(vsv_tmp(ctx, 1, &tot));
  const unsigned g = (unsigned)vsv_cdiv(n_read, 128);
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_sa_analysis, g, 128, 0, in, n_read, sv_size, max_size, max_parts, tmp, cnt);
  VSV_TRY(vsv_exclusive_scan_i32(ctx, cnt, n_read, excl, tot));
  VSV_LAUNCH(ctx, k_sa_compact, g, 128, 0, in.seg_off, cnt, excl, tmp, n_read, d_out, cap);
  VSV_TRY(vsv_t1(ctx));
  int64_t h_n = 0;
  VSV_CUDA(ctx, cudaMemcpyAsync(&h_n, tot, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n_out = h_n;
  if (h_n > cap) {
    vsv_end(ctx, true);
    VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_sa_analysis: output capacity too small");
  }
  VSV_TRY(vsv_out_done(ctx, out, d_out, (size_t)h_n));
  return vsv_end(ctx);
}
