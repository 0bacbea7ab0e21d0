// Stage 1: C-ABI entry points of the CIGAR signature scan (kernels: vsv_scan_kernels.cuh).
#include "vsv_scan_kernels.cuh"

// ---- host side ---------------------------------------------------------------------------
struct ScanCall {
  int mode;
  int min_len_a, min_mapq_a, min_len_b, min_mapq_b;
  const uint32_t* ops; const int64_t* op_off; const int32_t* pos; const uint8_t* mapq;
  const uint8_t* enable_a; const uint8_t* enable_b;
  int64_t n_rec;
  vsv_sig* out_a; int64_t cap_a; int64_t* n_a;
  vsv_sig* out_b; int64_t cap_b; int64_t* n_b;
  int32_t *ref_total, *qry_total, *aux_total;
};

// CTA shape of the streaming kernel: warps are independent workers (no block-level sync), so the
// shape only sets how many fit an SM - 8 KB of staging per warp, 64 registers per thread -> 28.
static constexpr size_t SCAN_SMEM = (size_t)SCAN_WARPS * (SCAN_STAGES * (WT_OPS * sizeof(uint32_t) + sizeof(uint64_t)) + 8 * sizeof(uint32_t));

template <int MODE>
static int run_scan_kernels(vsv_ctx* ctx, const ScanParams& p_in, int grid, TileState* d_excl, vsv_sig* out_a, int64_t cap_a,
                            vsv_sig* out_b, int64_t cap_b, int64_t* d_nout, uint32_t* rt, uint32_t* qt, uint32_t* xt) {
  typedef ModeTraits<MODE> M;
  constexpr uint32_t RARE = M::X1 | M::X2 | M::X3;
  ScanParams p = p_in;
  p.tab[0] = tab_word(M::REF, 0, 0xFFu, 0, 0); p.tab[1] = tab_word(M::REF, 4, 0xFFu, 0, 0);
  p.tab[2] = tab_word(M::QRY, 0, 0xFFu, 0, 0); p.tab[3] = tab_word(M::QRY, 4, 0xFFu, 0, 0);
  p.tab[4] = tab_word(6u, 0, 0xFFu, RARE, 0x01u); p.tab[5] = tab_word(6u, 4, 0xFFu, RARE, 0x01u);
  p.tab[6] = tab_word(6u, 0, 0xFFu, RARE, 0x80u); p.tab[7] = tab_word(6u, 4, 0xFFu, RARE, 0x80u);
  VSV_CUDA(ctx, cudaEventRecord(ctx->evk0, ctx->stream));
  VSV_CUDA(ctx, cudaFuncSetAttribute(k_scan_wt<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCAN_SMEM));
  VSV_LAUNCH(ctx, k_scan_wt<MODE>, (unsigned)grid, SCAN_THREADS, SCAN_SMEM, p);
  VSV_CUDA(ctx, cudaEventRecord(ctx->evk1, ctx->stream));
  ctx->main_open = true;
  VSV_LAUNCH(ctx, k_gate<MODE>, (unsigned)std::min<int64_t>(vsv_cdiv((int64_t)p.stage_cap, 256), (int64_t)ctx->sm_count * 8), 256, 0, p);
  {  // segmented exclusive scan of the tile aggregates (n_wt + 1 elements, last = identity)
    const int64_t n_el = (int64_t)p.n_wt + 1, nb = vsv_cdiv(n_el, TS_BLOCK);
    TileState* d_blk = nullptr;
    VSV_TRY(vsv_tmp(ctx, (size_t)nb, &d_blk));
    VSV_LAUNCH(ctx, k_ts_reduce, (unsigned)nb, TS_THREADS, 0, (const TileState*)p.wt_agg, n_el, d_blk);
    VSV_LAUNCH(ctx, k_ts_spine, 1, 1024, 0, d_blk, nb);
    VSV_LAUNCH(ctx, k_ts_scan, (unsigned)nb, TS_THREADS, 0, (const TileState*)p.wt_agg, n_el, (const TileState*)d_blk, d_excl);
  }
  VSV_LAUNCH(ctx, k_emit<MODE>, (unsigned)std::min<int64_t>(vsv_cdiv((int64_t)p.stage_cap, 256), (int64_t)ctx->sm_count * 8), 256, 0,
             p, d_excl, out_a, cap_a, out_b, cap_b, d_nout);
  if (rt) VSV_LAUNCH(ctx, k_rec_totals<MODE>, (unsigned)vsv_cdiv(p.n_rec, 256), 256, 0, p, d_excl, rt, qt, xt);
  return VSV_OK;
}

static int scan_impl(vsv_ctx* ctx, const ScanCall& c) {
  const char* fn = c.mode == VSV_MODE_FUSED ? "vsv_scan_reads_fused" : "vsv_scan_cigar";
  const bool fused = c.mode == VSV_MODE_FUSED;
  if (c.n_rec < 0 || c.cap_a < 0 || !c.n_a || (fused && (c.cap_b < 0 || !c.n_b)) ||
      (c.n_rec > 0 && (!c.op_off || !c.pos || !c.mapq)))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": bad argument");
  if (c.min_len_a < 0 || c.min_len_a >= (1 << 27) || c.min_len_b < 0 || c.min_len_b >= (1 << 27))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": min_len out of range");
  if (c.n_rec >= (int64_t(1) << 31) - 1) VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": more than 2^31-2 records per call");
  if (c.cap_a >= (int64_t(1) << 31) || c.cap_b >= (int64_t(1) << 31))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": capacity must be below 2^31 signatures per call");
  const bool want_tot = c.ref_total || c.qry_total || c.aux_total;
  if (want_tot && !(c.ref_total && c.qry_total && c.aux_total))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": pass all three totals or none");
  *c.n_a = 0;
  if (fused) *c.n_b = 0;
  if (c.n_rec == 0) return VSV_OK;

  // number of ops = op_off[n_rec] (host knows it in host-IO mode; fetch it in device-IO mode)
  int64_t n_ops = 0;
  if (ctx->device_io) {
    VSV_CUDA(ctx, cudaMemcpyAsync(&n_ops, c.op_off + c.n_rec, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    n_ops = c.op_off[c.n_rec];
  }
  if (n_ops < 0) VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": negative op count");
  const int64_t n_wt64 = vsv_cdiv(n_ops, WT_OPS);
  if (n_wt64 >= (int64_t(1) << 29)) VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": too many ops per call");  // tile index shares a word with 3 flag bits
  const int32_t n_wt = (int32_t)n_wt64;

  ScanParams p;
  memset(&p, 0, sizeof(p));
  const uint8_t *d_mapq = nullptr, *d_ena = nullptr, *d_enb = nullptr;
  VSV_TRY(vsv_in(ctx, c.ops, (size_t)n_ops, &p.ops));
  VSV_TRY(vsv_in(ctx, c.op_off, (size_t)c.n_rec + 1, &p.op_off));
  VSV_TRY(vsv_in(ctx, c.pos, (size_t)c.n_rec, &p.pos));
  VSV_TRY(vsv_in(ctx, c.mapq, (size_t)c.n_rec, &d_mapq));
  VSV_TRY(vsv_in(ctx, c.enable_a, (size_t)c.n_rec, &d_ena));
  VSV_TRY(vsv_in(ctx, c.enable_b, (size_t)c.n_rec, &d_enb));
  if (n_ops > 0 && (reinterpret_cast<uintptr_t>(p.ops) & 15u))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": ops must be 16-byte aligned");
  vsv_sig *d_out_a = nullptr, *d_out_b = nullptr;
  VSV_TRY(vsv_out(ctx, c.out_a, (size_t)c.cap_a, &d_out_a));
  if (fused) VSV_TRY(vsv_out(ctx, c.out_b, (size_t)c.cap_b, &d_out_b));
  if ((c.cap_a > 0 && (reinterpret_cast<uintptr_t>(d_out_a) & 15u)) ||
      (fused && c.cap_b > 0 && (reinterpret_cast<uintptr_t>(d_out_b) & 15u)))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": out must be 16-byte aligned");
  int32_t *d_rt = nullptr, *d_qt = nullptr, *d_xt = nullptr;
  if (want_tot) {
    VSV_TRY(vsv_out(ctx, c.ref_total, (size_t)c.n_rec, &d_rt));
    VSV_TRY(vsv_out(ctx, c.qry_total, (size_t)c.n_rec, &d_qt));
    VSV_TRY(vsv_out(ctx, c.aux_total, (size_t)c.n_rec, &d_xt));
  }
  const int grid = vsv_scan_grid(ctx, n_wt);
  // staging holds one entry per RAW candidate (before the record gates), reserved per warp in
  // chunks, so the exact need is only known after a run: start from the caller's capacities and
  // retry (rare) with the measured reservation if a warp found the staging full.
  const int64_t slack = (int64_t)grid * SCAN_WARPS * SCAN_CHUNK + SCAN_CHUNK;
  int64_t scap = (fused ? std::max(c.cap_a, c.cap_b) : c.cap_a) * 5 / 4 + slack;
  int64_t* d_nout = nullptr;  // [0..1] signature counts, [2] staging counter, [3] unused, [4] staging-full flag
  uint32_t* d_info = nullptr;
  int2* d_starts = nullptr;
  TileState *d_agg = nullptr, *d_excl = nullptr;
  uint2* d_wstage = nullptr;
  uint4* d_fstart = nullptr;
  VSV_TRY(vsv_tmp(ctx, 5, &d_nout));
  VSV_TRY(vsv_tmp(ctx, (size_t)c.n_rec, &d_info));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_wt + 1, &d_starts));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_wt + 1, &d_agg));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_wt + 1, &d_excl));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_wt + 1, &d_wstage));
  VSV_TRY(vsv_tmp(ctx, 2 * ((size_t)c.n_rec + 1), &d_fstart));

  p.rec_info = d_info;
  p.n_rec = c.n_rec; p.n_ops = n_ops; p.n_wt = n_wt;
  p.min_len_a = c.min_len_a; p.min_len_b = c.min_len_b;
  p.one = 1u;
  p.min_w = (uint32_t)(fused ? std::min(c.min_len_a, c.min_len_b) : c.min_len_a) << 4;
  p.wt_starts = d_starts; p.wt_agg = d_agg; p.wt_stage = d_wstage; p.fstart = d_fstart;
  p.stage_ctr = reinterpret_cast<unsigned long long*>(d_nout + 2);

  int64_t h_n[5] = {0, 0, 0, 0, 0};
  VSV_TRY(vsv_t0(ctx));
  if (n_wt > 0) {
    VSV_LAUNCH(ctx, k_rec_info, (unsigned)vsv_cdiv(c.n_rec, 256), 256, 0, p.ops, p.op_off, d_mapq, d_ena, d_enb, c.n_rec,
               c.min_mapq_a, c.min_mapq_b, d_info);
    int32_t* d_first = nullptr;
    VSV_TRY(vsv_tmp(ctx, (size_t)n_wt + 1, &d_first));
    VSV_LAUNCH(ctx, k_wt_first, (unsigned)vsv_cdiv(c.n_rec + 1, 256), 256, 0, p.op_off, c.n_rec, n_wt, d_first);
    VSV_LAUNCH(ctx, k_wt_starts, (unsigned)vsv_cdiv(n_wt, 256), 256, 0, (const int32_t*)d_first, n_wt, d_starts);
    VSV_CUDA(ctx, cudaMemsetAsync(d_agg + n_wt, 0, sizeof(TileState), ctx->stream));  // identity sentinel
  }
  for (int attempt = 0; attempt < 3; ++attempt) {
    if (scap >= (int64_t(1) << 32) - 1)
      VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": too many candidate signatures for one call; split the batch");
    StageEntry* d_stage = nullptr;
    VSV_TRY(vsv_tmp(ctx, (size_t)scap, &d_stage));
    p.stage = d_stage;
    p.stage_cap = (uint32_t)scap;
    VSV_CUDA(ctx, cudaMemsetAsync(d_nout, 0, 5 * sizeof(int64_t), ctx->stream));
    if (n_wt == 0) {
      if (want_tot) {
        VSV_CUDA(ctx, cudaMemsetAsync(d_rt, 0, sizeof(int32_t) * (size_t)c.n_rec, ctx->stream));
        VSV_CUDA(ctx, cudaMemsetAsync(d_qt, 0, sizeof(int32_t) * (size_t)c.n_rec, ctx->stream));
        VSV_CUDA(ctx, cudaMemsetAsync(d_xt, 0, sizeof(int32_t) * (size_t)c.n_rec, ctx->stream));
      }
      break;
    }
    uint32_t *rt = (uint32_t*)d_rt, *qt = (uint32_t*)d_qt, *xt = (uint32_t*)d_xt;
    switch (c.mode) {
      case VSV_MODE_CONTIG:
        VSV_TRY(run_scan_kernels<VSV_MODE_CONTIG>(ctx, p, grid, d_excl, d_out_a, c.cap_a, d_out_b, c.cap_b, d_nout, rt, qt, xt));
        break;
      case VSV_MODE_READS:
        VSV_TRY(run_scan_kernels<VSV_MODE_READS>(ctx, p, grid, d_excl, d_out_a, c.cap_a, d_out_b, c.cap_b, d_nout, rt, qt, xt));
        break;
      case VSV_MODE_CUTE:
        VSV_TRY(run_scan_kernels<VSV_MODE_CUTE>(ctx, p, grid, d_excl, d_out_a, c.cap_a, d_out_b, c.cap_b, d_nout, rt, qt, xt));
        break;
      default:
        VSV_TRY(run_scan_kernels<VSV_MODE_FUSED>(ctx, p, grid, d_excl, d_out_a, c.cap_a, d_out_b, c.cap_b, d_nout, rt, qt, xt));
        break;
    }
    // the counts are needed on the host to size the copy-out: one small sync
    VSV_CUDA(ctx, cudaMemcpyAsync(h_n, d_nout, 5 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h_n[4] == 0) break;           // staging held every candidate
    if (attempt == 2) VSV_FAIL(ctx, VSV_E_NOMEM, std::string(fn) + ": staging kept overflowing");
    scap = 2 * h_n[2] + slack;        // reservation of the failed run bounds the raw count
  }
  VSV_TRY(vsv_t1(ctx));
  *c.n_a = h_n[0];
  if (fused) *c.n_b = h_n[1];
  if (want_tot) {
    VSV_TRY(vsv_out_done(ctx, c.ref_total, d_rt, (size_t)c.n_rec));
    VSV_TRY(vsv_out_done(ctx, c.qry_total, d_qt, (size_t)c.n_rec));
    VSV_TRY(vsv_out_done(ctx, c.aux_total, d_xt, (size_t)c.n_rec));
  }
  if (h_n[0] > c.cap_a || (fused && h_n[1] > c.cap_b)) {
    vsv_end(ctx, true);
    VSV_FAIL(ctx, VSV_E_CAPACITY, std::string(fn) + ": output capacity too small");
  }
  VSV_TRY(vsv_out_done(ctx, c.out_a, d_out_a, (size_t)h_n[0]));
  if (fused) VSV_TRY(vsv_out_done(ctx, c.out_b, d_out_b, (size_t)h_n[1]));
  return vsv_end(ctx);
}

extern "C" int vsv_scan_cigar(vsv_ctx* ctx, int mode, int min_len, int min_mapq, const uint32_t* ops,
                              const int64_t* op_off, const int32_t* pos, const uint8_t* mapq,
                              const uint8_t* enable, int64_t n_rec, vsv_sig* out, int64_t cap,
                              int64_t* n_out, int32_t* ref_total, int32_t* qry_total,
                              int32_t* aux_total) {
  VSV_TRY(vsv_begin(ctx));
  if (mode < 0 || mode > 2) VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: bad mode");
  ScanCall c;
  memset(&c, 0, sizeof(c));
  c.mode = mode; c.min_len_a = min_len; c.min_mapq_a = min_mapq;
  c.ops = ops; c.op_off = op_off; c.pos = pos; c.mapq = mapq; c.enable_a = enable; c.n_rec = n_rec;
  c.out_a = out; c.cap_a = cap; c.n_a = n_out;
  c.ref_total = ref_total; c.qry_total = qry_total; c.aux_total = aux_total;
  return scan_impl(ctx, c);
}

extern "C" int vsv_scan_reads_fused(vsv_ctx* ctx, int min_len_reads, int min_mapq_reads, int min_len_cute,
                                    int min_mapq_cute, const uint32_t* ops, const int64_t* op_off,
                                    const int32_t* pos, const uint8_t* mapq, const uint8_t* enable_cute,
                                    int64_t n_rec, vsv_sig* out_reads, int64_t cap_reads, int64_t* n_reads,
                                    vsv_sig* out_cute, int64_t cap_cute, int64_t* n_cute, int32_t* ref_total,
                                    int32_t* qry_total, int32_t* aux_total) {
  VSV_TRY(vsv_begin(ctx));
  ScanCall c;
  memset(&c, 0, sizeof(c));
  c.mode = VSV_MODE_FUSED;
  c.min_len_a = min_len_reads; c.min_mapq_a = min_mapq_reads; c.min_len_b = min_len_cute; c.min_mapq_b = min_mapq_cute;
  c.ops = ops; c.op_off = op_off; c.pos = pos; c.mapq = mapq; c.enable_b = enable_cute; c.n_rec = n_rec;
  c.out_a = out_reads; c.cap_a = cap_reads; c.n_a = n_reads;
  c.out_b = out_cute; c.cap_b = cap_cute; c.n_b = n_cute;
  c.ref_total = ref_total; c.qry_total = qry_total; c.aux_total = aux_total;
  return scan_impl(ctx, c);
}
