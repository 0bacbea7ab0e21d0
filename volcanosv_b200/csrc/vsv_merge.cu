// K2: per-record merge (split out of vsv_scan.cu).
#include "vsv_common.cuh"

// ================================================================================
// K2: per-record merge.  cluster_ins_one_read / cluster_del_one_read (ECS_H:87-164),
// generate_combine_sigs (SE:373-435).  Signatures are sparse (<= a few per record), so
// this is one thread per record run walking its (already ordered) signatures; two
// passes (count, exclusive scan, write) keep the output order deterministic.
// ================================================================================
__global__ void k_type_flags(const vsv_sig* __restrict__ s, int64_t n, int32_t* __restrict__ is_del) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) is_del[i] = s[i].type == VSV_DEL ? 1 : 0;
}

__global__ void k_type_split(const vsv_sig* __restrict__ s, int64_t n, const int64_t* __restrict__ del_excl,
                             vsv_sig* __restrict__ dels, vsv_sig* __restrict__ inss) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  vsv_sig v = s[i];
  int64_t d = del_excl[i];
  if (v.type == VSV_DEL) dels[d] = v; else inss[i - d] = v;
}

// Walk one record's run [i0,i1) of same-type signatures; calls f(merged) per output.
template <typename F>
__device__ __forceinline__ void merge_walk(const vsv_sig* __restrict__ s, int64_t i0, int64_t i1, int mode,
                                           int type, F f) {
  vsv_sig cur = s[i0];
  cur.first = (int32_t)i0;
  cur.count = 1;
  if (mode == VSV_MODE_READS) {
    f(cur);
    for (int64_t k = i0 + 1; k < i1; ++k) {
      cur = s[k]; cur.first = (int32_t)k; cur.count = 1;
      f(cur);
    }
    return;
  }
  int32_t mark = (type == VSV_INS) ? cur.pos : cur.pos + cur.len;  // cute: last INS pos / DEL end
  for (int64_t k = i0 + 1; k < i1; ++k) {
    const vsv_sig n = s[k];
    bool hit;
    if (mode == VSV_MODE_CONTIG) {
      int32_t d = abs(n.pos - cur.pos);
      if (type == VSV_INS) {
        hit = (cur.len > 250 && n.len > 250 && d < 250) || (cur.len > 320 && n.len > 320 && d < 380) ||
              (cur.len > 100 && n.len > 100 && d < 250);           // ECS_H:112-135
        if (hit) { cur.q1 = n.q1; cur.len = cur.q1 - cur.q0; }     // ECS_H:91-97
      } else {
        hit = cur.len > 150 && n.len > 150 && d < 150;             // ECS_H:148-150
        if (hit) { cur.len = n.pos + n.len - cur.pos; cur.q1 = cur.q0 + 1; }  // ECS_H:99-106
      }
    } else {  // VSV_MODE_CUTE
      if (type == VSV_INS) {
        hit = n.pos - mark <= 100;                                 // SE:396-399
        if (hit) { cur.len += n.len; mark = n.pos; }
      } else {
        hit = n.pos - mark <= 0;                                   // SE:420
        if (hit) { cur.len += n.len; mark = n.pos + n.len; }
      }
    }
    if (hit) {
      cur.count++;
    } else {
      f(cur);
      cur = n; cur.first = (int32_t)k; cur.count = 1;
      mark = n.pos;                                                // SE:412-413 / SE:429-430 (end := start quirk)
    }
  }
  f(cur);
}

__global__ void k_merge_count(const vsv_sig* __restrict__ s, int64_t n, int mode, int type,
                              int32_t* __restrict__ cnt) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t c = 0;
  if (i == 0 || s[i - 1].rec != s[i].rec) {
    int64_t e = i + 1;
    const int32_t r = s[i].rec;
    while (e < n && s[e].rec == r) ++e;
    merge_walk(s, i, e, mode, type, [&](const vsv_sig&) { ++c; });
  }
  cnt[i] = c;
}

__global__ void k_merge_write(const vsv_sig* __restrict__ s, int64_t n, int mode, int type,
                              const int64_t* __restrict__ excl, vsv_sig* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (i == 0 || s[i - 1].rec != s[i].rec) {
    int64_t e = i + 1;
    const int32_t r = s[i].rec;
    while (e < n && s[e].rec == r) ++e;
    int64_t o = excl[i];
    merge_walk(s, i, e, mode, type, [&](const vsv_sig& v) { out[o++] = v; });
  }
}

extern "C" int vsv_merge_in_record(vsv_ctx* ctx, int mode, const vsv_sig* sigs, int64_t n, vsv_sig* out_del,
                                   int64_t* n_del, vsv_sig* out_ins, int64_t* n_ins) {
  VSV_TRY(vsv_begin(ctx));
  if (mode < 0 || mode > 2 || n < 0 || !n_del || !n_ins || n >= (int64_t(1) << 31))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_merge_in_record: bad argument");
  *n_del = *n_ins = 0;
  if (n == 0) return VSV_OK;
  const vsv_sig* d_in = nullptr;
  VSV_TRY(vsv_in(ctx, sigs, (size_t)n, &d_in));
  vsv_sig *d_od = nullptr, *d_oi = nullptr, *t_del = nullptr, *t_ins = nullptr;
  VSV_TRY(vsv_out(ctx, out_del, (size_t)n, &d_od));
  VSV_TRY(vsv_out(ctx, out_ins, (size_t)n, &d_oi));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &t_del));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &t_ins));
  int32_t* flags = nullptr;
  int64_t *excl = nullptr, *tot = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &flags));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &excl));
  VSV_TRY(vsv_tmp(ctx, 3, &tot));
  const unsigned g = (unsigned)vsv_cdiv(n, 256);
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_type_flags, g, 256, 0, d_in, n, flags);
  VSV_TRY(vsv_exclusive_scan_i32(ctx, flags, n, excl, tot));
  VSV_LAUNCH(ctx, k_type_split, g, 256, 0, d_in, n, excl, t_del, t_ins);
  int64_t h_tot[3] = {0, 0, 0};
  VSV_CUDA(ctx, cudaMemcpyAsync(h_tot, tot, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const int64_t nd = h_tot[0], ni = n - h_tot[0];
  for (int type = 0; type < 2; ++type) {
    const int64_t m = type == VSV_DEL ? nd : ni;
    if (m == 0) continue;
    const vsv_sig* src = type == VSV_DEL ? t_del : t_ins;
    vsv_sig* dst = type == VSV_DEL ? d_od : d_oi;
    const unsigned gm = (unsigned)vsv_cdiv(m, 128);
    VSV_LAUNCH(ctx, k_merge_count, gm, 128, 0, src, m, mode, type, flags);
    VSV_TRY(vsv_exclusive_scan_i32(ctx, flags, m, excl, tot + 1 + type));
    VSV_LAUNCH(ctx, k_merge_write, gm, 128, 0, src, m, mode, type, excl, dst);
  }
  VSV_TRY(vsv_t1(ctx));
  VSV_CUDA(ctx, cudaMemcpyAsync(h_tot + 1, tot + 1, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n_del = nd ? h_tot[1] : 0;
  *n_ins = ni ? h_tot[2] : 0;
  VSV_TRY(vsv_out_done(ctx, out_del, d_od, (size_t)*n_del));
  VSV_TRY(vsv_out_done(ctx, out_ins, d_oi, (size_t)*n_ins));
  return vsv_end(ctx);
}
