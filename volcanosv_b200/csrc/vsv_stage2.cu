// Stage 2: sort, split-pair signatures, leader clustering, haplotype pairing and the
// window joins (FP filter, coverage support).  All inputs here are the SPARSE outputs
// of stage 1 (signatures / calls), sorted by position, so every kernel is a sorted-
// array window join: binary search for the window, then a short scan.  The reference's
// sequential-greedy semantics (leader clustering ECS_H:196-288, hp pairing ECS_H:548-569)
// are kept exact by cutting the sorted list into independent segments (no interaction
// across a gap larger than the rule's max shift) and walking each segment in order.
#include <cub/device/device_radix_sort.cuh>

#include "vsv_rules.cuh"
#include "vsv_search.cuh"

// --------------------------------------------------------------------------------
// stable argsort (sort_sig, ECS_H:170-179, tie order pinned to stable)
// --------------------------------------------------------------------------------
__global__ void k_iota(int32_t* a, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = (int32_t)i;
}

static int argsort_dev(vsv_ctx* ctx, const int32_t* key, int64_t n, int32_t* order, int32_t* key_sorted) {
  int32_t* idx = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &idx));
  int32_t* ks = key_sorted;
  if (!ks) VSV_TRY(vsv_tmp(ctx, (size_t)n, &ks));
  VSV_LAUNCH(ctx, k_iota, (unsigned)vsv_cdiv(n, 256), 256, 0, idx, n);
  size_t tb = 0;
  VSV_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, ks, idx, order, (int)n, 0, 32, ctx->stream));
  void* tmp = nullptr;
  VSV_TRY(vsv_arena_alloc(ctx, tb, &tmp));
  VSV_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, ks, idx, order, (int)n, 0, 32, ctx->stream));
  ctx->launches += 3;
  return VSV_OK;
}

extern "C" int vsv_argsort_stable(vsv_ctx* ctx, const int32_t* key, int64_t n, int32_t* order) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0 || n >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_argsort_stable: bad n");
  if (n == 0) return VSV_OK;
  const int32_t* dk;
  int32_t* dord;
  VSV_TRY(vsv_in(ctx, key, (size_t)n, &dk));
  VSV_TRY(vsv_out(ctx, order, (size_t)n, &dord));
  VSV_TRY(vsv_t0(ctx));
  VSV_TRY(argsort_dev(ctx, dk, n, dord, nullptr));
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, order, dord, (size_t)n));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// split pairs
// --------------------------------------------------------------------------------
__global__ void k_select_flags(const uint8_t* __restrict__ sel, int64_t n, int32_t* __restrict__ f) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) f[i] = sel[i] ? 1 : 0;
}
__global__ void k_select_gather(const uint8_t* __restrict__ sel, const int64_t* __restrict__ excl,
                                const int32_t* __restrict__ name, int64_t n, int32_t* __restrict__ rec,
                                int32_t* __restrict__ key) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && sel[i]) { rec[excl[i]] = (int32_t)i; key[excl[i]] = name[i]; }
}

struct SplitIn {
  const int32_t *pos, *ref_end, *readlen, *name;
  const uint32_t *first_op, *last_op;
  const uint8_t *mapq, *is_rev;
};

__global__ void k_split_eval(SplitIn in, const int32_t* __restrict__ key_sorted, const int32_t* __restrict__ ord,
                             const int32_t* __restrict__ rec, int64_t m, int variant, int min_mapq,
                             int32_t* __restrict__ flag, vsv_sig* __restrict__ tmp) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= m) return;
  int32_t f = 0;
  if (k + 1 < m && key_sorted[k] == key_sorted[k + 1]) {
    const int32_t a = rec[ord[k]], b = rec[ord[k + 1]];
    vsv_sig o;
    o.rec = a; o.first = b; o.count = 1;
    if (split_rule(variant, min_mapq, in.ref_end[a], in.readlen[a], in.last_op[a], in.mapq[a], in.is_rev[a] != 0,
                   in.pos[b], in.first_op[b], in.mapq[b], in.is_rev[b] != 0, o)) {
      tmp[k] = o;
      f = 1;
    }
  }
  flag[k] = f;
}
__global__ void k_compact_sigs(const int32_t* __restrict__ flag, const int64_t* __restrict__ excl,
                               const vsv_sig* __restrict__ tmp, int64_t m, vsv_sig* __restrict__ out, int64_t cap) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < m && flag[k] && excl[k] < cap) out[excl[k]] = tmp[k];
}

extern "C" int vsv_split_pairs(vsv_ctx* ctx, int variant, int min_mapq, const int32_t* pos, const int32_t* ref_end,
                               const int32_t* readlen, const uint32_t* first_op, const uint32_t* last_op,
                               const uint8_t* mapq, const uint8_t* is_rev, const int32_t* name_id,
                               const uint8_t* select, int64_t n_rec, vsv_sig* out, int64_t cap, int64_t* n_out) {
  VSV_TRY(vsv_begin(ctx));
  if (variant < 0 || variant > 3 || n_rec < 0 || n_rec >= (int64_t(1) << 31) || !n_out || cap < 0)
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_split_pairs: bad argument");
  *n_out = 0;
  if (n_rec < 2) return VSV_OK;
  SplitIn in;
  const uint8_t* sel;
  VSV_TRY(vsv_in(ctx, pos, (size_t)n_rec, &in.pos));
  VSV_TRY(vsv_in(ctx, ref_end, (size_t)n_rec, &in.ref_end));
  VSV_TRY(vsv_in(ctx, readlen, (size_t)n_rec, &in.readlen));
  VSV_TRY(vsv_in(ctx, first_op, (size_t)n_rec, &in.first_op));
  VSV_TRY(vsv_in(ctx, last_op, (size_t)n_rec, &in.last_op));
  VSV_TRY(vsv_in(ctx, mapq, (size_t)n_rec, &in.mapq));
  VSV_TRY(vsv_in(ctx, is_rev, (size_t)n_rec, &in.is_rev));
  VSV_TRY(vsv_in(ctx, name_id, (size_t)n_rec, &in.name));
  VSV_TRY(vsv_in(ctx, select, (size_t)n_rec, &sel));
  vsv_sig *d_out, *tmp;
  VSV_TRY(vsv_out(ctx, out, (size_t)cap, &d_out));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &tmp));
  int32_t *flag, *rec, *key, *key_sorted, *ord;
  int64_t *excl, *tot;
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &flag));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &rec));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &key));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &key_sorted));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &ord));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_rec, &excl));
  VSV_TRY(vsv_tmp(ctx, 2, &tot));
  const unsigned g = (unsigned)vsv_cdiv(n_rec, 256);
  VSV_TRY(vsv_t0(ctx));
  if (sel) {
    VSV_LAUNCH(ctx, k_select_flags, g, 256, 0, sel, n_rec, flag);
    VSV_TRY(vsv_exclusive_scan_i32(ctx, flag, n_rec, excl, tot));
    VSV_LAUNCH(ctx, k_select_gather, g, 256, 0, sel, excl, in.name, n_rec, rec, key);
  }
  int64_t m = n_rec;
  if (sel) {
    VSV_CUDA(ctx, cudaMemcpyAsync(&m, tot, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    VSV_LAUNCH(ctx, k_iota, g, 256, 0, rec, n_rec);
    VSV_CUDA(ctx, cudaMemcpyAsync(key, in.name, sizeof(int32_t) * (size_t)n_rec, cudaMemcpyDeviceToDevice, ctx->stream));
  }
  int64_t h_n = 0;
  if (m >= 2) {
    VSV_TRY(argsort_dev(ctx, key, m, ord, key_sorted));
    const unsigned gm = (unsigned)vsv_cdiv(m, 256);
    VSV_LAUNCH(ctx, k_split_eval, gm, 256, 0, in, key_sorted, ord, rec, m, variant, min_mapq, flag, tmp);
    VSV_TRY(vsv_exclusive_scan_i32(ctx, flag, m, excl, tot + 1));
    VSV_LAUNCH(ctx, k_compact_sigs, gm, 256, 0, flag, excl, tmp, m, d_out, cap);
    VSV_TRY(vsv_t1(ctx));
    VSV_CUDA(ctx, cudaMemcpyAsync(&h_n, tot + 1, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    VSV_TRY(vsv_t1(ctx));
  }
  *n_out = h_n;
  if (h_n > cap) {
    vsv_end(ctx, true);
    VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_split_pairs: output capacity too small");
  }
  VSV_TRY(vsv_out_done(ctx, out, d_out, (size_t)h_n));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// leader clustering (cluster_del / cluster_ins, ECS_H:196-288)
// --------------------------------------------------------------------------------
__device__ __forceinline__ bool size_sim_half(int32_t a, int32_t b) {  // min/max >= 0.5
  const int64_t mn = min(a, b), mx = max(a, b);
  return 2 * mn >= mx;
}
__device__ __forceinline__ bool overlap_half(int32_t p1, int32_t l1, int32_t p2, int32_t l2) {
  // (min(e1,e2)-max(s1,s2))/minlen >= 0.5  (ECS_H:200-204)
  const int64_t ovl = min((int64_t)p1 + l1, (int64_t)p2 + l2) - max((int64_t)p1, (int64_t)p2);
  return 2 * ovl >= (int64_t)min(l1, l2);
}

// One thread per segment head (consecutive gaps <= max_shift chain a segment).
__global__ void k_cluster(const int32_t* __restrict__ pos, const int32_t* __restrict__ len, int64_t n, int is_del,
                          int max_shift, int32_t* __restrict__ cl, int32_t* __restrict__ best_of,
                          int32_t* __restrict__ is_leader) {
  int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i0 >= n) return;
  if (i0 != 0 && (int64_t)pos[i0] - (int64_t)pos[i0 - 1] <= max_shift) return;  // not a segment head
  int64_t e = i0 + 1;
  while (e < n && (int64_t)pos[e] - (int64_t)pos[e - 1] <= max_shift) ++e;
  for (int64_t i = i0; i < e; ++i) {
    if (cl[i] != -1) { is_leader[i] = 0; continue; }
    cl[i] = (int32_t)i;
    int32_t best = (int32_t)i;
    const int32_t pi = pos[i], li = len[i];
    for (int64_t j = i + 1; j < e && (int64_t)pos[j] - pi <= max_shift; ++j) {
      if (cl[j] != -1) continue;
      const int32_t lj = len[j];
      if (!size_sim_half(li, lj)) continue;
      if (is_del && !overlap_half(pi, li, pos[j], lj)) continue;
      cl[j] = (int32_t)i;
      if (lj > len[best]) best = (int32_t)j;
    }
    best_of[i] = best;
    is_leader[i] = 1;
  }
}
__global__ void k_compact_i32(const int32_t* __restrict__ flag, const int64_t* __restrict__ excl,
                              const int32_t* __restrict__ val, int64_t n, int32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flag[i]) out[excl[i]] = val[i];
}

extern "C" int vsv_cluster(vsv_ctx* ctx, int is_del, const int32_t* pos, const int32_t* len, int64_t n,
                           int32_t* keep, int64_t* n_keep) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0 || n >= (int64_t(1) << 31) || !n_keep) VSV_FAIL(ctx, VSV_E_ARG, "vsv_cluster: bad argument");
  *n_keep = 0;
  if (n == 0) return VSV_OK;
  const int32_t *dp, *dl;
  VSV_TRY(vsv_in(ctx, pos, (size_t)n, &dp));
  VSV_TRY(vsv_in(ctx, len, (size_t)n, &dl));
  int32_t *dk, *cl, *best, *lead;
  int64_t *excl, *tot;
  VSV_TRY(vsv_out(ctx, keep, (size_t)n, &dk));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &cl));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &best));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &lead));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &excl));
  VSV_TRY(vsv_tmp(ctx, 1, &tot));
  const unsigned g = (unsigned)vsv_cdiv(n, 128);
  VSV_TRY(vsv_t0(ctx));
  VSV_CUDA(ctx, cudaMemsetAsync(cl, 0xFF, sizeof(int32_t) * (size_t)n, ctx->stream));
  VSV_LAUNCH(ctx, k_cluster, g, 128, 0, dp, dl, n, is_del, 100, cl, best, lead);
  VSV_TRY(vsv_exclusive_scan_i32(ctx, lead, n, excl, tot));
  VSV_LAUNCH(ctx, k_compact_i32, g, 128, 0, lead, excl, best, n, dk);
  VSV_TRY(vsv_t1(ctx));
  VSV_CUDA(ctx, cudaMemcpyAsync(n_keep, tot, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  VSV_TRY(vsv_out_done(ctx, keep, dk, (size_t)*n_keep));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// haplotype pairing (pair_sig matching loop, ECS_H:548-569)
// --------------------------------------------------------------------------------
__global__ void k_pair_windows(const int32_t* __restrict__ pos1, int64_t n1, const int32_t* __restrict__ pos2,
                               int64_t n2, int max_shift, int32_t* __restrict__ lo, int32_t* __restrict__ hi) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n1) return;
  const int64_t p = pos1[i];
  const int64_t a = max(p - max_shift, (int64_t)INT32_MIN), b = min(p + max_shift, (int64_t)INT32_MAX);
  lo[i] = (int32_t)lower_bound_i32(pos2, 0, n2, (int32_t)a);
  hi[i] = (int32_t)upper_bound_i32(pos2, 0, n2, (int32_t)b);
}
// hp2 candidates of hp1[i] are pos2 in [pos1-200, pos1+200] (the reference scans from 0
// and breaks at dist>1000, ECS_H:554-559, but everything outside +-200 fails the shift
// test).  Windows are monotone in i, so hp1 splits into independent segments wherever
// hi[i-1] <= lo[i]; one thread walks a segment in hp1 order, taking the first
// still-unpaired match, exactly like the reference's greedy loop.
__global__ void k_pair(const int32_t* __restrict__ pos1, const int32_t* __restrict__ len1,
                       const int32_t* __restrict__ typ1, int64_t n1, const int32_t* __restrict__ pos2,
                       const int32_t* __restrict__ len2, const int32_t* __restrict__ typ2,
                       const int32_t* __restrict__ lo, const int32_t* __restrict__ hi, int32_t* __restrict__ p1,
                       int32_t* __restrict__ p2) {
  int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i0 >= n1) return;
  if (i0 != 0 && hi[i0 - 1] > lo[i0]) return;  // not a segment head
  for (int64_t i = i0; i < n1; ++i) {
    if (i != i0 && hi[i - 1] <= lo[i]) break;
    int32_t partner = -1;
    for (int32_t j = lo[i]; j < hi[i]; ++j) {
      if (typ1[i] != typ2[j] || p2[j] != -1) continue;
      if (!size_sim_half(len1[i], len2[j])) continue;
      if (typ1[i] == VSV_DEL && !overlap_half(pos1[i], len1[i], pos2[j], len2[j])) continue;
      partner = j;
      p2[j] = (int32_t)i;
      break;
    }
    p1[i] = partner;
  }
}

extern "C" int vsv_pair_haps(vsv_ctx* ctx, const int32_t* pos1, const int32_t* len1, const int32_t* typ1,
                             int64_t n1, const int32_t* pos2, const int32_t* len2, const int32_t* typ2,
                             int64_t n2, int32_t* partner1, int32_t* partner2) {
  VSV_TRY(vsv_begin(ctx));
  if (n1 < 0 || n2 < 0 || n1 >= (int64_t(1) << 31) || n2 >= (int64_t(1) << 31))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_pair_haps: bad argument");
  const int32_t *a1, *b1, *c1, *a2, *b2, *c2;
  VSV_TRY(vsv_in(ctx, pos1, (size_t)n1, &a1));
  VSV_TRY(vsv_in(ctx, len1, (size_t)n1, &b1));
  VSV_TRY(vsv_in(ctx, typ1, (size_t)n1, &c1));
  VSV_TRY(vsv_in(ctx, pos2, (size_t)n2, &a2));
  VSV_TRY(vsv_in(ctx, len2, (size_t)n2, &b2));
  VSV_TRY(vsv_in(ctx, typ2, (size_t)n2, &c2));
  int32_t *d1, *d2, *lo, *hi;
  VSV_TRY(vsv_out(ctx, partner1, (size_t)n1, &d1));
  VSV_TRY(vsv_out(ctx, partner2, (size_t)n2, &d2));
  VSV_TRY(vsv_tmp(ctx, (size_t)n1, &lo));
  VSV_TRY(vsv_tmp(ctx, (size_t)n1, &hi));
  VSV_TRY(vsv_t0(ctx));
  if (n2 > 0) VSV_CUDA(ctx, cudaMemsetAsync(d2, 0xFF, sizeof(int32_t) * (size_t)n2, ctx->stream));
  if (n1 > 0) {
    const unsigned g = (unsigned)vsv_cdiv(n1, 128);
    VSV_LAUNCH(ctx, k_pair_windows, g, 128, 0, a1, n1, a2, n2, 200, lo, hi);
    VSV_LAUNCH(ctx, k_pair, g, 128, 0, a1, b1, c1, n1, a2, b2, c2, lo, hi, d1, d2);
  }
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, partner1, d1, (size_t)n1));
  VSV_TRY(vsv_out_done(ctx, partner2, d2, (size_t)n2));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// window joins: FP filter support, INS / DEL coverage
// --------------------------------------------------------------------------------
#define WIN_GROUP 8
__global__ void k_fp_support(const int32_t* __restrict__ cpos, const int32_t* __restrict__ clen, int64_t n_call,
                             const int32_t* __restrict__ spos, const int32_t* __restrict__ slen, int64_t n_sig,
                             int max_comp, int max_shift, int32_t* __restrict__ support) {
  const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / WIN_GROUP;
  const int g = threadIdx.x % WIN_GROUP;
  const unsigned gmask = 0xFFu << ((threadIdx.x & 31) / WIN_GROUP * WIN_GROUP);
  if (k >= n_call) return;
  const int32_t cl = clen[k];
  int32_t s = 0;
  if (cl > max_comp) {                       // FPF:110-111
    s = (g == 0) ? 60 : 0;
  } else {
    // the +-1000 scan window (FPF:115-119) contains the +-500 shift test (FPF:99) on a
    // position-sorted signature file, so only [pos-500, pos+500] can match
    const int64_t p = cpos[k];
    const int64_t a = lower_bound_i32(spos, 0, n_sig, (int32_t)max(p - max_shift, (int64_t)INT32_MIN));
    const int64_t b = upper_bound_i32(spos, a, n_sig, (int32_t)min(p + max_shift, (int64_t)INT32_MAX));
    for (int64_t i = a + g; i < b; i += WIN_GROUP) {
      const int32_t sl = __ldg(slen + i);
      const int32_t mx = max(cl, sl);
      if (mx != 0 && 2 * (int64_t)min(cl, sl) >= mx) ++s;  // 0/0 -> similarity 0 (FPF:94-97)
    }
  }
#pragma unroll
  for (int d = WIN_GROUP / 2; d >= 1; d >>= 1) s += __shfl_xor_sync(gmask, s, d);
  if (g == 0) support[k] = s;
}

extern "C" int vsv_fp_support(vsv_ctx* ctx, const int32_t* call_pos, const int32_t* call_len, int64_t n_call,
                              const int32_t* sig_pos, const int32_t* sig_len, int64_t n_sig, int32_t* support) {
  VSV_TRY(vsv_begin(ctx));
  if (n_call < 0 || n_sig < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_fp_support: bad argument");
  if (n_call == 0) return VSV_OK;
  const int32_t *cp, *cl, *sp, *sl;
  VSV_TRY(vsv_in(ctx, call_pos, (size_t)n_call, &cp));
  VSV_TRY(vsv_in(ctx, call_len, (size_t)n_call, &cl));
  VSV_TRY(vsv_in(ctx, sig_pos, (size_t)n_sig, &sp));
  VSV_TRY(vsv_in(ctx, sig_len, (size_t)n_sig, &sl));
  int32_t* ds;
  VSV_TRY(vsv_out(ctx, support, (size_t)n_call, &ds));
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_fp_support, (unsigned)vsv_cdiv(n_call * WIN_GROUP, 256), 256, 0, cp, cl, n_call, sp, sl, n_sig, 250,
             500, ds);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, support, ds, (size_t)n_call));
  return vsv_end(ctx);
}

// INS coverage (CSS:81-125): sum of INS sig lengths with |sig_pos - call_pos| <= flanking.
__global__ void k_ins_cov(const int32_t* __restrict__ cpos, int64_t n_call, const int32_t* __restrict__ spos,
                          const int32_t* __restrict__ slen, int64_t n_sig, int flanking, int64_t* __restrict__ cov) {
  const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / WIN_GROUP;
  const int g = threadIdx.x % WIN_GROUP;
  const unsigned gmask = 0xFFu << ((threadIdx.x & 31) / WIN_GROUP * WIN_GROUP);
  if (k >= n_call) return;
  const int64_t p = cpos[k];
  const int64_t a = lower_bound_i32(spos, 0, n_sig, (int32_t)max(p - flanking, (int64_t)INT32_MIN));
  const int64_t b = upper_bound_i32(spos, a, n_sig, (int32_t)min(p + flanking, (int64_t)INT32_MAX));
  long long s = 0;
  for (int64_t i = a + g; i < b; i += WIN_GROUP) s += __ldg(slen + i);
#pragma unroll
  for (int d = WIN_GROUP / 2; d >= 1; d >>= 1) s += __shfl_xor_sync(gmask, s, d);
  if (g == 0) cov[k] = s;
}

extern "C" int vsv_ins_cov(vsv_ctx* ctx, const int32_t* call_pos, int64_t n_call, const int32_t* sig_pos,
                           const int32_t* sig_len, int64_t n_sig, int flanking, int64_t* cov) {
  VSV_TRY(vsv_begin(ctx));
  if (n_call < 0 || n_sig < 0 || flanking < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_ins_cov: bad argument");
  if (n_call == 0) return VSV_OK;
  const int32_t *cp, *sp, *sl;
  VSV_TRY(vsv_in(ctx, call_pos, (size_t)n_call, &cp));
  VSV_TRY(vsv_in(ctx, sig_pos, (size_t)n_sig, &sp));
  VSV_TRY(vsv_in(ctx, sig_len, (size_t)n_sig, &sl));
  int64_t* dc;
  VSV_TRY(vsv_out(ctx, cov, (size_t)n_call, &dc));
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_ins_cov, (unsigned)vsv_cdiv(n_call * WIN_GROUP, 256), 256, 0, cp, n_call, sp, sl, n_sig, flanking, dc);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, cov, dc, (size_t)n_call));
  return vsv_end(ctx);
}

// DEL coverage (CSS:138-253): the four sweeps collect the set of sigs whose closed
// interval [pos, pos+len] overlaps the call window [start-fl, end+fl]; cov = -sum(len).
__global__ void k_sig_end(const int32_t* __restrict__ spos, const int32_t* __restrict__ slen, int64_t n,
                          int32_t* __restrict__ send) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) send[i] = (int32_t)min((int64_t)spos[i] + slen[i], (int64_t)INT32_MAX);
}
__global__ void k_del_cov(const int32_t* __restrict__ cs, const int32_t* __restrict__ ce, int64_t n_call,
                          const int32_t* __restrict__ spos, const int32_t* __restrict__ slen,
                          const int32_t* __restrict__ send, const int32_t* __restrict__ pmax, int64_t n_sig,
                          int flanking, int64_t* __restrict__ cov) {
  const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / WIN_GROUP;
  const int g = threadIdx.x % WIN_GROUP;
  const unsigned gmask = 0xFFu << ((threadIdx.x & 31) / WIN_GROUP * WIN_GROUP);
  if (k >= n_call) return;
  const int64_t lb = (int64_t)cs[k] - flanking, rb = (int64_t)ce[k] + flanking;
  const int64_t hi = upper_bound_i32(spos, 0, n_sig, (int32_t)min(rb, (int64_t)INT32_MAX));  // start <= rb
  const int64_t lo = lower_bound_i32(pmax, 0, hi, (int32_t)max(lb, (int64_t)INT32_MIN));     // running max end >= lb
  long long s = 0;
  for (int64_t i = lo + g; i < hi; i += WIN_GROUP)
    if ((int64_t)__ldg(send + i) >= lb) s -= __ldg(slen + i);
#pragma unroll
  for (int d = WIN_GROUP / 2; d >= 1; d >>= 1) s += __shfl_xor_sync(gmask, s, d);
  if (g == 0) cov[k] = s;
}

extern "C" int vsv_del_cov(vsv_ctx* ctx, const int32_t* call_start, const int32_t* call_end, int64_t n_call,
                           const int32_t* sig_pos, const int32_t* sig_len, int64_t n_sig, int flanking,
                           int64_t* cov) {
  VSV_TRY(vsv_begin(ctx));
  if (n_call < 0 || n_sig < 0 || flanking < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_del_cov: bad argument");
  if (n_call == 0) return VSV_OK;
  const int32_t *cs, *ce, *sp, *sl;
  VSV_TRY(vsv_in(ctx, call_start, (size_t)n_call, &cs));
  VSV_TRY(vsv_in(ctx, call_end, (size_t)n_call, &ce));
  VSV_TRY(vsv_in(ctx, sig_pos, (size_t)n_sig, &sp));
  VSV_TRY(vsv_in(ctx, sig_len, (size_t)n_sig, &sl));
  int64_t* dc;
  int32_t *send, *pmax;
  VSV_TRY(vsv_out(ctx, cov, (size_t)n_call, &dc));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_sig, &send));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_sig, &pmax));
  VSV_TRY(vsv_t0(ctx));
  if (n_sig > 0) {
    VSV_LAUNCH(ctx, k_sig_end, (unsigned)vsv_cdiv(n_sig, 256), 256, 0, sp, sl, n_sig, send);
    VSV_TRY(inclusive_scan_i32(ctx, send, pmax, n_sig, MaxI32()));
  }
  VSV_LAUNCH(ctx, k_del_cov, (unsigned)vsv_cdiv(n_call * WIN_GROUP, 256), 256, 0, cs, ce, n_call, sp, sl, send, pmax, n_sig,
             flanking, dc);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, cov, dc, (size_t)n_call));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// signature sort + column extraction (device-resident pipeline plumbing)
// --------------------------------------------------------------------------------
// tie_mode 0: stable by pos (sort_sig, ECS_H:170-179).
// tie_mode 1: `.sigs` order (SE:637-638 `sort -k2,2 -k3,3n`, LC_ALL=C): pos numeric, ties by
//             the rest of the line, whose first field is len as a DECIMAL STRING
//             ("100" < "95", "10" < "100").
// After the stable sort by position, ties are in input order.  The `.sigs` tie order is a property
// of the runs of equal position only, and those runs are short (a handful of reads at one base),
// so they are fixed in place instead of paying a second full-width radix sort:
//   run <= TIE_SMALL : the thread at the run head does a stable insertion sort (registers/local);
//   longer runs      : queued, one CTA each, stable rank counting in k_tie_long.
#define TIE_SMALL 32
__global__ void k_tie_small(const vsv_sig* __restrict__ s, const int32_t* __restrict__ pos_sorted, int64_t n,
                            int32_t* __restrict__ order, int2* __restrict__ work, unsigned int* __restrict__ n_work) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t p = pos_sorted[i];
  if (i > 0 && pos_sorted[i - 1] == p) return;  // not a run head
  int64_t e = i + 1;
  while (e < n && e - i <= TIE_SMALL && pos_sorted[e] == p) ++e;
  const int L = (int)(e - i);
  if (L == 1) return;
  if (L > TIE_SMALL) {
    while (e < n && pos_sorted[e] == p) ++e;
    work[atomicAdd(n_work, 1u)] = make_int2((int32_t)i, (int32_t)(e - i));
    return;
  }
  int32_t id[TIE_SMALL];
  unsigned long long key[TIE_SMALL];
  for (int k = 0; k < L; ++k) {
    const int32_t x = order[i + k];
    const unsigned long long kx = len_string_key(s[x].len);
    int m = k;
    while (m > 0 && key[m - 1] > kx) { key[m] = key[m - 1]; id[m] = id[m - 1]; --m; }
    key[m] = kx; id[m] = x;
  }
  for (int k = 0; k < L; ++k) order[i + k] = id[k];
}
__global__ void __launch_bounds__(256) k_tie_long(const vsv_sig* __restrict__ s, int32_t* __restrict__ order,
                                                   const int2* __restrict__ work, const unsigned int* __restrict__ n_work,
                                                   unsigned long long* __restrict__ key_tmp, int32_t* __restrict__ ord_tmp) {
  const unsigned int nw = *n_work;
  for (unsigned int w = blockIdx.x; w < nw; w += gridDim.x) {
    const int64_t a = work[w].x;
    const int L = work[w].y;
    for (int k = threadIdx.x; k < L; k += blockDim.x) {
      const int32_t x = order[a + k];
      ord_tmp[a + k] = x;
      key_tmp[a + k] = len_string_key(s[x].len);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < L; k += blockDim.x) {
      const unsigned long long kk = key_tmp[a + k];
      int rank = 0;
      for (int m = 0; m < L; ++m) {
        const unsigned long long km = key_tmp[a + m];
        rank += (km < kk || (km == kk && m < k)) ? 1 : 0;
      }
      order[a + rank] = ord_tmp[a + k];
    }
    __syncthreads();
  }
}
__global__ void k_gather_pos(const vsv_sig* __restrict__ s, const int32_t* __restrict__ idx, int64_t n,
                             int32_t* __restrict__ pos) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pos[i] = s[idx ? idx[i] : i].pos;
}
__global__ void k_gather_sigs(const vsv_sig* __restrict__ s, const int32_t* __restrict__ idx, int64_t n,
                              vsv_sig* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4* src = reinterpret_cast<const int4*>(s + idx[i]);
  int4* dst = reinterpret_cast<int4*>(out + i);
  dst[0] = src[0];
  dst[1] = src[1];
}

extern "C" int vsv_sort_sigs(vsv_ctx* ctx, const vsv_sig* sigs, int64_t n, int tie_mode, vsv_sig* out) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0 || n >= (int64_t(1) << 31) || tie_mode < 0 || tie_mode > 1) VSV_FAIL(ctx, VSV_E_ARG, "vsv_sort_sigs: bad argument");
  if (n == 0) return VSV_OK;
  const vsv_sig* din;
  vsv_sig* dout;
  VSV_TRY(vsv_in(ctx, sigs, (size_t)n, &din));
  VSV_TRY(vsv_out(ctx, out, (size_t)n, &dout));
  int32_t *idx0 = nullptr, *idx1 = nullptr, *pos, *order;
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &pos));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &order));
  const unsigned g = (unsigned)vsv_cdiv(n, 256);
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_gather_pos, g, 256, 0, din, (const int32_t*)nullptr, n, pos);
  int32_t* ks;
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &ks));
  VSV_TRY(argsort_dev(ctx, pos, n, order, ks));
  if (tie_mode == 1) {
    unsigned long long* key_tmp;
    int32_t* ord_tmp;
    int2* work;
    unsigned int* n_work;
    VSV_TRY(vsv_tmp(ctx, (size_t)n, &key_tmp));
    VSV_TRY(vsv_tmp(ctx, (size_t)n, &ord_tmp));
    VSV_TRY(vsv_tmp(ctx, (size_t)(n / (TIE_SMALL + 1) + 1), &work));
    VSV_TRY(vsv_tmp(ctx, 1, &n_work));
    VSV_CUDA(ctx, cudaMemsetAsync(n_work, 0, sizeof(unsigned int), ctx->stream));
    VSV_LAUNCH(ctx, k_tie_small, g, 256, 0, din, (const int32_t*)ks, n, order, work, n_work);
    VSV_LAUNCH(ctx, k_tie_long, (unsigned)ctx->sm_count, 256, 0, din, order, (const int2*)work, (const unsigned int*)n_work,
               key_tmp, ord_tmp);
  }
  VSV_LAUNCH(ctx, k_gather_sigs, g, 256, 0, din, order, n, dout);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, out, dout, (size_t)n));
  return vsv_end(ctx);
}

__global__ void k_sig_columns(const vsv_sig* __restrict__ s, int64_t n, int32_t* __restrict__ pos,
                              int32_t* __restrict__ len, int32_t* __restrict__ rec) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4 a = *reinterpret_cast<const int4*>(s + i);  // rec, type, pos, len
  if (pos) pos[i] = a.z;
  if (len) len[i] = a.w;
  if (rec) rec[i] = a.x;
}

extern "C" int vsv_sig_columns(vsv_ctx* ctx, const vsv_sig* sigs, int64_t n, int32_t* pos, int32_t* len,
                               int32_t* rec) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_sig_columns: bad argument");
  if (n == 0) return VSV_OK;
  const vsv_sig* din;
  int32_t *dp, *dl, *dr;
  VSV_TRY(vsv_in(ctx, sigs, (size_t)n, &din));
  VSV_TRY(vsv_out(ctx, pos, (size_t)n, &dp));
  VSV_TRY(vsv_out(ctx, len, (size_t)n, &dl));
  VSV_TRY(vsv_out(ctx, rec, (size_t)n, &dr));
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_sig_columns, (unsigned)vsv_cdiv(n, 256), 256, 0, din, n, dp, dl, dr);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, pos, dp, (size_t)n));
  VSV_TRY(vsv_out_done(ctx, len, dl, (size_t)n));
  VSV_TRY(vsv_out_done(ctx, rec, dr, (size_t)n));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// elementwise glue for the device-resident pipeline (keeps everything on the ctx stream)
// --------------------------------------------------------------------------------
__global__ void k_add_i32(const int32_t* __restrict__ a, const int32_t* __restrict__ b, int64_t n,
                          int32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i] + b[i];
}
__global__ void k_ge_u8(const int32_t* __restrict__ a, int32_t thr, int64_t n, uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i] >= thr ? 1 : 0;
}
__global__ void k_select_i32(const int32_t* __restrict__ key, int32_t val, const int32_t* __restrict__ a,
                             const int32_t* __restrict__ b, int64_t n, int32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = key[i] == val ? a[i] : b[i];
}

extern "C" int vsv_add_i32(vsv_ctx* ctx, const int32_t* a, const int32_t* b, int64_t n, int32_t* out) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_add_i32: bad n");
  if (n == 0) return VSV_OK;
  const int32_t *da, *db;
  int32_t* dout;
  VSV_TRY(vsv_in(ctx, a, (size_t)n, &da));
  VSV_TRY(vsv_in(ctx, b, (size_t)n, &db));
  VSV_TRY(vsv_out(ctx, out, (size_t)n, &dout));
  VSV_LAUNCH(ctx, k_add_i32, (unsigned)vsv_cdiv(n, 256), 256, 0, da, db, n, dout);
  VSV_TRY(vsv_out_done(ctx, out, dout, (size_t)n));
  return vsv_end(ctx);
}
extern "C" int vsv_ge_u8(vsv_ctx* ctx, const int32_t* a, int32_t thr, int64_t n, uint8_t* out) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_ge_u8: bad n");
  if (n == 0) return VSV_OK;
  const int32_t* da;
  uint8_t* dout;
  VSV_TRY(vsv_in(ctx, a, (size_t)n, &da));
  VSV_TRY(vsv_out(ctx, out, (size_t)n, &dout));
  VSV_LAUNCH(ctx, k_ge_u8, (unsigned)vsv_cdiv(n, 256), 256, 0, da, thr, n, dout);
  VSV_TRY(vsv_out_done(ctx, out, dout, (size_t)n));
  return vsv_end(ctx);
}
extern "C" int vsv_select_i32(vsv_ctx* ctx, const int32_t* key, int32_t val, const int32_t* a, const int32_t* b,
                              int64_t n, int32_t* out) {
  VSV_TRY(vsv_begin(ctx));
  if (n < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_select_i32: bad n");
  if (n == 0) return VSV_OK;
  const int32_t *dk, *da, *db;
  int32_t* dout;
  VSV_TRY(vsv_in(ctx, key, (size_t)n, &dk));
  VSV_TRY(vsv_in(ctx, a, (size_t)n, &da));
  VSV_TRY(vsv_in(ctx, b, (size_t)n, &db));
  VSV_TRY(vsv_out(ctx, out, (size_t)n, &dout));
  VSV_LAUNCH(ctx, k_select_i32, (unsigned)vsv_cdiv(n, 256), 256, 0, dk, val, da, db, n, dout);
  VSV_TRY(vsv_out_done(ctx, out, dout, (size_t)n));
  return vsv_end(ctx);
}

// --------------------------------------------------------------------------------
// compact host->device staging of the packed CIGAR ops ("ops16")
// --------------------------------------------------------------------------------
// PCIe, not HBM, bounds the end-to-end path (4 bytes per op over a ~55 GB/s link), so the host
// packer may stage the ops as 16-bit words: op i = len << 4 | code when len < 4095, else the
// marker 0xFFF << 4 | code with the full 32-bit word in a sparse exception list.  This kernel
// widens them back to the BAM-native u32 array every scanner reads (8 ops per thread: one
// LDG.128 in, two STG.128 out); a second, tiny kernel patches the exceptions.
__global__ void __launch_bounds__(256) k_expand_ops16(const uint16_t* __restrict__ in, int64_t n, uint32_t* __restrict__ out) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + i));
    uint4 a, b;
    a.x = v.x & 0xFFFFu; a.y = v.x >> 16; a.z = v.y & 0xFFFFu; a.w = v.y >> 16;
    b.x = v.z & 0xFFFFu; b.y = v.z >> 16; b.z = v.w & 0xFFFFu; b.w = v.w >> 16;
    reinterpret_cast<uint4*>(out + i)[0] = a;
    reinterpret_cast<uint4*>(out + i)[1] = b;
  } else {
    for (int64_t k = i; k < n; ++k) out[k] = in[k];
  }
}
__global__ void k_patch_ops(const int64_t* __restrict__ idx, const uint32_t* __restrict__ val, int64_t n_exc, int64_t n,
                            uint32_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_exc && idx[i] >= 0 && idx[i] < n) out[idx[i]] = val[i];
}

// ops8: one byte per op.  The 254 codes are split by how long each op type runs in noisy long reads:
// 0..159 = M of 1..160 bases, 160..206 = I of 1..47, 207..253 = D of 1..47; 0xFF marks an op whose full
// word is in the exception list (anything else: S/H/N/P/=/X, longer or empty ops).  16 ops per thread:
// one LDG.128 in, four STG.128 out.
__device__ __forceinline__ uint32_t ops8_word(uint32_t b) {
  const uint32_t code = (b >= VSV_OPS8_M ? 1u : 0u) + (b >= VSV_OPS8_M + VSV_OPS8_I ? 1u : 0u);
  const uint32_t base = code == 0u ? 0u : (code == 1u ? (uint32_t)VSV_OPS8_M : (uint32_t)(VSV_OPS8_M + VSV_OPS8_I));
  return ((b - base + 1u) << 4) | code;
}
__global__ void __launch_bounds__(256) k_expand_ops8(const uint8_t* __restrict__ in, int64_t n, uint32_t* __restrict__ out) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (i + 16 <= n) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + i));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint4 o;
      o.x = ops8_word(w[k] & 0xFFu); o.y = ops8_word((w[k] >> 8) & 0xFFu);
      o.z = ops8_word((w[k] >> 16) & 0xFFu); o.w = ops8_word(w[k] >> 24);
      reinterpret_cast<uint4*>(out + i)[k] = o;
    }
  } else {
    for (int64_t k = i; k < n; ++k) out[k] = ops8_word(in[k]);
  }
}

extern "C" int vsv_expand_ops8(vsv_ctx* ctx, const uint8_t* ops8, int64_t n_ops, const int64_t* exc_idx,
                               const uint32_t* exc_val, int64_t n_exc, uint32_t* ops_out) {
  VSV_TRY(vsv_begin(ctx));
  if (n_ops < 0 || n_exc < 0 || (n_ops > 0 && (!ops8 || !ops_out)) || (n_exc > 0 && (!exc_idx || !exc_val)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_expand_ops8: bad argument");
  if (n_ops == 0) return VSV_OK;
  const uint8_t* din;
  const int64_t* dix;
  const uint32_t* dvl;
  uint32_t* dout;
  VSV_TRY(vsv_in(ctx, ops8, (size_t)n_ops, &din));
  VSV_TRY(vsv_in(ctx, exc_idx, (size_t)n_exc, &dix));
  VSV_TRY(vsv_in(ctx, exc_val, (size_t)n_exc, &dvl));
  VSV_TRY(vsv_out(ctx, ops_out, (size_t)n_ops, &dout));
  if ((reinterpret_cast<uintptr_t>(din) & 15u) || (reinterpret_cast<uintptr_t>(dout) & 15u))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_expand_ops8: buffers must be 16-byte aligned");
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_expand_ops8, (unsigned)vsv_cdiv(vsv_cdiv(n_ops, 16), 256), 256, 0, din, n_ops, dout);
  if (n_exc > 0) VSV_LAUNCH(ctx, k_patch_ops, (unsigned)vsv_cdiv(n_exc, 256), 256, 0, dix, dvl, n_exc, n_ops, dout);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, ops_out, dout, (size_t)n_ops));
  return vsv_end(ctx);
}

// ops12: 12 bits per PAIR of ops (0.75 byte per op).  minimap2's long-read CIGARs alternate a match run with
// a short indel, so the pair (ops[2k], ops[2k+1]) is almost always {M of 1..128, I/D of 1..8} in one of the two
// orders: token = mlen-1 | (xlen-1) << 7 | (D ? 1 : 0) << 10 | (indel first ? 1 : 0) << 11, tokens packed
// little-endian into the byte stream (token k = bits [12k, 12k+12)), the stream padded to a whole group of
// 8 tokens (12 bytes).  Any other pair has both full words in the exception list.  One thread per group:
// three LDG.32 in, four STG.128 out.  Pairing is positional over the whole ops array, so no prefix scan.
__device__ __forceinline__ uint2 ops12_pair(uint32_t t) {
  const uint32_t m = ((t & 127u) + 1u) << 4;
  const uint32_t x = ((((t >> 7) & 7u) + 1u) << 4) | (1u + ((t >> 10) & 1u));
  return (t & 0x800u) ? make_uint2(x, m) : make_uint2(m, x);
}
__global__ void __launch_bounds__(256) k_expand_ops12(const uint32_t* __restrict__ in, int64_t n_grp, int64_t n,
                                                      uint32_t* __restrict__ out) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_grp) return;
  const uint32_t w0 = __ldg(in + 3 * g), w1 = __ldg(in + 3 * g + 1), w2 = __ldg(in + 3 * g + 2);
  const uint32_t t[8] = {w0 & 0xFFFu, (w0 >> 12) & 0xFFFu, (w0 >> 24) | ((w1 & 0xFu) << 8), (w1 >> 4) & 0xFFFu,
                         (w1 >> 16) & 0xFFFu, (w1 >> 28) | ((w2 & 0xFFu) << 4), (w2 >> 8) & 0xFFFu, w2 >> 20};
  const int64_t i = g * 16;
  if (i + 16 <= n) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint2 a = ops12_pair(t[2 * k]), b = ops12_pair(t[2 * k + 1]);
      reinterpret_cast<uint4*>(out + i)[k] = make_uint4(a.x, a.y, b.x, b.y);
    }
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint2 a = ops12_pair(t[k]);
      if (i + 2 * k < n) out[i + 2 * k] = a.x;
      if (i + 2 * k + 1 < n) out[i + 2 * k + 1] = a.y;
    }
  }
}
__global__ void k_patch_pairs(const int64_t* __restrict__ pair, const uint32_t* __restrict__ val, int64_t n_exc, int64_t n,
                              uint32_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_exc) return;
  const int64_t o = 2 * pair[i];
  if (o < 0) return;
  if (o < n) out[o] = val[2 * i];
  if (o + 1 < n) out[o + 1] = val[2 * i + 1];
}

extern "C" int64_t vsv_ops12_bytes(int64_t n_ops) {
  return n_ops <= 0 ? 0 : vsv_cdiv(vsv_cdiv(n_ops, 2), 8) * 12;
}

extern "C" int vsv_expand_ops12(vsv_ctx* ctx, const uint8_t* ops12, int64_t n_ops, const int64_t* exc_pair,
                                const uint32_t* exc_val, int64_t n_exc, uint32_t* ops_out) {
  VSV_TRY(vsv_begin(ctx));
  if (n_ops < 0 || n_exc < 0 || (n_ops > 0 && (!ops12 || !ops_out)) || (n_exc > 0 && (!exc_pair || !exc_val)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_expand_ops12: bad argument");
  if (n_ops == 0) return VSV_OK;
  const int64_t n_bytes = vsv_ops12_bytes(n_ops), n_grp = n_bytes / 12;
  const uint8_t* din;
  const int64_t* dix;
  const uint32_t* dvl;
  uint32_t* dout;
  VSV_TRY(vsv_in(ctx, ops12, (size_t)n_bytes, &din));
  VSV_TRY(vsv_in(ctx, exc_pair, (size_t)n_exc, &dix));
  VSV_TRY(vsv_in(ctx, exc_val, (size_t)(2 * n_exc), &dvl));
  VSV_TRY(vsv_out(ctx, ops_out, (size_t)n_ops, &dout));
  if ((reinterpret_cast<uintptr_t>(din) & 3u) || (reinterpret_cast<uintptr_t>(dout) & 15u))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_expand_ops12: ops12 must be 4-byte aligned, ops_out 16-byte aligned");
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_expand_ops12, (unsigned)vsv_cdiv(n_grp, 256), 256, 0, reinterpret_cast<const uint32_t*>(din), n_grp,
             n_ops, dout);
  if (n_exc > 0) VSV_LAUNCH(ctx, k_patch_pairs, (unsigned)vsv_cdiv(n_exc, 256), 256, 0, dix, dvl, n_exc, n_ops, dout);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, ops_out, dout, (size_t)n_ops));
  return vsv_end(ctx);
}

extern "C" int vsv_expand_ops16(vsv_ctx* ctx, const uint16_t* ops16, int64_t n_ops, const int64_t* exc_idx,
                                const uint32_t* exc_val, int64_t n_exc, uint32_t* ops_out) {
  VSV_TRY(vsv_begin(ctx));
  if (n_ops < 0 || n_exc < 0 || (n_ops > 0 && (!ops16 || !ops_out)) || (n_exc > 0 && (!exc_idx || !exc_val)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_expand_ops16: bad argument");
  if (n_ops == 0) return VSV_OK;
  const uint16_t* din;
  const int64_t* dix;
  const uint32_t* dvl;
  uint32_t* dout;
  VSV_TRY(vsv_in(ctx, ops16, (size_t)n_ops, &din));
  VSV_TRY(vsv_in(ctx, exc_idx, (size_t)n_exc, &dix));
  VSV_TRY(vsv_in(ctx, exc_val, (size_t)n_exc, &dvl));
  VSV_TRY(vsv_out(ctx, ops_out, (size_t)n_ops, &dout));
  if ((reinterpret_cast<uintptr_t>(din) & 15u) || (reinterpret_cast<uintptr_t>(dout) & 15u))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_expand_ops16: buffers must be 16-byte aligned");
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_expand_ops16, (unsigned)vsv_cdiv(vsv_cdiv(n_ops, 8), 256), 256, 0, din, n_ops, dout);
  if (n_exc > 0) VSV_LAUNCH(ctx, k_patch_ops, (unsigned)vsv_cdiv(n_exc, 256), 256, 0, dix, dvl, n_exc, n_ops, dout);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, ops_out, dout, (size_t)n_ops));
  return vsv_end(ctx);
}
