// K2: per-record merge (split out of vsv_scan.cu).
#include "vsv_rules.cuh"

// ================================================================================
// K2: per-record merge.  cluster_ins_one_read / cluster_del_one_read (ECS_H:87-164),
// generate_combine_sigs (SE:373-435).  Signatures are sparse (<= a few per record), so
// this is one thread per record run walking its (already ordered) signatures, both types in
// one walk with a state per type; two passes (count, exclusive scan, write) keep the output
// order deterministic.
// ================================================================================
#include <cub/device/device_scan.cuh>

struct SumU4 {
  __host__ __device__ uint4 operator()(const uint4& a, const uint4& b) const {
    return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
};

// per record-run head: (DEL in, INS in, DEL out, INS out); zeros elsewhere
__global__ void k_merge2_count(const vsv_sig* __restrict__ s, int64_t n, int mode, uint4* __restrict__ cnt) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint4 c = make_uint4(0u, 0u, 0u, 0u);
  const int32_t r = s[i].rec;
  if (i == 0 || s[i - 1].rec != r) {
    int64_t e = i + 1;
    while (e < n && s[e].rec == r) ++e;
    for (int64_t k = i; k < e; ++k) {
      if (s[k].type == VSV_DEL) ++c.x; else ++c.y;
    }
    merge_walk2(s, i, e, mode, 0, 0, [&](const vsv_sig&) { ++c.z; }, [&](const vsv_sig&) { ++c.w; });
  }
  cnt[i] = c;
}

__global__ void k_merge2_write(const vsv_sig* __restrict__ s, int64_t n, int mode, const uint4* __restrict__ excl,
                               vsv_sig* __restrict__ out_del, vsv_sig* __restrict__ out_ins) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t r = s[i].rec;
  if (i == 0 || s[i - 1].rec != r) {
    int64_t e = i + 1;
    while (e < n && s[e].rec == r) ++e;
    const uint4 b = excl[i];
    int64_t od = b.z, oi = b.w;
    auto put = [](vsv_sig* dst, const vsv_sig& v) {
      reinterpret_cast<int4*>(dst)[0] = make_int4(v.rec, v.type, v.pos, v.len);
      reinterpret_cast<int4*>(dst)[1] = make_int4(v.q0, v.q1, v.first, v.count);
    };
    merge_walk2(s, i, e, mode, (int32_t)b.x, (int32_t)b.y, [&](const vsv_sig& v) { put(out_del + od++, v); },
                [&](const vsv_sig& v) { put(out_ins + oi++, v); });
  }
}

__global__ void k_merge2_total(const uint4* __restrict__ cnt, const uint4* __restrict__ excl, int64_t n, int64_t* __restrict__ tot) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    tot[0] = (int64_t)excl[n - 1].z + cnt[n - 1].z;
    tot[1] = (int64_t)excl[n - 1].w + cnt[n - 1].w;
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
  vsv_sig *d_od = nullptr, *d_oi = nullptr;
  VSV_TRY(vsv_out(ctx, out_del, (size_t)n, &d_od));
  VSV_TRY(vsv_out(ctx, out_ins, (size_t)n, &d_oi));
  uint4 *cnt = nullptr, *excl = nullptr;
  int64_t* tot = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &cnt));
  VSV_TRY(vsv_tmp(ctx, (size_t)n, &excl));
  VSV_TRY(vsv_tmp(ctx, 2, &tot));
  const unsigned g = (unsigned)vsv_cdiv(n, 128);
  VSV_TRY(vsv_t0(ctx));
  // one pass per record run with a state per type: count, one 4-lane exclusive scan, write
  VSV_LAUNCH(ctx, k_merge2_count, g, 128, 0, d_in, n, mode, cnt);
  {
    size_t tb = 0;
    VSV_CUDA(ctx, cub::DeviceScan::ExclusiveScan(nullptr, tb, cnt, excl, SumU4(), make_uint4(0u, 0u, 0u, 0u), n, ctx->stream));
    void* tmp = nullptr;
    VSV_TRY(vsv_arena_alloc(ctx, tb, &tmp));
    VSV_CUDA(ctx, cub::DeviceScan::ExclusiveScan(tmp, tb, cnt, excl, SumU4(), make_uint4(0u, 0u, 0u, 0u), n, ctx->stream));
    ctx->launches += 2;
  }
  VSV_LAUNCH(ctx, k_merge2_write, g, 128, 0, d_in, n, mode, (const uint4*)excl, d_od, d_oi);
  VSV_LAUNCH(ctx, k_merge2_total, 1, 32, 0, (const uint4*)cnt, (const uint4*)excl, n, tot);
  VSV_TRY(vsv_t1(ctx));
  int64_t h_tot[2] = {0, 0};
  VSV_CUDA(ctx, cudaMemcpyAsync(h_tot, tot, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n_del = h_tot[0];
  *n_ins = h_tot[1];
  VSV_TRY(vsv_out_done(ctx, out_del, d_od, (size_t)*n_del));
  VSV_TRY(vsv_out_done(ctx, out_ins, d_oi, (size_t)*n_ins));
  return vsv_end(ctx);
}
