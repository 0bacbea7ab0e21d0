// Stage 1: CIGAR signature scan (K1) and per-record merge (K2).
//
// K1 replaces extract_sig_from_cigar (LI/extract_contig_signature_Hifi.py:53-85,
// LI/extract_reads_signature.py:47-83) and the CIGAR walk of parse_read
// (LI/sig_extract.py:438-478).
//
// Design (B200-first, HBM-bound): the packed ops of ALL records form one flat u32
// array that is cut into fixed tiles of 4096 ops, independent of record boundaries,
// so every load is a coalesced, 16-byte aligned 128-bit access and the work per CTA is
// identical whether a record has 150 ops (HiFi) or 10^5 (contigs).  Reference/query
// offsets are prefix sums restarted at record starts: inside a tile they come from a
// thread-sequential pass over 16 ops + one block scan; across tiles the open record's
// running sums and the number of signatures emitted so far travel by a single-pass
// decoupled look-back (chained scan), which also gives every signature its final,
// deterministic (record, op)-ordered output slot: ordered stream compaction without a
// second pass over the ops and without unordered atomics.
#include "vsv_common.cuh"

#define SCAN_THREADS 256
#define SCAN_VPT 16
#define SCAN_TILE (SCAN_THREADS * SCAN_VPT)
#define SCAN_ROW 20  // 16 ops + 4 pad words per thread row: conflict-free LDS.128
#define SCAN_SMEM_WORDS (SCAN_THREADS * SCAN_ROW)

#define ST_AGG 1u
#define ST_INCL 2u
#define ST_BOUND 4u

template <int MODE>
struct ModeTraits;
// contig: only M,S,D,I are walked (ECS_H:72-85); aux = H
template <>
struct ModeTraits<VSV_MODE_CONTIG> {
  static constexpr uint32_t REF = (1u << 0) | (1u << 2);
  static constexpr uint32_t QRY = (1u << 0) | (1u << 1) | (1u << 4);
  static constexpr uint32_t AUX = (1u << 5);
  static constexpr int HSIGN = 1;
};
// reads: M,=,X both; S,I query; D,N reference (ERS:66-81); aux = H
template <>
struct ModeTraits<VSV_MODE_READS> {
  static constexpr uint32_t REF = (1u << 0) | (1u << 7) | (1u << 8) | (1u << 2) | (1u << 3);
  static constexpr uint32_t QRY = (1u << 0) | (1u << 7) | (1u << 8) | (1u << 4) | (1u << 1);
  static constexpr uint32_t AUX = (1u << 5);
  static constexpr int HSIGN = 1;
};
// cute: shift_del == shift_ins == sum{M,D,=,X}; shift_ins_read = every op but D
// (SE:461-478); N moves neither reference counter (quirk) -> aux = N
template <>
struct ModeTraits<VSV_MODE_CUTE> {
  static constexpr uint32_t REF = (1u << 0) | (1u << 2) | (1u << 7) | (1u << 8);
  static constexpr uint32_t QRY = 0xFFFFu & ~(1u << 2);
  static constexpr uint32_t AUX = (1u << 3);
  static constexpr int HSIGN = -1;
};

struct ScanParams {
  const uint32_t* ops;
  const int64_t* op_off;
  const int32_t* pos;
  const uint8_t* mapq;
  const uint8_t* enable;
  int64_t n_rec, n_ops;
  int32_t n_tiles;
  int min_len, min_mapq;
  // tile descriptors
  const int32_t* tile_rfirst;
  const int32_t* tile_rlast;
  uint32_t* status;
  uint4* agg;
  uint4* incl;
  uint32_t* ticket;
  // outputs
  vsv_sig* out;
  int64_t cap;
  int64_t* n_out;
  uint32_t* ref_total;
  uint32_t* qry_total;
  uint32_t* aux_total;
};

// last record r with op_off[r] <= x  (op_off[0]=0, op_off[n_rec]=n_ops > x)
__device__ __forceinline__ int32_t rec_of_op(const int64_t* __restrict__ op_off, int32_t lo, int32_t hi,
                                             int64_t x) {
  // invariant: op_off[lo] <= x, answer in [lo, hi]
  while (lo < hi) {
    int32_t mid = (lo + hi + 1) >> 1;
    if (__ldg(op_off + mid) <= x) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void k_tile_records(const int64_t* __restrict__ op_off, int64_t n_rec, int64_t n_ops,
                               int32_t n_tiles, int32_t* __restrict__ rfirst, int32_t* __restrict__ rlast) {
  int32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  int64_t x0 = (int64_t)t * SCAN_TILE;
  int64_t x1 = min(x0 + SCAN_TILE, n_ops) - 1;
  int32_t a = rec_of_op(op_off, 0, (int32_t)n_rec - 1, x0);
  rfirst[t] = a;
  rlast[t] = rec_of_op(op_off, a, (int32_t)n_rec - 1, x1);
}

template <int MODE>
__device__ __forceinline__ void op_contrib(uint32_t w, uint32_t& r, uint32_t& q, uint32_t& x) {
  typedef ModeTraits<MODE> M;
  uint32_t code = w & 15u, len = w >> 4;
  r += ((M::REF >> code) & 1u) ? len : 0u;
  q += ((M::QRY >> code) & 1u) ? len : 0u;
  x += ((M::AUX >> code) & 1u) ? len : 0u;
}

// prefix of (ref,qry,aux) over tile ops [0, j)
template <int MODE>
__device__ __forceinline__ void prefix_at(const uint32_t* s_ops, const uint32_t (*s_tx)[SCAN_THREADS],
                                          const uint32_t* s_tot, int j, uint32_t& r, uint32_t& q,
                                          uint32_t& x) {
  if (j >= SCAN_TILE) {
    r = s_tot[0]; q = s_tot[1]; x = s_tot[2];
    return;
  }
  int t = j >> 4, n = j & 15;
  r = s_tx[0][t]; q = s_tx[1][t]; x = s_tx[2][t];
  const uint32_t* row = s_ops + t * SCAN_ROW;
  for (int s = 0; s < n; ++s) op_contrib<MODE>(row[s], r, q, x);
}

template <int MODE>
__device__ __noinline__ void emit_sig(const ScanParams& p, const uint32_t* s_ops,
                                      const uint32_t (*s_tx)[SCAN_THREADS], const uint32_t* s_tot,
                                      int64_t tile_start, int32_t rfirst, int32_t rlast, int j,
                                      uint32_t pr, uint32_t pq, uint32_t w, uint32_t carry_r,
                                      uint32_t carry_q, int64_t out_idx) {
  typedef ModeTraits<MODE> M;
  if (out_idx >= p.cap) return;
  int32_t r = rec_of_op(p.op_off, rfirst, rlast, tile_start + j);
  int64_t off_r = __ldg(p.op_off + r);
  uint32_t ro, qo;
  if (off_r < tile_start) {
    ro = carry_r + pr;
    qo = carry_q + pq;
  } else {
    uint32_t br, bq, bx;
    prefix_at<MODE>(s_ops, s_tx, s_tot, (int)(off_r - tile_start), br, bq, bx);
    ro = pr - br;
    qo = pq - bq;
  }
  uint32_t f = __ldg(p.ops + off_r);
  int32_t hc = ((f & 15u) == 5u) ? (int32_t)(f >> 4) : 0;
  int32_t len = (int32_t)(w >> 4);
  int32_t type = ((w & 15u) == 1u) ? VSV_INS : VSV_DEL;
  int32_t q0 = (int32_t)qo + M::HSIGN * hc;
  int4 a = make_int4(r, type, __ldg(p.pos + r) + (int32_t)ro, len);
  int4 b = make_int4(q0, q0 + (type == VSV_INS ? len : 1), 0, 1);
  int4* o = reinterpret_cast<int4*>(p.out + out_idx);
  o[0] = a;
  o[1] = b;
}

template <int MODE>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_cigar(const ScanParams p) {
  __shared__ __align__(16) uint32_t s_ops[SCAN_SMEM_WORDS];
  __shared__ uint32_t s_tx[3][SCAN_THREADS];
  __shared__ uint32_t s_w[4][SCAN_THREADS / 32];
  __shared__ uint32_t s_tot[4];
  __shared__ uint32_t s_b[4];  // tile, carry_ref, carry_qry, cnt_in

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_b[0] = atomicAdd(p.ticket, 1u);
  __syncthreads();
  const int32_t tile = (int32_t)s_b[0];
  const int64_t tile_start = (int64_t)tile * SCAN_TILE;
  const int64_t remaining = p.n_ops - tile_start;

  // ---- stage the tile in shared memory (coalesced 128-bit loads, L1 bypass) -------
  {
    const uint4* g = reinterpret_cast<const uint4*>(p.ops + tile_start);
#pragma unroll
    for (int k = 0; k < SCAN_VPT / 4; ++k) {
      int i4 = k * SCAN_THREADS + tid;
      int i = i4 * 4;
      uint4 v;
      if (i + 4 <= remaining) {
        v = ld_cg_u4(g + i4);
      } else {
        const uint32_t* gs = p.ops + tile_start;
        v.x = (i + 0 < remaining) ? gs[i + 0] : 0u;
        v.y = (i + 1 < remaining) ? gs[i + 1] : 0u;
        v.z = (i + 2 < remaining) ? gs[i + 2] : 0u;
        v.w = (i + 3 < remaining) ? gs[i + 3] : 0u;
      }
      *reinterpret_cast<uint4*>(s_ops + i + ((i >> 4) << 2)) = v;
    }
  }
  const int32_t rfirst = __ldg(p.tile_rfirst + tile);
  const int32_t rlast = __ldg(p.tile_rlast + tile);
  __syncthreads();

  // ---- pass 1: 16 consecutive ops per thread -------------------------------------
  uint32_t w[SCAN_VPT];
  {
    const uint4* row = reinterpret_cast<const uint4*>(s_ops + tid * SCAN_ROW);
#pragma unroll
    for (int k = 0; k < SCAN_VPT / 4; ++k) {
      uint4 v = row[k];
      w[4 * k + 0] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
    }
  }
  uint32_t sr = 0, sq = 0, sx = 0, emask = 0;
  const uint32_t min_w = (uint32_t)p.min_len << 4;
#pragma unroll
  for (int s = 0; s < SCAN_VPT; ++s) {
    op_contrib<MODE>(w[s], sr, sq, sx);
    bool cand = (((w[s] & 15u) - 1u) < 2u) && (w[s] >= min_w);
    emask |= cand ? (1u << s) : 0u;
  }
  if (emask) {  // rare: apply the record gate (mapq / enable) now so counts are final
    uint32_t m = emask;
    while (m) {
      int s = __ffs(m) - 1;
      m &= m - 1;
      int32_t r = rec_of_op(p.op_off, rfirst, rlast, tile_start + tid * SCAN_VPT + s);
      bool ok = (int)__ldg(p.mapq + r) >= p.min_mapq && (p.enable == nullptr || __ldg(p.enable + r) != 0);
      if (!ok) emask &= ~(1u << s);
    }
  }
  uint32_t cnt = __popc(emask);

  // ---- block exclusive scan of (ref, qry, aux, cnt) -------------------------------
  uint32_t ir = sr, iq = sq, ix = sx, ic = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t a = __shfl_up_sync(0xffffffffu, ir, d);
    uint32_t b = __shfl_up_sync(0xffffffffu, iq, d);
    uint32_t c = __shfl_up_sync(0xffffffffu, ix, d);
    uint32_t e = __shfl_up_sync(0xffffffffu, ic, d);
    if (lane >= d) { ir += a; iq += b; ix += c; ic += e; }
  }
  if (lane == 31) { s_w[0][warp] = ir; s_w[1][warp] = iq; s_w[2][warp] = ix; s_w[3][warp] = ic; }
  __syncthreads();
  uint32_t wr = 0, wq = 0, wx = 0, wc = 0, tr = 0, tq = 0, tx = 0, tc = 0;
#pragma unroll
  for (int k = 0; k < SCAN_THREADS / 32; ++k) {
    uint32_t a = s_w[0][k], b = s_w[1][k], c = s_w[2][k], e = s_w[3][k];
    if (k < warp) { wr += a; wq += b; wx += c; wc += e; }
    tr += a; tq += b; tx += c; tc += e;
  }
  const uint32_t er = wr + ir - sr, eq = wq + iq - sq, ex = wx + ix - sx;  // thread exclusive
  const uint32_t ec = wc + ic - cnt;
  s_tx[0][tid] = er; s_tx[1][tid] = eq; s_tx[2][tid] = ex;
  if (tid == 0) { s_tot[0] = tr; s_tot[1] = tq; s_tot[2] = tx; s_tot[3] = tc; }
  __syncthreads();

  // ---- warp 0: publish the tile aggregate, then decoupled look-back ----------------
  if (warp == 0) {
    const int64_t last_start = __ldg(p.op_off + rlast);
    const bool bound = last_start >= tile_start;  // a record starts inside this tile
    uint32_t lr = tr, lq = tq;
    if (bound) {
      uint32_t br, bq, bx;
      prefix_at<MODE>(s_ops, s_tx, s_tot, (int)(last_start - tile_start), br, bq, bx);
      lr = tr - br;
      lq = tq - bq;
    }
    const uint32_t fl = bound ? ST_BOUND : 0u;
    uint32_t cr = 0, cq = 0, cc = 0;
    if (tile == 0) {
      if (lane == 0) {
        p.incl[0] = make_uint4(lr, lq, tc, fl);
        st_release_u32(p.status, ST_INCL | fl);
      }
    } else {
      if (lane == 0) {
        p.agg[tile] = make_uint4(lr, lq, tc, fl);
        st_release_u32(p.status + tile, ST_AGG | fl);
      }
      bool ref_open = true;
      int32_t pred = tile - 1 - lane;
      while (true) {
        uint32_t st = ST_INCL | ST_BOUND;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (pred >= 0) {
          do { st = ld_acquire_u32(p.status + pred); } while (st == 0u);
          v = ld_cg_u4((st & ST_INCL) ? (p.incl + pred) : (p.agg + pred));
        }
        const bool is_incl = (st & ST_INCL) != 0u;
        const bool closes = is_incl || (st & ST_BOUND) != 0u;
        const uint32_t incl_mask = __ballot_sync(0xffffffffu, is_incl);
        const uint32_t close_mask = __ballot_sync(0xffffffffu, closes);
        const int first_incl = incl_mask ? (__ffs(incl_mask) - 1) : 32;
        const int first_close = close_mask ? (__ffs(close_mask) - 1) : 32;
        uint32_t ac = (lane <= first_incl) ? v.z : 0u;
        const bool take = ref_open && lane <= first_close;
        uint32_t ar = take ? v.x : 0u, aq = take ? v.y : 0u;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
          ac += __shfl_xor_sync(0xffffffffu, ac, d);
          ar += __shfl_xor_sync(0xffffffffu, ar, d);
          aq += __shfl_xor_sync(0xffffffffu, aq, d);
        }
        cc += ac; cr += ar; cq += aq;
        if (first_close < 32) ref_open = false;
        if (first_incl < 32) break;
        pred -= 32;
      }
      if (lane == 0) {
        p.incl[tile] = make_uint4(bound ? lr : cr + tr, bound ? lq : cq + tq, cc + tc, fl);
        st_release_u32(p.status + tile, ST_INCL | fl);
      }
    }
    if (lane == 0) {
      s_b[1] = cr; s_b[2] = cq; s_b[3] = cc;
      if (tile == p.n_tiles - 1) *p.n_out = (int64_t)cc + tc;
    }
  } else if (p.ref_total != nullptr) {
    // ---- warps 1..7: per-record totals of the records that touch this tile ---------
    const int64_t tile_end = tile_start + SCAN_TILE;
    for (int32_t r = rfirst + (tid - 32); r <= rlast; r += SCAN_THREADS - 32) {
      int64_t a = __ldg(p.op_off + r), b = __ldg(p.op_off + r + 1);
      int j0 = (int)(max(a, tile_start) - tile_start);
      int j1 = (int)(min(b, tile_end) - tile_start);
      uint32_t r0, q0, x0, r1, q1, x1;
      prefix_at<MODE>(s_ops, s_tx, s_tot, j0, r0, q0, x0);
      prefix_at<MODE>(s_ops, s_tx, s_tot, j1, r1, q1, x1);
      if (a >= tile_start && b <= tile_end) {
        p.ref_total[r] = r1 - r0; p.qry_total[r] = q1 - q0; p.aux_total[r] = x1 - x0;
      } else {
        atomicAdd(p.ref_total + r, r1 - r0);
        atomicAdd(p.qry_total + r, q1 - q0);
        atomicAdd(p.aux_total + r, x1 - x0);
      }
    }
  }
  __syncthreads();

  // ---- pass 2: ordered emission ----------------------------------------------------
  if (emask) {
    const uint32_t carry_r = s_b[1], carry_q = s_b[2];
    int64_t out_idx = (int64_t)s_b[3] + ec;
    uint32_t pr = er, pq = eq, px = 0;
#pragma unroll
    for (int s = 0; s < SCAN_VPT; ++s) {
      if ((emask >> s) & 1u) {
        emit_sig<MODE>(p, s_ops, s_tx, s_tot, tile_start, rfirst, rlast, tid * SCAN_VPT + s, pr, pq, w[s],
                       carry_r, carry_q, out_idx);
        ++out_idx;
      }
      op_contrib<MODE>(w[s], pr, pq, px);
    }
  }
}

// --------------------------------------------------------------------------------
extern "C" int vsv_scan_cigar(vsv_ctx* ctx, int mode, int min_len, int min_mapq, const uint32_t* ops,
                              const int64_t* op_off, const int32_t* pos, const uint8_t* mapq,
                              const uint8_t* enable, int64_t n_rec, vsv_sig* out, int64_t cap,
                              int64_t* n_out, int32_t* ref_total, int32_t* qry_total,
                              int32_t* aux_total) {
  VSV_TRY(vsv_begin(ctx));
  if (mode < 0 || mode > 2 || n_rec < 0 || cap < 0 || !n_out || (n_rec > 0 && (!op_off || !pos || !mapq)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: bad argument");
  if (min_len < 0 || min_len >= (1 << 27)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: min_len out of range");
  if (n_rec >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: more than 2^31-1 records per call");
  const bool want_tot = ref_total || qry_total || aux_total;
  if (want_tot && !(ref_total && qry_total && aux_total))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: pass all three totals or none");
  *n_out = 0;
  if (n_rec == 0) return VSV_OK;

  // number of ops = op_off[n_rec] (host knows it in host-IO mode; fetch it in device-IO mode)
  int64_t n_ops = 0;
  if (ctx->device_io) {
    VSV_CUDA(ctx, cudaMemcpyAsync(&n_ops, op_off + n_rec, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    n_ops = op_off[n_rec];
  }
  if (n_ops < 0) VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: negative op count");
  const int64_t n_tiles64 = vsv_cdiv(n_ops, SCAN_TILE);
  if (n_tiles64 >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: too many ops per call");
  const int32_t n_tiles = (int32_t)n_tiles64;

  ScanParams p;
  memset(&p, 0, sizeof(p));
  VSV_TRY(vsv_in(ctx, ops, (size_t)n_ops, &p.ops));
  VSV_TRY(vsv_in(ctx, op_off, (size_t)n_rec + 1, &p.op_off));
  VSV_TRY(vsv_in(ctx, pos, (size_t)n_rec, &p.pos));
  VSV_TRY(vsv_in(ctx, mapq, (size_t)n_rec, &p.mapq));
  VSV_TRY(vsv_in(ctx, enable, (size_t)n_rec, &p.enable));
  if (n_ops > 0 && (reinterpret_cast<uintptr_t>(p.ops) & 15u))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: ops must be 16-byte aligned");
  vsv_sig* d_out = nullptr;
  VSV_TRY(vsv_out(ctx, out, (size_t)cap, &d_out));
  if (cap > 0 && (reinterpret_cast<uintptr_t>(d_out) & 15u))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_scan_cigar: out must be 16-byte aligned");
  int32_t *d_rt = nullptr, *d_qt = nullptr, *d_xt = nullptr;
  if (want_tot) {
    VSV_TRY(vsv_out(ctx, ref_total, (size_t)n_rec, &d_rt));
    VSV_TRY(vsv_out(ctx, qry_total, (size_t)n_rec, &d_qt));
    VSV_TRY(vsv_out(ctx, aux_total, (size_t)n_rec, &d_xt));
  }
  int64_t* d_nout = nullptr;
  VSV_TRY(vsv_tmp(ctx, 1, &d_nout));
  int32_t *d_rf = nullptr, *d_rl = nullptr;
  uint32_t *d_status = nullptr, *d_ticket = nullptr;
  uint4 *d_agg = nullptr, *d_incl = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_rf));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_rl));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles + 1, &d_status));  // [n_tiles] = ticket
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_agg));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_incl));
  d_ticket = d_status + n_tiles;

  p.n_rec = n_rec; p.n_ops = n_ops; p.n_tiles = n_tiles;
  p.min_len = min_len; p.min_mapq = min_mapq;
  p.tile_rfirst = d_rf; p.tile_rlast = d_rl;
  p.status = d_status; p.agg = d_agg; p.incl = d_incl; p.ticket = d_ticket;
  p.out = d_out; p.cap = cap; p.n_out = d_nout;
  p.ref_total = (uint32_t*)d_rt; p.qry_total = (uint32_t*)d_qt; p.aux_total = (uint32_t*)d_xt;

  VSV_TRY(vsv_t0(ctx));
  VSV_CUDA(ctx, cudaMemsetAsync(d_nout, 0, sizeof(int64_t), ctx->stream));
  if (n_tiles > 0) {
    VSV_CUDA(ctx, cudaMemsetAsync(d_status, 0, sizeof(uint32_t) * ((size_t)n_tiles + 1), ctx->stream));
    if (want_tot) {
      VSV_CUDA(ctx, cudaMemsetAsync(d_rt, 0, sizeof(int32_t) * (size_t)n_rec, ctx->stream));
      VSV_CUDA(ctx, cudaMemsetAsync(d_qt, 0, sizeof(int32_t) * (size_t)n_rec, ctx->stream));
      VSV_CUDA(ctx, cudaMemsetAsync(d_xt, 0, sizeof(int32_t) * (size_t)n_rec, ctx->stream));
    }
    VSV_LAUNCH(ctx, k_tile_records, (unsigned)vsv_cdiv(n_tiles, 256), 256, 0, p.op_off, n_rec, n_ops, n_tiles, d_rf,
               d_rl);
    VSV_CUDA(ctx, cudaEventRecord(ctx->evk0, ctx->stream));
    switch (mode) {
      case VSV_MODE_CONTIG: VSV_LAUNCH(ctx, k_scan_cigar<VSV_MODE_CONTIG>, (unsigned)n_tiles, SCAN_THREADS, 0, p); break;
      case VSV_MODE_READS: VSV_LAUNCH(ctx, k_scan_cigar<VSV_MODE_READS>, (unsigned)n_tiles, SCAN_THREADS, 0, p); break;
      default: VSV_LAUNCH(ctx, k_scan_cigar<VSV_MODE_CUTE>, (unsigned)n_tiles, SCAN_THREADS, 0, p); break;
    }
    VSV_CUDA(ctx, cudaEventRecord(ctx->evk1, ctx->stream));
    ctx->main_open = true;
  } else if (want_tot) {
    VSV_CUDA(ctx, cudaMemsetAsync(d_rt, 0, sizeof(int32_t) * (size_t)n_rec, ctx->stream));
    VSV_CUDA(ctx, cudaMemsetAsync(d_qt, 0, sizeof(int32_t) * (size_t)n_rec, ctx->stream));
    VSV_CUDA(ctx, cudaMemsetAsync(d_xt, 0, sizeof(int32_t) * (size_t)n_rec, ctx->stream));
  }
  VSV_TRY(vsv_t1(ctx));
  // the count is needed on the host to size the copy-out: one small sync
  VSV_CUDA(ctx, cudaMemcpyAsync(n_out, d_nout, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (want_tot) {
    VSV_TRY(vsv_out_done(ctx, ref_total, d_rt, (size_t)n_rec));
    VSV_TRY(vsv_out_done(ctx, qry_total, d_qt, (size_t)n_rec));
    VSV_TRY(vsv_out_done(ctx, aux_total, d_xt, (size_t)n_rec));
  }
  if (*n_out > cap) {
    vsv_end(ctx, true);
    VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_scan_cigar: output capacity too small");
  }
  VSV_TRY(vsv_out_done(ctx, out, d_out, (size_t)*n_out));
  return vsv_end(ctx);
}

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
