// Stage 1: CIGAR signature scan (K1).
//
// K1 replaces extract_sig_from_cigar (LI/extract_contig_signature_Hifi.py:53-85,
// LI/extract_reads_signature.py:47-83) and the CIGAR walk of parse_read
// (LI/sig_extract.py:438-478).  VSV_MODE_FUSED runs the two READ scanners (ERS + SE)
// in ONE pass over the packed ops and fills two output streams.
//
// Design v6 (B200-first; the streaming kernel has NO block-level synchronisation):
//  * the packed ops of ALL records are one flat u32 array cut into "warp tiles" of 1024
//    ops (4 KB), independent of record boundaries.  Every WARP is an independent worker: it
//    pulls its warp tiles round-robin into its own double-buffered shared-memory slot with
//    1-D TMA bulk copies (cp.async.bulk + mbarrier), folds 32 consecutive ops per lane
//    (bank-conflict-free XOR-swizzled LDS.128) into two running sums (reference / query
//    advance: PRMT byte-table look-up on the ALU pipe, IMAD accumulate on the FMA pipe)
//    and the running maximum of (op & table_ID); lanes whose maximum reaches the length
//    threshold are resolved afterwards by the whole warp (LDS + ballot -> candidate mask).
//    The sums are warp-scanned with shuffles.  Rare op codes (N, P, H, ...) and record
//    starts are handled warp-cooperatively (REDUX) only by warps that saw one.
//  * what the streaming kernel leaves behind per warp tile: a 32-byte aggregate (sums since
//    the last record start inside the tile), the in-tile prefix at every record start
//    (`fstart`, one 32-byte row per record), and one 32-byte staging entry per RAW
//    candidate (D/I >= min_len: well under 1 % of the ops) holding the lane's exclusive
//    prefix.  Staging slots are reserved per warp in chunks (one global atomic per ~128
//    candidates).
//  * everything that is sparse runs afterwards in small, lane-dense kernels: k_gate (one
//    thread per warp tile: record lookup, mapq/enable gate, exact in-record offsets),
//    a segmented exclusive scan over the tile aggregates (CUB, n_ops/1024 elements),
//    k_emit (final, deterministic (record, op)-ordered slots + carry of a record that
//    began in an earlier tile) and k_rec_totals (reference_end / read length per record).
//    The ops are read from HBM exactly once.
//
// History: v1 (single-pass chained look-back) was bound by the tile chain at 7 % of HBM
// (profiles/r01a_scan_v1.md); v2-v4 (CTA tiles, candidates finished by warp 0 behind
// __syncthreads) were bound by that serial phase at 23-25 %; v5 (warp-independent, per-op
// candidate test) reached 52 %; v6 is ALU-issue bound at 66 % (profiles/r01c_scan_v6.md).
#pragma once
#include <stdlib.h>

#include <algorithm>
#include "vsv_common.cuh"

#define WT_OPS 1024           // ops per warp tile
#define WT_VPT 32             // ops per lane
#define SCAN_WARPS 14         // warps (independent workers) per CTA: 2 CTAs x 14 warps x 8 KB fill the 227 KB of an SM
#define SCAN_THREADS (SCAN_WARPS * 32)
#define SCAN_STAGES 2
#define SCAN_CTAS_PER_SM 2
#define SCAN_CHUNK 128        // staging slots reserved per global atomic

#define VSV_MODE_FUSED 3      // internal: READS (stream A) + CUTE (stream B) in one pass

#define BIT(c) (1u << (c))

template <int MODE>
struct ModeTraits;
// contig: only M,S,D,I are walked (ECS_H:72-85); H is needed for readlen
template <>
struct ModeTraits<VSV_MODE_CONTIG> {
  static constexpr uint32_t REF = BIT(0) | BIT(2);
  static constexpr uint32_t QRY = BIT(0) | BIT(1) | BIT(4);
  static constexpr uint32_t X1 = BIT(5), X2 = 0, X3 = 0;
  static constexpr int NX = 1;
};
// reads: M,=,X both; S,I query; D,N reference (ERS:66-81)
template <>
struct ModeTraits<VSV_MODE_READS> {
  static constexpr uint32_t REF = BIT(0) | BIT(7) | BIT(8) | BIT(2) | BIT(3);
  static constexpr uint32_t QRY = BIT(0) | BIT(7) | BIT(8) | BIT(4) | BIT(1);
  static constexpr uint32_t X1 = BIT(5), X2 = 0, X3 = 0;
  static constexpr int NX = 1;
};
// cute: shift_del == shift_ins == sum{M,D,=,X}; shift_ins_read = every op but D
// (SE:461-478); N moves neither reference counter (quirk).  x1 = N, x2 = other non-D ops.
template <>
struct ModeTraits<VSV_MODE_CUTE> {
  static constexpr uint32_t REF = BIT(0) | BIT(2) | BIT(7) | BIT(8);
  static constexpr uint32_t QRY = BIT(0) | BIT(7) | BIT(8) | BIT(1) | BIT(4);
  static constexpr uint32_t X1 = BIT(3), X2 = 0xFFFFu & ~(REF | QRY | BIT(3)), X3 = 0;
  static constexpr int NX = 2;
};
// fused READS + CUTE: r = {M,D,=,X}, q = {M,=,X,I,S}, x1 = N, x2 = H, x3 = P and codes 9..15
template <>
struct ModeTraits<VSV_MODE_FUSED> {
  static constexpr uint32_t REF = BIT(0) | BIT(2) | BIT(7) | BIT(8);
  static constexpr uint32_t QRY = BIT(0) | BIT(7) | BIT(8) | BIT(1) | BIT(4);
  static constexpr uint32_t X1 = BIT(3), X2 = BIT(5), X3 = 0xFFFFu & ~(REF | QRY | BIT(3) | BIT(5));
  static constexpr int NX = 3;
};

struct Sums {
  uint32_t r, q, x1, x2, x3;
};
__device__ __forceinline__ Sums operator-(const Sums& a, const Sums& b) {
  return Sums{a.r - b.r, a.q - b.q, a.x1 - b.x1, a.x2 - b.x2, a.x3 - b.x3};
}

// Per (mode, stream): the reference scanner's reference offset, query offset, the sign
// of the leading hard clip in q0 and the per-record totals (ref, qry, aux).
template <int MODE, int STREAM>
__device__ __forceinline__ uint32_t ref_of(const Sums& s) {
  if (MODE == VSV_MODE_FUSED && STREAM == 0) return s.r + s.x1;  // READS counts N on the reference
  return s.r;
}
template <int MODE, int STREAM>
__device__ __forceinline__ uint32_t qry_of(const Sums& s) {
  if (MODE == VSV_MODE_CUTE) return s.q + s.x1 + s.x2;
  if (MODE == VSV_MODE_FUSED && STREAM == 1) return s.q + s.x1 + s.x2 + s.x3;
  return s.q;
}
template <int MODE, int STREAM>
__device__ __forceinline__ int hsign() {
  return (MODE == VSV_MODE_CUTE || (MODE == VSV_MODE_FUSED && STREAM == 1)) ? -1 : 1;
}
template <int MODE>
__device__ __forceinline__ uint32_t aux_of(const Sums& s) {  // CONTIG/READS/FUSED: H; CUTE: N
  if (MODE == VSV_MODE_FUSED) return s.x2;
  return s.x1;
}

// One tile's aggregate / scanned state (32 bytes).
struct __align__(16) TileState {
  uint32_t r, q, x1, x2, x3;  // sums since the last record start (or since the segment start)
  uint32_t cnt_a, cnt_b;      // signatures of stream A / B
  uint32_t flag;              // a record starts inside (aggregate) / before here (scanned)
};
struct TileStateOp {
  __host__ __device__ TileState operator()(const TileState& a, const TileState& b) const {
    TileState o;
    o.cnt_a = a.cnt_a + b.cnt_a;
    o.cnt_b = a.cnt_b + b.cnt_b;
    if (b.flag) {
      o.r = b.r; o.q = b.q; o.x1 = b.x1; o.x2 = b.x2; o.x3 = b.x3; o.flag = 1u;
    } else {
      o.r = a.r + b.r; o.q = a.q + b.q; o.x1 = a.x1 + b.x1; o.x2 = a.x2 + b.x2; o.x3 = a.x3 + b.x3;
      o.flag = a.flag;
    }
    return o;
  }
};

// Staging entry of one raw candidate (32 bytes).  Written by the streaming kernel as
// {j, ex.r, ex.q, ex.x1, ex.x2, ex.x3, -, -} (j = op index in the warp tile, ex = the lane's
// exclusive prefix in the warp tile); rewritten in place by k_gate as
// {rec, w, part.r, part.q, part.x1, part.x2, part.x3, flags} (part = sums since the record
// start or since the tile start when the record began earlier; flags: 1 stream A, 2 stream
// B, 4 carry pending).
struct __align__(16) StageEntry {
  uint32_t a, b, r, q, x1, x2, x3, f;
};

struct ScanParams {
  const uint32_t* ops;
  const int64_t* op_off;
  const int32_t* pos;
  const uint32_t* rec_info;  // hc << 2 | gate_b << 1 | gate_a
  int64_t n_rec, n_ops;
  int32_t n_wt;              // number of warp tiles
  uint32_t min_w;            // min candidate length << 4
  uint32_t one;              // 1 (opaque multiplier: keeps an accumulate on the FMA pipe)
  uint32_t tab[8];           // PRMT byte tables by op code: REF lo/hi, QRY lo/hi, candidate/rare lo/hi (v6), I/D -1 | rare -128 lo/hi (v8)
  int32_t min_len_a, min_len_b;
  const int2* wt_starts;     // [first,last] record (index in 0..n_rec) whose start lies in the tile
  TileState* wt_agg;         // [n_wt + 1] aggregates (last = identity), counts filled by k_gate
  uint2* wt_stage;           // staging base / raw candidate count of each tile
  unsigned long long* stage_ctr;  // [0] staging allocation counter, [2] staging-full flag
  StageEntry* stage;
  uint32_t stage_cap;
  uint4* fstart;             // [n_rec + 1][2] in-tile prefix at each record start
};

// ---- PTX wrappers: mbarrier + 1-D TMA bulk copy (sm_90+/sm_100a) ------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// last record r in [lo,hi] with op_off[r] <= x  (op_off[lo] <= x assumed)
__device__ __forceinline__ int32_t rec_of_op(const int64_t* __restrict__ op_off, int32_t lo, int32_t hi, int64_t x) {
  while (lo < hi) {
    int32_t mid = (lo + hi + 1) >> 1;
    if (__ldg(op_off + mid) <= x) lo = mid; else hi = mid - 1;
  }
  return lo;
}


// ---- small pre-kernels ---------------------------------------------------------------
// per warp tile: the records (indices in 0..n_rec, n_rec = end sentinel) whose START offset
// lies inside the tile: first(t) = lower_bound(op_off, t * WT_OPS), last(t) = first(t + 1) - 1;
// first > last when no record starts here.  Scatter from the records: record r (op_off[-1] = -1)
// is first(t) for the tiles t in (op_off[r-1] / WT_OPS, op_off[r] / WT_OPS]; first[n_wt] is
// n_rec + 1 unless the ops end exactly on a tile edge.
static __global__ void k_wt_first(const int64_t* __restrict__ op_off, int64_t n_rec, int32_t n_wt, int32_t* __restrict__ first) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n_rec) return;
  const int64_t cur = __ldg(op_off + r);
  const int64_t prev = r > 0 ? __ldg(op_off + r - 1) : -1;
  int64_t t0 = prev < 0 ? 0 : prev / WT_OPS + 1;
  const int64_t t1 = min(cur / WT_OPS, (int64_t)n_wt);
  for (; t0 <= t1; ++t0) first[t0] = (int32_t)r;
  if (r == n_rec && cur % WT_OPS != 0) first[n_wt] = (int32_t)n_rec + 1;  // no offset reaches the edge of the last tile
}
static __global__ void k_wt_starts(const int32_t* __restrict__ first, int32_t n_wt, int2* __restrict__ wt_starts) {
  const int32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_wt) return;
  wt_starts[t] = make_int2(first[t], first[t + 1] - 1);
}

// per record: leading hard clip and the record gates of both streams
static __global__ void k_rec_info(const uint32_t* __restrict__ ops, const int64_t* __restrict__ op_off,
                           const uint8_t* __restrict__ mapq, const uint8_t* __restrict__ enable_a,
                           const uint8_t* __restrict__ enable_b, int64_t n_rec, int min_mapq_a, int min_mapq_b,
                           uint32_t* __restrict__ info) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rec) return;
  const int64_t a = op_off[r], b = op_off[r + 1];
  uint32_t hc = 0;
  if (b > a) {
    const uint32_t f = __ldg(ops + a);
    if ((f & 15u) == 5u) hc = f >> 4;
  }
  const int mq = mapq[r];
  const uint32_t ga = (mq >= min_mapq_a && (enable_a == nullptr || enable_a[r] != 0)) ? 1u : 0u;
  const uint32_t gb = (mq >= min_mapq_b && (enable_b == nullptr || enable_b[r] != 0)) ? 2u : 0u;
  info[r] = (hc << 2) | gb | ga;
}

template <int MODE>
__device__ __forceinline__ void add_op(uint32_t w, Sums& s) {
  typedef ModeTraits<MODE> M;
  const uint32_t bit = 1u << (w & 15u), len = w >> 4;
  s.r += (bit & M::REF) ? len : 0u;
  s.q += (bit & M::QRY) ? len : 0u;
  if (bit & (M::X1 | M::X2 | M::X3)) {   // N / H / P / reserved codes: a branch that is almost never taken (k_gate spends its
    s.x1 += (bit & M::X1) ? len : 0u;    // issue slots in 32 of these per candidate)
    if (M::NX > 1) s.x2 += (bit & M::X2) ? len : 0u;
    if (M::NX > 2) s.x3 += (bit & M::X3) ? len : 0u;
  }
}

__device__ __forceinline__ void store_sums(uint4* dst, const Sums& s) {
  dst[0] = make_uint4(s.r, s.q, s.x1, s.x2);
  dst[1] = make_uint4(s.x3, 0u, 0u, 0u);
}
__device__ __forceinline__ Sums load_sums(const uint4* src) {
  const uint4 a = __ldg(src), b = __ldg(src + 1);
  return Sums{a.x, a.y, a.z, a.w, b.x};
}

// byte table for PRMT look-ups by op code: byte k (codes base..base+3) = v1 if code in set1,
// else v2 if code in set2, else 0
__host__ __device__ constexpr uint32_t tab_word(uint32_t set1, int base, uint32_t v1, uint32_t set2, uint32_t v2) {
  uint32_t wv = 0;
  for (int k = 0; k < 4; ++k) {
    const uint32_t c = (uint32_t)(base + k);
    const uint32_t b = ((set1 >> c) & 1u) ? v1 : (((set2 >> c) & 1u) ? v2 : 0u);
    wv |= b << (8 * k);
  }
  return wv;
}
__device__ __forceinline__ uint32_t prmt(uint32_t lo, uint32_t hi, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(lo), "r"(hi), "r"(sel));
  return d;
}

// CTAs of the streaming kernel: one per resident slot.  VSV_SCAN_GRID caps it (tests: many tiles per warp from a small input,
// which is what exercises the byte-path / v6 switching of a warp over consecutive tiles).
static inline int vsv_scan_grid(const vsv_ctx* ctx, int64_t n_wt) {
  int64_t g = std::max<int64_t>(1, std::min<int64_t>(vsv_cdiv(n_wt, SCAN_WARPS), (int64_t)ctx->sm_count * SCAN_CTAS_PER_SM));
  if (const char* e = getenv("VSV_SCAN_GRID")) {
    const long v = strtol(e, nullptr, 10);
    if (v > 0) g = std::min<int64_t>(g, v);
  }
  return (int)g;
}

// one step of three interleaved inclusive warp scans: the shuffle's own predicate (source lane in range) guards the adds
__device__ __forceinline__ void scan3_step(uint32_t& a, uint32_t& b, uint32_t& c, int d) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b32 t0, t1, t2;\n"
      "shfl.sync.up.b32 t0|p, %0, %3, 0, 0xffffffff;\n"
      "shfl.sync.up.b32 t1, %1, %3, 0, 0xffffffff;\n"
      "shfl.sync.up.b32 t2, %2, %3, 0, 0xffffffff;\n"
      "@p add.u32 %0, %0, t0;\n"
      "@p add.u32 %1, %1, t1;\n"
      "@p add.u32 %2, %2, t2;\n"
      "}\n"
      : "+r"(a), "+r"(b), "+r"(c)
      : "r"(d));
}

// d = c + sum of (unsigned byte i of a) * (signed byte i of b)
__device__ __forceinline__ uint32_t dp4a_us(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// ---- K1: the streaming kernel (every warp is an independent worker) --------------------
template <int MODE>
static __global__ void __launch_bounds__(SCAN_THREADS, SCAN_CTAS_PER_SM) k_scan_wt(const ScanParams p) {
  typedef ModeTraits<MODE> M;
  constexpr uint32_t RARE = M::X1 | M::X2 | M::X3;
  constexpr uint32_t FULL = 0xffffffffu;
  extern __shared__ __align__(128) unsigned char s_dyn[];
  uint32_t(*s_ops)[SCAN_STAGES][WT_OPS] = reinterpret_cast<uint32_t(*)[SCAN_STAGES][WT_OPS]>(s_dyn);
  uint64_t(*s_bar)[SCAN_STAGES] =
      reinterpret_cast<uint64_t(*)[SCAN_STAGES]>(s_dyn + (size_t)SCAN_WARPS * SCAN_STAGES * WT_OPS * sizeof(uint32_t));
  uint32_t(*s_tab)[8] = reinterpret_cast<uint32_t(*)[8]>(s_dyn + (size_t)SCAN_WARPS * SCAN_STAGES * (WT_OPS * sizeof(uint32_t) + sizeof(uint64_t)));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // tile indices are 32-bit (n_wt < 2^31 - 1, checked by the host; wt + 2 * TW cannot wrap in u32)
  const uint32_t TW = gridDim.x * SCAN_WARPS;
  const uint32_t gw = blockIdx.x * SCAN_WARPS + warp;
  const uint32_t n_wt = (uint32_t)p.n_wt;
  const uint32_t n_full = (uint32_t)(p.n_ops / WT_OPS);  // tiles below n_full are complete

  if (lane == 0) {
    for (int s = 0; s < SCAN_STAGES; ++s) mbar_init(&s_bar[warp][s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncwarp();
  if (lane == 0) {
    for (int s = 0; s < SCAN_STAGES; ++s) {
      const uint32_t t = gw + (uint32_t)s * TW;
      if (t < n_full) {
        mbar_expect_tx(&s_bar[warp][s], WT_OPS * 4);
        tma_load_1d(s_ops[warp][s], p.ops + (size_t)t * WT_OPS, WT_OPS * 4, &s_bar[warp][s]);
      }
    }
  }
  uint32_t tabs[8];
  if (lane < 8) s_tab[warp][lane] = p.tab[lane];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 8; ++k) tabs[k] = *reinterpret_cast<volatile uint32_t*>(&s_tab[warp][k]);
  int classic_left = 0;  // > 0: this warp runs the v6 loop for that many more tiles before probing the byte path again
  int2 cur = make_int2(0, -1), nxt = make_int2(0, -1);
  if (gw < n_wt) cur = __ldg(p.wt_starts + gw);
  if (gw + TW < n_wt) nxt = __ldg(p.wt_starts + gw + TW);
  unsigned long long ch_next = 0, ch_end = 0;  // this warp's staging chunk (lane 0)
  // record-start offsets of the CURRENT tile, loaded one tile ahead (the load's round trip hides behind a whole tile)
  int64_t off_cur = 0;
  if (lane <= cur.y - cur.x) off_cur = __ldg(p.op_off + cur.x + lane);

  int it = 0;
  for (uint32_t wt = gw; wt < n_wt; wt += TW, ++it) {
    const int stage = it % SCAN_STAGES;
    const uint32_t parity = (uint32_t)(it / SCAN_STAGES) & 1u;
    const int64_t wt_start = (int64_t)wt * WT_OPS;
    uint32_t* tl = s_ops[warp][stage];

    // record starts inside this tile (offsets prefetched during the previous tile); next tile's offsets are issued here
    const int nst = cur.y - cur.x + 1;
    const int32_t sx = lane < nst ? (int32_t)(off_cur - wt_start) : 0;
    int64_t off_nxt = 0;
    if (lane <= nxt.y - nxt.x) off_nxt = __ldg(p.op_off + nxt.x + lane);
    int2 nn = make_int2(0, -1);
    if (wt + 2 * TW < n_wt) nn = __ldg(p.wt_starts + wt + 2 * TW);

    if (wt < n_full) {
      mbar_wait(&s_bar[warp][stage], parity);
    } else {  // the one partial tile: guarded copy, zero filled (op 0 = "0M": contributes nothing)
      const int64_t rem = p.n_ops - wt_start;
      for (int i = lane; i < WT_OPS; i += 32) tl[i] = (i < rem) ? __ldg(p.ops + wt_start + i) : 0u;
      __syncwarp();
    }

    // -- fast path: 32 consecutive ops per lane, rotated so LDS.128 is conflict free -------
    // Per-op budget (v6).  The integer ALU pipe (LOP3/SHF/PRMT/VIMNMX) and the FMA pipe (IMAD)
    // each issue one warp instruction every two cycles per SM sub-partition, so the per-op work
    // is balanced across both: three byte-table look-ups by op code (PRMT: 0xFFFFFFFF = member,
    // 0 = not) feed IMADs that accumulate the NEGATED reference / query sums; the third table
    // marks I/D (-1) and rare codes (0x01010101) and is summed by an IMAD with an opaque 1.
    // Candidates (D/I >= min_len: < 1 % of the ops) are NOT located per op: the loop only keeps
    // max(w & tc), and lanes whose maximum reaches min_w are resolved afterwards by the whole
    // warp (one LDS + ballot per such lane).  Lanes that saw a rare code (N/H/P, reserved 9..15)
    // are recomputed exactly, also warp-cooperatively (REDUX).
    uint32_t r = 0, q = 0;
    Sums xs = {0, 0, 0, 0, 0};
    bool lane_rare, lane_cand;
    const uint32_t min_w = p.min_w;
    if constexpr (MODE == VSV_MODE_CONTIG) {  // contig BAMs are tiny: generic per-op path
      uint32_t bits = 0, mxk = 0;
      const int rot = lane & 7;
      const uint4* row = reinterpret_cast<const uint4*>(tl + lane * WT_VPT);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint4 v = row[k ^ rot];
        const uint32_t ww[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t w = ww[i];
          const uint32_t bit = 1u << (w & 15u), len = w >> 4;
          bits |= bit;
          r += (bit & M::REF) ? len : 0u;
          q += (bit & M::QRY) ? len : 0u;
          mxk = max(mxk, (bit & 6u) ? w : 0u);
        }
      }
      lane_rare = (bits & RARE) != 0u;
      lane_cand = mxk >= min_w;
    } else {
      // the six table words (filled by the host from tab_word) were bounced through shared memory
      // before the tile loop: as compile-time constants ptxas re-materialises them around every
      // PRMT, and as kernel parameters it keeps them in UNIFORM registers, which PRMT cannot read
      // (one IMAD.U32 copy per look-up) - a value loaded from shared memory stays in a plain register
      const uint32_t TR_LO = tabs[0], TR_HI = tabs[1], TQ_LO = tabs[2], TQ_HI = tabs[3], TC_LO = tabs[4], TC_HI = tabs[5];
      static_assert(((M::REF >> 8) & 1u) == (M::REF & 1u) && ((M::QRY >> 8) & 1u) == (M::QRY & 1u),
                    "code 8 (X) is looked up as the sign of code 0 (M)");
      // 128-byte rows: physical uint4 (k ^ rot) of the row = byte address (row ^ rot*16) ^ k*16
      const uint32_t row_x = (smem_u32(tl) + (uint32_t)lane * (WT_VPT * 4u)) ^ ((uint32_t)(lane & 7) << 4);
      const uint32_t one = p.one;  // == 1, opaque to the compiler: keeps `sc` on the FMA pipe
      uint32_t nr = 0, nq = 0;
      int32_t mnk = 0;
      // -- v8 byte path.  When every op of the tile has a code <= 7 and a length < 256 (M-alphabet long reads: all but a
      // few ops) the op word itself is the PRMT selector (its low nibble is the code; the other result bytes are garbage)
      // and dp4a(len, table bytes) multiplies the length, whose upper three bytes are zero, by the signed table byte of
      // the code: no `code`, no selector, no mask widening -- shr + 3 prmt + 3 dp4a + min per op instead of 12
      // instructions (tools/pipe_bench.cu: 5.8 vs 14.4 cycles per op slot).  The third table holds -1 for I/D and -128
      // for N/H/P, so one running minimum finds both the lanes with a candidate (<= -min_len) and the lanes that need the
      // exact recomputation (<= -128: a rare code, or an I/D of 128+ bases).  The OR of the lane's words tells whether
      // the assumptions held: a length >= 256 sends the lane to the exact path; a code >= 8 (X of an =/X alphabet, or
      // a reserved code), or more than six such lanes, sends the whole tile through the v6 loop below, and the warp
      // stays on v6 for the next 31 tiles before it probes again (=/X or HiFi-like files cost one probe in 32).
      bool classic = classic_left > 0;
#if !defined(VSV_ABLATE_LOOP) && !defined(VSV_ABLATE_DETECT) && !defined(VSV_NO_BYTE_PATH)
      if (!classic) {
        const uint32_t TF_LO = tabs[6], TF_HI = tabs[7];
        uint32_t orw = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          uint4 v;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(row_x ^ (uint32_t)(k << 4)));
          const uint32_t ww[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t w = ww[i];
            const uint32_t len = w >> 4;
            const uint32_t tr = prmt(TR_LO, TR_HI, w), tq = prmt(TQ_LO, TQ_HI, w), tf = prmt(TF_LO, TF_HI, w);
            nr = dp4a_us(len, tr, nr);
            nq = dp4a_us(len, tq, nq);
            mnk = min(mnk, (int32_t)dp4a_us(len, tf, 0u));
          }
          orw |= (v.x | v.y) | (v.z | v.w);
        }
        lane_rare = orw >= 4096u || mnk <= -128;
        const uint32_t xb = __ballot_sync(FULL, (orw & 8u) != 0u), rb = __ballot_sync(FULL, lane_rare);
        if (xb != 0u || __popc(rb) > 6) {
          classic = true;
          classic_left = 32;
          nr = 0; nq = 0; mnk = 0;
        }
      }
#else
      classic = true;
#endif
      if (classic) {
        --classic_left;
        uint32_t mxc = 0, sc = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          uint4 v;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(row_x ^ (uint32_t)(k << 4)));
          const uint32_t ww[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t w = ww[i];
            const uint32_t code = w & 15u, len = w >> 4;
            const uint32_t sel = code * 0x1111u;
#if defined(VSV_ABLATE_LOOP)   // profiling builds only (profiles/r02s_scan.md): stream + scans, no per-op work
            nr += w;
            (void)code; (void)len; (void)sel; (void)one;
#else
            const uint32_t tr = prmt(TR_LO, TR_HI, sel), tq = prmt(TQ_LO, TQ_HI, sel);
            nr = len * tr + nr;  // IMAD: subtracts len when the code is a member
            nq = len * tq + nq;
#if !defined(VSV_ABLATE_DETECT)  // profiling builds only: the two sums without candidate / rare-code detection
            const uint32_t tc = prmt(TC_LO, TC_HI, sel);
            sc = tc * one + sc;  // I/D: -1 each, rare: +0x01010101 each
            mnk = min(mnk, (int32_t)(len * tc));  // I/D: -len (an IMAD: the ALU pipe is the busier one)
            mxc = max(mxc, code);
#endif
#endif
          }
        }
        lane_rare = (int32_t)sc > 0 || mxc > 8u;  // a rare code (N/H/P...) or a reserved one (9..15)
      }
      r = 0u - nr;
      q = 0u - nq;
      lane_cand = mnk <= -(int32_t)(min_w >> 4) || lane_rare;  // rare lanes: mnk is not trustworthy
    }
    // -- rare lanes: exact sums of every class, by the whole warp ---------------------------------
    uint32_t rare_lanes = __ballot_sync(FULL, lane_rare);
    const bool anyrare = rare_lanes != 0u;
    while (rare_lanes) {
      const int L = __ffs(rare_lanes) - 1;
      rare_lanes &= rare_lanes - 1;
      Sums o = {0, 0, 0, 0, 0};
      add_op<MODE>(tl[L * WT_VPT + lane], o);
      o.r = __reduce_add_sync(FULL, o.r); o.q = __reduce_add_sync(FULL, o.q);
      o.x1 = __reduce_add_sync(FULL, o.x1);
      if (M::NX > 1) o.x2 = __reduce_add_sync(FULL, o.x2);
      if (M::NX > 2) o.x3 = __reduce_add_sync(FULL, o.x3);
      if (lane == L) { xs = o; r = o.r; q = o.q; }
    }
    // -- candidate lanes: which of the lane's 32 ops are D/I >= min_len ---------------------------
    uint32_t em = 0;
    {
      uint32_t cand_lanes = __ballot_sync(FULL, lane_cand);
      while (cand_lanes) {  // two candidate lanes per round: the two LDS are in flight together
        const int L0 = __ffs(cand_lanes) - 1;
        cand_lanes &= cand_lanes - 1;
        const int L1 = cand_lanes ? __ffs(cand_lanes) - 1 : L0;
        cand_lanes &= cand_lanes - 1;
        const uint32_t w0 = tl[L0 * WT_VPT + lane], w1 = tl[L1 * WT_VPT + lane];
        const uint32_t m0 = __ballot_sync(FULL, ((1u << (w0 & 15u)) & 6u) && w0 >= min_w);
        const uint32_t m1 = __ballot_sync(FULL, ((1u << (w1 & 15u)) & 6u) && w1 >= min_w);
        if (lane == L0) em = m0;
        if (lane == L1) em = m1;
      }
    }
    const uint32_t c = __popc(em);

    // -- warp scans (inclusive) ---------------------------------------------------------------
    uint32_t ir = r, iq = q, ic = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) scan3_step(ir, iq, ic, d);
    Sums ex = {ir - r, iq - q, 0, 0, 0};  // this lane's exclusive prefix inside the tile
    Sums tot = {__shfl_sync(FULL, ir, 31), __shfl_sync(FULL, iq, 31), 0, 0, 0};
    const uint32_t myoff = ic - c, coff = __shfl_sync(FULL, ic, 31);
    if (anyrare) {  // this warp saw N / H / P / ...: scan the rare sums too
      uint32_t i1 = xs.x1, i2 = xs.x2, i3 = xs.x3;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) scan3_step(i1, i2, i3, d);
      ex.x1 = i1 - xs.x1; ex.x2 = i2 - xs.x2; ex.x3 = i3 - xs.x3;
      tot.x1 = __shfl_sync(FULL, i1, 31); tot.x2 = __shfl_sync(FULL, i2, 31); tot.x3 = __shfl_sync(FULL, i3, 31);
    }

    // -- record starts in this tile: in-tile prefix at each start -> fstart[rec] -------------
    // start s begins at op x of the tile: prefix = exclusive prefix of lane x>>5 + the first x&31
    // ops of that lane's row, summed by the whole warp (one LDS + REDUX per class).
    Sums f_last = {0, 0, 0, 0, 0};
    for (int s = 0; s < nst; ++s) {
      int32_t x = (s < 32) ? __shfl_sync(FULL, sx, s) : (int32_t)(__ldg(p.op_off + cur.x + s) - wt_start);
      const int t = x >> 5, n = x & 31;
      Sums f;
      f.r = __shfl_sync(FULL, ex.r, t); f.q = __shfl_sync(FULL, ex.q, t);
      Sums o = {0, 0, 0, 0, 0};
      if (lane < n) add_op<MODE>(tl[t * WT_VPT + lane], o);
      f.r += __reduce_add_sync(FULL, o.r); f.q += __reduce_add_sync(FULL, o.q);
      f.x1 = 0; f.x2 = 0; f.x3 = 0;
      if (anyrare) {
        f.x1 = __shfl_sync(FULL, ex.x1, t) + __reduce_add_sync(FULL, o.x1);
        if (M::NX > 1) f.x2 = __shfl_sync(FULL, ex.x2, t) + __reduce_add_sync(FULL, o.x2);
        if (M::NX > 2) f.x3 = __shfl_sync(FULL, ex.x3, t) + __reduce_add_sync(FULL, o.x3);
      }
      if (lane == (s & 31)) store_sums(p.fstart + 2 * ((int64_t)cur.x + s), f);
      f_last = f;
    }

    // -- raw candidates -> staging entries (op order), chunked slot reservation ---------------
    const uint32_t ncand = coff;
    unsigned long long base = ~0ull;
    if (ncand) {
      if (lane == 0) {
        if (ch_next + ncand > ch_end) {
          const unsigned long long n = ncand > SCAN_CHUNK ? ncand : SCAN_CHUNK;
          ch_next = atomicAdd(p.stage_ctr, n);
          ch_end = ch_next + n;
        }
        base = ch_next;
        ch_next += ncand;
        if (base + ncand > p.stage_cap) {  // staging full: the host retries with a larger one
          base = ~0ull;
          p.stage_ctr[2] = 1ull;
        }
      }
      base = __shfl_sync(FULL, base, 0);
      if (em && base != ~0ull) {
        uint4* dst = reinterpret_cast<uint4*>(p.stage + base + myoff);
        uint32_t m = em;
        while (m) {
          const int s = __ffs(m) - 1;
          m &= m - 1;
          dst[0] = make_uint4((uint32_t)(lane * WT_VPT + s), (uint32_t)wt, ex.r, ex.q);
          dst[1] = make_uint4(ex.x1, ex.x2, ex.x3, 0u);
          dst += 2;
        }
      }
    }
    if (lane == 0) {
      Sums a = tot;
      if (nst > 0) a = a - f_last;
      TileState ts;
      ts.r = a.r; ts.q = a.q; ts.x1 = a.x1; ts.x2 = a.x2; ts.x3 = a.x3;
      ts.cnt_a = 0; ts.cnt_b = 0; ts.flag = nst > 0 ? 1u : 0u;
      p.wt_agg[wt] = ts;
      p.wt_stage[wt] = make_uint2(base == ~0ull ? 0xFFFFFFFFu : (uint32_t)base, ncand);
    }
    __syncwarp();  // every lane is done with this stage
    if (lane == 0) {
      const uint32_t t2 = wt + (uint32_t)SCAN_STAGES * TW;
      if (t2 < n_full) {
        fence_proxy_async();
        mbar_expect_tx(&s_bar[warp][stage], WT_OPS * 4);
        tma_load_1d(s_ops[warp][stage], p.ops + (size_t)t2 * WT_OPS, WT_OPS * 4, &s_bar[warp][stage]);
      }
    }
    cur = nxt;
    nxt = nn;
    off_cur = off_nxt;
  }
}

// ---- exclusive scan of the tile aggregates (segmented: a tile in which a record starts resets
// the sums; the signature counts are plain sums).  Three small launches over n_wt + 1 32-byte
// elements: per-CTA reduce, one-CTA spine, per-CTA scan with carry-in.
#define TS_THREADS 256
#define TS_ITEMS 4
#define TS_BLOCK (TS_THREADS * TS_ITEMS)
__device__ __forceinline__ TileState ts_identity() {
  TileState t;
  t.r = t.q = t.x1 = t.x2 = t.x3 = t.cnt_a = t.cnt_b = t.flag = 0u;
  return t;
}
__device__ __forceinline__ TileState ts_load(const TileState* p) {
  const uint4 a = reinterpret_cast<const uint4*>(p)[0], b = reinterpret_cast<const uint4*>(p)[1];
  TileState t;
  t.r = a.x; t.q = a.y; t.x1 = a.z; t.x2 = a.w; t.x3 = b.x; t.cnt_a = b.y; t.cnt_b = b.z; t.flag = b.w;
  return t;
}
__device__ __forceinline__ void ts_store(TileState* p, const TileState& t) {
  reinterpret_cast<uint4*>(p)[0] = make_uint4(t.r, t.q, t.x1, t.x2);
  reinterpret_cast<uint4*>(p)[1] = make_uint4(t.x3, t.cnt_a, t.cnt_b, t.flag);
}
__device__ __forceinline__ TileState ts_shfl_up(const TileState& t, int d) {
  TileState o;
  o.r = __shfl_up_sync(0xffffffffu, t.r, d); o.q = __shfl_up_sync(0xffffffffu, t.q, d);
  o.x1 = __shfl_up_sync(0xffffffffu, t.x1, d); o.x2 = __shfl_up_sync(0xffffffffu, t.x2, d);
  o.x3 = __shfl_up_sync(0xffffffffu, t.x3, d); o.cnt_a = __shfl_up_sync(0xffffffffu, t.cnt_a, d);
  o.cnt_b = __shfl_up_sync(0xffffffffu, t.cnt_b, d); o.flag = __shfl_up_sync(0xffffffffu, t.flag, d);
  return o;
}
// inclusive scan across the CTA of one value per thread; returns the thread's inclusive value and
// the CTA total in `total` (s_w: TS_THREADS / 32 slots)
__device__ __forceinline__ TileState ts_block_scan(TileState v, TileState* s_w, TileState& total) {
  const TileStateOp op;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const TileState u = ts_shfl_up(v, d);
    if (lane >= d) v = op(u, v);
  }
  if (lane == 31) s_w[warp] = v;
  __syncthreads();
  TileState pre = ts_identity();
  total = ts_identity();
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
    if (w == warp) pre = total;
    total = op(total, s_w[w]);
  }
  __syncthreads();
  return warp == 0 ? v : op(pre, v);
}
static __global__ void __launch_bounds__(TS_THREADS) k_ts_reduce(const TileState* __restrict__ in, int64_t n,
                                                            TileState* __restrict__ block_agg) {
  __shared__ TileState s_w[TS_THREADS / 32];
  const TileStateOp op;
  const int64_t base = (int64_t)blockIdx.x * TS_BLOCK + (int64_t)threadIdx.x * TS_ITEMS;
  TileState v = ts_identity();
#pragma unroll
  for (int k = 0; k < TS_ITEMS; ++k)
    if (base + k < n) v = op(v, ts_load(in + base + k));
  TileState total;
  ts_block_scan(v, s_w, total);
  if (threadIdx.x == 0) ts_store(block_agg + blockIdx.x, total);
}
static __global__ void __launch_bounds__(1024) k_ts_spine(TileState* __restrict__ block_agg, int64_t nb) {
  __shared__ TileState s_w[32];
  const TileStateOp op;
  const int64_t per = (nb + blockDim.x - 1) / blockDim.x;
  const int64_t a = (int64_t)threadIdx.x * per, b = min(a + per, nb);
  TileState v = ts_identity();
  for (int64_t i = a; i < b; ++i) v = op(v, ts_load(block_agg + i));
  TileState total;
  const TileState inc = ts_block_scan(v, s_w, total);
  // exclusive prefix of this thread's chunk = inclusive of the previous thread
  TileState pre = ts_shfl_up(inc, 1);
  __shared__ TileState s_last[32];
  if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) pre = threadIdx.x == 0 ? ts_identity() : s_last[(threadIdx.x >> 5) - 1];
  for (int64_t i = a; i < b; ++i) {
    const TileState cur = ts_load(block_agg + i);
    ts_store(block_agg + i, pre);
    pre = op(pre, cur);
  }
}
static __global__ void __launch_bounds__(TS_THREADS) k_ts_scan(const TileState* __restrict__ in, int64_t n,
                                                          const TileState* __restrict__ block_excl,
                                                          TileState* __restrict__ out) {
  __shared__ TileState s_w[TS_THREADS / 32];
  const TileStateOp op;
  const int64_t base = (int64_t)blockIdx.x * TS_BLOCK + (int64_t)threadIdx.x * TS_ITEMS;
  TileState it[TS_ITEMS];
  TileState v = ts_identity();
#pragma unroll
  for (int k = 0; k < TS_ITEMS; ++k) {
    it[k] = ts_identity();
    if (base + k < n) it[k] = ts_load(in + base + k);
    v = op(v, it[k]);
  }
  TileState total;
  const TileState inc = ts_block_scan(v, s_w, total);
  TileState pre = ts_shfl_up(inc, 1);  // inclusive value of the previous thread
  __shared__ TileState s_last[TS_THREADS / 32];
  if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) pre = threadIdx.x == 0 ? ts_identity() : s_last[(threadIdx.x >> 5) - 1];
  pre = op(ts_load(block_excl + blockIdx.x), pre);
#pragma unroll
  for (int k = 0; k < TS_ITEMS; ++k) {
    if (base + k < n) ts_store(out + base + k, pre);
    pre = op(pre, it[k]);
  }
}

// ---- the same scan in ONE launch: decoupled look-back over ticket-ordered blocks (resident shard engine) ----------------
// Block b (ticket order, so every predecessor is running or done) publishes its aggregate, warp 0 polls 32 predecessor
// descriptors at a time until it meets an inclusive prefix, folds them in index order (the operator is associative but not
// commutative: a tile in which a record starts resets the sums) and publishes the block's inclusive prefix.  Descriptor =
// flag (0 nothing, 1 aggregate, 2 inclusive prefix) + two 32-byte payloads; the host clears flags and ticket before the launch.
struct __align__(16) TsDesc {
  uint32_t flag, pad0, pad1, pad2;
  TileState agg, inc;
};
__device__ __forceinline__ TileState ts_load_volatile(const TileState* p) {
  uint4 a, b;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(p) : "memory");
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
               : "l"(reinterpret_cast<const uint4*>(p) + 1)
               : "memory");
  TileState t;
  t.r = a.x; t.q = a.y; t.x1 = a.z; t.x2 = a.w; t.x3 = b.x; t.cnt_a = b.y; t.cnt_b = b.z; t.flag = b.w;
  return t;
}
__device__ __forceinline__ TileState ts_shfl(const TileState& t, int src) {
  TileState o;
  o.r = __shfl_sync(0xffffffffu, t.r, src); o.q = __shfl_sync(0xffffffffu, t.q, src);
  o.x1 = __shfl_sync(0xffffffffu, t.x1, src); o.x2 = __shfl_sync(0xffffffffu, t.x2, src);
  o.x3 = __shfl_sync(0xffffffffu, t.x3, src); o.cnt_a = __shfl_sync(0xffffffffu, t.cnt_a, src);
  o.cnt_b = __shfl_sync(0xffffffffu, t.cnt_b, src); o.flag = __shfl_sync(0xffffffffu, t.flag, src);
  return o;
}
static __global__ void __launch_bounds__(TS_THREADS) k_ts_scan1(const TileState* __restrict__ in, int64_t n, TsDesc* __restrict__ desc,
                                                           unsigned int* __restrict__ ticket, TileState* __restrict__ out) {
  __shared__ TileState s_w[TS_THREADS / 32];
  __shared__ TileState s_last[TS_THREADS / 32];
  __shared__ TileState s_bex;
  __shared__ unsigned int s_tile;
  const TileStateOp op;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const unsigned int tile = s_tile;
  const int64_t base = (int64_t)tile * TS_BLOCK + (int64_t)threadIdx.x * TS_ITEMS;
  TileState it[TS_ITEMS];
  TileState v = ts_identity();
#pragma unroll
  for (int k = 0; k < TS_ITEMS; ++k) {
    it[k] = ts_identity();
    if (base + k < n) it[k] = ts_load(in + base + k);
    v = op(v, it[k]);
  }
  TileState total;
  const TileState inc = ts_block_scan(v, s_w, total);
  TileState pre = ts_shfl_up(inc, 1);  // inclusive value of the previous thread
  if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
  if (threadIdx.x < 32) {              // warp 0: the look-back
    const int lane = threadIdx.x;
    TileState excl = ts_identity();
    if (tile == 0) {
      if (lane == 0) {
        ts_store(&desc[0].inc, total);
        __threadfence();
        st_release_u32(&desc[0].flag, 2u);
      }
    } else {
      if (lane == 0) {
        ts_store(&desc[tile].agg, total);
        __threadfence();
        st_release_u32(&desc[tile].flag, 1u);
      }
      for (int64_t b0 = (int64_t)tile - 1;; b0 -= 32) {
        const int64_t j = b0 - lane;     // lane 0 looks at the nearest predecessor
        unsigned int f = 2u;
        TileState val = ts_identity();
        if (j >= 0) {
          do { f = ld_acquire_u32(&desc[j].flag); } while (f == 0u);
          val = ts_load_volatile(f == 2u ? &desc[j].inc : &desc[j].agg);
        }
        const unsigned int m2 = __ballot_sync(0xffffffffu, f == 2u);
        const int last = m2 ? __ffs(m2) - 1 : 31;
        for (int k = 0; k <= last; ++k) excl = op(ts_shfl(val, k), excl);   // older on the left
        if (m2) break;
      }
      if (lane == 0) {
        ts_store(&desc[tile].inc, op(excl, total));
        __threadfence();
        st_release_u32(&desc[tile].flag, 2u);
      }
    }
    if (lane == 0) s_bex = excl;
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) pre = threadIdx.x == 0 ? ts_identity() : s_last[(threadIdx.x >> 5) - 1];
  pre = op(s_bex, pre);
#pragma unroll
  for (int k = 0; k < TS_ITEMS; ++k) {
    if (base + k < n) ts_store(out + base + k, pre);
    pre = op(pre, it[k]);
  }
}

// ---- sparse follow-up kernels (lane-dense: one thread per candidate / warp tile / record) ----
// k_gate: one thread per staging slot.  A slot is live when the tile named in it (field b,
// written by the streaming kernel) reserved a range that contains the slot: chunk tails and
// entries left over from earlier calls fail that test, so the staging area is never cleared.
// Per live candidate: record lookup, gates, exact sums since the record start (or since the
// tile start when the record began earlier); per-tile counts of both streams by atomics.
template <int MODE>
__device__ __forceinline__ void gate_slot(const ScanParams& p, uint64_t i);

// persistent grid-stride launch: the number of reserved slots is only known on the device
template <int MODE>
static __global__ void __launch_bounds__(256) k_gate(const ScanParams p) {
  unsigned long long n_res = *reinterpret_cast<const volatile unsigned long long*>(p.stage_ctr);
  if (n_res > p.stage_cap) n_res = p.stage_cap;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_res; i += stride) gate_slot<MODE>(p, i);
}

template <int MODE>
__device__ __forceinline__ void gate_slot(const ScanParams& p, uint64_t i) {
  uint4* slot = reinterpret_cast<uint4*>(p.stage + i);
  const uint4 e0 = slot[0], e1 = slot[1];
  const uint32_t t = e0.y;
  if (t >= (uint32_t)p.n_wt) return;
  const uint2 st = __ldg(p.wt_stage + t);
  if (st.x == 0xFFFFFFFFu || i < st.x || i >= (uint64_t)st.x + st.y) return;
  const int j = (int)e0.x;
  if ((unsigned)j >= WT_OPS) return;
  const int64_t wt_start = (int64_t)t * WT_OPS;
  const int n = j & 31;
  // ops 0..n of the candidate's 128-byte row: every needed LDG.128 is issued before the first use
  const uint32_t* row = p.ops + wt_start + (j & ~31);
  const int64_t left = p.n_ops - (wt_start + (j & ~31));
  uint4 v[8];
#pragma unroll
  for (int g4 = 0; g4 < 8; ++g4) {
    v[g4] = make_uint4(0u, 0u, 0u, 0u);
    if (g4 * 4 <= n) {
      if ((int64_t)(g4 * 4 + 4) <= left) {
        v[g4] = __ldg(reinterpret_cast<const uint4*>(row) + g4);
      } else {
        v[g4].x = g4 * 4 < left ? __ldg(row + g4 * 4) : 0u;
        v[g4].y = g4 * 4 + 1 < left ? __ldg(row + g4 * 4 + 1) : 0u;
        v[g4].z = g4 * 4 + 2 < left ? __ldg(row + g4 * 4 + 2) : 0u;
      }
    }
  }
  const int2 ss = __ldg(p.wt_starts + t);
  const int32_t lo = max(ss.x - 1, 0);
  const int32_t hi = max(lo, min(ss.y, (int32_t)p.n_rec - 1));
  const int32_t rec = rec_of_op(p.op_off, lo, hi, wt_start + j);
  const bool carry = __ldg(p.op_off + rec) < wt_start;
  const uint32_t info = __ldg(p.rec_info + rec);
  Sums fs = {0, 0, 0, 0, 0};
  if (!carry) fs = load_sums(p.fstart + 2 * (int64_t)rec);
  Sums wk = {0, 0, 0, 0, 0};
  uint32_t w = 0;
#pragma unroll
  for (int g4 = 0; g4 < 8; ++g4) {
    const uint32_t vv[4] = {v[g4].x, v[g4].y, v[g4].z, v[g4].w};
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = g4 * 4 + u;
      if (idx < n) add_op<MODE>(vv[u], wk);
      if (idx == n) w = vv[u];
    }
  }
  const Sums f = {e0.z + wk.r, e0.w + wk.q, e1.x + wk.x1, e1.y + wk.x2, e1.z + wk.x3};
  const Sums part = f - fs;
  const int32_t len = (int32_t)(w >> 4);
  const bool ok_a = (info & 1u) && len >= p.min_len_a;
  const bool ok_b = (MODE == VSV_MODE_FUSED) && (info & 2u) && len >= p.min_len_b;
  if (ok_a) atomicAdd(&p.wt_agg[t].cnt_a, 1u);
  if (ok_b) atomicAdd(&p.wt_agg[t].cnt_b, 1u);
  slot[0] = make_uint4((uint32_t)rec, w, part.r, part.q);
  slot[1] = make_uint4(part.x1, part.x2, part.x3, (t << 3) | (ok_a ? 1u : 0u) | (ok_b ? 2u : 0u) | (carry ? 4u : 0u));
}

template <int MODE, int STREAM>
__device__ __forceinline__ void emit_sig(const ScanParams& p, vsv_sig* dst, uint32_t rec, uint32_t w, const Sums& s) {
  const int32_t len = (int32_t)(w >> 4);
  const int32_t type = ((w & 15u) == 1u) ? VSV_INS : VSV_DEL;
  const int32_t hc = (int32_t)(__ldg(p.rec_info + rec) >> 2);
  const int32_t q0 = (int32_t)qry_of<MODE, STREAM>(s) + hsign<MODE, STREAM>() * hc;
  int4* o = reinterpret_cast<int4*>(dst);
  o[0] = make_int4((int32_t)rec, type, __ldg(p.pos + rec) + (int32_t)ref_of<MODE, STREAM>(s), len);
  o[1] = make_int4(q0, q0 + (type == VSV_INS ? len : 1), 0, 1);
}

// k_emit: final ordered slots + carry of an open record.  One thread per staging slot (persistent
// grid-stride, liveness test as in k_gate with the tile index k_gate left in the flag word).  The
// output slot of a candidate = exclusive scan of its tile's counts + the gated candidates before
// it in the tile (a tile holds a handful, in op order).
template <int MODE>
static __global__ void __launch_bounds__(256) k_emit(const ScanParams p, const TileState* __restrict__ excl,
                                              vsv_sig* __restrict__ out_a, int64_t cap_a, vsv_sig* __restrict__ out_b,
                                              int64_t cap_b, int64_t* __restrict__ n_out) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // exclusive scan at the identity sentinel == totals
    n_out[0] = excl[p.n_wt].cnt_a;
    n_out[1] = excl[p.n_wt].cnt_b;
  }
  unsigned long long n_res = *reinterpret_cast<const volatile unsigned long long*>(p.stage_ctr);
  if (n_res > p.stage_cap) n_res = p.stage_cap;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_res; i += stride) {
    const uint4* slot = reinterpret_cast<const uint4*>(p.stage + i);
    const uint4 e1 = slot[1];
    if ((e1.w & 3u) == 0u) continue;
    const uint32_t t = e1.w >> 3;
    if (t >= (uint32_t)p.n_wt) continue;
    const uint2 st = __ldg(p.wt_stage + t);
    if (st.x == 0xFFFFFFFFu || i < st.x || i >= (uint64_t)st.x + st.y) continue;
    const uint4 e0 = slot[0];
    const TileState E = ts_load(excl + t);
    int64_t ia = E.cnt_a, ib = E.cnt_b;
    for (uint64_t k = st.x; k < i; ++k) {
      const uint32_t f = reinterpret_cast<const uint4*>(p.stage + k)[1].w;
      ia += f & 1u;
      ib += (f >> 1) & 1u;
    }
    Sums s = {e0.z, e0.w, e1.x, e1.y, e1.z};
    if (e1.w & 4u) { s.r += E.r; s.q += E.q; s.x1 += E.x1; s.x2 += E.x2; s.x3 += E.x3; }
    if ((e1.w & 1u) && ia < cap_a) emit_sig<MODE, 0>(p, out_a + ia, e0.x, e0.y, s);
    if (MODE == VSV_MODE_FUSED && (e1.w & 2u) && ib < cap_b) emit_sig<MODE, 1>(p, out_b + ib, e0.x, e0.y, s);
  }
}

// k_rec_totals: per-record sums of op lengths (reference_end, read length) from the
// in-tile prefixes at record starts and the scanned tile carries.
template <int MODE>
static __global__ void __launch_bounds__(256) k_rec_totals(const ScanParams p, const TileState* __restrict__ excl,
                                                    uint32_t* __restrict__ ref_total, uint32_t* __restrict__ qry_total,
                                                    uint32_t* __restrict__ aux_total) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= p.n_rec) return;
  const int64_t a = __ldg(p.op_off + r), b = __ldg(p.op_off + r + 1);
  Sums s = {0, 0, 0, 0, 0};
  if (b > a) {
    if ((b % WT_OPS) == 0) {  // ends on a tile edge: the next tile's carry is the whole record
      const TileState E = excl[b / WT_OPS];
      s = Sums{E.r, E.q, E.x1, E.x2, E.x3};
    } else {
      const Sums fb = load_sums(p.fstart + 2 * (r + 1));
      if (a / WT_OPS == b / WT_OPS) {
        s = fb - load_sums(p.fstart + 2 * r);
      } else {
        const TileState E = excl[b / WT_OPS];
        s = Sums{E.r + fb.r, E.q + fb.q, E.x1 + fb.x1, E.x2 + fb.x2, E.x3 + fb.x3};
      }
    }
  }
  ref_total[r] = ref_of<MODE, 0>(s);
  qry_total[r] = qry_of<MODE, 0>(s);
  aux_total[r] = aux_of<MODE>(s);
}

