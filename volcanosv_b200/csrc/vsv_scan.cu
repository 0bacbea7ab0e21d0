// Stage 1: CIGAR signature scan (K1).
//
// K1 replaces extract_sig_from_cigar (LI/extract_contig_signature_Hifi.py:53-85,
// LI/extract_reads_signature.py:47-83) and the CIGAR walk of parse_read
// (LI/sig_extract.py:438-478).  VSV_MODE_FUSED runs the two READ scanners (ERS + SE)
// in ONE pass over the packed ops and fills two output streams.
//
// Design v2 (B200-first, HBM/issue-bound streaming kernel, no inter-CTA dependency):
//  * the packed ops of ALL records are one flat u32 array cut into fixed tiles of 4096
//    ops, independent of record boundaries; a persistent CTA per SM slot walks tiles
//    round-robin and pulls each 16 KB tile into shared memory with ONE 1-D TMA bulk copy
//    (cp.async.bulk + mbarrier), double buffered, so loads of tile i+1 overlap the work on
//    tile i and no thread spends instructions on staging;
//  * fast path: every thread folds 16 consecutive ops (bank-conflict-free rotated
//    LDS.128) into two running sums (reference / query advance) plus a 16-bit candidate
//    mask; rare op codes (N, P, H, ...) are summed in a slow path only by warps that saw
//    one;
//  * candidates (D/I >= min_len: well under 1 % of the ops) are pushed to a per-warp
//    shared-memory list and finished densely by warp 0: record lookup in a prefetched
//    per-tile record table, mapq/enable gate, exact offsets from the block prefix, ordered
//    write to a STAGING buffer whose slots are reserved per CTA in chunks (one global
//    atomic per ~256 signatures);
//  * every tile publishes a 32-byte aggregate (sums since the last record start inside
//    the tile, signature counts).  A segmented exclusive scan over the tile aggregates
//    (n_ops/4096 elements) and a warp-per-tile fix-up kernel then move the staged
//    signatures to their final, deterministic (record, op)-ordered slot and add the carry
//    of a record that began in an earlier tile.  The ops are read from HBM exactly once.
//
// v1 (single-pass chained look-back) is described in profiles/r01a_scan_v1.md: it was
// bound by the tile chain at 7 % of HBM.
#include <cub/device/device_scan.cuh>

#include "vsv_common.cuh"

#define SCAN_THREADS 256
#define SCAN_WARPS (SCAN_THREADS / 32)
#define SCAN_VPT 16
#define SCAN_TILE (SCAN_THREADS * SCAN_VPT)
#define SCAN_STAGES 2
#define SCAN_RT 128           // records per tile kept in the shared-memory record table
#define SCAN_WCAND 64         // per-warp candidate list capacity (overflow: dense phase rescans the tile)
#define SCAN_CHUNK 256        // staging slots reserved per global atomic
#define SCAN_CTAS_PER_SM 4

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

struct ScanParams {
  const uint32_t* ops;
  const int64_t* op_off;
  const int32_t* pos;
  const uint32_t* rec_info;  // hc << 2 | gate_b << 1 | gate_a
  int64_t n_rec, n_ops;
  int32_t n_tiles;
  uint32_t min_w;            // min candidate length << 4
  int32_t min_len_a, min_len_b;
  const int2* tile_rec;      // first / last record touching each tile
  TileState* tile_agg;       // per tile aggregate (kernel output)
  uint2* tile_stage;         // staging base of the tile's signatures per stream
  unsigned long long* stage_ctr;  // [2] staging allocation counters
  vsv_sig* stage_a;
  vsv_sig* stage_b;
  uint32_t stage_cap_a, stage_cap_b;
  uint32_t* ref_total;
  uint32_t* qry_total;
  uint32_t* aux_total;
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

__global__ void k_tile_records(const int64_t* __restrict__ op_off, int64_t n_rec, int64_t n_ops, int32_t n_tiles,
                               int2* __restrict__ tile_rec) {
  int32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  int64_t x0 = (int64_t)t * SCAN_TILE;
  int64_t x1 = min(x0 + SCAN_TILE, n_ops) - 1;
  int32_t a = rec_of_op(op_off, 0, (int32_t)n_rec - 1, x0);
  tile_rec[t] = make_int2(a, rec_of_op(op_off, a, (int32_t)n_rec - 1, x1));
}

// per record: leading hard clip and the record gates of both streams
__global__ void k_rec_info(const uint32_t* __restrict__ ops, const int64_t* __restrict__ op_off,
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

// ---- shared memory of the streaming kernel -------------------------------------------
struct ScanSmem {
  uint32_t ops[SCAN_STAGES][SCAN_TILE];   // TMA destinations (linear tiles)
  uint32_t tx[5][SCAN_THREADS];           // warp-relative thread-exclusive prefixes (r,q,x1,x2,x3)
  uint16_t cand[SCAN_WARPS][SCAN_WCAND];  // per-warp candidate op indices, in op order
  int32_t t_off[2][SCAN_RT + 4];          // record table: op_off - tile_start, clamped to [-1, TILE+1]
  int32_t t_pos[2][SCAN_RT];
  uint32_t t_info[2][SCAN_RT];
  uint32_t P[5][SCAN_RT + 4];             // prefix sums at the (clamped) start of every table record
  uint32_t w[6][SCAN_WARPS];              // warp totals (r,q,x1,x2,x3,cand)
  uint32_t wx[6][SCAN_WARPS];             // warp exclusive offsets
  uint32_t tot[6];
  uint32_t wrare[SCAN_WARPS];
  uint32_t ovf;                           // a warp's candidate list overflowed in this tile
  uint64_t bar[SCAN_STAGES];
};

template <int MODE>
__device__ __forceinline__ void add_op(uint32_t w, Sums& s) {
  typedef ModeTraits<MODE> M;
  const uint32_t bit = 1u << (w & 15u), len = w >> 4;
  s.r += (bit & M::REF) ? len : 0u;
  s.q += (bit & M::QRY) ? len : 0u;
  s.x1 += (bit & M::X1) ? len : 0u;
  if (M::NX > 1) s.x2 += (bit & M::X2) ? len : 0u;
  if (M::NX > 2) s.x3 += (bit & M::X3) ? len : 0u;
}

// View of the records touching one tile: shared-memory table when it fits, else global.
struct RecTab {
  const ScanSmem* sm;
  const ScanParams* p;
  int64_t tile_start;
  int32_t rf, nrt, slot;
  bool fast;
  __device__ __forceinline__ int32_t off(int k) const {  // start of record rf+k relative to the tile
    if (fast) return sm->t_off[slot][k];
    const int64_t d = __ldg(p->op_off + rf + k) - tile_start;
    return (int32_t)max((int64_t)-1, min(d, (int64_t)SCAN_TILE + 1));
  }
  __device__ __forceinline__ int32_t pos(int k) const { return fast ? sm->t_pos[slot][k] : __ldg(p->pos + rf + k); }
  __device__ __forceinline__ uint32_t info(int k) const {
    return fast ? sm->t_info[slot][k] : __ldg(p->rec_info + rf + k);
  }
  __device__ __forceinline__ int lookup(int j) const {  // last k with off(k) <= j
    int lo = 0, hi = nrt - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (off(mid) <= j) lo = mid; else hi = mid - 1;
    }
    return lo;
  }
};

// sums over tile ops [0, j)
template <int MODE>
__device__ __forceinline__ Sums prefix_at(const ScanSmem& sm, const uint32_t* tile, int j) {
  typedef ModeTraits<MODE> M;
  Sums s;
  if (j >= SCAN_TILE) {
    s.r = sm.tot[0]; s.q = sm.tot[1]; s.x1 = sm.tot[2]; s.x2 = sm.tot[3]; s.x3 = sm.tot[4];
    return s;
  }
  const int t = j >> 4, n = j & 15, wp = t >> 5;
  s.r = sm.wx[0][wp] + sm.tx[0][t];
  s.q = sm.wx[1][wp] + sm.tx[1][t];
  s.x1 = sm.wx[2][wp]; s.x2 = sm.wx[3][wp]; s.x3 = sm.wx[4][wp];
  const uint32_t* row = tile + t * SCAN_VPT;
  if (sm.wrare[wp]) {  // this warp saw N / H / P ...: full walk
    s.x1 += sm.tx[2][t]; s.x2 += sm.tx[3][t]; s.x3 += sm.tx[4][t];
    for (int i = 0; i < n; ++i) add_op<MODE>(row[i], s);
  } else {
    uint32_t r = 0, q = 0;
    const uint4* row4 = reinterpret_cast<const uint4*>(row);
    const uint4 v0 = row4[0], v1 = row4[1], v2 = row4[2], v3 = row4[3];
    const uint32_t ww[16] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w, v3.x, v3.y, v3.z, v3.w};
#pragma unroll
    for (int i = 0; i < 15; ++i) {
      const uint32_t w = ww[i], bit = (i < n) ? (1u << (w & 15u)) : 0u, len = w >> 4;
      r += (bit & M::REF) ? len : 0u;
      q += (bit & M::QRY) ? len : 0u;
    }
    s.r += r; s.q += q;
  }
  return s;
}

template <int MODE, int STREAM>
__device__ __forceinline__ void write_sig(vsv_sig* dst, int32_t rec, uint32_t w, int32_t rpos, uint32_t hc,
                                          const Sums& part, bool carry) {
  const int32_t len = (int32_t)(w >> 4);
  const int32_t type = ((w & 15u) == 1u) ? VSV_INS : VSV_DEL;
  const int32_t q0 = (int32_t)qry_of<MODE, STREAM>(part) + hsign<MODE, STREAM>() * (int32_t)hc;
  int4* o = reinterpret_cast<int4*>(dst);
  o[0] = make_int4(rec, type, rpos + (int32_t)ref_of<MODE, STREAM>(part), len);
  o[1] = make_int4(q0, q0 + (type == VSV_INS ? len : 1), 0, carry ? 0 : 1);  // count==0: carry pending
}

// ---- K1: the streaming kernel ---------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(SCAN_THREADS, SCAN_CTAS_PER_SM) k_scan_tiles(const ScanParams p) {
  typedef ModeTraits<MODE> M;
  constexpr uint32_t RARE = M::X1 | M::X2 | M::X3;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  ScanSmem& sm = *reinterpret_cast<ScanSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x;
  const int32_t n_tiles = p.n_tiles;
  const int32_t n_full = (int32_t)(p.n_ops / SCAN_TILE);  // tiles below n_full are complete

  if (tid == 0) {
    sm.ovf = 0u;
    for (int s = 0; s < SCAN_STAGES; ++s) mbar_init(&sm.bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
  if (tid == 0) {
    for (int s = 0; s < SCAN_STAGES; ++s) {
      const int64_t t = (int64_t)blockIdx.x + (int64_t)s * G;
      if (t < n_full) {
        mbar_expect_tx(&sm.bar[s], SCAN_TILE * 4);
        tma_load_1d(sm.ops[s], p.ops + t * SCAN_TILE, SCAN_TILE * 4, &sm.bar[s]);
      }
    }
  }
  // record-table registers: desc of the current / next tile, prefetched one iteration ahead
  int2 cur = make_int2(0, -1), nxt = make_int2(0, -1);
  if ((int32_t)blockIdx.x < n_tiles) cur = __ldg(p.tile_rec + blockIdx.x);
  if ((int64_t)blockIdx.x + G < n_tiles) nxt = __ldg(p.tile_rec + blockIdx.x + G);
  {  // first tile's table, synchronously
    const int nrt = cur.y - cur.x + 1;
    if (nrt <= SCAN_RT && tid <= nrt) {
      const int64_t d = __ldg(p.op_off + cur.x + tid) - (int64_t)blockIdx.x * SCAN_TILE;
      sm.t_off[0][tid] = (int32_t)max((int64_t)-1, min(d, (int64_t)SCAN_TILE + 1));
      if (tid < nrt) { sm.t_pos[0][tid] = __ldg(p.pos + cur.x + tid); sm.t_info[0][tid] = __ldg(p.rec_info + cur.x + tid); }
    }
  }
  __syncthreads();

  // staging chunks of this CTA (live in warp 0 / lane 0 registers)
  unsigned long long ch_next[2] = {0, 0}, ch_end[2] = {0, 0};

  int it = 0;
  for (int64_t tile64 = blockIdx.x; tile64 < n_tiles; tile64 += G, ++it) {
    const int32_t tile = (int32_t)tile64;
    const int stage = it % SCAN_STAGES, slot = it & 1;
    const uint32_t parity = (uint32_t)(it / SCAN_STAGES) & 1u;
    const int64_t tile_start = (int64_t)tile * SCAN_TILE;
    uint32_t* tl = sm.ops[stage];

    // -- prefetch the next tile's record table into registers (stored at the bottom) ----
    const int nrt_next = nxt.y - nxt.x + 1;
    int32_t pf_off = 0, pf_pos = 0;
    uint32_t pf_info = 0;
    const bool pf_on = tid <= nrt_next && nrt_next <= SCAN_RT && (tile64 + G < n_tiles);
    if (pf_on) {
      const int64_t d = __ldg(p.op_off + nxt.x + tid) - (tile_start + (int64_t)G * SCAN_TILE);
      pf_off = (int32_t)max((int64_t)-1, min(d, (int64_t)SCAN_TILE + 1));
      if (tid < nrt_next) { pf_pos = __ldg(p.pos + nxt.x + tid); pf_info = __ldg(p.rec_info + nxt.x + tid); }
    }
    int2 nn = make_int2(0, -1);
    if (tile64 + 2 * (int64_t)G < n_tiles) nn = __ldg(p.tile_rec + tile64 + 2 * (int64_t)G);

    // -- the tile's ops: TMA (full tiles) or a guarded copy (the one partial tile) ------
    if (tile < n_full) {
      mbar_wait(&sm.bar[stage], parity);
    } else {
      const int64_t rem = p.n_ops - tile_start;
      for (int i = tid; i < SCAN_TILE; i += SCAN_THREADS) tl[i] = (i < rem) ? __ldg(p.ops + tile_start + i) : 0u;
      __syncthreads();
    }

    // -- fast path: 16 consecutive ops per thread, rotated so LDS.128 is conflict free ---
    uint32_t r = 0, q = 0, bits = 0, em = 0;
    {
      const int rot = (tid >> 1) & 3;
      const uint4* row = reinterpret_cast<const uint4*>(tl + tid * SCAN_VPT);
      const uint32_t min_w = p.min_w;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint4 v = row[(k + rot) & 3];
        const uint32_t ww[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t w = ww[i];
          const uint32_t bit = 1u << (w & 15u), len = w >> 4;
          bits |= bit;
          r += (bit & M::REF) ? len : 0u;
          q += (bit & M::QRY) ? len : 0u;
          em |= ((bit & 6u) && w >= min_w) ? (1u << (k * 4 + i)) : 0u;
        }
      }
      // un-rotate the candidate mask: chunk k of the loop is logical chunk (k+rot)&3
      em = ((em << (4 * rot)) | (em >> (16 - 4 * rot))) & 0xFFFFu;
    }
    const uint32_t c = __popc(em);

    // -- warp scans (inclusive); warp-relative exclusive prefixes go to shared memory ----
    uint32_t ir = r, iq = q, ic = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t a = __shfl_up_sync(0xffffffffu, ir, d);
      const uint32_t b = __shfl_up_sync(0xffffffffu, iq, d);
      const uint32_t e = __shfl_up_sync(0xffffffffu, ic, d);
      if (lane >= d) { ir += a; iq += b; ic += e; }
    }
    sm.tx[0][tid] = ir - r;
    sm.tx[1][tid] = iq - q;
    const bool anyrare = __any_sync(0xffffffffu, (bits & RARE) != 0u);
    uint32_t ix1 = 0, ix2 = 0, ix3 = 0;
    if (anyrare) {  // slow path: this warp saw N / H / P / ...
      Sums xs = {0, 0, 0, 0, 0};
      if (bits & RARE)
        for (int i = 0; i < SCAN_VPT; ++i) add_op<MODE>(tl[tid * SCAN_VPT + i], xs);
      ix1 = xs.x1; ix2 = xs.x2; ix3 = xs.x3;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, ix1, d);
        const uint32_t b = __shfl_up_sync(0xffffffffu, ix2, d);
        const uint32_t e = __shfl_up_sync(0xffffffffu, ix3, d);
        if (lane >= d) { ix1 += a; ix2 += b; ix3 += e; }
      }
      sm.tx[2][tid] = ix1 - xs.x1; sm.tx[3][tid] = ix2 - xs.x2; sm.tx[4][tid] = ix3 - xs.x3;
    }
    if (lane == 31) {
      sm.w[0][warp] = ir; sm.w[1][warp] = iq; sm.w[2][warp] = ix1; sm.w[3][warp] = ix2; sm.w[4][warp] = ix3;
      sm.w[5][warp] = ic;
      sm.wrare[warp] = anyrare ? 1u : 0u;
    }
    if (em) {  // push this thread's candidates, op order, warp-relative slot
      uint32_t idx = ic - c, m = em;
      while (m) {
        const int s = __ffs(m) - 1;
        m &= m - 1;
        if (idx < SCAN_WCAND) sm.cand[warp][idx] = (uint16_t)(tid * SCAN_VPT + s);
        else sm.ovf = 1u;
        ++idx;
      }
    }
    __syncthreads();  // B1: warp totals, candidate lists, tx visible

    if (warp == 0) {  // second level: exclusive offsets of the 8 warps
      uint32_t v[6], own[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) { own[k] = lane < SCAN_WARPS ? sm.w[k][lane] : 0u; v[k] = own[k]; }
#pragma unroll
      for (int d = 1; d < SCAN_WARPS; d <<= 1) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          const uint32_t a = __shfl_up_sync(0xffffffffu, v[k], d);
          if (lane >= d) v[k] += a;
        }
      }
      if (lane < SCAN_WARPS) {
#pragma unroll
        for (int k = 0; k < 6; ++k) sm.wx[k][lane] = v[k] - own[k];
      }
      if (lane == SCAN_WARPS - 1) {
#pragma unroll
        for (int k = 0; k < 6; ++k) sm.tot[k] = v[k];
      }
    }
    RecTab rt;
    rt.sm = &sm; rt.p = &p; rt.tile_start = tile_start; rt.rf = cur.x; rt.nrt = cur.y - cur.x + 1; rt.slot = slot;
    rt.fast = rt.nrt <= SCAN_RT;
    // B2 (block prefixes complete) is only needed when warps 1..7 take part in what follows
    if (!rt.fast) __syncthreads(); else __syncwarp();

    if (warp == 0) {
      // -- prefix sums at every record start (lanes = records), then the per-record totals --
      if (rt.fast) {
        for (int k = lane; k <= rt.nrt; k += 32) {
          const Sums ps = prefix_at<MODE>(sm, tl, min(max(rt.off(k), 0), SCAN_TILE));
          sm.P[0][k] = ps.r; sm.P[1][k] = ps.q; sm.P[2][k] = ps.x1; sm.P[3][k] = ps.x2; sm.P[4][k] = ps.x3;
        }
        __syncwarp();
        if (p.ref_total != nullptr) {
          for (int k = lane; k < rt.nrt; k += 32) {
            const Sums d = {sm.P[0][k + 1] - sm.P[0][k], sm.P[1][k + 1] - sm.P[1][k], sm.P[2][k + 1] - sm.P[2][k],
                            sm.P[3][k + 1] - sm.P[3][k], sm.P[4][k + 1] - sm.P[4][k]};
            const uint32_t tr = ref_of<MODE, 0>(d), tq = qry_of<MODE, 0>(d), tx = aux_of<MODE>(d);
            const int64_t rr = (int64_t)rt.rf + k;
            if (rt.off(k) >= 0 && rt.off(k + 1) <= SCAN_TILE) {
              p.ref_total[rr] = tr; p.qry_total[rr] = tq; p.aux_total[rr] = tx;
            } else {
              atomicAdd(p.ref_total + rr, tr); atomicAdd(p.qry_total + rr, tq); atomicAdd(p.aux_total + rr, tx);
            }
          }
        }
      }
      // -- dense candidate phase (one pass): staging slots are reserved for the RAW candidate
      //    count (an upper bound of what passes the record gates), then each group of 32
      //    candidates is looked up, gated and written in op order.
      const uint32_t ncand = sm.tot[5];
      const bool ovf = sm.ovf != 0u;
      uint32_t cnt_a = 0, cnt_b = 0;
      unsigned long long base_a = ~0ull, base_b = ~0ull;
      if (ncand) {
        if (lane == 0) {  // chunked reservation: one global atomic per ~SCAN_CHUNK signatures
#pragma unroll
          for (int s = 0; s < (MODE == VSV_MODE_FUSED ? 2 : 1); ++s) {
            if (ch_next[s] + ncand > ch_end[s]) {
              const unsigned long long n = ncand > SCAN_CHUNK ? ncand : SCAN_CHUNK;
              ch_next[s] = atomicAdd(p.stage_ctr + s, n);
              ch_end[s] = ch_next[s] + n;
            }
            unsigned long long b0 = ch_next[s];
            ch_next[s] += ncand;
            if (b0 + ncand > (s == 0 ? p.stage_cap_a : p.stage_cap_b)) {  // staging full: host retries
              b0 = ~0ull;
              p.stage_ctr[2] = 1ull;
            }
            if (s == 0) base_a = b0; else base_b = b0;
          }
        }
        base_a = __shfl_sync(0xffffffffu, base_a, 0);
        base_b = __shfl_sync(0xffffffffu, base_b, 0);
        const uint32_t ngroups = ovf ? SCAN_TILE / 32 : (ncand + 31) / 32;
        for (uint32_t g = 0; g < ngroups; ++g) {
          bool valid;
          int j;
          if (!ovf) {
            const uint32_t i = g * 32 + lane;
            valid = i < ncand;
            int wp = 0;
#pragma unroll
            for (int u = 1; u < SCAN_WARPS; ++u) wp += (sm.wx[5][u] <= i) ? 1 : 0;
            j = valid ? sm.cand[wp][i - sm.wx[5][wp]] : 0;
          } else {  // a list overflowed: enumerate the candidates straight from the tile
            j = g * 32 + lane;
            const uint32_t w0 = tl[j];
            valid = ((1u << (w0 & 15u)) & 6u) && w0 >= p.min_w;
          }
          bool ok_a = false, ok_b = false;
          int k = 0;
          uint32_t w = 0, info = 0;
          if (valid) {
            w = tl[j];
            k = rt.lookup(j);
            info = rt.info(k);
            const int32_t len = (int32_t)(w >> 4);
            ok_a = (info & 1u) && len >= p.min_len_a;
            if (MODE == VSV_MODE_FUSED) ok_b = (info & 2u) && len >= p.min_len_b;
          }
          const uint32_t ba = __ballot_sync(0xffffffffu, ok_a);
          const uint32_t bb = (MODE == VSV_MODE_FUSED) ? __ballot_sync(0xffffffffu, ok_b) : 0u;
          if (ok_a || ok_b) {
            const Sums pre = prefix_at<MODE>(sm, tl, j);
            const int32_t o = rt.off(k);
            const bool carry = o < 0;
            Sums part = pre;
            if (!carry) {
              if (rt.fast) part = pre - Sums{sm.P[0][k], sm.P[1][k], sm.P[2][k], sm.P[3][k], sm.P[4][k]};
              else part = pre - prefix_at<MODE>(sm, tl, o);
            }
            const uint32_t lt = (1u << lane) - 1u;
            if (ok_a && base_a != ~0ull)
              write_sig<MODE, 0>(p.stage_a + base_a + cnt_a + __popc(ba & lt), rt.rf + k, w, rt.pos(k), info >> 2, part,
                                 carry);
            if (MODE == VSV_MODE_FUSED && ok_b && base_b != ~0ull)
              write_sig<MODE, 1>(p.stage_b + base_b + cnt_b + __popc(bb & lt), rt.rf + k, w, rt.pos(k), info >> 2, part,
                                 carry);
          }
          cnt_a += __popc(ba);
          cnt_b += __popc(bb);
        }
      }
      if (lane == 0) {  // tile aggregate: sums since the last record start inside the tile
        Sums a = {sm.tot[0], sm.tot[1], sm.tot[2], sm.tot[3], sm.tot[4]};
        const int32_t o = rt.off(rt.nrt - 1);
        const bool bound = o >= 0;
        if (bound) {
          const int kl = rt.nrt - 1;
          if (rt.fast) a = a - Sums{sm.P[0][kl], sm.P[1][kl], sm.P[2][kl], sm.P[3][kl], sm.P[4][kl]};
          else a = a - prefix_at<MODE>(sm, tl, o);
        }
        TileState ts;
        ts.r = a.r; ts.q = a.q; ts.x1 = a.x1; ts.x2 = a.x2; ts.x3 = a.x3;
        ts.cnt_a = cnt_a; ts.cnt_b = cnt_b; ts.flag = bound ? 1u : 0u;
        p.tile_agg[tile] = ts;
        p.tile_stage[tile] = make_uint2(base_a == ~0ull ? 0xFFFFFFFFu : (uint32_t)base_a,
                                        base_b == ~0ull ? 0xFFFFFFFFu : (uint32_t)base_b);
        sm.ovf = 0u;  // for the next tile (pushers write it only after B3)
      }
    } else if (!rt.fast && p.ref_total != nullptr) {
      // -- warps 1..7, only for tiles with more than SCAN_RT records: per-record totals ----
      for (int k = tid - 32; k < rt.nrt; k += SCAN_THREADS - 32) {
        const int32_t a = rt.off(k), b = rt.off(k + 1);
        const Sums d = prefix_at<MODE>(sm, tl, min(b, SCAN_TILE)) - prefix_at<MODE>(sm, tl, max(a, 0));
        Sums rd = d;  // per-record totals follow stream A's semantics
        const uint32_t tr = ref_of<MODE, 0>(rd), tq = qry_of<MODE, 0>(rd), tx = aux_of<MODE>(rd);
        const int64_t rr = (int64_t)rt.rf + k;
        if (a >= 0 && b <= SCAN_TILE) {
          p.ref_total[rr] = tr; p.qry_total[rr] = tq; p.aux_total[rr] = tx;
        } else {
          atomicAdd(p.ref_total + rr, tr); atomicAdd(p.qry_total + rr, tq); atomicAdd(p.aux_total + rr, tx);
        }
      }
    }
    // -- store the prefetched record table of the next tile -------------------------------
    if (pf_on) {
      sm.t_off[slot ^ 1][tid] = pf_off;
      if (tid < nrt_next) { sm.t_pos[slot ^ 1][tid] = pf_pos; sm.t_info[slot ^ 1][tid] = pf_info; }
    }
    __syncthreads();  // B3: everyone is done with this stage and this table slot
    if (tid == 0) {
      const int64_t t2 = tile64 + (int64_t)SCAN_STAGES * G;
      if (t2 < n_full) {
        fence_proxy_async();
        mbar_expect_tx(&sm.bar[stage], SCAN_TILE * 4);
        tma_load_1d(sm.ops[stage], p.ops + t2 * SCAN_TILE, SCAN_TILE * 4, &sm.bar[stage]);
      }
    }
    cur = nxt;
    nxt = nn;
  }
}

// ---- fix-up: staged signatures -> final ordered slots (+ carry of an open record) -----
template <int MODE, int STREAM>
__device__ __forceinline__ void fix_stream(const vsv_sig* __restrict__ stage, uint32_t base, uint32_t cnt,
                                           uint64_t dst0, const Sums& carry, vsv_sig* __restrict__ out, int64_t cap,
                                           int lane) {
  if (base == 0xFFFFFFFFu) return;  // staging overflowed: the call reports VSV_E_CAPACITY
  const int32_t cr = (int32_t)ref_of<MODE, STREAM>(carry), cq = (int32_t)qry_of<MODE, STREAM>(carry);
  for (uint32_t i = lane; i < cnt; i += 32) {
    if ((int64_t)(dst0 + i) >= cap) break;
    const int4* s = reinterpret_cast<const int4*>(stage + base + i);
    int4 a = s[0], b = s[1];
    if (b.w == 0) { a.z += cr; b.x += cq; b.y += cq; b.w = 1; }
    int4* o = reinterpret_cast<int4*>(out + dst0 + i);
    o[0] = a;
    o[1] = b;
  }
}

template <int MODE>
__global__ void __launch_bounds__(256) k_scan_fixup(const TileState* __restrict__ agg, const TileState* __restrict__ excl,
                                                   const uint2* __restrict__ tile_stage, int32_t n_tiles,
                                                   const vsv_sig* __restrict__ stage_a, const vsv_sig* __restrict__ stage_b,
                                                   vsv_sig* __restrict__ out_a, int64_t cap_a, vsv_sig* __restrict__ out_b,
                                                   int64_t cap_b, int64_t* __restrict__ n_out) {
  const int lane = threadIdx.x & 31;
  const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (t >= n_tiles) return;
  const TileState a = agg[t], e = excl[t];
  if (t == n_tiles - 1 && lane == 0) {
    n_out[0] = (int64_t)e.cnt_a + a.cnt_a;
    n_out[1] = (int64_t)e.cnt_b + a.cnt_b;
  }
  if (a.cnt_a == 0 && a.cnt_b == 0) return;
  const uint2 st = tile_stage[t];
  const Sums carry = {e.r, e.q, e.x1, e.x2, e.x3};
  if (a.cnt_a) fix_stream<MODE, 0>(stage_a, st.x, a.cnt_a, e.cnt_a, carry, out_a, cap_a, lane);
  if (MODE == VSV_MODE_FUSED && a.cnt_b) fix_stream<MODE, 1>(stage_b, st.y, a.cnt_b, e.cnt_b, carry, out_b, cap_b, lane);
}

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

template <int MODE>
static int launch_scan(vsv_ctx* ctx, const ScanParams& p, int grid) {
  static bool attr_set = false;
  if (!attr_set) {
    VSV_CUDA(ctx, cudaFuncSetAttribute(k_scan_tiles<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)sizeof(ScanSmem)));
    attr_set = true;
  }
  VSV_LAUNCH(ctx, k_scan_tiles<MODE>, (unsigned)grid, SCAN_THREADS, sizeof(ScanSmem), p);
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
  if (c.n_rec >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": more than 2^31-1 records per call");
  if (c.cap_a >= (int64_t(1) << 30) || c.cap_b >= (int64_t(1) << 30))
    VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": capacity must be below 2^30 signatures per call");
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
  const int64_t n_tiles64 = vsv_cdiv(n_ops, SCAN_TILE);
  if (n_tiles64 >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": too many ops per call");
  const int32_t n_tiles = (int32_t)n_tiles64;

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
  const int grid = (int)std::min<int64_t>(n_tiles, (int64_t)ctx->sm_count * SCAN_CTAS_PER_SM);
  // staging: slots are reserved per tile for the RAW candidate count in chunks, so the exact
  // need is only known after a run; start from the caller's capacities and retry (rare) with
  // the measured reservation if a tile found the staging full.
  const int64_t slack = (int64_t)grid * SCAN_CHUNK + SCAN_CHUNK;
  int64_t scap_a = (fused ? std::max(c.cap_a, c.cap_b) : c.cap_a) * 5 / 4 + slack;
  int64_t scap_b = fused ? scap_a : 0;
  int64_t* d_nout = nullptr;  // [0..1] signature counts, [2..3] staging counters, [4] staging-full flag
  uint32_t* d_info = nullptr;
  int2* d_trec = nullptr;
  TileState *d_agg = nullptr, *d_excl = nullptr;
  uint2* d_tstage = nullptr;
  VSV_TRY(vsv_tmp(ctx, 5, &d_nout));
  VSV_TRY(vsv_tmp(ctx, (size_t)c.n_rec, &d_info));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_trec));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_agg));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_excl));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_tiles, &d_tstage));

  p.rec_info = d_info;
  p.n_rec = c.n_rec; p.n_ops = n_ops; p.n_tiles = n_tiles;
  p.min_len_a = c.min_len_a; p.min_len_b = c.min_len_b;
  p.min_w = (uint32_t)(fused ? std::min(c.min_len_a, c.min_len_b) : c.min_len_a) << 4;
  p.tile_rec = d_trec; p.tile_agg = d_agg; p.tile_stage = d_tstage;
  p.stage_ctr = reinterpret_cast<unsigned long long*>(d_nout + 2);
  p.ref_total = (uint32_t*)d_rt; p.qry_total = (uint32_t*)d_qt; p.aux_total = (uint32_t*)d_xt;

  int64_t h_n[5] = {0, 0, 0, 0, 0};
  VSV_TRY(vsv_t0(ctx));
  if (n_tiles > 0) {
    VSV_LAUNCH(ctx, k_rec_info, (unsigned)vsv_cdiv(c.n_rec, 256), 256, 0, p.ops, p.op_off, d_mapq, d_ena, d_enb, c.n_rec,
               c.min_mapq_a, c.min_mapq_b, d_info);
    VSV_LAUNCH(ctx, k_tile_records, (unsigned)vsv_cdiv(n_tiles, 256), 256, 0, p.op_off, c.n_rec, n_ops, n_tiles, d_trec);
  }
  for (int attempt = 0; attempt < 3; ++attempt) {
    if (scap_a >= (int64_t(1) << 32) - 1 || scap_b >= (int64_t(1) << 32) - 1)
      VSV_FAIL(ctx, VSV_E_ARG, std::string(fn) + ": too many candidate signatures for one call; split the batch");
    vsv_sig *d_sa = nullptr, *d_sb = nullptr;
    VSV_TRY(vsv_tmp(ctx, (size_t)scap_a, &d_sa));
    if (fused) VSV_TRY(vsv_tmp(ctx, (size_t)scap_b, &d_sb));
    p.stage_a = d_sa; p.stage_b = d_sb;
    p.stage_cap_a = (uint32_t)scap_a; p.stage_cap_b = (uint32_t)scap_b;
    VSV_CUDA(ctx, cudaMemsetAsync(d_nout, 0, 5 * sizeof(int64_t), ctx->stream));
    if (want_tot) {
      VSV_CUDA(ctx, cudaMemsetAsync(d_rt, 0, sizeof(int32_t) * (size_t)c.n_rec, ctx->stream));
      VSV_CUDA(ctx, cudaMemsetAsync(d_qt, 0, sizeof(int32_t) * (size_t)c.n_rec, ctx->stream));
      VSV_CUDA(ctx, cudaMemsetAsync(d_xt, 0, sizeof(int32_t) * (size_t)c.n_rec, ctx->stream));
    }
    if (n_tiles == 0) break;
    VSV_CUDA(ctx, cudaEventRecord(ctx->evk0, ctx->stream));
    switch (c.mode) {
      case VSV_MODE_CONTIG: VSV_TRY(launch_scan<VSV_MODE_CONTIG>(ctx, p, grid)); break;
      case VSV_MODE_READS: VSV_TRY(launch_scan<VSV_MODE_READS>(ctx, p, grid)); break;
      case VSV_MODE_CUTE: VSV_TRY(launch_scan<VSV_MODE_CUTE>(ctx, p, grid)); break;
      default: VSV_TRY(launch_scan<VSV_MODE_FUSED>(ctx, p, grid)); break;
    }
    VSV_CUDA(ctx, cudaEventRecord(ctx->evk1, ctx->stream));
    ctx->main_open = true;
    {  // segmented exclusive scan of the tile aggregates
      TileState init;
      memset(&init, 0, sizeof(init));
      size_t tb = 0;
      VSV_CUDA(ctx, cub::DeviceScan::ExclusiveScan(nullptr, tb, d_agg, d_excl, TileStateOp(), init, n_tiles, ctx->stream));
      void* tmp = nullptr;
      VSV_TRY(vsv_arena_alloc(ctx, tb, &tmp));
      VSV_CUDA(ctx, cub::DeviceScan::ExclusiveScan(tmp, tb, d_agg, d_excl, TileStateOp(), init, n_tiles, ctx->stream));
      ctx->launches += 2;
    }
    const unsigned gf = (unsigned)vsv_cdiv((int64_t)n_tiles * 32, 256);
    switch (c.mode) {
      case VSV_MODE_CONTIG:
        VSV_LAUNCH(ctx, k_scan_fixup<VSV_MODE_CONTIG>, gf, 256, 0, d_agg, d_excl, d_tstage, n_tiles, d_sa, d_sb, d_out_a,
                   c.cap_a, d_out_b, c.cap_b, d_nout);
        break;
      case VSV_MODE_READS:
        VSV_LAUNCH(ctx, k_scan_fixup<VSV_MODE_READS>, gf, 256, 0, d_agg, d_excl, d_tstage, n_tiles, d_sa, d_sb, d_out_a,
                   c.cap_a, d_out_b, c.cap_b, d_nout);
        break;
      case VSV_MODE_CUTE:
        VSV_LAUNCH(ctx, k_scan_fixup<VSV_MODE_CUTE>, gf, 256, 0, d_agg, d_excl, d_tstage, n_tiles, d_sa, d_sb, d_out_a,
                   c.cap_a, d_out_b, c.cap_b, d_nout);
        break;
      default:
        VSV_LAUNCH(ctx, k_scan_fixup<VSV_MODE_FUSED>, gf, 256, 0, d_agg, d_excl, d_tstage, n_tiles, d_sa, d_sb, d_out_a,
                   c.cap_a, d_out_b, c.cap_b, d_nout);
        break;
    }
    // the counts are needed on the host to size the copy-out: one small sync
    VSV_CUDA(ctx, cudaMemcpyAsync(h_n, d_nout, 5 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const bool fits = h_n[0] <= c.cap_a && (!fused || h_n[1] <= c.cap_b);
    if (h_n[4] == 0 || !fits) break;  // done, or the caller has to grow its buffers anyway
    if (attempt == 2) VSV_FAIL(ctx, VSV_E_NOMEM, std::string(fn) + ": staging kept overflowing");
    scap_a = 2 * h_n[2] + slack;      // reservation of the failed run bounds the raw count
    if (fused) scap_b = 2 * h_n[3] + slack;
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
