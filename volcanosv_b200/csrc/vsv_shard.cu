// Resident shard engine: the whole read-side hot path of one chromosome (RVC:83-96 ERS -> FPF and
// FGC:111-172 SE -> CSS -> GTD/GTI) enqueued on ONE stream with NO host synchronisation in between.
//
// Round 1 chained one C-ABI call per reference function; five of them returned a count through
// cudaStreamSynchronize and the step spent 1.3 ms of 2.0 ms outside the streaming scan (94 launches,
// 3 x 4 radix passes, host gaps).  Here every table has a fixed capacity, its row count lives on the
// device (ShardState), kernels read their bounds from there, and the host reads all counts ONCE after the
// step (vsv_shard_counts_get).  What changed in the data path:
//   * ONE radix sort (tag << 28 | position) orders four tables at once: the read-signature table
//     (sort_sig, ERS:254-258), the DEL and INS `.sigs` tables (SE:637-638) and the read ends that the
//     closed-form spanning depth needs;
//   * merged / split / SA rows are appended with block-aggregated atomics (their order is fixed by the sort);
//   * `sort -u` (SE:637-638), unique signatures (GTD:65-86, GTI:67-101) and the running maximum of the DEL
//     ends are one single-pass scan (decoupled look-back) per table;
//   * the per-candidate joins are one kernel (FP support, coverage, supporting signatures), one look-back
//     scan for the last_match_idx carry (GTD:94-133) and one kernel for depth + genotype.
#include <cub/device/device_radix_sort.cuh>

#include "vsv_depth.cuh"
#include "vsv_rules.cuh"
#include "vsv_scan_kernels.cuh"

#define POS_BITS 28
#define POS_MASK 0x0FFFFFFFu
#define KEY_SENTINEL 0xFFFFFFFFu
#define TAG_READS 0u
#define TAG_DEL 1u
#define TAG_INS 2u
#define TAG_END 3u
#define DUPLIST_CAP 65536
#define GROUP_THREADS 256
#define GROUP_ITEMS 8
#define GROUP_TILE (GROUP_THREADS * GROUP_ITEMS)
#define CHAIN_THREADS 1024
#define ROW_SA_REV 256

// ---- device-resident counters of one step (zeroed by one memset) -----------------------------------
struct ShardState {
  int64_t nout[2];                  // k_emit: CIGAR read signatures, raw cute signatures
  unsigned long long stage_ctr[3];  // staging counter, -, staging-full flag
  unsigned long long n_split, n_del, n_ins, n_sa, n_del_dup, n_ins_dup;
  unsigned long long n_u[2];        // unique DEL / INS signatures
  unsigned int overflow;            // VSV_OVF_*
  unsigned int n_work;              // long tie runs queued
  unsigned int ticket[2];           // tile tickets of the two group scans
  unsigned int pad[2];
};

// clamped row counts of the four sorted tables and their offsets in the sorted key array
struct Tables {
  int64_t n_cig, n_r, n_c, n_d, n_i;
  int64_t off[5];
};
struct Caps {
  int64_t r, c, d, i, n_rec;
};
__device__ __forceinline__ Tables tables_of(const ShardState* st, const Caps& cap) {
  Tables t;
  t.n_cig = min((int64_t)st->nout[0], cap.r);
  t.n_r = min(t.n_cig + (int64_t)st->n_split, cap.r);
  t.n_c = min((int64_t)st->nout[1], cap.c);
  t.n_d = min((int64_t)st->n_del, cap.d);
  t.n_i = min((int64_t)st->n_ins, cap.i);
  t.off[0] = 0; t.off[1] = t.n_r; t.off[2] = t.off[1] + t.n_d; t.off[3] = t.off[2] + t.n_i; t.off[4] = t.off[3] + cap.n_rec;
  return t;
}

// ---- single-pass scan across CTAs: decoupled look-back with ticket-ordered tiles ----------------------
// payload = two 32-bit lanes; flag 0 = nothing yet, 1 = tile aggregate in `agg`, 2 = inclusive prefix in `inc`
struct __align__(8) LbDesc {
  unsigned int flag, pad;
  int2 agg, inc;
};
__device__ __forceinline__ int2 ld_volatile_i2(const int2* p) {
  int2 v;
  asm volatile("ld.volatile.global.v2.s32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
// called by ONE thread of the tile: publishes the aggregate, walks back to the nearest inclusive prefix,
// publishes this tile's inclusive prefix; returns the exclusive prefix.  op(left, right).
template <typename Op>
__device__ __forceinline__ int2 lookback_exclusive(LbDesc* desc, unsigned int tile, int2 agg, int2 identity, Op op) {
  if (tile == 0) {
    desc[0].inc = agg;
    __threadfence();
    st_release_u32(&desc[0].flag, 2u);
    return identity;
  }
  desc[tile].agg = agg;
  __threadfence();
  st_release_u32(&desc[tile].flag, 1u);
  int2 excl = identity;
  for (int64_t j = (int64_t)tile - 1;; --j) {
    unsigned int f;
    do { f = ld_acquire_u32(&desc[j].flag); } while (f == 0u);
    if (f == 2u) {
      excl = op(ld_volatile_i2(&desc[j].inc), excl);
      break;
    }
    excl = op(ld_volatile_i2(&desc[j].agg), excl);
  }
  desc[tile].inc = op(excl, agg);
  __threadfence();
  st_release_u32(&desc[tile].flag, 2u);
  return excl;
}

// Same, called by all 32 lanes of ONE warp with identical arguments: 32 predecessor descriptors are polled at once
// (the serial walk above costs one global round trip per tile and was 16 % of the ONT step).  Returns the exclusive
// prefix to every lane.
template <typename Op>
__device__ __forceinline__ int2 lookback_exclusive_warp(LbDesc* desc, unsigned int tile, int2 agg, int2 identity, Op op) {
  const int lane = threadIdx.x & 31;
  if (tile == 0) {
    if (lane == 0) {
      desc[0].inc = agg;
      __threadfence();
      st_release_u32(&desc[0].flag, 2u);
    }
    return identity;
  }
  if (lane == 0) {
    desc[tile].agg = agg;
    __threadfence();
    st_release_u32(&desc[tile].flag, 1u);
  }
  int2 excl = identity;
  for (int64_t base = (int64_t)tile - 1;; base -= 32) {
    const int64_t j = base - lane;   // lane 0 looks at the nearest predecessor
    unsigned int f = 2u;
    int2 val = identity;
    if (j >= 0) {
      do { f = ld_acquire_u32(&desc[j].flag); } while (f == 0u);
      val = ld_volatile_i2(f == 2u ? &desc[j].inc : &desc[j].agg);
    }
    const unsigned int m2 = __ballot_sync(0xffffffffu, f == 2u);
    const int last = m2 ? __ffs(m2) - 1 : 31;
    for (int k = 0; k <= last; ++k) {
      int2 v;
      v.x = __shfl_sync(0xffffffffu, val.x, k); v.y = __shfl_sync(0xffffffffu, val.y, k);
      excl = op(v, excl);
    }
    if (m2) break;
  }
  if (lane == 0) {
    desc[tile].inc = op(excl, agg);
    __threadfence();
    st_release_u32(&desc[tile].flag, 2u);
  }
  return excl;
}

// ---- P0: per record gates + the first record of every warp tile ------------------------------------
static __global__ void k_shard_prep(const uint32_t* __restrict__ ops, const int64_t* __restrict__ op_off,
                                    const uint8_t* __restrict__ mapq, const int32_t* __restrict__ qlen, int64_t n_rec,
                                    int32_t n_wt, int min_mapq_a, int min_mapq_b, int min_read_len,
                                    uint32_t* __restrict__ info, int32_t* __restrict__ first) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n_rec) return;
  const int64_t cur = __ldg(op_off + r);
  if (r < n_rec) {
    const int64_t b = __ldg(op_off + r + 1);
    uint32_t hc = 0;
    if (b > cur) {
      const uint32_t f = __ldg(ops + cur);
      if ((f & 15u) == 5u) hc = f >> 4;
    }
    const int mq = mapq[r];
    const uint32_t ga = mq >= min_mapq_a ? 1u : 0u;                                 // ERS:120
    const uint32_t gb = (mq >= min_mapq_b && qlen[r] >= min_read_len) ? 2u : 0u;   // SE:439,456
    info[r] = (hc << 2) | gb | ga;
  }
  // k_wt_first: record r is first(t) for the tiles t in (op_off[r-1] / WT_OPS, op_off[r] / WT_OPS]
  const int64_t prev = r > 0 ? __ldg(op_off + r - 1) : -1;
  int64_t t0 = prev < 0 ? 0 : prev / WT_OPS + 1;
  const int64_t t1 = min(cur / WT_OPS, (int64_t)n_wt);
  for (; t0 <= t1; ++t0) first[t0] = (int32_t)r;
  if (r == n_rec && cur % WT_OPS != 0) first[n_wt] = (int32_t)n_rec + 1;
}

// ---- P8: reference_end, read length and the -include_bed gate of every record ----------------------
struct BedParams {
  const int32_t *b0, *b1;
  int64_t n;          // -1: no BED
  int64_t chrom_len;
  int64_t batch;
};
static __global__ void __launch_bounds__(256)
k_shard_rec_totals(const ScanParams p, const TileState* __restrict__ excl, BedParams bed, int32_t* __restrict__ ref_end,
                   int32_t* __restrict__ readlen, uint8_t* __restrict__ in_bed) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= p.n_rec) return;
  const int64_t a = __ldg(p.op_off + r), b = __ldg(p.op_off + r + 1);
  Sums s = {0, 0, 0, 0, 0};
  if (b > a) {
    if ((b % WT_OPS) == 0) {
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
  const int32_t start = __ldg(p.pos + r);
  const int32_t end = start + (int32_t)ref_of<VSV_MODE_FUSED, 0>(s);   // pysam reference_end (M,D,N,=,X)
  ref_end[r] = end;
  readlen[r] = (int32_t)(qry_of<VSV_MODE_FUSED, 0>(s) + aux_of<VSV_MODE_FUSED>(s));  // ERS:151: M,=,X,I,S + H
  uint8_t ok = 1;
  if (bed.n >= 0) {  // SE:58-80 (regions attached to the 10 Mb task they overlap) + SE:524-535
    ok = 0;
    int64_t t0 = 0, t1 = bed.chrom_len;
    if (bed.chrom_len >= bed.batch) {
      t0 = ((int64_t)start / bed.batch) * bed.batch;
      t1 = min(t0 + bed.batch, bed.chrom_len);
    }
    for (int64_t k = 0; k < bed.n; ++k) {
      const int64_t r0 = bed.b0[k], r1 = bed.b1[k];
      const bool attached = (t0 <= r0 && t1 > r0) || (r0 <= t0 && t0 < r1);
      if (attached && !((int64_t)end <= r0 || (int64_t)start >= r1)) { ok = 1; break; }
    }
  }
  in_bed[r] = ok;
}

// ---- P9: read split pairs (ERS:147-248): consecutive same-name records in BAM order ------------------
struct RecCols {
  const uint32_t* ops;
  const int64_t* op_off;
  const int32_t* pos;
  const uint8_t* mapq;
  const uint16_t* flag;
  const int32_t* qlen;
  const int32_t* name;
  const int32_t* next;
  int64_t n_rec;
};
__device__ __forceinline__ uint32_t first_op_of(const RecCols& R, int64_t r) {
  const int64_t a = __ldg(R.op_off + r), b = __ldg(R.op_off + r + 1);
  return b > a ? __ldg(R.ops + a) : 0u;
}
__device__ __forceinline__ uint32_t last_op_of(const RecCols& R, int64_t r) {
  const int64_t a = __ldg(R.op_off + r), b = __ldg(R.op_off + r + 1);
  return b > a ? __ldg(R.ops + b - 1) : 0u;
}
__device__ __forceinline__ void put_sig(vsv_sig* dst, const vsv_sig& v) {
  reinterpret_cast<int4*>(dst)[0] = make_int4(v.rec, v.type, v.pos, v.len);
  reinterpret_cast<int4*>(dst)[1] = make_int4(v.q0, v.q1, v.first, v.count);
}
static __global__ void __launch_bounds__(256)
k_shard_split(RecCols R, const int32_t* __restrict__ ref_end, const int32_t* __restrict__ readlen, ShardState* st,
              vsv_sig* __restrict__ sig_reads, int64_t cap_r) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R.n_rec) return;
  const int32_t nx = R.next[r];
  if (nx < 0 || nx >= R.n_rec) return;
  const int rv1 = R.flag ? ((R.flag[r] >> 4) & 1) : 0, rv2 = R.flag ? ((R.flag[nx] >> 4) & 1) : 0;
  vsv_sig o;
  o.rec = (int32_t)r; o.first = nx; o.count = 1; o.type = 0; o.pos = 0; o.len = 0; o.q0 = 0; o.q1 = 0;
  if (!split_rule(VSV_SPLIT_READS, 0, ref_end[r], readlen[r], last_op_of(R, r), R.mapq[r], rv1, R.pos[nx],
                  first_op_of(R, nx), R.mapq[nx], rv2, o))
    return;
  const int64_t base = min((int64_t)st->nout[0], cap_r);
  const int64_t slot = base + (int64_t)atomicAdd(&st->n_split, 1ull);
  if (slot < cap_r) put_sig(sig_reads + slot, o);
  else atomicOr(&st->overflow, (unsigned)VSV_OVF_READS);
}

// ---- P10: SA-tag analysis (SE:193-371) -------------------------------------------------------------
struct SaCols {
  const int32_t* rec;
  const int64_t* off;
  const int32_t *chrom, *pos, *first_s, *last_s, *bias;
  const uint8_t *strand, *mapq;
  int64_t n_read;
  int32_t self_chrom;
};
__device__ __forceinline__ void append_row(ShardState* st, int type, vsv_sig* rows_del, int64_t cap_d, vsv_sig* rows_ins,
                                           int64_t cap_i, const vsv_sig& v) {
  if (type == VSV_DEL) {
    const unsigned long long s = atomicAdd(&st->n_del, 1ull);
    if ((int64_t)s < cap_d) put_sig(rows_del + s, v); else atomicOr(&st->overflow, (unsigned)VSV_OVF_DEL);
  } else {
    const unsigned long long s = atomicAdd(&st->n_ins, 1ull);
    if ((int64_t)s < cap_i) put_sig(rows_ins + s, v); else atomicOr(&st->overflow, (unsigned)VSV_OVF_INS);
  }
}
static __global__ void __launch_bounds__(128)
k_shard_sa(RecCols R, SaCols A, const int32_t* __restrict__ ref_end, const uint8_t* __restrict__ in_bed, int min_mapq,
           int min_read_len, int sv_size, int max_size, int max_parts, ShardState* st, vsv_sig* __restrict__ rows_del,
           int64_t cap_d, vsv_sig* __restrict__ rows_ins, int64_t cap_i) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_read) return;
  const int32_t r = A.rec[i];
  if (r < 0 || r >= R.n_rec) return;
  const int fl = R.flag ? R.flag[r] : 0;
  const int32_t qlen = R.qlen[r];
  if ((fl != 0 && fl != 16) || qlen < min_read_len || !in_bed[r]) return;   // SE:439,495-511
  SaSeg sp[SA_MAXSEG];
  int k = 0;
  int mq_min = min_mapq;
  if ((int)R.mapq[r] >= min_mapq) {   // the primary alignment itself (SE:486-489,507-511)
    const uint32_t f = first_op_of(R, r), l = last_op_of(R, r);
    const int64_t left = ((f & 15u) == 4u || (f & 15u) == 5u) ? (int64_t)(f >> 4) : 0;
    const int64_t right = ((l & 15u) == 4u || (l & 15u) == 5u) ? (int64_t)(l >> 4) : 0;
    SaSeg s;
    s.rs = R.pos[r]; s.re = ref_end[r]; s.chr = A.self_chrom;
    if (fl == 0) { s.qs = left; s.qe = qlen - right; s.st = 0; }
    else { s.qs = right; s.qe = qlen - left; s.st = 1; }
    sa_insert(sp, k++, s);
    mq_min = 0;                       // SE:348-350
  }
  const int64_t a = A.off[i], b = A.off[i + 1];
  bool too_many = false;
  for (int64_t e = a; e < b; ++e) {
    if ((int)A.mapq[e] < mq_min) continue;
    if (k >= max_parts || k >= SA_MAXSEG) { too_many = true; break; }   // SE:370: more than max_parts -> nothing
    SaSeg s;
    s.rs = A.pos[e]; s.re = (int64_t)A.pos[e] + A.bias[e]; s.chr = A.chrom[e];
    if (A.strand[e] == 0) { s.qs = A.first_s[e]; s.qe = (int64_t)qlen - A.last_s[e]; s.st = 0; }
    else { s.qs = A.last_s[e]; s.qe = (int64_t)qlen - A.first_s[e]; s.st = 1; }
    sa_insert(sp, k++, s);
  }
  if (too_many || k < 2) return;
  vsv_sig loc[2 * SA_MAXSEG];
  SaEmit em;
  em.dst = loc; em.rec = r; em.n = 0; em.rev = 0; em.sv_size = sv_size; em.max_size = max_size;
  sa_analyze(sp, k, (int64_t)qlen, em);
  const int32_t uid0 = -1 - (int32_t)(2 * a + i);   // read i owns 2 * entries + 1 ids
  for (int e = 0; e < em.n; ++e) {
    vsv_sig v = loc[e];
    const int type = v.type;
    v.type = type | (v.first ? ROW_SA_REV : 0);
    v.first = uid0 - e;
    v.count = 0;
    append_row(st, type, rows_del, cap_d, rows_ins, cap_i, v);
    atomicAdd(&st->n_sa, 1ull);
  }
}

// ---- P11: in-read merge of the raw cute signatures (SE:373-435) -> `.sigs` rows ----------------------
// One thread per record run (head = first signature of a record); count, block-aggregated slot
// reservation (two atomics per CTA), write.  `first` of a row = index of its first raw signature.
#define MERGE_THREADS 256
#define MERGE_HALO 96
// the first half (rec, type, pos, len) of raw signature k: from the CTA's shared-memory window when it is inside
struct SigWindow {
  const int4 *s_a, *s_b;  // rows [lo, hi) staged: (rec, type, pos, len) and (q0, q1, first, count)
  int64_t lo, hi;
  const vsv_sig* g;
  __device__ __forceinline__ int4 a(int64_t k) const {
    return (k >= lo && k < hi) ? s_a[k - lo] : __ldg(reinterpret_cast<const int4*>(g + k));
  }
  __device__ __forceinline__ int4 b(int64_t k) const {
    return (k >= lo && k < hi) ? s_b[k - lo] : __ldg(reinterpret_cast<const int4*>(g + k) + 1);
  }
};
template <bool FULL, typename FD, typename FI>
__device__ __forceinline__ void merge_walk_rows(const SigWindow& W, int64_t i0, int64_t i1, FD fd, FI fi) {
  MState sd, si;
  sd.open = si.open = false;
  for (int64_t k = i0; k < i1; ++k) {
    const int4 a = W.a(k);
    int4 b = make_int4(0, 0, 0, 0);
    if (FULL) b = W.b(k);
    vsv_sig n;
    n.rec = a.x; n.type = a.y; n.pos = a.z; n.len = a.w; n.q0 = b.x; n.q1 = b.y; n.first = b.z; n.count = b.w;
    if (n.type == VSV_DEL) mstate_push(sd, n, (int32_t)k, VSV_MODE_CUTE, VSV_DEL, fd);
    else mstate_push(si, n, (int32_t)k, VSV_MODE_CUTE, VSV_INS, fi);
  }
  if (sd.open) fd(sd.cur);
  if (si.open) fi(si.cur);
}
// The CTA stages the decision fields of its 256 signatures (+ a halo: a run is a handful long) in shared memory with
// coalesced loads and COMPACTS the record heads (first signature of a record: ~1 in 8) into a dense list, so the threads
// that walk a run are the first n_heads threads of the CTA - full warps instead of ~4 active lanes per warp (the
// thread-per-signature version spent 200 us of issue slots on idle lanes).  Heads walk their run twice in shared
// memory: count, then write after a block-aggregated slot reservation (two atomics per CTA).
static __global__ void __launch_bounds__(MERGE_THREADS)
k_shard_merge(const vsv_sig* __restrict__ s, ShardState* st, Caps cap, const uint8_t* __restrict__ in_bed,
              vsv_sig* __restrict__ rows_del, vsv_sig* __restrict__ rows_ins) {
  __shared__ int4 s_a[MERGE_THREADS + MERGE_HALO + 1], s_b[MERGE_THREADS + MERGE_HALO + 1];
  __shared__ int32_t s_head[MERGE_THREADS];
  __shared__ unsigned int s_wd[MERGE_THREADS / 32], s_wi[MERGE_THREADS / 32], s_wh[MERGE_THREADS / 32];
  __shared__ unsigned long long s_base[2];
  __shared__ unsigned int s_nhead;
  const int64_t n = min((int64_t)st->nout[1], cap.c);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t base = (int64_t)blockIdx.x * MERGE_THREADS; base < n; base += (int64_t)gridDim.x * MERGE_THREADS) {
    SigWindow W;
    W.s_a = s_a; W.s_b = s_b; W.g = s;
    W.lo = max(base - 1, (int64_t)0);
    W.hi = min(base + MERGE_THREADS + MERGE_HALO, n);
    for (int64_t k = W.lo + threadIdx.x; k < W.hi; k += MERGE_THREADS) {
      s_a[k - W.lo] = __ldg(reinterpret_cast<const int4*>(s + k));
      s_b[k - W.lo] = __ldg(reinterpret_cast<const int4*>(s + k) + 1);
    }
    __syncthreads();
    // ---- compact the heads of this tile -----------------------------------------------------------------
    const int64_t i = base + threadIdx.x;
    bool head = false;
    if (i < n) {
      const int32_t r = W.a(i).x;
      head = (i == 0 || W.a(i - 1).x != r) && in_bed[r];
    }
    const unsigned int hm = __ballot_sync(0xffffffffu, head);
    if (lane == 0) s_wh[warp] = __popc(hm);
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int t = 0;
      for (int w = 0; w < MERGE_THREADS / 32; ++w) { const unsigned int c = s_wh[w]; s_wh[w] = t; t += c; }
      s_nhead = t;
    }
    __syncthreads();
    if (head) s_head[s_wh[warp] + __popc(hm & ((1u << lane) - 1u))] = (int32_t)threadIdx.x;
    __syncthreads();
    const unsigned int nh = s_nhead;
    // ---- dense walk: thread t owns head number t ------------------------------------------------------------
    unsigned int cd = 0, ci = 0;
    int64_t h = 0, e = 0;
    if (threadIdx.x < nh) {
      h = base + s_head[threadIdx.x];
      const int32_t r = W.a(h).x;
      e = h + 1;
      while (e < n && W.a(e).x == r) ++e;
      merge_walk_rows<false>(W, h, e, [&](const vsv_sig&) { ++cd; }, [&](const vsv_sig&) { ++ci; });
    }
    unsigned int xd = cd, xi = ci;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int a = __shfl_up_sync(0xffffffffu, xd, d), b = __shfl_up_sync(0xffffffffu, xi, d);
      if (lane >= d) { xd += a; xi += b; }
    }
    if (lane == 31) { s_wd[warp] = xd; s_wi[warp] = xi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int td = 0, ti = 0;
      for (int w = 0; w < MERGE_THREADS / 32; ++w) { const unsigned int a = s_wd[w], b = s_wi[w]; s_wd[w] = td; s_wi[w] = ti; td += a; ti += b; }
      // both reservations in flight together (the results are consumed after the second is issued)
      const unsigned long long bd = td ? atomicAdd(&st->n_del, (unsigned long long)td) : 0ull;
      const unsigned long long bi = ti ? atomicAdd(&st->n_ins, (unsigned long long)ti) : 0ull;
      s_base[0] = bd;
      s_base[1] = bi;
    }
    __syncthreads();
    if (cd | ci) {
      int64_t od = (int64_t)s_base[0] + s_wd[warp] + (xd - cd), oi = (int64_t)s_base[1] + s_wi[warp] + (xi - ci);
      bool ovd = false, ovi = false;
      merge_walk_rows<true>(W, h, e,
                            [&](const vsv_sig& v) { if (od < cap.d) put_sig(rows_del + od, v); else ovd = true; ++od; },
                            [&](const vsv_sig& v) { if (oi < cap.i) put_sig(rows_ins + oi, v); else ovi = true; ++oi; });
      if (ovd) atomicOr(&st->overflow, (unsigned)VSV_OVF_DEL);
      if (ovi) atomicOr(&st->overflow, (unsigned)VSV_OVF_INS);
    }
    __syncthreads();
  }
}

// ---- P12: keys of the combined sort -----------------------------------------------------------------
static __global__ void __launch_bounds__(256)
k_shard_keys(ShardState* st, Caps cap, const vsv_sig* __restrict__ sig_reads, const vsv_sig* __restrict__ rows_del,
             const vsv_sig* __restrict__ rows_ins, const int32_t* __restrict__ ref_end, uint32_t* __restrict__ keys,
             uint32_t* __restrict__ vals) {
  const Tables T = tables_of(st, cap);
  const int64_t total = cap.r + cap.d + cap.i + cap.n_rec;
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t key = KEY_SENTINEL, val = 0, tag = 0;
    int32_t p = 0;
    bool live = false;
    int64_t j = i;
    if (j < cap.r) {
      if (j < T.n_r) { p = sig_reads[j].pos; tag = TAG_READS; live = true; }
    } else if ((j -= cap.r) < cap.d) {
      if (j < T.n_d) { p = rows_del[j].pos; tag = TAG_DEL; live = true; }
    } else if ((j -= cap.d) < cap.i) {
      if (j < T.n_i) { p = rows_ins[j].pos; tag = TAG_INS; live = true; }
    } else {
      j -= cap.i;
      p = ref_end[j]; tag = TAG_END; live = true;
    }
    if (live) {
      if (p < 0) { bad = true; p = 0; }
      if ((uint32_t)p > POS_MASK) { bad = true; p = (int32_t)POS_MASK; }
      key = (tag << POS_BITS) | (uint32_t)p;
      val = (uint32_t)j;
    }
    keys[i] = key;
    vals[i] = val;
  }
  if (bad) atomicOr(&st->overflow, (unsigned)VSV_OVF_POS);
}

// ---- P14: sorted (key, value) -> struct-of-arrays columns of the four tables ----------------------------
struct Sorted {
  int32_t *r_pos, *r_len;                    // read-signature table
  int32_t *t_pos[2], *t_len[2], *t_idx[2];   // DEL / INS rows in `.sigs` order; idx = row in rows_del / rows_ins
  int32_t* end;                              // read ends ascending
};
static __global__ void __launch_bounds__(256)
k_shard_unpack(const ShardState* st, Caps cap, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
               const vsv_sig* __restrict__ sig_reads, const vsv_sig* __restrict__ rows_del,
               const vsv_sig* __restrict__ rows_ins, Sorted S) {
  const Tables T = tables_of(st, cap);
  const int64_t total = T.off[4];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t key = keys[i], v = vals[i];
    const uint32_t tag = key >> POS_BITS;
    const int32_t p = (int32_t)(key & POS_MASK);
    if (tag == TAG_READS) {
      S.r_pos[i] = p; S.r_len[i] = sig_reads[v].len;
    } else if (tag == TAG_DEL) {
      const int64_t j = i - T.off[1];
      S.t_pos[0][j] = p; S.t_idx[0][j] = (int32_t)v; S.t_len[0][j] = rows_del[v].len;
    } else if (tag == TAG_INS) {
      const int64_t j = i - T.off[2];
      S.t_pos[1][j] = p; S.t_idx[1][j] = (int32_t)v; S.t_len[1][j] = rows_ins[v].len;
    } else if (tag == TAG_END) {
      S.end[i - T.off[3]] = p;
    }
  }
}

// ---- P15/16: `.sigs` tie order inside runs of equal position (len as a decimal string, SE:637-638) ------
#define TIE_SMALL 32
struct TieWork { int32_t table, start, len, pad; };
static __global__ void __launch_bounds__(256)
k_shard_tie_small(ShardState* st, Caps cap, Sorted S, TieWork* __restrict__ work, unsigned int work_cap) {
  const Tables T = tables_of(st, cap);
  const int tb = blockIdx.y;
  const int64_t n = tb == 0 ? T.n_d : T.n_i;
  const int32_t* __restrict__ pos = S.t_pos[tb];
  int32_t* __restrict__ len = S.t_len[tb];
  int32_t* __restrict__ idx = S.t_idx[tb];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t p = pos[i];
    if (i > 0 && pos[i - 1] == p) continue;   // not a run head
    int64_t e = i + 1;
    while (e < n && e - i <= TIE_SMALL && pos[e] == p) ++e;
    const int L = (int)(e - i);
    if (L == 1) continue;
    if (L > TIE_SMALL) {
      while (e < n && pos[e] == p) ++e;
      const unsigned int w = atomicAdd(&st->n_work, 1u);
      if (w < work_cap) work[w] = TieWork{tb, (int32_t)i, (int32_t)(e - i), 0};
      continue;
    }
    int32_t id[TIE_SMALL], ln[TIE_SMALL];
    unsigned long long key[TIE_SMALL];
    for (int k = 0; k < L; ++k) {   // stable insertion sort
      const int32_t x = idx[i + k], l = len[i + k];
      const unsigned long long kx = len_string_key(l);
      int m = k;
      while (m > 0 && key[m - 1] > kx) { key[m] = key[m - 1]; id[m] = id[m - 1]; ln[m] = ln[m - 1]; --m; }
      key[m] = kx; id[m] = x; ln[m] = l;
    }
    for (int k = 0; k < L; ++k) { idx[i + k] = id[k]; len[i + k] = ln[k]; }
  }
}
static __global__ void __launch_bounds__(256)
k_shard_tie_long(const ShardState* st, Sorted S, const TieWork* __restrict__ work, unsigned int work_cap,
                 unsigned long long* __restrict__ key_tmp, int32_t* __restrict__ idx_tmp, int32_t* __restrict__ len_tmp,
                 int64_t tmp_off_ins) {
  const unsigned int nw = min(st->n_work, work_cap);
  for (unsigned int w = blockIdx.x; w < nw; w += gridDim.x) {
    const TieWork tw = work[w];
    int32_t* len = S.t_len[tw.table];
    int32_t* idx = S.t_idx[tw.table];
    const int64_t a = tw.start, t = (tw.table ? tmp_off_ins : 0) + a;
    const int L = tw.len;
    for (int k = threadIdx.x; k < L; k += blockDim.x) {
      idx_tmp[t + k] = idx[a + k];
      len_tmp[t + k] = len[a + k];
      key_tmp[t + k] = len_string_key(len[a + k]);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < L; k += blockDim.x) {
      const unsigned long long kk = key_tmp[t + k];
      int rank = 0;
      for (int m = 0; m < L; ++m) {
        const unsigned long long km = key_tmp[t + m];
        rank += (km < kk || (km == kk && m < k)) ? 1 : 0;
      }
      idx[a + rank] = idx_tmp[t + k];
      len[a + rank] = len_tmp[t + k];
    }
    __syncthreads();
  }
}

// ---- P17: `sort -u` (SE:637-638): a row equal to an EARLIER row in (pos, len, read) is a duplicate --------
// DEL lines carry nothing else, so the row is dropped here; INS lines also carry the sequence, which only the
// host has: the pair is reported (VSV_TAB_INS_DUPS) and dropped only when the host confirms it
// (vsv_shard_drop_ins).  Runs of equal (pos, len) are a handful of reads long.
struct GroupIn {
  const vsv_sig* rows[2];
  const int32_t* name;
  const int32_t* drop;   // confirmed INS duplicates: uids ascending
  int64_t n_drop;
  int32_t min_len[2];    // rows shorter than this are not grouped (GTI:81: 30 for INS; DEL: 0)
};
__device__ __forceinline__ bool uid_dropped(const GroupIn& G, int32_t uid) {
  int64_t lo = 0, hi = G.n_drop;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    const int32_t v = G.drop[mid];
    if (v == uid) return true;
    if (v < uid) lo = mid + 1; else hi = mid;
  }
  return false;
}
static __global__ void __launch_bounds__(256)
k_shard_keep(ShardState* st, Caps cap, Sorted S, GroupIn G, uint8_t* __restrict__ keep_del, uint8_t* __restrict__ keep_ins,
             int2* __restrict__ dup_list) {
  const Tables T = tables_of(st, cap);
  const int tb = blockIdx.y;
  const int64_t n = tb == 0 ? T.n_d : T.n_i;
  const int32_t* __restrict__ pos = S.t_pos[tb];
  const int32_t* __restrict__ len = S.t_len[tb];
  const int32_t* __restrict__ idx = S.t_idx[tb];
  const vsv_sig* __restrict__ rows = G.rows[tb];
  uint8_t* keep = tb == 0 ? keep_del : keep_ins;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint8_t k = 1;
    if (G.name != nullptr && i > 0 && pos[i - 1] == pos[i] && len[i - 1] == len[i]) {
      const int32_t p = pos[i], l = len[i];
      const vsv_sig me = rows[idx[i]];
      const int32_t nm = G.name[me.rec];
      for (int64_t j = i - 1; j >= 0 && pos[j] == p && len[j] == l; --j) {
        const vsv_sig o = rows[idx[j]];
        if (G.name[o.rec] != nm) continue;
        if (tb == 0) {
          k = 0;
          atomicAdd(&st->n_del_dup, 1ull);
        } else if (uid_dropped(G, me.first)) {
          k = 0;
        } else {
          const unsigned long long s = atomicAdd(&st->n_ins_dup, 1ull);
          if (s < DUPLIST_CAP) dup_list[s] = make_int2(me.first, o.first);
          else atomicOr(&st->overflow, (unsigned)VSV_OVF_DUPLIST);
        }
        break;
      }
    }
    keep[i] = k;
  }
}

// ---- P18: unique signatures (GTD:65-86 / GTI:67-101) + running maximum of the DEL ends: one single-pass scan
struct SumMax {
  __device__ __forceinline__ int2 operator()(const int2& a, const int2& b) const { return make_int2(a.x + b.x, max(a.y, b.y)); }
};
struct Unique {
  int32_t *pos[2], *len[2], *cnt[2];
};
static __global__ void __launch_bounds__(GROUP_THREADS)
k_shard_group(ShardState* st, Caps cap, Sorted S, GroupIn G, const uint8_t* __restrict__ keep_del,
              const uint8_t* __restrict__ keep_ins, LbDesc* __restrict__ desc_del, LbDesc* __restrict__ desc_ins, Unique U,
              int32_t* __restrict__ pmax_del) {
  __shared__ int2 s_w[GROUP_THREADS / 32];
  __shared__ unsigned int s_tile;
  __shared__ int2 s_excl;
  const Tables T = tables_of(st, cap);
  const int tb = blockIdx.y;
  const int64_t n = tb == 0 ? T.n_d : T.n_i;
  const int32_t* __restrict__ pos = S.t_pos[tb];
  const int32_t* __restrict__ len = S.t_len[tb];
  const uint8_t* __restrict__ keep = tb == 0 ? keep_del : keep_ins;
  LbDesc* desc = tb == 0 ? desc_del : desc_ins;
  if (threadIdx.x == 0) s_tile = atomicAdd(&st->ticket[tb], 1u);
  __syncthreads();
  const unsigned int tile = s_tile;
  const int64_t i0 = (int64_t)tile * GROUP_TILE + (int64_t)threadIdx.x * GROUP_ITEMS;   // GROUP_ITEMS consecutive rows per thread
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const SumMax op;
  const int2 ident = make_int2(0, INT32_MIN);
  int32_t p[GROUP_ITEMS], l[GROUP_ITEMS];
  int2 v[GROUP_ITEMS];
  int32_t pp = 0, pl = 0;
  bool have_prev = false;
  if (i0 > 0 && i0 < n) { pp = pos[i0 - 1]; pl = len[i0 - 1]; have_prev = true; }
  int2 agg = ident;
#pragma unroll
  for (int k = 0; k < GROUP_ITEMS; ++k) {
    const int64_t i = i0 + k;
    v[k] = ident;
    p[k] = 0; l[k] = 0;
    if (i < n) {
      p[k] = pos[i]; l[k] = len[i];
      const bool head = (!have_prev || pp != p[k] || pl != l[k]) && l[k] >= G.min_len[tb];
      v[k] = make_int2(head ? 1 : 0, (int32_t)min((int64_t)p[k] + l[k], (int64_t)INT32_MAX));
      pp = p[k]; pl = l[k]; have_prev = true;
    }
    agg = op(agg, v[k]);
  }
  int2 inc = agg;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int2 u;
    u.x = __shfl_up_sync(0xffffffffu, inc.x, d); u.y = __shfl_up_sync(0xffffffffu, inc.y, d);
    if (lane >= d) inc = op(u, inc);
  }
  int2 prevl;   // inclusive value of the previous lane == this lane's exclusive prefix inside the warp
  prevl.x = __shfl_up_sync(0xffffffffu, inc.x, 1); prevl.y = __shfl_up_sync(0xffffffffu, inc.y, 1);
  if (lane == 31) s_w[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int2 run = ident;
    if (lane == 0) {
      for (int w = 0; w < GROUP_THREADS / 32; ++w) { const int2 a = s_w[w]; s_w[w] = run; run = op(run, a); }
    }
    run.x = __shfl_sync(0xffffffffu, run.x, 0); run.y = __shfl_sync(0xffffffffu, run.y, 0);
    const int2 ex = lookback_exclusive_warp(desc, tile, run, ident, op);
    if (lane == 0) {
      s_excl = ex;
      // the last tile that holds rows publishes the unique count
      if (n > 0 && (int64_t)tile == (n - 1) / GROUP_TILE) st->n_u[tb] = (unsigned long long)(ex.x + run.x);
    }
  }
  __syncthreads();
  if (i0 >= n) return;
  int2 run = op(s_excl, s_w[warp]);   // tiles before + warps before + lanes before
  if (lane > 0) run = op(run, prevl);
#pragma unroll
  for (int k = 0; k < GROUP_ITEMS; ++k) {
    const int64_t i = i0 + k;
    if (i >= n) break;
    run = op(run, v[k]);
    if (tb == 0) pmax_del[i] = run.y;
    if (v[k].x) {
      int32_t c = 0;
      for (int64_t e = i; e < n && pos[e] == p[k] && len[e] == l[k]; ++e) c += keep[e];
      const int64_t u = (int64_t)run.x - 1;
      U.pos[tb][u] = p[k]; U.len[tb][u] = l[k]; U.cnt[tb][u] = c;
    }
  }
}

// ---- per-candidate joins -----------------------------------------------------------------------------------
#define JOIN_GROUP 8
struct JoinParams {
  int what;
  int32_t fp_max_comp, fp_max_shift, cov_flank;
  double min_size_sim, max_shift_ratio, min_shift;
};
struct CandOut {
  int32_t* fp;
  int64_t* cov;
  int32_t *sup, *lb, *ub;
};
// One WARP per candidate.  A candidate needs six binary searches (two window bounds in each of three tables); they run
// on six lanes at the same time and are broadcast, then all 32 lanes stride over the windows.  (One 8-lane group doing
// the searches one after the other was latency bound: 130 us for 3000 candidates.)
static __global__ void __launch_bounds__(256)
k_cand_joins(const ShardState* st, Caps cap, Sorted S, Unique U, const uint8_t* __restrict__ keep_del,
             const uint8_t* __restrict__ keep_ins, const int32_t* __restrict__ pmax_del, const int32_t* __restrict__ cpos,
             const int32_t* __restrict__ clen, const int32_t* __restrict__ ctype, int64_t n_cand, JoinParams jp,
             CandOut O) {
  const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (k >= n_cand) return;   // whole warps exit together
  const Tables T = tables_of(st, cap);
  const int64_t p = cpos[k];
  const int32_t cl = clen[k];
  const int tb = ctype[k] == VSV_INS ? 1 : 0;
  const int64_t nu = min((int64_t)st->n_u[tb], tb == 0 ? cap.d : cap.i);
  const double sv = (double)cl;
  const double shift = fmax(sv * jp.max_shift_ratio, jp.min_shift);
  const int64_t cov_lb = p - jp.cov_flank, cov_rb = (tb == 1 ? p : p + (int64_t)cl) + jp.cov_flank;
  const int32_t lo32 = (int32_t)max(cov_lb, (int64_t)INT32_MIN), hi32 = (int32_t)min(cov_rb, (int64_t)INT32_MAX);
  int64_t r = 0;
  switch (lane) {
    case 0: if (jp.what & VSV_JOIN_FP) r = lower_bound_i32(S.r_pos, 0, T.n_r, (int32_t)max(p - jp.fp_max_shift, (int64_t)INT32_MIN)); break;
    case 1: if (jp.what & VSV_JOIN_FP) r = upper_bound_i32(S.r_pos, 0, T.n_r, (int32_t)min(p + jp.fp_max_shift, (int64_t)INT32_MAX)); break;
    case 2: if (jp.what & VSV_JOIN_COV) r = tb == 1 ? lower_bound_i32(S.t_pos[1], 0, T.n_i, lo32) : lower_bound_i32(pmax_del, 0, T.n_d, lo32); break;
    case 3: if (jp.what & VSV_JOIN_COV) r = tb == 1 ? upper_bound_i32(S.t_pos[1], 0, T.n_i, hi32) : upper_bound_i32(S.t_pos[0], 0, T.n_d, hi32); break;
    case 4: if (jp.what & VSV_JOIN_SUPPORT) r = lower_bound_f64(U.pos[tb], 0, nu, (double)p - shift); break;
    case 5: if (jp.what & VSV_JOIN_SUPPORT) r = upper_bound_f64(U.pos[tb], 0, nu, (double)p + shift); break;
    default: break;
  }
  const int64_t fa = __shfl_sync(0xffffffffu, r, 0), fb = __shfl_sync(0xffffffffu, r, 1);
  const int64_t ca = __shfl_sync(0xffffffffu, r, 2), cb = __shfl_sync(0xffffffffu, r, 3);
  const int64_t sa = __shfl_sync(0xffffffffu, r, 4), sb = __shfl_sync(0xffffffffu, r, 5);
  if (jp.what & VSV_JOIN_FP) {      // eval_sig (FPF:106-123): any signature type counts
    int32_t s = 0;
    if (cl > jp.fp_max_comp) {
      s = (lane == 0) ? 60 : 0;
    } else {
      for (int64_t i = fa + lane; i < fb; i += 32) {
        const int32_t sl = __ldg(S.r_len + i);
        const int32_t mx = max(cl, sl);
        if (mx != 0 && 2 * (int64_t)min(cl, sl) >= mx) ++s;
      }
    }
    s = __reduce_add_sync(0xffffffffu, s);
    if (lane == 0) O.fp[k] = s;
  }
  if (jp.what & VSV_JOIN_COV) {
    long long s = 0;
    if (tb == 1) {                  // calc_ins_call_cov (CSS:81-125)
      for (int64_t i = ca + lane; i < cb; i += 32)
        if (keep_ins[i]) s += __ldg(S.t_len[1] + i);
    } else {                        // calc_del_call_cov (CSS:138-253): closed-interval overlap with the flanked call; rows before
      for (int64_t i = ca + lane; i < cb; i += 32) {   // `ca` (running max of the ends < lb) cannot reach the window
        const int32_t sl = __ldg(S.t_len[0] + i);
        if (keep_del[i] && (int64_t)__ldg(S.t_pos[0] + i) + sl >= cov_lb) s -= sl;
      }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if (lane == 0) O.cov[k] = s;
  }
  if (jp.what & VSV_JOIN_SUPPORT) {  // match_varlist_siglist window sum (GTD:92-137)
    const double max_size = sv / jp.min_size_sim, min_size = sv * jp.min_size_sim;
    int32_t s = 0;
    for (int64_t i = sa + lane; i < sb; i += 32) {
      const double L = (double)__ldg(U.len[tb] + i);
      if (min_size <= L && L <= max_size) s += __ldg(U.cnt[tb] + i);
    }
    s = __reduce_add_sync(0xffffffffu, s);
    if (lane == 0) { O.sup[k] = s; O.lb[k] = (int32_t)sa; O.ub[k] = (int32_t)max(sb, sa); }
  }
}

// last_match_idx (GTD:94-133): per type, the lower bound of the last earlier candidate with a non-empty window;
// single-pass scan over the candidates, then the double count at that index.
struct LastValid2 {
  __device__ __forceinline__ int2 operator()(const int2& a, const int2& b) const {
    return make_int2(b.x >= 0 ? b.x : a.x, b.y >= 0 ? b.y : a.y);
  }
};
static __global__ void __launch_bounds__(CHAIN_THREADS)
k_cand_chain(const ShardState* st, Caps cap, Unique U, const int32_t* __restrict__ clen, const int32_t* __restrict__ ctype,
             int64_t n_cand, int2 last_in, double min_size_sim, LbDesc* __restrict__ desc, unsigned int* __restrict__ ticket,
             int32_t* __restrict__ sup, const int32_t* __restrict__ lb, const int32_t* __restrict__ ub,
             int32_t* __restrict__ match_id, int2* __restrict__ last_out) {
  __shared__ int2 s_w[CHAIN_THREADS / 32];
  __shared__ unsigned int s_tile;
  __shared__ int2 s_excl;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const unsigned int tile = s_tile;
  const int64_t k = (int64_t)tile * CHAIN_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int tb = 0;
  int2 v = make_int2(-1, -1);
  if (k < n_cand) {
    tb = ctype[k] == VSV_INS ? 1 : 0;
    const int32_t vl = ub[k] > lb[k] ? lb[k] : -1;
    if (tb == 0) v.x = vl; else v.y = vl;
  }
  const LastValid2 op;
  int2 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int2 u;
    u.x = __shfl_up_sync(0xffffffffu, inc.x, d); u.y = __shfl_up_sync(0xffffffffu, inc.y, d);
    if (lane >= d) inc = op(u, inc);
  }
  if (lane == 31) s_w[warp] = inc;
  __syncthreads();
  // exclusive prefix of this candidate inside the warp = inclusive of the previous lane
  int2 prev;
  prev.x = __shfl_up_sync(0xffffffffu, inc.x, 1); prev.y = __shfl_up_sync(0xffffffffu, inc.y, 1);
  if (warp == 0) {
    int2 run = make_int2(-1, -1);
    if (lane == 0) {
      for (int w = 0; w < CHAIN_THREADS / 32; ++w) { const int2 a = s_w[w]; s_w[w] = run; run = op(run, a); }
    }
    run.x = __shfl_sync(0xffffffffu, run.x, 0); run.y = __shfl_sync(0xffffffffu, run.y, 0);
    const int2 ex = lookback_exclusive_warp(desc, tile, run, make_int2(-1, -1), op);
    if (lane == 0) {
      s_excl = ex;
      if (n_cand > 0 && (int64_t)tile == (n_cand - 1) / CHAIN_THREADS && last_out) *last_out = op(last_in, op(ex, run));
    }
  }
  __syncthreads();
  if (k >= n_cand) return;
  int2 pre = op(s_excl, s_w[warp]);
  if (lane > 0) pre = op(pre, prev);
  pre = op(last_in, pre);
  const int32_t idx0 = tb == 0 ? pre.x : pre.y;
  if (match_id) match_id[k] = idx0;
  const int64_t nu = min((int64_t)st->n_u[tb], tb == 0 ? cap.d : cap.i);
  if (idx0 >= 0 && idx0 >= lb[k] && idx0 < ub[k] && idx0 < nu) {   // both scans of the reference visit idx0 (GTD:112-130)
    const double sv = (double)clen[k];
    const double L = (double)U.len[tb][idx0];
    if (sv * min_size_sim <= L && L <= sv / min_size_sim) sup[k] += U.cnt[tb][idx0];
  }
}

// 8 lanes per candidate: the three binary searches of each of the (up to two) windows run on six lanes at once, the
// short scan of the reads that start inside a window is strided over the group (closed form: vsv_depth.cuh).
static __global__ void __launch_bounds__(256)
k_cand_genotype(const int32_t* __restrict__ rstart, const int32_t* __restrict__ rend, const int32_t* __restrict__ send,
                int64_t n_reads, const int32_t* __restrict__ pos, const int32_t* __restrict__ svlen,
                const int32_t* __restrict__ type, const int32_t* __restrict__ call_gt, const int32_t* __restrict__ nsup,
                int64_t n, GtParams gp, int32_t* __restrict__ depth2, double* __restrict__ ratio,
                int32_t* __restrict__ new_gt) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int g = threadIdx.x & 7;
  const int gbase = (threadIdx.x & 31) & ~7;
  const unsigned gmask = 0xFFu << gbase;
  if (i >= n) return;   // whole groups exit together
  const int32_t p = pos[i], L = svlen[i], tp = type[i];
  int32_t S[2], E[2];
  int nw = 1;
  if (tp == VSV_INS) { S[0] = p - gp.ins_flank; E[0] = p + gp.ins_flank; S[1] = 0; E[1] = 0; }
  else if (L <= gp.del_large) { S[0] = p; E[0] = p + L; S[1] = 0; E[1] = 0; }
  else { S[0] = p - gp.del_off; E[0] = S[0] + gp.del_win; S[1] = p + L + gp.del_off; E[1] = S[1] + gp.del_win; nw = 2; }
  // lane g: window g / 3, search g % 3 (0: #start < S, 1: #end <= E, 2: first start > E)
  int64_t r = 0;
  {
    const int w = g / 3, q = g % 3;
    if (w < nw && g < 6) {
      if (q == 0) r = lower_bound_i32(rstart, 0, n_reads, S[w]);
      else if (q == 1) r = upper_bound_i32(send, 0, n_reads, E[w]);
      else r = upper_bound_i32(rstart, 0, n_reads, E[w]);
    }
  }
  int32_t d[2] = {0, 0};
  for (int w = 0; w < 2; ++w) {
    const int64_t a = __shfl_sync(gmask, r, gbase + 3 * w), ue = __shfl_sync(gmask, r, gbase + 3 * w + 1),
                  b = __shfl_sync(gmask, r, gbase + 3 * w + 2);
    int32_t c = 0;
    if (w < nw) {
      for (int64_t j = a + g; j < b; j += 8) c += (__ldg(rend + j) <= E[w]) ? 1 : 0;   // empty when E < S (b <= a)
    }
#pragma unroll
    for (int s = 4; s >= 1; s >>= 1) c += __shfl_xor_sync(gmask, c, s);
    d[w] = w < nw ? (int32_t)(a - ue + c) : 0;
  }
  if (g != 0) return;
  const int32_t d2 = nw == 2 ? d[0] + d[1] : 2 * d[0];
  const int32_t ns = nsup[i];
  const double depth = (double)d2 / 2.0;  // GTD:168
  double rt;
  if (tp == VSV_DEL) rt = (d2 == 0) ? 1.0 : (double)ns / depth;                                     // GTD:213-217
  else rt = (d2 == 0) ? (ns == 0 ? CUDART_NAN : CUDART_INF) : (double)ns / depth;                   // GTI:228
  const int32_t gt = call_gt ? call_gt[i] : 2;
  int32_t out = gt;
  if (gt == 0 || gt == 1) {
    const double* t4 = tp == VSV_DEL ? gp.thr_del : gp.thr_ins;
    const bool large = L > gp.del_large;
    const double t = gt == 1 ? (large ? t4[0] : t4[1]) : (large ? t4[2] : t4[3]);
    if (!isnan(t)) {  // GTD:302-318
      if (rt > t) out = 1;
      else if (rt <= t) out = 0;
    }
  }
  depth2[i] = d2; ratio[i] = rt; new_gt[i] = out;
}

// ==================================================================================================
// host side
// ==================================================================================================
struct DBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct vsv_shard {
  vsv_ctx* ctx = nullptr;
  // inputs (device); `in_*` are owned copies in host-IO mode
  DBuf in_ops, in_raw, in_exc_idx, in_exc_val, in_op_off, in_pos, in_mapq, in_flag, in_qlen, in_name, in_next;
  DBuf in_sa_rec, in_sa_off, in_sa_chrom, in_sa_pos, in_sa_strand, in_sa_fs, in_sa_ls, in_sa_bias, in_sa_mapq;
  RecCols R{};
  SaCols A{};
  int64_t n_ops = 0;
  bool loaded = false, scanned = false;
  // bed
  DBuf bed0, bed1;
  int64_t n_bed = -1, chrom_len = 0;
  // capacities
  Caps cap{0, 0, 0, 0, 0};
  // scan temporaries
  DBuf info, first, starts, agg, excl, wstage, fstart, stage, blk;
  int64_t stage_cap = 0;
  // per record
  DBuf ref_end, readlen, in_bed;
  // tables
  DBuf sig_reads, sig_cute, rows_del, rows_ins;
  DBuf keys[2], vals[2], cub_tmp;
  DBuf r_pos, r_len, t_pos[2], t_len[2], t_idx[2], keep[2], pmax, send, u_pos[2], u_len[2], u_cnt[2];
  DBuf work, key_tmp, idx_tmp, len_tmp, dup_list, drop, desc;
  int64_t n_drop = 0;
  int64_t desc_tiles[2] = {0, 0};
  DBuf state;
  ShardState* h_state = nullptr;   // pinned copy read by vsv_shard_counts_get
  cudaEvent_t ev_state = nullptr;
  bool state_pending = false;
};

static int dbuf_ensure(vsv_ctx* ctx, DBuf& b, size_t bytes) {
  if (bytes == 0) bytes = 16;
  if (bytes <= b.cap) return VSV_OK;
  if (b.p) {
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
  }
  const size_t want = (bytes + 255) / 256 * 256;
  if (cudaMalloc(&b.p, want) != cudaSuccess) {
    cudaGetLastError();
    VSV_FAIL(ctx, VSV_E_NOMEM, "vsv_shard: device allocation of " + std::to_string(want) + " bytes failed");
  }
  b.cap = want;
  return VSV_OK;
}
static void dbuf_free(DBuf& b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
}

extern "C" int vsv_shard_create(vsv_ctx* ctx, vsv_shard** out) {
  if (!ctx || !out) return VSV_E_ARG;
  VSV_CUDA(ctx, cudaSetDevice(ctx->device));
  vsv_shard* s = new vsv_shard();
  s->ctx = ctx;
  if (cudaHostAlloc((void**)&s->h_state, sizeof(ShardState), cudaHostAllocDefault) != cudaSuccess ||
      cudaEventCreateWithFlags(&s->ev_state, cudaEventDisableTiming) != cudaSuccess) {
    delete s;
    VSV_FAIL(ctx, VSV_E_NOMEM, "vsv_shard_create: pinned state allocation failed");
  }
  memset(s->h_state, 0, sizeof(ShardState));
  if (!ctx->scan_attr_set) {
    VSV_CUDA(ctx, cudaFuncSetAttribute(k_scan_wt<VSV_MODE_FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)((size_t)SCAN_WARPS * (SCAN_STAGES * (WT_OPS * sizeof(uint32_t) + sizeof(uint64_t)) + 8 * sizeof(uint32_t)))));
    ctx->scan_attr_set = true;
  }
  *out = s;
  return VSV_OK;
}

extern "C" void vsv_shard_destroy(vsv_shard* s) {
  if (!s) return;
  cudaSetDevice(s->ctx->device);
  cudaStreamSynchronize(s->ctx->stream);
  DBuf* all[] = {&s->in_ops, &s->in_raw, &s->in_exc_idx, &s->in_exc_val, &s->in_op_off, &s->in_pos, &s->in_mapq, &s->in_flag,
                 &s->in_qlen, &s->in_name, &s->in_next, &s->in_sa_rec, &s->in_sa_off, &s->in_sa_chrom, &s->in_sa_pos,
                 &s->in_sa_strand, &s->in_sa_fs, &s->in_sa_ls, &s->in_sa_bias, &s->in_sa_mapq, &s->bed0, &s->bed1, &s->info,
                 &s->first, &s->starts, &s->agg, &s->excl, &s->wstage, &s->fstart, &s->stage, &s->blk, &s->ref_end, &s->readlen,
                 &s->in_bed, &s->sig_reads, &s->sig_cute, &s->rows_del, &s->rows_ins, &s->keys[0], &s->keys[1], &s->vals[0],
                 &s->vals[1], &s->cub_tmp, &s->r_pos, &s->r_len, &s->t_pos[0], &s->t_pos[1], &s->t_len[0], &s->t_len[1],
                 &s->t_idx[0], &s->t_idx[1], &s->keep[0], &s->keep[1], &s->pmax, &s->send, &s->u_pos[0], &s->u_pos[1],
                 &s->u_len[0], &s->u_len[1], &s->u_cnt[0], &s->u_cnt[1], &s->work, &s->key_tmp, &s->idx_tmp, &s->len_tmp,
                 &s->dup_list, &s->drop, &s->desc, &s->state};
  for (DBuf* b : all) dbuf_free(*b);
  if (s->h_state) cudaFreeHost(s->h_state);
  if (s->ev_state) cudaEventDestroy(s->ev_state);
  delete s;
}

extern "C" int vsv_shard_reserve(vsv_shard* s, int64_t cap_reads, int64_t cap_cute, int64_t cap_del, int64_t cap_ins) {
  if (!s) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  if (cap_reads < 0 || cap_cute < 0 || cap_del < 0 || cap_ins < 0 || cap_reads >= (int64_t(1) << 30) || cap_cute >= (int64_t(1) << 30) ||
      cap_del >= (int64_t(1) << 30) || cap_ins >= (int64_t(1) << 30))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_reserve: capacity out of range");
  if (cap_reads) s->cap.r = cap_reads;
  if (cap_cute) s->cap.c = cap_cute;
  if (cap_del) s->cap.d = cap_del;
  if (cap_ins) s->cap.i = cap_ins;
  return VSV_OK;
}

// copy (host-IO) or adopt (device-IO) one input column
template <typename T>
static int shard_in(vsv_shard* s, DBuf& b, const T* src, size_t n, const T** dev) {
  vsv_ctx* ctx = s->ctx;
  if (src == nullptr) { *dev = nullptr; return VSV_OK; }
  if (ctx->device_io) { *dev = src; return VSV_OK; }
  VSV_TRY(dbuf_ensure(ctx, b, n * sizeof(T)));
  if (n) VSV_CUDA(ctx, cudaMemcpyAsync(b.p, src, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  *dev = (const T*)b.p;
  return VSV_OK;
}

// expansion kernels of the staging formats live in vsv_stage2.cu
extern "C" int vsv_expand_ops16(vsv_ctx*, const uint16_t*, int64_t, const int64_t*, const uint32_t*, int64_t, uint32_t*);
extern "C" int vsv_expand_ops8(vsv_ctx*, const uint8_t*, int64_t, const int64_t*, const uint32_t*, int64_t, uint32_t*);
extern "C" int vsv_expand_ops12(vsv_ctx*, const uint8_t*, int64_t, const int64_t*, const uint32_t*, int64_t, uint32_t*);

extern "C" int vsv_shard_load(vsv_shard* s, const vsv_records* r) {
  if (!s || !r) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  VSV_TRY(vsv_begin(ctx));
  if (r->n_rec < 0 || r->n_ops < 0 || r->n_rec >= (int64_t(1) << 31) - 1 || (r->n_rec > 0 && (!r->op_off || !r->pos || !r->mapq || !r->qlen)) ||
      (r->n_ops > 0 && !r->ops) || r->n_exc < 0 || (r->n_exc > 0 && (!r->exc_idx || !r->exc_val)) || r->n_sa_read < 0 || r->n_sa_ent < 0)
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_load: bad argument");
  if (r->ops_format != 32 && r->ops_format != 16 && r->ops_format != 12 && r->ops_format != 8)
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_load: ops_format must be 32, 16, 12 or 8");
  if (vsv_cdiv(r->n_ops, WT_OPS) >= (int64_t(1) << 29)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_load: too many ops for one shard");
  if (r->n_sa_read > 0 && (!r->sa_rec || !r->sa_off || !r->flag || (r->n_sa_ent > 0 && (!r->sa_chrom || !r->sa_pos || !r->sa_strand ||
                                                                                         !r->sa_first_s || !r->sa_last_s || !r->sa_bias || !r->sa_mapq))))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_load: incomplete SA columns");
  s->loaded = s->scanned = false;
  const size_t n = (size_t)r->n_rec;
  RecCols R{};
  R.n_rec = r->n_rec;
  VSV_TRY(shard_in(s, s->in_op_off, r->op_off, n + 1, &R.op_off));
  VSV_TRY(shard_in(s, s->in_pos, r->pos, n, &R.pos));
  VSV_TRY(shard_in(s, s->in_mapq, r->mapq, n, &R.mapq));
  VSV_TRY(shard_in(s, s->in_flag, r->flag, n, &R.flag));
  VSV_TRY(shard_in(s, s->in_qlen, r->qlen, n, &R.qlen));
  VSV_TRY(shard_in(s, s->in_name, r->name_id, n, &R.name));
  VSV_TRY(shard_in(s, s->in_next, r->next_same_name, n, &R.next));
  if (r->ops_format == 32) {
    const uint32_t* src = (const uint32_t*)r->ops;
    VSV_TRY(shard_in(s, s->in_ops, src, (size_t)r->n_ops, &R.ops));
  } else {
    // staging format: raw bytes in, widened on the device into in_ops
    const int64_t raw_bytes = r->ops_format == 16 ? 2 * r->n_ops : (r->ops_format == 8 ? r->n_ops : vsv_ops12_bytes(r->n_ops));
    const uint8_t* raw = nullptr;
    const int64_t* ei = nullptr;
    const uint32_t* ev = nullptr;
    VSV_TRY(shard_in(s, s->in_raw, (const uint8_t*)r->ops, (size_t)raw_bytes, &raw));
    VSV_TRY(shard_in(s, s->in_exc_idx, r->exc_idx, (size_t)r->n_exc, &ei));
    VSV_TRY(shard_in(s, s->in_exc_val, r->exc_val, (size_t)r->n_exc * (r->ops_format == 12 ? 2 : 1), &ev));
    VSV_TRY(dbuf_ensure(ctx, s->in_ops, (size_t)r->n_ops * sizeof(uint32_t)));
    const bool was = ctx->device_io;
    ctx->device_io = true;   // the expansion entry points take the device copies made above
    int rc;
    if (r->ops_format == 16) rc = vsv_expand_ops16(ctx, (const uint16_t*)raw, r->n_ops, ei, ev, r->n_exc, (uint32_t*)s->in_ops.p);
    else if (r->ops_format == 8) rc = vsv_expand_ops8(ctx, raw, r->n_ops, ei, ev, r->n_exc, (uint32_t*)s->in_ops.p);
    else rc = vsv_expand_ops12(ctx, raw, r->n_ops, ei, ev, r->n_exc, (uint32_t*)s->in_ops.p);
    ctx->device_io = was;
    VSV_TRY(rc);
    R.ops = (const uint32_t*)s->in_ops.p;
  }
  if (r->n_ops > 0 && (reinterpret_cast<uintptr_t>(R.ops) & 15u)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_load: ops must be 16-byte aligned");
  SaCols A{};
  A.n_read = r->n_sa_read;
  A.self_chrom = r->sa_self_chrom;
  if (r->n_sa_read > 0) {
    const size_t m = (size_t)r->n_sa_read, e = (size_t)r->n_sa_ent;
    VSV_TRY(shard_in(s, s->in_sa_rec, r->sa_rec, m, &A.rec));
    VSV_TRY(shard_in(s, s->in_sa_off, r->sa_off, m + 1, &A.off));
    VSV_TRY(shard_in(s, s->in_sa_chrom, r->sa_chrom, e, &A.chrom));
    VSV_TRY(shard_in(s, s->in_sa_pos, r->sa_pos, e, &A.pos));
    VSV_TRY(shard_in(s, s->in_sa_strand, r->sa_strand, e, &A.strand));
    VSV_TRY(shard_in(s, s->in_sa_fs, r->sa_first_s, e, &A.first_s));
    VSV_TRY(shard_in(s, s->in_sa_ls, r->sa_last_s, e, &A.last_s));
    VSV_TRY(shard_in(s, s->in_sa_bias, r->sa_bias, e, &A.bias));
    VSV_TRY(shard_in(s, s->in_sa_mapq, r->sa_mapq, e, &A.mapq));
  }
  s->R = R;
  s->A = A;
  s->n_ops = r->n_ops;
  s->cap.n_rec = r->n_rec;
  s->loaded = true;
  return vsv_end(ctx);
}

extern "C" int vsv_shard_set_bed(vsv_shard* s, const int32_t* b0, const int32_t* b1, int64_t n, int64_t chrom_len) {
  if (!s) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  VSV_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n > 0 && (!b0 || !b1)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_set_bed: bad argument");
  s->n_bed = n < 0 ? -1 : n;
  s->chrom_len = chrom_len;
  if (n > 0) {
    VSV_TRY(dbuf_ensure(ctx, s->bed0, (size_t)n * sizeof(int32_t)));
    VSV_TRY(dbuf_ensure(ctx, s->bed1, (size_t)n * sizeof(int32_t)));
    VSV_CUDA(ctx, cudaMemcpyAsync(s->bed0.p, b0, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    VSV_CUDA(ctx, cudaMemcpyAsync(s->bed1.p, b1, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // the host arrays may be temporaries
  }
  return VSV_OK;
}

static Sorted sorted_of(vsv_shard* s) {
  Sorted S;
  S.r_pos = (int32_t*)s->r_pos.p; S.r_len = (int32_t*)s->r_len.p;
  for (int t = 0; t < 2; ++t) { S.t_pos[t] = (int32_t*)s->t_pos[t].p; S.t_len[t] = (int32_t*)s->t_len[t].p; S.t_idx[t] = (int32_t*)s->t_idx[t].p; }
  S.end = (int32_t*)s->send.p;
  return S;
}
static Unique unique_of(vsv_shard* s) {
  Unique U;
  for (int t = 0; t < 2; ++t) { U.pos[t] = (int32_t*)s->u_pos[t].p; U.len[t] = (int32_t*)s->u_len[t].p; U.cnt[t] = (int32_t*)s->u_cnt[t].p; }
  return U;
}
static GroupIn group_in_of(vsv_shard* s) {
  GroupIn G;
  G.rows[0] = (const vsv_sig*)s->rows_del.p; G.rows[1] = (const vsv_sig*)s->rows_ins.p;
  G.name = s->R.name;
  G.drop = (const int32_t*)s->drop.p; G.n_drop = s->n_drop;
  G.min_len[0] = 0; G.min_len[1] = s->ctx->params.gt_ins_min_len;
  return G;
}

// `sort -u` flags + unique signatures + DEL running max (re-run by vsv_shard_drop_ins)
static int enqueue_group(vsv_shard* s) {
  vsv_ctx* ctx = s->ctx;
  ShardState* st = (ShardState*)s->state.p;
  const Sorted S = sorted_of(s);
  const GroupIn G = group_in_of(s);
  const int64_t cap_max = std::max(s->cap.d, s->cap.i);
  const size_t desc_bytes = (size_t)(s->desc_tiles[0] + s->desc_tiles[1]) * sizeof(LbDesc);
  VSV_CUDA(ctx, cudaMemsetAsync(s->desc.p, 0, desc_bytes, ctx->stream));
  const unsigned gk = (unsigned)std::max<int64_t>(1, std::min<int64_t>(vsv_cdiv(cap_max, 256), (int64_t)ctx->sm_count * 8));
  VSV_LAUNCH(ctx, k_shard_keep, dim3(gk, 2), 256, 0, st, s->cap, S, G, (uint8_t*)s->keep[0].p, (uint8_t*)s->keep[1].p, (int2*)s->dup_list.p);
  const unsigned gt = (unsigned)std::max<int64_t>(1, vsv_cdiv(cap_max, GROUP_TILE));
  VSV_LAUNCH(ctx, k_shard_group, dim3(gt, 2), GROUP_THREADS, 0, st, s->cap, S, G, (const uint8_t*)s->keep[0].p, (const uint8_t*)s->keep[1].p,
             (LbDesc*)s->desc.p, (LbDesc*)s->desc.p + s->desc_tiles[0], unique_of(s), (int32_t*)s->pmax.p);
  VSV_CUDA(ctx, cudaMemcpyAsync(s->h_state, st, sizeof(ShardState), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaEventRecord(s->ev_state, ctx->stream));
  s->state_pending = true;
  return VSV_OK;
}

extern "C" int vsv_shard_scan(vsv_shard* s) {
  if (!s) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  VSV_TRY(vsv_begin(ctx));
  if (!s->loaded) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_scan: no records loaded");
  const vsv_params& P = ctx->params;
  const int64_t n_rec = s->R.n_rec, n_ops = s->n_ops;
  const int32_t n_wt = (int32_t)vsv_cdiv(n_ops, WT_OPS);
  // default capacities: signatures are well under 2 % of the ops
  if (s->cap.r == 0) s->cap.r = std::max<int64_t>(1 << 14, n_ops / 64);
  if (s->cap.c == 0) s->cap.c = std::max<int64_t>(1 << 14, n_ops / 64);
  if (s->cap.d == 0) s->cap.d = s->cap.c;
  if (s->cap.i == 0) s->cap.i = s->cap.c;
  const Caps cap = s->cap;
  const int64_t cap_total = cap.r + cap.d + cap.i + n_rec;
  if (cap_total >= (int64_t(1) << 31)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_scan: table capacities exceed 2^31 rows");
  const int grid = vsv_scan_grid(ctx, n_wt);
  const int64_t slack = (int64_t)grid * SCAN_WARPS * SCAN_CHUNK + SCAN_CHUNK;
  const int64_t scap = std::max(s->stage_cap, std::max(cap.r, cap.c) * 5 / 4 + slack);
  if (scap >= (int64_t(1) << 32) - 1) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_scan: staging capacity out of range");
  s->stage_cap = scap;
  // ---- buffers -----------------------------------------------------------------------------------
  const int64_t nb = vsv_cdiv((int64_t)n_wt + 1, TS_BLOCK);
  VSV_TRY(dbuf_ensure(ctx, s->state, sizeof(ShardState)));
  VSV_TRY(dbuf_ensure(ctx, s->info, (size_t)n_rec * 4));
  VSV_TRY(dbuf_ensure(ctx, s->first, ((size_t)n_wt + 2) * 4));
  VSV_TRY(dbuf_ensure(ctx, s->starts, ((size_t)n_wt + 1) * sizeof(int2)));
  VSV_TRY(dbuf_ensure(ctx, s->agg, ((size_t)n_wt + 1) * sizeof(TileState)));
  VSV_TRY(dbuf_ensure(ctx, s->excl, ((size_t)n_wt + 1) * sizeof(TileState)));
  VSV_TRY(dbuf_ensure(ctx, s->wstage, ((size_t)n_wt + 1) * sizeof(uint2)));
  VSV_TRY(dbuf_ensure(ctx, s->fstart, 2 * ((size_t)n_rec + 1) * sizeof(uint4)));
  VSV_TRY(dbuf_ensure(ctx, s->stage, (size_t)scap * sizeof(StageEntry)));
  VSV_TRY(dbuf_ensure(ctx, s->blk, (size_t)(nb + 1) * sizeof(TsDesc)));   // look-back descriptors; the slot after them holds the ticket
  VSV_TRY(dbuf_ensure(ctx, s->ref_end, (size_t)n_rec * 4));
  VSV_TRY(dbuf_ensure(ctx, s->readlen, (size_t)n_rec * 4));
  VSV_TRY(dbuf_ensure(ctx, s->in_bed, (size_t)n_rec));
  VSV_TRY(dbuf_ensure(ctx, s->sig_reads, (size_t)cap.r * sizeof(vsv_sig)));
  VSV_TRY(dbuf_ensure(ctx, s->sig_cute, (size_t)cap.c * sizeof(vsv_sig)));
  VSV_TRY(dbuf_ensure(ctx, s->rows_del, (size_t)cap.d * sizeof(vsv_sig)));
  VSV_TRY(dbuf_ensure(ctx, s->rows_ins, (size_t)cap.i * sizeof(vsv_sig)));
  for (int k = 0; k < 2; ++k) {
    VSV_TRY(dbuf_ensure(ctx, s->keys[k], (size_t)cap_total * 4));
    VSV_TRY(dbuf_ensure(ctx, s->vals[k], (size_t)cap_total * 4));
  }
  VSV_TRY(dbuf_ensure(ctx, s->r_pos, (size_t)cap.r * 4));
  VSV_TRY(dbuf_ensure(ctx, s->r_len, (size_t)cap.r * 4));
  const int64_t capt[2] = {cap.d, cap.i};
  for (int t = 0; t < 2; ++t) {
    VSV_TRY(dbuf_ensure(ctx, s->t_pos[t], (size_t)capt[t] * 4));
    VSV_TRY(dbuf_ensure(ctx, s->t_len[t], (size_t)capt[t] * 4));
    VSV_TRY(dbuf_ensure(ctx, s->t_idx[t], (size_t)capt[t] * 4));
    VSV_TRY(dbuf_ensure(ctx, s->keep[t], (size_t)capt[t]));
    VSV_TRY(dbuf_ensure(ctx, s->u_pos[t], (size_t)capt[t] * 4));
    VSV_TRY(dbuf_ensure(ctx, s->u_len[t], (size_t)capt[t] * 4));
    VSV_TRY(dbuf_ensure(ctx, s->u_cnt[t], (size_t)capt[t] * 4));
    s->desc_tiles[t] = vsv_cdiv(std::max(cap.d, cap.i), GROUP_TILE) + 1;
  }
  VSV_TRY(dbuf_ensure(ctx, s->pmax, (size_t)cap.d * 4));
  VSV_TRY(dbuf_ensure(ctx, s->send, (size_t)n_rec * 4));
  const unsigned work_cap = (unsigned)((cap.d + cap.i) / (TIE_SMALL + 1) + 2);
  VSV_TRY(dbuf_ensure(ctx, s->work, (size_t)work_cap * sizeof(TieWork)));
  VSV_TRY(dbuf_ensure(ctx, s->key_tmp, (size_t)(cap.d + cap.i) * 8));
  VSV_TRY(dbuf_ensure(ctx, s->idx_tmp, (size_t)(cap.d + cap.i) * 4));
  VSV_TRY(dbuf_ensure(ctx, s->len_tmp, (size_t)(cap.d + cap.i) * 4));
  VSV_TRY(dbuf_ensure(ctx, s->dup_list, (size_t)DUPLIST_CAP * sizeof(int2)));
  VSV_TRY(dbuf_ensure(ctx, s->drop, 16));
  VSV_TRY(dbuf_ensure(ctx, s->desc, (size_t)(s->desc_tiles[0] + s->desc_tiles[1]) * sizeof(LbDesc)));
  size_t cub_bytes = 0;
  VSV_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                                (uint32_t*)nullptr, (int)cap_total, 0, 32, ctx->stream));
  VSV_TRY(dbuf_ensure(ctx, s->cub_tmp, cub_bytes));
  s->n_drop = 0;

  ShardState* st = (ShardState*)s->state.p;
  ScanParams p;
  memset(&p, 0, sizeof(p));
  typedef ModeTraits<VSV_MODE_FUSED> M;
  constexpr uint32_t RARE = M::X1 | M::X2 | M::X3;
  p.ops = s->R.ops; p.op_off = s->R.op_off; p.pos = s->R.pos;
  p.rec_info = (const uint32_t*)s->info.p;
  p.n_rec = n_rec; p.n_ops = n_ops; p.n_wt = n_wt;
  p.min_len_a = P.reads_min_len; p.min_len_b = P.cute_min_len;
  p.min_w = (uint32_t)std::min(P.reads_min_len, P.cute_min_len) << 4;
  p.one = 1u;
  p.tab[0] = tab_word(M::REF, 0, 0xFFu, 0, 0); p.tab[1] = tab_word(M::REF, 4, 0xFFu, 0, 0);
  p.tab[2] = tab_word(M::QRY, 0, 0xFFu, 0, 0); p.tab[3] = tab_word(M::QRY, 4, 0xFFu, 0, 0);
  p.tab[4] = tab_word(6u, 0, 0xFFu, RARE, 0x01u); p.tab[5] = tab_word(6u, 4, 0xFFu, RARE, 0x01u);
  p.tab[6] = tab_word(6u, 0, 0xFFu, RARE, 0x80u); p.tab[7] = tab_word(6u, 4, 0xFFu, RARE, 0x80u);
  p.wt_starts = (const int2*)s->starts.p; p.wt_agg = (TileState*)s->agg.p; p.wt_stage = (uint2*)s->wstage.p;
  p.fstart = (uint4*)s->fstart.p;
  p.stage = (StageEntry*)s->stage.p; p.stage_cap = (uint32_t)scap;
  p.stage_ctr = st->stage_ctr;
  TileState* d_excl = (TileState*)s->excl.p;
  BedParams bed;
  bed.b0 = (const int32_t*)s->bed0.p; bed.b1 = (const int32_t*)s->bed1.p; bed.n = s->n_bed; bed.chrom_len = s->chrom_len;
  bed.batch = P.batch_size;

  // ---- the step: one stream, no host synchronisation ---------------------------------------------------
  VSV_TRY(vsv_t0(ctx));
  VSV_CUDA(ctx, cudaMemsetAsync(st, 0, sizeof(ShardState), ctx->stream));
  if (n_rec > 0 && n_wt > 0) {
    VSV_LAUNCH(ctx, k_shard_prep, (unsigned)vsv_cdiv(n_rec + 1, 256), 256, 0, s->R.ops, s->R.op_off, s->R.mapq, s->R.qlen, n_rec, n_wt,
               P.reads_min_mapq, P.cute_min_mapq, P.cute_min_read_len, (uint32_t*)s->info.p, (int32_t*)s->first.p);
    VSV_LAUNCH(ctx, k_wt_starts, (unsigned)vsv_cdiv(n_wt, 256), 256, 0, (const int32_t*)s->first.p, n_wt, (int2*)s->starts.p);
    VSV_CUDA(ctx, cudaMemsetAsync((TileState*)s->agg.p + n_wt, 0, sizeof(TileState), ctx->stream));
    VSV_CUDA(ctx, cudaEventRecord(ctx->evk0, ctx->stream));
    VSV_LAUNCH(ctx, k_scan_wt<VSV_MODE_FUSED>, (unsigned)grid, SCAN_THREADS,
               (size_t)SCAN_WARPS * (SCAN_STAGES * (WT_OPS * sizeof(uint32_t) + sizeof(uint64_t)) + 8 * sizeof(uint32_t)), p);
    VSV_CUDA(ctx, cudaEventRecord(ctx->evk1, ctx->stream));
    ctx->main_open = true;
    const unsigned gs = (unsigned)std::min<int64_t>(vsv_cdiv(scap, 256), (int64_t)ctx->sm_count * 8);
    VSV_LAUNCH(ctx, k_gate<VSV_MODE_FUSED>, gs, 256, 0, p);
    const int64_t n_el = (int64_t)n_wt + 1;
    // segmented exclusive scan of the tile aggregates in one launch (decoupled look-back)
    VSV_CUDA(ctx, cudaMemsetAsync(s->blk.p, 0, (size_t)(nb + 1) * sizeof(TsDesc), ctx->stream));
    VSV_LAUNCH(ctx, k_ts_scan1, (unsigned)nb, TS_THREADS, 0, (const TileState*)p.wt_agg, n_el, (TsDesc*)s->blk.p,
               reinterpret_cast<unsigned int*>((TsDesc*)s->blk.p + nb), d_excl);
    VSV_LAUNCH(ctx, k_emit<VSV_MODE_FUSED>, gs, 256, 0, p, (const TileState*)d_excl, (vsv_sig*)s->sig_reads.p, cap.r, (vsv_sig*)s->sig_cute.p,
               cap.c, st->nout);
    VSV_LAUNCH(ctx, k_shard_rec_totals, (unsigned)vsv_cdiv(n_rec, 256), 256, 0, p, (const TileState*)d_excl, bed, (int32_t*)s->ref_end.p,
               (int32_t*)s->readlen.p, (uint8_t*)s->in_bed.p);
  } else if (n_rec > 0) {   // records without a single op
    VSV_CUDA(ctx, cudaMemcpyAsync(s->ref_end.p, s->R.pos, (size_t)n_rec * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    VSV_CUDA(ctx, cudaMemsetAsync(s->readlen.p, 0, (size_t)n_rec * 4, ctx->stream));
    VSV_CUDA(ctx, cudaMemsetAsync(s->in_bed.p, 1, (size_t)n_rec, ctx->stream));
  }
  if (n_rec > 0) {
    if (s->R.next) VSV_LAUNCH(ctx, k_shard_split, (unsigned)vsv_cdiv(n_rec, 256), 256, 0, s->R, (const int32_t*)s->ref_end.p,
                              (const int32_t*)s->readlen.p, st, (vsv_sig*)s->sig_reads.p, cap.r);
    if (s->A.n_read > 0)
      VSV_LAUNCH(ctx, k_shard_sa, (unsigned)vsv_cdiv(s->A.n_read, 128), 128, 0, s->R, s->A, (const int32_t*)s->ref_end.p,
                 (const uint8_t*)s->in_bed.p, P.cute_min_mapq, P.cute_min_read_len, P.sa_sv_size, P.sa_max_size, P.sa_max_parts, st,
                 (vsv_sig*)s->rows_del.p, cap.d, (vsv_sig*)s->rows_ins.p, cap.i);
    const unsigned gm = (unsigned)std::max<int64_t>(1, std::min<int64_t>(vsv_cdiv(cap.c, 256), (int64_t)ctx->sm_count * 8));
    VSV_LAUNCH(ctx, k_shard_merge, gm, MERGE_THREADS, 0, (const vsv_sig*)s->sig_cute.p, st, cap, (const uint8_t*)s->in_bed.p, (vsv_sig*)s->rows_del.p,
               (vsv_sig*)s->rows_ins.p);
  }
  const unsigned gk = (unsigned)std::max<int64_t>(1, std::min<int64_t>(vsv_cdiv(cap_total, 256), (int64_t)ctx->sm_count * 16));
  VSV_LAUNCH(ctx, k_shard_keys, gk, 256, 0, st, cap, (const vsv_sig*)s->sig_reads.p, (const vsv_sig*)s->rows_del.p, (const vsv_sig*)s->rows_ins.p,
             (const int32_t*)s->ref_end.p, (uint32_t*)s->keys[0].p, (uint32_t*)s->vals[0].p);
  VSV_CUDA(ctx, cub::DeviceRadixSort::SortPairs(s->cub_tmp.p, cub_bytes, (const uint32_t*)s->keys[0].p, (uint32_t*)s->keys[1].p,
                                                (const uint32_t*)s->vals[0].p, (uint32_t*)s->vals[1].p, (int)cap_total, 0, 32, ctx->stream));
  ctx->launches += 6;
  const Sorted S = sorted_of(s);
  VSV_LAUNCH(ctx, k_shard_unpack, gk, 256, 0, (const ShardState*)st, cap, (const uint32_t*)s->keys[1].p, (const uint32_t*)s->vals[1].p,
             (const vsv_sig*)s->sig_reads.p, (const vsv_sig*)s->rows_del.p, (const vsv_sig*)s->rows_ins.p, S);
  const int64_t cap_max = std::max(cap.d, cap.i);
  const unsigned gtie = (unsigned)std::max<int64_t>(1, std::min<int64_t>(vsv_cdiv(cap_max, 256), (int64_t)ctx->sm_count * 8));
  VSV_LAUNCH(ctx, k_shard_tie_small, dim3(gtie, 2), 256, 0, st, cap, S, (TieWork*)s->work.p, work_cap);
  VSV_LAUNCH(ctx, k_shard_tie_long, (unsigned)ctx->sm_count, 256, 0, (const ShardState*)st, S, (const TieWork*)s->work.p, work_cap,
             (unsigned long long*)s->key_tmp.p, (int32_t*)s->idx_tmp.p, (int32_t*)s->len_tmp.p, cap.d);
  VSV_TRY(enqueue_group(s));
  VSV_TRY(vsv_t1(ctx));
  s->scanned = true;
  return vsv_end(ctx);
}

extern "C" int vsv_shard_counts_get(vsv_shard* s, vsv_shard_counts* c) {
  if (!s || !c) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  if (!s->scanned) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_counts_get: no scan");
  VSV_CUDA(ctx, cudaSetDevice(ctx->device));
  if (s->state_pending) {
    VSV_CUDA(ctx, cudaEventSynchronize(s->ev_state));
    s->state_pending = false;
  }
  const ShardState& h = *s->h_state;
  const Caps& cap = s->cap;
  memset(c, 0, sizeof(*c));
  c->n_reads_cigar = h.nout[0];
  c->n_reads_split = (int64_t)h.n_split;
  c->n_cute = h.nout[1];
  c->n_del = (int64_t)h.n_del;
  c->n_ins = (int64_t)h.n_ins;
  c->n_sa_rows = (int64_t)h.n_sa;
  c->n_del_dup = (int64_t)h.n_del_dup;
  c->n_ins_maybe_dup = (int64_t)h.n_ins_dup;
  c->n_udel = (int64_t)h.n_u[0];
  c->n_uins = (int64_t)h.n_u[1];
  c->n_stage = (int64_t)h.stage_ctr[0];
  int64_t ov = h.overflow;
  if (h.nout[0] + (int64_t)h.n_split > cap.r) ov |= VSV_OVF_READS;
  if (h.nout[1] > cap.c) ov |= VSV_OVF_CUTE;
  if ((int64_t)h.n_del > cap.d) ov |= VSV_OVF_DEL;
  if ((int64_t)h.n_ins > cap.i) ov |= VSV_OVF_INS;
  if (h.stage_ctr[2] != 0) {
    ov |= VSV_OVF_STAGE;
    s->stage_cap = 2 * (int64_t)h.stage_ctr[0] + (1 << 20);   // reservation of the failed run bounds the raw count
  }
  c->overflow = ov;
  return VSV_OK;
}

extern "C" int vsv_shard_drop_ins(vsv_shard* s, const int32_t* uid, int64_t n) {
  if (!s) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  VSV_TRY(vsv_begin(ctx));
  if (!s->scanned || n < 0 || (n > 0 && !uid)) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_drop_ins: bad argument");
  VSV_TRY(dbuf_ensure(ctx, s->drop, (size_t)std::max<int64_t>(n, 4) * 4));
  if (n) {
    VSV_CUDA(ctx, cudaMemcpyAsync(s->drop.p, uid, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  s->n_drop = n;
  // reset what the grouping pass accumulates
  ShardState* st = (ShardState*)s->state.p;
  VSV_CUDA(ctx, cudaMemsetAsync(&st->n_del_dup, 0, 4 * sizeof(unsigned long long), ctx->stream));   // n_del_dup, n_ins_dup, n_u[2]
  VSV_CUDA(ctx, cudaMemsetAsync(st->ticket, 0, 2 * sizeof(unsigned int), ctx->stream));
  VSV_TRY(enqueue_group(s));
  return vsv_end(ctx);
}

extern "C" int vsv_shard_scan_kernel_ms(vsv_shard* s, float* ms) {
  if (!s) return VSV_E_ARG;
  return vsv_get_main_kernel_ms(s->ctx, ms);
}

extern "C" int vsv_shard_fetch(vsv_shard* s, int table, void* out, int64_t cap_rows, int64_t* n) {
  if (!s || !n) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  if (!s->scanned) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_fetch: no scan");
  vsv_shard_counts c;
  VSV_TRY(vsv_shard_counts_get(s, &c));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const Caps& cap = s->cap;
  const int64_t n_cig = std::min(c.n_reads_cigar, cap.r);
  const int64_t n_r = std::min(n_cig + c.n_reads_split, cap.r);
  const void* src = nullptr;
  int64_t rows = 0;
  size_t row_bytes = 0;
  switch (table) {
    case VSV_TAB_READS: src = s->sig_reads.p; rows = n_r; row_bytes = sizeof(vsv_sig); break;
    case VSV_TAB_CUTE: src = s->sig_cute.p; rows = std::min(c.n_cute, cap.c); row_bytes = sizeof(vsv_sig); break;
    case VSV_TAB_DEL_ROWS: src = s->rows_del.p; rows = std::min(c.n_del, cap.d); row_bytes = sizeof(vsv_sig); break;
    case VSV_TAB_INS_ROWS: src = s->rows_ins.p; rows = std::min(c.n_ins, cap.i); row_bytes = sizeof(vsv_sig); break;
    case VSV_TAB_REF_END: src = s->ref_end.p; rows = s->R.n_rec; row_bytes = 4; break;
    case VSV_TAB_READ_LEN: src = s->readlen.p; rows = s->R.n_rec; row_bytes = 4; break;
    case VSV_TAB_IN_BED: src = s->in_bed.p; rows = s->R.n_rec; row_bytes = 1; break;
    case VSV_TAB_INS_DUPS: src = s->dup_list.p; rows = std::min<int64_t>(c.n_ins_maybe_dup, DUPLIST_CAP); row_bytes = sizeof(int2); break;
    case VSV_TAB_UDEL:
    case VSV_TAB_UINS: {
      const int t = table == VSV_TAB_UDEL ? 0 : 1;
      rows = std::min(t == 0 ? c.n_udel : c.n_uins, t == 0 ? cap.d : cap.i);
      *n = rows;
      if (rows > cap_rows) VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_shard_fetch: capacity too small");
      if (rows > 0 && !out) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_fetch: out is NULL");
      const void* cols[3] = {s->u_pos[t].p, s->u_len[t].p, s->u_cnt[t].p};
      for (int k = 0; k < 3 && rows > 0; ++k)
        VSV_CUDA(ctx, cudaMemcpy((int32_t*)out + (size_t)k * cap_rows, cols[k], (size_t)rows * 4, cudaMemcpyDeviceToHost));
      return VSV_OK;
    }
    default: VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_fetch: unknown table");
  }
  *n = rows;
  if (rows > cap_rows) VSV_FAIL(ctx, VSV_E_CAPACITY, "vsv_shard_fetch: capacity too small");
  if (rows > 0) {
    if (!out) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_fetch: out is NULL");
    VSV_CUDA(ctx, cudaMemcpy(out, src, (size_t)rows * row_bytes, cudaMemcpyDeviceToHost));
  }
  return VSV_OK;
}

extern "C" int vsv_shard_candidates(vsv_shard* s, int what, const int32_t* cand_pos, const int32_t* cand_svlen,
                                    const int32_t* cand_type, const int32_t* call_gt, int64_t n_cand, const int32_t* last_in,
                                    int32_t* fp_support, int64_t* cov, int32_t* n_support, int32_t* match_id, int32_t* depth2,
                                    double* ratio, int32_t* new_gt, int32_t* last_out) {
  if (!s) return VSV_E_ARG;
  vsv_ctx* ctx = s->ctx;
  VSV_TRY(vsv_begin(ctx));
  if (!s->scanned) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_candidates: no scan");
  if (n_cand < 0 || (what & ~15) || what == 0 || (n_cand > 0 && (!cand_pos || !cand_svlen || !cand_type)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_candidates: bad argument");
  if (((what & VSV_JOIN_FP) && !fp_support) || ((what & VSV_JOIN_COV) && !cov) || ((what & (VSV_JOIN_SUPPORT | VSV_JOIN_GENOTYPE)) && !n_support) ||
      ((what & VSV_JOIN_GENOTYPE) && (!depth2 || !ratio || !new_gt)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_candidates: an output of a selected stage is NULL");
  if (ctx->device_io && last_out) VSV_FAIL(ctx, VSV_E_ARG, "vsv_shard_candidates: last_out must be NULL on a device-IO ctx");
  if (last_out) { last_out[0] = last_in ? last_in[0] : 0; last_out[1] = last_in ? last_in[1] : 0; }
  if (n_cand == 0) return VSV_OK;
  const vsv_params& P = ctx->params;
  const size_t n = (size_t)n_cand;
  const int32_t *cp, *cl, *ct, *cg;
  VSV_TRY(vsv_in(ctx, cand_pos, n, &cp));
  VSV_TRY(vsv_in(ctx, cand_svlen, n, &cl));
  VSV_TRY(vsv_in(ctx, cand_type, n, &ct));
  VSV_TRY(vsv_in(ctx, call_gt, n, &cg));
  int32_t *d_fp = nullptr, *d_sup = nullptr, *d_mid = nullptr, *d_d2 = nullptr, *d_gt = nullptr, *d_lb = nullptr, *d_ub = nullptr;
  int64_t* d_cov = nullptr;
  double* d_ratio = nullptr;
  VSV_TRY(vsv_out(ctx, fp_support, n, &d_fp));
  VSV_TRY(vsv_out(ctx, cov, n, &d_cov));
  if (what & VSV_JOIN_SUPPORT) {
    VSV_TRY(vsv_out(ctx, n_support, n, &d_sup));
  } else if (what & VSV_JOIN_GENOTYPE) {   // n_support is an INPUT
    const int32_t* in_sup = nullptr;
    VSV_TRY(vsv_in(ctx, (const int32_t*)n_support, n, &in_sup));
    d_sup = const_cast<int32_t*>(in_sup);
  }
  VSV_TRY(vsv_out(ctx, match_id, n, &d_mid));
  VSV_TRY(vsv_out(ctx, depth2, n, &d_d2));
  VSV_TRY(vsv_out(ctx, ratio, n, &d_ratio));
  VSV_TRY(vsv_out(ctx, new_gt, n, &d_gt));
  ShardState* st = (ShardState*)s->state.p;
  const Sorted S = sorted_of(s);
  const Unique U = unique_of(s);
  VSV_TRY(vsv_t0(ctx));
  if (what & (VSV_JOIN_FP | VSV_JOIN_COV | VSV_JOIN_SUPPORT)) {
    if (what & VSV_JOIN_SUPPORT) {
      VSV_TRY(vsv_tmp(ctx, n, &d_lb));
      VSV_TRY(vsv_tmp(ctx, n, &d_ub));
    }
    JoinParams jp;
    jp.what = what; jp.fp_max_comp = P.fp_max_comp; jp.fp_max_shift = P.fp_max_shift; jp.cov_flank = P.cov_flanking;
    jp.min_size_sim = P.gt_min_size_sim; jp.max_shift_ratio = P.gt_max_shift_ratio; jp.min_shift = P.gt_min_shift;
    CandOut O;
    O.fp = d_fp; O.cov = d_cov; O.sup = d_sup; O.lb = d_lb; O.ub = d_ub;
    VSV_LAUNCH(ctx, k_cand_joins, (unsigned)vsv_cdiv(n_cand * 32, 256), 256, 0, (const ShardState*)st, s->cap, S, U,
               (const uint8_t*)s->keep[0].p, (const uint8_t*)s->keep[1].p, (const int32_t*)s->pmax.p, cp, cl, ct, n_cand, jp, O);
  }
  int2* d_last = nullptr;
  if (what & VSV_JOIN_SUPPORT) {
    const int64_t tiles = vsv_cdiv(n_cand, CHAIN_THREADS);
    LbDesc* desc = nullptr;
    unsigned int* ticket = nullptr;
    VSV_TRY(vsv_tmp(ctx, (size_t)tiles + 1, &desc));   // the slot after the descriptors holds the ticket
    ticket = reinterpret_cast<unsigned int*>(desc + tiles);
    VSV_CUDA(ctx, cudaMemsetAsync(desc, 0, ((size_t)tiles + 1) * sizeof(LbDesc), ctx->stream));
    if (last_out) VSV_TRY(vsv_tmp(ctx, 1, &d_last));
    const int2 li = make_int2(last_in ? last_in[0] : 0, last_in ? last_in[1] : 0);
    VSV_LAUNCH(ctx, k_cand_chain, (unsigned)tiles, CHAIN_THREADS, 0, (const ShardState*)st, s->cap, U, cl, ct, n_cand, li, P.gt_min_size_sim, desc,
               ticket, d_sup, (const int32_t*)d_lb, (const int32_t*)d_ub, d_mid, d_last);
  }
  if (what & VSV_JOIN_GENOTYPE) {
    const GtParams gp = gt_params_of(P);
    VSV_LAUNCH(ctx, k_cand_genotype, (unsigned)vsv_cdiv(n_cand * 8, 256), 256, 0, s->R.pos, (const int32_t*)s->ref_end.p, (const int32_t*)s->send.p,
               s->R.n_rec, cp, cl, ct, cg, (const int32_t*)d_sup, n_cand, gp, d_d2, d_ratio, d_gt);
  }
  VSV_TRY(vsv_t1(ctx));
  if (what & VSV_JOIN_FP) VSV_TRY(vsv_out_done(ctx, fp_support, d_fp, n));
  if (what & VSV_JOIN_COV) VSV_TRY(vsv_out_done(ctx, cov, d_cov, n));
  if (what & VSV_JOIN_SUPPORT) {
    VSV_TRY(vsv_out_done(ctx, n_support, d_sup, n));
    VSV_TRY(vsv_out_done(ctx, match_id, d_mid, n));
  }
  if (what & VSV_JOIN_GENOTYPE) {
    VSV_TRY(vsv_out_done(ctx, depth2, d_d2, n));
    VSV_TRY(vsv_out_done(ctx, ratio, d_ratio, n));
    VSV_TRY(vsv_out_done(ctx, new_gt, d_gt, n));
  }
  if (last_out && d_last) {
    int2 h = make_int2(0, 0);
    VSV_CUDA(ctx, cudaMemcpyAsync(&h, d_last, sizeof(int2), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    last_out[0] = h.x; last_out[1] = h.y;
  }
  return vsv_end(ctx);
}

// ---- ECS_C:53-70: per-record sums for the CLR contig gate ----------------------------------------------
static __global__ void __launch_bounds__(256)
k_class_sums(const uint32_t* __restrict__ ops, const int64_t* __restrict__ op_off, int64_t n_rec, int64_t* __restrict__ ins_sum,
             int64_t* __restrict__ match_sum, int64_t* __restrict__ match_cnt) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;   // one warp per record
  const int lane = threadIdx.x & 31;
  if (r >= n_rec) return;
  const int64_t a = op_off[r], b = op_off[r + 1];
  long long si = 0, sm = 0, cm = 0;
  for (int64_t k = a + lane; k < b; k += 32) {
    const uint32_t w = __ldg(ops + k);
    const uint32_t c = w & 15u;
    if (c == 1u) si += w >> 4;
    if (c == 0u) { sm += w >> 4; ++cm; }
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    si += __shfl_xor_sync(0xffffffffu, si, d); sm += __shfl_xor_sync(0xffffffffu, sm, d); cm += __shfl_xor_sync(0xffffffffu, cm, d);
  }
  if (lane == 0) { ins_sum[r] = si; match_sum[r] = sm; match_cnt[r] = cm; }
}
extern "C" int vsv_cigar_class_sums(vsv_ctx* ctx, const uint32_t* ops, const int64_t* op_off, int64_t n_rec, int64_t* ins_sum,
                                    int64_t* match_sum, int64_t* match_cnt) {
  VSV_TRY(vsv_begin(ctx));
  if (n_rec < 0 || (n_rec > 0 && (!op_off || !ins_sum || !match_sum || !match_cnt))) VSV_FAIL(ctx, VSV_E_ARG, "vsv_cigar_class_sums: bad argument");
  if (n_rec == 0) return VSV_OK;
  int64_t n_ops = 0;
  if (ctx->device_io) {
    VSV_CUDA(ctx, cudaMemcpyAsync(&n_ops, op_off + n_rec, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    n_ops = op_off[n_rec];
  }
  const uint32_t* d_ops;
  const int64_t* d_off;
  int64_t *d_i, *d_m, *d_c;
  VSV_TRY(vsv_in(ctx, ops, (size_t)n_ops, &d_ops));
  VSV_TRY(vsv_in(ctx, op_off, (size_t)n_rec + 1, &d_off));
  VSV_TRY(vsv_out(ctx, ins_sum, (size_t)n_rec, &d_i));
  VSV_TRY(vsv_out(ctx, match_sum, (size_t)n_rec, &d_m));
  VSV_TRY(vsv_out(ctx, match_cnt, (size_t)n_rec, &d_c));
  VSV_TRY(vsv_t0(ctx));
  VSV_LAUNCH(ctx, k_class_sums, (unsigned)vsv_cdiv(n_rec * 32, 256), 256, 0, d_ops, d_off, n_rec, d_i, d_m, d_c);
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, ins_sum, d_i, (size_t)n_rec));
  VSV_TRY(vsv_out_done(ctx, match_sum, d_m, (size_t)n_rec));
  VSV_TRY(vsv_out_done(ctx, match_cnt, d_c, (size_t)n_rec));
  return vsv_end(ctx);
}
