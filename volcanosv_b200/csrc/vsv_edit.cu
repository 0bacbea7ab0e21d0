// K7: batched global edit distance (unit-cost Levenshtein / Needleman-Wunsch), the quantity
// remove_redundancy.py takes from edlib (`edlib.align(a, b)["editDistance"]`, RR:75-81) to
// decide whether two nearby insertion calls carry the same sequence.  The value is unique, so
// any exact algorithm is bit-exact; this one is Myers' bit-vector algorithm in Hyyro's block
// formulation (64 pattern rows per machine word, horizontal delta carried between blocks).
//
// Mapping: one warp per pair.  The shorter string is the pattern, cut into W = ceil(m/64) blocks;
// lane l owns K = ceil(W/32) consecutive blocks and the warp runs as a systolic pipeline over
// the text: at step s lane l advances its blocks through column s - l and hands the horizontal
// delta of its last block to lane l + 1 with one shuffle.  The pattern is padded to 64 W rows
// with rows that match everything and the text with as many match-all columns, so the global
// distance is read at the bottom-right corner of the padded matrix.  Per-pair tables (match
// masks of the <= 16 distinct pattern symbols, vertical deltas) live in a global scratch that
// stays in L1/L2.
#include "vsv_common.cuh"

#define ED_CODES 18        // 0..15 pattern symbols, 16 = symbol absent from the pattern, 17 = padding column
#define ED_MAXSYM 16

struct EdIn {
  const uint8_t* chars;
  const int64_t* str_off;
  const int32_t *ia, *ib;
};

__global__ void k_ed_blocks(EdIn in, int64_t n_pair, int32_t* __restrict__ nblk) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pair) return;
  const int64_t la = in.str_off[in.ia[p] + 1] - in.str_off[in.ia[p]];
  const int64_t lb = in.str_off[in.ib[p] + 1] - in.str_off[in.ib[p]];
  const int64_t m = min(la, lb);
  nblk[p] = (int32_t)((m + 63) / 64);
}

// one block update (Myers 1999, Hyyro 2003): vertical deltas Pv/Mv of 64 rows, match mask eq,
// horizontal delta hin in {-1,0,+1} entering at the top row; returns the delta leaving the bottom row
__device__ __forceinline__ int ed_block(unsigned long long& Pv, unsigned long long& Mv, unsigned long long eq, int hin) {
  const unsigned long long hneg = hin < 0 ? 1ull : 0ull;
  const unsigned long long Xv = eq | Mv;
  eq |= hneg;
  const unsigned long long Xh = (((eq & Pv) + Pv) ^ Pv) | eq;
  unsigned long long Ph = Mv | ~(Xh | Pv);
  unsigned long long Mh = Pv & Xh;
  const int hout = (int)(Ph >> 63) - (int)(Mh >> 63);
  Ph <<= 1;
  Mh <<= 1;
  Mh |= hneg;
  Ph |= hin > 0 ? 1ull : 0ull;
  Pv = Mh | ~(Xv | Ph);
  Mv = Ph & Xv;
  return hout;
}

__global__ void __launch_bounds__(32) k_edit_distance(EdIn in, int64_t n_pair, const int64_t* __restrict__ blk_off,
                                                      unsigned long long* __restrict__ scratch, int32_t* __restrict__ dist,
                                                      int32_t* __restrict__ bad) {
  __shared__ uint32_t s_present[8];
  __shared__ uint8_t s_code[256];
  const int64_t p = blockIdx.x;
  const int lane = threadIdx.x;
  const uint8_t* A = in.chars + in.str_off[in.ia[p]];
  const uint8_t* B = in.chars + in.str_off[in.ib[p]];
  int64_t la = in.str_off[in.ia[p] + 1] - in.str_off[in.ia[p]];
  int64_t lb = in.str_off[in.ib[p] + 1] - in.str_off[in.ib[p]];
  if (la > lb) {  // pattern = the shorter string
    const uint8_t* t = A; A = B; B = t;
    const int64_t tl = la; la = lb; lb = tl;
  }
  const int64_t m = la, n = lb;
  if (m == 0) {
    if (lane == 0) dist[p] = (int32_t)n;
    return;
  }
  const int64_t W = (m + 63) / 64;
  const int64_t pad = 64 * W - m;
  unsigned long long* peq = scratch + blk_off[p] * (ED_CODES + 2);  // [ED_CODES][W]
  unsigned long long* pv = peq + ED_CODES * W;
  unsigned long long* mv = pv + W;
  // -- symbol codes of the pattern's alphabet ---------------------------------------------------
  if (lane < 8) s_present[lane] = 0u;
  __syncwarp();
  for (int64_t i = lane; i < m; i += 32) atomicOr(&s_present[A[i] >> 5], 1u << (A[i] & 31));
  __syncwarp();
  for (int v = lane; v < 256; v += 32) {
    int rank = 0;
    for (int k = 0; k < (v >> 5); ++k) rank += __popc(s_present[k]);
    rank += __popc(s_present[v >> 5] & ((1u << (v & 31)) - 1u));
    const bool on = (s_present[v >> 5] >> (v & 31)) & 1u;
    if (on && rank >= ED_MAXSYM) *bad = 1;
    s_code[v] = on ? (uint8_t)min(rank, ED_MAXSYM - 1) : (uint8_t)ED_MAXSYM;
  }
  // -- match masks: pattern rows, then the padding rows (match every symbol) and the padding column code
  for (int64_t i = lane; i < (ED_CODES + 2) * W; i += 32) peq[i] = 0ull;
  __syncwarp();
  for (int64_t i = lane; i < m; i += 32) atomicOr(&peq[(int64_t)s_code[A[i]] * W + (i >> 6)], 1ull << (i & 63));
  __syncwarp();
  if (pad > 0 && lane < ED_CODES - 1) peq[(int64_t)lane * W + (W - 1)] |= ~0ull << (64 - pad);
  for (int64_t w = lane; w < W; w += 32) {
    peq[(int64_t)(ED_CODES - 1) * W + w] = ~0ull;
    pv[w] = ~0ull;
    mv[w] = 0ull;
  }
  __syncwarp();
  // -- systolic sweep over the n + pad columns -----------------------------------------------------
  const int64_t K = (W + 31) / 32;
  const int64_t w0 = (int64_t)lane * K, w1 = min(W, w0 + K);
  const int last_lane = (int)((W - 1) / K);
  const int64_t ncol = n + pad;
  int64_t score = 64 * W;  // D[64 W][0]
  int carry = 0;           // horizontal delta handed to the next lane
  for (int64_t s = 0; s < ncol + last_lane; ++s) {
    const int from_prev = __shfl_up_sync(0xffffffffu, carry, 1);
    const int64_t c = s - lane;
    int hout = 0;
    if (c >= 0 && c < ncol && w0 < W) {
      const int code = c < n ? (int)s_code[B[c]] : ED_CODES - 1;
      int h = lane == 0 ? 1 : from_prev;  // top row of the matrix: D[0][j] = j
      const unsigned long long* row = peq + (int64_t)code * W;
      for (int64_t w = w0; w < w1; ++w) {
        unsigned long long P = pv[w], M = mv[w];
        h = ed_block(P, M, row[w], h);
        pv[w] = P;
        mv[w] = M;
      }
      hout = h;
      if (lane == last_lane) score += h;
    }
    carry = hout;
  }
  if (lane == last_lane) dist[p] = (int32_t)score;
}

extern "C" int vsv_edit_distance_batch(vsv_ctx* ctx, const uint8_t* chars, const int64_t* str_off, int64_t n_str,
                                       const int32_t* ia, const int32_t* ib, int64_t n_pair, int32_t* dist) {
  VSV_TRY(vsv_begin(ctx));
  if (n_str < 0 || n_pair < 0 || (n_pair > 0 && (!str_off || !ia || !ib || !dist)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: bad argument");
  if (n_pair == 0) return VSV_OK;
  if (ctx->device_io) VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: host buffers only");
  for (int64_t p = 0; p < n_pair; ++p)
    if (ia[p] < 0 || ia[p] >= n_str || ib[p] < 0 || ib[p] >= n_str)
      VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: string index out of range");
  const int64_t n_chars = str_off[n_str];
  EdIn in;
  VSV_TRY(vsv_in(ctx, chars, (size_t)n_chars, &in.chars));
  VSV_TRY(vsv_in(ctx, str_off, (size_t)n_str + 1, &in.str_off));
  VSV_TRY(vsv_in(ctx, ia, (size_t)n_pair, &in.ia));
  VSV_TRY(vsv_in(ctx, ib, (size_t)n_pair, &in.ib));
  int32_t *d_dist, *nblk, *bad;
  int64_t *blk_off, *tot;
  VSV_TRY(vsv_out(ctx, dist, (size_t)n_pair, &d_dist));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_pair, &nblk));
  VSV_TRY(vsv_tmp(ctx, (size_t)n_pair, &blk_off));
  VSV_TRY(vsv_tmp(ctx, 1, &tot));
  VSV_TRY(vsv_tmp(ctx, 1, &bad));
  VSV_TRY(vsv_t0(ctx));
  VSV_CUDA(ctx, cudaMemsetAsync(bad, 0, sizeof(int32_t), ctx->stream));
  VSV_LAUNCH(ctx, k_ed_blocks, (unsigned)vsv_cdiv(n_pair, 256), 256, 0, in, n_pair, nblk);
  VSV_TRY(vsv_exclusive_scan_i32(ctx, nblk, n_pair, blk_off, tot));
  int64_t h_tot = 0;
  VSV_CUDA(ctx, cudaMemcpyAsync(&h_tot, tot, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  unsigned long long* scratch;
  VSV_TRY(vsv_tmp(ctx, (size_t)(h_tot + 1) * (ED_CODES + 2), &scratch));
  VSV_LAUNCH(ctx, k_edit_distance, (unsigned)n_pair, 32, 0, in, n_pair, blk_off, scratch, d_dist, bad);
  VSV_TRY(vsv_t1(ctx));
  int32_t h_bad = 0;
  VSV_CUDA(ctx, cudaMemcpyAsync(&h_bad, bad, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  VSV_TRY(vsv_out_done(ctx, dist, d_dist, (size_t)n_pair));
  VSV_TRY(vsv_end(ctx, true));
  if (h_bad) VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: a string uses more than 16 distinct symbols");
  return VSV_OK;
}
