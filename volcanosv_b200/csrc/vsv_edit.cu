// K7: batched global edit distance (unit-cost Levenshtein / Needleman-Wunsch), the quantity
// remove_redundancy.py takes from edlib (`edlib.align(a, b)["editDistance"]`, RR:75-81) to
// decide whether two nearby insertion calls carry the same sequence.  The value is unique, so
// any exact algorithm is bit-exact; this one is Myers' bit-vector algorithm in Hyyro's block
// formulation (64 pattern rows per machine word) run inside an Ukkonen band that widens until
// the result is proven exact -- the scheme edlib itself uses for its NW mode.
//
// Layout of one pair: the shorter string is the pattern, cut into W = ceil(m/64) row blocks; the
// text is cut into chunks of 32 columns.  Block w is owned by ring lane w mod L (L = 32 lanes of
// a warp for short pairs, the 512 threads of a CTA for long ones).  The sweep is systolic at
// TILE granularity: at macro step T the owner of block w advances it through the 32 columns of
// chunk T - w with Pv/Mv in registers, consuming the 32 horizontal deltas block w-1 produced one
// macro step earlier (two 32-bit masks) and leaving its own for block w+1.  The blocks a lane
// owns at one macro step sit on different chunks and are independent, which gives the lane ILP.
// One warp-/CTA-level barrier per macro step, i.e. per 32 columns.
//
// Band: with threshold k and delta = n - m, a cell (i, j) can lie on a path of cost <= k only if
// -(k-delta)/2 <= j - i <= (k+delta)/2.  A block is computed in the chunks that intersect that
// stripe; when it enters the stripe its vertical deltas start at +1 (an upper bound) and its
// bottom-row score at (score of the block above) + rows, when the block above has left the
// stripe the incoming horizontal deltas are +1.  Every computed value is >= the true one and
// values on in-band optimal paths are exact, so a result <= k is the distance; otherwise k grows
// 8x (k >= n always terminates).  First k = max(delta + 64, (m+n)/64): near-identical pairs --
// what the redundancy stage is looking for -- finish in the first, narrow round.
//
// State (Pv, Mv, per-symbol match masks Peq, the carry ring) lives in shared memory: 40 + 8*nsym
// bytes per block.  Pairs whose state exceeds the CTA's shared memory use a global scratch
// through the same pointers.  Short pairs (m <= 8192) run one per warp, four per CTA.
#include <algorithm>

#include "vsv_common.cuh"

#define ED_SHORT_BLOCKS 128          // W <= 128: one warp per pair
#define ED_LONG_WARPS 16             // warps cooperating on a long pair
#define ED_LONG_SMEM (200 * 1024)    // shared-memory budget of a long pair's CTA
#define ED_MAX_LEN (1 << 28)
#define ED_MAX_SYM 32                // 4 short pairs x 128 blocks x (5 + 32) words = 151 KB of shared memory at most

struct EdIn {
  const uint8_t* codes;      // strings translated to symbol codes 0..nsym-1 (host side, one table for the call)
  const int64_t* str_off;
  const int32_t *ia, *ib;
};

template <int NW>
__device__ __forceinline__ void ed_sync() {
  if (NW == 1) __syncwarp(); else __syncthreads();
}

// words of state per block: Pv, Mv, Peq[nsym], carry ring 2 x {hp, hm, score} (3 words)
__host__ __device__ __forceinline__ int64_t ed_words(int64_t W, int nsym) { return W * (5 + nsym); }

template <int NW>
__global__ void __launch_bounds__(NW == 1 ? 128 : NW * 32)
k_edit_banded(EdIn in, const int32_t* __restrict__ list, int n_list, int nsym, int words_per_group,
              unsigned long long* __restrict__ gscratch, const int64_t* __restrict__ goff, int32_t* __restrict__ dist) {
  extern __shared__ unsigned long long ed_smem[];
  constexpr int L = NW * 32;
  const int l = NW == 1 ? (threadIdx.x & 31) : threadIdx.x;
  const int g = NW == 1 ? blockIdx.x * 4 + (threadIdx.x >> 5) : blockIdx.x;
  if (g >= n_list) return;                      // NW == 1: whole warps leave, no CTA barrier is used in that case
  const int p = list[g];
  const uint8_t* A = in.codes + in.str_off[in.ia[p]];
  const uint8_t* B = in.codes + in.str_off[in.ib[p]];
  int m = (int)(in.str_off[in.ia[p] + 1] - in.str_off[in.ia[p]]);
  int n = (int)(in.str_off[in.ib[p] + 1] - in.str_off[in.ib[p]]);
  if (m > n) {                                  // pattern = the shorter string
    const uint8_t* t = A; A = B; B = t;
    const int tl = m; m = n; n = tl;
  }
  if (m == 0) {
    if (l == 0) dist[p] = n;
    return;
  }
  const int W = (m + 63) >> 6, Q = (n + 31) >> 5;
  unsigned long long* st = (goff != nullptr && goff[g] >= 0) ? gscratch + goff[g]
                                                             : ed_smem + (NW == 1 ? (size_t)(threadIdx.x >> 5) * words_per_group : 0);
  unsigned long long* pv = st;
  unsigned long long* mv = st + W;
  unsigned long long* peq = st + 2 * (size_t)W;                       // [nsym][W]
  int32_t* carry = (int32_t*)(peq + (size_t)nsym * W);                // [2][W][3] = {hp, hm, score}
  for (int64_t i = l; i < (int64_t)nsym * W; i += L) peq[i] = 0ull;
  ed_sync<NW>();
  for (int i = l; i < m; i += L) atomicOr(&peq[(size_t)A[i] * W + (i >> 6)], 1ull << (i & 63));
  ed_sync<NW>();

  const int delta = n - m;
  int64_t k = max((int64_t)delta + 64, (int64_t)(m + n) >> 6);
  const int Tn = Q + W - 1;
  int result = 0;
  for (;;) {
    const int64_t lo = (k - delta) >> 1, hi = (k + delta) >> 1;
    for (int T = 0; T < Tn; ++T) {
      const int par = T & 1;
      for (int w = l; w < W; w += L) {
        const int q = T - w;
        if (q < 0) break;
        if (q >= Q) continue;
        const int row0 = w << 6, rows_end = min(row0 + 64, m);
        // chunks of this block that intersect the stripe (columns are 1-based: chunk q holds 32q+1 .. 32q+32)
        const int64_t jlo = (int64_t)row0 + 1 - lo, jhi = (int64_t)rows_end + hi;
        const int64_t qlo = jlo <= 1 ? 0 : (jlo - 1) >> 5, qhi = min((int64_t)Q - 1, (jhi - 1) >> 5);
        if (q < qlo || q > qhi) continue;
        bool above = false;                     // is block w-1 computed in chunk q?
        if (w > 0) {
          const int64_t ajlo = jlo - 64, ajhi = (int64_t)row0 + hi;
          const int64_t aqlo = ajlo <= 1 ? 0 : (ajlo - 1) >> 5, aqhi = min((int64_t)Q - 1, (ajhi - 1) >> 5);
          above = q >= aqlo && q <= aqhi;
        }
        const int cmax = min(32, n - (q << 5));
        uint32_t hp_in = 0xffffffffu, hm_in = 0u;   // top boundary D[0][j] = j, or an upper bound below a block that left the stripe
        int above_after = 0;
        if (above) {
          const int32_t* c = carry + ((size_t)(par ^ 1) * W + (w - 1)) * 3;
          hp_in = (uint32_t)c[0]; hm_in = (uint32_t)c[1]; above_after = c[2];
        }
        unsigned long long P, M;
        int score;
        if (q == qlo) {                         // the block enters the stripe: vertical deltas +1
          P = ~0ull; M = 0ull;
          const uint32_t cm = cmax == 32 ? 0xffffffffu : ((1u << cmax) - 1u);
          const int top = q == 0 ? row0 : above_after - (__popc(hp_in & cm) - __popc(hm_in & cm));
          score = top + (rows_end - row0);
        } else {
          P = pv[w]; M = mv[w];
          score = carry[((size_t)(par ^ 1) * W + w) * 3 + 2];
        }
        const int topbit = rows_end - row0 - 1;
        const uint8_t* tc = B + (q << 5);
        uint32_t hp_out = 0u, hm_out = 0u;
#pragma unroll 4
        for (int c = 0; c < cmax; ++c) {
          const unsigned long long eq0 = peq[(size_t)tc[c] * W + w];
          const unsigned long long hin_p = (hp_in >> c) & 1u, hin_n = (hm_in >> c) & 1u;
          const unsigned long long Xv = eq0 | M;
          const unsigned long long eq = eq0 | hin_n;
          const unsigned long long Xh = (((eq & P) + P) ^ P) | eq;
          unsigned long long Ph = M | ~(Xh | P);
          unsigned long long Mh = P & Xh;
          hp_out |= (uint32_t)((Ph >> topbit) & 1ull) << c;
          hm_out |= (uint32_t)((Mh >> topbit) & 1ull) << c;
          Ph = (Ph << 1) | hin_p;
          Mh = (Mh << 1) | hin_n;
          P = Mh | ~(Xv | Ph);
          M = Ph & Xv;
        }
        score += __popc(hp_out) - __popc(hm_out);
        pv[w] = P; mv[w] = M;
        int32_t* o = carry + ((size_t)par * W + w) * 3;
        o[0] = (int32_t)hp_out; o[1] = (int32_t)hm_out; o[2] = score;
      }
      ed_sync<NW>();
    }
    result = carry[((size_t)((Tn - 1) & 1) * W + (W - 1)) * 3 + 2];   // block W-1 is in the stripe at the last chunk (hi >= delta)
    if ((int64_t)result <= k) break;
    k <<= 3;
    ed_sync<NW>();                              // everyone has read the result before the ring is reused
  }
  if (l == 0) dist[p] = result;
}

extern "C" int vsv_edit_distance_batch(vsv_ctx* ctx, const uint8_t* chars, const int64_t* str_off, int64_t n_str,
                                       const int32_t* ia, const int32_t* ib, int64_t n_pair, int32_t* dist) {
  VSV_TRY(vsv_begin(ctx));
  if (n_str < 0 || n_pair < 0 || (n_pair > 0 && (!str_off || !ia || !ib || !dist)))
    VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: bad argument");
  if (n_pair == 0) return VSV_OK;
  if (ctx->device_io) VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: host buffers only");
  if (n_pair > INT32_MAX) VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: too many pairs");
  const int64_t n_chars = str_off[n_str];
  // ---- one symbol table for the call: the distinct byte values of the strings, in byte order -----------------
  bool present[256] = {false};
  for (int64_t i = 0; i < n_chars; ++i) present[chars[i]] = true;
  uint8_t code_of[256];
  int nsym = 0;
  for (int v = 0; v < 256; ++v) {
    code_of[v] = (uint8_t)nsym;
    if (present[v]) ++nsym;
  }
  if (nsym > ED_MAX_SYM) VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: more than 32 distinct byte values in the string table");
  if (nsym == 0) nsym = 1;
  std::vector<uint8_t> codes((size_t)n_chars + 1);
  for (int64_t i = 0; i < n_chars; ++i) codes[(size_t)i] = code_of[chars[i]];
  // ---- classes: short pairs one per warp, long pairs one per CTA; longest first -------------------------------
  struct Item { int32_t p; int32_t W; };
  std::vector<Item> shorts, longs;
  for (int64_t p = 0; p < n_pair; ++p) {
    if (ia[p] < 0 || ia[p] >= n_str || ib[p] < 0 || ib[p] >= n_str)
      VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: string index out of range");
    const int64_t la = str_off[ia[p] + 1] - str_off[ia[p]], lb = str_off[ib[p] + 1] - str_off[ib[p]];
    if (la < 0 || lb < 0 || la >= ED_MAX_LEN || lb >= ED_MAX_LEN)
      VSV_FAIL(ctx, VSV_E_ARG, "vsv_edit_distance_batch: string length out of range");
    const int32_t W = (int32_t)((std::min(la, lb) + 63) / 64);
    (W <= ED_SHORT_BLOCKS ? shorts : longs).push_back({(int32_t)p, W});
  }
  auto by_size = [](const Item& a, const Item& b) { return a.W > b.W; };
  std::stable_sort(shorts.begin(), shorts.end(), by_size);
  std::stable_sort(longs.begin(), longs.end(), by_size);
  std::vector<int32_t> list_s(shorts.size()), list_l(longs.size());
  std::vector<int64_t> goff(longs.size() + 1, -1);
  int64_t words_s = 0, words_l = 0, gwords = 0;
  for (size_t i = 0; i < shorts.size(); ++i) {
    list_s[i] = shorts[i].p;
    words_s = std::max(words_s, ed_words(shorts[i].W, nsym));
  }
  for (size_t i = 0; i < longs.size(); ++i) {
    list_l[i] = longs[i].p;
    const int64_t w = ed_words(longs[i].W, nsym);
    if (w * 8 > ED_LONG_SMEM) {
      goff[i] = gwords;
      gwords += w;
    } else {
      words_l = std::max(words_l, w);
    }
  }
  EdIn in;
  VSV_TRY(vsv_in(ctx, codes.data(), (size_t)n_chars + 1, &in.codes));
  VSV_TRY(vsv_in(ctx, str_off, (size_t)n_str + 1, &in.str_off));
  VSV_TRY(vsv_in(ctx, ia, (size_t)n_pair, &in.ia));
  VSV_TRY(vsv_in(ctx, ib, (size_t)n_pair, &in.ib));
  const int32_t *d_ls = nullptr, *d_ll = nullptr;
  const int64_t* d_goff = nullptr;
  if (!list_s.empty()) VSV_TRY(vsv_in(ctx, list_s.data(), list_s.size(), &d_ls));
  if (!list_l.empty()) {
    VSV_TRY(vsv_in(ctx, list_l.data(), list_l.size(), &d_ll));
    VSV_TRY(vsv_in(ctx, goff.data(), goff.size(), &d_goff));
  }
  int32_t* d_dist;
  VSV_TRY(vsv_out(ctx, dist, (size_t)n_pair, &d_dist));
  unsigned long long* scratch = nullptr;
  VSV_TRY(vsv_tmp(ctx, (size_t)gwords + 1, &scratch));
  // the staged host vectors must outlive the async copies: the call synchronises before it returns (vsv_end below)
  VSV_TRY(vsv_t0(ctx));
  if (!list_s.empty()) {
    const size_t smem = (size_t)words_s * 8 * 4;
    VSV_CUDA(ctx, cudaFuncSetAttribute(k_edit_banded<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    VSV_LAUNCH(ctx, k_edit_banded<1>, (unsigned)vsv_cdiv((int64_t)list_s.size(), 4), 128, smem, in, d_ls, (int)list_s.size(),
               nsym, (int)words_s, scratch, (const int64_t*)nullptr, d_dist);
  }
  if (!list_l.empty()) {
    const size_t smem = (size_t)words_l * 8;
    VSV_CUDA(ctx, cudaFuncSetAttribute(k_edit_banded<ED_LONG_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    VSV_LAUNCH(ctx, k_edit_banded<ED_LONG_WARPS>, (unsigned)list_l.size(), ED_LONG_WARPS * 32, smem, in, d_ll, (int)list_l.size(),
               nsym, (int)words_l, scratch, d_goff, d_dist);
  }
  VSV_TRY(vsv_t1(ctx));
  VSV_TRY(vsv_out_done(ctx, dist, d_dist, (size_t)n_pair));
  VSV_TRY(vsv_end(ctx, true));
  return VSV_OK;
}
