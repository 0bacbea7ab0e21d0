/* TEST INFRASTRUCTURE -- C twin of oracle/vsv_oracle.py for the volume stages.
 *
 * Used only by tests/ (large-size parity), __graft_entry__.smoke() and bench.py's
 * CPU-baseline legs.  Never linked into the product library.  Each function restates
 * the reference loop it cites (LI/ = bin/VolcanoSV-vc/Large_INDEL/) on the packed
 * layout of include/volcanosv_b200.h; it is checked against oracle/vsv_oracle.py, which
 * is pinned by the reference-generated fixtures in tests/golden/.
 *
 * Build: make -C oracle   (gcc -O2 -pthread -shared; libgomp is absent from the image,
 * so multi-threading is a small pthread fork-join)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

typedef struct { int32_t rec, type, pos, len, q0, q1, first, count; } sig_t;

static const uint32_t REF_MASK[3] = {(1u<<0)|(1u<<2), (1u<<0)|(1u<<7)|(1u<<8)|(1u<<2)|(1u<<3), (1u<<0)|(1u<<2)|(1u<<7)|(1u<<8)};
static const uint32_t QRY_MASK[3] = {(1u<<0)|(1u<<1)|(1u<<4), (1u<<0)|(1u<<7)|(1u<<8)|(1u<<4)|(1u<<1), 0xFFFFu & ~(1u<<2)};
static const uint32_t AUX_MASK[3] = {(1u<<5), (1u<<5), (1u<<3)};
static const int HSIGN[3] = {1, 1, -1};

int orc_num_threads(void) {
  long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n > 0 ? (int)n : 1;
}

/* fork-join: fn(t, threads, arg) for t in [0, threads) */
typedef void (*par_fn)(int t, int threads, void* arg);
typedef struct { par_fn fn; int t, threads; void* arg; } par_job;
static void* par_entry(void* p) { par_job* j = (par_job*)p; j->fn(j->t, j->threads, j->arg); return NULL; }
static void run_parallel(int threads, par_fn fn, void* arg) {
  if (threads <= 1) { fn(0, 1, arg); return; }
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  par_job* jobs = (par_job*)malloc(sizeof(par_job) * (size_t)threads);
  for (int t = 0; t < threads; ++t) {
    jobs[t].fn = fn; jobs[t].t = t; jobs[t].threads = threads; jobs[t].arg = arg;
    pthread_create(&th[t], NULL, par_entry, &jobs[t]);
  }
  for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
  free(th); free(jobs);
}

/* one record: ECS_H:53-85 / ERS:47-83 / SE:438-478 (walk only) */
static int64_t scan_record(int mode, int min_len, const uint32_t* ops, int64_t a, int64_t b, int32_t pos, int emit,
                           int32_t rec, sig_t* out, int64_t cap, int64_t k, int32_t* rt, int32_t* qt, int32_t* xt) {
  int64_t rs = 0, qs = 0, xs = 0;
  int32_t hc = 0;
  if (b > a && (ops[a] & 15u) == 5u) hc = (int32_t)(ops[a] >> 4);
  for (int64_t i = a; i < b; ++i) {
    uint32_t op = ops[i] & 15u;
    int64_t ln = ops[i] >> 4;
    if (emit && ln >= min_len && (op == 1u || op == 2u)) {
      if (out && k < cap) {
        sig_t s;
        s.rec = rec; s.type = op == 1u; s.pos = (int32_t)(pos + rs); s.len = (int32_t)ln;
        s.q0 = (int32_t)(qs + HSIGN[mode] * hc); s.q1 = s.q0 + (op == 1u ? (int32_t)ln : 1);
        s.first = 0; s.count = 1;
        out[k] = s;
      }
      ++k;
    }
    if ((REF_MASK[mode] >> op) & 1u) rs += ln;
    if ((QRY_MASK[mode] >> op) & 1u) qs += ln;
    if ((AUX_MASK[mode] >> op) & 1u) xs += ln;
  }
  if (rt) { rt[rec] = (int32_t)rs; qt[rec] = (int32_t)qs; xt[rec] = (int32_t)xs; }
  return k;
}

typedef struct {
  int mode, min_len, min_mapq;
  const uint32_t* ops; const int64_t* op_off; const int32_t* pos; const uint8_t* mapq; const uint8_t* enable;
  int64_t n_rec; sig_t* out; int64_t cap; int32_t *rt, *qt, *xt; int64_t* cnt; int pass;
} scan_job;
static void scan_part(int t, int threads, void* arg) {
  scan_job* j = (scan_job*)arg;
  int64_t r0 = j->n_rec * t / threads, r1 = j->n_rec * (t + 1) / threads, k = j->pass ? j->cnt[t] : 0;
  for (int64_t r = r0; r < r1; ++r) {
    int emit = j->mapq[r] >= j->min_mapq && (!j->enable || j->enable[r]);
    if (j->pass == 0)
      k = scan_record(j->mode, j->min_len, j->ops, j->op_off[r], j->op_off[r + 1], j->pos[r], emit, (int32_t)r, NULL, 0, k,
                      j->rt, j->qt, j->xt);
    else
      k = scan_record(j->mode, j->min_len, j->ops, j->op_off[r], j->op_off[r + 1], j->pos[r], emit, (int32_t)r, j->out,
                      j->cap, k, NULL, NULL, NULL);
  }
  if (j->pass == 0) j->cnt[t + 1] = k;
}

/* threads<=1: plain loop.  threads>1: records are cut into `threads` contiguous chunks,
 * counted, then written at the chunk's offset (order preserved). */
int64_t orc_scan_cigar(int mode, int min_len, int min_mapq, const uint32_t* ops, const int64_t* op_off,
                       const int32_t* pos, const uint8_t* mapq, const uint8_t* enable, int64_t n_rec, sig_t* out,
                       int64_t cap, int32_t* rt, int32_t* qt, int32_t* xt, int threads) {
  if (threads <= 1) {
    int64_t k = 0;
    for (int64_t r = 0; r < n_rec; ++r) {
      int emit = mapq[r] >= min_mapq && (!enable || enable[r]);
      k = scan_record(mode, min_len, ops, op_off[r], op_off[r + 1], pos[r], emit, (int32_t)r, out, cap, k, rt, qt, xt);
    }
    return k;
  }
  scan_job j = {mode, min_len, min_mapq, ops, op_off, pos, mapq, enable, n_rec, out, cap, rt, qt, xt, NULL, 0};
  j.cnt = (int64_t*)calloc((size_t)threads + 1, sizeof(int64_t));
  run_parallel(threads, scan_part, &j);
  for (int t = 0; t < threads; ++t) j.cnt[t + 1] += j.cnt[t];
  j.pass = 1;
  run_parallel(threads, scan_part, &j);
  int64_t total = j.cnt[threads];
  free(j.cnt);
  return total;
}

/* generate_combine_sigs (SE:373-435) over a type-separated, record-ordered list */
int64_t orc_merge_cute(const sig_t* s, int64_t n, int type, sig_t* out) {
  int64_t k = 0, i = 0;
  while (i < n) {
    int64_t e = i + 1;
    while (e < n && s[e].rec == s[i].rec) ++e;
    sig_t cur = s[i];
    cur.first = (int32_t)i; cur.count = 1;
    int32_t mark = type == 1 ? cur.pos : cur.pos + cur.len;
    for (int64_t j = i + 1; j < e; ++j) {
      int hit = type == 1 ? (s[j].pos - mark <= 100) : (s[j].pos - mark <= 0);
      if (hit) {
        cur.len += s[j].len; cur.count++;
        mark = type == 1 ? s[j].pos : s[j].pos + s[j].len;
      } else {
        out[k++] = cur;
        cur = s[j]; cur.first = (int32_t)j; cur.count = 1;
        mark = s[j].pos;  /* SE:429-430 end := start */
      }
    }
    out[k++] = cur;
    i = e;
  }
  return k;
}

/* count_reads_span_region (GTD:140-146): the reference fetches the records overlapping
 * [S,E) through the BAM index and tests start<S && end>E.  Port: reads sorted by start,
 * candidates = reads with start in (S-max_len, S). */
static int32_t span_one(const int32_t* rs, const int32_t* re, int64_t n, int32_t max_len, int32_t S, int32_t E) {
  int64_t lo = 0, hi = n;
  while (lo < hi) { int64_t m = (lo + hi) >> 1; if (rs[m] < S) lo = m + 1; else hi = m; }
  int32_t c = 0;
  for (int64_t i = lo - 1; i >= 0 && (int64_t)rs[i] > (int64_t)S - max_len - 1; --i)
    if (re[i] > E) ++c;
  return c;
}

typedef struct { const int32_t *rs, *re; int64_t n; const int32_t *S, *E; int64_t nw; int32_t* depth; int32_t max_len; } depth_job;
static void depth_part(int t, int threads, void* arg) {
  depth_job* j = (depth_job*)arg;
  for (int64_t w = j->nw * t / threads; w < j->nw * (t + 1) / threads; ++w)
    j->depth[w] = span_one(j->rs, j->re, j->n, j->max_len, j->S[w], j->E[w]);
}
void orc_span_depth(const int32_t* rs, const int32_t* re, int64_t n, const int32_t* S, const int32_t* E, int64_t nw,
                    int32_t* depth, int threads) {
  int32_t max_len = 0;
  for (int64_t i = 0; i < n; ++i) if (re[i] - rs[i] > max_len) max_len = re[i] - rs[i];
  depth_job j = {rs, re, n, S, E, nw, depth, max_len};
  run_parallel(threads, depth_part, &j);
}

/* check_full_cover_reads + ratio + correct_gt_eval (GTD:149-170,206-217,291-318; GTI:188-196,228,258-283) */
typedef struct {
  const int32_t *rs, *re; int64_t n; const int32_t *cpos, *clen, *ctype, *cgt, *nsup; int64_t nc;
  const double *thr_del, *thr_ins; int32_t* depth2; double* ratio; int32_t* new_gt; int32_t max_len;
} gt_job;
static void gt_part(int t, int threads, void* arg) {
  gt_job* j = (gt_job*)arg;
  const int32_t *rs = j->rs, *re = j->re;
  const int64_t n = j->n;
  const int32_t max_len = j->max_len;
  for (int64_t i = j->nc * t / threads; i < j->nc * (t + 1) / threads; ++i) {
    int32_t p = j->cpos[i], L = j->clen[i], d2;
    if (j->ctype[i] == 1) d2 = 2 * span_one(rs, re, n, max_len, p - 100, p + 100);
    else if (L <= 1000) d2 = 2 * span_one(rs, re, n, max_len, p, p + L);
    else d2 = span_one(rs, re, n, max_len, p - 150, p - 50) + span_one(rs, re, n, max_len, p + L + 150, p + L + 250);
    double depth = d2 / 2.0, r;
    int32_t ns = j->nsup[i];
    if (j->ctype[i] == 0) r = d2 == 0 ? 1.0 : ns / depth;
    else r = d2 == 0 ? (ns == 0 ? NAN : INFINITY) : ns / depth;
    int32_t g = j->cgt[i], o = g;
    if (g == 0 || g == 1) {
      const double* t4 = j->ctype[i] == 0 ? j->thr_del : j->thr_ins;
      double th = t4[(L > 1000 ? 0 : 1) + (g == 1 ? 0 : 2)];
      if (!isnan(th)) { if (r > th) o = 1; else if (r <= th) o = 0; }
    }
    j->depth2[i] = d2; j->ratio[i] = r; j->new_gt[i] = o;
  }
}
void orc_genotype(const int32_t* rs, const int32_t* re, int64_t n, const int32_t* cpos, const int32_t* clen,
                  const int32_t* ctype, const int32_t* cgt, const int32_t* nsup, int64_t nc, const double* thr_del,
                  const double* thr_ins, int32_t* depth2, double* ratio, int32_t* new_gt, int threads) {
  int32_t max_len = 0;
  for (int64_t i = 0; i < n; ++i) if (re[i] - rs[i] > max_len) max_len = re[i] - rs[i];
  gt_job j = {rs, re, n, cpos, clen, ctype, cgt, nsup, nc, thr_del, thr_ins, depth2, ratio, new_gt, max_len};
  run_parallel(threads, gt_part, &j);
}

/* match_varlist_siglist (GTD:92-137), literal: forward and backward scans from
 * last_match_idx, breaks only on same-chromosome rows (single chromosome here). */
void orc_sig_support(const int32_t* cpos, const int32_t* clen, int64_t nc, const int32_t* sp, const int32_t* sl,
                     const int32_t* sn, int64_t ns, int32_t* support) {
  int64_t last = 0;
  for (int64_t k = 0; k < nc; ++k) {
    double sv = clen[k], max_size = sv / 0.6, min_size = sv * 0.6, shift = fmax(sv * 2.3, 500.0);
    double lo = cpos[k] - shift, hi = cpos[k] + shift;
    int64_t mn = -1;
    int32_t s = 0;
    for (int64_t i = last; i < ns; ++i) {
      if (lo <= sp[i] && sp[i] <= hi) { if (mn < 0 || i < mn) mn = i; if (min_size <= sl[i] && sl[i] <= max_size) s += sn[i]; }
      else if (sp[i] > hi) break;
    }
    for (int64_t i = last < ns ? last : ns - 1; i >= 0; --i) {
      if (lo <= sp[i] && sp[i] <= hi) { if (mn < 0 || i < mn) mn = i; if (min_size <= sl[i] && sl[i] <= max_size) s += sn[i]; }
      else if (sp[i] < lo) break;
    }
    if (mn >= 0) last = mn;
    support[k] = s;
  }
}

/* eval_sig (FPF:106-123) on a position-sorted signature list */
void orc_fp_support(const int32_t* cpos, const int32_t* clen, int64_t nc, const int32_t* sp, const int32_t* sl,
                    int64_t ns, int32_t* support) {
  for (int64_t k = 0; k < nc; ++k) {
    if (clen[k] > 250) { support[k] = 60; continue; }
    int64_t lo = 0, hi = ns;
    while (lo < hi) { int64_t m = (lo + hi) >> 1; if ((int64_t)sp[m] < (int64_t)cpos[k] - 1000) lo = m + 1; else hi = m; }
    int32_t s = 0;
    for (int64_t i = lo; i < ns && (int64_t)sp[i] - cpos[k] <= 1000; ++i) {
      int64_t sh = llabs((int64_t)sp[i] - cpos[k]);
      int32_t mx = clen[k] > sl[i] ? clen[k] : sl[i], mi = clen[k] < sl[i] ? clen[k] : sl[i];
      if (sh <= 500 && mx != 0 && 2 * (int64_t)mi >= mx) ++s;
    }
    support[k] = s;
  }
}

/* calc_ins_call_cov (CSS:81-125) / calc_del_call_cov (CSS:138-253) */
void orc_cov(int is_del, const int32_t* cs, const int32_t* ce, int64_t nc, const int32_t* sp, const int32_t* sl,
             int64_t ns, int64_t* cov) {
  int32_t max_len = 0;
  for (int64_t i = 0; i < ns; ++i) if (sl[i] > max_len) max_len = sl[i];
  for (int64_t k = 0; k < nc; ++k) {
    int64_t lb = (int64_t)cs[k] - 1000, rb = (int64_t)(is_del ? ce[k] : cs[k]) + 1000, tot = 0;
    int64_t lo = 0, hi = ns, from = lb - (is_del ? max_len : 0);
    while (lo < hi) { int64_t m = (lo + hi) >> 1; if ((int64_t)sp[m] < from) lo = m + 1; else hi = m; }
    for (int64_t i = lo; i < ns && sp[i] <= rb; ++i) {
      if (is_del) { if ((int64_t)sp[i] + sl[i] >= lb) tot -= sl[i]; }
      else if (sp[i] >= lb) tot += sl[i];
    }
    cov[k] = tot;
  }
}
