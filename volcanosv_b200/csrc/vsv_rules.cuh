// Device-side restatements of the reference's per-record / per-pair RULES, shared by the per-function
// C-ABI files (vsv_merge.cu, vsv_stage2.cu, vsv_sa.cu) and the resident shard engine (vsv_shard.cu):
//   mstate_push / merge_walk2   cluster_*_one_read (ECS_H:87-164), generate_combine_sigs (SE:373-435)
//   split_rule                  extract_sig_from_split (ECS_H:307-371, ECS_O:307-382, ECS_C:328-402, ERS:147-197)
//   len_string_key              `.sigs` tie order (SE:637-638, LC_ALL=C)
//   sa_insert / sa_analyze      organize_split_signal + analysis_split_read (SE:193-371)
#pragma once
#include "vsv_common.cuh"

// State of one type's merge inside a record (the reference keeps one list per type, so DEL merging
// never sees the INS signatures in between and vice versa).
struct MState {
  vsv_sig cur;
  int32_t mark;  // cute: last INS pos / DEL end
  bool open;
};

// Feeds the next same-type signature `n` (index `idx` in the type's own list); calls f(merged)
// whenever a merged signature is complete.
template <typename F>
__device__ __forceinline__ void mstate_push(MState& st, const vsv_sig& n, int32_t idx, int mode, int type, F f) {
  vsv_sig& cur = st.cur;
  if (!st.open) {
    cur = n; cur.first = idx; cur.count = 1;
    st.mark = (type == VSV_INS) ? n.pos : n.pos + n.len;
    st.open = true;
    return;
  }
  bool hit = false;
  if (mode == VSV_MODE_CONTIG) {
    const int32_t d = abs(n.pos - cur.pos);
    if (type == VSV_INS) {
      hit = (cur.len > 250 && n.len > 250 && d < 250) || (cur.len > 320 && n.len > 320 && d < 380) ||
            (cur.len > 100 && n.len > 100 && d < 250);           // ECS_H:112-135
      if (hit) { cur.q1 = n.q1; cur.len = cur.q1 - cur.q0; }     // ECS_H:91-97
    } else {
      hit = cur.len > 150 && n.len > 150 && d < 150;             // ECS_H:148-150
      if (hit) { cur.len = n.pos + n.len - cur.pos; cur.q1 = cur.q0 + 1; }  // ECS_H:99-106
    }
  } else if (mode == VSV_MODE_CUTE) {
    if (type == VSV_INS) {
      hit = n.pos - st.mark <= 100;                              // SE:396-399
      if (hit) { cur.len += n.len; st.mark = n.pos; }
    } else {
      hit = n.pos - st.mark <= 0;                                // SE:420
      if (hit) { cur.len += n.len; st.mark = n.pos + n.len; }
    }
  }                                                              // VSV_MODE_READS: never merged
  if (hit) {
    cur.count++;
  } else {
    f(cur);
    cur = n; cur.first = idx; cur.count = 1;
    st.mark = n.pos;                                             // SE:412-413 / SE:429-430 (end := start quirk)
  }
}

// Walks one record's run [i0,i1) of signatures (op order, both types interleaved) with one state per
// type.  in_del / in_ins: how many DEL / INS signatures precede the run (the `first` indices count
// inside each type's own list).  fd / fi receive the merged DEL / INS signatures in order.
template <typename FD, typename FI>
__device__ __forceinline__ void merge_walk2(const vsv_sig* __restrict__ s, int64_t i0, int64_t i1, int mode, int32_t in_del,
                                            int32_t in_ins, FD fd, FI fi) {
  MState sd, si;
  sd.open = si.open = false;
  for (int64_t k = i0; k < i1; ++k) {
    const int4 a = reinterpret_cast<const int4*>(s + k)[0], b = reinterpret_cast<const int4*>(s + k)[1];
    vsv_sig n;
    n.rec = a.x; n.type = a.y; n.pos = a.z; n.len = a.w; n.q0 = b.x; n.q1 = b.y; n.first = b.z; n.count = b.w;
    if (n.type == VSV_DEL) mstate_push(sd, n, in_del++, mode, VSV_DEL, fd);
    else mstate_push(si, n, in_ins++, mode, VSV_INS, fi);
  }
  if (sd.open) fd(sd.cur);
  if (si.open) fi(si.cur);
}


__device__ __forceinline__ bool is_clip(uint32_t op) { return (op & 15u) == 4u || (op & 15u) == 5u; }

// One consecutive same-name record pair -> at most one signature.
// ECS_H:307-371 / ECS_O:307-382 / ECS_C:328-402 / ERS:147-197.
__device__ __forceinline__ bool split_rule(int variant, int min_mapq, int32_t end1, int32_t rl1, uint32_t last1, int mq1, int rv1,
                           int32_t pos2, uint32_t first2, int mq2, int rv2, vsv_sig& o) {
  if (rv1 != rv2 || mq1 < min_mapq || mq2 < min_mapq) return false;
  if (!is_clip(last1) || !is_clip(first2)) return false;
  const int64_t Ref1e = end1, Ref2s = pos2;
  const int64_t Read1e = (int64_t)rl1 - (int64_t)(last1 >> 4);
  const int64_t Read2s = (int64_t)(first2 >> 4);
  const int64_t D = (Ref2s - Ref1e) - (Read2s - Read1e);
  if (llabs(D) > 50000) return false;
  if (variant == VSV_SPLIT_READS) {
    const int64_t O = Ref1e - Ref2s;
    if (O < 30 && D >= 30) {
      o.type = VSV_DEL; o.pos = (int32_t)Ref1e; o.len = (int32_t)D; o.q0 = (int32_t)Read1e; o.q1 = (int32_t)Read2s;
      return true;
    }
    if (O < 30 && D <= -30) {
      o.type = VSV_INS; o.pos = (int32_t)((Ref1e + Ref2s) / 2); o.len = (int32_t)llabs(D);
      o.q0 = (int32_t)Read1e; o.q1 = (int32_t)Read2s;
      return true;
    }
    return false;
  }
  if (variant == VSV_SPLIT_HIFI) {
    if (D >= 30) {
      const int64_t O = Read1e - Read2s;
      if (llabs(O) <= 3000) {
        const int64_t h = O / 2;  // int(O/2): truncation toward zero, as C
        o.type = VSV_DEL; o.pos = (int32_t)(Ref1e - h); o.len = (int32_t)D;
        o.q0 = (int32_t)(Read1e - h); o.q1 = o.q0 + 1;
        return true;
      }
      return false;
    }
    if (D <= -30) {
      const int64_t O = Ref1e - Ref2s;
      if (O < 3000) {
        o.type = VSV_INS; o.len = (int32_t)llabs(Read2s - Read1e + O);
        o.pos = (int32_t)(llabs(O) > 400 ? (Ref1e + Ref2s) / 2 : Ref2s);
        o.q0 = (int32_t)(Read1e - O); o.q1 = (int32_t)Read2s;
        return true;
      }
    }
    return false;
  }
  const double r = variant == VSV_SPLIT_ONT ? 0.5 : 0.3;
  const double lo = variant == VSV_SPLIT_ONT ? 0.8 : r;
  if (D >= 30) {
    const int64_t O = Read1e - Read2s;
    const double dr = (double)D * r;
    if (-dr <= (double)O && (double)O <= dr) {
      o.type = VSV_DEL; o.pos = (int32_t)Ref1e; o.len = (int32_t)D; o.q0 = (int32_t)Read1e; o.q1 = (int32_t)Read2s;
      return true;
    }
    return false;
  }
  const int64_t O = Ref1e - Ref2s;
  if ((double)D * lo <= (double)O && (double)O <= (double)llabs(D) * r && D <= -30) {
    o.type = VSV_INS; o.len = (int32_t)llabs(Read2s - Read1e + O);
    o.pos = (int32_t)(llabs(O) > 400 ? (Ref1e + Ref2s) / 2 : Ref2s);
    o.q0 = (int32_t)(Read1e - O); o.q1 = (int32_t)Read2s;
    return true;
  }
  return false;
}


__device__ __forceinline__ unsigned long long len_string_key(int32_t len) {
  unsigned long long v = (unsigned long long)(uint32_t)max(len, 0);
  int digits = 1;
  for (unsigned long long t = v; t >= 10ull; t /= 10ull) ++digits;
  unsigned long long scaled = v;
  for (int d = digits; d < 10; ++d) scaled *= 10ull;  // left-justify to 10 digits
  return (scaled << 4) | (unsigned long long)digits;
}

#define SA_MAXSEG 16

struct SaSeg {
  int64_t qs, qe, rs, re;
  int32_t chr, st;
};

struct SaIn {
  const int64_t* seg_off;
  const int32_t *q_s, *q_e, *r_s, *r_e, *chrom, *read_len;
  const uint8_t* strand;
};

__device__ __forceinline__ SaSeg sa_mirror(const SaSeg& e, int64_t L) {  // SE: '-' strand -> query mirrored
  SaSeg o = e;
  o.qs = L - e.qe;
  o.qe = L - e.qs;
  return o;
}

struct SaEmit {
  vsv_sig* dst;
  int32_t rec, n, rev;
  int64_t sv_size, max_size;
  __device__ __forceinline__ void put(int type, int64_t pos, int64_t len, int64_t q0, int64_t q1) {
    vsv_sig o;
    o.rec = rec; o.type = type; o.pos = (int32_t)pos; o.len = (int32_t)len;
    o.q0 = (int32_t)q0; o.q1 = (int32_t)q1; o.first = rev; o.count = 1;
    dst[n++] = o;
  }
  // SE:217-232 (pair) / SE:255-293 (triple): INS then DEL test of one ordered pair
  __device__ __forceinline__ void ins_del(const SaSeg& e1, const SaSeg& e2, bool guard) {
    if (e1.re - e2.rs < sv_size) {
      const int64_t il = e2.qs + e1.re - e2.rs - e1.qe;
      if (il >= sv_size && e2.rs - e1.re <= 100 && (il <= max_size || max_size == -1) && guard) {
        const int64_t h = (e2.rs - e1.re) / 2;  // int(x/2): truncation toward zero == C
        put(VSV_INS, (e2.rs + e1.re) / 2, il, e1.qe + h, e2.qs - h);
      }
      const int64_t dl = e2.rs - e2.qs + e1.qe - e1.re;
      if (dl >= sv_size && e2.qs - e1.qe <= 100 && (dl <= max_size || max_size == -1) && guard)
        put(VSV_DEL, e1.re, dl, 0, 0);
    }
  }
};

// stable insertion of segment s as the i-th element (sorted by query start == Python's sorted, SE:366)
__device__ __forceinline__ void sa_insert(SaSeg* sp, int i, const SaSeg& s) {
  int j = i;
  while (j > 0 && sp[j - 1].qs > s.qs) { sp[j] = sp[j - 1]; --j; }
  sp[j] = s;
}

// analysis_split_read (SE:193-326) on k query-sorted segments of a read of length L
__device__ __forceinline__ void sa_analyze(const SaSeg* sp, int k, int64_t L, SaEmit& em) {
  bool trig = false;
  if (k == 2) {
    if (sp[0].chr == sp[1].chr) {
      if (sp[0].st == sp[1].st) {
        SaSeg e1 = sp[0], e2 = sp[1];
        if (e1.st == 1) { e1 = sa_mirror(sp[1], L); e2 = sa_mirror(sp[0], L); em.rev ^= 1; }
        em.ins_del(e1, e2, true);
      }
    } else {
      trig = true;
    }
  } else {
    for (int i = 0; i + 2 < k; ++i) {
      if (sp[i].chr == sp[i + 1].chr) {
        if (sp[i + 1].chr == sp[i + 2].chr && sp[i].st == sp[i + 2].st && sp[i].st == sp[i + 1].st) {
          SaSeg e1 = sp[i], e2 = sp[i + 1], e3 = sp[i + 2];
          if (e1.st == 1) {
            e1 = sa_mirror(sp[i + 2], L); e2 = sa_mirror(sp[i + 1], L); e3 = sa_mirror(sp[i], L);
            em.rev ^= 1;
          }
          em.ins_del(e1, e2, e3.rs >= e2.re);
          if (k - 3 == i) em.ins_del(e2, e3, true);
        }
      } else {
        trig = true;
      }
    }
  }
  if (k >= 3 && trig && sp[0].chr == sp[k - 1].chr && sp[0].st == sp[k - 1].st) {  // SE:303-326
    SaSeg e1 = sp[0], e2 = sp[k - 1];
    if (sp[0].st == 1) { e1 = sa_mirror(sp[k - 1], L); e2 = sa_mirror(sp[0], L); em.rev ^= 1; }
    const int64_t dis_ref = e2.rs - e1.re, dis_read = e2.qs - e1.qe, d = dis_read - dis_ref;
    if (dis_ref < 100 && d >= em.sv_size && (d <= em.max_size || em.max_size == -1)) {
      const int64_t h = dis_ref / 2;
      em.put(VSV_INS, min(e2.rs, e1.re), d, e1.qe + h, e2.qs - h);
    }
  }
}
