// Host-side BAM feeder: BGZF inflate + BAM record -> the packed
// struct-of-arrays the scanners read, without pysam/htslib.  Replaces, for whole-chromosome
// scans, `pysam.AlignmentFile(path).fetch(chr)` + the per-record attribute reads at
// LI/extract_reads_signature.py:107-145, LI/sig_extract.py:514-560 and
// LI/extract_contig_signature_Hifi.py:386-419 (read.cigar, .pos, .mapq, .flag, .qname,
// .query_length, .seq, get_tag("SA")).  Pure host code (zlib + std::thread); no CUDA calls.
//
// Format notes (SAM/BAM specification v1.6): a BAM file is a series of BGZF blocks (gzip members
// with a "BC" extra field holding the block size); the inflated stream is "BAM\1", header text,
// reference list, then records: block_size, refID, pos, l_read_name, mapq, bin, n_cigar_op, flag,
// l_seq, next_refID, next_pos, tlen, read_name, cigar (u32 = len<<4|op: the layout the kernels
// use), seq (4 bit), qual, aux.  CIGARs with more than 65535 ops live in a CG:B,I aux array behind
// a <l_seq>S<ref_len>N placeholder.
#include <sys/types.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/volcanosv_b200.h"

namespace {

// std::vector that does not zero-fill on resize (the bytes are overwritten right away by fread /
// inflate / memcpy; the zero fill was a serial first touch of every page)
template <class T>
struct NoInit : std::allocator<T> {
  template <class U> struct rebind { using other = NoInit<U>; };
  template <class U, class... A>
  void construct(U* p, A&&... a) {
    if constexpr (sizeof...(A) == 0) ::new ((void*)p) U;
    else ::new ((void*)p) U(std::forward<A>(a)...);
  }
};
typedef std::vector<uint8_t, NoInit<uint8_t>> Bytes;

struct BamRec {
  uint64_t off;      // offset of the record body (after block_size) in the inflated stream
  uint32_t size;     // block_size
};

struct RefInfo {
  std::string name;
  int64_t len;
  std::vector<uint64_t> recs;  // indices into vsv_bam::recs, file order
};

inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline int32_t rd32s(const uint8_t* p) { int32_t v; memcpy(&v, p, 4); return v; }
inline uint16_t rd16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }

// size of one aux value starting at p (type byte at p[0]); 0 on malformed input
size_t aux_size(const uint8_t* p, const uint8_t* end) {
  if (p >= end) return 0;
  const size_t room = (size_t)(end - p);
  switch (*p) {
    case 'A': case 'c': case 'C': return room >= 2 ? 2 : 0;
    case 's': case 'S': return room >= 3 ? 3 : 0;
    case 'i': case 'I': case 'f': return room >= 5 ? 5 : 0;
    case 'Z': case 'H': {
      const uint8_t* q = p + 1;
      while (q < end && *q) ++q;
      return q < end ? (size_t)(q - p) + 1 : 0;
    }
    case 'B': {
      if (p + 6 > end) return 0;
      size_t es = 0;
      switch (p[1]) {
        case 'c': case 'C': es = 1; break;
        case 's': case 'S': es = 2; break;
        case 'i': case 'I': case 'f': es = 4; break;
        default: return 0;
      }
      const size_t sz = 6 + es * (size_t)rd32(p + 2);
      return sz <= room ? sz : 0;   // a count that runs past the record is malformed, not a size
    }
    default: return 0;
  }
}

struct RecView {
  const uint8_t* body;
  int32_t ref_id, pos, l_seq;
  uint32_t l_name, mapq, n_cigar, flag;
  const char* name;
  const uint8_t* cigar;      // n_cigar little-endian u32 words, NOT 4-byte aligned (may point into the CG tag): read with memcpy
  const uint8_t* seq;
  const uint8_t* aux;
  const uint8_t* end;
};

// `has_qual`: records as they are in the file; the kept copies (vsv_bam::data) drop the qualities
bool view(const uint8_t* body, uint32_t size, RecView* v, bool has_qual = false) {
  if (size < 32) return false;
  v->body = body;
  v->end = body + size;
  v->ref_id = rd32s(body);
  v->pos = rd32s(body + 4);
  v->l_name = body[8];
  v->mapq = body[9];
  v->n_cigar = rd16(body + 12);
  v->flag = rd16(body + 14);
  v->l_seq = rd32s(body + 16);
  if (v->l_seq < 0) return false;
  const uint8_t* p = body + 32;
  v->name = reinterpret_cast<const char*>(p);
  p += v->l_name;
  v->cigar = p;
  if ((size_t)(v->end - body) < 32 + (size_t)v->l_name + 4 * (size_t)v->n_cigar) return false;
  p += 4 * (size_t)v->n_cigar;
  v->seq = p;
  if ((size_t)(v->end - p) < ((size_t)v->l_seq + 1) / 2 + (has_qual ? (size_t)v->l_seq : 0)) return false;
  p += ((size_t)v->l_seq + 1) / 2 + (has_qual ? (size_t)v->l_seq : 0);
  v->aux = p;
  // long CIGAR: <l_seq>S<n>N placeholder + CG:B,I
  if (v->n_cigar == 2) {
    uint32_t c0, c1;
    memcpy(&c0, v->cigar, 4);
    memcpy(&c1, v->cigar + 4, 4);
    if ((c0 & 15u) == 4u && (int32_t)(c0 >> 4) == v->l_seq && (c1 & 15u) == 3u) {
      for (const uint8_t* a = v->aux; a + 4 <= v->end;) {
        const size_t sz = aux_size(a + 2, v->end);
        if (!sz) return false;   // a malformed tag in a record that needs its CG tag: reject the record
        if (a[0] == 'C' && a[1] == 'G' && a[2] == 'B' && a[3] == 'I') {
          // aux_size has checked that the whole B array (8 + 4 * count bytes from `a`) lies inside the record
          v->n_cigar = rd32(a + 4);
          v->cigar = a + 8;
          break;
        }
        a += 2 + sz;
      }
    }
  }
  return true;
}

const char* find_sa(const RecView& v, size_t* len) {
  for (const uint8_t* a = v.aux; a + 3 <= v.end;) {
    const size_t sz = aux_size(a + 2, v.end);
    if (!sz) break;
    if (a[0] == 'S' && a[1] == 'A' && a[2] == 'Z') {
      *len = sz - 2;  // without type byte and NUL
      return reinterpret_cast<const char*>(a + 3);
    }
    a += 2 + sz;
  }
  *len = 0;
  return nullptr;
}

template <class F>
void parallel_for(int64_t n, int n_threads, F f) {
  n_threads = (int)std::max<int64_t>(1, std::min<int64_t>(n_threads, n));
  if (n_threads == 1) { f(0, n, 0); return; }
  std::vector<std::thread> th;
  const int64_t per = (n + n_threads - 1) / n_threads;
  for (int t = 0; t < n_threads; ++t) {
    const int64_t a = t * per, b = std::min(n, a + per);
    if (a < b) th.emplace_back([=] { f(a, b, t); });
  }
  for (auto& x : th) x.join();
}

}  // namespace

struct vsv_bam {
  Bytes data;                  // kept record bodies (qualities dropped), back to back
  std::vector<BamRec> recs;
  std::vector<RefInfo> refs;
  std::string header_text;
  std::string err;
  int n_threads = 1;
};

static thread_local std::string g_bam_err;

extern "C" const char* vsv_bam_last_error(void) { return g_bam_err.c_str(); }

namespace {

// Streaming BGZF reader: fill() reads the next stretch of compressed bytes, inflates the complete
// blocks in it in parallel and appends the result; memory is bounded by the window, not the file.
class BgzfReader {
 public:
  BgzfReader(FILE* f, int n_threads) : f_(f), nt_(n_threads) {}
  void seek(uint64_t coffset) {
    fseeko(f_, (off_t)coffset, SEEK_SET);
    raw_.clear();
    eof_ = false;
  }
  // 1 = appended something, 0 = end of file, -1 = error (err set)
  int fill(Bytes& out, size_t want, std::string& err) {
    for (;;) {
      if (eof_ && raw_.empty()) return 0;
      const size_t have = raw_.size();
      if (!eof_) {
        raw_.resize(have + want);
        const size_t got = fread(raw_.data() + have, 1, want, f_);
        raw_.resize(have + got);
        if (got < want) eof_ = true;
      }
      struct Blk { size_t in_off, in_len, out_off; uint32_t out_len; };
      std::vector<Blk> blks;
      size_t p = 0, total = 0;
      while (p + 18 <= raw_.size()) {
        const uint8_t* h = raw_.data() + p;
        if (h[0] != 31 || h[1] != 139 || h[2] != 8 || !(h[3] & 4)) { err = "not a BGZF file"; return -1; }
        const uint16_t xlen = rd16(h + 10);
        if (p + 12 + (size_t)xlen > raw_.size()) break;
        int bsize = -1;
        for (size_t x = 12; x + 4 <= 12 + (size_t)xlen;) {
          const uint16_t slen = rd16(h + x + 2);
          if (h[x] == 'B' && h[x + 1] == 'C' && slen == 2 && x + 6 <= 12 + (size_t)xlen) bsize = rd16(h + x + 4);
          x += 4 + slen;
        }
        if (bsize < 0 || (size_t)bsize + 1 < 12 + (size_t)xlen + 8) { err = "malformed BGZF block header"; return -1; }
        const size_t blen = (size_t)bsize + 1;
        if (p + blen > raw_.size()) break;
        const uint32_t isize = rd32(h + blen - 4);
        blks.push_back({p + 12 + xlen, blen - 12 - xlen - 8, total, isize});
        total += isize;
        p += blen;
      }
      if (blks.empty()) {
        if (eof_) {
          if (!raw_.empty()) { err = "truncated BGZF block"; return -1; }
          return 0;
        }
        continue;  // a block larger than what was read so far: read more
      }
      const size_t base = out.size();
      out.resize(base + total);
      const auto t_inf0 = std::chrono::steady_clock::now();
      std::atomic<int> bad(0);
      uint8_t* dst = out.data() + base;
      const uint8_t* src = raw_.data();
      parallel_for((int64_t)blks.size(), nt_, [&](int64_t a, int64_t e, int) {
        z_stream zs;
        for (int64_t i = a; i < e; ++i) {
          if (blks[i].out_len == 0) continue;
          memset(&zs, 0, sizeof(zs));
          if (inflateInit2(&zs, -15) != Z_OK) { bad = 1; return; }
          zs.next_in = const_cast<uint8_t*>(src) + blks[i].in_off;
          zs.avail_in = (uInt)blks[i].in_len;
          zs.next_out = dst + blks[i].out_off;
          zs.avail_out = blks[i].out_len;
          const int rc = inflate(&zs, Z_FINISH);
          inflateEnd(&zs);
          if (rc != Z_STREAM_END || zs.avail_out != 0) { bad = 1; return; }
        }
      });
      inflate_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t_inf0).count();
      if (bad) { err = "inflate failed (corrupt BGZF block)"; return -1; }
      raw_.erase(raw_.begin(), raw_.begin() + (ptrdiff_t)p);
      if (total == 0) continue;  // only empty blocks (the EOF marker): look further
      return 1;
    }
  }

 private:
  FILE* f_;
  int nt_;
  Bytes raw_;
  bool eof_ = false;

 public:
  double inflate_s = 0;
};

// smallest chunk start (virtual offset) of reference `tid` in a BAI index; false if the file is
// absent / unreadable; *has = false when the reference has no records
bool bai_ref_start(const std::string& bam_path, int tid, int n_ref_bam, uint64_t* voff, bool* has) {
  std::vector<std::string> cand = {bam_path + ".bai"};
  if (bam_path.size() > 4 && bam_path.compare(bam_path.size() - 4, 4, ".bam") == 0)
    cand.push_back(bam_path.substr(0, bam_path.size() - 4) + ".bai");
  for (const std::string& p : cand) {
    FILE* f = fopen(p.c_str(), "rb");
    if (!f) continue;
    std::vector<uint8_t> d;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    d.resize(sz > 0 ? (size_t)sz : 0);
    const bool ok = sz > 0 && fread(d.data(), 1, (size_t)sz, f) == (size_t)sz;
    fclose(f);
    if (!ok || d.size() < 8 || memcmp(d.data(), "BAI\1", 4) != 0) continue;
    size_t q = 4;
    const int32_t n_ref = rd32s(d.data() + q);
    q += 4;
    if (n_ref != n_ref_bam) continue;   // an index of another file
    bool bad = false;
    for (int32_t r = 0; r < n_ref && !bad; ++r) {
      if (q + 4 > d.size()) { bad = true; break; }
      const int32_t n_bin = rd32s(d.data() + q);
      q += 4;
      uint64_t best = ~0ull;
      for (int32_t k = 0; k < n_bin; ++k) {
        if (q + 8 > d.size()) { bad = true; break; }
        const uint32_t bin = rd32(d.data() + q);
        const int32_t n_chunk = rd32s(d.data() + q + 4);
        q += 8;
        if (n_chunk < 0 || q + 16 * (size_t)n_chunk > d.size()) { bad = true; break; }
        for (int32_t c = 0; c < n_chunk; ++c) {
          uint64_t beg;
          memcpy(&beg, d.data() + q + 16 * (size_t)c, 8);
          if (bin != 37450u && beg < best) best = beg;  // 37450 = the metadata pseudo-bin
        }
        q += 16 * (size_t)n_chunk;
      }
      if (bad || q + 4 > d.size()) { bad = true; break; }
      const int32_t n_intv = rd32s(d.data() + q);
      q += 4;
      if (n_intv < 0 || q + 8 * (size_t)n_intv > d.size()) { bad = true; break; }
      q += 8 * (size_t)n_intv;
      if (r == tid) {
        *has = best != ~0ull;
        *voff = best;
        return true;
      }
    }
  }
  return false;
}

int bam_load(const char* path, int n_threads, const char* ref_name, vsv_bam** out) {
  const char* fn = ref_name ? "vsv_bam_open_ref" : "vsv_bam_open";
  if (!path || !out) { g_bam_err = std::string(fn) + ": null argument"; return VSV_E_ARG; }
  *out = nullptr;
  FILE* f = fopen(path, "rb");
  if (!f) { g_bam_err = std::string(fn) + ": cannot open " + path; return VSV_E_ARG; }
  vsv_bam* b = new vsv_bam();
  b->n_threads = n_threads > 0 ? n_threads : (int)std::max(1u, std::thread::hardware_concurrency());
  BgzfReader rd(f, b->n_threads);
  Bytes buf;
  size_t pos = 0;
  std::string err;
  size_t WINDOW = size_t(32) << 20;                       // compressed bytes per read; VSV_BAM_WINDOW overrides (tests)
  if (const char* e = getenv("VSV_BAM_WINDOW")) { const long long w = atoll(e); if (w >= 64) WINDOW = (size_t)w; }
  int rc_fill = 1;
  auto fail = [&](const std::string& m) {
    g_bam_err = std::string(fn) + ": " + m;
    fclose(f);
    delete b;
    return VSV_E_ARG;
  };
  // makes `n` bytes available at buf[pos..]; false at end of file (or error: err set)
  auto need = [&](size_t n) -> bool {
    while (buf.size() - pos < n) {
      if (pos > (size_t(1) << 20)) { buf.erase(buf.begin(), buf.begin() + (ptrdiff_t)pos); pos = 0; }
      rc_fill = rd.fill(buf, WINDOW, err);
      if (rc_fill <= 0) return false;
    }
    return true;
  };
  // ---- BAM header -------------------------------------------------------------------------
  if (!need(12)) return fail(err.empty() ? "not a BAM stream (CRAM/SAM are not read here)" : err);
  if (memcmp(buf.data() + pos, "BAM\1", 4) != 0) return fail("not a BAM stream (CRAM/SAM are not read here)");
  const uint32_t l_text = rd32(buf.data() + pos + 4);
  pos += 8;
  if (l_text > (1u << 30)) return fail("implausible header text length");
  if (!need((size_t)l_text + 4)) return fail("truncated header");
  b->header_text.assign(reinterpret_cast<const char*>(buf.data() + pos), l_text);
  pos += l_text;
  const uint32_t n_ref = rd32(buf.data() + pos);
  pos += 4;
  if (n_ref > (1u << 24)) return fail("implausible reference count");   // each reference needs >= 9 bytes of stream anyway
  b->refs.resize(n_ref);
  for (uint32_t i = 0; i < n_ref; ++i) {
    if (!need(4)) return fail("truncated reference list");
    const uint32_t l_name = rd32(buf.data() + pos);
    pos += 4;
    if (l_name == 0 || l_name > (1u << 20) || !need((size_t)l_name + 4)) return fail("truncated reference list");
    b->refs[i].name.assign(reinterpret_cast<const char*>(buf.data() + pos), l_name - 1);
    pos += l_name;
    b->refs[i].len = rd32(buf.data() + pos);
    pos += 4;
  }
  // ---- which records to keep ------------------------------------------------------------------
  int tid = -1;
  bool stop_after = false;  // coordinate-sorted: nothing of `tid` follows a later reference
  if (ref_name && !*ref_name) {  // "" = header only
    fclose(f);
    *out = b;
    return VSV_OK;
  }
  if (ref_name) {
    for (uint32_t i = 0; i < n_ref; ++i)
      if (b->refs[i].name == ref_name) tid = (int)i;
    if (tid < 0) return fail(std::string("reference ") + ref_name + " is not in the BAM header");
    stop_after = b->header_text.find("SO:coordinate") != std::string::npos;
    uint64_t voff = 0;
    bool has = false;
    if (bai_ref_start(path, tid, (int)n_ref, &voff, &has)) {
      if (!has) {  // indexed and empty
        fclose(f);
        *out = b;
        return VSV_OK;
      }
      // trust the index only if the record it points at belongs to `tid` (a stale or foreign .bai would make the
      // feeder seek mid-stream and return part of a chromosome); otherwise read the file from its start
      fseek(f, 0, SEEK_END);
      const long fsz = ftell(f);
      bool ok = fsz > 0 && (voff >> 16) < (uint64_t)fsz;
      if (ok) {
        rd.seek(voff >> 16);
        buf.clear();
        pos = 0;
        ok = need((size_t)(voff & 0xFFFFu) + 40);
        if (ok) {
          const uint8_t* r = buf.data() + (voff & 0xFFFFu);
          ok = rd32(r) >= 32 && rd32s(r + 4) == tid;
        }
        if (ok) {
          pos = (size_t)(voff & 0xFFFFu);
          stop_after = true;
        } else {   // back to the linear scan after the header
          err.clear();
          rd.seek(0);
          buf.clear();
          pos = 0;
          // re-read the header bytes to reposition behind them
          size_t skip = 12 + (size_t)l_text;
          if (!need(skip)) return fail("truncated header");
          for (uint32_t i = 0; i < n_ref; ++i) {
            if (!need(skip + 4)) return fail("truncated reference list");
            skip += 8 + (size_t)rd32(buf.data() + skip);
          }
          if (!need(skip)) return fail("truncated reference list");
          pos = skip;
        }
      }
    }
  }
  // ---- records: per window, a serial walk of the block_size chain, then sizes and copies in parallel ----
  struct Keep { size_t off; uint32_t bs, head, aux; int32_t rid; };
  std::vector<Keep> keep;
  bool stop = false;
  while (!stop) {
    if (!need(4)) {
      if (rc_fill < 0) return fail(err);
      if (buf.size() != pos) return fail("truncated record");
      break;
    }
    {
      const uint32_t bs0 = rd32(buf.data() + pos);
      if (bs0 < 32) return fail("truncated record");
      if (!need(4 + (size_t)bs0)) return fail(rc_fill < 0 ? err : "truncated record");
    }
    keep.clear();
    size_t p = pos;
    while (p + 4 <= buf.size()) {
      const uint32_t bs = rd32(buf.data() + p);
      if (bs < 32) return fail("truncated record");
      if (p + 4 + (size_t)bs > buf.size()) break;
      const int32_t rid = rd32s(buf.data() + p + 4);
      if (tid >= 0 && stop_after && (rid > tid || rid < 0)) { stop = true; break; }  // sorted: later references, then the unmapped tail
      if (rid >= 0 && (uint32_t)rid < n_ref && (tid < 0 || rid == tid)) keep.push_back({p + 4, bs, 0u, 0u, rid});
      p += 4 + (size_t)bs;
    }
    std::atomic<int> bad(0);
    parallel_for((int64_t)keep.size(), b->n_threads, [&](int64_t a0, int64_t e0, int) {
      for (int64_t i = a0; i < e0; ++i) {
        RecView v;
        const uint8_t* body = buf.data() + keep[i].off;
        if (!view(body, keep[i].bs, &v, true)) { bad = 1; return; }
        keep[i].head = (uint32_t)((size_t)(v.aux - body) - (size_t)v.l_seq);   // fixed part, name, cigar, 4-bit seq
        keep[i].aux = (uint32_t)(v.end - v.aux);
      }
    });
    if (bad) return fail("malformed record");
    std::vector<size_t> dst(keep.size() + 1);
    dst[0] = b->data.size();
    for (size_t i = 0; i < keep.size(); ++i) dst[i + 1] = dst[i] + keep[i].head + keep[i].aux;
    b->data.resize(dst[keep.size()]);
    parallel_for((int64_t)keep.size(), b->n_threads, [&](int64_t a0, int64_t e0, int) {
      for (int64_t i = a0; i < e0; ++i) {
        const uint8_t* body = buf.data() + keep[i].off;
        memcpy(b->data.data() + dst[i], body, keep[i].head);
        memcpy(b->data.data() + dst[i] + keep[i].head, body + keep[i].bs - keep[i].aux, keep[i].aux);
      }
    });
    for (size_t i = 0; i < keep.size(); ++i) {
      b->refs[keep[i].rid].recs.push_back(b->recs.size());
      b->recs.push_back({dst[i], keep[i].head + keep[i].aux});
    }
    pos = p;
  }
  fclose(f);
  if (getenv("VSV_BAM_DEBUG")) fprintf(stderr, "vsv_bam: inflate %.3f s (threads %d), kept %zu bytes\n", rd.inflate_s, b->n_threads, b->data.size());
  *out = b;
  return VSV_OK;
}

}  // namespace

// bad_alloc / length_error from a crafted size field must not cross the C ABI
static int bam_load_guarded(const char* path, int n_threads, const char* ref_name, vsv_bam** out) {
  try {
    return bam_load(path, n_threads, ref_name, out);
  } catch (const std::bad_alloc&) {
    g_bam_err = "vsv_bam_open: out of memory (or a corrupt size field)";
    if (out) *out = nullptr;
    return VSV_E_NOMEM;
  } catch (const std::exception& e) {
    g_bam_err = std::string("vsv_bam_open: ") + e.what();
    if (out) *out = nullptr;
    return VSV_E_ARG;
  }
}

extern "C" int vsv_bam_open(const char* path, int n_threads, vsv_bam** out) { return bam_load_guarded(path, n_threads, nullptr, out); }

extern "C" int vsv_bam_open_ref(const char* path, const char* ref_name, int n_threads, vsv_bam** out) {
  if (!ref_name) { g_bam_err = "vsv_bam_open_ref: null argument"; return VSV_E_ARG; }
  return bam_load_guarded(path, n_threads, ref_name, out);
}

extern "C" void vsv_bam_close(vsv_bam* b) { delete b; }

extern "C" int vsv_bam_n_ref(const vsv_bam* b) { return b ? (int)b->refs.size() : 0; }

extern "C" int vsv_bam_ref(const vsv_bam* b, int i, const char** name, int64_t* len, int64_t* n_rec) {
  if (!b || i < 0 || i >= (int)b->refs.size()) { g_bam_err = "vsv_bam_ref: bad index"; return VSV_E_ARG; }
  if (name) *name = b->refs[i].name.c_str();
  if (len) *len = b->refs[i].len;
  if (n_rec) *n_rec = (int64_t)b->refs[i].recs.size();
  return VSV_OK;
}

// sizes[0..4] = n_rec, n_ops, name bytes, SEQ bytes, SA bytes of the records mapped to reference `tid`
extern "C" int vsv_bam_chrom_sizes(vsv_bam* b, int tid, int64_t* sizes) {
  if (!b || !sizes || tid < 0 || tid >= (int)b->refs.size()) { g_bam_err = "vsv_bam_chrom_sizes: bad argument"; return VSV_E_ARG; }
  const std::vector<uint64_t>& rr = b->refs[tid].recs;
  const int T = b->n_threads;
  std::vector<int64_t> part(5 * (size_t)std::max(1, T), 0);
  std::atomic<int> bad(0);
  parallel_for((int64_t)rr.size(), T, [&](int64_t a, int64_t e, int t) {
    int64_t* s = part.data() + 5 * t;
    for (int64_t i = a; i < e; ++i) {
      RecView v;
      const BamRec& r = b->recs[rr[i]];
      if (!view(b->data.data() + r.off, r.size, &v)) { bad = 1; return; }
      size_t sl;
      find_sa(v, &sl);
      s[1] += v.n_cigar;
      s[2] += v.l_name ? v.l_name - 1 : 0;
      s[3] += v.l_seq;
      s[4] += (int64_t)sl;
    }
  });
  if (bad) { g_bam_err = "vsv_bam_chrom_sizes: malformed record"; return VSV_E_ARG; }
  for (int k = 0; k < 5; ++k) sizes[k] = 0;
  sizes[0] = (int64_t)rr.size();
  for (int t = 0; t < std::max(1, T); ++t)
    for (int k = 1; k < 5; ++k) sizes[k] += part[5 * t + k];
  return VSV_OK;
}

// Fills the packed arrays of one chromosome (file order = coordinate order of a sorted BAM).
// Any output pointer may be NULL to skip that column; `ops` (u32) and `ops16` (+ exception list)
// are alternative encodings of the same CIGAR ops and may both be requested.
extern "C" int vsv_bam_chrom_fill(vsv_bam* b, int tid, int32_t* pos, uint8_t* mapq, uint16_t* flag, int32_t* qlen,
                                  int64_t* op_off, uint32_t* ops, uint16_t* ops16, uint8_t* ops8, int64_t* exc_idx,
                                  uint32_t* exc_val, int64_t exc_cap, int64_t* n_exc, int64_t* name_off, char* names,
                                  int64_t* seq_off, char* seq, int64_t* sa_off, char* sa) {
  if (!b || tid < 0 || tid >= (int)b->refs.size()) { g_bam_err = "vsv_bam_chrom_fill: bad argument"; return VSV_E_ARG; }
  if ((ops16 || ops8) && (!exc_idx || !exc_val || !n_exc)) { g_bam_err = "vsv_bam_chrom_fill: ops16 / ops8 need the exception list"; return VSV_E_ARG; }
  if (ops16 && ops8) { g_bam_err = "vsv_bam_chrom_fill: ask for ops16 or ops8, not both"; return VSV_E_ARG; }
  const std::vector<uint64_t>& rr = b->refs[tid].recs;
  const int64_t n = (int64_t)rr.size();
  // pass 1 (serial, pointer hops only): offsets of the variable-length columns
  std::vector<int64_t> o_ops((size_t)n + 1), o_name((size_t)n + 1), o_seq((size_t)n + 1), o_sa((size_t)n + 1);
  {
    std::atomic<int> bad(0);
    parallel_for(n, b->n_threads, [&](int64_t a, int64_t e, int) {
      for (int64_t i = a; i < e; ++i) {
        RecView v;
        const BamRec& r = b->recs[rr[i]];
        if (!view(b->data.data() + r.off, r.size, &v)) { bad = 1; return; }
        size_t sl;
        find_sa(v, &sl);
        o_ops[i + 1] = v.n_cigar;
        o_name[i + 1] = v.l_name ? v.l_name - 1 : 0;
        o_seq[i + 1] = v.l_seq;
        o_sa[i + 1] = (int64_t)sl;
      }
    });
    if (bad) { g_bam_err = "vsv_bam_chrom_fill: malformed record"; return VSV_E_ARG; }
    o_ops[0] = o_name[0] = o_seq[0] = o_sa[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
      o_ops[i + 1] += o_ops[i]; o_name[i + 1] += o_name[i]; o_seq[i + 1] += o_seq[i]; o_sa[i + 1] += o_sa[i];
    }
  }
  if (op_off) memcpy(op_off, o_ops.data(), sizeof(int64_t) * ((size_t)n + 1));
  if (name_off) memcpy(name_off, o_name.data(), sizeof(int64_t) * ((size_t)n + 1));
  if (seq_off) memcpy(seq_off, o_seq.data(), sizeof(int64_t) * ((size_t)n + 1));
  if (sa_off) memcpy(sa_off, o_sa.data(), sizeof(int64_t) * ((size_t)n + 1));
  std::atomic<long long> exc_n(0);
  static const char NT16[] = "=ACMGRSVTWYHKDBN";
  parallel_for(n, b->n_threads, [&](int64_t a, int64_t e, int) {
    for (int64_t i = a; i < e; ++i) {
      RecView v;
      const BamRec& r = b->recs[rr[i]];
      view(b->data.data() + r.off, r.size, &v);
      if (pos) pos[i] = v.pos;
      if (mapq) mapq[i] = (uint8_t)v.mapq;
      if (flag) flag[i] = (uint16_t)v.flag;
      if (qlen) qlen[i] = v.l_seq;
      const int64_t o = o_ops[i];
      if (ops) memcpy(ops + o, v.cigar, 4 * (size_t)v.n_cigar);
      if (ops16) {
        for (uint32_t k = 0; k < v.n_cigar; ++k) {
          uint32_t w;
          memcpy(&w, v.cigar + 4 * (size_t)k, 4);
          if (w < 0xFFF0u) {
            ops16[o + k] = (uint16_t)w;
          } else {  // len >= 4095: marker 0xFFF<<4|op + (index, full word) in the exception list (vsv_expand_ops16)
            ops16[o + k] = (uint16_t)(0xFFF0u | (w & 15u));
            const long long s = exc_n.fetch_add(1);
            if (s < exc_cap) { exc_idx[s] = o + k; exc_val[s] = w; }
          }
        }
      }
      if (ops8) {
        for (uint32_t k = 0; k < v.n_cigar; ++k) {
          uint32_t w;
          memcpy(&w, v.cigar + 4 * (size_t)k, 4);
          const uint32_t code = w & 15u, len = w >> 4;
          static const uint32_t cnt[3] = {VSV_OPS8_M, VSV_OPS8_I, VSV_OPS8_D};
          static const uint32_t base[3] = {0, VSV_OPS8_M, VSV_OPS8_M + VSV_OPS8_I};
          if (code <= 2u && len >= 1u && len <= cnt[code]) {
            ops8[o + k] = (uint8_t)(base[code] + len - 1u);
          } else {  // not a short M/I/D run: marker + (index, full word) in the exception list (vsv_expand_ops8)
            ops8[o + k] = 0xFFu;
            const long long s = exc_n.fetch_add(1);
            if (s < exc_cap) { exc_idx[s] = o + k; exc_val[s] = w; }
          }
        }
      }
      if (names) memcpy(names + o_name[i], v.name, (size_t)(o_name[i + 1] - o_name[i]));
      if (seq) {
        char* dst = seq + o_seq[i];
        for (int32_t k = 0; k < v.l_seq; ++k) dst[k] = NT16[(v.seq[k >> 1] >> ((~k & 1) << 2)) & 15];
      }
      if (sa) {
        size_t sl;
        const char* s = find_sa(v, &sl);
        if (sl) memcpy(sa + o_sa[i], s, sl);
      }
    }
  });
  if (n_exc) *n_exc = exc_n.load();
  if ((ops16 || ops8) && exc_n.load() > exc_cap) {
    g_bam_err = "vsv_bam_chrom_fill: exception list too small";
    return VSV_E_CAPACITY;
  }
  if ((ops16 || ops8) && exc_n.load() > 1) {  // threads appended out of order: vsv_expand_ops16 wants ascending indices
    const long long m = exc_n.load();
    std::vector<std::pair<int64_t, uint32_t>> tmp((size_t)m);
    for (long long k = 0; k < m; ++k) tmp[k] = {exc_idx[k], exc_val[k]};
    std::sort(tmp.begin(), tmp.end());
    for (long long k = 0; k < m; ++k) { exc_idx[k] = tmp[k].first; exc_val[k] = tmp[k].second; }
  }
  return VSV_OK;
}

// Host packer of the ops12 staging format (vsv_expand_ops12): 12 bits per {M 1..128, I/D 1..8} pair in either
// order, both words of every other pair in the exception list.  Threads take contiguous ranges of 8-token
// groups (12 output bytes each, so no two threads share a byte), collect their exceptions locally and copy
// them out in range order: the list is ascending and the result does not depend on the thread count.
extern "C" int vsv_pack_ops12(const uint32_t* ops, int64_t n_ops, int n_threads, uint8_t* ops12, int64_t* exc_pair,
                              uint32_t* exc_val, int64_t exc_cap, int64_t* n_exc) {
  if (n_ops < 0 || exc_cap < 0 || !n_exc || (n_ops > 0 && (!ops || !ops12)) || (exc_cap > 0 && (!exc_pair || !exc_val))) {
    g_bam_err = "vsv_pack_ops12: bad argument";
    return VSV_E_ARG;
  }
  *n_exc = 0;
  if (n_ops == 0) return VSV_OK;
  const int64_t n_pair = (n_ops + 1) / 2, n_grp = (n_pair + 7) / 8;
  const int T = (int)std::max<int64_t>(1, std::min<int64_t>(n_threads > 0 ? n_threads : (int)std::max(1u, std::thread::hardware_concurrency()),
                                                            (n_grp + 4095) / 4096));
  std::vector<std::vector<int64_t>> found((size_t)T);
  parallel_for(n_grp, T, [&](int64_t a, int64_t e, int t) {
    std::vector<int64_t>& mine = found[(size_t)t];
    for (int64_t g = a; g < e; ++g) {
      uint32_t tok[8];
      for (int k = 0; k < 8; ++k) {
        const int64_t p = g * 8 + k;
        tok[k] = 0;
        if (p >= n_pair) continue;
        const uint32_t w0 = ops[2 * p], w1 = 2 * p + 1 < n_ops ? ops[2 * p + 1] : 0u;
        const bool indel_first = (w0 & 15u) != 0u;
        const uint32_t m = indel_first ? w1 : w0, x = indel_first ? w0 : w1;
        const uint32_t lm = m >> 4, lx = x >> 4, cx = x & 15u;
        if ((m & 15u) == 0u && lm >= 1u && lm <= (uint32_t)VSV_OPS12_M && (cx == 1u || cx == 2u) && lx >= 1u &&
            lx <= (uint32_t)VSV_OPS12_X)
          tok[k] = (lm - 1u) | ((lx - 1u) << 7) | ((cx == 2u ? 1u : 0u) << 10) | ((indel_first ? 1u : 0u) << 11);
        else
          mine.push_back(p);
      }
      uint8_t* o = ops12 + g * 12;
      for (int k = 0; k < 4; ++k) {
        const uint32_t t0 = tok[2 * k], t1 = tok[2 * k + 1];
        o[3 * k] = (uint8_t)(t0 & 0xFFu);
        o[3 * k + 1] = (uint8_t)((t0 >> 8) | ((t1 & 15u) << 4));
        o[3 * k + 2] = (uint8_t)(t1 >> 4);
      }
    }
  });
  std::vector<int64_t> start((size_t)T + 1, 0);
  for (int t = 0; t < T; ++t) start[(size_t)t + 1] = start[(size_t)t] + (int64_t)found[(size_t)t].size();
  *n_exc = start[(size_t)T];
  if (*n_exc > exc_cap) {
    g_bam_err = "vsv_pack_ops12: exception list too small";
    return VSV_E_CAPACITY;
  }
  parallel_for(T, T, [&](int64_t a, int64_t e, int) {
    for (int64_t t = a; t < e; ++t) {
      int64_t o = start[(size_t)t];
      for (int64_t p : found[(size_t)t]) {
        exc_pair[o] = p;
        exc_val[2 * o] = ops[2 * p];
        exc_val[2 * o + 1] = 2 * p + 1 < n_ops ? ops[2 * p + 1] : 0u;
        ++o;
      }
    }
  });
  return VSV_OK;
}

// SEQ of record i (file order) of reference `tid`, decoded to ASCII: the sparse accesses of
// `read.seq` / `read.query_sequence` (SE:452, ECS_H:642-700) without materialising every read's bases.
extern "C" int vsv_bam_record_seq(vsv_bam* b, int tid, int64_t i, char* out, int64_t cap, int64_t* len) {
  if (!b || !len || tid < 0 || tid >= (int)b->refs.size() || i < 0 || i >= (int64_t)b->refs[tid].recs.size()) {
    g_bam_err = "vsv_bam_record_seq: bad argument";
    return VSV_E_ARG;
  }
  RecView v;
  const BamRec& r = b->recs[b->refs[tid].recs[i]];
  if (!view(b->data.data() + r.off, r.size, &v)) { g_bam_err = "vsv_bam_record_seq: malformed record"; return VSV_E_ARG; }
  *len = v.l_seq;
  if (!out || cap < v.l_seq) return v.l_seq ? VSV_E_CAPACITY : VSV_OK;
  static const char NT16[] = "=ACMGRSVTWYHKDBN";
  for (int32_t k = 0; k < v.l_seq; ++k) out[k] = NT16[(v.seq[k >> 1] >> ((~k & 1) << 2)) & 15];
  return VSV_OK;
}
