/* volcanosv_b200 -- C-ABI of the B200-native large-indel hot path of VolcanoSV-vc.
 *
 * The reference (maiziezhoulab/VolcanoSV) is pure Python: it has no FFI for this path.
 * Each entry point below replaces the body of one reference FUNCTION; the Python
 * operators of volcanosv_b200/ keep the reference scripts' argv and file contracts
 * and bind these symbols through ctypes (INTEGRATION.md shows the stub a maintainer
 * of the reference would add).  Paths: LI/ = bin/VolcanoSV-vc/Large_INDEL/.
 *
 * Conventions
 *  - every function returns 0 on success, <0 on error (VSV_E_*); vsv_last_error(ctx)
 *    returns the message.  No exceptions cross the ABI.
 *  - one vsv_ctx per (thread, GPU); a ctx is not thread-safe.
 *  - buffers are caller-owned.  By default they are HOST pointers (pinned memory from
 *    vsv_alloc_pinned makes the copies asynchronous); after
 *    vsv_ctx_set_device_io(ctx,1) every data pointer is a DEVICE pointer on the ctx's
 *    GPU and nothing is copied (inputs stay resident in HBM; used by the pipeline
 *    driver and by bench.py's resident timing).  Scalar outputs (counts) are always
 *    host pointers.
 *  - variable-size outputs use caller-provided capacity; if the result does not fit,
 *    the call returns VSV_E_CAPACITY and *n_out holds the required count.
 *  - there is NO CPU fallback: without a CUDA device vsv_ctx_create fails.
 */
#ifndef VOLCANOSV_B200_H
#define VOLCANOSV_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VSV_OK 0
#define VSV_E_CUDA (-1)
#define VSV_E_ARG (-2)
#define VSV_E_CAPACITY (-3)
#define VSV_E_NOMEM (-4)

/* scan modes (which reference scanner is reproduced) */
#define VSV_MODE_CONTIG 0 /* LI/extract_contig_signature_{Hifi,ONT,CLR}.py:53-85  */
#define VSV_MODE_READS 1  /* LI/extract_reads_signature.py:47-83                  */
#define VSV_MODE_CUTE 2   /* LI/sig_extract.py:438-478                            */

#define VSV_DEL 0
#define VSV_INS 1

/* split-pair rule sets */
#define VSV_SPLIT_HIFI 0  /* LI/extract_contig_signature_Hifi.py:307-371 */
#define VSV_SPLIT_ONT 1   /* LI/extract_contig_signature_ONT.py:307-382  */
#define VSV_SPLIT_CLR 2   /* LI/extract_contig_signature_CLR.py:328-402  */
#define VSV_SPLIT_READS 3 /* LI/extract_reads_signature.py:147-197       */

/* One signature, 32 bytes.  Mirrors the reference's 10-field contig sig / 8-field
 * read sig (ECS_H:80,84; ERS:74,78): chrom/qname/strand/mapq are looked up by the
 * host from `rec`. */
typedef struct vsv_sig {
  int32_t rec;   /* record index in the scanned batch (split pairs: first record)  */
  int32_t type;  /* VSV_DEL / VSV_INS                                               */
  int32_t pos;   /* reference position                                              */
  int32_t len;   /* SV length                                                       */
  int32_t q0;    /* query start (contig: offset_contig+hardclip; cute: index in SEQ) */
  int32_t q1;    /* query end                                                       */
  int32_t first; /* merge: index of first pre-merge sig of this type folded in;
                    split pairs: second record index                               */
  int32_t count; /* merge: number of pre-merge sigs folded in (>=1)                 */
} vsv_sig;

typedef struct vsv_ctx vsv_ctx;

/* ---- context ---------------------------------------------------------------- */
int vsv_ctx_create(int device, vsv_ctx** out);
void vsv_ctx_destroy(vsv_ctx* ctx);
const char* vsv_last_error(vsv_ctx* ctx);
int vsv_ctx_set_device_io(vsv_ctx* ctx, int on);
int vsv_sync(vsv_ctx* ctx);
int vsv_alloc_pinned(void** p, size_t bytes);
int vsv_free_pinned(void* p);
/* CUDA-event time (ms) of the kernels of the last call on this ctx, and number of
 * kernel launches since the ctx was created. */
int vsv_get_timing(vsv_ctx* ctx, float* kernel_ms, int64_t* launches);
const char* vsv_version(void);
/* CUDA-event stopwatch on the ctx stream (bench.py brackets K steps with it), and the
 * event time of the dominant kernel (k_scan_cigar) of the last vsv_scan_cigar call. */
int vsv_timer_start(vsv_ctx* ctx);
int vsv_timer_stop(vsv_ctx* ctx, float* ms);
int vsv_get_main_kernel_ms(vsv_ctx* ctx, float* ms);

/* ---- stage 1: signature scan -------------------------------------------------- */
/* Replaces extract_sig_from_cigar (ECS_H:53-85, ERS:47-83) and the CIGAR walk of
 * parse_read (SE:438-478).  ops[i] = len<<4|op (BAM-native), op_off[n_rec+1].
 * Records with mapq<min_mapq or enable[r]==0 (enable may be NULL) emit nothing.
 * out: signatures in (record, op) order.  ref/qry/aux_total (each may be NULL):
 * per-record sums of op lengths over the mode's reference / query / auxiliary op
 * sets (aux = H for CONTIG/READS so readlen = qry+aux; aux = N for CUTE so
 * reference_end = pos+ref+aux). */
int vsv_scan_cigar(vsv_ctx* ctx, int mode, int min_len, int min_mapq, const uint32_t* ops,
                   const int64_t* op_off, const int32_t* pos, const uint8_t* mapq,
                   const uint8_t* enable, int64_t n_rec, vsv_sig* out, int64_t cap,
                   int64_t* n_out, int32_t* ref_total, int32_t* qry_total, int32_t* aux_total);

/* Compact host->device staging of the packed ops.  ops16[i] = len << 4 | op when len < 4095,
 * else 0xFFF << 4 | op with the full BAM-native word in the sparse exception list
 * (exc_idx ascending op indices, exc_val).  Writes the u32 array vsv_scan_* read (ops_out,
 * 16-byte aligned).  PCIe bytes per op drop from 4 to 2; the scanners are unchanged.  The
 * host side of the format is volcanosv_b200.lib.pack_ops16. */
int vsv_expand_ops16(vsv_ctx* ctx, const uint16_t* ops16, int64_t n_ops, const int64_t* exc_idx,
                     const uint32_t* exc_val, int64_t n_exc, uint32_t* ops_out);
/* Denser staging for noisy long reads (ONT/CLR: ~99.9 % of the ops are short M/I/D runs):
 * one byte per op, the 254 codes split by op type: 0..159 = M of 1..160 bases, 160..206 = I of
 * 1..47, 207..253 = D of 1..47; 0xFF = the full word is in the exception list (S/H/N/P/=/X,
 * longer or empty ops).  1 byte per op over PCIe (+ 12 bytes per exception).  Host side:
 * volcanosv_b200.lib.pack_ops8, or vsv_bam_chrom_fill straight from the BAM records. */
#define VSV_OPS8_M 160
#define VSV_OPS8_I 47
#define VSV_OPS8_D 47
int vsv_expand_ops8(vsv_ctx* ctx, const uint8_t* ops8, int64_t n_ops, const int64_t* exc_idx,
                    const uint32_t* exc_val, int64_t n_exc, uint32_t* ops_out);
/* Densest staging for noisy long reads: 12 bits per PAIR of ops (0.75 byte per op).  The CIGARs minimap2
 * writes for ONT/CLR reads alternate a match run with a short indel, so (ops[2k], ops[2k+1]) is nearly
 * always {M of 1..128, I or D of 1..8} in either order: token = (mlen-1) | (xlen-1) << 7 | is_D << 10 |
 * indel_first << 11; token k occupies bits [12k, 12k+12) of the little-endian byte stream `ops12`, which is
 * padded to a whole group of 8 tokens: vsv_ops12_bytes(n_ops) bytes, 4-byte aligned.  Every other pair has
 * both full words in the exception list (exc_pair[i] = k ascending, exc_val[2i], exc_val[2i+1]).  Pairing
 * is positional over the whole array (it ignores record boundaries).  Host side: volcanosv_b200.lib.pack_ops12. */
#define VSV_OPS12_M 128
#define VSV_OPS12_X 8
int64_t vsv_ops12_bytes(int64_t n_ops);
int vsv_expand_ops12(vsv_ctx* ctx, const uint8_t* ops12, int64_t n_ops, const int64_t* exc_pair,
                     const uint32_t* exc_val, int64_t n_exc, uint32_t* ops_out);
/* Host packer of that format (no CUDA; threads over groups of 8 tokens, n_threads <= 0 = all cores): fills
 * ops12 (vsv_ops12_bytes(n_ops) bytes) and the ascending exception list; *n_exc = pairs listed.  VSV_E_CAPACITY
 * with *n_exc = the capacity needed when exc_cap is too small (ops12 is complete then).  Errors:
 * vsv_bam_last_error().  Same bytes as volcanosv_b200.lib.pack_ops12 (the numpy statement of the format). */
int vsv_pack_ops12(const uint32_t* ops, int64_t n_ops, int n_threads, uint8_t* ops12, int64_t* exc_pair,
                   uint32_t* exc_val, int64_t exc_cap, int64_t* n_exc);

/* The two READ-BAM scanners in ONE pass over the packed ops (the ops are the bulk of the
 * bytes, so this halves HBM traffic of stage 1): stream "reads" == vsv_scan_cigar(
 * VSV_MODE_READS, min_len_reads, min_mapq_reads, enable=NULL) (ERS:47-83,107-145) and
 * stream "cute" == vsv_scan_cigar(VSV_MODE_CUTE, min_len_cute, min_mapq_cute,
 * enable_cute) (SE:438-478), bit for bit.  Totals follow VSV_MODE_READS. */
int vsv_scan_reads_fused(vsv_ctx* ctx, int min_len_reads, int min_mapq_reads, int min_len_cute,
                         int min_mapq_cute, const uint32_t* ops, const int64_t* op_off,
                         const int32_t* pos, const uint8_t* mapq, const uint8_t* enable_cute,
                         int64_t n_rec, vsv_sig* out_reads, int64_t cap_reads, int64_t* n_reads,
                         vsv_sig* out_cute, int64_t cap_cute, int64_t* n_cute, int32_t* ref_total,
                         int32_t* qry_total, int32_t* aux_total);

/* Replaces cluster_ins_one_read / cluster_del_one_read (ECS_H:87-164) for
 * VSV_MODE_CONTIG and generate_combine_sigs (SE:373-435) for VSV_MODE_CUTE
 * (VSV_MODE_READS: identity split by type).  Input: scan output.  out_del/out_ins
 * need capacity n each. */
int vsv_merge_in_record(vsv_ctx* ctx, int mode, const vsv_sig* sigs, int64_t n, vsv_sig* out_del,
                        int64_t* n_del, vsv_sig* out_ins, int64_t* n_ins);

/* Replaces extract_sig_from_split + its driver loop (ECS_H:307-371,421-457;
 * ERS:147-237).  select[r] = the reference's membership filter; name_id = interned
 * qname in first-appearance order.  Output order = (name first appearance, pair). */
int vsv_split_pairs(vsv_ctx* ctx, int variant, int min_mapq, const int32_t* pos,
                    const int32_t* ref_end, const int32_t* readlen, const uint32_t* first_op,
                    const uint32_t* last_op, const uint8_t* mapq, const uint8_t* is_rev,
                    const int32_t* name_id, const uint8_t* select, int64_t n_rec, vsv_sig* out,
                    int64_t cap, int64_t* n_out);

/* analysis_split_read (SE:193-326) on the segment lists organize_split_signal (SE:345-371)
 * builds (primary first, then the SA entries that pass the mapq filter; SA positions are
 * kept 1-based, SE:354).  Read r owns segments [seg_off[r], seg_off[r+1]) given as columns
 * q_s,q_e,r_s,r_e,chrom,strand(0 '+',1 '-'); read_len[r] = query length incl. clips.  Reads
 * with more than max_parts segments (reference: 7, SE:370) emit nothing; max_parts must be
 * in [2,16].  Output rows in (read, emission) order: rec = read index, type, pos (already
 * truncated as '%d' prints the reference's float, SE:562-565), len; for INS q0,q1 = slice of
 * the query sequence (Python slice semantics) after reversing it `first` times (0/1).  TRA
 * rows are not produced (SE:637-638 greps them away). */
int vsv_sa_analysis(vsv_ctx* ctx, int sv_size, int max_size, int max_parts, const int64_t* seg_off,
                    const int32_t* q_s, const int32_t* q_e, const int32_t* r_s, const int32_t* r_e,
                    const int32_t* chrom, const uint8_t* strand, const int32_t* read_len,
                    int64_t n_read, int64_t n_seg, vsv_sig* out, int64_t cap, int64_t* n_out);

/* ---- stage 2: sort / cluster / pair / window joins ---------------------------- */
/* sort_sig (ECS_H:170-179) with the tie order pinned to stable: order[i] = index of
 * the i-th smallest key. */
int vsv_argsort_stable(vsv_ctx* ctx, const int32_t* key, int64_t n, int32_t* order);

/* Sort a signature array.  tie_mode 0: stable by pos (sort_sig, ECS_H:170-179);
 * tie_mode 1: the `.sigs` file order of SE:637-638 (`sort -k2,2 -k3,3n`, C locale): pos
 * numeric, ties by len compared as a decimal STRING. */
int vsv_sort_sigs(vsv_ctx* ctx, const vsv_sig* sigs, int64_t n, int tie_mode, vsv_sig* out);
/* Struct-of-arrays view of a signature array (any output may be NULL). */
int vsv_sig_columns(vsv_ctx* ctx, const vsv_sig* sigs, int64_t n, int32_t* pos, int32_t* len,
                    int32_t* rec);

/* Elementwise glue so a device-resident pipeline never leaves the ctx stream:
 * out = a + b (reference_end = pos + ref_total); out = (a >= thr) (SE:439 read-length
 * gate); out = key == val ? a : b (DEL/INS support select). */
int vsv_add_i32(vsv_ctx* ctx, const int32_t* a, const int32_t* b, int64_t n, int32_t* out);
int vsv_ge_u8(vsv_ctx* ctx, const int32_t* a, int32_t thr, int64_t n, uint8_t* out);
int vsv_select_i32(vsv_ctx* ctx, const int32_t* key, int32_t val, const int32_t* a, const int32_t* b,
                   int64_t n, int32_t* out);

/* cluster_del / cluster_ins (ECS_H:196-288) on a pos-sorted list: kept indices in
 * leader order. */
int vsv_cluster(vsv_ctx* ctx, int is_del, const int32_t* pos, const int32_t* len, int64_t n,
                int32_t* keep, int64_t* n_keep);

/* pair_sig matching part (ECS_H:548-569): partner index or -1 for each side. */
int vsv_pair_haps(vsv_ctx* ctx, const int32_t* pos1, const int32_t* len1, const int32_t* typ1,
                  int64_t n1, const int32_t* pos2, const int32_t* len2, const int32_t* typ2,
                  int64_t n2, int32_t* partner1, int32_t* partner2);

/* eval_sig (FPF:106-123): read-signature support per raw call (60 if len>250). */
int vsv_fp_support(vsv_ctx* ctx, const int32_t* call_pos, const int32_t* call_len, int64_t n_call,
                   const int32_t* sig_pos, const int32_t* sig_len, int64_t n_sig, int32_t* support);

/* calc_ins_call_cov (CSS:81-125) / calc_del_call_cov (CSS:138-253), flanking 1000.
 * cov[i] = sum of sig lengths (DEL: negated) in the call's window. */
int vsv_ins_cov(vsv_ctx* ctx, const int32_t* call_pos, int64_t n_call, const int32_t* sig_pos,
                const int32_t* sig_len, int64_t n_sig, int flanking, int64_t* cov);
int vsv_del_cov(vsv_ctx* ctx, const int32_t* call_start, const int32_t* call_end, int64_t n_call,
                const int32_t* sig_pos, const int32_t* sig_len, int64_t n_sig, int flanking,
                int64_t* cov);

/* ---- stage 3: genotyping ------------------------------------------------------ */
/* load_sig (GTD:65-86) / load_sig_file (GTI:67-101) on a (chrom,pos)-sorted table:
 * unique (chrom,pos,len) rows with read counts.  Outputs need capacity n.  chrom may be
 * NULL (single chromosome: u_chrom is then filled with 0).  Rows with len < min_len are
 * dropped first (GTI:81 keeps INS signatures >= 30 bp; DEL: pass 0). */
int vsv_group_sigs(vsv_ctx* ctx, const int32_t* chrom, const int32_t* pos, const int32_t* len,
                   int64_t n, int min_len, int32_t* u_chrom, int32_t* u_pos, int32_t* u_len,
                   int32_t* u_cnt, int64_t* n_unique);

/* match_varlist_siglist (GTD:92-137) / extract_sig_support (GTI:107-166), including
 * the double count at last_match_idx.  last_in/last_out carry last_match_idx across
 * shards (0 for the first).  match_id may be NULL.  last_out may be NULL when both chromosome
 * arrays are NULL (one chromosome): the call then returns without waiting for the stream. */
int vsv_sig_support(vsv_ctx* ctx, const int32_t* cand_chrom, const int32_t* cand_pos,
                    const int32_t* cand_svlen, int64_t n_cand, const int32_t* sig_chrom,
                    const int32_t* sig_pos, const int32_t* sig_len, const int32_t* sig_cnt,
                    int64_t n_sig, int32_t last_in, int32_t* support, int32_t* match_id,
                    int32_t* last_out);

/* count_reads_span_region (GTD:140-146; GTI:188-196): per window, number of records
 * with start < S and end > E.  Reads sorted by start. */
int vsv_span_depth(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end,
                   int64_t n_reads, const int32_t* S, const int32_t* E, int64_t n_win,
                   int32_t* depth);

/* check_full_cover_reads + ratio + correct_gt_eval (GTD:149-170,206-217,291-318;
 * GTI:188-196,228,258-283) for candidates on one chromosome.  vtype per call:
 * VSV_DEL / VSV_INS array; call_gt 0='0/1', 1='1/1', other = pass through.
 * thr_del/thr_ins = {t_large_11,t_small_11,t_large_01,t_small_01} (NaN = leave).
 * Outputs: depth2 (= 2*depth, exact), ratio (double), new_gt. */
int vsv_genotype(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, int64_t n_reads,
                 const int32_t* cand_pos, const int32_t* cand_svlen, const int32_t* cand_type,
                 const int32_t* call_gt, const int32_t* n_support, int64_t n_cand,
                 const double* thr_del, const double* thr_ins, int32_t* depth2, double* ratio,
                 int32_t* new_gt);

/* ---- redundancy: batched global edit distance ------------------------------------ */
/* Replaces edlib.align(a, b)["editDistance"] at LI/remove_redundancy.py:75-81 (global /
 * Needleman-Wunsch, unit costs) for a batch of string pairs: strings are rows of a table
 * (chars, str_off[n_str+1]); pair p compares strings ia[p] and ib[p].  Exact (Myers/Hyyro
 * bit-vector blocks inside an Ukkonen band that widens until the result is proven, as edlib's
 * NW mode does).  The string table may use at most 32 distinct byte values (DNA alphabets) and
 * strings shorter than 2^28, else VSV_E_ARG.  Host buffers only. */
int vsv_edit_distance_batch(vsv_ctx* ctx, const uint8_t* chars, const int64_t* str_off, int64_t n_str,
                            const int32_t* ia, const int32_t* ib, int64_t n_pair, int32_t* dist);

/* ---- parameters: every constant of the reference's operators --------- */
/* Defaults = the reference's (argparse defaults, literals in the scripts, LI/para/*.txt, LI/filter_para.csv).
 * A ctx starts with vsv_params_default(dtype Hifi); vsv_ctx_set_params replaces them.  dtype: 0 Hifi, 1 ONT,
 * 2 CLR (the -dtype flag of volcanosv-vc-large-indel.py, MAIN:13). */
#define VSV_DTYPE_HIFI 0
#define VSV_DTYPE_ONT 1
#define VSV_DTYPE_CLR 2
typedef struct vsv_params {
  int32_t dtype;
  /* read scanners */
  int32_t reads_min_len;      /* 30   ERS --min_siglen (ERS:40)                         */
  int32_t reads_min_mapq;     /* 50   ERS:120                                           */
  int32_t cute_min_len;       /* 10   SE -sl / --min_siglength                          */
  int32_t cute_min_mapq;      /* 20   SE -q                                             */
  int32_t cute_min_read_len;  /* 500  SE -r  (SE:439)                                   */
  int32_t sa_sv_size;         /* 30   SE -l  (min_size)                                 */
  int32_t sa_max_size;        /* 100000 SE -L                                           */
  int32_t sa_max_parts;       /* 7    SE -p  (SE:370)                                   */
  int32_t batch_size;         /* 10000000 SE -b: read-start ownership of a task (SE:603-612,535) */
  /* contig side */
  int32_t contig_min_len;     /* 30   ECS min_siglen                                    */
  int32_t contig_min_mapq;    /* 50   ECS_H:392-394                                     */
  int32_t cluster_max_shift;  /* 100  ECS_H:196-288                                     */
  int32_t pair_max_shift;     /* 200  ECS_H:548-569                                     */
  double clr_max_ins_pct;     /* 0.13 ECS_C:425-427                                     */
  double clr_min_var_dist;    /* 200                                                    */
  /* FP filter */
  int32_t fp_max_comp;        /* 250  FPF:110                                           */
  int32_t fp_max_shift;       /* 500  FPF:99                                            */
  /* coverage support / band filter */
  int32_t cov_flanking;       /* 1000 CSS:28                                            */
  int32_t cov_min_size;       /* 30   CSS:29                                            */
  double cov_lb_ins, cov_rb_ins, cov_lb_del, cov_rb_del; /* LI/filter_para.csv row (volcano, dtype) */
  /* genotype */
  int32_t gt_ins_min_len;     /* 30   GTI:81                                            */
  double gt_min_size_sim;     /* 0.6  GTD:103-104                                       */
  double gt_max_shift_ratio;  /* 2.3  GTD:105                                           */
  double gt_min_shift;        /* 500                                                    */
  int32_t gt_ins_flank;       /* 100  GTI:327                                           */
  int32_t gt_del_large;       /* 1000 GTD:150                                           */
  int32_t gt_del_off;         /* 150  GTD:157-168                                       */
  int32_t gt_del_win;         /* 100                                                    */
  double thr_del[4];          /* LI/para/GT_correction_para_<dtype>_DEL.txt: large_11, small_11, large_01, small_01 */
  double thr_ins[4];          /* ..._INS.txt (NaN = leave the genotype)                  */
} vsv_params;
int vsv_params_default(vsv_params* p, int dtype);
int vsv_ctx_set_params(vsv_ctx* ctx, const vsv_params* p);
int vsv_ctx_get_params(vsv_ctx* ctx, vsv_params* p);

/* ECS_C:53-70: per contig record, sum of I lengths, sum of M lengths and number of M ops (the CLR gate
 * ins_pct = I / (M + I) <= 0.13 or var_dist = M / #M >= 200 is then two fp64 divisions on the host). */
int vsv_cigar_class_sums(vsv_ctx* ctx, const uint32_t* ops, const int64_t* op_off, int64_t n_rec,
                         int64_t* ins_sum, int64_t* match_sum, int64_t* match_cnt);

/* ---- resident shard engine: the whole read-side hot path of one chromosome in one enqueue --------- */
/* A vsv_shard owns the device copy of one chromosome's packed records and every table derived from them.
 * vsv_shard_scan enqueues, WITHOUT any host synchronisation in between:
 *   both read scanners in one pass over the ops (ERS:47-145 + SE:438-478), reference_end / read length per
 *   record, the read split pairs (ERS:147-248), the SA-tag analysis (SE:193-371), the in-read merge
 *   (SE:373-435), ONE radix sort that orders the read-signature table (sort_sig), the DEL and INS `.sigs`
 *   tables (SE:637-638) and the read ends (for the spanning-depth join), `sort -u` de-duplication, unique
 *   signatures (GTD:65-86, GTI:67-101).
 * Counts stay on the device; vsv_shard_counts reads them once.  vsv_shard_candidates runs the per-call joins
 * (FPF:106-123, CSS:81-253, GTD:92-170, GTI:107-196) and the genotype decision against the resident tables.
 * Reference call chain replaced: RVC:83-96 and FGC:111-172 for one chromosome. */
typedef struct vsv_shard vsv_shard;

typedef struct vsv_records {
  int64_t n_rec, n_ops;
  int32_t ops_format;            /* 32: BAM-native u32 words; 16 / 12 / 8: the staging formats above          */
  const void* ops;               /* 16-byte aligned                                                          */
  const int64_t* exc_idx;        /* exception list of the staging format (NULL for 32)                        */
  const uint32_t* exc_val;
  int64_t n_exc;
  const int64_t* op_off;         /* [n_rec + 1]                                                               */
  const int32_t* pos;            /* 0-based leftmost, non-decreasing (coordinate-sorted BAM)                  */
  const uint8_t* mapq;
  const uint16_t* flag;          /* may be NULL: every record is treated as primary forward (no SA analysis)  */
  const int32_t* qlen;           /* query_length: SEQ length, 0 for '*' (SE:439)                              */
  const int32_t* name_id;        /* interned qname; NULL: no `sort -u` de-duplication                         */
  const int32_t* next_same_name; /* next record in BAM order with the same qname, -1 = none (ERS:227-237);
                                    NULL: no split pairs                                                      */
  /* SA tags of the records that carry one (SE:495-511), parsed by the host as acquire_clip_pos does
   * (SE:328-358): entry e of read i in [sa_off[i], sa_off[i+1]) */
  int64_t n_sa_read, n_sa_ent;
  const int32_t* sa_rec;         /* record index of read i (ascending)                                        */
  const int64_t* sa_off;         /* [n_sa_read + 1]                                                           */
  const int32_t* sa_chrom;       /* interned chromosome of the entry; sa_self_chrom = id of this chromosome   */
  const int32_t* sa_pos;         /* POS field as written (1-based, kept: SE:354)                              */
  const uint8_t* sa_strand;      /* 0 '+', 1 '-'                                                              */
  const int32_t* sa_first_s;     /* leading / trailing soft clip (S only), sum of M,D,=,X                     */
  const int32_t* sa_last_s;
  const int32_t* sa_bias;
  const uint8_t* sa_mapq;
  int32_t sa_self_chrom;
} vsv_records;

typedef struct vsv_shard_counts {
  int64_t n_reads_cigar;  /* ERS CIGAR signatures                                     */
  int64_t n_reads_split;  /* ERS split-pair signatures (stored after the CIGAR ones)  */
  int64_t n_cute;         /* cuteSV-style signatures before the in-read merge         */
  int64_t n_del, n_ins;   /* `.sigs` rows (merged CIGAR rows + SA rows), before sort -u */
  int64_t n_sa_rows;      /* of which from the SA analysis                            */
  int64_t n_del_dup;      /* DEL rows removed by sort -u                              */
  int64_t n_ins_maybe_dup;/* INS rows equal in (pos,len,read) to an earlier one: the host compares sequences */
  int64_t n_udel, n_uins; /* unique signatures                                        */
  int64_t overflow;       /* bit mask of VSV_OVF_*: a capacity was too small, counts hold the need */
  int64_t n_stage;        /* staging slots reserved by the scan (sizing)              */
} vsv_shard_counts;
#define VSV_OVF_READS 1
#define VSV_OVF_CUTE 2
#define VSV_OVF_DEL 4
#define VSV_OVF_INS 8
#define VSV_OVF_STAGE 16
#define VSV_OVF_POS 32   /* a position >= 2^28: the chromosome is too long for the combined sort key */
#define VSV_OVF_DUPLIST 64

int vsv_shard_create(vsv_ctx* ctx, vsv_shard** out);
void vsv_shard_destroy(vsv_shard* s);
/* Capacities (rows) of the read-signature, raw cute, DEL and INS tables; 0 keeps the current one. */
int vsv_shard_reserve(vsv_shard* s, int64_t cap_reads, int64_t cap_cute, int64_t cap_del, int64_t cap_ins);
/* Copies (host-IO ctx) or adopts (device-IO ctx: the pointers must stay valid) one chromosome's records and
 * widens a staging format.  Asynchronous on the ctx stream with pinned host memory. */
int vsv_shard_load(vsv_shard* s, const vsv_records* r);
/* -include_bed regions of this chromosome, already widened by 1000 bp as load_bed does (SE:58-80), and the
 * chromosome length; n = -1: no BED (every record is in).  Host pointers. */
int vsv_shard_set_bed(vsv_shard* s, const int32_t* b0, const int32_t* b1, int64_t n, int64_t chrom_len);
int vsv_shard_scan(vsv_shard* s);
/* Waits for the stream and returns the counts of the last vsv_shard_scan. */
int vsv_shard_counts_get(vsv_shard* s, vsv_shard_counts* c);
/* INS rows the host found to be exact duplicates (same sequence too): their uids (field `first` of the
 * row), ascending.  Re-runs the grouping.  Host pointer. */
int vsv_shard_drop_ins(vsv_shard* s, const int32_t* uid, int64_t n);

/* tables a host formatter can fetch (host pointers, capacity in rows; returns rows in *n) */
#define VSV_TAB_READS 0      /* vsv_sig: CIGAR signatures in (record, op) order, then the split-pair ones */
#define VSV_TAB_CUTE 1       /* vsv_sig: raw cuteSV-style signatures in (record, op) order               */
#define VSV_TAB_DEL_ROWS 2   /* vsv_sig: `.sigs` rows in no particular order: rec, type (| 256 for a reversed
                                query in an SA row), pos, len, q0, q1 (SA rows: slice of the query), first = uid
                                (merged rows: index of the first folded raw signature in VSV_TAB_CUTE; SA rows:
                                negative), count (merged rows: raw signatures folded; SA rows: 0)           */
#define VSV_TAB_INS_ROWS 3
#define VSV_TAB_REF_END 4    /* int32 per record: reference_end                                            */
#define VSV_TAB_READ_LEN 5   /* int32 per record: read length incl. hard clips (ERS:151)                   */
#define VSV_TAB_IN_BED 6     /* uint8 per record                                                           */
#define VSV_TAB_INS_DUPS 7   /* int32 pairs (uid of the later row, uid of the earlier row)                 */
#define VSV_TAB_UDEL 8       /* int32 triples (pos, len, count) of the unique DEL signatures, file order   */
#define VSV_TAB_UINS 9
int vsv_shard_fetch(vsv_shard* s, int table, void* out, int64_t cap_rows, int64_t* n);

/* Per-call joins against the resident tables.  Candidates in VCF order on this chromosome: pos, svlen (DEL:
 * |SVLEN|, INS: SVLEN), type (VSV_DEL / VSV_INS), call_gt (0 '0/1', 1 '1/1', other = pass through).  `what`
 * selects the stages; outputs of unselected stages may be NULL:
 *   VSV_JOIN_FP       fp_support  eval_sig (FPF:106-123) against the read-signature table
 *   VSV_JOIN_COV      cov         calc_ins_call_cov / calc_del_call_cov (CSS:81-253), flanking cov_flanking
 *   VSV_JOIN_SUPPORT  n_support, match_id  match_varlist_siglist / extract_sig_support incl. the double count
 *                     at last_match_idx; DEL and INS candidates carry separate indices (the reference runs
 *                     GTD and GTI separately): last_in[2] / last_out[2] = {DEL, INS} (-1 = none; NULL = {0,0})
 *   VSV_JOIN_GENOTYPE depth2, ratio, new_gt  (GTD:140-170,206-217,291-318; GTI:188-196,228,258-283); needs
 *                     n_support from VSV_JOIN_SUPPORT in the same call or as input in `n_support`
 * Host-IO ctx: synchronous.  Device-IO ctx: asynchronous, last_out must be NULL. */
#define VSV_JOIN_FP 1
#define VSV_JOIN_COV 2
#define VSV_JOIN_SUPPORT 4
#define VSV_JOIN_GENOTYPE 8
int vsv_shard_candidates(vsv_shard* s, int what, const int32_t* cand_pos, const int32_t* cand_svlen,
                         const int32_t* cand_type, const int32_t* call_gt, int64_t n_cand, const int32_t* last_in,
                         int32_t* fp_support, int64_t* cov, int32_t* n_support, int32_t* match_id, int32_t* depth2,
                         double* ratio, int32_t* new_gt, int32_t* last_out);
/* CUDA-event time (ms) of the streaming scan kernel of the last vsv_shard_scan. */
int vsv_shard_scan_kernel_ms(vsv_shard* s, float* ms);

/* Closed-form spanning depth (GTD:140-146): reads sorted by start; `sorted_end` = the read ends sorted
 * ascending.  depth = #{start < S} - #{end <= E} + #{S <= start <= E and end <= E}. */
int vsv_span_depth_sorted(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, const int32_t* sorted_end,
                          int64_t n_reads, const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth);

/* vsv_span_depth for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as
 * vsv_span_depth; its temporaries (the ascending copy of the read ends and the sort's work space) live in one fixed
 * file-scope device array (17 MiB, held by every GPU that loads the library until exit), so the call performs no
 * cudaMalloc / cudaFree and touches no arena.  Requirements: a device-IO context (every pointer is a DEVICE pointer;
 * else VSV_E_ARG); n_reads <= VSV_NOALLOC_MAX_READS (else VSV_E_CAPACITY); the first call on a device resolves the
 * array's address and should be made outside a graph capture.  The array is shared by all calls on a device and keeps
 * the last call's contents: calls that overlap on the GPU race on it -- serialise them.  Asynchronous on the ctx's stream. */
#define VSV_NOALLOC_MAX_READS (1 << 21)
int vsv_span_depth_noalloc(vsv_ctx* ctx, const int32_t* read_start, const int32_t* read_end, int64_t n_reads,
                           const int32_t* S, const int32_t* E, int64_t n_win, int32_t* depth);

/* ---- host-side BAM feeder (no CUDA; zlib + threads) ------------------------------------- */
/* Replaces `pysam.AlignmentFile(path).fetch(chr)` and the per-record attribute reads of the
 * whole-chromosome scans (LI/extract_reads_signature.py:107-145, LI/sig_extract.py:514-560,
 * LI/extract_contig_signature_Hifi.py:386-419): BGZF inflate (blocks in parallel) + BAM record
 * -> the packed columns of this header.  The file is streamed in 32 MB windows (blocks of a
 * window inflate in parallel); what stays in memory are the kept records without their qualities.
 * vsv_bam_open keeps every mapped record; vsv_bam_open_ref keeps one reference ("" = header only),
 * seeks to it through <path>.bai / <path minus .bam>.bai when present and stops after it in a
 * coordinate-sorted file.  Unmapped records (refID < 0) are skipped, CG:B,I long CIGARs are
 * resolved.  Errors: VSV_E_ARG + vsv_bam_last_error() (thread-local message). */
typedef struct vsv_bam vsv_bam;
int vsv_bam_open(const char* path, int n_threads, vsv_bam** out);
int vsv_bam_open_ref(const char* path, const char* ref_name, int n_threads, vsv_bam** out);
void vsv_bam_close(vsv_bam* b);
const char* vsv_bam_last_error(void);
int vsv_bam_n_ref(const vsv_bam* b);
int vsv_bam_ref(const vsv_bam* b, int i, const char** name, int64_t* len, int64_t* n_rec);
/* sizes[5] = records, CIGAR ops, qname bytes, SEQ bytes, SA-tag bytes of reference `tid` */
int vsv_bam_chrom_sizes(vsv_bam* b, int tid, int64_t* sizes);
/* Columns of reference `tid` in file order; NULL skips a column.  `ops` = BAM-native u32,
 * `ops16` or `ops8` (not both) + (exc_idx, exc_val, exc_cap, n_exc) = the staging formats of
 * vsv_expand_ops16 / vsv_expand_ops8 (VSV_E_CAPACITY with *n_exc = needed when the list is too small).  *_off arrays have n_rec + 1
 * entries; names / seq (ASCII) / sa are concatenated without terminators. */
int vsv_bam_chrom_fill(vsv_bam* b, int tid, int32_t* pos, uint8_t* mapq, uint16_t* flag, int32_t* qlen,
                       int64_t* op_off, uint32_t* ops, uint16_t* ops16, uint8_t* ops8, int64_t* exc_idx,
                       uint32_t* exc_val, int64_t exc_cap, int64_t* n_exc, int64_t* name_off, char* names,
                       int64_t* seq_off, char* seq, int64_t* sa_off, char* sa);
/* SEQ (ASCII) of record i of reference `tid`: sparse `read.seq` accesses (SE:452, ECS_H:642-700).
 * VSV_E_CAPACITY with *len = needed when cap is too small. */
int vsv_bam_record_seq(vsv_bam* b, int tid, int64_t i, char* out, int64_t cap, int64_t* len);

#ifdef __cplusplus
}
#endif
#endif
