"""One chromosome of the read alignments on the GPU, scanned ONCE, serving every read-side operator.

The reference reads the read BAM five times per chromosome (ERS:113,202,215; SE:522; GTD:143 / GTI:193 per
candidate).  `ReadShard` packs the records once, loads them into a `vsv_shard` (csrc/vsv_shard.cu) and runs the
fused step once; `extract_reads_signature`, `fp_filter`, `sig_extract`, `signature_support` and `correct_gt`
(volcanosv_b200.operators) are formatters / per-call joins over that resident result.  Shards stay cached per
(alignment file, chromosome) with their CIGAR ops released, so the whole-genome driver keeps all 22 resident.
"""
from __future__ import annotations

import re
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import lib as L

_CIG = re.compile(r"(\d+)([MIDNSHP=X])")


def intern_names(qname) -> Tuple[np.ndarray, np.ndarray]:
    """(name_id per record, first record index of every name).  name ids are arbitrary but stable."""
    a = qname if isinstance(qname, np.ndarray) else np.asarray(list(qname))
    n = len(a)
    if n == 0:
        return np.zeros(0, np.int32), np.zeros(0, np.int64)
    _, inv = np.unique(a, return_inverse=True)
    inv = inv.reshape(-1)
    first = np.full(int(inv.max()) + 1, n, np.int64)
    np.minimum.at(first, inv, np.arange(n, dtype=np.int64))
    return inv.astype(np.int32), first


def parse_sa(entry: str):
    """acquire_clip_pos + the SA fields organize_split_signal uses (SE:328-358)."""
    chrom, pos, strand, cg, mq, _nm = entry.split(",")
    items = [(int(n), ch) for n, ch in _CIG.findall(cg)]
    fs = items[0][0] if items[0][1] == "S" else 0
    ls = items[-1][0] if items[-1][1] == "S" else 0
    bias = sum(n for n, ch in items if ch in "MD=X")
    return chrom, int(pos), 0 if strand == "+" else 1, fs, ls, bias, int(mq)


def pack_sa(R, chrom: str):
    """SA columns of vsv_records for the records that carry an SA tag (any flag: the engine applies SE:495-511)."""
    ids: Dict[str, int] = {chrom: 0}
    rec, off = [], [0]
    cols: List[list] = [[] for _ in range(7)]
    for r in range(R.n):
        s = R.sa[r]
        if s is None:
            continue
        for e in s.split(";")[:-1]:
            c, p, st, fs, ls, bias, mq = parse_sa(e)
            for k, v in enumerate((ids.setdefault(c, len(ids)), p, st, fs, ls, bias, min(mq, 255))):
                cols[k].append(v)
        rec.append(r)
        off.append(len(cols[0]))
    if not rec:
        return None
    return dict(rec=np.asarray(rec, np.int32), off=np.asarray(off, np.int64), chrom=np.asarray(cols[0], np.int32),
                pos=np.asarray(cols[1], np.int32), strand=np.asarray(cols[2], np.uint8), first_s=np.asarray(cols[3], np.int32),
                last_s=np.asarray(cols[4], np.int32), bias=np.asarray(cols[5], np.int32), mapq=np.asarray(cols[6], np.uint8))


class ReadShard:
    """The resident tables of one chromosome of a read alignment file."""

    def __init__(self, ctx: L.Context, R, chrom: str, chrom_len: int):
        self.ctx, self.R, self.chrom, self.chrom_len = ctx, R, chrom, int(chrom_len)
        key = getattr(R, "name_key", None)
        if key is not None and isinstance(key, np.ndarray) and key.dtype.kind in "iu":
            # a packed batch whose names are interned already (ids from the decoder): nothing to hash
            self.name_id, self._name_first = np.ascontiguousarray(key, dtype=np.int32), None
        else:
            self.name_id, self._name_first = intern_names(key if key is not None else R.qname)
        self.shard = L.Shard(ctx)
        if R.n:
            # a packed batch may bring the links and the parsed SA columns with it (they are part of the host packing)
            nxt = getattr(R, "next_same_name", None)
            sa = getattr(R, "sa_packed", None)
            self.shard.load(R.ops, R.op_off, R.pos, R.mapq, R.qlen, R.flag.astype(np.uint16), self.name_id,
                            nxt if nxt is not None else L.next_same_name(self.name_id),
                            sa=sa if sa is not None else pack_sa(R, chrom), self_chrom=0)
        self.bed = "unset"
        self.counts: Optional[dict] = None
        self._cache: Dict[int, np.ndarray] = {}

    @property
    def name_first(self):
        """First record index of every name id (the first-appearance order of ERS:227-237), built when a formatter asks."""
        if self._name_first is None:
            n = len(self.name_id)
            first = np.full(int(self.name_id.max()) + 1 if n else 0, n, np.int64)
            np.minimum.at(first, self.name_id, np.arange(n, dtype=np.int64))
            self._name_first = first
        return self._name_first

    # ---- the one scan ----------------------------------------------------------------------------------
    def scan(self, bed: Optional[List[Tuple[int, int]]] = None):
        """Runs the fused step (again only if the -include_bed regions changed).  bed: regions of THIS chromosome
        already widened by 1000 bp (SE:69), or None."""
        if bed is not None and any(r0 <= 0 and r1 >= self.chrom_len for r0, r1 in bed):
            # a region that covers the whole chromosome (the driver's sample.bed, FGC:71-75) is attached to every task and
            # overlaps every record: same result as no BED, and the scan that already ran without one is reused
            bed = None
        key = None if bed is None else tuple(bed)
        if self.counts is not None and key == self.bed:
            return self.counts
        self._cache = {}
        if self.R.n == 0:
            self.counts = dict(n_reads_cigar=0, n_reads_split=0, n_cute=0, n_del=0, n_ins=0, n_sa_rows=0, n_del_dup=0,
                               n_ins_maybe_dup=0, n_udel=0, n_uins=0, overflow=0, n_stage=0)
            self.bed = key
            return self.counts
        self.shard.set_bed(bed, self.chrom_len)
        self.counts = self.shard.scan()
        self.bed = key
        if self.counts["n_ins_maybe_dup"]:
            self._resolve_ins_duplicates()
        return self.counts

    def fetch(self, table):
        if table not in self._cache:
            if self.R.n == 0:
                empty = {L.TAB_REF_END: np.zeros(0, np.int32), L.TAB_READ_LEN: np.zeros(0, np.int32), L.TAB_IN_BED: np.zeros(0, np.uint8),
                         L.TAB_INS_DUPS: np.zeros((0, 2), np.int32), L.TAB_UDEL: (np.zeros(0, np.int32),) * 3, L.TAB_UINS: (np.zeros(0, np.int32),) * 3}
                self._cache[table] = empty.get(table, np.zeros(0, L.SIG_DTYPE))
            else:
                self._cache[table] = self.shard.fetch(table)
        return self._cache[table]

    # ---- INS rows: sequences (host only) ---------------------------------------------------------------
    def _ins_piece_index(self):
        """For the merged INS rows: raw INS signatures in (record, op) order and the rank of each raw index."""
        cute = self.fetch(L.TAB_CUTE)
        is_ins = cute["type"] == L.INS
        rank = np.cumsum(is_ins) - 1                 # rank of raw signature k among the INS ones (valid where is_ins)
        return cute, np.flatnonzero(is_ins), rank

    def ins_sequence(self, row, pieces=None) -> str:
        """The sequence column of an INS `.sigs` row (SE:394-415 concatenated pieces; SA rows: slice of the possibly
        reversed query, SE:220-232)."""
        seq = self.R.seq[int(row["rec"])]
        if int(row["count"]) == 0:                   # SA row
            q = seq[::-1] if int(row["type"]) & L.ROW_SA_REV else seq
            return q[int(row["q0"]):int(row["q1"])]
        cute, ins_idx, rank = pieces if pieces is not None else self._ins_piece_index()
        r0 = int(rank[int(row["first"])])
        return "".join(seq[int(cute["q0"][k]):int(cute["q1"][k])] for k in ins_idx[r0: r0 + int(row["count"])])

    def _resolve_ins_duplicates(self):
        """`sort -u` on INS lines compares the sequence too (A-10): the engine reports rows equal in (pos, len, read);
        the ones whose sequences are equal as well are dropped."""
        dups = self.fetch(L.TAB_INS_DUPS)
        rows = self.fetch(L.TAB_INS_ROWS)
        by_uid = {int(u): i for i, u in enumerate(rows["first"])}
        pieces = self._ins_piece_index()
        seq_of: Dict[int, str] = {}

        def s(uid):
            if uid not in seq_of:
                seq_of[uid] = self.ins_sequence(rows[by_uid[uid]], pieces)
            return seq_of[uid]
        # a later row is a duplicate if ANY earlier row of its (pos, len, read) group has the same sequence; the engine
        # reports one earlier row per later row, so compare within the reported groups
        groups: Dict[Tuple[int, int, int], List[int]] = {}
        for later, earlier in dups:
            for u in (int(earlier), int(later)):
                r = rows[by_uid[u]]
                k = (int(r["pos"]), int(r["len"]), int(self.name_id[r["rec"]]))
                if u not in groups.setdefault(k, []):
                    groups[k].append(u)
        drop = []
        for uids in groups.values():
            seen = set()
            for u in uids:
                q = s(u)
                if q in seen:
                    drop.append(u)
                seen.add(q)
        self._cache = {}
        self.counts = self.shard.drop_ins(drop)

    # ---- joins -------------------------------------------------------------------------------------------
    def candidates(self, what, pos, svlen, typ, gt=None, last_in=(0, 0), n_support=None):
        n = len(pos)
        if self.R.n == 0 or n == 0:
            z32 = np.zeros(n, np.int32)
            nan = np.full(n, np.nan)
            sup = z32 if n_support is None else np.asarray(n_support, np.int32)
            is_del = np.asarray(typ) == L.DEL
            return dict(fp_support=np.where(np.asarray(svlen) > self.ctx.get_params().fp_max_comp, 60, 0).astype(np.int32), cov=np.zeros(n, np.int64),
                        n_support=sup.copy(), match_id=np.full(n, last_in[0], np.int32), depth2=z32.copy(),
                        ratio=np.where(is_del, 1.0, np.where(sup > 0, np.inf, nan)), new_gt=None, last_out=tuple(last_in))
        return self.shard.candidates(what, pos, svlen, typ, gt, last_in, n_support)

    def release_inputs(self):
        """Keeps the tables, drops the host copy of the ops (the device copy goes with the next load)."""
        self.R.ops = None

    def close(self):
        self.shard.close()


# ---- process-wide registry: (alignment path, chromosome) -> ReadShard, and which files were written from it ----------
_shards: Dict[Tuple[str, str], ReadShard] = {}
_file_owner: Dict[str, Dict[str, ReadShard]] = {}      # path of a reads_sig.txt / .sigs file -> {chrom: shard}


def get(ctx, source, path: str, chrom: str, chrom_len: int) -> ReadShard:
    key = (path, chrom)
    if key not in _shards:
        _shards[key] = ReadShard(ctx, source.records(chrom), chrom, chrom_len)
    return _shards[key]


def register_file(path: str, chrom: str, shard: ReadShard):
    import os
    _file_owner.setdefault(os.path.abspath(path), {})[chrom] = shard


def owner_of(path: str) -> Optional[Dict[str, ReadShard]]:
    """The resident shards a file was written from in THIS process, if it has not been modified since."""
    import os
    p = os.path.abspath(path)
    ent = _file_owner.get(p)
    if ent is None or not os.path.exists(p):
        return None
    return ent


def clear():
    for s in _shards.values():
        s.close()
    _shards.clear()
    _file_owner.clear()


def _at_exit():
    """Destroy the engines while the CUDA runtime is still up (Python's atexit runs before the C-level handlers)."""
    try:
        clear()
    except Exception:  # noqa: BLE001
        pass


import atexit  # noqa: E402

atexit.register(_at_exit)
