"""Host-side multi-GPU logic (volcanosv_b200/shard.py): LPT assignment, ordering rules, the
last_match_idx carry rule against the oracle run genome-wide, and a world_size-2 gloo gather."""
import os
import socket

import numpy as np

import oracle as O
from volcanosv_b200 import shard as S


def test_lpt_assignment_is_balanced_and_deterministic():
    hg19 = [249250621, 243199373, 198022430, 191154276, 180915260, 171115067, 159138663, 146364022, 141213431,
            135534747, 135006516, 133851895, 115169878, 107349540, 102531392, 90354753, 81195210, 78077248,
            59128983, 63025520, 48129895, 51304566]                          # LI/header:5-26
    w = {"chr%d" % (i + 1): float(n) for i, n in enumerate(hg19)}
    for n in (1, 2, 4, 8):
        a = S.assign_lpt(w, n)
        assert sorted(sum(a, [])) == sorted(w)
        loads = [sum(w[c] for c in r) for r in a]
        assert max(loads) <= sum(loads) / n * 1.12 + 1
        assert a == S.assign_lpt(w, n)
    assert S.assign_lpt({}, 3) == [[], [], []]


def test_orders():
    names = ["chr%d" % i for i in range(1, 23)]
    assert S.chrom_numeric_order(reversed(names)) == names
    d = S.chrom_dictionary_order(names)
    assert d[:3] == ["chr1", "chr10", "chr11"] and d[-1] == "chr9"


def test_support_carry_rule_matches_genome_wide_oracle():
    """Per-chromosome shards with support_carry_in() == match_varlist_siglist run genome-wide."""
    rng = np.random.default_rng(5)
    chroms = ["chr1", "chr2", "chr10"]
    for trial in range(30):
        sig = {c: None for c in chroms}
        if trial % 5 == 0:
            sig.pop("chr1")                                    # first block is then chr10
        for c in sig:
            n = int(rng.integers(1, 40))
            p = np.sort(rng.integers(1000, 20000, n)).astype(np.int32)
            sig[c] = (p, rng.integers(30, 400, n).astype(np.int32), rng.integers(1, 5, n).astype(np.int32))
        cand = {}
        for c in chroms:
            n = int(rng.integers(0, 12))
            cand[c] = (np.sort(rng.integers(1000, 20000, n)).astype(np.int32), rng.integers(30, 400, n).astype(np.int32))
        cid = {c: i for i, c in enumerate(S.chrom_dictionary_order(chroms))}
        so = S.chrom_dictionary_order(list(sig))
        g_sc = np.concatenate([np.full(len(sig[c][0]), cid[c], np.int32) for c in so])
        g_sp, g_sl, g_sn = (np.concatenate([sig[c][k] for c in so]) for k in range(3))
        co = S.chrom_numeric_order(chroms)
        g_cc = np.concatenate([np.full(len(cand[c][0]), cid[c], np.int32) for c in co])
        g_cp, g_cl = (np.concatenate([cand[c][k] for c in co]) for k in range(2))
        exp, _ = O.sig_support(g_cc, g_cp, g_cl, g_sc, g_sp, g_sl, g_sn)
        got, matched = [], False
        for c in co:
            if c not in sig:
                got += [0] * len(cand[c][0])
                continue
            z = np.zeros(len(cand[c][0]), np.int32)
            zs = np.zeros(len(sig[c][0]), np.int32)
            last_in = S.support_carry_in(c, co, list(sig), matched)
            s, _, last = O.sig_support(z, cand[c][0], cand[c][1], zs, *sig[c], last_in=last_in, return_last=True)
            got += list(s)
            matched = matched or (last != last_in)
        assert list(exp) == got, trial


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = {"chr%d" % i: float(30 - i) for i in range(1, 9)}
    mine = S.assign_lpt(w, world)[rank]
    local = {c: "%s:body-from-rank%d\n" % (c, rank) for c in mine}
    allr = S.gather_to_root(local)
    if rank == 0:
        q.put(S.concat_in_order(allr, S.chrom_numeric_order(list(w))))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_gather_concat():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    out = q.get(timeout=120)
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    assert [x.split(":")[0] for x in out] == ["chr%d" % i for i in range(1, 9)]
    assert {x.strip()[-5:] for x in out} == {"rank0", "rank1"}


def test_split_records_balanced_and_complete():
    import numpy as np
    from volcanosv_b200 import shard as SH
    rng = np.random.default_rng(3)
    n_ops = rng.integers(0, 5000, 1000)
    n_ops[::50] = 0                                            # records without ops
    off = np.concatenate([[0], np.cumsum(n_ops)])
    for parts in (1, 2, 3, 8, 2000):
        r = SH.split_records(off, parts)
        assert len(r) == parts and r[0][0] == 0 and r[-1][1] == 1000
        assert all(a[1] == b[0] for a, b in zip(r, r[1:])) and all(lo <= hi for lo, hi in r)
        if parts <= 8:
            load = [off[hi] - off[lo] for lo, hi in r]
            assert max(load) - min(load) <= 2 * 5000
    b = dict(ops=np.arange(off[-1], dtype=np.uint32), op_off=off, pos=np.arange(1000, dtype=np.int32))
    lo, hi = SH.split_records(off, 3)[1]
    t = SH.take_records(b, lo, hi)
    assert t["op_off"][0] == 0 and len(t["ops"]) == t["op_off"][-1] and t["ops"][0] == off[lo]
    assert np.array_equal(t["pos"], b["pos"][lo:hi])


def test_bind_host_to_gpu_is_best_effort():
    """No driver here: the helper reports why it did nothing and leaves the affinity mask alone."""
    import os
    from volcanosv_b200 import shard as SH
    before = os.sched_getaffinity(0)
    info = SH.bind_host_to_gpu(0)
    assert isinstance(info, dict) and "bound" in info
    if not info["bound"]:
        assert os.sched_getaffinity(0) == before
    else:
        assert os.sched_getaffinity(0) <= before
        os.sched_setaffinity(0, before)


def test_write_fai_variants(tmp_path):
    """the samtools-faidx stand-in on FASTA variants the real tool accepts (empty record, CRLF, short last line)
    and the one it refuses (ragged lines)."""
    import pytest
    from volcanosv_b200 import operators as OP
    p = tmp_path / "a.fa"
    p.write_bytes(b">chr1 desc\nACGTAC\nACGTAC\nAC\n>empty\n>chr2\nAAAA\n")
    assert OP._write_fai(str(p)) == [("chr1", 14), ("empty", 0), ("chr2", 4)]
    assert open(str(p) + ".fai").read() == "chr1\t14\t11\t6\t7\nempty\t0\t35\t0\t0\nchr2\t4\t41\t4\t5\n"
    q = tmp_path / "crlf.fa"
    q.write_bytes(b">chr1\r\nACGT\r\nAC\r\n")
    assert OP._write_fai(str(q)) == [("chr1", 6)]
    assert open(str(q) + ".fai").read() == "chr1\t6\t7\t4\t6\n"
    r = tmp_path / "ragged.fa"
    r.write_bytes(b">chr1\nACGT\nAC\nACGT\n")
    with pytest.raises(ValueError):
        OP._write_fai(str(r))


def test_assemble_support_chain_matches_genome_wide_oracle():
    """Rank 0's assembly of the sharded supporting-signature count (operators._fgc_sharded): every shard computes its
    candidates with the local carry-in -1 (and 0 for the block at global index 0), `assemble_support_chain` picks the variant
    and rebuilds the GLOBAL match_id column == match_varlist_siglist run over the whole genome (GTD:92-137)."""
    rng = np.random.default_rng(11)
    chroms = ["chr1", "chr2", "chr3", "chr10", "chr11"]
    for trial in range(40):
        sig = {}
        for c in chroms:
            if (trial % 4 == 0 and c == "chr1") or rng.random() < 0.15:
                continue                                       # a chromosome without signatures (first block moves to chr10)
            n = int(rng.integers(1, 30))
            sig[c] = (np.sort(rng.integers(1000, 20000, n)).astype(np.int32), rng.integers(30, 400, n).astype(np.int32),
                      rng.integers(1, 5, n).astype(np.int32))
        cand = {c: (np.sort(rng.integers(1000, 20000, int(rng.integers(0, 10)))).astype(np.int32),) for c in chroms}
        cand = {c: (p[0], rng.integers(30, 400, len(p[0])).astype(np.int32)) for c, p in cand.items()}
        cid = {c: i for i, c in enumerate(S.chrom_dictionary_order(chroms))}
        so = S.chrom_dictionary_order(list(sig))
        co = [c for c in S.chrom_numeric_order(chroms) if len(cand[c][0])]
        if not so or not co:
            continue
        g_sc = np.concatenate([np.full(len(sig[c][0]), cid[c], np.int32) for c in so])
        g_sp, g_sl, g_sn = (np.concatenate([sig[c][k] for c in so]) for k in range(3))
        g_cc = np.concatenate([np.full(len(cand[c][0]), cid[c], np.int32) for c in co])
        g_cp, g_cl = (np.concatenate([cand[c][k] for c in co]) for k in range(2))
        exp_sup, exp_mid = O.sig_support(g_cc, g_cp, g_cl, g_sc, g_sp, g_sl, g_sn)
        blocks, off = {}, 0
        for c in so:
            blocks[c] = (off, len(sig[c][0]))
            off += len(sig[c][0])
        runs = []
        for c in co:                                           # what each rank computes for its chromosome
            n = len(cand[c][0])
            z = np.zeros(n, np.int32)
            run = dict(chrom=c)
            for tag, li in (("a", -1), ("b", 0)):
                if tag == "b" and blocks.get(c, (1, 0))[0] != 0:
                    continue
                if c in sig:
                    s, m, last = O.sig_support(z, cand[c][0], cand[c][1], np.zeros(len(sig[c][0]), np.int32), *sig[c], last_in=li, return_last=True)
                else:
                    s, m, last = [0] * n, [li] * n, li
                run[tag] = (np.asarray(s), np.asarray(m), last)
            runs.append(run)
        chosen, _ = S.assemble_support_chain(runs, blocks)
        got_sup = np.concatenate([x[1] for x in chosen])
        got_mid = np.concatenate([x[2] for x in chosen])
        assert list(got_sup) == list(exp_sup), trial
        assert list(got_mid) == list(exp_mid), trial
