"""Developer helper: a sequential Python model of k_edit_banded's schedule (csrc/vsv_edit.cu) -- same stripe arithmetic,
activation rule and carry ring -- checked against the oracle DP.  Used to debug the band logic without a GPU."""
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as O  # noqa: E402

M64 = (1 << 64) - 1


def banded(a, b, k0=None, grow=8, L=32):
    if len(a) > len(b):
        a, b = b, a
    m, n = len(a), len(b)
    if m == 0:
        return n, 0
    W, Q = (m + 63) >> 6, (n + 31) >> 5
    peq = {}
    for i, ch in enumerate(a):
        peq[(ch, i >> 6)] = peq.get((ch, i >> 6), 0) | (1 << (i & 63))
    delta = n - m
    k = max(delta + 64, (m + n) >> 6) if k0 is None else k0
    Tn = Q + W - 1
    rounds = 0
    while True:
        rounds += 1
        lo, hi = (k - delta) >> 1, (k + delta) >> 1
        carry = [[None] * W, [None] * W]
        pv, mv = [None] * W, [None] * W
        for T in range(Tn):
            par = T & 1
            for w in range(W):
                q = T - w
                if q < 0 or q >= Q:
                    continue
                row0, rows_end = w << 6, min((w << 6) + 64, m)
                jlo, jhi = row0 + 1 - lo, rows_end + hi
                qlo = 0 if jlo <= 1 else (jlo - 1) >> 5
                qhi = min(Q - 1, (jhi - 1) >> 5)
                if q < qlo or q > qhi:
                    continue
                above = False
                if w > 0:
                    ajlo, ajhi = jlo - 64, row0 + hi
                    aqlo = 0 if ajlo <= 1 else (ajlo - 1) >> 5
                    aqhi = min(Q - 1, (ajhi - 1) >> 5)
                    above = aqlo <= q <= aqhi
                cmax = min(32, n - (q << 5))
                hp_in, hm_in, above_after = 0xFFFFFFFF, 0, 0
                if above:
                    hp_in, hm_in, above_after = carry[par ^ 1][w - 1]
                if q == qlo:
                    P, Mv = M64, 0
                    cm = (1 << cmax) - 1
                    if q > 0:
                        assert above
                    top = row0 if q == 0 else above_after - (bin(hp_in & cm).count("1") - bin(hm_in & cm).count("1"))
                    score = top + rows_end - row0
                else:
                    P, Mv = pv[w], mv[w]
                    score = carry[par ^ 1][w][2]
                topbit = rows_end - row0 - 1
                hp_out = hm_out = 0
                for c in range(cmax):
                    eq0 = peq.get((b[(q << 5) + c], w), 0)
                    hin_p, hin_n = (hp_in >> c) & 1, (hm_in >> c) & 1
                    Xv = eq0 | Mv
                    eq = eq0 | hin_n
                    Xh = ((((eq & P) + P) & M64) ^ P) | eq
                    Ph = (Mv | ~(Xh | P)) & M64
                    Mh = P & Xh
                    hp_out |= ((Ph >> topbit) & 1) << c
                    hm_out |= ((Mh >> topbit) & 1) << c
                    Ph = ((Ph << 1) | hin_p) & M64
                    Mh = ((Mh << 1) | hin_n) & M64
                    P = (Mh | ~(Xv | Ph)) & M64
                    Mv = Ph & Xv
                score += bin(hp_out).count("1") - bin(hm_out).count("1")
                pv[w], mv[w] = P, Mv
                carry[par][w] = (hp_out, hm_out, score)
        result = carry[(Tn - 1) & 1][W - 1][2]
        if result <= k:
            return result, rounds
        k <<= 3


if __name__ == "__main__":
    rng = random.Random(3)
    n_multi = 0
    for trial in range(400):
        m = rng.choice([1, 2, 31, 63, 64, 65, 127, 128, 129, 200, 333, 700])
        a = "".join(rng.choice("ACGT") for _ in range(m))
        mode = rng.random()
        if mode < 0.4:
            b = list(a)
            for _ in range(rng.randint(0, max(1, m // 8))):
                kk = rng.randrange(len(b)) if b else 0
                u = rng.random()
                if u < 0.3 and len(b) > 1:
                    b.pop(kk)
                elif u < 0.6:
                    b.insert(kk, rng.choice("ACGTN"))
                elif b:
                    b[kk] = rng.choice("ACGT")
            b = "".join(b)
        else:
            b = "".join(rng.choice("ACGTN") for _ in range(rng.choice([m, m // 2 + 1, 2 * m + 3, 5 * m + 1, max(1, m - 1)])))
        k0 = rng.choice([None, None, abs(len(a) - len(b)) + 64])
        got, rounds = banded(a, b, k0=k0, grow=8)
        n_multi += rounds > 1
        exp = O.edit_distance(a, b)
        assert got == exp, (trial, len(a), len(b), got, exp, rounds)
    print("ok; pairs that needed more than one round:", n_multi)
