"""Shared helpers of the parity tests: workloads named after BASELINE.json's configs."""
import numpy as np
import ora

BSW_FIELDS = ora.BSW_OUT
KSW_FIELDS = ("score", "te", "qe", "score2", "te2", "tb", "qb")


def diff_fields(a, b, fields):
    bad = np.zeros(len(a), dtype=bool)
    for f in fields:
        bad |= a[f] != b[f]
    return bad


def describe_first(a, b, bad, extra=None):
    k = int(np.nonzero(bad)[0][0])
    return f"{int(bad.sum())} mismatching pairs; first at {k}: oracle={a[k]} got={b[k]}" + (f" {extra[k]}" if extra is not None else "")


def bsw_c1(n_reads, seed=0xC1, read_len=150, div=0.01):
    """C1/C2-shaped extension pairs; right h0 chained from the oracle's left scores."""
    d = ora.gen_bsw(n_reads, read_len=read_len, seed=seed, div=div)
    L, lr, lq = d["left"]
    R, rr, rq = d["right"]
    prm = ora.make_params()
    Lo = L.copy()
    ora.bsw_batch(Lo, lr, lq, prm, 100, 8)
    ora.chain_h0(Lo, R)
    return (L, lr, lq), (R, rr, rq)


def concat_workloads(parts):
    """Concatenate (pairs, ref, qer) triples into one call's buffers (offsets rebased)."""
    pairs, refs, qers = [], [], []
    ro = qo = 0
    for p, r, q in parts:
        p = p.copy()
        p["idr"] += ro
        p["idq"] += qo
        pairs.append(p); refs.append(r); qers.append(q)
        ro += len(r); qo += len(q)
    out = np.concatenate(pairs)
    out["id"] = np.arange(len(out))
    return out, np.concatenate(refs), np.concatenate(qers)


def edge_case_pairs():
    """The fork's documented edge cases (R:GRAVITON4_SVE2_RESULTS.md:398-413,
    R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:280-294): empty, max-length, all-N,
    repetitive, unequal lengths, all-match, all-mismatch."""
    rng = np.random.default_rng(42)
    cases = []

    def add(q, t, h0):
        cases.append((np.asarray(q, np.uint8), np.asarray(t, np.uint8), int(h0)))
    q = rng.integers(0, 4, 100)
    add(q, q, 30)                                   # all match, equal length
    add(q, np.concatenate([q, rng.integers(0, 4, 60)]), 50)   # all match, longer target
    add(q, (q + 1) % 4, 40)                         # all mismatch
    add([], q, 20)                                  # empty query
    add(q, [], 20)                                  # empty target
    add([], [], 7)                                  # both empty
    add(np.full(80, 4), np.full(90, 4), 60)         # all N
    add(np.tile([0, 1], 60), np.tile([0, 1], 64), 25)          # repetitive
    add(np.zeros(127, int), np.zeros(127, int), 1)  # max length 8-bit tier, homopolymer
    add(q[:10], np.concatenate([q[:10], rng.integers(0, 4, 110)]), 100)  # very unequal
    add(q, np.concatenate([q[:50], q[53:]]), 70)    # 3-base deletion in target
    add(q, np.concatenate([q[:50], [3, 3, 3, 3], q[50:]]), 70)  # 4-base insertion in target
    add(q[:1], q[:1], 1)                            # single cell, h0 = 1
    qn = q.copy(); qn[::7] = 4
    add(qn, q, 33)                                  # N sprinkled in the query
    add(rng.integers(0, 4, 120), rng.integers(0, 4, 125), 120)  # unrelated: z-drop / zero row
    add(rng.integers(0, 4, 120), rng.integers(0, 4, 125), 3)    # unrelated, tiny h0: dies at once
    n = len(cases)
    pairs = np.zeros(n, dtype=ora.SEQPAIR)
    refs, qers = [], []
    ro = qo = 0
    for i, (qq, tt, h0) in enumerate(cases):
        pairs[i]["idr"], pairs[i]["idq"], pairs[i]["id"] = ro, qo, i
        pairs[i]["len1"], pairs[i]["len2"], pairs[i]["h0"] = len(tt), len(qq), h0
        refs.append(tt); qers.append(qq)
        ro += len(tt); qo += len(qq)
    ref = np.concatenate(refs + [np.zeros(64, np.uint8)])
    qer = np.concatenate(qers + [np.zeros(64, np.uint8)])
    return pairs, ref, qer
