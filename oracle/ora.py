"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE, see oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  "Parity unpinned": see oracle.h.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

SEQPAIR = np.dtype([(n, "<i4") for n in (
    "idr", "idq", "id", "len1", "len2", "h0", "seqid", "regid",
    "score", "tle", "gtle", "qle", "gscore", "max_off")])
KSWR = np.dtype([(n, "<i4") for n in ("score", "te", "qe", "score2", "te2", "tb", "qb")])
assert SEQPAIR.itemsize == 56 and KSWR.itemsize == 28

BSW_OUT = ("score", "tle", "gtle", "qle", "gscore", "max_off")
GLOBALJOB = np.dtype([(n, "<i4") for n in ("idr", "idq", "tlen", "qlen", "w", "score", "n_cigar", "cigar_off")])
assert GLOBALJOB.itemsize == 32
KSW_XBYTE, KSW_XSTOP, KSW_XSUBO, KSW_XSTART = 0x10000, 0x20000, 0x40000, 0x80000


class Params(C.Structure):
    _fields_ = [("o_del", C.c_int32), ("e_del", C.c_int32), ("o_ins", C.c_int32), ("e_ins", C.c_int32),
                ("zdrop", C.c_int32), ("end_bonus", C.c_int32), ("mat", C.c_int8 * 25)]


def fill_scmat(a=1, b=4, ambig=-1):
    """bwa bwa_fill_scmat: a on the diagonal, -b off it, `ambig` in row/col 4."""
    m = np.full((5, 5), ambig, dtype=np.int8)
    for i in range(4):
        for j in range(4):
            m[i, j] = a if i == j else -b
    return m.reshape(-1)


def make_params(o_del=6, e_del=1, o_ins=6, e_ins=1, zdrop=100, end_bonus=5, a=1, b=4, mat=None):
    p = Params()
    p.o_del, p.e_del, p.o_ins, p.e_ins, p.zdrop, p.end_bonus = o_del, e_del, o_ins, e_ins, zdrop, end_bonus
    m = fill_scmat(a, b) if mat is None else np.asarray(mat, dtype=np.int8).reshape(-1)
    for i in range(25):
        p.mat[i] = int(m[i])
    return p


class BswCfg(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_reads", C.c_int32), ("read_len", C.c_int32),
                ("seed_min", C.c_int32), ("seed_max", C.c_int32), ("div", C.c_double), ("p_n", C.c_double),
                ("a", C.c_int32), ("o_del", C.c_int32), ("e_del", C.c_int32), ("o_ins", C.c_int32),
                ("e_ins", C.c_int32), ("w", C.c_int32)]


class KswvCfg(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_pairs", C.c_int32), ("qlen", C.c_int32), ("tmin", C.c_int32),
                ("tmax", C.c_int32), ("p_contain", C.c_double), ("div", C.c_double),
                ("indel_frac", C.c_double), ("p_n", C.c_double), ("xtra", C.c_int32)]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle.c", "workload.c", "bsw_avx2.c", "global.c", "striped.c", "oracle.h", "Makefile")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = build()
        try:
            _LIB = C.CDLL(so)
        except OSError:
            _LIB = C.CDLL(build(force=True))
        L = _LIB
        vp = C.c_void_p
        L.ora_bsw_extend.restype = C.c_int
        L.ora_bsw_extend.argtypes = [C.c_int, vp, C.c_int, vp, C.POINTER(Params), C.c_int, C.c_int,
                                     vp, vp, vp, vp, vp, vp]
        L.ora_bsw_batch.restype = C.c_int64
        L.ora_bsw_batch.argtypes = [vp, vp, vp, C.c_int32, C.POINTER(Params), C.c_int32, C.c_int, vp]
        L.ora_bsw_batch_avx2.restype = C.c_int64
        L.ora_bsw_batch_avx2.argtypes = [vp, vp, vp, C.c_int32, C.POINTER(Params), C.c_int32, C.c_int]
        L.ora_kswv_batch.restype = C.c_int64
        L.ora_kswv_batch.argtypes = [vp, vp, vp, vp, C.c_int32, C.POINTER(Params), C.c_int, C.c_int, C.c_int, vp]
        L.ora_ksw_align2.restype = None
        L.ora_ksw_align2.argtypes = [C.c_int, vp, C.c_int, vp, C.POINTER(Params), C.c_int, vp, vp]
        L.ora_kswv_batch_avx2.restype = C.c_int64
        L.ora_kswv_batch_avx2.argtypes = [vp, vp, vp, vp, C.c_int32, C.POINTER(Params), C.c_int, C.c_int, C.c_int, vp]
        L.ora_kswv_batch_striped.restype = None
        L.ora_kswv_batch_striped.argtypes = [vp, vp, vp, vp, C.c_int32, C.POINTER(Params), C.c_int]
        L.ora_ksw_global2.restype = C.c_int
        L.ora_ksw_global2.argtypes = [C.c_int, vp, C.c_int, vp, C.POINTER(Params), C.c_int, vp, vp, vp]
        L.ora_global2_batch.restype = C.c_int64
        L.ora_global2_batch.argtypes = [vp, vp, vp, C.c_int32, C.POINTER(Params), vp, C.c_int64, C.c_int]
        L.wl_global_fill.restype = None
        L.wl_global_fill.argtypes = [C.c_uint64, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_int32] + [C.c_int32] * 5 + [vp] * 3
        L.wl_bsw_sizes.restype = None
        L.wl_bsw_sizes.argtypes = [C.POINTER(BswCfg), vp, vp]
        L.wl_bsw_fill.restype = None
        L.wl_bsw_fill.argtypes = [C.POINTER(BswCfg)] + [vp] * 12 + [C.c_int]
        L.wl_kswv_fill.restype = None
        L.wl_kswv_fill.argtypes = [C.POINTER(KswvCfg), vp, vp, vp]
        L.wl_random_pairs.restype = None
        L.wl_random_pairs.argtypes = [C.c_uint64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_double, C.c_double, vp, vp, vp]
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def bsw_batch(pairs, ref, qer, params, w=100, nthreads=1, want_cells=False):
    """Runs the oracle over pairs (in place). Returns total cells (and per-pair cells)."""
    assert pairs.dtype == SEQPAIR and pairs.flags.c_contiguous
    cells = np.zeros(len(pairs), dtype=np.int64) if want_cells else None
    tot = lib().ora_bsw_batch(_p(pairs), _p(ref), _p(qer), len(pairs), C.byref(params), w, nthreads, _p(cells))
    return (tot, cells) if want_cells else tot


def bsw_batch_avx2(pairs, ref, qer, params, w=100, nthreads=1):
    """Inter-sequence AVX2 port (bsw_avx2.c: 16 pairs in the 16-bit lanes of a register, as the reference's
    smithWaterman256_16 organises them); same contract and results as bsw_batch.  Returns total cells."""
    assert pairs.dtype == SEQPAIR and pairs.flags.c_contiguous
    return lib().ora_bsw_batch_avx2(_p(pairs), _p(ref), _p(qer), len(pairs), C.byref(params), w, nthreads)


def kswv_batch(pairs, ref, qer, aln, params, is8, phase, nthreads=1, want_cells=False):
    assert pairs.dtype == SEQPAIR and aln.dtype == KSWR
    cells = np.zeros(len(pairs), dtype=np.int64) if want_cells else None
    tot = lib().ora_kswv_batch(_p(pairs), _p(ref), _p(qer), _p(aln), len(pairs), C.byref(params),
                               int(is8), int(phase), nthreads, _p(cells))
    return (tot, cells) if want_cells else tot


def kswv_batch_avx2(pairs, ref, qer, aln, params, is8, phase, nthreads=1, want_cells=False):
    """The inter-sequence AVX2 port (oracle/kswv_avx2.c): same contract and results as kswv_batch."""
    assert pairs.dtype == SEQPAIR and aln.dtype == KSWR
    cells = np.zeros(len(pairs), dtype=np.int64) if want_cells else None
    tot = lib().ora_kswv_batch_avx2(_p(pairs), _p(ref), _p(qer), _p(aln), len(pairs), C.byref(params),
                                    int(is8), int(phase), nthreads, _p(cells))
    return (tot, cells) if want_cells else tot


def kswv_batch_striped(pairs, ref, qer, aln, params, is8):
    """Phase 0 through the lane-by-lane emulation of bwa's striped ksw_u8 / ksw_i16 (oracle/striped.c; analysis only)."""
    assert pairs.dtype == SEQPAIR and aln.dtype == KSWR
    lib().ora_kswv_batch_striped(_p(pairs), _p(ref), _p(qer), _p(aln), len(pairs), C.byref(params), int(is8))


def ksw_align2(query, target, params, xtra):
    q = np.ascontiguousarray(query, dtype=np.uint8)
    t = np.ascontiguousarray(target, dtype=np.uint8)
    r = np.zeros(1, dtype=KSWR)
    c = np.zeros(1, dtype=np.int64)
    lib().ora_ksw_align2(len(q), _p(q), len(t), _p(t), C.byref(params), xtra, _p(r), _p(c))
    return r[0], int(c[0])


def bsw_extend(query, target, params, w, h0):
    q = np.ascontiguousarray(query, dtype=np.uint8)
    t = np.ascontiguousarray(target, dtype=np.uint8)
    out = np.zeros(5, dtype=np.int32)
    cells = np.zeros(1, dtype=np.int64)
    ptrs = [C.c_void_p(out.ctypes.data + 4 * i) for i in range(5)]
    sc = lib().ora_bsw_extend(len(q), _p(q), len(t), _p(t), C.byref(params), w, h0, *ptrs, _p(cells))
    return dict(score=sc, qle=int(out[0]), tle=int(out[1]), gtle=int(out[2]), gscore=int(out[3]),
                max_off=int(out[4]), cells=int(cells[0]))


# --------------------------------------------------------------------- workloads

def gen_bsw(n_reads, read_len=150, seed=0xC1, div=0.01, p_n=1e-4, seed_min=19, seed_max=100,
            a=1, o_del=6, e_del=1, o_ins=6, e_ins=1, w=100, nthreads=8):
    """C1/C2/C3-shaped extension pairs. Returns dict with
    left/right = (pairs, ref, qer).  right.regid = index of its left partner (or -1);
    right.h0 must be patched to the partner's score by the caller (chain_h0)."""
    cfg = BswCfg(seed, n_reads, read_len, seed_min, seed_max, div, p_n, a, o_del, e_del, o_ins, e_ins, w)
    len2 = np.zeros(2 * n_reads, dtype=np.int32)
    len1 = np.zeros(2 * n_reads, dtype=np.int32)
    lib().wl_bsw_sizes(C.byref(cfg), _p(len2), _p(len1))
    out = {}
    slots, offs, bufs, arrs = [], [], [], []
    for side in (0, 1):
        l2, l1 = len2[side::2], len1[side::2]
        emit = l2 > 0
        slot = np.where(emit, np.cumsum(emit) - 1, -1).astype(np.int64)
        l2e, l1e = l2[emit].astype(np.int64), l1[emit].astype(np.int64)
        off_q = np.concatenate(([0], np.cumsum(l2e)[:-1])).astype(np.int64) if len(l2e) else np.zeros(0, np.int64)
        off_r = np.concatenate(([0], np.cumsum(l1e)[:-1])).astype(np.int64) if len(l1e) else np.zeros(0, np.int64)
        qer = np.zeros(int(l2e.sum()) + 64, dtype=np.uint8)
        ref = np.zeros(int(l1e.sum()) + 64, dtype=np.uint8)
        pairs = np.zeros(int(emit.sum()), dtype=SEQPAIR)
        slots.append(slot); offs.append((off_r, off_q)); bufs.append((ref, qer)); arrs.append(pairs)
    lib().wl_bsw_fill(C.byref(cfg), _p(arrs[0]), _p(arrs[1]), _p(slots[0]), _p(slots[1]),
                      _p(offs[0][0]), _p(offs[0][1]), _p(offs[1][0]), _p(offs[1][1]),
                      _p(bufs[0][0]), _p(bufs[0][1]), _p(bufs[1][0]), _p(bufs[1][1]), nthreads)
    out["left"] = (arrs[0], bufs[0][0], bufs[0][1])
    out["right"] = (arrs[1], bufs[1][0], bufs[1][1])
    return out


def chain_h0(left_pairs, right_pairs):
    """right-extension h0 = score of the read's left extension."""
    has = right_pairs["regid"] >= 0
    right_pairs["h0"][has] = left_pairs["score"][right_pairs["regid"][has]]


def classify(pairs, a=1):
    """Upstream tier rule (SURVEY.md A.1 note 2): 0 = 8-bit, 1 = 16-bit, 2 = scalar."""
    mn = pairs["h0"] + np.minimum(pairs["len1"], pairs["len2"]) * a
    t8 = (pairs["len1"] < 128) & (pairs["len2"] < 128) & (mn < 128)
    t16 = (pairs["len1"] < 32768) & (pairs["len2"] < 32768) & (mn < 32768)
    return np.where(t8, 0, np.where(t16, 1, 2)).astype(np.int8)


def gen_kswv(n_pairs, qlen=150, tmin=300, tmax=500, seed=0xC4, p_contain=0.7, div=0.02,
             indel_frac=0.1, p_n=1e-4, xtra=None, min_seed_len=19, a=1):
    if xtra is None:
        xtra = KSW_XSUBO | KSW_XSTART | (KSW_XBYTE if qlen * a < 250 else 0) | (min_seed_len * a)
    cfg = KswvCfg(seed, n_pairs, qlen, tmin, tmax, p_contain, div, indel_frac, p_n, xtra)
    pairs = np.zeros(n_pairs, dtype=SEQPAIR)
    ref = np.zeros(n_pairs * tmax + 64, dtype=np.uint8)
    qer = np.zeros(n_pairs * qlen + 64, dtype=np.uint8)
    lib().wl_kswv_fill(C.byref(cfg), _p(pairs), _p(ref), _p(qer))
    return pairs, ref, qer


def gen_random(n, qmin, qmax, tmin, tmax, h0min=1, h0max=60, sim=0.8, p_n=0.01, seed=7):
    pairs = np.zeros(n, dtype=SEQPAIR)
    ref = np.zeros(n * max(tmax, 1) + 64, dtype=np.uint8)
    qer = np.zeros(n * max(qmax, 1) + 64, dtype=np.uint8)
    lib().wl_random_pairs(seed, n, qmin, qmax, tmin, tmax, h0min, h0max, sim, p_n, _p(pairs), _p(ref), _p(qer))
    return pairs, ref, qer


def kswv_phase1_inputs(pairs, ref, qer, aln, a=1):
    """Host hand-off of mem_sam_pe_batch: for hits passing
    XSTART && !(XSUBO && score < minsc) build reversed prefixes, h0 = KSW_XSTOP|score,
    regid = index of the aln entry.  Returns (pairs1, ref1, qer1)."""
    xtra = pairs["h0"]
    ok = ((xtra & KSW_XSTART) != 0) & ~(((xtra & KSW_XSUBO) != 0) & (aln["score"] < (xtra & 0xffff)))
    idx = np.nonzero(ok)[0]
    n = len(idx)
    p1 = np.zeros(n, dtype=SEQPAIR)
    l1 = (aln["te"][idx] + 1).astype(np.int64)
    l2 = (aln["qe"][idx] + 1).astype(np.int64)
    off_r = np.concatenate(([0], np.cumsum(l1)[:-1])).astype(np.int64) if n else np.zeros(0, np.int64)
    off_q = np.concatenate(([0], np.cumsum(l2)[:-1])).astype(np.int64) if n else np.zeros(0, np.int64)
    ref1 = np.zeros(int(l1.sum()) + 64, dtype=np.uint8)
    qer1 = np.zeros(int(l2.sum()) + 64, dtype=np.uint8)
    for k, i in enumerate(idx):
        r0, q0 = int(pairs["idr"][i]), int(pairs["idq"][i])
        ref1[off_r[k]:off_r[k] + l1[k]] = ref[r0:r0 + l1[k]][::-1]
        qer1[off_q[k]:off_q[k] + l2[k]] = qer[q0:q0 + l2[k]][::-1]
    p1["idr"], p1["idq"], p1["id"] = off_r, off_q, np.arange(n)
    p1["len1"], p1["len2"] = l1, l2
    p1["h0"] = KSW_XSTOP | aln["score"][idx]
    p1["seqid"], p1["regid"] = idx, idx
    return p1, ref1, qer1


# --------------------------------------------------------------------- CIGAR generation

def global2_batch(jobs, ref, qer, params, nthreads=8):
    """ksw_global2 over jobs (in place: score, n_cigar, cigar_off).  Returns (dense cigar array, cells)."""
    assert jobs.dtype == GLOBALJOB and jobs.flags.c_contiguous
    cap = int((jobs["tlen"].astype(np.int64) + jobs["qlen"]).sum()) + 16
    cig = np.zeros(cap, dtype=np.uint32)
    cells = lib().ora_global2_batch(_p(jobs), _p(ref), _p(qer), len(jobs), C.byref(params), _p(cig), cap, nthreads)
    assert cells >= 0
    tot = int(jobs["n_cigar"].astype(np.int64).sum())
    return cig[:tot].copy(), cells


def gen_global(n, qmin=30, qmax=150, div=0.03, p_n=1e-3, w_opt=100, seed=0xC6, a=1, o_del=6, e_del=1, o_ins=6, e_ins=1):
    """CIGAR jobs shaped like bwa_gen_cigar2's calls (oracle/workload.c wl_global_fill).  Returns (jobs, ref, qer)."""
    jobs = np.zeros(n, dtype=GLOBALJOB)
    ref = np.zeros(n * (2 * qmax + 8) + 64, dtype=np.uint8)
    qer = np.zeros(n * qmax + 64, dtype=np.uint8)
    lib().wl_global_fill(seed, n, qmin, qmax, div, p_n, w_opt, a, o_del, e_del, o_ins, e_ins, _p(jobs), _p(ref), _p(qer))
    return jobs, ref, qer


def cigar_string(ops):
    return "".join(f"{int(o) >> 4}{'MID'[int(o) & 15]}" for o in ops)


# --------------------------------------------------------------------- caller protocol on the oracle

def protocol_side(pairs, ref, qer, params, w=100, tries=2, nthreads=8, runner=None):
    """One side of mem_chain2aln_across_reads_V2's band loop: attempt 0 with band w on all pairs, then (MAX_BAND_TRY = 2)
    the pairs with score != score before the attempt and max_off >= (w >> 1) + (w >> 2) again with 2w.  `runner(sub, w)`
    computes a sub-array in place (default: the oracle).  Returns (results, pairs submitted over all attempts, cells)."""
    out = pairs.copy()
    submitted, cells = 0, 0
    prev = None
    for attempt in range(tries):
        wa = w << attempt
        if attempt == 0:
            todo = np.arange(len(out))
        else:
            wp = w << (attempt - 1)
            todo = np.nonzero((out["score"] != prev) & (out["max_off"] >= (wp >> 1) + (wp >> 2)))[0]
            if len(todo) == 0:
                break
        prev = out["score"].copy()
        sub = np.ascontiguousarray(out[todo])
        if runner is None:
            cells += bsw_batch(sub, ref, qer, params, wa, nthreads)
        else:
            runner(sub, wa)
        out[todo] = sub
        submitted += len(todo)
    return out, submitted, cells


def kswv_align2_batch(pairs, ref, qer, params, nthreads=8, simd=False):
    """mem_sam_pe_batch's sequence on the oracle: phase 0 by precision, hand-off, phase 1.  Returns (aln, cells).
    simd: the passes on the inter-sequence AVX2 port (identical results)."""
    kswv_batch = kswv_batch_avx2 if simd else globals()["kswv_batch"]
    aln = np.zeros(len(pairs), KSWR)
    is8 = (pairs["h0"] & KSW_XBYTE) != 0
    cells = 0
    for sel, flag in ((np.nonzero(is8)[0], 1), (np.nonzero(~is8)[0], 0)):
        if len(sel) == 0:
            continue
        sub = np.ascontiguousarray(pairs[sel]); a = np.zeros(len(sel), KSWR)
        cells += kswv_batch(sub, ref, qer, a, params, flag, 0, nthreads)
        p1, r1, q1 = kswv_phase1_inputs(sub, ref, qer, a)
        if len(p1):
            cells += kswv_batch(p1, r1, q1, a, params, flag, 1, nthreads)
        aln[sel] = a
    return aln, cells


STREAM_MAGIC = 0x35434d4145525453


def write_stream_workload(path, mode, L, R, lref, lqer, rref, rqer, ltask, rtask):
    """Binary workload of tools/stream_c5 (mode 0: BSW left/right extension tasks, 1: kswv tasks in L)."""
    nT = len(ltask) - 1
    hdr = np.array([STREAM_MAGIC, mode, len(L), len(R), nT, len(lref), len(lqer), len(rref), len(rqer)], dtype=np.int64)
    with open(path, "wb") as f:
        for a in (hdr, np.asarray(ltask, np.int64), np.asarray(rtask, np.int64), L, R, lref, lqer, rref, rqer):
            f.write(np.ascontiguousarray(a).tobytes())


def task_bounds(seqid, n_reads, task_reads):
    """Pair-index boundaries of the tasks of `task_reads` consecutive reads (pairs are in read order)."""
    edges = list(range(0, n_reads, task_reads)) + [n_reads]
    return np.searchsorted(seqid, edges, side="left").astype(np.int64)


# --------------------------------------------------------------------- reference-based targets

def pac_pack(bases):
    """bwa .pac layout: base x in byte x >> 2, bits ((~x & 3) << 1)."""
    b = np.asarray(bases, dtype=np.uint8)
    pad = (-len(b)) % 4
    if pad:
        b = np.concatenate([b, np.zeros(pad, np.uint8)])
    q = b.reshape(-1, 4)
    return ((q[:, 0] << 6) | (q[:, 1] << 4) | (q[:, 2] << 2) | q[:, 3]).astype(np.uint8)


def ref_slice(fwd, rpos, length):
    """Target bases of one pair from the forward strand `fwd` (values 0..3) in bwa's doubled coordinates:
    rpos >= 0 reads upwards from rpos, rpos < 0 downwards from ~rpos; x >= l_pac is the reverse complement."""
    l_pac = len(fwd)
    k = np.arange(length, dtype=np.int64)
    x = (~rpos - k) if rpos < 0 else (rpos + k)
    rc = x >= l_pac
    f = np.where(rc, 2 * l_pac - 1 - x, x)
    b = fwd[f]
    return np.where(rc, 3 - b, b).astype(np.uint8)
