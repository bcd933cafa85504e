"""CPU check of the CUDA kernels' per-thread DP (bsw_core.cuh / kswv_core.cuh compiled for
the host with emulated DPX ops) against the oracle.  This validates the packed-lane logic
(masks, band trimming, key packing) in a container without a GPU; the real parity tests,
through the C ABI on a B200, are in test_gpu_parity.py."""
import numpy as np
import pytest

import common
import emu


def _check_bsw(ora, pairs, ref, qer, **kw):
    prm = ora.make_params(**{k: v for k, v in kw.items() if k != "w"})
    o = pairs.copy()
    _, oc = ora.bsw_batch(o, ref, qer, prm, kw.get("w", 100), 4, True)
    for fg, count in ((0, True), (0, False), (1, True), (2, False)):      # 2: packed path, general (N-capable) selectors
        e, ec, _ = emu.bsw(pairs, ref, qer, force_general=fg, count=count, **kw)
        bad = common.diff_fields(o, e, common.BSW_FIELDS)
        if count:
            bad |= oc != ec
        assert not bad.any(), common.describe_first(o, e, bad)
    # the library sorts the pairs by (len2, len1, h0) before it deals two of them to a thread: same order here, so that
    # the two lanes of the emulated thread are alike (the fast full-group sweep) instead of arbitrary neighbours
    order = np.lexsort((pairs["h0"], pairs["len1"] >> 1, pairs["len2"]))[::-1]
    ps = np.ascontiguousarray(pairs[order])
    e, _, _ = emu.bsw(ps, ref, qer, force_general=0, count=False, **kw)
    bad = common.diff_fields(o[order], e, common.BSW_FIELDS)
    assert not bad.any(), common.describe_first(o[order], e, bad)


def test_bsw_edge_cases(ora):
    p, r, q = common.edge_case_pairs()
    for w in (100, 7, 1):
        _check_bsw(ora, p, r, q, w=w)


def test_bsw_c1_shape(ora):
    (L, lr, lq), (R, rr, rq) = common.bsw_c1(4000)
    _check_bsw(ora, L, lr, lq)
    _check_bsw(ora, R, rr, rq)


def test_bsw_c3_shape_zdrop(ora):
    d = ora.gen_bsw(1500, read_len=250, seed=0xC3, div=0.05)
    L, lr, lq = d["left"]
    _check_bsw(ora, L, lr, lq)
    _check_bsw(ora, L, lr, lq, w=10, zdrop=20)


@pytest.mark.parametrize("cfg", [(0, 40, 0, 60, 0.8, 0.02), (1, 200, 1, 300, 0.6, 0.05),
                                 (50, 255, 50, 400, 0.9, 0.001), (1, 30, 1, 30, 0.3, 0.2)])
def test_bsw_random(ora, cfg):
    q0, q1, t0, t1, sim, pn = cfg
    P, r, q = ora.gen_random(1500, q0, q1, t0, t1, h0min=1, h0max=50, sim=sim, p_n=pn, seed=q1)
    _check_bsw(ora, P, r, q)
    _check_bsw(ora, P, r, q, w=5, zdrop=10)
    _check_bsw(ora, P, r, q, w=1, zdrop=0, end_bonus=0)


def test_bsw_scoring_variants(ora):
    P, r, q = ora.gen_random(1500, 1, 120, 1, 200, h0min=1, h0max=60, sim=0.85, p_n=0.02, seed=31)
    for (od, ed, oi, ei, zd, eb, a, b) in ((4, 2, 7, 1, 30, 0, 2, 3), (1, 1, 1, 1, 0, 9, 1, 9), (12, 3, 5, 2, 50, 5, 1, 1), (5, 2, 5, 2, 40, 3, 3, 2)):
        _check_bsw(ora, P, r, q, w=37, o_del=od, e_del=ed, o_ins=oi, e_ins=ei, zdrop=zd, end_bonus=eb, a=a, b=b)


def _check_kswv(ora, pairs, ref, qer, is8):
    prm = ora.make_params()
    for fg in (0, 1):
        ao = np.zeros(len(pairs), ora.KSWR)
        ae = np.zeros(len(pairs), ora.KSWR)
        _, oc = ora.kswv_batch(pairs, ref, qer, ao, prm, is8, 0, 4, True)
        ec, _ = emu.kswv(pairs, ref, qer, ae, is8, 0, force_general=fg)
        assert (ao == ae).all() and (oc == ec).all()
        p1, r1, q1 = ora.kswv_phase1_inputs(pairs, ref, qer, ao)
        ora.kswv_batch(p1, r1, q1, ao, prm, is8, 1, 4)
        emu.kswv(p1, r1, q1, ae, is8, 1, force_general=fg)
        assert (ao == ae).all()


def _check_kswv_sys(ora, pairs, ref, qer, is8, mode=0, **sc):
    """kswv_sys.cuh (DP state in registers, rows skewed over the 8 lanes of a group) on an emulated warp."""
    prm = ora.make_params(**sc) if sc else ora.make_params()
    ao = np.zeros(len(pairs), ora.KSWR)
    ae = np.zeros(len(pairs), ora.KSWR)
    _, oc = ora.kswv_batch(pairs, ref, qer, ao, prm, is8, 0, 4, True)
    ec, npk = emu.kswv(pairs, ref, qer, ae, is8, 0, force_general=mode, **sc)
    assert (ao == ae).all() and (oc == ec).all()
    p1, r1, q1 = ora.kswv_phase1_inputs(pairs, ref, qer, ao)
    ora.kswv_batch(p1, r1, q1, ao, prm, is8, 1, 4)
    emu.kswv(p1, r1, q1, ae, is8, 1, force_general=mode, **sc)
    assert (ao == ae).all()
    return npk


def test_kswv_sys_kernel(ora):
    P, r, q = ora.gen_kswv(240)
    assert _check_kswv_sys(ora, P, r, q, 1) == len(P)
    _check_kswv_sys(ora, P, r, q, 1, mode=1 << 16)          # every group through the general score table (QN)
    _check_kswv_sys(ora, P, r, q, 1, mode=26 << 8)          # a larger shape than needed: more pad columns
    _check_kswv_sys(ora, P, r, q, 1, mode=1 << 17)          # 16 lanes per group (small-batch shape)
    _check_kswv_sys(ora, P, r, q, 1, mode=2 << 17)          # 32 lanes per group
    P, r, q = ora.gen_kswv(100, qlen=250, tmin=400, tmax=700, seed=5)
    _check_kswv_sys(ora, P, r, q, 0)
    # ragged lengths (neighbouring pairs of one group differ), empty sequences, N in queries and targets, early stops
    P, r, q = ora.gen_random(800, 0, 40, 0, 80, sim=0.7, p_n=0.05, seed=3)
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 5
    _check_kswv_sys(ora, P, r, q, 1)
    P, r, q = ora.gen_random(400, 0, 200, 0, 300, sim=0.8, p_n=0.002, seed=4)
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | 9
    _check_kswv_sys(ora, P, r, q, 0)
    _check_kswv_sys(ora, P, r, q, 1)
    _check_kswv_sys(ora, P, r, q, 0, mode=2 << 17)
    _check_kswv_sys(ora, P, r, q, 1, mode=1 << 17)
    # 8-bit overflow rule, tandem repeats (second-best list), scoring variants (asymmetric gap sums: the non-SYM variants)
    P, r, q = ora.gen_kswv(80, qlen=255, tmin=300, tmax=400, seed=9, div=0.0,
                           xtra=ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 19)
    _check_kswv_sys(ora, P, r, q, 1)
    P, r, q = ora.gen_kswv(200, qlen=60, tmin=200, tmax=400, seed=11, p_contain=0.9, div=0.05)
    for i in range(len(P)):
        o, l = int(P["idr"][i]), int(P["len1"][i])
        h = l // 2
        r[o + h:o + 2 * h] = r[o:o + h]
    _check_kswv_sys(ora, P, r, q, 1)
    P, r, q = ora.gen_random(300, 0, 120, 0, 300, sim=0.85, p_n=0.004, seed=14)
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | 9
    for sc in (dict(o_del=4, e_del=2, o_ins=7, e_ins=1, a=2, b=3), dict(o_del=1, e_del=1, o_ins=1, e_ins=1, a=1, b=9),
               dict(o_del=12, e_del=3, o_ins=5, e_ins=2, a=1, b=1)):
        _check_kswv_sys(ora, P, r, q, 0, **sc)


def test_kswv_sys_16bit_tier(ora):
    """The 16-bit shapes of kswv_sys.cuh (W: key = max << 5 | column within the lane, the lane beside it): scores up to 2047,
    queries up to 512 bases, 8 and 16 lanes per group; and every byte-key pair forced through them."""
    P, r, q = ora.gen_kswv(120, qlen=300, tmin=400, tmax=700, seed=21)                   # 2 x 300 bp: 16 lanes x 20 columns
    assert _check_kswv_sys(ora, P, r, q, 0) == len(P)
    P, r, q = ora.gen_kswv(60, qlen=500, tmin=600, tmax=900, seed=22, p_n=0.002)          # 16 x 32
    assert _check_kswv_sys(ora, P, r, q, 0) == len(P)
    P, r, q = ora.gen_kswv(160, qlen=150, tmin=300, tmax=500, seed=23, p_n=0.002)         # a = 2: scores to 300 on 150 columns
    assert _check_kswv_sys(ora, P, r, q, 0, o_del=6, e_del=1, o_ins=6, e_ins=1, a=2, b=4) == len(P)
    _check_kswv_sys(ora, P, r, q, 0, o_del=4, e_del=2, o_ins=7, e_ins=1, a=3, b=3)        # asymmetric gap sums (non-SYM)
    P, r, q = ora.gen_random(500, 0, 330, 0, 500, sim=0.85, p_n=0.004, seed=24)           # ragged, empty, every configuration
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | 9
    _check_kswv_sys(ora, P, r, q, 0, mode=1 << 19)
    # the 8-bit entry point keeps ksw_u8's saturation rule: nothing of it may reach the 16-bit shapes
    P, r, q = ora.gen_kswv(60, qlen=300, tmin=400, tmax=500, seed=25, div=0.0, xtra=ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 19)
    assert _check_kswv_sys(ora, P, r, q, 1) == 0


def test_kswv_c4_shape(ora):
    P, r, q = ora.gen_kswv(1200)
    _check_kswv(ora, P, r, q, 1)
    P, r, q = ora.gen_kswv(400, qlen=250, tmin=400, tmax=700, seed=5)
    _check_kswv(ora, P, r, q, 0)


def test_kswv_overflow_and_tandem(ora):
    P, r, q = ora.gen_kswv(200, qlen=255, tmin=300, tmax=400, seed=9, div=0.0,
                           xtra=ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 19)
    _check_kswv(ora, P, r, q, 1)
    P, r, q = ora.gen_kswv(800, qlen=60, tmin=200, tmax=400, seed=11, p_contain=0.9, div=0.05)
    for i in range(len(P)):
        o, l = int(P["idr"][i]), int(P["len1"][i])
        h = l // 2
        r[o + h:o + 2 * h] = r[o:o + h]
    _check_kswv(ora, P, r, q, 1)
    P, r, q = ora.gen_random(1000, 0, 40, 0, 80, sim=0.7, p_n=0.05, seed=3)
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 5
    _check_kswv(ora, P, r, q, 1)


def test_bsw_count_tiny_queries_cheap_gaps(ora):
    """the cell-counting variant once failed for a short or unclean lane next to a longer query (tiny
    queries, h0 up to 250, cheap or free gap opens, unsorted pairing).  The counting variant is now the same row loop
    (bsw_core2.cuh, COUNT) and must agree on results AND cells."""
    for seed, (q1, t1) in enumerate(((12, 130), (5, 40), (12, 12))):
        P, r, q = ora.gen_random(1200, 1, q1, 1, t1, h0min=1, h0max=250, sim=0.7, p_n=0.02, seed=100 + seed)
        _check_bsw(ora, P, r, q, w=100, a=3, b=5, o_del=2, e_del=2, o_ins=3, e_ins=1)
        _check_bsw(ora, P, r, q, w=2, o_del=0, o_ins=0)
        _check_bsw(ora, P, r, q, w=100, zdrop=0, o_del=0, e_del=1, o_ins=0, e_ins=3, a=2, b=1)


def _check_wide(ora, pairs, ref, qer, **kw):
    prm = ora.make_params(**{k: v for k, v in kw.items() if k not in ("w", "min_cfg", "latency")})
    o = pairs.copy()
    ora.bsw_batch(o, ref, qer, prm, kw.get("w", 100), 4)
    e, taken = emu.bsw_wide(pairs, ref, qer, **kw)
    bad = common.diff_fields(o, e, common.BSW_FIELDS) & taken
    assert not bad.any(), common.describe_first(o, e, bad)
    return taken


def test_bsw_wide_lanes_per_pair(ora):
    """bsw_wide.cuh (G lanes per pair, 16-bit register state, prefix-max f) on an emulated warp: edge cases, bands, z-drop,
    scoring variants, unclean initial rows (h0 up to 250 against short bands), forced larger configurations."""
    p, r, q = common.edge_case_pairs()
    for w in (100, 7, 1):
        assert _check_wide(ora, p, r, q, w=w).all()
    P, r, q = ora.gen_random(160, 0, 140, 0, 300, h0min=1, h0max=250, sim=0.85, p_n=0.02, seed=11)
    assert _check_wide(ora, P, r, q).all()
    _check_wide(ora, P, r, q, w=5, zdrop=10)
    _check_wide(ora, P, r, q, w=1, zdrop=0, end_bonus=0)
    _check_wide(ora, P, r, q, w=30, min_cfg=2)
    _check_wide(ora, P, r, q, w=60, min_cfg=3)
    _check_wide(ora, P, r, q, w=3, zdrop=12, min_cfg=5)
    _check_wide(ora, P, r, q, w=37, o_del=4, e_del=2, o_ins=7, e_ins=1, zdrop=30, end_bonus=0, a=2, b=3)
    _check_wide(ora, P, r, q, w=2, o_del=0, o_ins=0)
    # the shapes small batches run on: (8, 5) for up to 80, (16, 5) for up to 160 and (32, 5) for up to 320 columns
    assert _check_wide(ora, P, r, q, latency=True).all()
    _check_wide(ora, P, r, q, w=9, zdrop=15, latency=True)
    P, r, q = ora.gen_random(60, 100, 300, 100, 400, h0min=1, h0max=200, sim=0.9, p_n=0.01, seed=15)
    assert _check_wide(ora, P, r, q, latency=True).all()
    _check_wide(ora, P, r, q, w=33, o_del=4, e_del=2, o_ins=7, e_ins=1, zdrop=30, a=2, b=3, latency=True)


def test_bsw_wide_16bit_tier(ora):
    """Scores > 255 and queries up to 1023 columns (row a4): every configuration, plus C3-shaped pairs with z-drop."""
    P, r, q = ora.gen_random(50, 200, 600, 200, 900, h0min=100, h0max=600, sim=0.95, p_n=0.002, seed=12)
    assert _check_wide(ora, P, r, q).all()
    _check_wide(ora, P, r, q, w=20, zdrop=50)
    P, r, q = ora.gen_random(12, 900, 1023, 900, 1400, h0min=10, h0max=900, sim=0.97, p_n=0.001, seed=13)
    assert _check_wide(ora, P, r, q).all()
    d = ora.gen_bsw(100, read_len=250, seed=0xC3, div=0.05)
    L, lr, lq = d["left"]
    assert _check_wide(ora, L, lr, lq).all()
    _check_wide(ora, L, lr, lq, w=10, zdrop=20)
    P, r, q = ora.gen_random(4, 1100, 1200, 1100, 1200, h0min=10, h0max=20, sim=0.9, seed=14)       # too long: not admitted
    assert not _check_wide(ora, P, r, q).any()


def _global_cases(ora):
    """CIGAR jobs: bwa_gen_cigar2-shaped ones at three divergences, plus hand-made corners (empty sides, band 0, band
    equal to the length difference, all-N, a band wider than the query).  w >= |tlen - qlen| throughout: below that the
    reference's backtrack leaves its matrix (bwa_gen_cigar2 never asks for it)."""
    cases = [ora.gen_global(400, div=0.03, seed=0xC6), ora.gen_global(300, qmin=1, qmax=60, div=0.15, seed=0xC7),
             ora.gen_global(100, qmin=200, qmax=400, div=0.08, seed=0xC8, w_opt=100)]
    rng = np.random.default_rng(3)
    q = rng.integers(0, 4, 400, dtype=np.uint8); r = rng.integers(0, 4, 400, dtype=np.uint8)
    r[100:160] = q[100:160]; q[300:310] = 4
    J = np.zeros(9, dtype=ora.GLOBALJOB)
    rows = [(0, 0, 0, 0, 3), (0, 0, 10, 0, 10), (0, 0, 0, 10, 12), (100, 100, 60, 60, 0), (100, 100, 60, 40, 20), (100, 100, 30, 60, 31),
            (0, 0, 50, 50, 100), (300, 300, 10, 10, 4), (90, 100, 80, 60, 25)]
    for k, (a, b, tl, ql, w) in enumerate(rows):
        J[k]["idr"], J[k]["idq"], J[k]["tlen"], J[k]["qlen"], J[k]["w"] = a, b, tl, ql, w
    cases.append((J, r, q))
    return cases


def test_global2_cigar_generation(ora):
    """global_core.cuh (row f.4: ksw_global2 with backtrack) on the host with the library's scratch layout (interleaved
    rolling rows modulo 2 w_max + 2, packed direction words, count + emit passes) == the oracle: scores, op counts, ops."""
    for kw in (dict(), dict(o_del=4, e_del=2, o_ins=7, e_ins=1, a=2, b=3)):
        prm = ora.make_params(**kw)
        for J, r, q in _global_cases(ora):
            o = J.copy(); oc, _ = ora.global2_batch(o, r, q, prm, 4)
            g, gc = emu.global2(J, r, q, **kw)
            for f in ("score", "n_cigar", "cigar_off"):
                assert (o[f] == g[f]).all(), (f, np.nonzero(o[f] != g[f])[0][:5])
            assert len(oc) == len(gc) and (oc == gc).all()
            # every CIGAR consumes exactly its query and target
            for i in range(len(o)):
                ops = oc[o["cigar_off"][i]:o["cigar_off"][i] + o["n_cigar"][i]]
                assert sum(int(x) >> 4 for x in ops if int(x) & 15 in (0, 1)) == o["qlen"][i]
                assert sum(int(x) >> 4 for x in ops if int(x) & 15 in (0, 2)) == o["tlen"][i]
