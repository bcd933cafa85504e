"""Parity of the CUDA path (through the C ABI) with the CPU oracle -- bit-exact on every
result field.  Runs on the B200 box (`pytest -m gpu`); nothing here reads /root/reference."""
import json
import os

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


# Host-pointer calls take one of two routes (bsw_api.cu): below BSW_CUDA_DIRECT_MIN pairs the staged engine (merged
# batches, compact records), above it the chunked DMA path.  Every test that goes through these fixtures runs on both.
# Each route runs with both DP kernels: "wide" = lanes-per-pair 16-bit kernel for every pair it admits (the default for calls /
# batches of up to 32768 pairs), "thread" = thread-owns-two-pairs kernel (large calls), wide only for what it alone admits.
@pytest.fixture(scope="module", params=["staged-wide", "staged-thread", "direct-wide", "direct-thread"])
def sw(bsw, request):
    assert bsw.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    obj = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    route, kern = request.param.split("-")
    if route == "direct":
        obj.set_direct_min(0)
    obj.set_wide_max(1 << 30 if kern == "wide" else 0)
    obj.route = route
    yield obj
    obj.close()


@pytest.fixture(scope="module", params=["staged", "direct"])
def ks(bsw, request):
    obj = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    if request.param == "direct":
        obj.set_direct_min(0)
    obj.route = request.param
    yield obj
    obj.close()


def _oracle_bsw(ora, pairs, ref, qer, w=100, **kw):
    o = pairs.copy()
    _, cells = ora.bsw_batch(o, ref, qer, ora.make_params(**kw), w, 8, True)
    return o, cells


def _assert_same(o, g, extra=None):
    bad = common.diff_fields(o, g, common.BSW_FIELDS)
    assert not bad.any(), common.describe_first(o, g, bad, extra)
    for f in ("idr", "idq", "id", "len1", "len2", "h0", "seqid", "regid"):      # caller bookkeeping preserved
        assert (o[f] == g[f]).all(), f


def test_edge_cases(ora, sw):
    p, r, q = common.edge_case_pairs()
    for w in (100, 7, 1):
        o, _ = _oracle_bsw(ora, p, r, q, w)
        g = p.copy()
        sw.getScores16(g, r, q, len(g), 1, w)
        _assert_same(o, g)


def test_golden_fixtures(bsw, sw, ks):
    g = json.load(open(os.path.join(HERE, "golden", "bsw_golden.json")))
    for w in sorted({c["w"] for c in g["cases"]}):
        cases = [c for c in g["cases"] if c["w"] == w]
        pairs = np.zeros(len(cases), bsw.SEQPAIR)
        ref, qer = [], []
        ro = qo = 0
        for i, c in enumerate(cases):
            pairs[i]["idr"], pairs[i]["idq"], pairs[i]["len1"], pairs[i]["len2"], pairs[i]["h0"] = ro, qo, len(c["target"]), len(c["query"]), c["h0"]
            ref += c["target"]; qer += c["query"]
            ro += len(c["target"]); qo += len(c["query"])
        ref = np.array(ref + [0] * 16, np.uint8); qer = np.array(qer + [0] * 16, np.uint8)
        sw.getScores16(pairs, ref, qer, len(pairs), 1, w)
        for i, c in enumerate(cases):
            got = [int(pairs[i][k]) for k in ("score", "tle", "gtle", "qle", "gscore", "max_off")]
            assert got == c["expect"], c["name"]
    g = json.load(open(os.path.join(HERE, "golden", "kswv_golden.json")))
    for is8 in (1, 0):
        cases = [c for c in g["cases"] if bool(c["xtra"] & bsw.KSW_XBYTE) == bool(is8)]
        pairs = np.zeros(len(cases), bsw.SEQPAIR)
        ref, qer = [], []
        ro = qo = 0
        for i, c in enumerate(cases):
            pairs[i]["idr"], pairs[i]["idq"], pairs[i]["len1"], pairs[i]["len2"], pairs[i]["h0"] = ro, qo, len(c["target"]), len(c["query"]), c["xtra"]
            pairs[i]["regid"] = i
            ref += c["target"]; qer += c["query"]
            ro += len(c["target"]); qo += len(c["query"])
        ref = np.array(ref + [0] * 16, np.uint8); qer = np.array(qer + [0] * 16, np.uint8)
        aln = np.zeros(len(cases), bsw.KSWR)
        f = ks.getScores8 if is8 else ks.getScores16
        f(pairs, ref, qer, aln, len(pairs), 1, 0)
        import ora as _ora
        p1, r1, q1 = _ora.kswv_phase1_inputs(pairs, ref, qer, aln)
        if len(p1):
            f(p1, r1, q1, aln, len(p1), 1, 1)
        for i, c in enumerate(cases):
            assert [int(aln[i][k]) for k in common.KSW_FIELDS] == c["expect"], c["name"]


@pytest.mark.parametrize("n", [1, 15, 31, 32, 33, 50, 64, 100, 129])
def test_partial_batches(ora, sw, n):
    """The fork's partial-batch list (R:GRAVITON4_SVE2_RESULTS.md:398-413)."""
    P, r, q = ora.gen_random(n, 50, 127, 50, 127, h0min=1, h0max=60, sim=0.9, p_n=0.001, seed=n)
    o, _ = _oracle_bsw(ora, P, r, q)
    g = P.copy()
    sw.getScores8(g, r, q, n, 1, 100)
    _assert_same(o, g)


def test_c1_protocol(ora, sw):
    """C1-shaped trace through the caller protocol: tier routing,
    left before right with h0 chaining, band retry w then 2w."""
    d = ora.gen_bsw(20000, seed=0xC1)
    prm = ora.make_params()

    def run_side(pairs, ref, qer, dev):
        out = pairs.copy()
        for attempt in range(2):                         # MAX_BAND_TRY = 2
            w = 100 << attempt
            if attempt == 0:
                todo = np.arange(len(out))
            else:
                todo = np.nonzero((out["score"] != prev) & (out["max_off"] >= (w0 >> 1) + (w0 >> 2)))[0]
                if len(todo) == 0:
                    break
            prev = out["score"].copy()
            w0 = w
            sub = out[todo].copy()
            if dev:
                tiers = ora.classify(sub)
                for tier, fn in ((0, sw.getScores8), (1, sw.getScores16)):
                    sel = np.nonzero(tiers == tier)[0]
                    if len(sel):
                        part = np.ascontiguousarray(sub[sel])
                        fn(part, ref, qer, len(part), 1, w)
                        sub[sel] = part
            else:
                ora.bsw_batch(sub, ref, qer, prm, w, 8)
            out[todo] = sub
        return out

    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    Lo, Lg = run_side(L, lr, lq, False), run_side(L, lr, lq, True)
    _assert_same(Lo, Lg)
    Ro_in, Rg_in = R.copy(), R.copy()
    ora.chain_h0(Lo, Ro_in); ora.chain_h0(Lg, Rg_in)
    Ro, Rg = run_side(Ro_in, rr, rq, False), run_side(Rg_in, rr, rq, True)
    _assert_same(Ro, Rg)


def test_c3_long_reads_zdrop(ora, sw):
    """Config 3: 250 bp reads at 5% divergence -> getScores16 path, z-drop and zero-row exits."""
    d = ora.gen_bsw(8000, read_len=250, seed=0xC3, div=0.05, seed_max=60)
    for side in ("left",):
        P, r, q = d[side]
        for zdrop, w in ((100, 100), (20, 100), (20, 10)):
            o, cells = _oracle_bsw(ora, P, r, q, w, zdrop=zdrop)
            obj = sw if zdrop == 100 else None
            if obj is None:
                import bswb200
                obj = bswb200.BandedPairWiseSW(6, 1, 6, 1, zdrop, 5, None, 1, 4, 1)
            g = P.copy()
            obj.getScores16(g, r, q, len(g), 1, w)
            _assert_same(o, g)
            if obj is not sw:
                obj.close()
            full = (P["len1"].astype(np.int64) * P["len2"]).sum()
            if zdrop == 20:
                assert cells.sum() < 0.8 * full           # early exits really happen


def test_16bit_tier_and_general_path(ora, sw):
    """Pairs the byte-packed kernel does not admit (score bound > 255 or query > 252 columns) run on the 16-bit
    lanes-per-pair kernel up to 1023 columns / scores of 16383 -- 400 bp near-identical pairs, h0 up to 300 -- and
    only what is beyond that on the 32-bit kernel: a 1500 x 2000 pair."""
    P, r, q = ora.gen_random(300, 256, 400, 256, 500, h0min=100, h0max=300, sim=0.97, p_n=0.001, seed=99)
    o, _ = _oracle_bsw(ora, P, r, q)
    g = P.copy()
    sw.reset_stats()
    sw.getScores16(g, r, q, len(g), 1, 100)
    _assert_same(o, g)
    st = sw.stats()
    assert st["pairs_general"] == 0 and st["pairs_wide"] > 0     # the 16-bit tier stays on chip (row a4)
    P, r, q = ora.gen_random(3, 1500, 1500, 2000, 2000, h0min=40, h0max=40, sim=0.98, p_n=0.0, seed=5)
    o, _ = _oracle_bsw(ora, P, r, q)
    g = P.copy()
    sw.reset_stats()
    sw.scalarBandedSWAWrapper(g, r, q, len(g), 1, 100)
    _assert_same(o, g)
    assert sw.stats()["pairs_general"] == 3                       # beyond 1023 columns: the 32-bit kernel


def test_mixed_lengths_and_cells(ora, sw):
    """Ragged batch (all bins, odd counts) + the library's DP-cell counter equals the oracle's."""
    parts = [ora.gen_random(n, q0, q1, t0, t1, h0min=1, h0max=40, sim=s, p_n=0.01, seed=q1 + n)
             for n, q0, q1, t0, t1, s in ((701, 0, 20, 0, 50, 0.8), (333, 10, 127, 5, 200, 0.9),
                                          (97, 100, 200, 100, 300, 0.95), (1, 0, 0, 0, 0, 0.5), (45, 200, 255, 1, 600, 0.6))]
    P, r, q = common.concat_workloads(parts)
    o, cells = _oracle_bsw(ora, P, r, q)
    for counting in (True, False):           # the counting variant tracks the exact trimmed beg as well
        sw.set_cell_counting(counting)
        sw.reset_stats()
        g = P.copy()
        sw.getScores16(g, r, q, len(g), 1, 100)
        _assert_same(o, g)
        st = sw.stats()
        assert st["pairs"] == len(P)
        if counting:
            assert st["cells"] == int(cells.sum())
    sw.set_cell_counting(False)


def test_scoring_variants(ora, bsw):
    P, r, q = ora.gen_random(3000, 1, 120, 1, 200, h0min=1, h0max=60, sim=0.85, p_n=0.02, seed=31)
    for (od, ed, oi, ei, zd, eb, a, b) in ((6, 1, 6, 1, 100, 5, 1, 4), (4, 2, 7, 1, 30, 0, 2, 3), (1, 1, 1, 1, 0, 9, 1, 9), (12, 3, 5, 2, 50, 5, 1, 1)):
        prm = ora.make_params(o_del=od, e_del=ed, o_ins=oi, e_ins=ei, zdrop=zd, end_bonus=eb, a=a, b=b)
        o = P.copy()
        ora.bsw_batch(o, r, q, prm, 37, 8)
        obj = bsw.BandedPairWiseSW(od, ed, oi, ei, zd, eb, bsw.fill_scmat(a, b), a, b, 1)
        g = P.copy()
        obj.getScores16(g, r, q, len(g), 1, 37)
        obj.close()
        _assert_same(o, g)
        obj = bsw.BandedPairWiseSW(od, ed, oi, ei, zd, eb, bsw.fill_scmat(a, b), a, -b, 1)   # fork passes -b
        g2 = P.copy()
        obj.getScores16(g2, r, q, len(g2), 1, 37)
        obj.close()
        _assert_same(o, g2)


def test_kswv_c4(ora, ks, bsw):
    """Config 4: mate rescue, phase 0 (8-bit then 16-bit) and phase 1 (16 then 8), SURVEY.md 3.2."""
    prm = ora.make_params()
    P8, r8, q8 = ora.gen_kswv(20000)
    P16, r16, q16 = ora.gen_kswv(3000, qlen=250, tmin=400, tmax=700, seed=0xC5)
    for P, r, q, is8 in ((P8, r8, q8, 1), (P16, r16, q16, 0)):
        ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
        _, cells = ora.kswv_batch(P, r, q, ao, prm, is8, 0, 8, True)
        f = ks.getScores8 if is8 else ks.getScores16
        ks.reset_stats()
        f(P, r, q, ag, len(P), 1, 0)
        assert (ao == ag).all(), np.nonzero(ao != ag)[0][:5]
        if ks.route == "direct":                         # the handle's cell counter covers the direct / resident routes
            assert ks.stats()["cells"] == int(cells.sum())
        p1, r1, q1 = ora.kswv_phase1_inputs(P, r, q, ao)
        ora.kswv_batch(p1, r1, q1, ao, prm, is8, 1, 8)
        f(p1, r1, q1, ag, len(p1), 1, 1)
        assert (ao == ag).all()
        assert (ao["tb"] >= 0).mean() > 0.5


def test_kswv_dev_align2(ora, bsw):
    """kswv_dev_align2: forward pass, hand-off and reverse pass on the device in one submission equal
    ksw_align2 per pair -- 8-bit and 16-bit batches, hits below minsc (no second pass), queries with N, the 32-bit general
    kernel for long queries, a batch in which no hit qualifies, an empty batch."""
    import torch
    prm = ora.make_params()
    ks = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    st = torch.cuda.current_stream().cuda_stream
    cases = []
    P, r, q = ora.gen_kswv(20000); cases.append((P, r, q, 1))
    P, r, q = ora.gen_kswv(3000, qlen=250, tmin=400, tmax=700, seed=0xC5); cases.append((P, r, q, 0))
    P, r, q = ora.gen_random(3000, 0, 120, 0, 250, sim=0.75, p_n=0.03, seed=31)
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 30; cases.append((P, r, q, 1))
    P, r, q = ora.gen_kswv(200, qlen=400, tmin=600, tmax=900, seed=10, xtra=ora.KSW_XSUBO | ora.KSW_XSTART | 19); cases.append((P, r, q, 0))
    P, r, q = ora.gen_kswv(500, seed=77, xtra=ora.KSW_XSUBO | ora.KSW_XBYTE | 19); cases.append((P, r, q, 1))       # no KSW_XSTART
    for P, r, q, is8 in cases:
        flag = P["h0"].copy()
        if is8:
            assert ((flag & ora.KSW_XBYTE) != 0).all()
        ao, _ = ora.kswv_align2_batch(P, r, q, prm, 8)
        dp = torch.from_numpy(P.view(np.uint8).reshape(-1).copy()).cuda()
        dr, dq = torch.from_numpy(r).cuda(), torch.from_numpy(q).cuda()
        da = torch.zeros(len(P) * 28, dtype=torch.uint8, device="cuda")
        ks.dev_align2(dp, dr, dq, da, len(P), is8, st)
        torch.cuda.synchronize()
        ag = da.cpu().numpy().view(bsw.KSWR)
        assert (ao == ag).all(), np.nonzero(ao != ag)[0][:5]
        assert (dp.cpu().numpy().view(bsw.SEQPAIR) == P).all()          # inputs untouched
    ks.dev_align2(dp, dr, dq, da, 0, 1, st)
    ks.close()


def test_global2_cigar_generation(ora, bsw, monkeypatch):
    """Row f.4: bsw_global2 / bsw_dev_global2 (batched ksw_global2 with backtrack) == the oracle: score, n_cigar, cigar_off
    and every op -- bwa_gen_cigar2-shaped jobs at three divergences, two scoring sets, corner jobs, a call cut into
    several launches, a too-small op buffer."""
    import torch
    from test_kernel_logic_cpu import _global_cases
    for kw, sc in ((dict(), (6, 1, 6, 1, 100, 5, None, 1, 4, 1)), (dict(o_del=4, e_del=2, o_ins=7, e_ins=1, a=2, b=3), (4, 2, 7, 1, 100, 5, None, 2, 3, 1))):
        prm = ora.make_params(**kw)
        h = bsw.BandedPairWiseSW(*sc)
        for J, r, q in _global_cases(ora) + [ora.gen_global(30000, div=0.04, seed=0xD1)]:
            o = J.copy(); oc, _ = ora.global2_batch(o, r, q, prm, 8)
            g = J.view(bsw.GLOBALJOB).copy()
            gc = h.global2(g, r, q)
            assert (g == o.view(bsw.GLOBALJOB)).all() and len(gc) == len(oc) and (gc == oc).all()
        h.close()
    # resident buffers, several launches per call
    monkeypatch.setenv("BSW_CUDA_GLOBAL_CHUNK", "4096")
    h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    J, r, q = ora.gen_global(20000, div=0.06, seed=0xD2)
    prm = ora.make_params()
    o = J.copy(); oc, _ = ora.global2_batch(o, r, q, prm, 8)
    dj = torch.from_numpy(J.view(np.uint8).reshape(-1).copy()).cuda()
    dr, dq = torch.from_numpy(r).cuda(), torch.from_numpy(q).cuda()
    dc = torch.zeros(len(oc) + 8, dtype=torch.int32, device="cuda")
    tot = h.dev_global2(dj, len(J), dr, dq, dc, len(oc) + 8, int(J["w"].max()), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert tot == len(oc)
    assert (dj.cpu().numpy().view(bsw.GLOBALJOB) == o.view(bsw.GLOBALJOB)).all()
    assert (dc.cpu().numpy().view(np.uint32)[:tot] == oc).all()
    # op buffer too small: the call reports the size it needs
    small = np.zeros(100, np.uint32)
    with pytest.raises(RuntimeError, match=str(len(oc))):
        h.global2(J.view(bsw.GLOBALJOB).copy(), r, q, small)
    h.close()


def test_guard_bands_untouched(ora, bsw):
    """No resident entry point writes outside the caller's buffers: every buffer sits between two 4 KiB guard bands that
    must come back unchanged (compute-sanitizer is closed on this pool; this covers the caller-visible side of memcheck)."""
    import torch
    G = 4096

    def guarded(a):
        raw = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
        t = torch.full((len(raw) + 2 * G,), 0xA5, dtype=torch.uint8, device="cuda")
        t[G:G + len(raw)] = torch.from_numpy(raw.copy()).cuda()
        return t, t[G:G + len(raw)]

    def intact(*ts):
        for t, body in ts:
            assert bool((t[:G] == 0xA5).all()) and bool((t[G + len(body):] == 0xA5).all())

    st = torch.cuda.current_stream().cuda_stream
    prm = ora.make_params()
    # BSW: thread kernel (large call), lanes-per-pair kernel (small call, 16-bit tier), general kernel
    for P, r, q in (ora.gen_bsw(60000, seed=0x91)["left"], ora.gen_bsw(700, seed=0x92)["left"],
                    ora.gen_bsw(600, read_len=300, seed=0x93, div=0.02, seed_max=150)["right"],
                    ora.gen_random(40, 1100, 1300, 1100, 1300, h0min=10, h0max=20, sim=0.9, seed=14)):
        o, _ = _oracle_bsw(ora, P, r, q)
        tp, tr, tq = guarded(P), guarded(r), guarded(q)
        h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        h.dev_getScores(tp[1], tr[1], tq[1], len(P), 100, st)
        torch.cuda.synchronize()
        _assert_same(o, tp[1].cpu().numpy().view(bsw.SEQPAIR))
        intact(tp, tr, tq)
        h.close()
    # kswv: both passes and the hand-off
    K, kr, kq = ora.gen_kswv(5000, seed=0x94)
    ao, _ = ora.kswv_align2_batch(K, kr, kq, prm, 8)
    tp, tr, tq, ta = guarded(K), guarded(kr), guarded(kq), guarded(np.zeros(len(K), bsw.KSWR))
    ks = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    ks.dev_align2(tp[1], tr[1], tq[1], ta[1], len(K), 1, st)
    torch.cuda.synchronize()
    assert (ta[1].cpu().numpy().view(bsw.KSWR) == ao).all()
    intact(tp, tr, tq, ta)
    ks.close()
    # CIGAR generation
    J, jr, jq = ora.gen_global(5000, seed=0x95)
    jo = J.copy(); oc, _ = ora.global2_batch(jo, jr, jq, prm, 8)
    tj, tr, tq, tc = guarded(J), guarded(jr), guarded(jq), guarded(np.zeros(len(oc), np.uint32))
    h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    tot = h.dev_global2(tj[1], len(J), tr[1], tq[1], tc[1], len(oc), int(J["w"].max()), st)
    torch.cuda.synchronize()
    assert tot == len(oc) and (tc[1].cpu().numpy().view(np.uint32) == oc).all()
    intact(tj, tr, tq, tc)
    h.close()


def test_kswv_align2_host(ora, bsw):
    """kswv_align2: ksw_align2 on host buffers in one call (records and bases up once, hand-off and both passes on the device)
    equals the CPU restatement per pair -- several chunks (ramped), 8- and 16-bit, N, hits below minsc, 16-bit shapes, the
    32-bit kernel, a batch with no second pass, an empty batch; inputs untouched."""
    import os
    prm = ora.make_params()
    os.environ["BSW_CUDA_KSWV_CHUNK"] = "8192"
    try:
        ks = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
        cases = []
        P, r, q = ora.gen_kswv(60000); cases.append((P, r, q, 1))
        P, r, q = ora.gen_kswv(3000, qlen=250, tmin=400, tmax=700, seed=0xC5); cases.append((P, r, q, 0))
        P, r, q = ora.gen_random(3000, 0, 120, 0, 250, sim=0.75, p_n=0.03, seed=31)
        P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 30; cases.append((P, r, q, 1))
        P, r, q = ora.gen_kswv(2000, qlen=300, tmin=400, tmax=700, seed=12, p_n=0.001); cases.append((P, r, q, 0))
        P, r, q = ora.gen_kswv(200, qlen=600, tmin=700, tmax=900, seed=10, xtra=ora.KSW_XSUBO | ora.KSW_XSTART | 19); cases.append((P, r, q, 0))
        P, r, q = ora.gen_kswv(500, seed=77, xtra=ora.KSW_XSUBO | ora.KSW_XBYTE | 19); cases.append((P, r, q, 1))       # no KSW_XSTART
        for P, r, q, is8 in cases:
            P0 = P.copy()
            ao, _ = ora.kswv_align2_batch(P, r, q, prm, 8)
            ag = np.zeros(len(P), bsw.KSWR)
            ks.align2(P, r, q, ag, len(P), is8)
            assert (ao == ag).all(), np.nonzero(ao != ag)[0][:5]
            assert (P == P0).all()
        ks.align2(P, r, q, ag, 0, 1)
        ks.close()
    finally:
        os.environ.pop("BSW_CUDA_KSWV_CHUNK")


def test_kswv_edge(ora, ks, bsw):
    prm = ora.make_params()
    P, r, q = ora.gen_random(2000, 0, 60, 0, 150, sim=0.7, p_n=0.05, seed=3)
    P["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 5
    P, r, q2 = P, r, q
    ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
    ora.kswv_batch(P, r, q, ao, prm, 1, 0, 8)
    ks.getScores8(P, r, q, ag, len(P), 1, 0)
    assert (ao == ag).all()
    # 8-bit overflow rule and the general (32-bit) kernel for long queries
    P, r, q = ora.gen_kswv(300, qlen=255, tmin=300, tmax=400, seed=9, div=0.0,
                           xtra=ora.KSW_XSUBO | ora.KSW_XSTART | ora.KSW_XBYTE | 19)
    ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
    ora.kswv_batch(P, r, q, ao, prm, 1, 0, 8)
    ks.getScores8(P, r, q, ag, len(P), 1, 0)
    assert (ao == ag).all() and (ao["score"] == 255).any()
    P, r, q = ora.gen_kswv(200, qlen=600, tmin=700, tmax=900, seed=10, xtra=ora.KSW_XSUBO | ora.KSW_XSTART | 19)
    ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
    ora.kswv_batch(P, r, q, ao, prm, 0, 0, 8)
    ks.getScores16(P, r, q, ag, len(P), 1, 0)
    assert (ao == ag).all()
    assert ks.stats()["pairs_general"] > 0


def test_kswv_long_targets(ora, ks, bsw):
    """Targets around and beyond the registers kernel's row limit (2046: it sizes the `b` lists in shared memory): the
    longer ones take the 32-bit kernel, results equal either way."""
    prm = ora.make_params()
    P, r, q = ora.gen_kswv(600, qlen=100, tmin=2030, tmax=2070, seed=0xD9)
    ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
    ora.kswv_batch(P, r, q, ao, prm, 1, 0, 8)
    ks.reset_stats()
    ks.getScores8(P, r, q, ag, len(P), 1, 0)
    assert (ao == ag).all(), np.nonzero(ao != ag)[0][:5]
    st = ks.stats()
    assert st["pairs_general"] == int((P["len1"] > 2046).sum()) and 0 < st["pairs_general"] < len(P)


def test_kswv_16bit_tier(ora, ks, bsw):
    """kswv::getScores16 proper: scores above 255 and queries above 255 bases stay on the registers kernel (its 16-bit
    shapes, kswv_sys.cuh W) -- nothing below 2048 / 512 bases may reach the 32-bit kernel.  2 x 300 bp and 2 x 500 bp mates,
    phase 0 and 1, with N; match score 2 on 150 bp mates; asymmetric gap sums."""
    prm = ora.make_params()
    for n, ql, t0, t1 in ((6000, 300, 400, 700), (1500, 500, 600, 900)):
        P, r, q = ora.gen_kswv(n, qlen=ql, tmin=t0, tmax=t1, seed=0xD0 + ql, p_n=0.001)
        ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
        ora.kswv_batch(P, r, q, ao, prm, 0, 0, 8)
        ks.reset_stats()
        ks.getScores16(P, r, q, ag, len(P), 1, 0)
        assert (ao == ag).all(), np.nonzero(ao != ag)[0][:5]
        assert (ao["score"] > 255).any()
        p1, r1, q1 = ora.kswv_phase1_inputs(P, r, q, ao)
        ora.kswv_batch(p1, r1, q1, ao, prm, 0, 1, 8)
        ks.getScores16(p1, r1, q1, ag, len(p1), 1, 1)
        assert (ao == ag).all()
        assert ks.stats()["pairs_general"] == 0
    for sc in (dict(o_del=6, e_del=1, o_ins=6, e_ins=1, a=2, b=4), dict(o_del=4, e_del=2, o_ins=7, e_ins=1, a=3, b=3)):
        h = bsw.kswv(sc["o_del"], sc["e_del"], sc["o_ins"], sc["e_ins"], sc["a"], sc["b"], 1, 2048, 256)
        if ks.route == "direct":
            h.set_direct_min(0)
        P, r, q = ora.gen_kswv(5000, seed=0xD5, a=sc["a"])
        ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
        ora.kswv_batch(P, r, q, ao, ora.make_params(**sc), 0, 0, 8)
        h.getScores16(P, r, q, ag, len(P), 1, 0)
        assert (ao == ag).all(), np.nonzero(ao != ag)[0][:5]
        assert h.stats()["pairs_general"] == 0 and (ao["score"] > 255).any()
        h.close()


def test_device_resident_api(ora, sw, bsw):
    """bsw_dev_getScores on torch CUDA tensors == host-pointer call == oracle."""
    import torch
    (L, lr, lq), _ = common.bsw_c1(5000)
    o, _ = _oracle_bsw(ora, L, lr, lq)
    dp = torch.from_numpy(L.view(np.uint8).copy()).cuda()
    dr = torch.from_numpy(lr).cuda(); dq = torch.from_numpy(lq).cuda()
    sw.dev_getScores(dp, dr, dq, len(L), 100, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    g = dp.cpu().numpy().view(bsw.SEQPAIR)
    _assert_same(o, g)


def test_device_resident_calls_overlap_on_streams(ora, sw, bsw):
    """Device-resident calls issued on different caller streams overlap inside the library (slot ring of
    bsw_dev_getScores): more calls in flight than slots, no host sync in between, then a host-pointer call
    straight after -- every call's results equal the oracle's."""
    import torch
    streams = [torch.cuda.Stream(priority=-1) for _ in range(3)]
    jobs = []
    for i in range(7):                                   # > BSW_NSLOT calls, different sizes and seeds
        (L, lr, lq), _ = common.bsw_c1(1500 + 700 * i, seed=100 + i)
        o, _ = _oracle_bsw(ora, L, lr, lq)
        dp = torch.from_numpy(L.view(np.uint8).copy()).cuda()
        jobs.append((o, dp, torch.from_numpy(lr).cuda(), torch.from_numpy(lq).cuda(), len(L)))
    torch.cuda.synchronize()
    for rep in range(2):                                 # second round: the calls rewrite the same arrays in place
        for i, (o, dp, dr, dq, n) in enumerate(jobs):
            sw.dev_getScores(dp, dr, dq, n, 100, streams[i % len(streams)].cuda_stream)
    (L, lr, lq), _ = common.bsw_c1(3000, seed=7)         # host-pointer call while the device calls are in flight
    oh, _ = _oracle_bsw(ora, L, lr, lq)
    gh = L.copy()
    sw.getScores16(gh, lr, lq, len(gh), 1, 100)
    _assert_same(oh, gh)
    torch.cuda.synchronize()
    for o, dp, _, _, _ in jobs:
        _assert_same(o, dp.cpu().numpy().view(bsw.SEQPAIR))


def _oracle_protocol_side(ora, pairs, ref, qer, w, tries):
    """Host statement of the caller protocol of one side on the oracle."""
    prm = ora.make_params()
    out = pairs.copy()
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
        sub = out[todo].copy()
        ora.bsw_batch(sub, ref, qer, prm, wa, 8)
        out[todo] = sub
    return out, (0 if attempt == 0 else len(todo))


@pytest.mark.parametrize("w,div,tries", [(100, 0.01, 2), (4, 0.05, 2), (2, 0.08, 2), (2, 0.08, 3), (3, 0.08, 1)])
def test_device_side_caller_protocol(ora, sw, bsw, w, div, tries):
    """bsw_dev_extendChained: left attempts with band retry, right h0 chained from the left score by
    regid, right attempts -- one submission on resident pairs == the host protocol on the oracle."""
    import torch
    d = ora.gen_bsw(6000, seed=0xF1, div=div)
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    Lo, nl = _oracle_protocol_side(ora, L, lr, lq, w, tries)
    Rin = R.copy(); ora.chain_h0(Lo, Rin)
    Ro, nr = _oracle_protocol_side(ora, Rin, rr, rq, w, tries)
    if w <= 2 and tries > 1:
        assert nl > 0 and nr > 0                       # the retry path is really exercised
    dev = lambda a: torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).cuda()
    dL, dR = dev(L), dev(R)
    sw.dev_extendChained(dL, dev(lr), dev(lq), len(L), dR, dev(rr), dev(rq), len(R), w, tries,
                         torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    _assert_same(Lo, dL.cpu().numpy().view(bsw.SEQPAIR))
    Rg = dR.cpu().numpy().view(bsw.SEQPAIR)
    assert (Rg["h0"] == Rin["h0"]).all()
    Ro["h0"] = Rin["h0"]
    _assert_same(Ro, Rg)


def test_properties_at_scale(ora, sw, bsw):
    """Size-independent properties on 1M C2-shaped pairs: idempotence, order independence
    (results do not depend on where a pair sits in the batch), all-match closed form."""
    d = ora.gen_bsw(520000, seed=0xC2)
    P, r, q = d["left"]
    a = P.copy(); sw.getScores16(a, r, q, len(a), 1, 100)
    b = a.copy(); sw.getScores16(b, r, q, len(b), 1, 100)
    assert not common.diff_fields(a, b, common.BSW_FIELDS).any()
    perm = np.random.default_rng(0).permutation(len(P))
    c = np.ascontiguousarray(P[perm]); sw.getScores16(c, r, q, len(c), 1, 100)
    assert not common.diff_fields(a[perm], c, common.BSW_FIELDS).any()
    sample = np.random.default_rng(1).choice(len(P), 20000, replace=False)
    o = np.ascontiguousarray(P[sample]); ora.bsw_batch(o, r, q, ora.make_params(), 100, 8)
    assert not common.diff_fields(o, a[sample], common.BSW_FIELDS).any()
    # exact-copy targets: score = h0 + len2, query fully consumed
    d0 = ora.gen_bsw(100000, seed=0xC6, div=0.0, p_n=0.0)
    P0, r0, q0 = d0["right"]
    g = P0.copy(); sw.getScores16(g, r0, q0, len(g), 1, 100)
    assert (g["score"] == g["h0"] + g["len2"]).all() and (g["qle"] == g["len2"]).all() and (g["gscore"] == g["score"]).all()


def test_contract_violation_exits_like_the_reference(bsw):
    """Reference convention (R:docs-archive/WEEK1_WRAPPER_COMPLETE.md:239-242): fprintf + exit(EXIT_FAILURE).
    getScores8 with a 200-base pair, and h0 <= 0, must terminate the process with a message -- checked in a child."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import sys
        sys.path[:0] = %r
        import numpy as np, bswb200
        sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        p = np.zeros(1, bswb200.SEQPAIR); p['len1'] = %d; p['len2'] = 50; p['h0'] = %d
        buf = np.zeros(400, np.uint8)
        sw.getScores8(p, buf, buf, 1, 1, 100)
        print('survived')
    """)
    paths = [p for p in sys.path if p.endswith(("bwa-mem2-arm_b200", "oracle", "tests"))]
    for len1, h0 in ((200, 10), (60, 0)):
        r = subprocess.run([sys.executable, "-c", code % (paths, len1, h0)], capture_output=True, text=True, timeout=120)
        assert r.returncode != 0 and "survived" not in r.stdout
        assert "violate the call contract" in r.stderr


def test_small_and_repeated_calls(ora, sw):
    """Many small calls on one handle (the host pipeline's real pattern: a few thousand pairs per call),
    buffers reused, results independent of call history."""
    prm = ora.make_params()
    for k in range(12):
        P, r, q = ora.gen_random(257 + 97 * k, 1, 127, 1, 127, h0min=1, h0max=100, sim=0.9, p_n=0.002, seed=500 + k)
        o = P.copy(); ora.bsw_batch(o, r, q, prm, 100, 4)
        g = P.copy(); sw.getScores8(g, r, q, len(g), 1, 100)
        _assert_same(o, g)


def test_concurrent_handles_from_host_threads(ora, bsw):
    """The boundary is called concurrently from every worker thread of the host pipeline, each with its own
    object (SURVEY.md 8b "Threading"): four threads, four handles, same device, different batches."""
    import threading
    prm = ora.make_params()
    jobs = []
    for k in range(4):
        d = ora.gen_bsw(6000, seed=0x700 + k)
        P, r, q = d["left"]
        o = P.copy(); ora.bsw_batch(o, r, q, prm, 100, 4)
        jobs.append((P, r, q, o))
    out = [None] * 4

    def work(k):
        P, r, q, _ = jobs[k]
        sw = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        g = None
        for _ in range(3):                       # repeated calls, as a worker does per task
            g = P.copy(); sw.getScores16(g, r, q, len(g), 1, 100)
        sw.close()
        out[k] = g

    th = [threading.Thread(target=work, args=(k,)) for k in range(4)]
    for t in th: t.start()
    for t in th: t.join()
    for k in range(4):
        _assert_same(jobs[k][3], out[k])


def test_kswv_call_in_parts(ora, bsw, monkeypatch):
    """Large kswv calls are served in chunks and, when the row-maximum scratch would exceed its budget, in several
    launches per bin: same answers, phase 0 and 1."""
    monkeypatch.setenv("BSW_CUDA_KSWV_CHUNK", "3000")
    monkeypatch.setenv("BSW_CUDA_ROWMAX_MB", "1")            # floor is 16 MB: 500 rows x 2 B -> <= 16 k slots per launch
    prm = ora.make_params()
    K, kr, kq = ora.gen_kswv(40000, seed=0x77)
    ao = np.zeros(len(K), ora.KSWR); ag = np.zeros(len(K), bsw.KSWR)
    ora.kswv_batch(K, kr, kq, ao, prm, 1, 0, 8)
    for chunk in ("3000", "1000000"):
        monkeypatch.setenv("BSW_CUDA_KSWV_CHUNK", chunk)
        ks2 = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
        ks2.set_direct_min(0)
        ag[:] = 0
        ks2.getScores8(K, kr, kq, ag, len(K), 1, 0)
        assert (ao == ag).all()
        a1 = ao.copy()
        p1, r1, q1 = ora.kswv_phase1_inputs(K, kr, kq, a1)
        ora.kswv_batch(p1, r1, q1, a1, prm, 1, 1, 8)
        ks2.getScores8(p1, r1, q1, ag, len(p1), 1, 1)
        ks2.close()
        assert (a1 == ag).all()


def test_staged_calls_from_many_threads(ora, bsw):
    """The drop-in pattern (SURVEY.md 8b "Threading"): 12 host threads, each with per-task objects, small calls of
    mixed kinds at once -- BSW with two bands and two scoring sets, kswv phase 0 and 1 -- merged into shared device
    batches by the engine.  Every result equals the oracle's, whatever shared its batch."""
    import threading
    rng = np.random.default_rng(5)
    jobs = []
    for k in range(12):
        kind = k % 4
        if kind in (0, 1):
            w = 100 if kind == 0 else 11
            d = ora.gen_bsw(3000 + 211 * k, seed=0x900 + k)
            parts = []
            for side in ("left", "right"):
                P, r, q = d[side]
                o, _, _ = ora.protocol_side(P, r, q, ora.make_params(), w, 1, 4)
                parts.append((P, r, q, o))
            jobs.append(("bsw", (6, 1, 6, 1, 100, 5, 1, 4), w, parts))
        elif kind == 2:
            P, r, q = ora.gen_random(2500, 1, 127, 1, 200, h0min=1, h0max=60, sim=0.85, p_n=0.02, seed=40 + k)
            prm = ora.make_params(o_del=4, e_del=2, o_ins=7, e_ins=1, zdrop=30, end_bonus=0, a=2, b=3)
            o = P.copy(); ora.bsw_batch(o, r, q, prm, 37, 4)
            jobs.append(("bsw", (4, 2, 7, 1, 30, 0, 2, 3), 37, [(P, r, q, o)]))
        else:
            K, kr, kq = ora.gen_kswv(2500, seed=0x50 + k)
            ao, _ = ora.kswv_align2_batch(K, kr, kq, ora.make_params(), 4)
            jobs.append(("kswv", None, 0, [(K, kr, kq, ao)]))
    out = [None] * len(jobs)

    def work(k):
        kind, sc, w, parts = jobs[k]
        res = []
        for rep in range(3):                                   # per-task objects, repeated
            res = []
            if kind == "bsw":
                for P, r, q, _ in parts:
                    h = bsw.BandedPairWiseSW(sc[0], sc[1], sc[2], sc[3], sc[4], sc[5], None, sc[6], sc[7], 1)
                    g = P.copy()
                    step = 700 + 13 * k                        # several calls per object, ragged sizes
                    for a0 in range(0, len(g), step):
                        part = np.ascontiguousarray(g[a0:a0 + step])
                        h.getScores16(part, r, q, len(part), 1, w)
                        g[a0:a0 + step] = part
                    h.close()
                    res.append(g)
            else:
                K, kr, kq, _ = parts[0]
                h = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
                ag = np.zeros(len(K), bsw.KSWR)
                h.getScores8(K, kr, kq, ag, len(K), 1, 0)
                p1, r1, q1 = ora.kswv_phase1_inputs(K, kr, kq, ag.view(ora.KSWR))
                h.getScores8(p1, r1, q1, ag, len(p1), 1, 1)
                h.close()
                res.append(ag)
        out[k] = res

    th = [threading.Thread(target=work, args=(k,)) for k in range(len(jobs))]
    for x in th: x.start()
    for x in th: x.join()
    for k, (kind, _, _, parts) in enumerate(jobs):
        for (P, r, q, o), g in zip(parts, out[k]):
            if kind == "bsw":
                _assert_same(o, g)
            else:
                assert (o == g).all()
    es = bsw.engine_stats(0)
    assert es["batches"] > 0 and es["pairs"] > 0
    assert es["calls"] >= es["batches"]                       # calls were merged (or at worst one per batch)


def test_staged_batch_size_regimes(ora, bsw):
    """engine.cuh submit(): a staged BSW batch runs on the lanes-per-pair kernel alone (small), on both kernels side by side
    (long bins lanes-per-pair, short bins thread-owns-two-pairs) or on the thread kernel alone (large).  Same results."""
    d = ora.gen_bsw(60000, seed=0x7A1)
    P, r, q = d["left"]
    o, _ = _oracle_bsw(ora, P, r, q)
    for n, wide_max, wide_lat, wide_split in ((1500, 32768, 32768, 0), (9000, 32768, 4096, 5), (9000, 32768, 4096, 0),
                                              (40000, 1 << 30, 8192, 9), (40000, 16384, 8192, 0)):
        h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        h.set_direct_min(1 << 30); h.set_wide_policy(wide_max, wide_lat, wide_split)
        bsw.engine_reset_stats(0)
        g = np.ascontiguousarray(P[:n])
        h.getScores16(g, r, q, n, 1, 100)
        _assert_same(o[:n], g)
        es = bsw.engine_stats(0)
        assert es["pairs"] == n and es["batches"] == 1
        assert es["wide_batches"] == (1 if n <= wide_max else 0)
        h.close()


def test_handle_creation_is_cheap(bsw):
    """The reference builds its objects per task: creating + destroying a handle must not
    create streams or pinned memory (0.57 s per handle in round 1)."""
    import time
    h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1); h.close()          # builds the per-device runtime
    t0 = time.perf_counter()
    for _ in range(200):
        h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1); h.close()
        k = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256); k.close()
    us = (time.perf_counter() - t0) / 400 * 1e6
    assert us < 200, us


def test_arena_upload_once(ora, bsw):
    """bsw_arena_begin / bsw_arena_end: the tier-8 and tier-16 calls of a batch share one pair of sequence arenas
    (bwamem.cpp passes the same buffers to both); declared, they go up once.  Same results, fewer H2D bytes."""
    (L, lr, lq), _ = common.bsw_c1(30000, seed=0x6A)
    o, _ = _oracle_bsw(ora, L, lr, lq)
    tiers = ora.classify(L)
    res = {}
    for arena in (False, True):
        h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        h.set_direct_min(0)
        g = L.copy()
        for rep in range(2):                                  # a second batch on the same handle re-declares the arenas
            if arena:
                h.arena_begin(lr, lq)
            for tier, fn in ((0, h.getScores8), (1, h.getScores16)):
                sel = np.nonzero(tiers == tier)[0]
                part = np.ascontiguousarray(g[sel])
                fn(part, lr, lq, len(part), 1, 100)
                g[sel] = part
            if arena:
                h.arena_end()
        res[arena] = (g, h.stats()["h2d_bytes"])
        h.close()
    _assert_same(o, res[False][0])
    _assert_same(o, res[True][0])
    assert res[True][1] < 0.75 * res[False][1]


def test_arena_deferred_completion(ora, bsw):
    """bsw_arena_defer: the large calls of an arena bracket are recorded and run as one interleaved chunk pipeline by
    bsw_arena_end; every result has landed when it returns.  Small chunks so that the two calls really interleave."""
    import os
    (L, lr, lq), _ = common.bsw_c1(40000, seed=0x6B)
    o, _ = _oracle_bsw(ora, L, lr, lq)
    tiers = ora.classify(L)
    os.environ["BSW_CUDA_CHUNK_PAIRS"] = "4096"
    try:
        h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    finally:
        os.environ.pop("BSW_CUDA_CHUNK_PAIRS")
    h.set_direct_min(0)
    h.arena_defer(True)
    for rep in range(2):
        parts = []
        h.arena_begin(lr, lq)
        for tier, fn in ((0, h.getScores8), (1, h.getScores16)):
            sel = np.nonzero(tiers == tier)[0]
            part = np.ascontiguousarray(L[sel])
            fn(part, lr, lq, len(part), 1, 100)
            parts.append((sel, part))
        h.arena_end()
        g = L.copy()
        for sel, part in parts:
            g[sel] = part
        _assert_same(o, g)
    h.arena_defer(False)
    h.close()


def _ref_workload(ora, n, seed, l_pac=200000):
    """Pairs whose targets are slices of a reference: both strands, both read directions."""
    rng = np.random.default_rng(seed)
    fwd = rng.integers(0, 4, l_pac).astype(np.uint8)
    pairs = np.zeros(n, ora.SEQPAIR); rpos = np.zeros(n, np.int64)
    refs, qers = [], []
    ro = qo = 0
    for i in range(n):
        ql = int(rng.integers(1, 131)); tl = ql + int(rng.integers(0, 60))
        down = bool(rng.integers(0, 2))
        x0 = int(rng.integers(tl, 2 * l_pac - tl))
        if (x0 < l_pac) != ((x0 - tl if down else x0 + tl) < l_pac):      # keep a slice on one strand, like bns_fetch_seq
            x0 = int(rng.integers(tl, l_pac - tl))
        rp = ~x0 if down else x0
        t = ora.ref_slice(fwd, rp, tl)
        q = t[:ql].copy()
        m = rng.random(ql) < 0.04
        q[m] = (q[m] + rng.integers(1, 4, m.sum())) % 4
        if rng.random() < 0.05:
            q[rng.integers(0, ql)] = 4
        pairs[i]["idr"], pairs[i]["idq"], pairs[i]["id"] = ro, qo, i
        pairs[i]["len1"], pairs[i]["len2"], pairs[i]["h0"] = tl, ql, int(rng.integers(1, 100))
        rpos[i] = rp
        refs.append(t); qers.append(q); ro += tl; qo += ql
    ref = np.concatenate(refs + [np.zeros(64, np.uint8)]); qer = np.concatenate(qers + [np.zeros(64, np.uint8)])
    return fwd, pairs, rpos, ref, qer


def test_reference_resident_targets(ora, bsw, monkeypatch):
    """bsw_ref_upload + bsw_getScoresRef: targets fetched on the device from the 2-bit reference at
    (rpos, len1, direction, strand) == the same pairs with the target bases sent as bytes == the oracle."""
    monkeypatch.setenv("BSW_CUDA_CHUNK_PAIRS", "2048")        # several chunks, the sampled size estimate at work
    fwd, P, rpos, ref, qer = _ref_workload(ora, 9000, 77)
    o, _ = _oracle_bsw(ora, P, ref, qer)
    R = bsw.Reference(bsw.pack_pac(fwd), len(fwd))
    h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    g = P.copy()
    g["idr"] = 123456789                                      # ignored by the call, and preserved
    h.getScoresRef(g, R, rpos, qer, len(g), 100)
    assert (g["idr"] == 123456789).all()
    g["idr"] = P["idr"]
    _assert_same(o, g)
    st = h.stats()
    assert st["h2d_bytes"] < len(P) * 56 + len(qer) + len(P) * 8 + 4096       # no target bases went up
    h.close(); R.close()


def _kswv_ref_workload(ora, n, seed, qlo, qhi, l_pac=300000, skew=False):
    """Mate-rescue pairs whose windows are slices of a reference, as mem_matesw cuts them (bns_fetch_seq): both strands
    (and, beyond what bwa issues, both directions); the mate is a mutated piece of its window."""
    rng = np.random.default_rng(seed)
    fwd = rng.integers(0, 4, l_pac).astype(np.uint8)
    pairs = np.zeros(n, ora.SEQPAIR); rpos = np.zeros(n, np.int64)
    refs, qers = [], []
    ro = qo = 0
    for i in range(n):
        ql = int(rng.integers(qlo, qhi + 1)); tl = ql + int(rng.integers(50, 300))
        if skew:                                                # every 32nd pair (the ones the size estimate samples) short, the rest long
            tl = ql + 20 if i % 32 == 0 else ql + int(rng.integers(400, 600))
        down = bool(rng.integers(0, 4) == 0)
        x0 = int(rng.integers(tl, 2 * l_pac - tl))
        if (x0 < l_pac) != ((x0 - tl if down else x0 + tl) < l_pac):
            x0 = int(rng.integers(tl, l_pac - tl))
        rp = ~x0 if down else x0
        t = ora.ref_slice(fwd, rp, tl)
        o = int(rng.integers(0, tl - ql + 1))
        q = t[o:o + ql].copy()
        m = rng.random(ql) < 0.05
        q[m] = (q[m] + rng.integers(1, 4, m.sum())) % 4
        if rng.random() < 0.05:
            q[rng.integers(0, ql)] = 4
        if rng.random() < 0.1:                                  # an unrelated mate: low score, no second pass
            q = rng.integers(0, 4, ql).astype(np.uint8)
        pairs[i]["idr"], pairs[i]["idq"], pairs[i]["id"] = ro, qo, i
        pairs[i]["len1"], pairs[i]["len2"] = tl, ql
        pairs[i]["h0"] = ora.KSW_XSUBO | ora.KSW_XSTART | (ora.KSW_XBYTE if ql < 250 else 0) | 19
        rpos[i] = rp
        refs.append(t); qers.append(q); ro += tl; qo += ql
    ref = np.concatenate(refs + [np.zeros(64, np.uint8)]); qer = np.concatenate(qers + [np.zeros(64, np.uint8)])
    return fwd, pairs, rpos, ref, qer


def test_kswv_align2_reference_resident(ora, bsw, monkeypatch):
    """kswv_align2Ref (SURVEY.md 8(f).2 applied to mate rescue): windows fetched on the device from the resident 2-bit
    reference at rpos == kswv_align2 on the same windows sent as bytes == the CPU restatement; several ramped chunks on
    two workers (the sampled estimate of the target scratch at work), 8- and 16-bit arithmetic; no target base goes up."""
    monkeypatch.setenv("BSW_CUDA_KSWV_CHUNK", "2048")
    prm = ora.make_params()
    ks = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    # third case: the sampled estimate of the target scratch is far too small -> the exact retry of the chunk
    for n, seed, qlo, qhi, is8, skew in ((12000, 5, 60, 150, 1, False), (3000, 6, 200, 300, 0, False), (6000, 7, 60, 150, 1, True)):
        fwd, P, rpos, ref, qer = _kswv_ref_workload(ora, n, seed, qlo, qhi, skew=skew)
        R = bsw.Reference(bsw.pack_pac(fwd), len(fwd))
        ao, _ = ora.kswv_align2_batch(P, ref, qer, prm, 8)
        assert (ao["tb"] >= 0).sum() > n // 2
        a1 = np.zeros(n, bsw.KSWR)
        ks.align2(P, ref, qer, a1, n, is8)
        assert (ao == a1).all(), np.nonzero(ao != a1)[0][:5]
        g = P.copy(); g["idr"] = 987654321                    # ignored by the call
        a2 = np.full(n, -7, np.int32).repeat(7).view(bsw.KSWR)[:n].copy()
        ks.reset_stats()
        ks.align2Ref(g, R, rpos, qer, a2, n, is8)
        assert (ao == a2).all(), np.nonzero(ao != a2)[0][:5]
        assert (g["idr"] == 987654321).all()
        assert ks.stats()["h2d_bytes"] < (2 if skew else 1) * (n * 56 + n * 8 + len(qer)) + 4096
        R.close()
    ks.close()


def test_shim_classes_cpp_harness(ora, bsw, tmp_path):
    """bandedSWA_b200.h compiled into tools/bin/stream_c5 (g++, links libbswcuda.so): worker threads with per-task
    BandedPairWiseSW / kswv objects replay the caller protocol of both call sites; all pairs equal the oracle."""
    import subprocess
    exe = os.path.join(os.path.dirname(HERE), "tools", "bin", "stream_c5")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(os.path.dirname(HERE), "tools"), "bin/stream_c5"])
    prm = ora.make_params()
    n_reads = 12000
    d = ora.gen_bsw(n_reads, seed=0xC5)
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    Lo, _, _ = ora.protocol_side(L, lr, lq, prm, 100, 2, 8)
    Rin = R.copy(); ora.chain_h0(Lo, Rin)
    Ro, _, _ = ora.protocol_side(Rin, rr, rq, prm, 100, 2, 8)
    w_in, w_out = str(tmp_path / "w.bin"), str(tmp_path / "r.bin")
    ora.write_stream_workload(w_in, 0, L, R, lr, lq, rr, rq, ora.task_bounds(L["seqid"], n_reads, 512),
                              ora.task_bounds(R["seqid"], n_reads, 512))
    for pinned in (0, 1):
        r = subprocess.run([exe, "--in", w_in, "--out", w_out, "--threads", "6", "--pinned", str(pinned)],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        info = json.loads(r.stdout.strip().splitlines()[-1])
        assert info["handles_created"] >= 2 * ((n_reads + 511) // 512)
        raw = np.fromfile(w_out, dtype=bsw.SEQPAIR)
        _assert_same(Lo, raw[:len(L)])
        Ro2 = Ro.copy(); Ro2["h0"] = Rin["h0"]
        _assert_same(Ro2, raw[len(L):])
    K, kr, kq = ora.gen_kswv(9000, seed=0xC7)
    K16, kr16, kq16 = ora.gen_kswv(1000, qlen=250, tmin=400, tmax=700, seed=0xC8)
    Kall, krall, kqall = common.concat_workloads([(K, kr, kq), (K16, kr16, kq16)])
    perm = np.random.default_rng(3).permutation(len(Kall)); Kall = np.ascontiguousarray(Kall[perm])
    ao, _ = ora.kswv_align2_batch(Kall, krall, kqall, prm, 8)
    tb = np.arange(0, len(Kall) + 2499, 2500).clip(max=len(Kall)).astype(np.int64)
    ora.write_stream_workload(w_in, 1, Kall, Kall[:0], krall, kqall, krall[:0], kqall[:0], tb, np.zeros(len(tb), np.int64))
    r = subprocess.run([exe, "--in", w_in, "--out", w_out, "--threads", "4"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    ag = np.fromfile(w_out, dtype=bsw.KSWR)
    assert (ao == ag.view(ora.KSWR)).all()


def test_multi_device_handle(ora, bsw, monkeypatch):
    """In-process multi-GPU (BSW_CUDA_DEVICES=0,1,...): the route an unchanged single-process bwa-mem2 would use.  Large
    calls deal their chunks round-robin, small calls go to the least-loaded engine; results equal the oracle's."""
    import threading
    nd = bsw.device_count()
    if nd < 2:
        pytest.skip("needs 2 or more GPUs")
    monkeypatch.setenv("BSW_CUDA_DEVICES", ",".join(str(i) for i in range(nd)))
    monkeypatch.setenv("BSW_CUDA_CHUNK_PAIRS", "8192")
    (L, lr, lq), _ = common.bsw_c1(60000, seed=0x2D)
    o, _ = _oracle_bsw(ora, L, lr, lq)
    h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    h.set_direct_min(0)
    g = L.copy(); h.getScores16(g, lr, lq, len(g), 1, 100)
    _assert_same(o, g)
    h.close()
    out = [None] * 8

    def work(k):
        hh = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        gg = np.ascontiguousarray(L[k::8])
        for a0 in range(0, len(gg), 1500):
            part = np.ascontiguousarray(gg[a0:a0 + 1500]); hh.getScores16(part, lr, lq, len(part), 1, 100); gg[a0:a0 + 1500] = part
        hh.close(); out[k] = gg

    th = [threading.Thread(target=work, args=(k,)) for k in range(8)]
    for x in th: x.start()
    for x in th: x.join()
    for k in range(8):
        _assert_same(o[k::8], out[k])
    used = sum(1 for dv in range(nd) if bsw.engine_stats(dv)["batches"] > 0)
    assert used >= 2
    # kswv on every device of the handle: the two boundary calls and the one-call ksw_align2
    monkeypatch.setenv("BSW_CUDA_KSWV_CHUNK", "4096")
    prm = ora.make_params()
    P, r, q = ora.gen_kswv(40000, seed=0x2E)
    ao, _ = ora.kswv_align2_batch(P, r, q, prm, 8)
    ks = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    ks.set_direct_min(0)
    ag = np.zeros(len(P), bsw.KSWR)
    ks.align2(P, r, q, ag, len(P), 1)
    assert (ao == ag).all()
    a0 = np.zeros(len(P), ora.KSWR); ora.kswv_batch(P, r, q, a0, prm, 1, 0, 8)
    ag[...] = 0
    ks.getScores8(P, r, q, ag, len(P), 1, 0)
    assert (a0 == ag).all()
    # ... and with the windows cut from the reference copy each device holds
    fwd, P, rpos, ref, qer = _kswv_ref_workload(ora, 20000, 9, 60, 150)
    Rf = bsw.Reference(bsw.pack_pac(fwd), len(fwd))
    ao, _ = ora.kswv_align2_batch(P, ref, qer, prm, 8)
    ag = np.zeros(len(P), bsw.KSWR)
    ks.align2Ref(P, Rf, rpos, qer, ag, len(P), 1)
    assert (ao == ag).all()
    Rf.close()
    ks.close()
