"""Parity of the CUDA path (through the C ABI) with the CPU oracle -- bit-exact on every
result field.  Runs on the B200 box (`pytest -m gpu`); nothing here reads /root/reference."""
import json
import os

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def sw(bsw):
    assert bsw.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    obj = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    yield obj
    obj.close()


@pytest.fixture(scope="module")
def ks(bsw):
    obj = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
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


def test_general_path_long_and_high_scores(ora, sw):
    """Pairs the packed kernel does not admit (score bound > 255 or query > 255) go to the
    32-bit kernel: 400 bp near-identical pairs, and a 1500 x 2000 pair."""
    P, r, q = ora.gen_random(300, 256, 400, 256, 500, h0min=100, h0max=300, sim=0.97, p_n=0.001, seed=99)
    o, _ = _oracle_bsw(ora, P, r, q)
    g = P.copy()
    sw.getScores16(g, r, q, len(g), 1, 100)
    _assert_same(o, g)
    assert sw.stats()["pairs_general"] > 0
    P, r, q = ora.gen_random(3, 1500, 1500, 2000, 2000, h0min=40, h0max=40, sim=0.98, p_n=0.0, seed=5)
    o, _ = _oracle_bsw(ora, P, r, q)
    g = P.copy()
    sw.scalarBandedSWAWrapper(g, r, q, len(g), 1, 100)
    _assert_same(o, g)


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
        assert ks.stats()["cells"] == int(cells.sum())
        p1, r1, q1 = ora.kswv_phase1_inputs(P, r, q, ao)
        ora.kswv_batch(p1, r1, q1, ao, prm, is8, 1, 8)
        f(p1, r1, q1, ag, len(p1), 1, 1)
        assert (ao == ag).all()
        assert (ao["tb"] >= 0).mean() > 0.5


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
    P, r, q = ora.gen_kswv(200, qlen=400, tmin=600, tmax=900, seed=10, xtra=ora.KSW_XSUBO | ora.KSW_XSTART | 19)
    ao = np.zeros(len(P), ora.KSWR); ag = np.zeros(len(P), bsw.KSWR)
    ora.kswv_batch(P, r, q, ao, prm, 0, 0, 8)
    ks.getScores16(P, r, q, ag, len(P), 1, 0)
    assert (ao == ag).all()
    assert ks.stats()["pairs_general"] > 0


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


@pytest.mark.parametrize("threads", [1, 5])
def test_gathered_upload_of_host_calls(ora, bsw, monkeypatch, threads):
    """Host-pointer calls with BSW_CUDA_GATHER_THREADS: the bases the pairs name are gathered into pinned staging by
    helper threads and uploaded densely (the two tier calls share ONE sequence arena, so a byte-range upload carries the
    other tier's bases too).  Same results, caller's idr / idq untouched, fewer H2D bytes."""
    (L, lr, lq), _ = common.bsw_c1(30000, seed=0x6A)
    o, _ = _oracle_bsw(ora, L, lr, lq)
    tiers = ora.classify(L)
    res = {}
    for mode in (0, threads):
        monkeypatch.setenv("BSW_CUDA_GATHER_THREADS", str(mode))
        monkeypatch.setenv("BSW_CUDA_CHUNK_PAIRS", "4096")
        h = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        g = L.copy()
        for tier, fn in ((0, h.getScores8), (1, h.getScores16)):
            sel = np.nonzero(tiers == tier)[0]
            part = np.ascontiguousarray(g[sel])
            fn(part, lr, lq, len(part), 1, 100)
            g[sel] = part
        res[mode] = (g, h.stats()["h2d_bytes"])
        h.close()
    _assert_same(o, res[0][0])
    _assert_same(o, res[threads][0])
    assert res[threads][1] < res[0][1]


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
    """Large kswv calls are served in parts (bounded row-maximum scratch): same answers, phase 0 and 1."""
    monkeypatch.setenv("BSW_CUDA_KSWV_PART", "3000")
    prm = ora.make_params()
    K, kr, kq = ora.gen_kswv(10000, seed=0x77)
    ao = np.zeros(len(K), ora.KSWR); ag = np.zeros(len(K), bsw.KSWR)
    ora.kswv_batch(K, kr, kq, ao, prm, 1, 0, 8)
    ks2 = bsw.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    ks2.getScores8(K, kr, kq, ag, len(K), 1, 0)
    assert (ao == ag).all()
    p1, r1, q1 = ora.kswv_phase1_inputs(K, kr, kq, ao)
    ora.kswv_batch(p1, r1, q1, ao, prm, 1, 1, 8)
    ks2.getScores8(p1, r1, q1, ag, len(p1), 1, 1)
    ks2.close()
    assert (ao == ag).all()
