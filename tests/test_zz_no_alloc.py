"""bsw_dev_getScores_static -- the allocation-free entry point (include/bsw_cuda.h; no counterpart in the reference)
against the existing bsw_dev_getScores, through the C ABI.

Collected last on purpose (`pytest --collect-only`): a failure here must not hide
the parity tests.  Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def _device_copies(torch, P, r, q):
    pad = np.zeros(16, np.uint8)                       # the kernels of the existing entry point fetch bases as aligned words
    return (torch.from_numpy(P.view(np.uint8).copy()).cuda(),
            torch.from_numpy(np.concatenate([r, pad])).cuda(), torch.from_numpy(np.concatenate([q, pad])).cuda())


def _both(torch, bsw, sw, P, r, q, w=100):
    """(results of bsw_dev_getScores, results of bsw_dev_getScores_static, device free bytes before / after the static call)"""
    stream = torch.cuda.current_stream().cuda_stream
    dp, dr, dq = _device_copies(torch, P, r, q)
    sw.dev_getScores(dp, dr, dq, len(P), w, stream)
    torch.cuda.synchronize()
    want = dp.cpu().numpy().view(bsw.SEQPAIR).copy()
    dp2, _, _ = _device_copies(torch, P, r, q)          # fresh records: the outputs start as the caller left them
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = sw.dev_getScores_static(dp2, dr, dq, len(P), int(P["len2"].max()), w, stream)
    free1 = torch.cuda.mem_get_info()[0]
    assert rc == 0
    got = dp2.cpu().numpy().view(bsw.SEQPAIR).copy()
    return want, got, free0, free1


def test_static_scratch_entry_point_matches_dev_getScores(ora, bsw):
    import torch
    sw = bsw.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    # call 1: C1-shaped left extensions + the documented edge cases (empty, all-N, maximum 8-bit length, ...)
    (L, lr, lq), (R, rr, rq) = common.bsw_c1(2500, seed=0x5A)
    E, er, eq = common.edge_case_pairs()
    P1, r1, q1 = common.concat_workloads([(L, lr, lq), (E, er, eq)])
    want, got, _, _ = _both(torch, bsw, sw, P1, r1, q1)
    bad = common.diff_fields(want, got, common.BSW_FIELDS)
    assert not bad.any(), common.describe_first(want, got, bad)
    for f in ("idr", "idq", "id", "len1", "len2", "h0", "seqid", "regid"):
        assert (got[f] == P1[f]).all()
    # the oracle agrees too (bit-exact: integer work, tolerance 0)
    o = P1.copy()
    ora.bsw_batch(o, r1, q1, ora.make_params(), 100, 8)
    bad = common.diff_fields(o, got, common.BSW_FIELDS)
    assert not bad.any(), common.describe_first(o, got, bad)

    # call 2, straight after, FEWER pairs and other data: the shared array still holds call 1's rows and must not matter
    P2, r2, q2 = R[:1500].copy(), rr, rq
    want, got, free0, free1 = _both(torch, bsw, sw, P2, r2, q2, w=37)
    bad = common.diff_fields(want, got, common.BSW_FIELDS)
    assert not bad.any(), common.describe_first(want, got, bad)
    # (>=: handles of earlier tests may be collected, and their device buffers freed, at any moment)
    assert free1 >= free0, "the static-scratch call took device memory"

    # beyond the static scratch: refused with a code, nothing launched, records untouched
    dp, dr, dq = _device_copies(torch, P2, r2, q2)
    before = dp.clone()
    assert sw.dev_getScores_static(dp, dr, dq, len(P2), (1 << 20) // len(P2) + 1, 100, 0) == -2
    assert sw.dev_getScores_static(dp, dr, dq, (1 << 20) + 1, 0, 100, 0) == -2
    assert sw.dev_getScores_static(dp, dr, dq, -1, 10, 100, 0) == -1
    assert sw.dev_getScores_static(dp, dr, dq, 0, 10, 100, 0) == 0
    torch.cuda.synchronize()
    assert torch.equal(dp, before)
    # a declared maximum below a pair's query: those pairs are marked, the others computed, -4 returned
    short = int(P2["len2"].max()) - 1
    assert sw.dev_getScores_static(dp, dr, dq, len(P2), short, 37, 0) == -4
    g = dp.cpu().numpy().view(bsw.SEQPAIR)
    over = P2["len2"] > short
    assert over.any() and (g["score"][over] == -1).all()
    bad = common.diff_fields(want[~over], g[~over], common.BSW_FIELDS)
    assert not bad.any()
    # and the next call is clean again (the status word is cleared by every call)
    dp3, _, _ = _device_copies(torch, P2, r2, q2)
    assert sw.dev_getScores_static(dp3, dr, dq, len(P2), int(P2["len2"].max()), 37, 0) == 0
    g = dp3.cpu().numpy().view(bsw.SEQPAIR)
    assert not common.diff_fields(want, g, common.BSW_FIELDS).any()
    # recorded into a CUDA graph (what the entry point is for): memset + kernel are captured, nothing is waited for;
    # replayed twice on different records in the same device buffers
    maxq = int(P2["len2"].max())
    sp = torch.from_numpy(P2.view(np.uint8).copy()).cuda()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
        rc = sw.dev_getScores_static(sp, dr, dq, len(P2), maxq, 37, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    graph.replay()
    torch.cuda.synchronize()
    g = sp.cpu().numpy().view(bsw.SEQPAIR)
    assert not common.diff_fields(want, g, common.BSW_FIELDS).any()
    P3 = P2[::-1].copy()                                     # other records, same buffers, same graph
    sp.copy_(torch.from_numpy(P3.view(np.uint8).copy()))
    graph.replay()
    torch.cuda.synchronize()
    g = sp.cpu().numpy().view(bsw.SEQPAIR)
    assert not common.diff_fields(want[::-1], g, common.BSW_FIELDS).any()
    del graph
    sw.close()
