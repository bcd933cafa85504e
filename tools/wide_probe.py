"""Crossover of the two BSW kernels on resident calls: thread-owns-two-pairs (throughput) vs lanes-per-pair (latency),
and the 16-bit tier (2 x 300 bp reads, h0 up to 300) on the lanes-per-pair kernel vs the 32-bit general kernel."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, bswb200, ora, common

def timed(sw, dp, dr, dq, n, reps):
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2): sw.dev_getScores(dp, dr, dq, n, 100, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): sw.dev_getScores(dp, dr, dq, n, 100, st)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

out = []
peak = bswb200.measure_dpx_peak(0, 2) * 2 / 8 / 1e9
d = ora.gen_bsw(300000, seed=0xC2, nthreads=16)
L, lr, lq = d["left"]
dr, dq = torch.from_numpy(lr).cuda(), torch.from_numpy(lq).cuda()
for n in (256, 1024, 2500, 8192, 16384, 32768, 65536, 131072, 262144):
    P = np.ascontiguousarray(L[:n])
    o = P.copy(); cells = ora.bsw_batch(o, lr, lq, ora.make_params(), 100, 16)
    row = {"pairs": n, "cells": int(cells)}
    for name, wm in (("thread", 0), ("wide", 1 << 30)):
        sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1); sw.set_wide_max(wm)
        dp = torch.from_numpy(P.view(np.uint8).reshape(-1).copy()).cuda()
        ms = timed(sw, dp, dr, dq, n, 10 if n <= 65536 else 4)
        sw.set_timing(True); sw.reset_stats()
        for _ in range(4): sw.dev_getScores(dp, dr, dq, n, 100, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize(); kms = sw.stats()["kernel_ms"] / 4; sw.set_timing(False)
        g = dp.cpu().numpy().view(bswb200.SEQPAIR)
        ok = not common.diff_fields(o, g, common.BSW_FIELDS).any()
        row[name] = {"us": ms * 1e3, "dp_window_us": kms * 1e3, "gcups": cells / ms / 1e6, "frac": cells / ms / 1e6 / peak, "parity_ok": bool(ok)}
        sw.close()
    out.append(row); print(json.dumps(row), flush=True)
# 16-bit tier: 2 x 300 bp reads, seeds up to 150 -> h0 up to 300, queries up to 281
d = ora.gen_bsw(150000, read_len=300, seed=0xC6, div=0.02, seed_max=150, nthreads=16)
L, lr, lq = d["left"]; R, rr, rq = d["right"]
Lo = L.copy(); ora.bsw_batch(Lo, lr, lq, ora.make_params(), 100, 16); ora.chain_h0(Lo, R)
for name, (P, r, q) in (("left", (L, lr, lq)), ("right (h0 = left score)", (R, rr, rq))):
    o = P.copy(); cells = ora.bsw_batch(o, r, q, ora.make_params(), 100, 16)
    sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    dp = torch.from_numpy(P.view(np.uint8).reshape(-1).copy()).cuda(); dr, dq = torch.from_numpy(r).cuda(), torch.from_numpy(q).cuda()
    sw.reset_stats(); ms = timed(sw, dp, dr, dq, len(P), 4); s = sw.stats()
    g = dp.cpu().numpy().view(bswb200.SEQPAIR)
    ok = not common.diff_fields(o, g, common.BSW_FIELDS).any()
    row = {"workload": "2x300bp reads " + name, "pairs": len(P), "cells": int(cells), "ms": ms, "gcups": cells / ms / 1e6,
           "frac": cells / ms / 1e6 / peak, "pairs_packed": s["pairs_packed"] // 6, "pairs_wide": s["pairs_wide"] // 6,
           "pairs_general": s["pairs_general"] // 6, "parity_all_pairs_ok": bool(ok), "tiers": np.bincount(ora.classify(P), minlength=3).tolist()}
    out.append(row); print(json.dumps(row), flush=True)
    sw.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "wide_probe.json"), "w"), indent=1)
