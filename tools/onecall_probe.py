"""How much of a bench step is call structure?  Times the bench workload (resident inputs) as 4 calls (left/right x
tier 8/16, what bench.py does), as 2 calls (left, right) and as ONE call over all pairs, with the DP-kernel time of each."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
n_reads = int(os.environ.get("PROBE_READS", "5060000"))
d = ora.gen_bsw(n_reads, seed=0xC2, nthreads=16); L, lr, lq = d["left"]; R, rr, rq = d["right"]
sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
st = torch.cuda.current_stream().cuda_stream
dev = lambda a: torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).cuda()
dL = dev(L); dlr, dlq, drr, drq = dev(lr), dev(lq), dev(rr), dev(rq)
sw.dev_getScores(dL, dlr, dlq, len(L), 100, st); torch.cuda.synchronize()
ora.chain_h0(dL.cpu().numpy().view(bswb200.SEQPAIR), R)
# one buffer pair for everything: right sequences appended to the left ones
ref = np.concatenate([lr, rr]); qer = np.concatenate([lq, rq])
R2 = R.copy(); R2["idr"] += len(lr); R2["idq"] += len(lq)
ALL = np.concatenate([L, R2]); dref, dqer = dev(ref), dev(qer)
def tiers(p):
    t = ora.classify(p); return [np.ascontiguousarray(p[t == 0]), np.ascontiguousarray(p[t != 0])]
plans = {"4 calls": [(x, dref, dqer) for x in tiers(L) + tiers(R2)], "2 calls": [(L, dref, dqer), (R2, dref, dqer)], "1 call": [(ALL, dref, dqer)]}
sw.set_cell_counting(True); sw.reset_stats(); dA = dev(ALL); sw.dev_getScores(dA, dref, dqer, len(ALL), 100, st); torch.cuda.synchronize()
cells = sw.stats()["cells"]; sw.set_cell_counting(False)
for name, calls in plans.items():
    dps = [dev(c[0]) for c in calls]
    def step():
        for c, dp in zip(calls, dps): sw.dev_getScores(dp, c[1], c[2], len(c[0]), 100, st)
    for _ in range(2): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): step()
    e1.record(); torch.cuda.synchronize(); ms = e0.elapsed_time(e1) / 3
    sw.set_timing(True); sw.reset_stats(); step(); torch.cuda.synchronize(); kms = sw.stats()["kernel_ms"]; sw.set_timing(False)
    print(json.dumps(dict(plan=name, pairs=len(ALL), cells=cells, ms=ms, gcups=cells / ms / 1e6, kernel_ms=kms, kernel_gcups=cells / kms / 1e6)), flush=True)
