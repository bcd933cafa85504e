"""Device-side caller protocol (bsw_dev_extendChained) on the bench workload: one submission for left + chain + right
(with the band-retry pass) against the host-driven sequence left call -> D2H scores -> host chain -> H2D h0 -> right call.
Checks that both give identical results; prints ms and pairs/s."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
n_reads = int(os.environ.get("PROBE_READS", "5060000"))
d = ora.gen_bsw(n_reads, seed=0xC2, nthreads=16); L, lr, lq = d["left"]; R, rr, rq = d["right"]
sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
st = torch.cuda.current_stream().cuda_stream
dev = lambda a: torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).cuda()
dlr, dlq, drr, drq = dev(lr), dev(lq), dev(rr), dev(rq)
# host-driven: left, read back, chain on the host, upload, right (retry pass never fires at w=100 on this workload)
dL, dR = dev(L), dev(R)
def host_driven():
    sw.dev_getScores(dL, dlr, dlq, len(L), 100, st)
    Lres = dL.cpu().numpy().view(bswb200.SEQPAIR)
    Rin = R.copy(); ora.chain_h0(Lres, Rin)
    dR.copy_(torch.from_numpy(Rin.view(np.uint8).reshape(-1)))
    sw.dev_getScores(dR, drr, drq, len(R), 100, st)
    torch.cuda.synchronize()
host_driven()
t0 = time.perf_counter(); host_driven(); host_ms = (time.perf_counter() - t0) * 1e3
refL, refR = dL.cpu().numpy().view(bswb200.SEQPAIR).copy(), dR.cpu().numpy().view(bswb200.SEQPAIR).copy()
# one submission
dL2, dR2 = dev(L), dev(R)
def chained():
    sw.dev_extendChained(dL2, dlr, dlq, len(L), dR2, drr, drq, len(R), 100, 2, st)
chained(); torch.cuda.synchronize()
gL, gR = dL2.cpu().numpy().view(bswb200.SEQPAIR), dR2.cpu().numpy().view(bswb200.SEQPAIR)
same = all((refL[f] == gL[f]).all() and (refR[f] == gR[f]).all() for f in ("score", "tle", "gtle", "qle", "gscore", "max_off", "h0"))
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); chained(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
npairs = len(L) + len(R)
print(json.dumps(dict(pairs=npairs, same_as_host_driven=bool(same), chained_ms=round(best, 3),
                      chained_pairs_per_s=npairs / best * 1e3, host_driven_ms=round(host_ms, 3),
                      host_driven_pairs_per_s=npairs / host_ms * 1e3)))
