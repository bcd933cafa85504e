"""A/B timing of library builds on the bench workload (resident inputs, 4 calls per step like bench.py, and one call).
usage: python tools/ab_bench.py bwa-mem2-arm_b200/ab/a.so bwa-mem2-arm_b200/ab/b.so ...   (each lib in its own subprocess)"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] != "--child":
    for rnd in range(int(os.environ.get("AB_ROUNDS", "2"))):
        for lib in sys.argv[1:]:
            env = dict(os.environ, BSW_CUDA_LIB=os.path.abspath(lib))
            out = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True, text=True)
            print(os.path.basename(lib), out.stdout.strip().split("\n")[-1] if out.returncode == 0 else out.stderr[-400:], flush=True)
    sys.exit(0)
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
n_reads = int(os.environ.get("PROBE_READS", "5060000"))
d = ora.gen_bsw(n_reads, seed=0xC2, nthreads=16); L, lr, lq = d["left"]; R, rr, rq = d["right"]
sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
st = torch.cuda.current_stream().cuda_stream
dev = lambda a: torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).cuda()
dL = dev(L); dlr, dlq, drr, drq = dev(lr), dev(lq), dev(rr), dev(rq)
sw.dev_getScores(dL, dlr, dlq, len(L), 100, st); torch.cuda.synchronize()
Lres = dL.cpu().numpy().view(bswb200.SEQPAIR); ora.chain_h0(Lres, R)
def tiers(p):
    t = ora.classify(p); return [np.ascontiguousarray(p[t == 0]), np.ascontiguousarray(p[t != 0])]
calls = [(x, dlr, dlq) for x in tiers(L)] + [(x, drr, drq) for x in tiers(R)]
dps = [dev(c[0]) for c in calls]
def step():
    for c, dp in zip(calls, dps): sw.dev_getScores(dp, c[1], c[2], len(c[0]), 100, st)
for _ in range(3): step()
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): step()
    e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) / 3)
sw.set_timing(True); sw.reset_stats(); step(); torch.cuda.synchronize(); kms = sw.stats()["kernel_ms"]; sw.set_timing(False)
# checksum of the results: every variant must agree
chk = 0
for dp in dps:
    a = dp.cpu().numpy().view(bswb200.SEQPAIR)
    for f in ("score", "tle", "gtle", "qle", "gscore", "max_off"): chk = (chk * 1000003 + int(a[f].astype(np.int64).sum())) % (1 << 61)
print(json.dumps(dict(ms=round(best, 3), kernel_ms=round(kms, 3), checksum=chk)))
