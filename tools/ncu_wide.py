"""Small resident-input run of the lanes-per-pair kernel for ncu captures: 2 x 300 bp right extensions (16-bit tier,
throughput shapes) and a 2.5k-pair call of 150 bp reads (latency shapes)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
st = torch.cuda.current_stream().cuda_stream
d = ora.gen_bsw(int(os.environ.get("NCU_READS", "40000")), read_len=300, seed=0xC6, div=0.02, seed_max=150)
Lo = d["left"][0].copy(); ora.bsw_batch_avx2(Lo, d["left"][1], d["left"][2], ora.make_params(), 100, 8)
R, rr, rq = d["right"]; ora.chain_h0(Lo, R)
sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
dp = torch.from_numpy(R.view(np.uint8).copy()).cuda(); dr = torch.from_numpy(rr).cuda(); dq = torch.from_numpy(rq).cuda()
for _ in range(2): sw.dev_getScores(dp, dr, dq, len(R), 100, st)
d = ora.gen_bsw(1300, seed=0xC2); L, lr, lq = d["left"]
dp2 = torch.from_numpy(L.view(np.uint8).copy()).cuda(); dr2 = torch.from_numpy(lr).cuda(); dq2 = torch.from_numpy(lq).cuda()
for _ in range(2): sw.dev_getScores(dp2, dr2, dq2, len(L), 100, st)
torch.cuda.synchronize()
print("ok", sw.stats())
