"""Small resident-input BSW run for ncu captures (one warm-up call + one profiled call)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
n_reads = int(os.environ.get("NCU_READS", "150000"))
d = ora.gen_bsw(n_reads, seed=0xC2); L, lr, lq = d["left"]
dp = torch.from_numpy(L.view(np.uint8).copy()).cuda(); dr = torch.from_numpy(lr).cuda(); dq = torch.from_numpy(lq).cuda()
sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
st = torch.cuda.current_stream().cuda_stream
for _ in range(int(os.environ.get("NCU_CALLS", "2"))):
    sw.dev_getScores(dp, dr, dq, len(L), 100, st)
torch.cuda.synchronize()
print("ok", sw.stats())
