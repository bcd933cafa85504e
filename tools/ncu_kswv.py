"""Small resident-input kswv run for ncu captures (phase 0 of config-4 shaped pairs, twice)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
n = int(os.environ.get("NCU_PAIRS", "200000"))
K, kr, kq = ora.gen_kswv(n)
ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
dp = torch.from_numpy(K.view(np.uint8).copy()).cuda(); dr = torch.from_numpy(kr).cuda(); dq = torch.from_numpy(kq).cuda()
da = torch.zeros(n * 28, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(2): ks.dev_getScores(dp, dr, dq, da, n, 1, 0, st)
torch.cuda.synchronize()
print("ok", ks.stats())
