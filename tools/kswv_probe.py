"""Config 4 probe: kswv batched mate rescue on 1M synthetic pairs (resident timing + parity on a sample)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
n = int(os.environ.get("KSWV_PAIRS", "1000000"))
K, kr, kq = ora.gen_kswv(n)
ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
st = torch.cuda.current_stream().cuda_stream
dk = torch.from_numpy(K.view(np.uint8).copy()).cuda(); dr = torch.from_numpy(kr).cuda(); dq = torch.from_numpy(kq).cuda()
da = torch.zeros(n * 28, dtype=torch.uint8, device="cuda")
for _ in range(2): ks.dev_getScores(dk, dr, dq, da, n, 1, 0, st)
torch.cuda.synchronize(); ks.reset_stats()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): ks.dev_getScores(dk, dr, dq, da, n, 1, 0, st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3; s = ks.stats(); cells = s["cells"] / 3
aln = da.cpu().numpy().view(bswb200.KSWR)
ns = min(n, 50000)
ao = np.zeros(ns, ora.KSWR); ora.kswv_batch(K[:ns], kr, kq, ao, ora.make_params(), 1, 0, 16)
ok = bool((ao == aln[:ns]).all())
peak = bswb200.measure_dpx_peak(0, 2) * 2 / 7 / 1e9
# e2e through the C ABI, pinned
def pin(a):
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True); o = t.numpy().view(a.dtype).reshape(a.shape); o[...] = a; return o, t
hK, t1 = pin(K); hr, t2 = pin(kr); hq, t3 = pin(kq); ha, t4 = pin(np.zeros(n, bswb200.KSWR))
ks.getScores8(hK, hr, hq, ha, n, 1, 0)
t0 = time.perf_counter(); ks.getScores8(hK, hr, hq, ha, n, 1, 0); e2e = time.perf_counter() - t0
out = dict(pairs=n, ms=ms, cells=cells, gcups=cells / ms / 1e6, pairs_per_s=n / ms * 1e3, parity_sample_ok=ok,
           peak_gcups=peak, frac=cells / ms / 1e6 / peak, e2e_ms=e2e * 1e3, e2e_gcups=cells / e2e / 1e9, packed=s["pairs_packed"] / 3)
print(json.dumps(out))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "kswv_probe.json"), "w"))
