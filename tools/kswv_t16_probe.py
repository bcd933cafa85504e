"""The 16-bit tier of kswv (kswv_sys.cuh, W): resident GCUPS and all-pairs parity for mates of 300 and 500 bp and for
match score 2 on 150 bp mates; what the 32-bit thread-per-pair kernel does on the same pairs (BSW_CUDA_KSWV_NO_W=1)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, torch, bswb200, ora
st = torch.cuda.current_stream().cuda_stream
peak = bswb200.measure_dpx_peak(0, 2) * 2 / 7 / 1e9
for tag, n, kw, a in (("2x300bp", 200000, dict(qlen=300, tmin=400, tmax=700), 1), ("2x500bp", 80000, dict(qlen=500, tmin=600, tmax=900), 1),
                      ("150bp, a=2", 400000, dict(), 2)):
    K, kr, kq = ora.gen_kswv(n, seed=0xE0, a=a, **kw)
    ao = np.zeros(n, ora.KSWR); cells = ora.kswv_batch(K, kr, kq, ao, ora.make_params(a=a), 0, 0, 16)
    ks = bswb200.kswv(6, 1, 6, 1, a, 4, 1, 2048, 256)
    dk = torch.from_numpy(K.view(np.uint8).copy()).cuda(); dr = torch.from_numpy(kr).cuda(); dq = torch.from_numpy(kq).cuda()
    da = torch.zeros(n * 28, dtype=torch.uint8, device="cuda")
    for _ in range(6): ks.dev_getScores(dk, dr, dq, da, n, 0, 0, st)       # every slot of the context has grown its buffers
    torch.cuda.synchronize(); ks.reset_stats()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(6): ks.dev_getScores(dk, dr, dq, da, n, 0, 0, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 6; s = ks.stats()
    ok = int((da.cpu().numpy().view(bswb200.KSWR) != ao).sum())
    print(json.dumps({"workload": tag, "pairs": n, "ms": ms, "gcups": cells / ms / 1e6, "frac_of_dpx_peak": cells / ms / 1e6 / peak, "mismatches": ok,
                      "max_score": int(ao["score"].max()), "pairs_packed": s["pairs_packed"] // 6, "pairs_wide": s["pairs_wide"] // 6, "pairs_general": s["pairs_general"] // 6}), flush=True)
    ks.close()
