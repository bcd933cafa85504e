"""Where does the host-pointer (e2e) kswv step spend its time?  Times the phase-0 and phase-1 calls of config 4 separately,
over the direct path's chunk size / look-ahead."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, bswb200, ora

keep = []
def pin(a):
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
    o = t.numpy().view(a.dtype).reshape(a.shape); o[...] = a
    keep.append(t)
    return o

n = int(os.environ.get("PROBE_PAIRS", "1000000"))
prm = ora.make_params()
K, kr, kq = ora.gen_kswv(n)
a0 = np.zeros(n, ora.KSWR); c0 = ora.kswv_batch(K, kr, kq, a0, prm, 1, 0, 16)
p1, r1, q1 = ora.kswv_phase1_inputs(K, kr, kq, a0)
a1 = a0.copy(); c1 = ora.kswv_batch(p1, r1, q1, a1, prm, 1, 1, 16)
hK, hkr, hkq, hp1, hr1, hq1 = pin(K), pin(kr), pin(kq), pin(p1), pin(r1), pin(q1)
ha = pin(np.zeros(n, bswb200.KSWR))
n16 = n // 20
K16, kr16, kq16 = ora.gen_kswv(n16, qlen=250, tmin=400, tmax=700, seed=0xC5)
b0 = np.zeros(n16, ora.KSWR); ora.kswv_batch(K16, kr16, kq16, b0, prm, 0, 0, 16)
p16, r16, q16 = ora.kswv_phase1_inputs(K16, kr16, kq16, b0)
b1 = b0.copy(); ora.kswv_batch(p16, r16, q16, b1, prm, 0, 1, 16)
hK16, hkr16, hkq16, hp16, hr16, hq16 = pin(K16), pin(kr16), pin(kq16), pin(p16), pin(r16), pin(q16)
hb = pin(np.zeros(n16, bswb200.KSWR))
for dmin in (1 << 17, 1 << 14):
    os.environ["BSW_CUDA_DIRECT_MIN"] = str(dmin)
    ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    t = [0.0, 0.0]
    for rep in range(5):
        if rep == 2: t = [0.0, 0.0]
        t0 = time.perf_counter(); ks.getScores16(hK16, hkr16, hkq16, hb, n16, 1, 0); t[0] += time.perf_counter() - t0
        t0 = time.perf_counter(); ks.getScores16(hp16, hr16, hq16, hb, len(hp16), 1, 1); t[1] += time.perf_counter() - t0
    print(json.dumps({"slice16_pairs": n16, "direct_min": dmin, "phase0_ms": t[0] / 3 * 1e3, "phase1_ms": t[1] / 3 * 1e3, "equal": bool((hb.view(ora.KSWR) == b1).all())}), flush=True)
    ks.close()
os.environ.pop("BSW_CUDA_DIRECT_MIN")
for chunk, ahead in ((1 << 18, 2), (1 << 17, 2), (1 << 16, 2), (1 << 19, 2), (1 << 18, 1), (1 << 17, 1)):
    os.environ["BSW_CUDA_KSWV_CHUNK"] = str(chunk); os.environ["BSW_CUDA_AHEAD"] = str(ahead)
    ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    t = [0.0, 0.0]
    for rep in range(5):
        if rep == 2: t = [0.0, 0.0]; ks.reset_stats()
        t0 = time.perf_counter(); ks.getScores8(hK, hkr, hkq, ha, n, 1, 0); t[0] += time.perf_counter() - t0
        t0 = time.perf_counter(); ks.getScores8(hp1, hr1, hq1, ha, len(hp1), 1, 1); t[1] += time.perf_counter() - t0
    st = ks.stats()
    ok = bool((ha.view(ora.KSWR) == a1).all())
    print(json.dumps({"chunk_pairs": chunk, "ahead": ahead, "phase0_ms": t[0] / 3 * 1e3, "phase1_ms": t[1] / 3 * 1e3,
                      "phase0_gcups": c0 / (t[0] / 3) / 1e9, "phase1_gcups": c1 / (t[1] / 3) / 1e9, "h2d_gb_per_step": st["h2d_bytes"] / 3 / 1e9, "equal": ok}), flush=True)
    ks.close()
