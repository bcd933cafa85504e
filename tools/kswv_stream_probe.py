"""Mate-rescue calls at the reference's size: T host threads, each issuing kswv getScores8 calls of ~2.5 k pairs (the fork's
run shows 958,581 rescue pairs in 391 calls, R:docs-archive/WEEK1_BENCHMARK_RESULTS.md:88-95) on its own kswv object, pageable
buffers, phase 0.  Prints pairs/s, GCUPS and us per call; every result is compared with the CPU restatement."""
import json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, bswb200, ora

n = int(os.environ.get("PROBE_PAIRS", "400000")); per_call = int(os.environ.get("PROBE_CALL", "2452"))
prm = ora.make_params()
K, kr, kq = ora.gen_kswv(n)
a0 = np.zeros(n, ora.KSWR); cells = ora.kswv_batch(K, kr, kq, a0, prm, 1, 0, 16)
calls = [(f, min(per_call, n - f)) for f in range(0, n, per_call)]
for T in [int(x) for x in os.environ.get("PROBE_THREADS", "1,4,8,16,32").split(",")]:
    out = np.zeros(n, bswb200.KSWR)
    objs = [bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256) for _ in range(T)]
    def work(t, rep):
        for ci in range(t, len(calls), T):
            f, m = calls[ci]
            part = np.ascontiguousarray(K[f:f + m]); res = np.zeros(m, bswb200.KSWR)
            objs[t].getScores8(part, kr, kq, res, m, 1, 0)
            out[f:f + m] = res
    best = None
    bswb200.engine_reset_stats(0)
    for rep in range(3):
        th = [threading.Thread(target=work, args=(t, rep)) for t in range(T)]
        t0 = time.perf_counter()
        for x in th: x.start()
        for x in th: x.join()
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best
    ok = bool((out.view(ora.KSWR) == a0).all())
    es = bswb200.engine_stats(0); nc = max(1, es["calls"]); nb = max(1, es["batches"])
    prof = {k: round(es[k + "_ns"] / nc / 1e3, 1) for k in ("scan", "reserve", "pack", "wait", "unpack")}
    prof.update({"batches": es["batches"], "pairs_per_batch": es["pairs"] // nb, "submit_us_per_batch": round(es["submit_ns"] / nb / 1e3, 1), "flight_us_per_batch": round(es["flight_ns"] / nb / 1e3, 1)})
    print(json.dumps({"threads": T, "pairs_per_call": per_call, "calls": len(calls), "wall_ms": best * 1e3, "pairs_per_s": n / best,
                      "gcups": cells / best / 1e9, "us_per_call_per_thread": best / (len(calls) / T) * 1e6, "equal": ok, "engine_us_per_call": prof}), flush=True)
    for o in objs: o.close()
