"""Where does the host-pointer (e2e) step of the c2 workload spend its time?  Sweeps the direct path's chunk size / look-ahead
and prints the per-call host trace (BSW_CUDA_TRACE) for the default setting."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, bswb200, ora

def pinned(a):
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
    o = t.numpy().view(a.dtype).reshape(a.shape); o[...] = a
    return o, t

n_reads = int(os.environ.get("PROBE_READS", "5060000"))
d = ora.gen_bsw(n_reads, seed=0xC2, nthreads=16)
L, lr, lq = d["left"]; R, rr, rq = d["right"]
Lo = L.copy(); cells = ora.bsw_batch_avx2(Lo, lr, lq, ora.make_params(), 100, 16)
Rin = R.copy(); ora.chain_h0(Lo, Rin)
Ro = Rin.copy(); cells += ora.bsw_batch_avx2(Ro, rr, rq, ora.make_params(), 100, 16)
keep = []
def pin(a):
    o, t = pinned(a); keep.append(t); return o
h_lr, h_lq, h_rr, h_rq = pin(lr), pin(lq), pin(rr), pin(rq)
calls = []
for P, hr, hq in ((L, h_lr, h_lq), (Rin, h_rr, h_rq)):
    t = ora.classify(P)
    for tier in (0, 1):
        sub = np.ascontiguousarray(P[(t == 0) if tier == 0 else (t != 0)])
        if len(sub): calls.append((pin(sub), hr, hq, 8 if tier == 0 else 16))
hL, hR = pin(L), pin(Rin)

def step(sw, mode):
    if mode == "tiers":
        cur = None
        for hp, hr, hq, tier in calls:
            if cur is not hr:
                if cur is not None: sw.arena_end()
                sw.arena_begin(hr, hq); cur = hr
            (sw.getScores8 if tier == 8 else sw.getScores16)(hp, hr, hq, len(hp), 1, 100)
        sw.arena_end()
    else:                       # one call per side (both tiers share one arithmetic here)
        sw.getScores16(hL, h_lr, h_lq, len(hL), 1, 100)
        sw.getScores16(hR, h_rr, h_rq, len(hR), 1, 100)

rows = []
for chunk, ahead, mode, trace in ((1 << 19, 2, "tiers", 1), (1 << 19, 2, "tiers", 0), (1 << 18, 2, "tiers", 0), (1 << 20, 2, "tiers", 0),
                                  (1 << 19, 1, "tiers", 0), (1 << 17, 2, "tiers", 0), (1 << 19, 2, "sides", 0), (1 << 18, 2, "sides", 0)):
    os.environ["BSW_CUDA_CHUNK_PAIRS"] = str(chunk); os.environ["BSW_CUDA_AHEAD"] = str(ahead); os.environ["BSW_CUDA_TRACE"] = str(trace)
    sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    step(sw, mode); step(sw, mode)
    torch.cuda.synchronize(); sw.reset_stats()
    t0 = time.perf_counter()
    for _ in range(3): step(sw, mode)
    ms = (time.perf_counter() - t0) / 3 * 1e3
    st = sw.stats()
    row = {"chunk_pairs": chunk, "ahead": ahead, "mode": mode, "ms_per_step": ms, "gcups": cells / ms / 1e6, "h2d_gb_per_step": st["h2d_bytes"] / 3 / 1e9,
           "d2h_gb_per_step": st["d2h_bytes"] / 3 / 1e9, "h2d_gbps_if_serial": st["h2d_bytes"] / 3 / 1e6 / ms}
    print(json.dumps(row), flush=True); rows.append(row)
    sw.close()
