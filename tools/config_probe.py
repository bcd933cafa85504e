"""Resident-input timing + sampled parity of the BASELINE.json configs other than the bench's config[1]:
C1 (100k PE reads, 150 bp, caller protocol shape), C3 (250 bp reads, 5 % divergence, getScores16 + z-drop)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, bswb200, ora, common

def run(name, pairs, ref, qer, reps=5):
    sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    st = torch.cuda.current_stream().cuda_stream
    dp = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).cuda(); dr = torch.from_numpy(ref).cuda(); dq = torch.from_numpy(qer).cuda()
    sw.set_cell_counting(True); sw.reset_stats(); sw.dev_getScores(dp, dr, dq, len(pairs), 100, st); torch.cuda.synchronize()
    s0 = sw.stats(); cells = s0["cells"]; sw.set_cell_counting(False)
    for _ in range(2): sw.dev_getScores(dp, dr, dq, len(pairs), 100, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): sw.dev_getScores(dp, dr, dq, len(pairs), 100, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    g = dp.cpu().numpy().view(bswb200.SEQPAIR)
    ns = min(len(pairs), 40000)
    o = np.ascontiguousarray(pairs[:ns]); ocells = ora.bsw_batch(o, ref, qer, ora.make_params(), 100, 16)
    ok = not common.diff_fields(o, g[:ns], common.BSW_FIELDS).any()
    full = int((pairs["len1"].astype(np.int64) * pairs["len2"]).sum())
    r = dict(config=name, pairs=len(pairs), cells=cells, ms=ms, gcups=cells / ms / 1e6, pairs_per_s=len(pairs) / ms * 1e3,
             packed=s0["pairs_packed"], general=s0["pairs_general"], parity_sample_ok=bool(ok),
             cells_over_full_matrix=cells / max(full, 1), tiers=np.bincount(ora.classify(pairs), minlength=3).tolist())
    sw.close(); print(json.dumps(r), flush=True); return r

out = []
d = ora.gen_bsw(200000, seed=0xC1)                      # C1: 100k paired-end reads = 200k reads
L, lr, lq = d["left"]; R, rr, rq = d["right"]
Lo = L.copy(); ora.bsw_batch(Lo, lr, lq, ora.make_params(), 100, 16); ora.chain_h0(Lo, R)
out.append(run("C1 left (100k PE reads, 150bp)", L, lr, lq)); out.append(run("C1 right (h0 = left score)", R, rr, rq))
d = ora.gen_bsw(1000000, read_len=250, seed=0xC3, div=0.05)   # C3: 2M pairs
L, lr, lq = d["left"]
out.append(run("C3 left (250bp, 5% divergence, getScores16, z-drop)", L, lr, lq, reps=3))
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "config_probe.json"), "w"), indent=1)
