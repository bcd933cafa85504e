"""Exploratory GPU probe (not the bench): DPX issue-rate microbenchmarks and a rough
resident-input timing of the BSW / kswv kernels under a few launch shapes."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import torch
import bswb200
import ora

out = {}
names = {0: "vimax3_s16x2", 1: "viaddmax_s16x2", 2: "dpx_mix", 3: "lop3+iadd (x2)", 4: "viadd_16x2"}
for which in range(5):
    p = bswb200.measure_dpx_peak(0, which)
    out[names[which]] = p
    print(f"peak[{names[which]}] = {p/1e12:.3f} T thread-instr/s  -> per SM per clk @1.965GHz: {p/148/1.965e9:.1f}", flush=True)

n_reads = int(os.environ.get("PROBE_READS", "600000"))
d = ora.gen_bsw(n_reads, seed=0xC2)
L, lr, lq = d["left"]
prm = ora.make_params()
samp = np.random.default_rng(0).choice(len(L), 50000, replace=False)
so = np.ascontiguousarray(L[samp]); tot, cells = ora.bsw_batch(so, lr, lq, prm, 100, 8, True)
print("pairs", len(L), "mean cells/pair (sample)", cells.mean(), flush=True)

dp0 = torch.from_numpy(L.view(np.uint8).copy()).cuda()
dr = torch.from_numpy(lr).cuda(); dq = torch.from_numpy(lq).cuda()
stream = torch.cuda.current_stream().cuda_stream

def run_cfg(env):
    code = f"""
import os, sys, json, time
sys.path[:0] = {[os.path.join(ROOT, 'bwa-mem2-arm_b200'), os.path.join(ROOT, 'oracle')]!r}
import numpy as np, torch, bswb200, ora
d = ora.gen_bsw({n_reads}, seed=0xC2); L, lr, lq = d['left']
dp = torch.from_numpy(L.view(np.uint8).copy()).cuda(); dr = torch.from_numpy(lr).cuda(); dq = torch.from_numpy(lq).cuda()
sw = bswb200.BandedPairWiseSW(6,1,6,1,100,5,None,1,4,1)
st = torch.cuda.current_stream().cuda_stream
for _ in range(2): sw.dev_getScores(dp, dr, dq, len(L), 100, st)
torch.cuda.synchronize(); sw.reset_stats()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): sw.dev_getScores(dp, dr, dq, len(L), 100, st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)/3
s = sw.stats()
print(json.dumps(dict(ms=ms, cells=s['cells']/3, pairs=len(L), gcups=s['cells']/3/ms/1e6, packed=s['pairs_packed']/3)))
"""
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=600)
    line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-500:]
    print(env, line, flush=True)
    return line

for env in ({}, {"BSW_CUDA_BLOCKS_PER_SM": "2", "BSW_CUDA_BLOCK_WARPS": "8"},
            {"BSW_CUDA_BLOCKS_PER_SM": "4", "BSW_CUDA_BLOCK_WARPS": "4"},
            {"BSW_CUDA_BLOCKS_PER_SM": "1", "BSW_CUDA_BLOCK_WARPS": "8"}):
    out[json.dumps(env)] = run_cfg(env)

# kswv rough timing
K, kr, kq = ora.gen_kswv(200000)
ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
dk = torch.from_numpy(K.view(np.uint8).copy()).cuda(); dkr = torch.from_numpy(kr).cuda(); dkq = torch.from_numpy(kq).cuda()
da = torch.zeros(len(K) * 28, dtype=torch.uint8, device="cuda")
for _ in range(2): ks.dev_getScores(dk, dkr, dkq, da, len(K), 1, 0, stream)
torch.cuda.synchronize(); ks.reset_stats()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); ks.dev_getScores(dk, dkr, dkq, da, len(K), 1, 0, stream); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1); s = ks.stats()
print("kswv", dict(ms=ms, cells=s["cells"], gcups=s["cells"] / ms / 1e6), flush=True)
out["kswv"] = dict(ms=ms, cells=s["cells"], gcups=s["cells"] / ms / 1e6)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w"), indent=1)
