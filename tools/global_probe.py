"""Row f.4 (CIGAR generation, batched ksw_global2): resident and host-pointer rates of bsw_global2 against the CPU
restatement on all host threads.  One JSON line."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch, bswb200, ora

n = int(os.environ.get("PROBE_JOBS", "2000000"))
cores = len(os.sched_getaffinity(0))
J, r, q = ora.gen_global(n, div=0.03, seed=0xC6)
prm = ora.make_params()
o = J.copy(); t0 = time.perf_counter(); oc, cells = ora.global2_batch(o, r, q, prm, cores); cpu_s = time.perf_counter() - t0
h = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
st = torch.cuda.current_stream().cuda_stream
dj = torch.from_numpy(J.view(np.uint8).reshape(-1).copy()).cuda(); dr, dq = torch.from_numpy(r).cuda(), torch.from_numpy(q).cuda()
dc = torch.zeros(len(oc) + 8, dtype=torch.int32, device="cuda")
maxw = int(J["w"].max())
for _ in range(2): h.dev_global2(dj, n, dr, dq, dc, len(oc) + 8, maxw, st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): tot = h.dev_global2(dj, n, dr, dq, dc, len(oc) + 8, maxw, st)
e1.record(); torch.cuda.synchronize()
dev_ms = e0.elapsed_time(e1) / 3
ok_dev = tot == len(oc) and bool((dj.cpu().numpy().view(bswb200.GLOBALJOB) == o.view(bswb200.GLOBALJOB)).all()) and bool((dc.cpu().numpy().view(np.uint32)[:tot] == oc).all())
g = J.view(bswb200.GLOBALJOB).copy(); cig = np.zeros(len(oc) + 8, np.uint32)
h.global2(g, r, q, cig)
t0 = time.perf_counter()
for _ in range(3): gc = h.global2(g, r, q, cig)
host_ms = (time.perf_counter() - t0) / 3 * 1e3
ok_host = bool((g == o.view(bswb200.GLOBALJOB)).all()) and bool((gc == oc).all())
print(json.dumps({"workload": f"{n} CIGAR jobs shaped like bwa_gen_cigar2's calls (30-150 bp, 3 % divergence, band by bwa's rule)",
                  "cells": int(cells), "ops": int(len(oc)), "parity_all_jobs": {"resident": ok_dev, "host_pointer": ok_host},
                  "resident": {"ms": dev_ms, "jobs_per_s": n / dev_ms * 1e3, "gcups": cells / dev_ms / 1e6},
                  "host_pointer": {"ms": host_ms, "jobs_per_s": n / host_ms * 1e3, "gcups": cells / host_ms / 1e6},
                  "cpu_oracle": {"threads": cores, "s": cpu_s, "jobs_per_s": n / cpu_s, "gcups": cells / cpu_s / 1e9}}))
