"""DESIGN.md section 0, open point (i): would a kernel that reproduces bwa's STRIPED ksw_u8 / ksw_i16 (zero-score pad
columns, E from the pre-lazy-F H) answer differently from the exact-row statement that the oracle and the CUDA
kernels implement?  Runs config 4 (phase 0) through both CPU statements and counts the pairs that differ.  CPU only."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, ora

prm = ora.make_params()
out = []
for name, n, kw, is8 in (("C4: 150 bp mates x 300-500 bp windows, 8-bit (ksw_u8)", int(os.environ.get("PROBE_PAIRS", "1000000")), dict(), 1),
                         ("C4 16-bit slice: 250 bp mates x 400-700 bp windows (ksw_i16)", 50000, dict(qlen=250, tmin=400, tmax=700, seed=0xC5), 0),
                         ("143 bp mates (qlen % 16 = 15: one pad column), 8-bit", 100000, dict(qlen=143, seed=0xC9), 1),
                         ("129 bp mates (15 pad columns), 8-bit", 100000, dict(qlen=129, seed=0xCA), 1)):
    K, kr, kq = ora.gen_kswv(n, **kw)
    a = np.zeros(n, ora.KSWR); ora.kswv_batch(K, kr, kq, a, prm, is8, 0, os.cpu_count())
    s = np.zeros(n, ora.KSWR); t0 = time.perf_counter(); ora.kswv_batch_striped(K, kr, kq, s, prm, is8); dt = time.perf_counter() - t0
    per = {f: int((a[f] != s[f]).sum()) for f in ("score", "te", "qe", "score2", "te2")}
    anyd = int(np.any([a[f] != s[f] for f in ("score", "te", "qe", "score2", "te2")], axis=0).sum())
    gained = int(((a["score2"] < 0) & (s["score2"] >= 0)).sum())
    row = {"workload": name, "pairs": n, "pairs_differing": anyd, "per_field": per,
           "striped_reports_a_second_best_where_exact_rows_have_none": gained, "striped_emulation_s": dt}
    print(json.dumps(row), flush=True); out.append(row)
json.dump(out, open(os.path.join(ROOT, "profiles", "r02_striped_vs_exact.json"), "w"), indent=1)
