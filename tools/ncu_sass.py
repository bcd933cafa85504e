"""Prints the SASS of an address range with executed counts, samples and top stall reasons.
usage: NCU_SKIP=<launch> python tools/ncu_sass.py <report.ncu-rep> <hex from> <hex to>"""
import csv, os, subprocess, sys
rep, lo, hi_ = sys.argv[1], int(sys.argv[2], 16), int(sys.argv[3], 16)
skip = os.environ.get("NCU_SKIP", "0")
out = subprocess.run(f"ncu -i {rep} --launch-skip {skip} --launch-count 1 --page source --csv", shell=True, capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"]
base = int(data[0][0], 16) if data[0][0].startswith("0x") else int(data[0][0])
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in data:
    a = (int(r[0], 16) if r[0].startswith("0x") else int(r[0])) - base
    if lo <= a < hi_:
        st = sorted(((int(r[ix[h]]), h[6:]) for h in stalls), reverse=True)[:2]
        print(f"{a:6x} {int(r[ix['Instructions Executed']]):9d} {int(r[ix['# Samples']]):6d}  {r[ix['Source']][:70]:70s} {st}")
