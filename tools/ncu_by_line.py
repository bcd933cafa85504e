"""Aggregates an ncu report's per-SASS-instruction counters by CUDA source line.
usage: python tools/ncu_by_line.py <report.ncu-rep> [kernel-mangled-prefix] [top]
Joins `ncu --page source --csv` (SASS order) with `nvdisasm -g` line markers of the built library."""
import collections, csv, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = sys.argv[1]
prefix = sys.argv[2] if len(sys.argv) > 2 else "_ZN4bswk17bsw_packed_kernelILb0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
tmp = tempfile.mkdtemp()
subprocess.run(f"cd {tmp} && cuobjdump -xelf all {ROOT}/bwa-mem2-arm_b200/libbswcuda.so > /dev/null && nvdisasm -g -c *.cubin > dis.txt", shell=True, check=True)
txt = open(os.path.join(tmp, "dis.txt")).read().split("\n")
start = [i for i, l in enumerate(txt) if l.startswith(".text." + prefix)][0]
ins, cur = [], None
for l in txt[start + 1:]:
    if (l.startswith("\t.section") or l.startswith("//-----")) and ins:
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((m.group(2).strip(), cur))
skip = os.environ.get("NCU_SKIP", "0")
out = subprocess.run(f"ncu -i {rep} --launch-skip {skip} --launch-count 1 --page source --csv", shell=True, capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"][:len(ins)]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for k, r in enumerate(data):
    a = agg[ins[k][1]]
    a[0] += int(r[ix["Instructions Executed"]]); a[1] += int(r[ix["# Samples"]])
    for h in stalls:
        a[2][h[6:]] += int(r[ix[h]])
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
src = {f: open(os.path.join(ROOT, "bwa-mem2-arm_b200", "csrc", f)).read().split("\n") for f in ("bsw_core.cuh", "bsw_kernels.cuh", "kswv_core.cuh")}
print(f"instructions {ti}, samples {ts}  ({len(ins)} SASS instr, {len(data)} profiled)")
for key, (e, s_, st) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    f, ln = key if key else ("?", 0)
    line = src[f][ln - 1].strip()[:64] if f in src and 0 < ln <= len(src[f]) else ""
    tops = ",".join(f"{n}:{c}" for n, c in st.most_common(2))
    print(f"{s_/ts*100:5.1f}% time {e/ti*100:5.1f}% inst  {f}:{ln:<4d} {line:64s} [{tops}]")
