"""Aggregates an ncu report's per-SASS-instruction counters by SASS basic block (label to label).
usage: NCU_SKIP=<launch> python tools/ncu_by_block.py <report.ncu-rep> [kernel-mangled-prefix]"""
import collections, csv, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = sys.argv[1]
prefix = sys.argv[2] if len(sys.argv) > 2 else "_ZN4bswk17bsw_packed_kernelILb0ELb1"
tmp = tempfile.mkdtemp()
subprocess.run(f"cd {tmp} && cuobjdump -xelf all {os.environ.get("BSW_CUDA_LIB") or ROOT + "/bwa-mem2-arm_b200/libbswcuda.so"} > /dev/null && nvdisasm -g -c *.cubin > dis.txt", shell=True, check=True)
txt = open(os.path.join(tmp, "dis.txt")).read().split("\n")
start = [i for i, l in enumerate(txt) if l.startswith(".text." + prefix)][0]
ins, cur, blk, nblk = [], None, "entry", 0
for l in txt[start + 1:]:
    if (l.startswith("\t.section") or l.startswith("//-----")) and ins:
        break
    m = re.match(r"(\.L_x_\d+):", l)
    if m:
        blk = m.group(1); continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = int(m.group(2)) if "bsw_core" in m.group(1) else cur; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((m.group(2).strip(), cur, blk, int(m.group(1), 16)))
        if re.search(r"\bBRA\b|\bBSYNC\b|\bEXIT\b", m.group(2)):
            nblk += 1; blk = f"{blk}+"
skip = os.environ.get("NCU_SKIP", "0")
out = subprocess.run(f"ncu -i {rep} --launch-skip {skip} --launch-count 1 --page source --csv", shell=True, capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"][:len(ins)]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.OrderedDict()
for k, r in enumerate(data):
    a = agg.setdefault(ins[k][2], [0, 0, 0, collections.Counter(), set(), ins[k][3], 0, 0])
    opc = ins[k][0].split()[1] if ins[k][0].startswith("@") else ins[k][0].split()[0]
    if re.match(r"(PRMT|LOP3|VIADD|VIMNMX|VIADDMNMX|ISETP|SEL|SHF|IADD3|IABS|FSETP|LEA|PLOP3|VABSDIFF|SGXT|BMSK|FLO|POPC|I2F|F2I)", opc) and not opc.startswith("IMAD"): a[7] += 1
    a[0] += 1; a[1] += int(r[ix["Instructions Executed"]]); a[2] += int(r[ix["# Samples"]]); a[4].add(ins[k][1])
    a[6] += int(r[ix["Thread Instructions Executed"]]) if "Thread Instructions Executed" in ix else 0
    for h in stalls:
        a[3][h[6:]] += int(r[ix[h]])
ti = sum(v[1] for v in agg.values()); ts = sum(v[2] for v in agg.values())
print(f"warp instructions {ti}, samples {ts}")
thr = float(os.environ.get("MIN_PCT", "0.4"))
for b, (n, e, s_, st, lines, addr, te, nalu) in agg.items():
    if s_ / ts * 100 < thr: continue
    ls = sorted(x for x in lines if x)
    tops = ",".join(f"{n_}:{c}" for n_, c in st.most_common(7))
    print(f"{addr:6x} {b:12s} n={n:4d} alu={nalu:3d} exec/inst={e/n:12.0f} {e/ti*100:5.1f}% inst {s_/ts*100:5.1f}% time lanes={te/max(e,1):4.1f} lines {ls[0] if ls else 0}-{ls[-1] if ls else 0} [{tops}]")
