"""Prints the hot group loop (the shortest backward-branch body with 4 VIMNMX3.S16x2) of a kernel in a library build.
usage: python tools/sass_loop.py <lib.so> [mangled kernel prefix] [-v]"""
import collections, re, subprocess, sys, tempfile, os
lib = os.path.abspath(sys.argv[1]); prefix = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("-") else "_ZN4bswk17bsw_packed_kernelILb0ELb1ELb0"
tmp = tempfile.mkdtemp()
subprocess.run(f"cd {tmp} && cuobjdump -xelf all {lib} > /dev/null && nvdisasm -c *.cubin > dis.txt", shell=True, check=True)
txt = open(os.path.join(tmp, "dis.txt")).read().split("\n")
start = [i for i, l in enumerate(txt) if l.startswith(".text." + prefix)][0]
labels, ins = {}, []
for l in txt[start + 1:]:
    if l.startswith(".text.") or l.startswith("\t.section"):
        if ins: break
    m = re.match(r"(\.L_x_\d+):", l)
    if m: labels[m.group(1)] = len(ins); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: ins.append(m.group(2).strip())
op = lambda b: (b.split()[1] if b.startswith("@") else b.split()[0])
best, cands = None, []
for i, x in enumerate(ins):
    m = re.search(r"BRA `\((\.L_x_\d+)\)", x)
    if m and m.group(1) in labels and labels[m.group(1)] < i:
        body = ins[labels[m.group(1)]:i + 1]
        nv = sum("VIMNMX3.S16x2" in b for b in body)
        if nv: cands.append((len(body), nv))
        if nv and nv % 4 == 0 and not any(op(b) in ("SEL", "BSSY.RECONVERGENT") for b in body) and (best is None or len(body) / nv < len(best) / bestn): best, bestn = body, nv
print("loops with cells (length, VIMNMX3):", cands)
print(len(ins), "instructions in the kernel; loop:", len(best))
print(sorted(collections.Counter(op(b) for b in best).items(), key=lambda x: -x[1]))
if "-v" in sys.argv:
    print("\n".join(best))
