"""Static issue-cycle estimate of the hot group loop: sums the stall counts in the control bits of each instruction
(Volta+ encoding: bits 105..108 of the 128-bit word = bits 41..44 of the second 64-bit half).
usage: python tools/sass_stalls.py <lib.so> [mangled kernel prefix]"""
import re, subprocess, sys, os
lib = os.path.abspath(sys.argv[1]); prefix = sys.argv[2] if len(sys.argv) > 2 else "_ZN4bswk17bsw_packed_kernelILb0ELb1ELb0"
out = subprocess.run(f"cuobjdump -sass {lib}", shell=True, capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(out) if "Function : " + prefix in l][0]
ins = []   # (addr, text, hi64)
i = start + 1
while i < len(out) and "Function :" not in out[i]:
    m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/", out[i])
    if m and i + 1 < len(out):
        m2 = re.match(r"\s+/\* (0x[0-9a-f]{16}) \*/", out[i + 1])
        if m2: ins.append((int(m.group(1), 16), m.group(2).strip(), int(m2.group(1), 16))); i += 1
    i += 1
# loops: backward branches
best = None
for k, (a, t, h) in enumerate(ins):
    m = re.search(r"BRA (0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < a:
            body = [x for x in ins if tgt <= x[0] <= a]
            nv = sum("VIMNMX3.S16x2" in x[1] for x in body)
            if nv == 4 and (best is None or len(body) < len(best)): best = body
tot = 0
for a, t, h in best:
    stall = (h >> 41) & 0xf; yld = (h >> 45) & 1; wb = (h >> 46) & 7; rb = (h >> 49) & 7; wait = (h >> 52) & 0x3f
    tot += max(stall, 1)
    if "-v" in sys.argv: print(f"{a:5x} stall={stall:2d} y={yld} wb={wb} rb={rb} wait={wait:02x}  {t}")
print(os.path.basename(lib), "loop instrs", len(best), "sum of stall counts", tot)
