"""Static register-bank check of the hot group loop: counts instructions whose three register sources fall in one bank
(bank = register index mod 2, Volta-style two-bank file; sources served by the operand reuse cache are skipped).
usage: python tools/sass_banks.py <lib.so> [-v]"""
import re, subprocess, sys, os
sys.argv = [a for a in sys.argv]
lib = os.path.abspath(sys.argv[1]); prefix = "_ZN4bswk17bsw_packed_kernelILb0ELb1ELb0"
out = subprocess.run(f"python {os.path.dirname(os.path.abspath(__file__))}/sass_loop.py {lib} {prefix} -v", shell=True, capture_output=True, text=True).stdout.split("\n")[3:]
prev_reuse = {}   # slot -> reg kept in the reuse cache by the previous instruction
tot3 = conf = 0
for line in out:
    line = line.strip()
    if not line: continue
    body = re.sub(r"^@!?P\d\s+", "", line)
    parts = body.split(None, 1)
    if len(parts) < 2: continue
    ops = [o.strip() for o in parts[1].split(",")]
    srcs = ops[1:]                               # first operand = destination (good enough for ALU/IMAD forms)
    regs = []
    for slot, o in enumerate(srcs):
        m = re.match(r"-?\|?R(\d+)(\.reuse)?", o)
        if m:
            reg = int(m.group(1))
            cached = prev_reuse.get(slot) == reg
            regs.append((reg, cached, bool(m.group(2)), slot))
    new_reuse = {slot: reg for reg, c, ru, slot in regs if ru}
    live = [r for r, cached, ru, slot in regs if not cached]
    if len(regs) >= 3:
        tot3 += 1
        if len(live) >= 3 and len({r % 2 for r in live}) == 1 and len(set(live)) == 3:
            conf += 1
            if "-v" in sys.argv: print("  conflict:", line)
    prev_reuse = new_reuse
print(os.path.basename(lib), "3-source instrs", tot3, "same-bank triples", conf)
