"""Lists the backward branches (loops) of a kernel's SASS with their length and opcode mix.
Usage: python tools/sass_loops.py file.o [min_len]"""
import re
import subprocess
import sys
from collections import Counter

out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
min_len = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ins = []
for l in out.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
    elif "Function :" in l:
        ins.append((-1, l.strip()))
addr_idx = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    if a < 0:
        print(t)
        continue
    m = re.search(r"BRA\S*\s+(?:\S+,\s+)*0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in addr_idx:
        j = addr_idx[int(m.group(1), 16)]
        n = i - j + 1
        if n < min_len:
            continue
        ops = Counter()
        for _, tt in ins[j:i + 1]:
            tt = re.sub(r"^@!?U?P\d+\s+", "", tt)
            ops[tt.split()[0].split(".")[0]] += 1
        f64 = sum(ops[k] for k in ("DFMA", "DMUL", "DADD", "DSETP"))
        print(f"loop 0x{ins[j][0]:05x}..0x{a:05x}: {n} instr, fp64 {f64}, LDL {ops['LDL']}, STL {ops['STL']}, LDS {ops['LDS']}, LDG {ops['LDG']+ops['LD']}, "
              f"MUFU {ops['MUFU']}, BRA {ops['BRA']}")
