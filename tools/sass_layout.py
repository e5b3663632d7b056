"""Prints a one-character-per-instruction map of a kernel's SASS between the first LDTM.x16 and the first UTCHMMA
(the epi-1 role of segsum_tc_kernel): M FMNMX, F FADD, B branch, R RED, S ST, L LDTM.  usage: sass_layout.py lib.so mangled_name"""
import re
import subprocess
import sys

out = subprocess.run(["cuobjdump", "-sass", "-fun", sys.argv[2], sys.argv[1]], capture_output=True, text=True).stdout
ins = [l for l in out.split("\n") if re.search(r"^\s+/\*[0-9a-f]{4,6}\*/", l)]
idx = [i for i, l in enumerate(ins) if "LDTM.x16" in l]
mma = [i for i, l in enumerate(ins) if "UTCHMMA" in l]
a, b = idx[0] - 10, mma[0]
seq = []
for i in range(a, b):
    m = re.search(r"\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", ins[i])
    seq.append(m.group(2) if m else "?")
marks = "".join("R" if "RED" in k else "B" if k.startswith("BRA") else "F" if k.startswith("FADD") else "M" if k.startswith("FMNMX")
                else "L" if k.startswith("LDTM") else "S" if k.startswith("ST.") else "." for k in seq)
print("instructions", len(ins), "epi-1 region", a, b)
for i in range(0, len(marks), 130):
    print(a + i, marks[i:i + 130])
