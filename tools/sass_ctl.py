"""Decode the scheduling control fields (stall, write/read barrier, wait mask) of a kernel's SASS.
usage: sass_ctl.py <object.o> <mangled-kernel-substring> <first> <last>"""
import re, subprocess, sys
obj, kern, a, b = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
L = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout.splitlines()
s = next(i for i, l in enumerate(L) if "Function :" in l and kern in l)
ins = []
i = s + 1
while i < len(L) and "Function :" not in L[i]:
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* (0x[0-9a-f]+) \*/', L[i])
    if m and i + 1 < len(L):
        m2 = re.match(r'\s*/\* (0x[0-9a-f]+) \*/', L[i + 1])
        c = (int(m2.group(1), 16) >> 41) & 0x1fffff
        ins.append((m.group(1), m.group(2), c & 0xf, (c >> 5) & 7, (c >> 8) & 7, (c >> 11) & 0x3f))
        i += 2
    else:
        i += 1
for k in range(a, min(b, len(ins))):
    ad, t, st, wb, rb, wm = ins[k]
    print(f"{k:4d} {ad} {t[:62]:62s} st={st:2d} wb={wb if wb != 7 else '-'} rb={rb if rb != 7 else '-'} wait={wm:06b}")
