"""Attribute ncu warp-state samples (ncu --page source --csv of one kernel) to source lines, using nvdisasm -g line
annotations of the same kernel in the built object; prints the hottest SASS rows with their line and the hottest lines.
usage: stall_by_line.py <src.csv> <object.o> <mangled-kernel-substring> [top]"""
import csv, collections, re, subprocess, sys, tempfile, os, glob

src_csv, obj, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l and l.rstrip().endswith(":"))
lines, cur, stack = [], ("?", 0), ""
for l in dis[start + 1:]:
    if l.startswith("//---------------------") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); stack = m.group(3)
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        lines.append((cur, stack, m.group(2)))
rows = list(csv.reader(open(src_csv)))
H = rows[1]; D = [r for r in rows[2:] if len(r) > 5]
ss, ii = H.index("# Samples"), H.index("Instructions Executed")
assert len(D) == len(lines), (len(D), len(lines))
tot = sum(int(r[ss]) for r in D if r[ss].isdigit())
print("total samples", tot)
order = sorted(range(len(D)), key=lambda k: -(int(D[k][ss]) if D[k][ss].isdigit() else 0))[:top]
for k in order:
    (f, ln), stk, sass = lines[k]
    prev = lines[k - 1][2] if k else ""
    print(f"{100 * int(D[k][ss]) / tot:5.1f}%  #{k:<5d} {f}:{ln:<4d} {sass[:60]:60s} <- prev: {prev[:50]} {stk[:60]}")
by = collections.Counter()
for k, r in enumerate(D):
    if r[ss].isdigit():
        by[lines[k][0]] += int(r[ss])
print("---- by line")
for (f, ln), n in by.most_common(top):
    print(f"{100 * n / tot:5.1f}%  {f}:{ln}")
