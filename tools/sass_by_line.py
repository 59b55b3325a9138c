"""Attribute executed warp instructions (ncu --page source --csv of one kernel) to source lines,
using nvdisasm -g line annotations of the same kernel in the built object.
usage: sass_by_line.py <src.csv> <object.o> <mangled-kernel-substring> [top]"""
import csv, collections, re, subprocess, sys, tempfile, os, glob

src_csv, obj, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# locate the kernel's section
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l and l.rstrip().endswith(":"))
lines, cur, inl = [], ("?", 0), ""
for l in dis[start + 1:]:
    if l.startswith("//---------------------") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); inl = m.group(3)
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        lines.append((cur, m.group(2)))
rows = list(csv.reader(open(src_csv)))
H = rows[1]; D = [r for r in rows[2:] if len(r) > 5]
ii, si = H.index("Instructions Executed"), H.index("Source")
assert len(D) == len(lines), (len(D), len(lines))
by = collections.Counter(); byop = collections.defaultdict(collections.Counter); tot = 0
for (loc, sass), r in zip(lines, D):
    n = int(r[ii]) if r[ii].isdigit() else 0
    by[loc] += n; tot += n
    p = sass.split(); op = (p[1] if p[0].startswith('@') else p[0]).split('.')[0]
    byop[loc][op] += n
print("total warp instructions", tot)
srcs = {}
for (f, ln), n in by.most_common(top):
    if f not in srcs:
        cand = glob.glob(os.path.join(os.path.dirname(os.path.abspath(obj)), "..", f))
        srcs[f] = open(cand[0]).read().splitlines() if cand else []
    text = srcs[f][ln - 1].strip()[:90] if 0 < ln <= len(srcs[f]) else ""
    ops = ", ".join(f"{o} {c * 100 // max(n, 1)}%" for o, c in byop[(f, ln)].most_common(3))
    print(f"{100 * n / tot:5.1f}%  {f}:{ln:<4d} {text}   [{ops}]")
