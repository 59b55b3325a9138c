"""Summarise an ncu report: key raw metrics, opcode mix, stall reasons, hottest SASS lines."""
import csv, collections, itertools, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H = rows[0]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_sector_hit_rate.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__grid_size', 'launch__occupancy_limit_registers',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active']
print("kernel:", rows[2][H.index('Kernel Name')][:60])
for w in want:
    if w in H:
        i = H.index(w)
        print(f"  {w:58s} {rows[1][i]:12s} {[r[i] for r in rows[2:]]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
ks = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name'] + [len(rows)]
seg = rows[ks[0] + 1: ks[1]]
H, D = seg[0], seg[1:]
si, ii, ss = H.index('Source'), H.index('Instructions Executed'), H.index('# Samples')
byop, tot = collections.Counter(), 0
for r in D:
    if not r[ii].isdigit():
        continue
    p = r[si].split()
    op = (p[1] if p[0].startswith('@') else p[0]).split('.')[0]
    byop[op] += int(r[ii]); tot += int(r[ii])
print("  warp instructions", tot, " top opcodes:", ", ".join(f"{o} {100*n/tot:.1f}%" for o, n in byop.most_common(14)))
stall = [(i, h) for i, h in enumerate(H) if h.startswith('stall_') and 'Not Issued' not in h]
st = collections.Counter()
for r in D:
    for i, h in stall:
        if r[i].isdigit():
            st[h] += int(r[i])
s = sum(st.values())
print("  stalls:", ", ".join(f"{h[6:]} {100*n/s:.1f}%" for h, n in st.most_common(8)))
top = sorted(D, key=lambda r: -int(r[ss]) if r[ss].isdigit() else 0)[:14]
for r in top:
    print(f"    {r[ss]:>6s} samples  x{r[ii]:>9s}  {r[si][:80]}")
