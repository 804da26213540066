"""Aggregate an `ncu --page source --csv --print-source=cuda,sass` dump by CUDA source line:
warp instructions, thread instructions, avg lanes. usage: ncu_lines.py dump.csv [top N]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = collections.OrderedDict()
cur_file = None; hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    iI = hdr.index("Instructions Executed"); iT = hdr.index("Thread Instructions Executed")
    if not r[iI].isdigit(): continue
    k = (cur_file, ln)
    a = agg.setdefault(k, [0, 0, r[1].strip()[:70]])
    a[0] += int(r[iI]); a[1] += int(r[iT])
tot_i = sum(a[0] for a in agg.values()); tot_t = sum(a[1] for a in agg.values())
print("total warp instr", tot_i, "thread instr", tot_t, "lanes", tot_t / max(1, tot_i))
# by file
byf = collections.Counter(); byft = collections.Counter()
for (f, ln), a in agg.items(): byf[f] += a[0]; byft[f] += a[1]
for f, c in byf.most_common(): print(f"  {f}: {100*c/tot_i:.1f}% warp-instr, lanes {byft[f]/max(1,c):.1f}")
for (f, ln), a in sorted(agg.items(), key=lambda x: -x[1][0])[:N]:
    print(f"{f}:{ln} {100*a[0]/tot_i:5.1f}% lanes {a[1]/max(1,a[0]):5.1f}  {a[2]}")
