"""Per CUDA source line totals from `ncu --page source --print-source cuda,sass --csv`: python tools/src_lines.py file.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
cur_file = ""
hdr = None
agg = {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; iI = hdr.index("Instructions Executed"); iS = hdr.index("# Samples"); continue
    if hdr and len(r) == len(hdr) and r[0].isdigit() and r[2] == "-":   # a CUDA source line row (no SASS address)
        try:
            ii, ss = int(r[iI]), int(r[iS])
        except ValueError:
            continue
        k = (cur_file, int(r[0]))
        a = agg.setdefault(k, [0, 0, r[1].strip()[:90]])
        a[0] += ii; a[1] += ss
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print("total instr %d samples %d" % (ti, ts))
for k in sorted(agg):
    a = agg[k]
    if 100.0 * a[0] / ti >= minp or 100.0 * a[1] / ts >= minp:
        print("%-16s %5d  i %5.1f%%  s %5.1f%%  %s" % (k[0], k[1], 100.0 * a[0] / ti, 100.0 * a[1] / ts, a[2]))
