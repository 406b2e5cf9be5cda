"""Sum ncu --page source --csv (SASS view) columns over address ranges: python tools/sass_regions.py file.csv off1 off2 ... (hex offsets)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iA, iS, iI, iSm = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = rows[2:]
base = int(body[0][iA], 16)
cuts = [int(x, 16) for x in sys.argv[2:]] + [1 << 40]
tot_i = sum(int(r[iI]) for r in body); tot_s = sum(int(r[iSm]) for r in body)
print("total warp-instr %d  samples %d" % (tot_i, tot_s))
lo = 0
for c in cuts:
    sel = [r for r in body if lo <= int(r[iA], 16) - base < c]
    if sel:
        ii = sum(int(r[iI]) for r in sel); ss = sum(int(r[iSm]) for r in sel)
        print("  [%5x, %5x) n=%4d  instr %10d (%5.1f %%)  samples %7d (%5.1f %%)" % (lo, min(c, int(body[-1][iA], 16) - base + 16), len(sel), ii, 100 * ii / tot_i, ss, 100 * ss / tot_s))
    lo = c
