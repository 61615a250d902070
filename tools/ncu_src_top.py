"""Top stall lines / per-region sample shares of an `ncu --page source --csv` dump (gzip)."""
import csv, gzip, io, sys
rows = list(csv.reader(io.TextIOWrapper(gzip.open(sys.argv[1]))))
hdr = rows[1]; data = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
base = int(data[0][0], 16)
tot = sum(int(r[ci['# Samples']]) for r in data)
toti = sum(int(r[ci['Instructions Executed']]) for r in data)
print('total samples', tot, 'warp-instructions', toti)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
for r in sorted(data, key=lambda r: -int(r[ci['# Samples']]))[:n]:
    print('%7s %11s  %05x  %s' % (r[ci['# Samples']], r[ci['Instructions Executed']], int(r[0], 16) - base, r[1].strip()[:90]))
if len(sys.argv) > 3:
    lo, hi = int(sys.argv[3], 16), int(sys.argv[4], 16)
    for r in data:
        off = int(r[0], 16) - base
        if lo <= off <= hi:
            print('%05x %6s %10s %s' % (off, r[ci['# Samples']], r[ci['Instructions Executed']], r[1].strip()[:100]))
