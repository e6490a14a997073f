"""Top stall instructions of an `ncu --page source --csv` dump:  ncu_src_top.py <csv> [n] [lo hi]  (row range optional)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix, isrc, isamp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [(int(r[isamp] or 0), int(r[ix] or 0), r[isrc], r, n) for n, r in enumerate(rows[2:]) if len(r) > ix]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
if len(sys.argv) > 4:
    data = data[int(sys.argv[3]):int(sys.argv[4])]
tot = sum(d[0] for d in data)
print("samples", tot, "instructions", sum(d[1] for d in data))
agg = {}
for d in data:
    for i in stalls:
        agg[hdr[i]] = agg.get(hdr[i], 0) + int(d[3][i] or 0)
print(sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
for d in sorted(data, key=lambda d: -d[0])[:n]:
    st = sorted(((int(d[3][i] or 0), hdr[i]) for i in stalls), reverse=True)[:2]
    print("%5d %5.1f%% row %4d inst=%9d  %-58s %s" % (d[0], 100.0 * d[0] / max(tot, 1), d[4], d[1], d[2][:58], st))
