"""Top stall lines from an `ncu --page source --csv` dump: python tools/ncu_stalls.py file.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
print(rows[0][1][:100])
hdr = rows[1]
si = hdr.index('# Samples'); so = hdr.index('Source'); ie = hdr.index('Instructions Executed')
stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = rows[2:]
tot = sum(int(r[si] or 0) for r in data)
print('total samples', tot)
agg = {}
for r in data:
    for j in stalls:
        agg[hdr[j]] = agg.get(hdr[j], 0) + int(r[j] or 0)
print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
top = sorted(range(len(data)), key=lambda i: -int(data[i][si] or 0))[:topn]
for i in sorted(top):
    r = data[i]
    st = sorted(((int(r[j] or 0), hdr[j][6:]) for j in stalls), reverse=True)[:2]
    print(i, r[si], r[ie], r[so][:80], st)
