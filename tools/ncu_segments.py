"""Summarise an `ncu --page source --csv` export: stall samples per barrier-delimited SASS segment."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if '# Samples' in r][0]
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) >= len(hdr)]
si = hdr.index('# Samples')
names = ['stall_barrier', 'stall_long_sb', 'stall_math', 'stall_short_sb', 'stall_wait', 'stall_mio']
idx = [hdr.index(n) for n in names]
tot = sum(int(r[si]) for r in data)
print('total samples', tot)
cur = [0] * (len(names) + 1); start = 0; label = 'start'
def flush(end):
    if cur[0] > tot * 0.002:
        print(f"{start:5d}-{end:5d} {cur[0]/tot*100:5.1f}% | " + ' '.join(f"{n[6:]}={c/tot*100:4.1f}" for n, c in zip(names, cur[1:])) + f" | after {label[:30]}")
for n, r in enumerate(data):
    ins = r[1].strip()
    cur[0] += int(r[si])
    for k, i in enumerate(idx): cur[k + 1] += int(r[i] or 0)
    if 'BAR.SYNC' in ins or 'EXIT' in ins:
        flush(n); cur = [0] * (len(names) + 1); start = n; label = ins
flush(len(data))
if len(sys.argv) > 2:
    top = sorted(data, key=lambda r: -int(r[si]))[:int(sys.argv[2])]
    for r in top: print(r[si], r[1].strip()[:70])
