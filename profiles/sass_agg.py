"""Aggregate an `ncu --page source --csv` dump (SASS view) by opcode: warp instructions, lane efficiency, stall samples.
usage: python profiles/sass_agg.py dump.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ie, te, s = hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('# Samples')
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
nk = 0
for r in rows[2:]:
    if r and r[0] == 'Kernel Name':
        nk += 1
        if nk >= 1:
            break
    if len(r) <= te or not r[ie].isdigit():
        continue
    op = r[1].strip().split()
    if not op:
        continue
    o = op[1] if op[0].startswith('@') else op[0]
    o = o.split('.')[0]
    a = agg[o]
    a[0] += int(r[ie]); a[1] += int(r[te]); a[2] += int(r[s])
    tot[0] += int(r[ie]); tot[1] += int(r[te]); tot[2] += int(r[s])
print("warp-inst %.2fM thread-inst %.2fM lanes/inst %.2f samples %d" % (tot[0] / 1e6, tot[1] / 1e6, tot[1] / tot[0], tot[2]))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:28]:
    print(f"{k:12s} inst={v[0]/1e6:8.2f}M share={v[0]/tot[0]:.3f} lanes/inst={v[1]/max(v[0],1):5.1f} stall-samples={v[2]/max(tot[2],1):.3f}")
