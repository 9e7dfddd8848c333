"""Join an `ncu --page source --csv` SASS dump with `nvdisasm -g` line info and aggregate executed warp instructions
per source function. usage: python profiles/sass_lines.py dump.csv cubin kernel_substring [top]"""
import collections
import csv
import re
import subprocess
import sys

dump, cubin, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
out = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
lines, cur, active = [], None, False
for ln in out:
    if ln.startswith("//---------------------") and ".text." in ln:
        active = kname in ln
        continue
    if not active:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
rows = list(csv.reader(open(dump)))
hdr = rows[1]
ie, s = hdr.index('Instructions Executed'), hdr.index('# Samples')
data = []
for r in rows[2:]:
    if r and r[0] == 'Kernel Name':
        break
    if len(r) > s and r[ie].isdigit():
        data.append((int(r[ie]), int(r[s])))
print("sass instructions: nvdisasm %d, ncu %d" % (len(lines), len(data)))
n = min(len(lines), len(data))
funcs = {}


def func_of(fl):
    if fl is None:
        return "?"
    f, l = fl
    if f not in funcs:
        tbl = []
        try:
            for i, t in enumerate(open(f), 1):
                m = re.match(r"\s*(?:HRG_HD|__global__|__device__|static|inline|template).*?\b(\w+)\s*\(", t)
                if m and not t.strip().startswith("//"):
                    tbl.append((i, m.group(1)))
        except OSError:
            pass
        funcs[f] = tbl
    name = "?"
    for i, nm in funcs[f]:
        if i <= l:
            name = nm
        else:
            break
    return f.split("/")[-1] + ":" + name


agg = collections.defaultdict(lambda: [0, 0])
tot = [0, 0]
for i in range(n):
    k = func_of(lines[i])
    agg[k][0] += data[i][0]
    agg[k][1] += data[i][1]
    tot[0] += data[i][0]
    tot[1] += data[i][1]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k:55s} inst={v[0]/1e6:8.2f}M share={v[0]/max(tot[0],1):.3f} stall-samples={v[1]/max(tot[1],1):.3f}")
if len(sys.argv) > 5:  # per source line detail: python sass_lines.py dump cubin kernel top by_line
    agl = collections.defaultdict(lambda: [0, 0])
    for i in range(n):
        fl = lines[i]
        k = (fl[0].split("/")[-1], fl[1]) if fl else ("?", 0)
        agl[k][0] += data[i][0]
        agl[k][1] += data[i][1]
    print("--- by source line")
    for k, v in sorted(agl.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{k[0]}:{k[1]:<5d} inst={v[0]/1e6:8.2f}M share={v[0]/max(tot[0],1):.3f} stall-samples={v[1]/max(tot[1],1):.3f}")
