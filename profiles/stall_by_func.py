"""Per source function: executed warp instructions and warp-stall samples by reason (joins ncu SASS view with nvdisasm -g)."""
import csv,re,subprocess,glob,collections,sys
# usage: python profiles/stall_by_func.py dump.csv cubin kernel_substring [mangled template token, e.g. ILi1E / ILb0E] [section]
# dump.csv = `ncu -i rep --page source --csv --kernel-name regex:<kernel>`; one section per profiled launch.
dump,cubin,k=sys.argv[1:4]
inst=sys.argv[4] if len(sys.argv)>4 else ''
want=int(sys.argv[5]) if len(sys.argv)>5 else 0
allrows=list(csv.reader(open(dump)))
secs=[];curs=None
for r in allrows:
    if r and r[0]=='Kernel Name':
        curs=[r]; secs.append(curs)
    elif curs is not None: curs.append(r)
tok=({'i':'(int)','b':'(bool)'}[inst[2]]+inst[3:-1]+'>') if inst else ''  # ILi1E -> (int)1>
secs=[x for x in secs if tok in x[0][1]]
rows=secs[want]
print('#',rows[0][1])
hdr=rows[1]
ie=hdr.index('Instructions Executed')
cols=['# Samples','stall_long_sb','stall_short_sb','stall_wait','stall_no_inst','stall_branch_resolving','stall_barrier','stall_lg']
ci=[hdr.index(c) for c in cols]
data=[r for r in rows[2:] if len(r)>46 and r[ie].isdigit()]
cub=cubin
out=subprocess.run(["nvdisasm","-g","-c",cub],capture_output=True,text=True).stdout.splitlines()
lines=[];cur=None;active=False
for ln in out:
    if ln.startswith("//---------------------") and ".text." in ln:
        active=k in ln and inst in ln; continue
    if not active: continue
    m=re.search(r'//## File "([^"]+)", line (\d+)',ln)
    if m: cur=(m.group(1),int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/",ln): lines.append(cur)
funcs={}
def func_of(fl):
    if fl is None: return "?"
    f,l=fl
    if f not in funcs:
        tbl=[]
        try:
            for i,t in enumerate(open(f),1):
                m=re.match(r"\s*(?:HRG_HD|__global__|__device__|static|inline|template).*?\b(\w+)\s*\(",t)
                if m and not t.strip().startswith("//"): tbl.append((i,m.group(1)))
        except OSError: pass
        funcs[f]=tbl
    name="?"
    for i,nm in funcs[f]:
        if i<=l: name=nm
        else: break
    return f.split("/")[-1]+":"+name
agg=collections.defaultdict(lambda:[0]*(len(cols)+1))
n=min(len(lines),len(data))
for i in range(n):
    a=agg[func_of(lines[i])]
    a[0]+=int(data[i][ie])
    for q,c in enumerate(ci): a[q+1]+=int(data[i][c])
tot=sum(v[1] for v in agg.values())
print('func'.ljust(44),'inst(M)','samples%',' '.join(c.replace('stall_','')[:8].rjust(8) for c in cols[1:]))
for kk,v in sorted(agg.items(),key=lambda kv:-kv[1][1])[:18]:
    print(kk.ljust(44), ('%.2f'%(v[0]/1e6)).rjust(7), ('%.3f'%(v[1]/tot)).rjust(8), ' '.join(str(x).rjust(8) for x in v[2:]))
