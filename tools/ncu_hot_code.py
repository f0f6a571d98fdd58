"""Hot instruction footprint of a kernel by source range, from an ncu report (the ensemble kernel is
bound by instruction fetch: what matters is how many distinct SASS instructions a block's loop body
executes, against the 32 KB L1.5 instruction cache).

usage: ncu -i report.ncu-rep --page source --csv > src.csv
       ncu_hot_code.py src.csv <object.o> <ranges.txt> [min executions]
ranges.txt holds a Python dict {"name": (first line, last line), ...} over the .cu file; the object
must be the -lineinfo build the report was taken from."""
import csv,re,subprocess,os,sys,tempfile,collections,io
src_csv,obj,rangesf=sys.argv[1:4]
rows=list(csv.reader(open(src_csv)))
hi=next(i for i,r in enumerate(rows) if "Source" in r and "# Samples" in r)
h=rows[hi]; ie=h.index("Instructions Executed")
body=[r for r in rows[hi+1:] if len(r)==len(h)]
base=int(body[0][0],16)
tmp=tempfile.mkdtemp()
subprocess.run(["cuobjdump","-xelf","all",os.path.abspath(obj)],cwd=tmp,capture_output=True)
cubin=[f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass=subprocess.run(["nvdisasm","-g","-c",os.path.join(tmp,cubin)],capture_output=True,text=True).stdout
ranges=eval(open(rangesf).read())
fn=None;cur=None;loc={}
for ln in sass.splitlines():
    m=re.match(r"\s*\.text\.(\S+):",ln)
    if m: fn=m.group(1); continue
    if '//## File' in ln:
        m=re.search(r'File "([^"]*)", line (\d+)',ln); cur=(os.path.basename(m.group(1)),int(m.group(2))); continue
    m=re.match(r"\s+/\*([0-9a-f]+)\*/\s+\S",ln)
    if m and fn and 'k_ensembleILi128' in fn: loc[int(m.group(1),16)]=cur
thr=float(sys.argv[4]) if len(sys.argv)>4 else 1184*311*0.5
cnt=collections.Counter(); files=collections.Counter()
for r in body:
    a=int(r[0],16)-base; c=int(r[ie])
    if c<thr: continue
    l=loc.get(a); key='other'
    if l and l[0]=='ensemble.cu':
        for name,(x,y) in ranges.items():
            if x<=l[1]<=y: key=name;break
    elif l: key='other:'+l[0]
    cnt[key]+=1
print("hot instrs",sum(cnt.values()))
for k,v in cnt.most_common(): print(f"   {k:40s} {v:5d} instr {v*16/1024:5.1f} KB")
