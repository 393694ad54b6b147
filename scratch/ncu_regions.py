# aggregate ncu source-page instruction/sample counts by function-ish region (line ranges given by markers in the source)
import csv, subprocess, sys, re
rep = sys.argv[1]; srcfile = sys.argv[2] if len(sys.argv) > 2 else '/root/repo/flac_raster_b200/csrc/frx_encode.cu'
src = subprocess.run(['ncu','-i',rep,'--page','source','--print-source','cuda,sass','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
cur=None; H=None; per={}
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>2 and r[0]=='Line No': H=r; continue
    if H is None or len(r)<10 or r[2]!='-': continue
    try: ln=int(r[0]); samples=int(r[4]); inst=int(r[7])
    except: continue
    per[(cur,ln)]=(samples,inst)
lines=open(srcfile).read().split('\n')
# regions: start at lines matching a function definition or a '// ----' stage comment inside the kernel
marks=[]
for i,l in enumerate(lines,1):
    if re.match(r'^(template|__device__|__global__|static|struct)', l) or re.search(r'// ---- ', l) or re.search(r'compute_sync\(\);\s*//', l):
        marks.append((i,l.strip()[:90]))
ts=sum(v[0] for v in per.values()); ti=sum(v[1] for v in per.values())
fn=srcfile.split('/')[-1]
import bisect
starts=[m[0] for m in marks]
agg={}
other=[0,0]
for (f,ln),(s,i) in per.items():
    if f!=fn: other[0]+=s; other[1]+=i; continue
    k=bisect.bisect_right(starts,ln)-1
    key=marks[k] if k>=0 else (0,'<top>')
    a=agg.setdefault(key,[0,0]); a[0]+=s; a[1]+=i
print('total samples',ts,'inst',ti)
for key,(s,i) in sorted(agg.items()):
    if s*1000<ts and i*1000<ti: continue
    print(f"{key[0]:5d} samp {100*s/ts:5.1f}% inst {100*i/ti:5.1f}%  {key[1]}")
print(f"other files samp {100*other[0]/ts:5.1f}% inst {100*other[1]/ti:5.1f}%")
