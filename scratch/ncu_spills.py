import csv, subprocess, sys
rep=sys.argv[1]
src = subprocess.run(['ncu','-i',rep,'--page','source','--print-source','cuda,sass','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
cur=None; H=None; curline=None; agg={}
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>2 and r[0]=='Line No': H=r; continue
    if H is None or len(r)<10: continue
    if r[2]=='-':
        curline=(cur,r[0],r[1]); continue
    sass=r[3]
    if 'STL' in sass or 'LDL' in sass:
        try: inst=int(r[7])
        except: inst=0
        k=curline
        a=agg.setdefault(k,[0,0,0,0])
        if 'STL' in sass: a[0]+=1; a[1]+=inst
        else: a[2]+=1; a[3]+=inst
for k,v in sorted(agg.items(), key=lambda kv:-(kv[1][1]+kv[1][3]))[:30]:
    print(f"{k[0][:14]}:{k[1]:>5} STL {v[0]:3d} ({v[1]:8d}) LDL {v[2]:3d} ({v[3]:8d})  {k[2][:90]}")
