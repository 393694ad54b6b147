"""per-source-line instruction / sample shares for a line range of a file: ncu_range.py rep file lo hi"""
import csv, subprocess, sys
rep, fname, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
cur = None; H = None; tot_i = tot_s = 0; out = []; nsass = {}
curline = None
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) > 2 and r[0] == 'Line No': H = r; continue
    if H is None or len(r) < 10: continue
    if r[2] == '-':
        try: ln = int(r[0]); samples = int(r[4]); inst = int(r[7])
        except Exception: continue
        tot_i += inst; tot_s += samples
        curline = (cur, ln)
        if cur == fname and lo <= ln <= hi: out.append((ln, inst, samples, r[1]))
    else:
        nsass[curline] = nsass.get(curline, 0) + 1
si = ss = 0
for ln, inst, samples, text in out:
    si += inst; ss += samples
    print(f"{ln:5d} inst {100*inst/tot_i:5.2f}% samp {100*samples/tot_s:5.2f}% sass {nsass.get((fname, ln), 0):4d}  {text[:110]}")
print(f"range total inst {100*si/tot_i:.1f}% samples {100*ss/tot_s:.1f}%")
