"""Instruction and stall-sample share per source function from an .ncu-rep (source page), for files under csrc/."""
import csv, re, subprocess, sys
from pathlib import Path
rep = sys.argv[1]
root = Path(__file__).resolve().parents[1] / "flac_raster_b200" / "csrc"
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
fn_of = {}
def load(fname):
    p = root / fname
    if not p.exists():
        return None
    names, cur = [], "?"
    for line in p.read_text().split("\n"):
        m = re.match(r"^(?:template.*>\s*)?(?:static\s+)?(?:__device__|__global__|__host__)[^(]*?([A-Za-z_][A-Za-z0-9_]*)\s*\(", line)
        if m and not line.startswith(" "):
            cur = m.group(1)
        m2 = re.match(r"^struct\s+([A-Za-z0-9_]+)", line)
        if m2:
            cur = m2.group(1)
        names.append(cur)
    return names
cur = None; H = None; agg = {}
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]
        if cur not in fn_of: fn_of[cur] = load(cur)
        continue
    if len(r) > 2 and r[0] == 'Line No': H = r; continue
    if H is None or len(r) < 10 or r[2] != '-': continue
    try: ln = int(r[0]); samples = int(r[4]); inst = int(r[7])
    except Exception: continue
    names = fn_of.get(cur)
    fn = names[ln - 1] if names and 0 < ln <= len(names) else cur
    a = agg.setdefault(fn, [0, 0])
    a[0] += samples; a[1] += inst
ts = sum(v[0] for v in agg.values()); ti = sum(v[1] for v in agg.values())
print(f"total samples {ts} inst {ti}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{k:28s} inst {100*v[1]/ti:5.1f}%  samples {100*v[0]/ts:5.1f}%  cpi-ish {v[0]/max(1,v[1])*ti/ts:5.2f}")
