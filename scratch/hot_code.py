"""Hot code footprint per source function: executed-instruction counts of an ncu capture (sass csv) joined with nvdisasm -g
line info of the matching cubin.  usage: hot_code.py <sass.csv> <cubin> <kernel substring>"""
import csv, re, subprocess, sys
from pathlib import Path
csvf, cubin, kern = sys.argv[1:4]
root = Path(__file__).resolve().parents[1] / "flac_raster_b200" / "csrc"


def load(p):
    names, cur = [], "?"
    for line in Path(p).read_text().split("\n"):
        m = re.match(r"^(?:template.*>\s*)?(?:static\s+)?(?:__device__|__global__|__host__)[^(]*?([A-Za-z_][A-Za-z0-9_]*)\s*\(", line)
        if m and not line.startswith(" "):
            cur = m.group(1)
        m2 = re.match(r"^struct\s+([A-Za-z0-9_]+)", line)
        if m2:
            cur = m2.group(1)
        names.append(cur)
    return names


out = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
files, fn_at, active, fn = {}, {}, False, "?"
for line in out.split("\n"):
    m = re.match(r"^\s*\.text\.(\S+):", line)
    if m:
        active = kern in m.group(1)
        continue
    if not active:
        continue
    m = re.match(r'^\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        base, ln = m.group(1).split("/")[-1], int(m.group(2))
        if base not in files:
            p = root / base
            files[base] = load(p) if p.exists() else None
        names = files[base]
        fn = names[ln - 1] if names and 0 < ln <= len(names) else base
        continue
    m = re.match(r"^\s+/\*([0-9a-f]{4,6})\*/\s+\S", line)
    if m:
        fn_at[int(m.group(1), 16)] = fn
rows = list(csv.reader(open(csvf)))
H = rows[1]
ia = H.index("Instructions Executed")
isamp = H.index("# Samples")
base = None
agg = {}
for r in rows[2:]:
    try:
        addr = int(r[0], 16)
        n = int(r[ia])
        s = int(r[isamp])
    except Exception:
        continue
    if base is None:
        base = addr
    f = fn_at.get(addr - base, "?")
    a = agg.setdefault(f, [0, 0, 0, 0])
    a[0] += 1
    a[1] += n > 0
    a[2] += n
    a[3] += s
tot = sum(v[2] for v in agg.values())
th = sum(v[1] for v in agg.values())
ts = sum(v[3] for v in agg.values())
print(f"executed code {th * 16 / 1024:.0f} KB of {sum(v[0] for v in agg.values()) * 16 / 1024:.0f} KB")
print(f"{'function':28s} {'static KB':>9s} {'hot KB':>7s} {'dyn %':>6s} {'samp %':>6s} {'exec/inst':>10s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{k:28s} {v[0] * 16 / 1024:9.1f} {v[1] * 16 / 1024:7.1f} {100 * v[2] / tot:6.1f} {100 * v[3] / ts:6.1f} {v[2] / max(1, v[1]):10.0f}")
