"""Static SASS size per source function of one kernel: nvdisasm -g line annotations mapped to the functions of csrc files.
usage: sass_static.py <cubin> <kernel substring>"""
import re, subprocess, sys
from pathlib import Path
cubin, kern = sys.argv[1], sys.argv[2]
root = Path(__file__).resolve().parents[1] / "flac_raster_b200" / "csrc"
out = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
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
files = {}
agg = {}
active = False
fn = "?"
for line in out.split("\n"):
    m = re.match(r"^\s*\.text\.(\S+):", line)
    if m:
        active = kern in m.group(1)
        continue
    if not active:
        continue
    m = re.match(r'^\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        path, ln = m.group(1), int(m.group(2))
        base = path.split("/")[-1]
        if base not in files:
            p = root / base
            files[base] = load(p) if p.exists() else None
        names = files[base]
        fn = names[ln - 1] if names and 0 < ln <= len(names) else base
        continue
    if re.match(r"^\s+/\*[0-9a-f]{4,6}\*/\s+\S", line):
        agg[fn] = agg.get(fn, 0) + 1
tot = sum(agg.values())
print(f"{kern}: {tot} instructions, {tot * 16 / 1024:.0f} KB")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:45]:
    print(f"{k:28s} {v:7d} inst {v * 16 / 1024:7.1f} KB {100 * v / tot:5.1f}%")
