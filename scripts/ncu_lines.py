#!/usr/bin/env python
"""Stall samples of one kernel of an .ncu-rep aggregated by CUDA source line (no GPU needed).
  python scripts/ncu_lines.py <rep> <object.o> <mangled-name-substring> [top]
The SASS page of the report and `nvdisasm -g` of the object list the kernel's instructions in the same order."""
import csv, io, re, subprocess, sys
from collections import defaultdict

rep, obj, pat = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; data = rows[2:]
isamp, isrc = hdr.index("# Samples"), hdr.index("Source")
import glob, os, tempfile
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
dis = subprocess.run(["nvdisasm", "-g", "-c", glob.glob(tmp + "/*.cubin")[0]], capture_output=True, text=True).stdout.splitlines()
lines, cur, infn = [], None, False
for ln in dis:
    if ln.startswith("\t.section") or ln.startswith(".section"):
        infn = pat in ln and ".text." in ln
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", ln):
        lines.append(cur)
print(f"sass rows {len(data)}  disasm instructions {len(lines)}")
agg = defaultdict(int)
for i, r in enumerate(data):
    s = int(r[isamp] or 0)
    if s and i < len(lines):
        agg[lines[i]] += s
tot = sum(agg.values())
src_cache = {}
def text(f, n):
    if f not in src_cache:
        try:
            import glob
            p = glob.glob(f"sglang_custom_b200/csrc/{f}")[0]
            src_cache[f] = open(p).read().splitlines()
        except Exception:
            src_cache[f] = []
    L = src_cache[f]
    return L[n - 1].strip()[:110] if 0 < n <= len(L) else ""
for (k, s) in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    f, n = k if k else ("?", 0)
    print(f"{s / tot * 100:5.1f}%  {f}:{n:<5d} {text(f, n)}")
