"""Stall samples of one kernel per CUDA source line: joins the SASS page of an .ncu-rep with nvdisasm's line table.

    python scripts/ncu_lines.py gpurun_out/prof.ncu-rep _ZN3ldc12k_stream_tmaENS_8S2ParamsE [top]

Needs the library built from the same sources (-lineinfo).  Prints the lines with most samples and writes nothing."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, sym = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
lib = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "02689-advancednumericalalgorithmp3_b200", "ldc_b200",
                   "libldc_b200.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(f".text.{sym}:"))
line_of = {}   # instruction offset -> (file, line)
cur = ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("\t.section") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        # keep the outermost non-inlined position? the first annotation before an instruction is the innermost
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        line_of[int(m.group(1), 16)] = cur
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ia, isrc, ist, ie = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
data = rows[2:]
base = int(data[0][ia], 16)
per_line, inst_line = collections.Counter(), collections.Counter()
ops_at = collections.defaultdict(collections.Counter)
for r in data:
    off = int(r[ia], 16) - base
    key = line_of.get(off, ("?", 0))
    n = int(r[ist] or 0)
    per_line[key] += n
    inst_line[key] += int(r[ie] or 0)
    op = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[isrc])
    ops_at[key][op.group(2).split(".")[0] if op else "?"] += n
tot = sum(per_line.values())
print(f"total stall samples {tot}")
for (f, ln), n in per_line.most_common(top):
    print(f"{100 * n / tot:6.2f}%  {f}:{ln}  inst {inst_line[(f, ln)]}  {dict(ops_at[(f, ln)].most_common(3))}")
