"""Summarise an .ncu-rep (raw + source pages) into the CSV / text kept under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof_eo.ncu-rep profiles/r01_ncu_eo_N64_B256
"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "pipe_fp64", "sm__throughput.avg.pct", "bank_conflicts",
        "smsp__issue_active.avg.pct", "sm__inst_executed_pipe", "smsp__average_warps_issue_stalled", "sm__inst_executed.sum",
        "sm__cycles", "l1tex__data_pipe_lsu_wavefronts_mem_shared", "lts__t_bytes.sum", "smsp__inst_executed.sum"]
with open(out + "_raw.csv", "w") as f:
    w = csv.writer(f)
    w.writerow(["metric", "unit", "value"])
    for h, u, v in zip(hdr, units, vals):
        if any(k in h for k in keep):
            w.writerow([h, u, v])
rows = list(csv.reader(io.StringIO(src)))
h2, data = rows[1], rows[2:]
isrc, ist, ie = h2.index("Source"), h2.index("Warp Stall Sampling (All Samples)"), h2.index("Instructions Executed")
iw, iwi = h2.index("L1 Wavefronts Shared"), h2.index("L1 Wavefronts Shared Ideal")
ops, st, wf, wfi = (collections.Counter() for _ in range(4))
for r in data:
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[isrc])
    op = m.group(2).split(".")[0] if m else "?"
    ops[op] += int(r[ie] or 0); st[op] += int(r[ist] or 0); wf[op] += int(r[iw] or 0); wfi[op] += int(r[iwi] or 0)
tot, tote = sum(st.values()), sum(ops.values())
d = dict(zip(hdr, vals))
with open(out + "_summary.txt", "w") as f:
    f.write(f"kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}\n")
    for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
              "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
              "smsp__issue_active.avg.pct_of_peak_sustained_active"):
        f.write(f"{k}: {d.get(k)} {units[hdr.index(k)] if k in hdr else ''}\n")
    for k in hdr:
        if "smsp__average_warps_issue_stalled" in k and "not_issued" not in k and float(d[k] or 0) > 0.2:
            f.write(f"{k}: {d[k]}\n")
    f.write(f"stall samples {tot}, warp instructions {tote}\nopcode  stall%  inst%  smem_wavefronts  ideal\n")
    for op, c in st.most_common(14):
        f.write(f"{op:8s} {100 * c / tot:6.2f} {100 * ops[op] / tote:6.2f} {wf[op]:12d} {wfi[op]:12d}\n")
print(open(out + "_summary.txt").read())
