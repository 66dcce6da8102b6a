"""Per-opcode share of warp-stall samples and executed instructions of a kernel, from the SASS source page of an
.ncu-rep (development tool; summaries go to profiles/).   usage: python tools/ncu_hot_sass.py <file.ncu-rep>"""
import collections
import csv
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
print("kernel:", rows[0][1])
hdr = rows[1]
i_src, i_smp, i_exe = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
smp, exe = collections.Counter(), collections.Counter()
for r in rows[2:]:
    if len(r) <= max(i_src, i_smp, i_exe):
        continue
    toks = r[i_src].split()
    if not toks:
        continue
    op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    op = ".".join(op.split(".")[:2]) if op.startswith(("MUFU", "IMAD", "FFMA", "FMUL", "LDS", "SHFL", "I2FP", "F2I")) else op.split(".")[0]
    smp[op] += int(r[i_smp] or 0)
    exe[op] += int(r[i_exe] or 0)
ts, te = sum(smp.values()), sum(exe.values())
print(f"{'opcode':16s} {'executed %':>10s} {'stall samples %':>16s}   (samples land on the instruction a warp is WAITING TO ISSUE)")
for op, n in exe.most_common(14):
    print(f"{op:16s} {100.0 * n / te:10.1f} {100.0 * smp[op] / max(ts, 1):16.1f}")
