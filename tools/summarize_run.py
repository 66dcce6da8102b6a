"""Summarise gpurun_out/ of `bash tools/gpu_r02_run.sh` (+ the captures of tools/gpu_r02_capture.sh) into one text file for profiles/.
usage: python tools/summarize_run.py "<header line>" > profiles/<name>.txt"""
import collections
import csv
import json
import os
import subprocess
import sys

G = "gpurun_out"
print("#", sys.argv[1] if len(sys.argv) > 1 else "GPU run")
print("# commands: gpurun -- 'bash tools/gpu_r02_run.sh', then one gpurun -- 'bash tools/gpu_r02_capture.sh <mode>' per ncu --set full capture\n")
print("## pytest -m gpu")
print("".join(open(f"{G}/pytest_gpu.log").readlines()[-2:]))
print("## smoke")
print(open(f"{G}/smoke.log").read())
print("## bench.py, the driver's default command (device-timed value | e2e through smmc_run with host buffers | kernel ms | frac by restated I_alg | pipe floor | clocks)")
d = json.loads(open(f"{G}/bench_c2.json").read().strip().splitlines()[-1])
r, c = d["roofline"], d["clocks"]
pf = r.get("pipe_floor")
print(f"c2        value {d['value']:.4e} pm/s  e2e {d['e2e']['value']:.4e}  ms/step {d['ms_per_step']:.3f}  kernel_ms {r['kernel_ms']:.3f}  "
      f"i_alg {r['i_alg']}  frac {r['frac']:.3f}  frac_survey_budget {r['frac_survey_budget'] and round(r['frac_survey_budget'], 3)}  "
      f"pipe_floor {pf and (pf['pipe'], round(pf['frac'], 3))}  traffic {r.get('traffic')}  sm_mhz {c['sm_mhz']} reasons {c['reasons']} power_w_max {c['power_w_max']}")
print("   cpu_baseline:", d.get("cpu_baseline"))
print("   gpu_launches:", d.get("gpu_launches"), " result:", d.get("result"))
ex = d.get("extras") or {}
print("\n## extras.workloads (5 timed steps each, same process, after the headline)")
for k, w in (ex.get("workloads") or {}).items():
    rf = w.get("roofline", {})
    print(f"{k:12s} value {w['value']:.4e} {w['unit']}  kernel_ms {w.get('kernel_ms'):.4f}  frac {rf.get('frac'):.3f} ({rf.get('bound')}"
          + (f", {rf.get('achieved'):.0f} of {rf.get('peak')} GB/s, traffic {rf.get('traffic')}" if rf.get("bound") == "hbm" else f", i_alg {rf.get('i_alg')}")
          + f")  pipe_floor_frac {rf.get('pipe_floor_frac')}  e2e {w.get('e2e', {}).get('value')}  vs_ref_gpu_same_box {w.get('vs_ref_gpu_same_box')}  sm_mhz {w.get('sm_mhz')} {w.get('clock_reasons')}")
print("\n## extras.ref_gpu_same_box (the reference's src/simulations.cu compiled for sm_100a, same process, same box)")
for k, v in (ex.get("ref_gpu_same_box") or {}).items():
    print(f"   {k}: {v}")
print("\n## extras.c5 (10^10 paths x 360 months of ONE run)")
print("  ", ex.get("c5"))
print("\n## extras quartiles")
for k in ("quartiles_min_q1_med_q3_max", "quartiles_device_ms", "quartiles_how", "quartiles_riding_along_extra_ms", "quartiles_riding_along_equal"):
    print(f"   {k}: {ex.get(k)}")
for aux in ("reduce_mean", "vector_add"):
    if not os.path.exists(f"{G}/bench_{aux}.json"):
        continue
    d = json.loads(open(f"{G}/bench_{aux}.json").read().strip().splitlines()[-1])
    r = d["roofline"]
    print(f"{aux} value {d['value']:.4e} elements/s  kernel_ms {r['kernel_ms']:.5f}  {r['achieved']:.1f} GB/s of {r['peak']} (hbm, measured copy)  "
          f"frac {r['frac']:.3f}  traffic {r.get('traffic')} vs algorithmic {r.get('algorithmic_bytes')}  e2e {d['e2e']['value']:.4e} elements/s  exact {d['result']['exact']}  "
          f"sm_mhz {d['clocks']['sm_mhz']} {d['clocks']['reasons']}")
d = json.loads(open(f"{G}/bench_ref.json").read().strip().splitlines()[-1])
print("--impl reference:", d["value"], d["cpu_baseline"])
print("\n## ncu launch list of: python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-extras  (gpu__time_duration.sum)")
rows = list(csv.reader(l for l in open(f"{G}/launches_c2.csv") if l.startswith('"')))
h = rows[0]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
    agg.setdefault(r[ki], []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print(f"{len(v):3d} x {k[:90]:90s} avg {sum(v) / len(v):10.1f} us  share {100 * sum(v) / tot:5.2f}%")
traffic = {}
try:  # keep entries of captures that are not part of this run (e.g. reduce_mean)
    traffic = dict(json.load(open("profiles/ncu_traffic.json"))["bytes_per_launch"])
except Exception:
    pass
for m in ["gaussian", "bootstrap", "bootstrap_compat", "sobol_gaussian", "sobol_bootstrap", "reduce_mean", "vector_add"]:
    if not os.path.exists(f"{G}/prof_{m}.ncu-rep"):
        continue
    print(f"\n## ncu --set full --clock-control none: " + (f"python bench.py --workload {m}" if m in ("reduce_mean", "vector_add") else
          f"python tools/run_one.py {m} (the bench launch: 100M x 360; sobol 5M x 1000)"))
    sys.stdout.flush()
    subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "ncu_summary.py"), f"{G}/prof_{m}.ncu-rep"])
    raw = subprocess.run(["ncu", "-i", f"{G}/prof_{m}.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    tot = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = hdr.index(k)
        v = float(vals[i].replace(",", ""))
        tot += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
    traffic[m] = int(tot)
# dram bytes (read + write) of the simulation kernel per launch at the bench size: bench.py's roofline.traffic
json.dump({"source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, one launch of the bench workload",
           "bytes_per_launch": traffic}, open("profiles/ncu_traffic.json", "w"), indent=1)
