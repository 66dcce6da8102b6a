#!/bin/bash
# Experiment call on the GPU box: GPU tests, then A/B bench lines (value only) for the knobs under test.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
run() { # label, env..., workload
  local label=$1; shift; local w=$1; shift
  env "$@" python bench.py --workload $w --steps 30 --no-cpu-baseline --no-e2e 2> gpurun_out/exp_$label.err | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$label', '%.4e' % d['value'], 'ms', round(d['ms_per_step'],3), 'frac', round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])" | tee -a gpurun_out/exp.log
}
: > gpurun_out/exp.log
run c3_dn c3 A=1
run c3_lea c3 SMMC_BOOT_ADDR_LEA=1
run c3_dn_ilp8 c3 SMMC_ILP=8
run c3_dn_ilp2 c3 SMMC_ILP=2
run c4_auto c4 A=1
run c4_w32 c4 SMMC_WARPS=32
run c4boot_auto c4boot A=1
run c4boot_w32 c4boot SMMC_WARPS=32
run c2 c2 A=1
run c2_w28 c2 SMMC_WARPS=28
run c2_w24 c2 SMMC_WARPS=24
run c3compat c3compat A=1
