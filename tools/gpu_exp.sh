#!/bin/bash
# Experiment call on the GPU box: GPU tests, then bench lines (value only) for the workloads / knobs under test.
# usage: bash tools/gpu_exp.sh "label:workload[:ENV=V ...]" ...
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/exp.log
for spec in "$@"; do
  IFS=: read -r label w envs <<< "$spec"
  env $envs python bench.py --workload $w --steps 30 --no-cpu-baseline --no-e2e 2> gpurun_out/exp_$label.err | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$label', '%.4e' % d['value'], 'ms', round(d['ms_per_step'],3), 'frac', round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])" | tee -a gpurun_out/exp.log
done
