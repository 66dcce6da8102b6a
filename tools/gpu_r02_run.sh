#!/bin/bash
# Round-2 evidence run, one gpurun call: smoke, GPU tests, the default bench (with extras), the reference arm, then the ncu
# launch list of the bench command (the only profiler pass of this call; full captures: tools/gpu_r02_capture.sh).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_c2.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-400 gpurun_out/bench_ref.json
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_bench.log 2>&1
echo "ncu rc=$?"
