#!/bin/bash
# Iteration call on the GPU box: GPU tests, ILP sweep, benches, ncu of the two Philox kernels.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for m in gaussian bootstrap; do for ilp in 1 2 4 8; do SMMC_ILP=$ilp python tools/run_one.py $m 100000000 360; done; done 2>&1 | tee gpurun_out/ilp_sweep.log
for w in c2 c3 c3compat; do
  python bench.py --workload $w --steps 40 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; cat gpurun_out/bench_$w.json
done
for m in gaussian bootstrap; do
  python tools/run_one.py $m 20000000 360 > gpurun_out/plain_$m.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:mc_kernel -s 1 -c 1 -f -o gpurun_out/prof_$m \
      python tools/run_one.py $m 20000000 360 > gpurun_out/ncu_$m.log 2>&1
done
