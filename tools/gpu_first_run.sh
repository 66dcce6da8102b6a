#!/bin/bash
# One gpurun call: smoke, GPU tests, GPU golden fixture, benches, then ncu (launch list + full captures).
# usage (from the repo root, on the GPU box): bash tools/gpu_first_run.sh
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python tests/golden/make_golden_gpu.py gpurun_out/ref_gpu_compat.npz > gpurun_out/golden.log 2>&1; echo "golden rc=$?"
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cat gpurun_out/bench_c2.json
for w in c3 c3compat c4 c4boot; do
  python bench.py --workload $w --steps 30 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; cat gpurun_out/bench_$w.json
done
python bench.py --workload reduce_mean > gpurun_out/bench_reduce_mean.json 2> gpurun_out/bench_reduce_mean.err; cat gpurun_out/bench_reduce_mean.json
python bench.py --workload vector_add > gpurun_out/bench_vector_add.json 2> gpurun_out/bench_vector_add.err; cat gpurun_out/bench_vector_add.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
# ---- ncu: launch list of the bench command, then one full capture of the top kernel per mode
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_bench.log 2>&1
for m in gaussian bootstrap bootstrap_compat; do
  python tools/run_one.py $m 100000000 360 > gpurun_out/plain_$m.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:mc_kernel -s 1 -c 1 -f -o gpurun_out/prof_$m \
      python tools/run_one.py $m 100000000 360 > gpurun_out/ncu_$m.log 2>&1
done
for m in sobol_gaussian sobol_bootstrap; do
  python tools/run_one.py $m 5000000 1000 > gpurun_out/plain_$m.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:mc_kernel -s 1 -c 1 -f -o gpurun_out/prof_$m \
      python tools/run_one.py $m 5000000 1000 > gpurun_out/ncu_$m.log 2>&1
done
python bench.py --workload reduce_mean --steps 5 --warmup 3 > gpurun_out/plain_reduce_mean.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:reduce_sum_tma -s 2 -c 1 -f -o gpurun_out/prof_reduce_mean \
    python bench.py --workload reduce_mean --steps 5 --warmup 3 > gpurun_out/ncu_reduce_mean.log 2>&1
ls -la gpurun_out
