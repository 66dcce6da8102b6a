#!/bin/bash
# One ncu --set full capture per gpurun call.  usage: bash tools/gpu_r02_capture.sh <mode|reduce_mean|vector_add> [run_one args...]
m=$1; shift
mkdir -p gpurun_out
case $m in
  reduce_mean|vector_add)
    k=$([ $m = reduce_mean ] && echo reduce_sum || echo vector_add)
    python bench.py --workload $m --steps 5 --warmup 3 > gpurun_out/plain_$m.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/prof_$m \
        python bench.py --workload $m --steps 5 --warmup 3 > gpurun_out/ncu_$m.log 2>&1 ;;
  *)
    case $m in sobol_*) n=5000000; p=1000;; *) n=100000000; p=360;; esac
    tag=$m$(echo "$*" | tr -c 'a-z0-9=\n' '_' | sed 's/_*$//; s/^\(.\)/_\1/')
    python tools/run_one.py $m $n $p "$@" > gpurun_out/plain_$tag.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:mc_kernel -s 1 -c 1 -f -o gpurun_out/prof_$tag \
        python tools/run_one.py $m $n $p "$@" > gpurun_out/ncu_$tag.log 2>&1 ;;
esac
echo "capture rc=$?"; ls -la gpurun_out/prof_* | tail -3; tail -2 gpurun_out/plain_*.log | tail -4
