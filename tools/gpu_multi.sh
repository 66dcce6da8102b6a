#!/bin/bash
# usage: bash tools/gpu_multi.sh <n_gpus>   (on a box with that many GPUs)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_L.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_multi.log
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then
      python bench.py --gpus 1 --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n --steps 40 --warmup 5 > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    fi
    echo "N=$n rc=$?"; tail -c 1500 gpurun_out/scale_$n.json; tail -3 gpurun_out/scale_$n.err
  fi
done
