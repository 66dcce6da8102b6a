#!/bin/bash
# usage: bash tools/gpu_multi.sh <n_gpus>   (on a box with that many GPUs)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_L.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_multi.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29507 \
  tools/multi_check.py > gpurun_out/multi_check_$N.log 2>&1; echo "multi_check rc=$?"; grep multi_check gpurun_out/multi_check_$N.log | tail -4
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then
      python bench.py --gpus 1 --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n --steps 40 --warmup 5 > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    fi
    echo "N=$n rc=$?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/scale_$n.json').read().strip().splitlines()[-1])
print('n_gpus',d['n_gpus'],'value %.4e'%d['value'],'ms/step',round(d['ms_per_step'],3),'e2e %.4e'%d['e2e']['value'],'clocks',d['clocks']['sm_mhz'],d['clocks']['reasons'])"
  fi
done
# C5 (360 months x 10^10 paths of ONE run, strong scaling): 1 GPU and N GPUs
python bench.py --gpus 1 --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c5_1.json 2> gpurun_out/c5_1.err; echo "c5 N=1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
  bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c5_$N.json 2> gpurun_out/c5_$N.err; echo "c5 N=$N rc=$?"
for n in 1 $N; do python -c "
import json
d=json.loads(open('gpurun_out/c5_$n.json').read().strip().splitlines()[-1])
print('c5 n_gpus',d['n_gpus'],'value %.4e'%d['value'],'ms/step',round(d['ms_per_step'],3),'scaling',d['scaling'],'result',d.get('result'))"; done
