#!/bin/bash
# 2-GPU bench only: NCCL INIT lines on stderr, JSON on stdout
mkdir -p gpurun_out; env | grep -i "^NCCL\|^TORCH" 
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/i_bench2.json 2> gpurun_out/i_bench2.err; echo "bench2 rc=$?"
echo "--- stdout lines:"; wc -l gpurun_out/i_bench2.json; grep -v '^{' gpurun_out/i_bench2.json | head
echo "--- nccl lines on stderr:"; grep -i "nranks\|Init COMPLETE\|NCCL version\|NVLS" gpurun_out/i_bench2.err | head -12
python - <<'PY'
import json
d=json.loads(open('gpurun_out/i_bench2.json').read().strip().splitlines()[-1])
print('value %.4e ms %.3f e2e %.4e n_gpus %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['n_gpus']))
print(d['config'].get('collective'))
print('c5', d['extras'].get('c5'))
PY
ls -la /tmp/smmc_nccl* 2>&1 | head; head -c 1500 /tmp/smmc_nccl*.log | head -30; head -20 gpurun_out/i_bench2.err
