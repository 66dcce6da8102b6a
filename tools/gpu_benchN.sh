#!/bin/bash
# usage: bash tools/gpu_benchN.sh N   -- the driver's launch line at N GPUs, JSON -> gpurun_out/benchN_<N>.json
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/benchN_$N.json 2> gpurun_out/benchN_$N.err; echo "bench rc=$?"
wc -l gpurun_out/benchN_$N.json
grep -c "Init COMPLETE" gpurun_out/benchN_$N.err
python - $N <<'PY'
import json, sys
n = sys.argv[1]
d=json.loads(open(f'gpurun_out/benchN_{n}.json').read().strip().splitlines()[-1])
print('value %.4e ms %.3f e2e %.4e n_gpus %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['n_gpus']))
c5 = d['extras'].get('c5') or {}
print('c5 ms', c5.get('ms_per_step'), 'value', c5.get('value'), 'sha', c5.get('hist_sha256'), c5.get('closed_form'))
print('sharded', d['extras'].get('sharded_equals_single'))
print('clocks', d['clocks'])
PY
