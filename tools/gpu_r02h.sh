#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "two_gpus or run_multi or multi or sharded or device_guard" > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/h_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/h_bench2.json 2> gpurun_out/h_bench2.err; echo "bench2 rc=$?"
grep -i "nranks\|Init COMPLETE" gpurun_out/h_bench2.err | head -8
python - <<'PY'
import json
d=json.loads(open('gpurun_out/h_bench2.json').read().strip().splitlines()[-1])
print('value %.4e ms %.3f e2e %.4e n_gpus %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['n_gpus']))
print(d['config']['collective'])
print('c5', d['extras'].get('c5'))
print('sharded', d['extras'].get('sharded_equals_single'))
PY
tail -5 gpurun_out/h_bench2.err
