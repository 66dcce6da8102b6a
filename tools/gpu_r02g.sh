#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/g_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/g_bench.json').read().strip().splitlines()[-1])
print('value %.4e ms %.3f e2e %.4e frac %.3f pipe %s clocks %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline'].get('pipe_floor'), d['clocks']))
ex=d['extras']
for k,v in ex.get('workloads',{}).items():
    print(k, '%.4e' % v['value'], 'kernel_ms', round(v['kernel_ms'],3), 'frac', round(v['roofline']['frac'],3), v.get('vs_ref_gpu_same_box'), v.get('e2e'))
print('ref', json.dumps(ex.get('ref_gpu_same_box'))[:900])
print('c5', ex.get('c5'))
print('quart', ex.get('quartiles_device_ms'), ex.get('quartiles_riding_along_extra_ms'), ex.get('quartiles_riding_along_equal'))
print('cpu', d['cpu_baseline'])
PY
