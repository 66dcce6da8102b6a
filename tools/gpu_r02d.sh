#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/d_pytest.log
for w in c2 c3 c3compat c4 c4boot; do timeout 300 python bench.py --workload $w --steps 30 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w', '%.4e' % d['value'], 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'frac', round(d['roofline']['frac'],3))"; done
python tools/run_one.py gaussian 100000000 360 rounds=7
python tools/run_one.py bootstrap 100000000 360 rounds=7
