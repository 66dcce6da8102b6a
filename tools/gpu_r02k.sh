#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/k_pytest.log
timeout 300 python tools/exp_e2e.py 2>&1 | tail -4
python bench.py --steps 30 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('value %.4e e2e %.4e ratio %.4f' % (d['value'], d['e2e']['value'], d['e2e']['value']/d['value']))"
