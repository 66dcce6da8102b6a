#!/bin/bash
mkdir -p gpurun_out
SMMC_LIB=$PWD/stock_market_monte_carlo_b200/lib_exp/libsmmc.so timeout 900 python tools/exp_variants.py sobol > gpurun_out/b_variants.log 2>&1; echo "variants rc=$?"
cat gpurun_out/b_variants.log
for w in c4 c4boot; do timeout 300 python bench.py --workload $w --steps 30 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w', '%.4e' % d['value'], 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'frac', round(d['roofline']['frac'],3))"; done
