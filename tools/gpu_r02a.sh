#!/bin/bash
# r02 call A: GPU tests on the production library, then kernel-variant timings on the experiment library.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/a_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/a_pytest.log
SMMC_LIB=$PWD/stock_market_monte_carlo_b200/lib_exp/libsmmc.so timeout 900 python tools/exp_variants.py > gpurun_out/a_variants.log 2>&1; echo "variants rc=$?"
cat gpurun_out/a_variants.log
timeout 600 python tests/ref_gpu_time.py > gpurun_out/a_refgpu.log 2>&1; echo "refgpu rc=$?"; cat gpurun_out/a_refgpu.log
