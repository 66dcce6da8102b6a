#!/bin/bash
# experiment build: parity tests of the touched modes, then variant timings.  usage: bash tools/gpu_exp.sh "<pytest -k expr>" <groups...>
mkdir -p gpurun_out
k=$1; shift
export SMMC_LIB=stock_market_monte_carlo_b200/lib_exp/libsmmc.so
timeout 900 python -m pytest tests -m gpu -x -q -k "$k" > gpurun_out/exp4_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/exp4_pytest.log
timeout 900 python tools/exp_variants.py "$@" > gpurun_out/exp4.log 2>&1; echo "rc=$?"; tail -40 gpurun_out/exp4.log
