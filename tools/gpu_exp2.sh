#!/bin/bash
# usage: bash tools/gpu_exp2.sh <groups...>  -- experiment build, variant timings
mkdir -p gpurun_out
SMMC_LIB=stock_market_monte_carlo_b200/lib_exp/libsmmc.so timeout 900 python tools/exp_variants.py "$@" > gpurun_out/exp2.log 2>&1; echo "rc=$?"; cat gpurun_out/exp2.log | tail -60
