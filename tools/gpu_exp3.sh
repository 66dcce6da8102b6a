#!/bin/bash
# usage: bash tools/gpu_exp3.sh <libdir> <groups...>
mkdir -p gpurun_out
lib=$1; shift
SMMC_LIB=stock_market_monte_carlo_b200/$lib/libsmmc.so timeout 900 python tools/exp_variants.py "$@" > gpurun_out/exp3_$lib.log 2>&1; echo "rc=$?"; tail -40 gpurun_out/exp3_$lib.log
