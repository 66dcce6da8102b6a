#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/exp_finals.py > gpurun_out/e_finals.log 2>&1; echo "finals rc=$?"; cat gpurun_out/e_finals.log
timeout 600 python tests/run_cxx_finals_time.py > gpurun_out/e_cxx.log 2>&1; echo "cxx rc=$?"; cat gpurun_out/e_cxx.log
