#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/j_pytest.log
timeout 600 python tools/exp_variants.py gauss ticket > gpurun_out/j_exp.log 2>&1; echo "exp rc=$?"; cat gpurun_out/j_exp.log
