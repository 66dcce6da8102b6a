#!/bin/bash
mkdir -p gpurun_out
for a in "hist=100" "hist=0"; do python tools/run_one.py gaussian 100000000 360 $a; done
python tools/run_one.py sobol_gaussian 5000000 1000 hist=0
python tools/run_one.py sobol_gaussian 5000000 1000 hist=100
prof() { # name, args...
  name=$1; shift
  python tools/run_one.py "$@" > gpurun_out/plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:mc_kernel -s 1 -c 1 -f -o gpurun_out/c_$name python tools/run_one.py "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
}
prof sobol_boot sobol_bootstrap 5000000 1000
prof sobol_gauss_tbl sobol_gaussian 5000000 1000
prof gauss_tbl5 gaussian 100000000 360
prof gauss_tbl0 gaussian 100000000 360 gauss_tbl=0
ls -la gpurun_out | tail -12
