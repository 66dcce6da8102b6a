"""Builds tests/cxx/finals_time.cpp against the in-tree libsmmc.so and runs it (GPU box; timing driver, not a test).
usage: python tests/run_cxx_finals_time.py [N] [P]"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
libdir = os.path.join(ROOT, "stock_market_monte_carlo_b200", "lib")
with tempfile.TemporaryDirectory() as d:
    exe = os.path.join(d, "finals_time")
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", exe, os.path.join(ROOT, "tests", "cxx", "finals_time.cpp"),
                    f"-I{ROOT}/include", f"-L{libdir}", "-lsmmc", f"-Wl,-rpath,{libdir}"], check=True)
    subprocess.run([exe] + sys.argv[1:], check=True)
