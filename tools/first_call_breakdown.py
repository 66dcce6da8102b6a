"""Where the first call of a fresh process spends its time: run in a fresh interpreter on the GPU box.
usage: python tools/first_call_breakdown.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

t = time.perf_counter()
import stock_market_monte_carlo_b200 as S  # noqa: E402
lib = S.load()
print("dlopen libsmmc.so            %8.1f ms" % (1e3 * (time.perf_counter() - t)))
t = time.perf_counter()
n = lib.smmc_device_count()
print("smmc_device_count (cuInit)   %8.1f ms  devices=%d" % (1e3 * (time.perf_counter() - t), n))
tbl = S.synthetic_sp500_table(1127, seed=7)
t = time.perf_counter()
r = S.simulate("bootstrap_compat", 256, 1, returns=tbl)
print("tiny first smmc_run          %8.1f ms  (primary context, per-device Context, module load of one kernel)  device_ms=%.3f" % (1e3 * (time.perf_counter() - t), r.device_ms))
t = time.perf_counter()
r = S.simulate("bootstrap_compat", 256, 1, returns=tbl)
print("tiny second smmc_run         %8.1f ms" % (1e3 * (time.perf_counter() - t)))
t = time.perf_counter()
r = S.simulate("gaussian", 256, 1)
print("tiny first call, other mode  %8.1f ms  device_ms=%.3f" % (1e3 * (time.perf_counter() - t), r.device_ms))
N, P = 100_000_000, 360
for k in range(3):
    t = time.perf_counter()
    r = S.simulate("bootstrap_compat", N, P, returns=tbl, want_finals=True)
    print("10^8 x 360 all finals, call %d %8.1f ms  device_ms=%.2f" % (k + 1, 1e3 * (time.perf_counter() - t), r.device_ms))
    del r
