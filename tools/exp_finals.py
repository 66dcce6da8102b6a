"""Development driver (GPU box): the all-finals path through the C ABI -- pipelined vs monolithic, bit-identity, wall time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stock_market_monte_carlo_b200 as S

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
P = 360
tbl = S.synthetic_sp500_table(1127, seed=7)
def run(chunks, mode="bootstrap_compat", **kw):
    S.set_variant("finals_chunks", chunks)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        r = S.simulate(mode, N, P, returns=tbl if "bootstrap" in mode else None, want_finals=True, hist_bins=100, hist_range=(0.0, 200000.0), **kw)
        ts.append(1e3 * (time.perf_counter() - t0))
    return r, ts
mono, t1 = run(1)
print("monolithic   wall ms", [round(t, 2) for t in t1], "device_ms", round(mono.device_ms, 3), flush=True)
for ch in (0, 8, 32):
    r, t = run(ch)
    same = np.array_equal(r.finals.view(np.uint32), mono.finals.view(np.uint32))
    print(f"chunks={ch:2d}    wall ms", [round(x, 2) for x in t], "device_ms", round(r.device_ms, 3), "finals identical", same,
          "hist identical", np.array_equal(r.hist, mono.hist), "below", r.below_count == mono.below_count,
          "mean rel diff", abs(r.mean - mono.mean) / mono.mean, "launches", r.gpu_launches, flush=True)
for th in (4, 12, 16):
    S.set_variant("finals_threads", th)
    r, t = run(0)
    print(f"threads={th:2d}   wall ms", [round(x, 2) for x in t], "device_ms", round(r.device_ms, 3), flush=True)
S.set_variant()
import subprocess
print(open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip(), "| hw threads", os.cpu_count())
r2 = S.simulate("bootstrap_compat", N, P, returns=tbl, want_finals=True, want_block_partials=True)
m2 = S.simulate("bootstrap_compat", N, P, returns=tbl, want_block_partials=True)
print("block partials identical with/without pipelined finals:", np.array_equal(r2.block_means, m2.block_means), np.array_equal(r2.block_vars, m2.block_vars))
