"""Development driver: one simulate() call (after a small warm-up call), for ncu captures.
usage: python tools/run_one.py <mode> <N> <P> [finals]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stock_market_monte_carlo_b200 as S

mode, N, P = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
kw = dict(seed=1234)
if "bootstrap" in mode:
    kw["returns"] = S.synthetic_sp500_table(1200)
if mode == "bootstrap_compat":
    kw["seed"] = 0
S.simulate(mode, 256 * 148 * 32, P, **kw)
r = S.simulate(mode, N, P, hist_bins=100, want_block_partials=not mode.startswith("sobol"), want_finals=len(sys.argv) > 4, **kw)
print(mode, N, P, "ms", r.device_ms, "Tpm/s", N * P / r.device_ms / 1e9, "mean", r.mean, "std", r.std)
