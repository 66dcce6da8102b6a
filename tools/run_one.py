"""Development driver: one simulate() call (after a small warm-up call), for ncu captures.
usage: python tools/run_one.py <mode> <N> <P> [finals] [rounds=7] [hist=0] [key=value ...]   (key=value: smmc_debug_set_variant)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stock_market_monte_carlo_b200 as S

mode, N, P = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
kw = dict(seed=1234)
bins = 100
for a in sys.argv[4:]:
    if "=" in a:
        k, v = a.split("=")
        if k == "rounds":
            kw["philox_rounds"] = int(v)
        elif k == "hist":
            bins = int(v)
        else:
            S.set_variant(k, int(v))
S.set_variant("finals_chunks", 1)  # one monolithic launch: the capture is of the whole workload (no split partials download)
if "bootstrap" in mode:
    kw["returns"] = S.synthetic_sp500_table(1200)
if mode == "bootstrap_compat":
    kw["seed"] = 0
rng = {"sobol_gaussian": (0.0, 500000.0), "sobol_bootstrap": (0.0, 1.0e8)}.get(mode, (0.0, 50000.0 if mode == "gaussian" else 200000.0))
S.simulate(mode, 256 * 148 * 32, P, **kw)
r = S.simulate(mode, N, P, hist_bins=bins, hist_range=rng if bins else None, want_block_partials=not mode.startswith("sobol"), want_finals="finals" in sys.argv[4:], **kw)
print(mode, N, P, sys.argv[4:], "ms", r.device_ms, "Tpm/s", N * P / r.device_ms / 1e9, "mean", r.mean, "std", r.std)
