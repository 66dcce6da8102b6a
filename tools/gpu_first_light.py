"""Development script (GPU box): first-light parity + timing sweep over modes and ILP variants."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stock_market_monte_carlo_b200 as S
from oracle import binding as O

tbl = S.synthetic_sp500_table(1200)
tbl1127 = S.synthetic_sp500_table(1127, seed=7)
print(S.load().smmc_version().decode(), "devices", S.load().smmc_device_count())

kat = O.philox([[0,0,0,0],[0xffffffff]*4,[0x243f6a88,0x85a308d3,0x13198a2e,0x03707344]][2:3], [0xa4093822,0x299f31d0])
dev = S.debug_philox([[0x243f6a88,0x85a308d3,0x13198a2e,0x03707344]], [0xa4093822,0x299f31d0])
print("philox dev", [hex(int(x)) for x in dev[0]], "oracle", [hex(int(x)) for x in kat[0]])

def cmp(mode, N, P, **kw):
    g = S.simulate(mode, N, P, hist_bins=100, want_finals=True, want_block_partials=True, **kw)
    o = O.simulate(mode, N, P, hist_bins=100, want_finals=True, want_block_partials=True, **kw)
    fin_eq = np.array_equal(g.finals, o.finals)
    rel = np.max(np.abs(g.finals.astype(np.float64) - o.finals) / np.abs(o.finals)) if N else 0
    print(f"{mode:18s} N={N:7d} P={P:4d} finals_bitexact={fin_eq} maxrel={rel:.3e} "
          f"bmeans_eq={np.array_equal(g.block_means, o.block_means)} bvars_eq={np.array_equal(g.block_vars, o.block_vars)} "
          f"hist_eq={np.array_equal(g.hist, o.hist)} hist_l1={int(np.abs(g.hist.astype(np.int64)-o.hist.astype(np.int64)).sum())} "
          f"below {g.below_count}/{o.below_count} mean {g.mean:.6f}/{o.mean:.6f} relm2 {abs(g.m2-o.m2)/max(o.m2,1e-300):.2e} ms={g.device_ms:.3f}")

for mode, kw in (("bootstrap_compat", dict(returns=tbl)), ("bootstrap", dict(returns=tbl, seed=1234)), ("gaussian", dict(seed=1234))):
    for N, P in ((0, 360), (1, 1), (255, 7), (256, 360), (1000, 361), (65536 + 77, 360), (200000, 120)):
        cmp(mode, N, P, **kw)
    d = S.debug_draws(mode, 64, 50, returns=kw.get("returns"), seed=kw.get("seed", 0))
    o = O.draws(mode, 64, 50, returns=kw.get("returns"), seed=kw.get("seed", 0))
    print(mode, "draws equal:", np.array_equal(d, o))

# reference GPU kernel compiled for sm_100a (1127-row table)
if O.ref_available("libref_gpu.so"):
    N = 100000
    ref = O.ref_mc_simulations_gpu(N, 360, 1000.0, tbl1127)
    g = S.simulate("bootstrap_compat", N, 360, returns=tbl1127, want_finals=True, want_block_partials=True)
    neq = int((ref != g.finals).sum())
    print("vs reference .cu on this GPU: finals mismatches", neq, "of", N)
    rm, rv = O.ref_mc_simulations_gpu_reduceBlock(N, 360, 1000.0, tbl1127)
    print("   reduceBlock means mismatches", int((rm != g.block_means).sum()), "vars mismatches", int((rv != g.block_vars).sum()), "of", rm.size)

# timing sweep
N, P = 100_000_000, 360
for mode, kw in (("gaussian", dict(seed=1234)), ("bootstrap", dict(returns=tbl, seed=1234)), ("bootstrap_compat", dict(returns=tbl))):
    for ilp in (1, 2, 4, 8):
        os.environ["SMMC_ILP"] = str(ilp)
        S.simulate(mode, 1_000_000, P, **kw)
        best = 1e9
        for _ in range(3):
            r = S.simulate(mode, N, P, hist_bins=100, want_block_partials=True, **kw)
            best = min(best, r.device_ms)
        print(f"{mode:18s} ILP={ilp} {best:8.3f} ms  {N*P/best/1e9:8.2f} G path-months/s  mean={r.mean:.3f} std={r.std:.3f}")

print("---- sobol")
for mode, kw in (("sobol_bootstrap", dict(returns=tbl, seed=0)), ("sobol_bootstrap", dict(returns=tbl, seed=99)),
                 ("sobol_gaussian", dict(seed=0)), ("sobol_gaussian", dict(seed=99))):
    for N, P in ((0, 360), (1, 1), (255, 7), (256, 360), (1000, 361), (65536 + 77, 360), (20000, 1000)):
        cmp(mode, N, P, **kw)
    d = S.debug_draws(mode, 300, 50, returns=kw.get("returns"), seed=kw.get("seed", 0))
    o = O.draws(mode, 300, 50, returns=kw.get("returns"), seed=kw.get("seed", 0))
    print(mode, "draws equal:", np.array_equal(d, o))
for mode, kw in (("sobol_bootstrap", dict(returns=tbl, seed=0)), ("sobol_gaussian", dict(seed=0))):
    N, P = 5_000_000, 1000
    S.simulate(mode, 100000, P, **kw)
    best = 1e9
    for _ in range(3):
        r = S.simulate(mode, N, P, hist_bins=100, **kw)
        best = min(best, r.device_ms)
    print(f"{mode:18s} {best:8.3f} ms  {N*P/best/1e9:8.3f} T path-months/s  mean={r.mean:.3f} std={r.std:.3f} p_below={r.p_below_target:.6f}")
