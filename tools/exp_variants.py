"""Development driver (GPU box): kernel-only times of the compiled variants of the simulation kernel, plus a correctness
check of every variant against the default one / the oracle tolerance.  Variants are chosen through the test hook
smmc_debug_set_variant.  Use with the experiment build:  SMMC_LIB=stock_market_monte_carlo_b200/lib_exp/libsmmc.so

usage: python tools/exp_variants.py [group ...]     groups: gauss boot sobol finals (default: all)
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import stock_market_monte_carlo_b200 as S  # noqa: E402

groups = [a for a in sys.argv[1:] if "=" not in a] or ["gauss", "boot", "sobol"]
MASKS = [int(x) for a in sys.argv[1:] if a.startswith("masks=") for x in a[6:].split(",")] or [0, 1, 5, 7, 15]
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
tbl = S.synthetic_sp500_table(1200)
tbl_dev = torch.from_numpy(tbl).to(dev)
stats = torch.zeros(S.stats_len(100), dtype=torch.float64, device=dev)
stream = torch.cuda.current_stream(dev).cuda_stream
PEAK = 148 * 128 * 1.965e9


def timed(mode, N, P, reps=5, warm=2, partials=True, hist_range=(0.0, 50000.0), **kw):
    bm = torch.empty((N + 255) // 256, dtype=torch.float32, device=dev)
    bv = torch.empty_like(bm)
    extra = dict(kw)
    if "bootstrap" in mode:
        extra.update(returns_ptr=tbl_dev.data_ptr(), n_returns=tbl.size)
    for i in range(warm + reps):
        S.simulate_device(mode, N, P, stats.data_ptr(), seed=0 if mode == "bootstrap_compat" else 1234,
                          stream_offset=((i * N + 255) // 256 * 256) % (1 << 30), hist_bins=100, hist_range=hist_range,
                          block_means_ptr=bm.data_ptr() if partials else None, block_vars_ptr=bv.data_ptr() if partials else None,
                          stream=stream, **extra)
    torch.cuda.synchronize()
    ms = S.kernel_times(0, reps)
    r = S.finalize_stats(stats.cpu().numpy(), 100, 1000.0)
    return float(np.mean(ms)), float(np.min(ms)), r


def line(label, mode, N, P, ms, mn, r):
    rate = N * P / (ms * 1e-3)
    print(f"{label:44s} {mode:17s} ms {ms:8.3f} (min {mn:8.3f})  {rate:.4e} pm/s  cyc/warp-month {PEAK / rate:6.2f}  mean {r.mean:.4f} std {r.std:.4f}",
          flush=True)


def check_gauss(mode, **kw):
    """finals of this variant vs the all-MUFU variant: same contract, different sin/cos evaluation"""
    N, P = 20000, 360
    S.set_variant("gauss_tbl", 0)
    S.set_variant("sobol_tbl", 0)
    a = S.simulate(mode, N, P, seed=7, want_finals=True, **kw).finals
    return a


if "gauss" in groups:
    N, P = 100_000_000, 360
    for rounds in (10, 7):
        for tblmask in MASKS:
            S.set_variant()
            S.set_variant("gauss_tbl", tblmask)
            ref = None
            small = S.simulate("gaussian", 20000, 360, seed=7, want_finals=True, philox_rounds=rounds).finals
            S.set_variant("gauss_tbl", 0)
            base = S.simulate("gaussian", 20000, 360, seed=7, want_finals=True, philox_rounds=rounds).finals
            S.set_variant("gauss_tbl", tblmask)
            err = float(np.max(np.abs(small / base - 1.0)))
            ms, mn, r = timed("gaussian", N, P, philox_rounds=rounds)
            line(f"gauss rounds={rounds} tbl_mask={tblmask:2d} maxrel_vs_mufu={err:.1e}", "gaussian", N, P, ms, mn, r)
    S.set_variant()

if "boot" in groups:
    N, P = 100_000_000, 360
    for rounds in (10, 7):
        for ilp in (4, 8):
            S.set_variant()
            S.set_variant("ilp", ilp)
            ms, mn, r = timed("bootstrap", N, P, philox_rounds=rounds)
            line(f"bootstrap rounds={rounds} ilp={ilp}", "bootstrap", N, P, ms, mn, r)
    S.set_variant()
    ms, mn, r = timed("bootstrap_compat", N, P)
    line("compat", "bootstrap_compat", N, P, ms, mn, r)

if "sobol" in groups:
    N, P = 5_000_000, 1000
    for t in (0x77,):
        S.set_variant()
        S.set_variant("sobol_tbl", t)
        small = S.simulate("sobol_gaussian", 20000, 360, seed=7, want_finals=True).finals
        S.set_variant("sobol_tbl", 0)
        base = S.simulate("sobol_gaussian", 20000, 360, seed=7, want_finals=True).finals
        S.set_variant("sobol_tbl", t)
        err = float(np.max(np.abs(small / base - 1.0)))
        for w in (0,):
            S.set_variant("warps", w)
            ms, mn, r = timed("sobol_gaussian", N, P, reps=10, partials=False, hist_range=(0.0, 500000.0))
            line(f"sobol_gaussian tbl={t} warps={w} maxrel_vs_mufu={err:.1e}", "sobol_gaussian", N, P, ms, mn, r)
    for u in (2, 100, 101, 102, 103, 2):
        S.set_variant()
        S.set_variant("sobol_unroll", u)
        for w in (0, 32):
            S.set_variant("warps", w)
            ms, mn, r = timed("sobol_bootstrap", N, P, reps=10, partials=False, hist_range=(0.0, 1.0e8))
            line(f"sobol_bootstrap unroll={u} warps={w}", "sobol_bootstrap", N, P, ms, mn, r)
    S.set_variant()

if "ticket" in groups:
    # block assignment: static vs ticket (block-order merge), warps per CTA sweep
    for mode, N, P, rng, ws in (("gaussian", 100_000_000, 360, (0.0, 50000.0), (0,)), ("bootstrap", 100_000_000, 360, (0.0, 200000.0), (0,)),
                                ("bootstrap", 20_000_000, 360, (0.0, 200000.0), (0, 32)), ("gaussian", 5_000_000, 360, (0.0, 50000.0), (0, 28, 32)),
                                ("sobol_gaussian", 5_000_000, 1000, (0.0, 500000.0), (0, 24, 27, 28, 32)),
                                ("sobol_bootstrap", 5_000_000, 1000, (0.0, 1.0e8), (0, 24, 27, 28, 32))):
        for t in (0, 1, -1):
            for w in ws:
                S.set_variant()
                S.set_variant("ticket", t)
                S.set_variant("warps", w)
                ms, mn, r = timed(mode, N, P, reps=5 if N > 10_000_000 else 10, partials=not mode.startswith("sobol"), hist_range=rng)
                line(f"ticket={t} warps={w}", mode, N, P, ms, mn, r)
    S.set_variant()
