"""B200-native (sm_100a) Monte Carlo portfolio hot path: drop-in for the GPU entry points of
matthijsvk/stock_market_monte_carlo (reference src/simulations.cu, src/gpu.cpp) behind the C ABI of include/smmc.h.
"""
from ._native import (MODES, MODE_BOOTSTRAP, MODE_BOOTSTRAP_COMPAT, MODE_GAUSSIAN, MODE_SOBOL_BOOTSTRAP,
                      MODE_SOBOL_GAUSSIAN, SmmcError, load)
from .simulations import (AtomicLong, FloatVector, Result, debug_draws, debug_philox, finalize_stats, kernel_times, kernel_kinds,
                          mc_simulations_gpu, mc_simulations_gpu_reduceBlock, reduce_mean_gpu, reduce_mean_device, reduce_mean_device_noalloc, simulate,
                          simulate_device, stats_len, vector_add, vector_add_gpu, vector_add_device, quantiles, quartile_ranks,
                          update_quartiles, simulate_with_quantiles, SelectState, trajectories, mc_simulations_keepdata, read_historical_returns,
                          write_data_file, plan_launch, make_config, set_variant, Comm, rank_range)
from .data import synthetic_sp500_table

__all__ = [
    "MODES", "SmmcError", "load", "AtomicLong", "FloatVector", "Result", "simulate", "simulate_device",
    "finalize_stats", "kernel_times", "kernel_kinds", "stats_len", "mc_simulations_gpu", "mc_simulations_gpu_reduceBlock", "reduce_mean_gpu", "reduce_mean_device", "reduce_mean_device_noalloc",
    "vector_add", "vector_add_gpu", "vector_add_device", "debug_draws", "debug_philox", "synthetic_sp500_table", "quantiles", "quartile_ranks",
    "update_quartiles", "simulate_with_quantiles", "SelectState", "trajectories", "mc_simulations_keepdata", "read_historical_returns", "write_data_file",
    "plan_launch", "make_config", "set_variant", "Comm", "rank_range",
]
