"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so) and of the compiled reference (oracle/_ref/*.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  Nothing under stock_market_monte_carlo_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from stock_market_monte_carlo_b200 import _native as N
from stock_market_monte_carlo_b200.data import sobol_directions
from stock_market_monte_carlo_b200.simulations import Result, make_config

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "_build", "liboracle.so")
REF_DIR = os.path.join(_HERE, "_ref")

_lib = None
_sobol_keep = None


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)


def load():
    global _lib, _sobol_keep
    if _lib is not None:
        return _lib
    if not os.path.exists(ORACLE_SO):
        build()
    lib = C.CDLL(ORACLE_SO)
    P = C.POINTER
    lib.orc_rand_pcg.argtypes = [C.c_uint32]
    lib.orc_rand_pcg.restype = C.c_uint32
    lib.orc_xorshift.argtypes = [C.c_uint32]
    lib.orc_xorshift.restype = C.c_uint32
    lib.orc_compat_index.argtypes = [C.c_uint32, C.c_int32]
    lib.orc_compat_index.restype = C.c_int32
    lib.orc_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_philox4x32_10.restype = None
    lib.orc_philox4x32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.orc_philox4x32.restype = None
    lib.orc_sobol_set_table.argtypes = [C.c_void_p, C.c_int32]
    lib.orc_sobol_point.argtypes = [C.c_uint64, C.c_int32]
    lib.orc_sobol_point.restype = C.c_uint32
    lib.orc_hist_bin.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int32]
    lib.orc_hist_bin.restype = C.c_int32
    lib.orc_hist_inv_width.argtypes = [C.c_float, C.c_float, C.c_int32]
    lib.orc_hist_inv_width.restype = C.c_float
    lib.orc_run.argtypes = [P(N.SmmcConfig), P(N.SmmcResult)]
    lib.orc_run.restype = C.c_int
    lib.orc_draws.argtypes = [P(N.SmmcConfig), C.c_int32, C.c_void_p]
    lib.orc_draws.restype = C.c_int
    lib.orc_reduce_mean.argtypes = [C.c_void_p, C.c_int64]
    lib.orc_reduce_mean.restype = C.c_float
    lib.orc_num_threads.restype = C.c_int
    lib.orc_gauss_pairs.argtypes = [C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_void_p]
    lib.orc_gauss_pairs.restype = None
    lib.orc_sobol_gauss_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_void_p]
    lib.orc_sobol_gauss_pairs.restype = None
    lib.orc_trajectories.argtypes = [P(N.SmmcConfig), C.c_int32, C.c_void_p, C.c_void_p]
    lib.orc_trajectories.restype = C.c_int
    lib.orc_quantiles.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    lib.orc_quantiles.restype = C.c_int
    lib.orc_quartile_ranks.argtypes = [C.c_int64, C.c_void_p]
    lib.orc_quartile_ranks.restype = None
    _sobol_keep = np.ascontiguousarray(sobol_directions())
    lib.orc_sobol_set_table(_sobol_keep.ctypes.data, 1111)
    _lib = lib
    return lib


def simulate(mode, n_simulations, n_periods, initial_capital=1000.0, returns=None, seed=0, stream_offset=0,
             gauss_mean=0.5, gauss_std=10.0 / 12.0, target=None, hist_bins=0, hist_range=None,
             want_finals=False, want_block_partials=False, **withdrawals) -> Result:
    """Same signature and result as stock_market_monte_carlo_b200.simulate, computed by the CPU oracle."""
    lib = load()
    cfg, keep = make_config(mode, n_simulations, n_periods, initial_capital, returns, seed, stream_offset,
                            gauss_mean, gauss_std, target, hist_bins, hist_range, **withdrawals)
    res = N.SmmcResult()
    out = Result()
    nb = (int(n_simulations) + N.SMMC_BLOCK - 1) // N.SMMC_BLOCK
    if want_finals:
        out.finals = np.empty(int(n_simulations), dtype=np.float32)
        res.finals = out.finals.ctypes.data
    if want_block_partials:
        out.block_means = np.zeros(nb, dtype=np.float32)
        out.block_vars = np.zeros(nb, dtype=np.float32)
        res.block_means = out.block_means.ctypes.data
        res.block_vars = out.block_vars.ctypes.data
    if hist_bins:
        out.hist = np.zeros(int(hist_bins) + 2, dtype=np.uint64)
        res.hist = out.hist.ctypes.data
    rc = lib.orc_run(C.byref(cfg), C.byref(res))
    if rc != 0:
        raise ValueError("oracle rejected the configuration")
    del keep
    out.count, out.below_count = int(res.count), int(res.below_count)
    out.mean, out.m2 = float(res.mean), float(res.m2)
    return out


def draws(mode, n_paths, n_periods, returns=None, seed=0, stream_offset=0, philox_rounds=0):
    lib = load()
    cfg, keep = make_config(mode, n_paths, n_periods, returns=returns, seed=seed, stream_offset=stream_offset,
                            philox_rounds=philox_rounds)
    out = np.zeros((int(n_paths), int(n_periods)), dtype=np.uint32)
    if lib.orc_draws(C.byref(cfg), int(n_paths), out.ctypes.data) != 0:
        raise ValueError("oracle rejected the configuration")
    del keep
    return out


def trajectories(mode, n_paths, n_periods, initial_capital=1000.0, returns=None, seed=0, stream_offset=0,
                 gauss_mean=0.5, gauss_std=10.0 / 12.0, want_returns=False, **withdrawals):
    """Same contract as stock_market_monte_carlo_b200.trajectories, computed by the CPU oracle."""
    lib = load()
    cfg, keep = make_config(mode, n_paths, n_periods, initial_capital, returns, seed, stream_offset, gauss_mean,
                            gauss_std, **withdrawals)
    values = np.zeros((int(n_paths), int(n_periods) + 1), dtype=np.float32)
    rets = np.zeros((int(n_paths), int(n_periods)), dtype=np.float32) if want_returns else None
    if lib.orc_trajectories(C.byref(cfg), int(n_paths), values.ctypes.data,
                            rets.ctypes.data if rets is not None else None) != 0:
        raise ValueError("oracle rejected the configuration")
    del keep
    return (values, rets) if want_returns else values


def quartile_ranks(n):
    r = np.zeros(5, dtype=np.uint64)
    load().orc_quartile_ranks(int(n), r.ctypes.data)
    return r


def quantiles_of(finals, ranks):
    """the reference's way (update_quartiles): sort all finals, read the value at each rank"""
    lib = load()
    f = np.ascontiguousarray(np.asarray(finals, dtype=np.float32))
    r = np.ascontiguousarray(np.asarray(ranks, dtype=np.uint64))
    out = np.zeros(r.size, dtype=np.float32)
    if lib.orc_quantiles(f.ctypes.data, int(f.size), int(r.size), r.ctypes.data, out.ctypes.data) != 0:
        raise ValueError("bad ranks")
    return out


def gauss_pairs(words, mean_pct=0.0, std_pct=100.0):
    """growth factors (g0, g1) of each 32-bit word; with mean 0 % and std 100 % they are the normals (z0, z1)"""
    lib = load()
    w = np.ascontiguousarray(np.asarray(words, dtype=np.uint32))
    g = np.zeros((w.size, 2), dtype=np.float32)
    lib.orc_gauss_pairs(w.ctypes.data, int(w.size), float(mean_pct), float(std_pct), g.ctypes.data)
    return g


def sobol_gauss_pairs(xr, xa, mean_pct=0.0, std_pct=100.0):
    """SOBOL_GAUSSIAN growth factors (g0, g1) of each (radius word, angle word) pair of Sobol words; with mean 0 % and
    std 100 % they are the normals (z0, z1)"""
    lib = load()
    xr = np.ascontiguousarray(np.asarray(xr, dtype=np.uint32))
    xa = np.ascontiguousarray(np.asarray(xa, dtype=np.uint32))
    assert xr.shape == xa.shape
    g = np.zeros((xr.size, 2), dtype=np.float32)
    lib.orc_sobol_gauss_pairs(xr.ctypes.data, xa.ctypes.data, int(xr.size), float(mean_pct), float(std_pct), g.ctypes.data)
    return g


def philox(ctr, key, rounds=10):
    lib = load()
    ctr = np.ascontiguousarray(np.asarray(ctr, dtype=np.uint32).reshape(-1, 4))
    key = np.ascontiguousarray(np.asarray(key, dtype=np.uint32).reshape(2))
    out = np.zeros_like(ctr)
    for i in range(ctr.shape[0]):
        lib.orc_philox4x32(ctr[i].ctypes.data, key.ctypes.data, int(rounds), out[i].ctypes.data)
    return out


# ---- the compiled reference (oracle/_ref) ---------------------------------------------------------------------------

def ref_available(name):
    return os.path.exists(os.path.join(REF_DIR, name))


_ref = {}


def ref_lib(name):
    if name not in _ref:
        _ref[name] = C.CDLL(os.path.join(REF_DIR, name))
    return _ref[name]


def ref_mc_simulations(n, n_periods, initial_capital, returns, quiet=True):
    """reference CPU v2 OpenMP bootstrap (src/simulations.cpp:204-266) -> (finals, threads)"""
    lib = ref_lib("libref_cpu.so")
    lib.ref_mc_simulations.argtypes = [C.c_long, C.c_uint, C.c_float, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.ref_mc_simulations.restype = C.c_int
    r = np.ascontiguousarray(np.asarray(returns, dtype=np.float32))
    out = np.empty(int(n), dtype=np.float32)
    threads = lib.ref_mc_simulations(int(n), int(n_periods), float(initial_capital), r.ctypes.data, int(r.size),
                                     out.ctypes.data, 1 if quiet else 0)
    return out, threads


def ref_gaussian_finals(n, n_periods, initial_capital, mean_pct, std_pct):
    """reference sample_returns_gaussian + many_updates per path under OpenMP -> (finals, threads)"""
    lib = ref_lib("libref_cpu.so")
    lib.ref_gaussian_finals.argtypes = [C.c_long, C.c_uint, C.c_float, C.c_float, C.c_float, C.c_void_p]
    lib.ref_gaussian_finals.restype = C.c_int
    out = np.empty(int(n), dtype=np.float32)
    threads = lib.ref_gaussian_finals(int(n), int(n_periods), float(initial_capital), float(mean_pct),
                                      float(std_pct), out.ctypes.data)
    return out, threads


def ref_read_historical_returns(path, max_n=1 << 20):
    lib = ref_lib("libref_cpu.so")
    lib.ref_read_historical_returns.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
    lib.ref_read_historical_returns.restype = C.c_long
    out = np.zeros(int(max_n), dtype=np.float32)
    n = lib.ref_read_historical_returns(str(path).encode(), out.ctypes.data, int(max_n))
    return out[:n].copy()


def ref_write_data_file(fname, returns, values):
    """reference write_data_file: writes ./outputs/<fname> under the CURRENT working directory"""
    lib = ref_lib("libref_cpu.so")
    lib.ref_write_data_file.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.ref_write_data_file.restype = None
    r = np.ascontiguousarray(np.asarray(returns, dtype=np.float32))
    v = np.ascontiguousarray(np.asarray(values, dtype=np.float32))
    lib.ref_write_data_file(str(fname).encode(), r.ctypes.data, int(r.size), v.ctypes.data, int(v.size))


def ref_many_updates(fund, returns):
    lib = ref_lib("libref_cpu.so")
    lib.ref_many_updates.argtypes = [C.c_float, C.c_void_p, C.c_uint, C.c_void_p]
    lib.ref_many_updates.restype = None
    r = np.ascontiguousarray(np.asarray(returns, dtype=np.float32))
    out = np.zeros(r.size + 1, dtype=np.float32)
    lib.ref_many_updates(float(fund), r.ctypes.data, int(r.size), out.ctypes.data)
    return out


def ref_update_fund(fund, ret):
    lib = ref_lib("libref_cpu.so")
    lib.ref_update_fund.argtypes = [C.c_float, C.c_float]
    lib.ref_update_fund.restype = C.c_float
    return float(lib.ref_update_fund(float(fund), float(ret)))


def ref_i4_sobol(dim_num, point):
    lib = ref_lib("libref_sobol.so")
    lib.ref_i4_sobol.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_void_p]
    q = np.zeros(int(dim_num), dtype=np.float32)
    seed = C.c_int(int(point))
    lib.ref_i4_sobol(int(dim_num), C.byref(seed), q.ctypes.data)
    return q


def ref_mc_simulations_gpu(n, n_periods, initial_capital, returns, n_gpus=1):
    """the reference's own GPU library compiled for sm_100a (GPU box only) -> totals[N]"""
    lib = ref_lib("libref_gpu.so")
    lib.ref_mc_simulations_gpu.argtypes = [C.c_long, C.c_int, C.c_float, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.ref_mc_simulations_gpu.restype = C.c_int
    r = np.ascontiguousarray(np.asarray(returns, dtype=np.float32))
    out = np.zeros(int(n), dtype=np.float32)
    rc = lib.ref_mc_simulations_gpu(int(n), int(n_periods), float(initial_capital), r.ctypes.data, int(r.size),
                                    out.ctypes.data, int(n_gpus))
    if rc != 0:
        raise RuntimeError(f"reference GPU run failed rc={rc}")
    return out


def ref_mc_simulations_gpu_reduceBlock(n, n_periods, initial_capital, returns):
    lib = ref_lib("libref_gpu.so")
    lib.ref_mc_simulations_gpu_reduceBlock.argtypes = [C.c_long, C.c_int, C.c_float, C.c_void_p, C.c_int,
                                                       C.c_void_p, C.c_void_p]
    lib.ref_mc_simulations_gpu_reduceBlock.restype = C.c_int
    r = np.ascontiguousarray(np.asarray(returns, dtype=np.float32))
    nb = (int(n) + 255) // 256
    means = np.zeros(nb, dtype=np.float32)
    variances = np.zeros(nb, dtype=np.float32)
    rc = lib.ref_mc_simulations_gpu_reduceBlock(int(n), int(n_periods), float(initial_capital), r.ctypes.data,
                                                int(r.size), means.ctypes.data, variances.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"reference GPU reduceBlock run failed rc={rc}")
    return means, variances


def ref_gpu_last_kernel_seconds():
    """the reference library's own "GPU kernel itself took X s" clock of the last ref_mc_simulations_gpu* call (1 ms steps)"""
    lib = ref_lib("libref_gpu.so")
    lib.ref_last_kernel_seconds.restype = C.c_double
    return float(lib.ref_last_kernel_seconds())


def ref_kernel_event_ms(which, n, n_periods, initial_capital, returns, reps=5):
    """CUDA-event duration (ms, mean of `reps` launches after a warm-up) of the reference's own kernel, compiled unmodified for
    sm_100a: which = "reduceBlock" (src/simulations.cu:185-247) or "finals" (:112-152), launched as its launcher does."""
    lib = ref_lib("libref_gpu_events.so")
    lib.ref_kernel_event_ms.argtypes = [C.c_int, C.c_long, C.c_int, C.c_float, C.c_void_p, C.c_int, C.c_int]
    lib.ref_kernel_event_ms.restype = C.c_float
    r = np.ascontiguousarray(np.asarray(returns, dtype=np.float32))
    ms = float(lib.ref_kernel_event_ms(0 if which == "reduceBlock" else 1, int(n), int(n_periods), float(initial_capital),
                                       r.ctypes.data, int(r.size), int(reps)))
    if ms <= 0.0:
        raise RuntimeError(f"reference kernel event timing failed ({ms})")
    return ms
