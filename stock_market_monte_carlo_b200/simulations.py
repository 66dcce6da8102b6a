"""Host-side mirror of the reference's GPU entry points (reference include/stock_market_monte_carlo/simulations.h:69-86)
over the C ABI of libsmmc.so.  Same names, argument order, argument meaning and error behaviour:

    mc_simulations_gpu(n_simulations, max_n_simulations, n_periods, initial_capital, returns, totals, n_gpus)
    mc_simulations_gpu_reduceBlock(n_simulations, max_n_simulations, n_periods, initial_capital, returns,
                                   means, variances, n_gpus)
    reduce_mean_gpu(vec, n)            vector_add(out, a, b, n)            vector_add_gpu(out, a, b, n)

`std::atomic<long>&` becomes an `AtomicLong`, `std::vector<float>&` a `FloatVector` (a resizable wrapper around a
float32 numpy array).  `simulate()` is the general entry point (seed, mode, fused statistics, histogram,
P(target)); `simulate_device()` is the asynchronous variant over device pointers (torch tensors).
"""
import ctypes as C
import threading
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import _native as N


class AtomicLong:
    """Mirror of the `std::atomic<long>& n_simulations` progress counter of the reference API."""

    def __init__(self, value=0):
        self._v = int(value)
        self._lock = threading.Lock()

    def load(self):
        with self._lock:
            return self._v

    def store(self, v):
        with self._lock:
            self._v = int(v)

    value = property(load, store)

    def __int__(self):
        return self.load()


class FloatVector:
    """Mirror of `std::vector<float>`: `.array` is the float32 storage, replaced/resized by the callee."""

    def __init__(self, data=()):
        self.array = np.ascontiguousarray(np.asarray(data, dtype=np.float32))

    def size(self):
        return int(self.array.size)

    def resize(self, n, value=0.0):
        old = self.array
        self.array = np.full(int(n), value, dtype=np.float32)
        k = min(old.size, int(n))
        self.array[:k] = old[:k]

    def __len__(self):
        return self.size()

    def __getitem__(self, i):
        return self.array[i]


def _as_f32(x):
    if isinstance(x, FloatVector):
        return x.array
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32))


@dataclass
class Result:
    count: int = 0
    below_count: int = 0
    mean: float = 0.0
    m2: float = 0.0
    finals: Optional[np.ndarray] = None
    block_means: Optional[np.ndarray] = None
    block_vars: Optional[np.ndarray] = None
    hist: Optional[np.ndarray] = None
    select_hist: Optional[np.ndarray] = None  # first radix-select digit histogram (simulate(select=...))
    device_ms: float = 0.0
    gpu_launches: int = 0

    @property
    def variance(self):
        return self.m2 / self.count if self.count else 0.0

    @property
    def std(self):
        return self.variance ** 0.5

    @property
    def p_below_target(self):
        return self.below_count / self.count if self.count else 0.0


def make_config(mode, n_simulations, n_periods, initial_capital=1000.0, returns=None, seed=0, stream_offset=0,
                gauss_mean=0.5, gauss_std=10.0 / 12.0, target=None, hist_bins=0, hist_range=None, device=0,
                flags=0, stream=0, returns_ptr=None, n_returns=None, withdraw_fixed=0.0, withdraw_pct=0.0,
                withdraw_pct_schedule=None, philox_rounds=0):
    cfg = N.SmmcConfig()
    cfg.abi_version = N.SMMC_ABI_VERSION
    cfg.mode = N.MODES[mode] if isinstance(mode, str) else int(mode)
    cfg.n_simulations = int(n_simulations)
    cfg.n_periods = int(n_periods)
    keep = None
    if returns_ptr is not None:
        cfg.returns = int(returns_ptr)
        cfg.n_returns = int(n_returns)
    elif returns is not None:
        keep = _as_f32(returns)
        cfg.returns = keep.ctypes.data
        cfg.n_returns = int(keep.size)
    cfg.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    cfg.stream_offset = int(stream_offset)
    cfg.initial_capital = float(initial_capital)
    cfg.gauss_mean = float(gauss_mean)
    cfg.gauss_std = float(gauss_std)
    cfg.target = float(initial_capital if target is None else target)
    cfg.hist_bins = int(hist_bins)
    if hist_bins:
        lo, hi = hist_range if hist_range is not None else (0.0, 50.0 * float(initial_capital))
        cfg.hist_lo, cfg.hist_hi = float(lo), float(hi)
    cfg.device = int(device)
    cfg.flags = int(flags)
    cfg.stream = int(stream) if stream else None
    cfg.philox_rounds = int(philox_rounds)
    cfg.withdraw_fixed = float(withdraw_fixed)
    cfg.withdraw_pct = float(withdraw_pct)
    if withdraw_pct_schedule is not None:
        sched = _as_f32(withdraw_pct_schedule)
        if sched.size != int(n_periods):
            raise ValueError("withdraw_pct_schedule needs one entry per period")
        cfg.withdraw_pct_schedule = sched.ctypes.data
        keep = (keep, sched)
    return cfg, keep


def simulate(mode, n_simulations, n_periods, initial_capital=1000.0, returns=None, seed=0, stream_offset=0,
             gauss_mean=0.5, gauss_std=10.0 / 12.0, target=None, hist_bins=0, hist_range=None,
             want_finals=False, want_block_partials=False, device=0, n_gpus=1, select=None, **withdrawals) -> Result:
    """One blocking run through `smmc_run` / `smmc_run_multi` with HOST buffers (the call a user makes).
    withdrawals: withdraw_fixed=, withdraw_pct=, withdraw_pct_schedule= (include/smmc.h, ABI 2) and philox_rounds= (ABI 3).
    select: a fresh SelectState -- the first radix-select pass rides along (`smmc_run_select`, one GPU) and its digit
    histogram is returned as Result.select_hist (see simulate_with_quantiles)."""
    lib = N.load()
    cfg, keep = make_config(mode, n_simulations, n_periods, initial_capital, returns, seed, stream_offset,
                            gauss_mean, gauss_std, target, hist_bins, hist_range, device, **withdrawals)
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
    if select is not None:
        if n_gpus != 1:
            raise ValueError("select rides along on one GPU only; use quantiles(..., n_gpus=) or distributed.quantiles_sharded")
        out.select_hist = np.zeros(N.SELECT_MAX_TARGETS * N.SELECT_BINS, dtype=np.uint64)
        N.check(lib.smmc_run_select(C.byref(cfg), C.byref(res), C.byref(select.st), out.select_hist.ctypes.data))
    elif n_gpus == 1:
        N.check(lib.smmc_run(C.byref(cfg), C.byref(res)))
    else:
        N.check(lib.smmc_run_multi(C.byref(cfg), int(n_gpus), C.byref(res)))
    del keep
    out.count, out.below_count = int(res.count), int(res.below_count)
    out.mean, out.m2 = float(res.mean), float(res.m2)
    out.device_ms, out.gpu_launches = float(res.device_ms), int(res.gpu_launches)
    return out


def stats_len(hist_bins):
    return int(N.load().smmc_stats_len(int(hist_bins)))


def simulate_device(mode, n_simulations, n_periods, stats_ptr, initial_capital=1000.0, returns_ptr=None,
                    n_returns=0, seed=0, stream_offset=0, gauss_mean=0.5, gauss_std=10.0 / 12.0, target=None,
                    hist_bins=0, hist_range=None, finals_ptr=None, block_means_ptr=None, block_vars_ptr=None,
                    device=0, stream=0, **withdrawals):
    """Asynchronous run over DEVICE pointers on `stream` (e.g. torch tensors' data_ptr() and torch's current
    stream).  Nothing is synchronised; `stats_ptr` (stats_len(hist_bins) float64 on the device) receives the
    allreduce-ready statistics vector described in include/smmc.h.  Returns the number of kernels launched."""
    lib = N.load()
    cfg, _ = make_config(mode, n_simulations, n_periods, initial_capital, None, seed, stream_offset, gauss_mean,
                         gauss_std, target, hist_bins, hist_range, device,
                         flags=N.FLAG_DEVICE_POINTERS | N.FLAG_NO_SYNC, stream=stream,
                         returns_ptr=returns_ptr, n_returns=n_returns, **withdrawals)
    res = N.SmmcResult()
    res.stats_dev = int(stats_ptr)
    res.finals = int(finals_ptr) if finals_ptr else None
    res.block_means = int(block_means_ptr) if block_means_ptr else None
    res.block_vars = int(block_vars_ptr) if block_vars_ptr else None
    N.check(lib.smmc_run(C.byref(cfg), C.byref(res)))
    return int(res.gpu_launches)


def kernel_times(device=0, max_n=256):
    """CUDA-event durations (ms, oldest first) of the simulation kernel of the most recent launches on `device`."""
    lib = N.load()
    out = np.zeros(int(max_n), dtype=np.float32)
    n = C.c_int32()
    N.check(lib.smmc_kernel_times(int(device), out.ctypes.data, int(max_n), C.byref(n)))
    return out[:int(n.value)].copy()


def kernel_kinds(device=0, max_n=256):
    """Which kernel each entry of kernel_times() was (oldest first): N.KERNEL_SIMULATION or N.KERNEL_SELECT_CANDIDATES."""
    lib = N.load()
    out = np.zeros(int(max_n), dtype=np.int32)
    n = C.c_int32()
    N.check(lib.smmc_kernel_kinds(int(device), out.ctypes.data, int(max_n), C.byref(n)))
    return [int(k) for k in out[:int(n.value)]]


def finalize_stats(stats, hist_bins, initial_capital) -> Result:
    """Turn a (possibly all-reduced) host statistics vector into count / mean / M2 / below / histogram."""
    lib = N.load()
    stats = np.ascontiguousarray(np.asarray(stats, dtype=np.float64))
    assert stats.size == stats_len(hist_bins)
    count, below = C.c_uint64(), C.c_uint64()
    mean, m2 = C.c_double(), C.c_double()
    hist = np.zeros(int(hist_bins) + 2, dtype=np.uint64) if hist_bins else None
    N.check(lib.smmc_stats_finalize(stats.ctypes.data, int(hist_bins), float(initial_capital), C.byref(count),
                                    C.byref(below), C.byref(mean), C.byref(m2),
                                    hist.ctypes.data if hist is not None else None))
    return Result(count=int(count.value), below_count=int(below.value), mean=float(mean.value), m2=float(m2.value),
                  hist=hist)


# ---- reference-named entry points ---------------------------------------------------------------------------------

def mc_simulations_gpu(n_simulations, max_n_simulations, n_periods, initial_capital, returns, totals, n_gpus):
    """reference src/simulations.cu:661-680: totals <- fresh vector of the N final values; n_simulations <- N."""
    r = simulate("bootstrap_compat", max_n_simulations, n_periods, initial_capital, returns, want_finals=True,
                 n_gpus=n_gpus)
    totals.array = r.finals
    n_simulations.store(max_n_simulations)


def mc_simulations_gpu_reduceBlock(n_simulations, max_n_simulations, n_periods, initial_capital, returns, means,
                                   variances, n_gpus):
    """reference src/simulations.cu:682-697: per-256-path-block mean / population variance; 1 GPU only."""
    if n_gpus != 1:
        # the reference throws std::invalid_argument (src/simulations.cu:693)
        raise ValueError(">1 GPU not yet supported for reduceBlock version!!")
    r = simulate("bootstrap_compat", max_n_simulations, n_periods, initial_capital, returns,
                 want_block_partials=True)
    means.array = r.block_means
    variances.array = r.block_vars
    n_simulations.store(max_n_simulations)


def reduce_mean_gpu(vec, n):
    """reference src/simulations.cu:269-341."""
    lib = N.load()
    a = _as_f32(vec)
    out = C.c_float()
    N.check(lib.smmc_reduce_mean(a.ctypes.data, int(n), 0, C.byref(out)))
    return float(out.value)


def reduce_mean_device(ptr, n, device=0, stream=0):
    """Mean of n floats resident on `device` (e.g. a torch tensor's data_ptr()); returns (mean, kernel_ms): the
    single-pass HBM-bound reduction behind reduce_mean_gpu without the host round trip."""
    lib = N.load()
    out, ms = C.c_float(), C.c_float()
    N.check(lib.smmc_reduce_mean_device(int(ptr), int(n), int(device), int(stream) or None, C.byref(out), C.byref(ms)))
    return float(out.value), float(ms.value)


def reduce_mean_device_noalloc(ptr, n, device=0, stream=0):
    """Mean of n floats resident on `device`, through smmc_reduce_mean_device_noalloc: for processes in which the library
    may not allocate device memory (the GPU is fully owned by the host's own allocator, other streams must not be stalled
    by cudaMalloc, or the memory budget is fixed at load time).  Scratch is a fixed array inside the library, so
    n <= SMMC_NOALLOC_MAX_N (2**30) and calls on one device must not overlap.  No counterpart in the reference."""
    lib = N.load()
    out = C.c_float()
    N.check(lib.smmc_reduce_mean_device_noalloc(int(ptr), int(n), int(device), int(stream) or None, C.byref(out)))
    return float(out.value)


def vector_add_device(out_ptr, a_ptr, b_ptr, n, device=0, stream=0):
    """out = a + b over device-resident float vectors (data_ptr()s); returns the kernel's CUDA-event time in ms."""
    lib = N.load()
    ms = C.c_float()
    N.check(lib.smmc_vector_add_device(int(out_ptr), int(a_ptr), int(b_ptr), int(n), int(device), int(stream) or None,
                                       C.byref(ms)))
    return float(ms.value)


def vector_add(out, a, b, n):
    """reference src/gpu.cpp:7-15 (host loop)."""
    out[:n] = np.asarray(a[:n], dtype=np.float32) + np.asarray(b[:n], dtype=np.float32)


def vector_add_gpu(out, a, b, n):
    """reference src/gpu.cu:17-47."""
    lib = N.load()
    a32, b32 = _as_f32(a), _as_f32(b)
    assert isinstance(out, np.ndarray) and out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
    N.check(lib.smmc_vector_add(out.ctypes.data, a32.ctypes.data, b32.ctypes.data, int(n), 0))


# ---- order statistics (reference update_quartiles), trajectories (mc_simulations_keepdata), data-format edges -----

def quartile_ranks(n):
    """ranks of {min, Q1, median, Q3, max} in the sorted finals, as the reference indexes them
    (examples/visualize_returns_gpu.cpp:84-86, :109: v[0], v[N/4], v[N/2], v[N/4 + N/2], v[N-1])"""
    n = int(n)
    q1, q2 = n // 4, n // 2
    return [0, q1, q2, q1 + q2, n - 1]


def quantiles(mode, n_simulations, n_periods, ranks=None, initial_capital=1000.0, returns=None, seed=0,
              stream_offset=0, gauss_mean=0.5, gauss_std=10.0 / 12.0, device=0, n_gpus=1, return_ms=False,
              **withdrawals):
    """sorted(finals)[ranks] of the run, exactly, without materialising the finals: three replays of the
    counter-based simulation drive a radix select in shared memory (`smmc_quantiles`).  ranks=None gives the
    reference's five quartile ranks."""
    lib = N.load()
    cfg, keep = make_config(mode, n_simulations, n_periods, initial_capital, returns, seed, stream_offset,
                            gauss_mean, gauss_std, device=device, **withdrawals)
    r = np.ascontiguousarray(np.asarray(quartile_ranks(n_simulations) if ranks is None else ranks, dtype=np.uint64))
    out = np.zeros(r.size, dtype=np.float32)
    ms = C.c_float()
    N.check(lib.smmc_quantiles(C.byref(cfg), int(n_gpus), int(r.size), r.ctypes.data, out.ctypes.data, C.byref(ms)))
    del keep
    return (out, float(ms.value)) if return_ms else out


def simulate_with_quantiles(mode, n_simulations, n_periods, ranks=None, **kw):
    """Statistics AND exact order statistics of one run in three replays instead of four: the first radix-select pass
    rides along with the statistics launch (`smmc_run_select`), two more replays fix the remaining key bits.
    Returns (Result, values, device_ms of the two extra replays); kw as for simulate()."""
    st = SelectState(quartile_ranks(n_simulations) if ranks is None else ranks)
    r = simulate(mode, n_simulations, n_periods, select=st, **kw)
    st.advance(r.select_hist)
    skw = {k: v for k, v in kw.items() if k not in ("target", "hist_bins", "hist_range", "want_finals", "want_block_partials", "n_gpus")}
    cfg, keep = make_config(mode, n_simulations, n_periods, kw.get("initial_capital", 1000.0), kw.get("returns"),
                            kw.get("seed", 0), kw.get("stream_offset", 0), kw.get("gauss_mean", 0.5),
                            kw.get("gauss_std", 10.0 / 12.0), device=kw.get("device", 0),
                            **{k: v for k, v in skw.items() if k.startswith("withdraw") or k == "philox_rounds"})
    extra_ms = 0.0
    while not st.done:
        st.advance(st.run_pass(cfg))
        extra_ms += float(kernel_times(kw.get("device", 0), 1)[-1])
    del keep
    return r, st.values(), extra_ms


def update_quartiles(quartiles, mode, n_simulations, n_periods, **kw):
    """Mirror of the reference helper of the same name (examples/visualize_returns_gpu.cpp:83-110), fed by the
    simulation instead of a host vector of finals: quartiles <- {min, Q1, median, Q3, max}."""
    quartiles.array = quantiles(mode, n_simulations, n_periods, None, **kw)


class SelectState:
    """Pass-by-pass protocol of the radix select, for one-process-per-GPU jobs (distributed.quantiles_sharded)."""

    def __init__(self, ranks):
        self.lib = N.load()
        self.st = N.SmmcSelectState()
        r = np.ascontiguousarray(np.asarray(ranks, dtype=np.uint64))
        N.check(self.lib.smmc_select_begin(C.byref(self.st), int(r.size), r.ctypes.data))

    @property
    def done(self):
        return int(self.st.bits_done) >= 32

    def run_pass(self, cfg, hist_ptr=None):
        """one replay; host histogram returned unless `hist_ptr` (device, with a DEVICE_POINTERS cfg) is given"""
        if hist_ptr is not None:
            N.check(self.lib.smmc_select_pass(C.byref(cfg), C.byref(self.st), int(hist_ptr)))
            return None
        hist = np.zeros(N.SELECT_MAX_TARGETS * N.SELECT_BINS, dtype=np.uint64)
        N.check(self.lib.smmc_select_pass(C.byref(cfg), C.byref(self.st), hist.ctypes.data))
        return hist

    def advance(self, hist):
        h = np.ascontiguousarray(np.asarray(hist, dtype=np.uint64).reshape(-1))
        assert h.size == N.SELECT_MAX_TARGETS * N.SELECT_BINS
        N.check(self.lib.smmc_select_advance(C.byref(self.st), h.ctypes.data))

    def values(self):
        out = np.zeros(int(self.st.n_targets), dtype=np.float32)
        N.check(self.lib.smmc_select_values(C.byref(self.st), out.ctypes.data))
        return out


def trajectories(mode, n_paths, n_periods, initial_capital=1000.0, returns=None, seed=0, stream_offset=0,
                 gauss_mean=0.5, gauss_std=10.0 / 12.0, want_returns=False, device=0, **withdrawals):
    """values[n_paths][n_periods+1] (and the monthly returns in percent [n_paths][n_periods]) of paths
    stream_offset .. stream_offset+n_paths-1: the reference's `mc_data` of mc_simulations_keepdata
    (src/simulations.cpp:139-202), from the same RNG streams as the statistics kernel."""
    lib = N.load()
    cfg, keep = make_config(mode, n_paths, n_periods, initial_capital, returns, seed, stream_offset, gauss_mean,
                            gauss_std, device=device, **withdrawals)
    values = np.zeros((int(n_paths), int(n_periods) + 1), dtype=np.float32)
    rets = np.zeros((int(n_paths), int(n_periods)), dtype=np.float32) if want_returns else None
    N.check(lib.smmc_trajectories(C.byref(cfg), int(n_paths), values.ctypes.data,
                                  rets.ctypes.data if rets is not None else None))
    del keep
    return (values, rets) if want_returns else values


def mc_simulations_keepdata(n_simulations, max_n_simulations, n_periods, initial_capital, historical_returns, mc_data,
                            final_values):
    """reference src/simulations.cpp:139-202: mc_data[i] <- the n_periods+1 values of path i, final_values[i] <- the
    last one; n_simulations <- N.  (GPU bootstrap from the reference's RNG stream.)"""
    v = trajectories("bootstrap_compat", max_n_simulations, n_periods, initial_capital, historical_returns)
    mc_data[:] = [row for row in v]
    final_values.array = np.ascontiguousarray(v[:, -1])
    n_simulations.store(max_n_simulations)


def read_historical_returns(csv_fpath):
    """reference src/simulations.cpp:83-93: the "returns" column of a CSV with a header line, as float32"""
    lib = N.load()
    n = C.c_int32()
    path = str(csv_fpath).encode()
    N.check(lib.smmc_read_returns_csv(path, None, 0, C.byref(n)))
    out = np.zeros(int(n.value), dtype=np.float32)
    N.check(lib.smmc_read_returns_csv(path, out.ctypes.data, int(out.size), C.byref(n)))
    return out


def write_data_file(fname, returns, values):
    """reference src/helpers.cpp:23-39 ("Returns,,r0,...,\nValues,v0,...,"); `fname` is the full path here (the
    reference prepends ./outputs/)"""
    lib = N.load()
    r, v = _as_f32(returns), _as_f32(values)
    N.check(lib.smmc_write_data_file(str(fname).encode(), r.ctypes.data, int(r.size), v.ctypes.data, int(v.size)))


def plan_launch(mode, n_simulations, n_periods, n_returns=0, hist_bins=0, select_targets=0, n_sms=148,
                smem_optin=232448, full=False):
    """(grid, threads, table_copies, smem_bytes) smmc_run would use on a device with `n_sms` SMs (host only);
    full=True: a dict that also says whether the blocks are handed out by ticket"""
    lib = N.load()
    cfg, _ = make_config(mode, n_simulations, n_periods, hist_bins=hist_bins)
    cfg.n_returns = int(n_returns)
    plan = N.SmmcPlan()
    N.check(lib.smmc_plan_launch(C.byref(cfg), int(select_targets), int(n_sms), int(smem_optin), C.byref(plan)))
    if full:
        return dict(grid=int(plan.grid), threads=int(plan.threads), table_copies=int(plan.table_copies),
                    smem_bytes=int(plan.smem_bytes), ticket=bool(plan.ticket))
    return int(plan.grid), int(plan.threads), int(plan.table_copies), int(plan.smem_bytes)


def debug_draws(mode, n_paths, n_periods, returns=None, seed=0, stream_offset=0, device=0, philox_rounds=0):
    lib = N.load()
    cfg, keep = make_config(mode, n_paths, n_periods, returns=returns, seed=seed, stream_offset=stream_offset,
                            device=device, philox_rounds=philox_rounds)
    out = np.zeros((int(n_paths), int(n_periods)), dtype=np.uint32)
    N.check(lib.smmc_debug_draws(C.byref(cfg), int(n_paths), out.ctypes.data))
    del keep
    return out


def debug_philox(ctr, key, device=0, rounds=10):
    lib = N.load()
    ctr = np.ascontiguousarray(np.asarray(ctr, dtype=np.uint32).reshape(-1, 4))
    key = np.ascontiguousarray(np.asarray(key, dtype=np.uint32).reshape(2))
    out = np.zeros_like(ctr)
    N.check(lib.smmc_debug_philox_rounds(ctr.ctypes.data, key.ctypes.data, int(rounds), int(ctr.shape[0]), int(device),
                                         out.ctypes.data))
    return out


def set_variant(key=None, value=-1):
    """Test hook (`smmc_debug_set_variant`): pick a compiled kernel variant; set_variant() resets everything."""
    N.check(N.load().smmc_debug_set_variant(key.encode() if key is not None else None, int(value)))


# ---- one rank per GPU: the NCCL collective behind the C ABI (include/smmc.h, smmc_comm_* / smmc_run_rank) -----------

def rank_range(n_simulations, n_ranks, rank):
    """(first path relative to the run's stream_offset, number of paths) of `rank` -- the partition of smmc_run_multi /
    smmc_run_rank: contiguous ranges of whole 256-path blocks.  Host only."""
    first, count = C.c_int64(), C.c_int64()
    N.check(N.load().smmc_rank_range(int(n_simulations), int(n_ranks), int(rank), C.byref(first), C.byref(count)))
    return int(first.value), int(count.value)


class Comm:
    """One rank's communicator (ncclCommInitRank inside libsmmc.so).  `Comm.unique_id()` on rank 0, the 128 bytes to every
    rank by any means (distributed.init_comm uses torch.distributed), then `Comm(uid, n_ranks, rank, device)` on every rank."""

    def __init__(self, uid, n_ranks, rank, device):
        self.lib = N.load()
        buf = C.create_string_buffer(bytes(uid), N.COMM_ID_BYTES)
        h = C.c_void_p()
        N.check(self.lib.smmc_comm_init_rank(buf, int(n_ranks), int(rank), int(device), C.byref(h)))
        self.handle, self.n_ranks, self.rank, self.device = h, int(n_ranks), int(rank), int(device)

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(N.COMM_ID_BYTES)
        N.check(N.load().smmc_comm_unique_id(buf))
        return bytes(buf.raw)

    def info(self):
        a, b, c, d = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        N.check(self.lib.smmc_comm_info(self.handle, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(n_ranks=a.value, rank=b.value, device=c.value, nccl_version=d.value)

    def allreduce_count(self):
        """ncclAllReduce calls smmc_run_rank has issued through this communicator (exactly one per call)"""
        n = C.c_uint64()
        N.check(self.lib.smmc_comm_allreduce_count(self.handle, C.byref(n)))
        return int(n.value)

    def destroy(self):
        if self.handle:
            self.lib.smmc_comm_destroy(self.handle)
            self.handle = None

    def simulate(self, mode, n_simulations, n_periods, initial_capital=1000.0, returns=None, seed=0, stream_offset=0,
                 gauss_mean=0.5, gauss_std=10.0 / 12.0, target=None, hist_bins=0, hist_range=None, want_finals=False,
                 want_block_partials=False, **extra) -> Result:
        """This rank's part of the GLOBAL run + the all-reduce (smmc_run_rank, host buffers, blocking).  count / mean / m2 /
        below_count / hist describe the global run on every rank; finals / block partials cover this rank's range."""
        cfg, keep = make_config(mode, n_simulations, n_periods, initial_capital, returns, seed, stream_offset, gauss_mean,
                                gauss_std, target, hist_bins, hist_range, self.device, **extra)
        _, n_local = rank_range(n_simulations, self.n_ranks, self.rank)
        nb = (n_local + N.SMMC_BLOCK - 1) // N.SMMC_BLOCK
        res, out = N.SmmcResult(), Result()
        if want_finals:
            out.finals = np.empty(n_local, dtype=np.float32)
            res.finals = out.finals.ctypes.data
        if want_block_partials:
            out.block_means, out.block_vars = np.zeros(nb, dtype=np.float32), np.zeros(nb, dtype=np.float32)
            res.block_means, res.block_vars = out.block_means.ctypes.data, out.block_vars.ctypes.data
        if hist_bins:
            out.hist = np.zeros(int(hist_bins) + 2, dtype=np.uint64)
            res.hist = out.hist.ctypes.data
        N.check(self.lib.smmc_run_rank(C.byref(cfg), C.byref(res), self.handle))
        del keep
        out.count, out.below_count = int(res.count), int(res.below_count)
        out.mean, out.m2 = float(res.mean), float(res.m2)
        out.device_ms, out.gpu_launches = float(res.device_ms), int(res.gpu_launches)
        return out

    def simulate_device(self, mode, n_simulations, n_periods, stats_ptr, stream=0, returns_ptr=None, n_returns=0,
                        block_means_ptr=None, block_vars_ptr=None, finals_ptr=None, **kw):
        """Asynchronous variant over device pointers on `stream`: kernel -> ncclAllReduce of the statistics vector at
        `stats_ptr`, nothing synchronised.  Returns the number of kernels launched on this rank."""
        cfg, _ = make_config(mode, n_simulations, n_periods, device=self.device, flags=N.FLAG_DEVICE_POINTERS | N.FLAG_NO_SYNC,
                             stream=stream, returns_ptr=returns_ptr, n_returns=n_returns, **kw)
        res = N.SmmcResult()
        res.stats_dev = int(stats_ptr)
        res.finals = int(finals_ptr) if finals_ptr else None
        res.block_means = int(block_means_ptr) if block_means_ptr else None
        res.block_vars = int(block_vars_ptr) if block_vars_ptr else None
        N.check(self.lib.smmc_run_rank(C.byref(cfg), C.byref(res), self.handle))
        return int(res.gpu_launches)

    def quantiles(self, mode, n_simulations, n_periods, ranks=None, return_ms=False, **kw):
        """sorted(finals of the GLOBAL run)[ranks] on every rank (smmc_quantiles_rank)."""
        cfg, keep = make_config(mode, n_simulations, n_periods, device=self.device, **kw)
        r = np.ascontiguousarray(np.asarray(quartile_ranks(n_simulations) if ranks is None else ranks, dtype=np.uint64))
        out = np.zeros(r.size, dtype=np.float32)
        ms = C.c_float()
        N.check(self.lib.smmc_quantiles_rank(C.byref(cfg), self.handle, int(r.size), r.ctypes.data, out.ctypes.data, C.byref(ms)))
        del keep
        return (out, float(ms.value)) if return_ms else out
