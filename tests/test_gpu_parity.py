"""Parity of the CUDA path (through the C ABI, libsmmc.so) against the oracle, the golden fixtures and -- when
oracle/_ref/libref_gpu.so travelled with the snapshot -- the reference's own kernels on the same GPU.

Bar (BASELINE.json north_star): bootstrap index sequences, finals of the table-driven modes, per-block partials,
histogram bin counts and below-target counts are BIT-EXACT; Gaussian finals agree within a stated fp32 relative
tolerance (device MUFU lg2/sqrt/sin/cos vs libm in the oracle) and moments within 1e-5 relative.
"""
import os
import subprocess
import threading

import numpy as np
import pytest

from conftest import ALL_MODES, EXACT_MODES, ROOT, golden, mode_kwargs

pytestmark = pytest.mark.gpu

# fp32 tolerance of the Gaussian modes: the per-month growth factor differs by <= ~2 ulp between MUFU and libm,
# errors accumulate like a random walk over P months: ~ sqrt(P) * 2^-24 * few.  Measured max 2.7e-6 at P = 1000.
GAUSS_RTOL = 2e-5
MOMENT_RTOL = 1e-5

SHAPES = [(0, 360), (1, 1), (1, 0), (31, 5), (255, 7), (256, 360), (257, 13), (1000, 361), (1000, 362), (1000, 363),
          (65536 + 77, 360), (200_000, 120), (20_000, 1000), (777, 1111), (513, 1110), (300, 33)]


def _compare(S, O, mode, N, P, table, **extra):
    kw = mode_kwargs(mode, table)
    kw.update(extra)
    g = S.simulate(mode, N, P, hist_bins=100, want_finals=True, want_block_partials=True, **kw)
    o = O.simulate(mode, N, P, hist_bins=100, want_finals=True, want_block_partials=True, **kw)
    assert g.count == o.count == N and g.gpu_launches >= 1
    assert int(g.hist.sum()) == N
    if mode in EXACT_MODES:
        assert np.array_equal(g.finals, o.finals)
        assert np.array_equal(g.block_means, o.block_means)
        assert np.array_equal(g.block_vars, o.block_vars)
        assert np.array_equal(g.hist, o.hist)
        assert g.below_count == o.below_count
        assert abs(g.mean - o.mean) <= 1e-12 * abs(o.mean) + 0.0
        assert abs(g.m2 - o.m2) <= 1e-9 * abs(o.m2)
    else:
        np.testing.assert_allclose(g.finals, o.finals, rtol=GAUSS_RTOL)
        np.testing.assert_allclose(g.block_means, o.block_means, rtol=GAUSS_RTOL)
        np.testing.assert_allclose(g.block_vars, o.block_vars, rtol=1e-3, atol=1e-6)
        # a final within 2e-5 of a bin edge / the target may land on the other side
        near_edges = int(np.abs(g.hist.astype(np.int64) - o.hist.astype(np.int64)).sum())
        assert near_edges <= max(4, N // 2000)
        assert abs(int(g.below_count) - int(o.below_count)) <= max(2, N // 5000)
        if N:
            assert abs(g.mean - o.mean) <= MOMENT_RTOL * abs(o.mean)
            assert abs(g.m2 - o.m2) <= MOMENT_RTOL * abs(o.m2) + 1e-9
    return g, o


def test_device_philox_known_answers(S):
    from test_oracle_pins import PHILOX_KAT
    for ctr, key, want in PHILOX_KAT:
        assert [int(x) for x in S.debug_philox([ctr], key)[0]] == want


@pytest.mark.parametrize("mode", ALL_MODES)
@pytest.mark.parametrize("N,P", SHAPES)
def test_modes_against_oracle(S, O, table, mode, N, P):
    _compare(S, O, mode, N, P, table)


@pytest.mark.parametrize("mode", ALL_MODES)
@pytest.mark.parametrize("seed", [0, 99])
def test_draw_sequences_bit_exact(S, O, table, mode, seed):
    """bootstrap modes: the table index of every (path, month); gaussian: the bit field behind every month"""
    kw = mode_kwargs(mode, table, seed=seed)
    for off in (0, 512, 1 << 20):
        d = S.debug_draws(mode, 300, 50, returns=kw.get("returns"), seed=kw["seed"], stream_offset=off)
        o = O.draws(mode, 300, 50, returns=kw.get("returns"), seed=kw["seed"], stream_offset=off)
        assert np.array_equal(d, o)


@pytest.mark.parametrize("mode", ["bootstrap", "gaussian"])
def test_64_bit_path_ids(S, O, table, mode):
    """native modes take 64-bit stream offsets (the 10^10-path config needs ids >= 2^32)"""
    off = (1 << 33) + 256 * 12345
    _compare(S, O, mode, 1000, 48, table, stream_offset=off)
    kw = mode_kwargs(mode, table)
    d = S.debug_draws(mode, 64, 20, returns=kw.get("returns"), seed=kw["seed"], stream_offset=off)
    o = O.draws(mode, 64, 20, returns=kw.get("returns"), seed=kw["seed"], stream_offset=off)
    assert np.array_equal(d, o)
    lo = S.debug_draws(mode, 64, 20, returns=kw.get("returns"), seed=kw["seed"], stream_offset=256 * 12345)
    assert not np.array_equal(d, lo)  # the high word of the path id reaches the counter


def test_golden_reference_kernel_fixture(S, table1127):
    """fixture = output of the reference's kernels on a B200 (tests/golden/make_golden_gpu.py)"""
    g = golden("ref_gpu_compat.npz")
    n = int(g["n"])
    r = S.simulate("bootstrap_compat", n, 360, returns=table1127, want_finals=True, want_block_partials=True)
    assert np.array_equal(r.finals, g["totals"])
    assert np.array_equal(r.block_means, g["means"]) and np.array_equal(r.block_vars, g["variances"])


def test_against_reference_kernels_on_this_gpu(S, O, table1127):
    """the reference's src/simulations.cu compiled unmodified for sm_100a (oracle/_ref/libref_gpu.so)"""
    if not O.ref_available("libref_gpu.so"):
        pytest.skip("oracle/_ref/libref_gpu.so did not travel with the snapshot")
    for N, P in ((100_000, 360), (1000, 1), (777, 50)):
        ref = O.ref_mc_simulations_gpu(N, P, 1000.0, table1127)
        g = S.simulate("bootstrap_compat", N, P, returns=table1127, want_finals=True, want_block_partials=True)
        assert np.array_equal(ref, g.finals)
        rm, rv = O.ref_mc_simulations_gpu_reduceBlock(N, P, 1000.0, table1127)
        assert np.array_equal(rm, g.block_means) and np.array_equal(rv, g.block_vars)
    # the reference-named Python wrappers produce the same layout
    n, totals = S.AtomicLong(0), S.FloatVector()
    S.mc_simulations_gpu(n, 5000, 36, 1000.0, S.FloatVector(table1127), totals, 1)
    assert n.load() == 5000 and np.array_equal(totals.array, O.ref_mc_simulations_gpu(5000, 36, 1000.0, table1127))
    means, variances = S.FloatVector(), S.FloatVector()
    S.mc_simulations_gpu_reduceBlock(n, 5000, 36, 1000.0, S.FloatVector(table1127), means, variances, 1)
    rm, rv = O.ref_mc_simulations_gpu_reduceBlock(5000, 36, 1000.0, table1127)
    assert means.size() == 20 and np.array_equal(means.array, rm) and np.array_equal(variances.array, rv)


def test_against_reference_kernels_at_baseline_size(S, O, table1127):
    """config C3's shape (100 M paths x 360 months) through the reference's own kernels compiled for sm_100a: every final
    and every block mean / variance bit for bit -- except where the reference's result is undefined: a state >= 2^32 - 128
    maps to idx == n_returns and the reference then reads bufferReturns[1127], one float past its
    shared array (src/simulations.cu:123, :148-149), i.e. stale shared memory; this build reads a zero pad there.
    Expected 10^8 * 360 * 128 / 2^32 = 1073 such paths; every differing path must be one of them (checked with the
    oracle's draws)."""
    if not O.ref_available("libref_gpu.so"):
        pytest.skip("oracle/_ref/libref_gpu.so did not travel with the snapshot")
    N, P, n = 100_000_000, 360, 1127
    g = S.simulate("bootstrap_compat", N, P, returns=table1127, want_finals=True, want_block_partials=True)
    ref = O.ref_mc_simulations_gpu(N, P, 1000.0, table1127)
    bad = np.flatnonzero(ref != g.finals)
    del ref
    expected = N * P * 128 / 2.0 ** 32
    assert bad.size < expected + 6 * np.sqrt(expected), bad.size
    for b in bad:
        blk = (int(b) // 256) * 256
        d = O.draws("bootstrap_compat", 256, P, returns=table1127, stream_offset=blk).reshape(256, P)
        assert (d[int(b) - blk] == n).any(), f"path {b} differs from the reference kernel without an idx == n draw"
    rm, rv = O.ref_mc_simulations_gpu_reduceBlock(N, P, 1000.0, table1127)
    assert rm.size == (N + 255) // 256
    clean = np.ones(rm.size, dtype=bool)
    clean[bad // 256] = False
    # the reduceBlock kernel's own out-of-bounds read lands in different stale memory: its affected blocks are the same
    assert np.array_equal(rm[clean], g.block_means[clean]) and np.array_equal(rv[clean], g.block_vars[clean])


@pytest.mark.parametrize("mode", ALL_MODES)
def test_sharding_identity_on_device(S, table, mode):
    kw = mode_kwargs(mode, table)
    N = 256 * 700 + 19
    full = S.simulate(mode, N, 30, hist_bins=64, want_finals=True, want_block_partials=True, **kw)
    cut = 256 * 300
    a = S.simulate(mode, cut, 30, hist_bins=64, want_finals=True, want_block_partials=True, **kw)
    b = S.simulate(mode, N - cut, 30, hist_bins=64, want_finals=True, want_block_partials=True, stream_offset=cut, **kw)
    assert np.array_equal(np.concatenate([a.finals, b.finals]), full.finals)
    assert np.array_equal(np.concatenate([a.block_means, b.block_means]), full.block_means)
    assert np.array_equal(a.hist + b.hist, full.hist) and a.below_count + b.below_count == full.below_count
    n, s1 = a.count + b.count, a.mean * a.count + b.mean * b.count
    assert n == N and abs(s1 / n - full.mean) <= 1e-12 * full.mean


def test_statistics_do_not_depend_on_requested_outputs(S, table):
    for mode in ALL_MODES:
        kw = mode_kwargs(mode, table)
        a = S.simulate(mode, 50_000, 40, hist_bins=10, **kw)
        b = S.simulate(mode, 50_000, 40, hist_bins=10, want_finals=True, want_block_partials=True, **kw)
        assert (a.count, a.below_count, a.mean, a.m2) == (b.count, b.below_count, b.mean, b.m2)
        assert np.array_equal(a.hist, b.hist) and a.finals is None and a.block_means is None
        c = S.simulate(mode, 50_000, 40, **kw)
        assert c.hist is None and (c.mean, c.m2) == (a.mean, a.m2)


@pytest.mark.parametrize("n_returns", [1, 2, 7, 1127, 1200, 1777, 5000, 40_000])
def test_table_sizes_and_replication(S, O, n_returns):
    """the table is replicated up to 32x (bank-private rows) while it fits in shared memory; results never depend on it"""
    tbl = S.synthetic_sp500_table(n_returns, seed=3)
    for mode in EXACT_MODES:
        _compare(S, O, mode, 3000, 24, tbl)


def test_table_too_large_is_reported(S):
    tbl = S.synthetic_sp500_table(70_000, seed=3)
    with pytest.raises(S.SmmcError) as e:
        S.simulate("bootstrap", 1000, 12, returns=tbl, seed=1)
    assert e.value.status == 4


def test_ilp_variants_agree(S, table):
    """kernel variants are chosen through the test hook smmc_debug_set_variant (never the environment)"""
    base = {}
    try:
        for ilp in (2, 4, 8):
            S.set_variant("ilp", ilp)
            for mode in ("bootstrap_compat", "bootstrap", "gaussian"):
                kw = mode_kwargs(mode, table)
                r = S.simulate(mode, 10_000, 41, hist_bins=10, want_finals=True, **kw)
                if mode in base:
                    assert np.array_equal(r.finals, base[mode].finals) and r.mean == base[mode].mean
                else:
                    base[mode] = r
    finally:
        S.set_variant()


def test_environment_does_not_select_kernels(S, table, monkeypatch):
    """round-1 development knobs (SMMC_ILP / SMMC_WARPS / SMMC_REDUCE) are gone: the library never reads the environment"""
    base = S.plan_launch("gaussian", 148 * 27 * 256 + 300, 12)
    for k, v in (("SMMC_ILP", "2"), ("SMMC_WARPS", "20"), ("SMMC_REDUCE", "ldg")):
        monkeypatch.setenv(k, v)
    assert S.plan_launch("gaussian", 148 * 27 * 256 + 300, 12) == base
    try:
        S.set_variant("warps", 20)
        assert S.plan_launch("gaussian", 148 * 27 * 256 + 300, 12)[1] == 20 * 32
    finally:
        S.set_variant()


def test_histogram_range_target_and_seed(S, O, table):
    r = S.simulate("gaussian", 100_000, 120, seed=5, hist_bins=7, hist_range=(1500.0, 2000.0), target=1800.0,
                   want_finals=True)
    o_bins = np.zeros(9, dtype=np.uint64)
    lib = O.load()
    inv = lib.orc_hist_inv_width(1500.0, 2000.0, 7)
    for x in r.finals:
        o_bins[lib.orc_hist_bin(float(x), 1500.0, inv, 7)] += 1
    assert np.array_equal(o_bins, r.hist)  # bin rule applied to the device's own finals: bit-exact
    assert r.hist[0] > 0 and r.hist[-1] > 0
    assert r.below_count == int((r.finals < np.float32(1800.0)).sum())
    r2 = S.simulate("gaussian", 100_000, 120, seed=6, want_finals=True)
    assert not np.array_equal(r.finals, r2.finals)
    r3 = S.simulate("gaussian", 100_000, 120, seed=5, gauss_mean=0.0, gauss_std=0.0, want_finals=True)
    assert np.all(r3.finals == 1000.0)


def test_device_pointer_async_entry(S, table):
    import torch
    dev = torch.device("cuda", 0)
    bins = 50
    for mode in ALL_MODES:
        kw = mode_kwargs(mode, table)
        host = S.simulate(mode, 100_000, 36, hist_bins=bins, want_finals=True, **kw)
        stats = torch.zeros(S.stats_len(bins), dtype=torch.float64, device=dev)
        finals = torch.zeros(100_000, dtype=torch.float32, device=dev)
        tdev = torch.from_numpy(table).to(dev) if "returns" in kw else None
        launches = S.simulate_device(mode, 100_000, 36, stats.data_ptr(), seed=kw["seed"], hist_bins=bins,
                                     returns_ptr=tdev.data_ptr() if tdev is not None else None,
                                     n_returns=table.size if tdev is not None else 0,
                                     finals_ptr=finals.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert launches >= 1
        r = S.finalize_stats(stats.cpu().numpy(), bins, 1000.0)
        assert np.array_equal(finals.cpu().numpy(), host.finals) and np.array_equal(r.hist, host.hist)
        assert r.count == host.count and r.below_count == host.below_count
        assert abs(r.mean - host.mean) <= 1e-12 * host.mean


def test_callable_from_a_worker_thread(S, table):
    """the reference's GUI calls the entry point from a std::thread and polls n_simulations
    (examples/visualize_returns_gpu.cpp:174-182, :362)"""
    n, totals = S.AtomicLong(0), S.FloatVector()
    t = threading.Thread(target=S.mc_simulations_gpu, args=(n, 100_000, 120, 1000.0, S.FloatVector(table), totals, 1))
    t.start()
    t.join()
    assert n.load() == 100_000 and totals.size() == 100_000
    ref = S.simulate("bootstrap_compat", 100_000, 120, returns=table, want_finals=True)
    assert np.array_equal(totals.array, ref.finals)


def test_reduce_mean_device_pointer(S):
    """smmc_reduce_mean_device: device-resident input, any 4-byte alignment, 64-bit counts, exact fp64 accumulation"""
    import torch
    rng = np.random.default_rng(5)
    h = (rng.standard_normal(3_000_003) * 1000.0 + 5000.0).astype(np.float32)
    d = torch.from_numpy(h).cuda()
    try:
        for kernel in (1, 2):  # plain streaming loads / the TMA-pipelined kernel (forced here; by size otherwise)
            S.set_variant("reduce", kernel)
            for off, n in ((0, h.size), (1, h.size - 1), (2, 1), (3, 4097), (5, 2_000_000), (0, 8192 * 296 + 5), (3, 8192 * 300 + 1)):
                got, ms = S.reduce_mean_device(d.data_ptr() + 4 * off, n)
                assert got == np.float32(h[off:off + n].astype(np.float64).mean()), (kernel, off, n)
                assert ms > 0.0
    finally:
        S.set_variant()
    big = torch.full((1 << 28,), 1.5, dtype=torch.float32, device="cuda")  # 1 GiB: larger than L2
    got, ms = S.reduce_mean_device(big.data_ptr(), big.numel(), stream=torch.cuda.current_stream().cuda_stream)
    assert got == 1.5


def test_vector_add_device_pointer(S):
    """smmc_vector_add_device: vector path (16-byte aligned), scalar path (misaligned views), tails"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(3)
    a = torch.randn(1_000_003, device="cuda", generator=g)
    b = torch.randn(1_000_003, device="cuda", generator=g)
    out = torch.empty_like(a)
    for off, n in ((0, a.numel()), (0, 4096), (0, 5), (1, 999_999), (3, 1), (2, 123_457)):
        out.zero_()
        ms = S.vector_add_device(out.data_ptr() + 4 * off, a.data_ptr() + 4 * off, b.data_ptr() + 4 * off, n)
        assert torch.equal(out[off:off + n], a[off:off + n] + b[off:off + n]) and ms > 0.0
        assert not out[off + n:].any() and not out[:off].any()


def test_reduce_mean_and_vector_add(S):
    rng = np.random.default_rng(0)
    for n in (1, 3, 1000, 1_000_003, 20_000_000):
        v = rng.normal(5.0, 2.0, n).astype(np.float32)
        got = S.reduce_mean_gpu(v, n)
        assert abs(got - v.astype(np.float64).mean()) <= 1e-6 * 5.0
    # the reference's own check: vec[i] = i (examples/benchmark_reduce_mean.cpp:28-43)
    v = np.arange(1_000_000, dtype=np.float32)
    assert S.reduce_mean_gpu(v, v.size) == np.float32(v.astype(np.float64).mean())
    assert S.reduce_mean_gpu(v, 1000) == np.float32(499.5)  # first n elements only
    a, b = rng.normal(size=100_001).astype(np.float32), rng.normal(size=100_001).astype(np.float32)
    out = np.zeros_like(a)
    S.vector_add_gpu(out, a, b, a.size)
    assert np.array_equal(out, a + b)


def test_multi_gpu_single_process(S, table):
    """smmc_run_multi (the reference's one-process driver model, src/simulations.cu:576-655)"""
    ndev = S.load().smmc_device_count()
    if ndev < 2:
        pytest.skip("needs >= 2 GPUs")
    for mode in ("bootstrap_compat", "gaussian"):
        kw = mode_kwargs(mode, table)
        one = S.simulate(mode, 1_000_003, 60, hist_bins=100, want_finals=True, want_block_partials=True, **kw)
        for g in range(2, min(ndev, 8) + 1):
            m = S.simulate(mode, 1_000_003, 60, hist_bins=100, want_finals=True, want_block_partials=True, n_gpus=g, **kw)
            assert np.array_equal(m.finals, one.finals) and np.array_equal(m.block_means, one.block_means)
            assert np.array_equal(m.hist, one.hist) and m.below_count == one.below_count
            assert abs(m.mean - one.mean) <= 1e-12 * one.mean and abs(m.m2 - one.m2) <= 1e-9 * one.m2


# ---- BASELINE.json sizes: size-independent properties -------------------------------------------------------------
def _gauss_closed_form(P):
    mu, sigma = float(np.float32(0.5) * np.float32(0.01)), float(np.float32(10.0 / 12.0) * np.float32(0.01))
    mean = 1000.0 * (1.0 + mu) ** P
    return mean, np.sqrt(1000.0 ** 2 * ((1.0 + mu) ** 2 + sigma ** 2) ** P - mean ** 2)


def test_full_size_gaussian_100m(S, O):
    """config C2: 360 x 100 M Gaussian with per-block mean/variance partials"""
    N, P = 100_000_000, 360
    r = S.simulate("gaussian", N, P, seed=1234, hist_bins=100, want_block_partials=True)
    mean, std = _gauss_closed_form(P)
    assert r.count == N and int(r.hist.sum()) == N
    assert abs(r.mean - mean) < 5 * std / np.sqrt(N)           # SE_rel ~ 1.6e-5
    assert abs(r.std / std - 1.0) < 2e-3
    assert r.block_means.size == 390_625
    assert abs(r.block_means.astype(np.float64).mean() - r.mean) < 1e-6 * r.mean
    # pooled variance = mean within-block variance + variance of block means (Chan merge, exact)
    pooled = r.block_vars.astype(np.float64).mean() + r.block_means.astype(np.float64).var()
    assert abs(pooled / r.variance - 1.0) < 1e-5
    # sharded by RNG offset: two halves == whole, bit-identical integer partials
    h = (N // 2) // 256 * 256  # shard boundaries are whole 256-path logical blocks
    a = S.simulate("gaussian", h, P, seed=1234, hist_bins=100)
    b = S.simulate("gaussian", N - h, P, seed=1234, hist_bins=100, stream_offset=h)
    assert np.array_equal(a.hist + b.hist, r.hist) and a.below_count + b.below_count == r.below_count
    assert abs((a.mean * a.count + b.mean * b.count) / N - r.mean) < 1e-12 * r.mean
    # spot-check far blocks against the oracle
    for off in (256 * 390_000, 256 * 200_001):
        g = S.simulate("gaussian", 512, P, seed=1234, want_finals=True, stream_offset=off)
        o = O.simulate("gaussian", 512, P, seed=1234, want_finals=True, stream_offset=off)
        np.testing.assert_allclose(g.finals, o.finals, rtol=GAUSS_RTOL)


@pytest.mark.parametrize("mode", ["bootstrap", "bootstrap_compat"])
def test_full_size_bootstrap_100m(S, O, table, mode):
    """config C3: 360 x 100 M bootstrap from the ~1200-month table"""
    from test_oracle_pins import _closed_form_bootstrap
    N, P = 100_000_000, 360
    kw = mode_kwargs(mode, table)
    r = S.simulate(mode, N, P, hist_bins=100, hist_range=(0.0, 200_000.0), **kw)
    mean, std = _closed_form_bootstrap(table, P)
    assert r.count == N and int(r.hist.sum()) == N
    assert abs(r.mean - mean) < 5 * std / np.sqrt(N)
    assert abs(r.std / std - 1.0) < 0.03  # heavy right tail: the sample std converges slowly
    off = 256 * 390_000
    g = S.simulate(mode, 512, P, want_finals=True, stream_offset=off, **kw)
    o = O.simulate(mode, 512, P, want_finals=True, stream_offset=off, **kw)
    assert np.array_equal(g.finals, o.finals)


def test_full_size_sobol_5m(S, O, table):
    """config C4: 1000 x 5 M Sobol with histogram + P(target)"""
    N, P = 5_000_000, 1000
    mean, std = _gauss_closed_form(P)
    r = S.simulate("sobol_gaussian", N, P, seed=0, hist_bins=100, hist_range=(0.0, 500_000.0), target=100_000.0)
    assert r.count == N and int(r.hist.sum()) == N
    assert abs(r.mean / mean - 1.0) < 1e-4          # QMC: well inside the MC standard error std/sqrt(N)/mean = 1.2e-4
    assert 0.0 < r.p_below_target < 1.0
    g = S.simulate("sobol_gaussian", 512, P, seed=0, want_finals=True, stream_offset=256 * 19_000)
    o = O.simulate("sobol_gaussian", 512, P, seed=0, want_finals=True, stream_offset=256 * 19_000)
    np.testing.assert_allclose(g.finals, o.finals, rtol=GAUSS_RTOL)
    rb = S.simulate("sobol_bootstrap", N, P, returns=table, seed=0, hist_bins=100, hist_range=(0.0, 1e8))
    assert rb.count == N and int(rb.hist.sum()) == N
    gb = S.simulate("sobol_bootstrap", 512, P, returns=table, seed=0, want_finals=True, stream_offset=256 * 19_000)
    ob = O.simulate("sobol_bootstrap", 512, P, returns=table, seed=0, want_finals=True, stream_offset=256 * 19_000)
    assert np.array_equal(gb.finals, ob.finals)


def test_distribution_vs_reference_cpu_generators(S, table):
    """KS agreement of the device finals with the reference's own (unseeded) CPU generators"""
    from test_oracle_pins import _ks_against_quantiles
    g = golden("ref_cpu_quantiles.npz")
    for mode, key in (("bootstrap", "boot_q"), ("bootstrap_compat", "boot_q"), ("gaussian", "gauss_q")):
        r = S.simulate(mode, 1_000_000, 360, want_finals=True, **mode_kwargs(mode, table, seed=42))
        d, crit = _ks_against_quantiles(r.finals, g["qs"], g[key], g["n"])
        assert d < crit, (mode, d, crit)


# ---- the C++ drop-in boundary -------------------------------------------------------------------------------------
CALLER = r'''
#include <cstdio>
#include <stdexcept>
#include "stock_market_monte_carlo/simulations.h"
int main(int argc, char **argv) {
  std::vector<float> returns;
  { FILE *f = fopen(argv[1], "rb"); float x; while (fread(&x, 4, 1, f) == 1) returns.push_back(x); fclose(f); }
  std::atomic<long> n_sim{0};
  std::vector<float> totals, means, variances;
  mc_simulations_gpu(n_sim, 5000L, 36, 1000.f, returns, totals, 1);
  if (n_sim.load() != 5000 || totals.size() != 5000) return 2;
  mc_simulations_gpu_reduceBlock(n_sim, 5000L, 36, 1000.f, returns, means, variances, 1);
  if (means.size() != 20 || variances.size() != 20) return 3;
  bool threw = false;
  try { mc_simulations_gpu_reduceBlock(n_sim, 5000L, 36, 1000.f, returns, means, variances, 2); }
  catch (const std::invalid_argument &) { threw = true; }
  if (!threw) return 4;
  std::vector<float> v(1000); for (int i = 0; i < 1000; i++) v[i] = (float)i;
  if (reduce_mean_gpu(v, 1000) != 499.5f) return 5;
  float a[3] = {1, 2, 3}, b[3] = {10, 20, 30}, o[3], o2[3];
  vector_add(o, a, b, 3); vector_add_gpu(o2, a, b, 3);
  for (int i = 0; i < 3; i++) if (o[i] != o2[i]) return 6;
  FILE *f = fopen(argv[2], "wb");
  fwrite(totals.data(), 4, totals.size(), f); fwrite(means.data(), 4, means.size(), f);
  fwrite(variances.data(), 4, variances.size(), f); fclose(f);
  return 0;
}
'''


def test_cxx_drop_in_program(S, O, table1127, tmp_path):
    """a C++ caller written against the reference's header signatures, linked to libsmmc.so"""
    from stock_market_monte_carlo_b200 import _native
    libdir = os.path.dirname(_native.LIB_PATH)
    (tmp_path / "caller.cpp").write_text(CALLER)
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", str(tmp_path / "caller"), str(tmp_path / "caller.cpp"),
                    f"-I{ROOT}/include", f"-L{libdir}", "-lsmmc", f"-Wl,-rpath,{libdir}"], check=True)
    table1127.tofile(tmp_path / "table.bin")
    subprocess.run([str(tmp_path / "caller"), str(tmp_path / "table.bin"), str(tmp_path / "out.bin")], check=True)
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float32)
    o = O.simulate("bootstrap_compat", 5000, 36, returns=table1127, want_finals=True, want_block_partials=True)
    assert np.array_equal(out[:5000], o.finals)
    assert np.array_equal(out[5000:5020], o.block_means) and np.array_equal(out[5020:], o.block_vars)


def _run_relinked(tmp_path, exe, args, table):
    """run one of the reference's own example programs (unmodified source, relinked against libsmmc.so by
    `make -C oracle relink`) in a scratch directory holding the CSV it expects"""
    import shutil
    (tmp_path / "data").mkdir(exist_ok=True)
    with open(tmp_path / "data" / "SP500_monthly_returns.csv", "w") as f:  # reference src/simulations.cpp:83-93
        f.write("Date,returns\n")
        for x in table:
            f.write(f"1900-01-01,{float(x)!r}\n")
    env = dict(os.environ)
    if os.path.isdir("/usr/lib/locale/C.utf8"):  # the examples ask for en_US.UTF-8, which this image does not ship
        loc = tmp_path / "locale"
        if not loc.exists():
            shutil.copytree("/usr/lib/locale/C.utf8", loc / "en_US.UTF-8")
        env["LOCPATH"] = str(loc)
    p = subprocess.run([exe] + [str(a) for a in args], cwd=tmp_path, env=env, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-2000:]
    return p.stdout


def test_reference_examples_relinked_against_libsmmc(S, O, table1127, tmp_path):
    """the reference's benchmark_mc_gpu / benchmark_mc_gpu_reduceBlock, sources untouched,
    linked to the product library instead of the reference's src/simulations.cu"""
    import re
    exe = os.path.join(ROOT, "oracle", "_ref", "benchmark_mc_gpu")
    exe_rb = os.path.join(ROOT, "oracle", "_ref", "benchmark_mc_gpu_reduceBlock")
    if not (os.path.exists(exe) and os.path.exists(exe_rb)):
        pytest.skip("relinked reference examples did not travel with the snapshot (make -C oracle relink)")
    N, P = 200_000, 360
    out = _run_relinked(tmp_path, exe, [1, P, N], table1127)
    o = O.simulate("bootstrap_compat", N, P, returns=table1127, want_finals=True)
    mean, std = map(float, re.search(r"mean: ([\d.]+) \| std: ([\d.]+)", out).groups())
    fin = o.finals.astype(np.float64)
    # the example's own host statistics (examples/benchmark_mc_gpu.cpp:7-41) over the finals we returned
    assert abs(mean - fin.mean()) <= 0.02 + 1e-6 * fin.mean() and abs(std - fin.std()) <= 0.02 + 1e-5 * fin.std()
    below = int(re.search(r"count_below [\d.,]+: ([\d,]+)", out).group(1).replace(",", ""))
    assert below == int((o.finals < np.float32(1000.0)).sum())
    out = _run_relinked(tmp_path, exe_rb, [1, P, N], table1127)
    m = re.search(r"mean: ([\d.]+) \| std: ([\d.]+)", out)
    assert m, out[-500:]
    ob = O.simulate("bootstrap_compat", N, P, returns=table1127, want_block_partials=True)
    # examples/benchmark_mc_gpu_reduceBlock.cpp:7-26: mean of block means, sqrt(mean of block variances)
    assert abs(float(m.group(1)) - ob.block_means.astype(np.float64).mean()) <= 0.02 + 1e-6 * fin.mean()
