"""SURVEY 8(f) rows on the GPU, through the C ABI: exact order statistics by replay radix select (the reference sorts
the finals on the host: update_quartiles, examples/visualize_returns_gpu.cpp:83-110), trajectory mode (the
reference's mc_simulations_keepdata, src/simulations.cpp:139-202), and launches with fewer than 32 warps per CTA.

Bar: order statistics are EXACTLY sorted(finals)[rank] -- against the device's own finals in every mode (a replay
reproduces every path bit for bit) and against the oracle's finals in the table-driven modes; trajectories are
bit-exact against the oracle in the table-driven modes, within the stated fp32 tolerance in the Gaussian modes,
and their last column equals the statistics kernel's finals bit for bit in every mode."""
import numpy as np
import pytest

from conftest import ALL_MODES, EXACT_MODES, mode_kwargs
from test_gpu_parity import GAUSS_RTOL

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ALL_MODES)
@pytest.mark.parametrize("N,P", [(1, 12), (257, 7), (10_000, 120), (300_001, 36)])
def test_quartiles_equal_sorted_finals(S, O, table, mode, N, P):
    kw = mode_kwargs(mode, table)
    q, ms = S.quantiles(mode, N, P, return_ms=True, **kw)
    finals = S.simulate(mode, N, P, want_finals=True, **kw).finals
    ranks = S.quartile_ranks(N)
    want = np.sort(finals)[ranks]
    assert np.array_equal(q.view(np.uint32), want.view(np.uint32)) and ms > 0.0
    if mode in EXACT_MODES:  # the reference's way on the oracle's finals: std::sort, then index
        o = O.simulate(mode, N, P, want_finals=True, **kw).finals
        assert np.array_equal(q, O.quantiles_of(o, O.quartile_ranks(N)))
    quart = S.FloatVector()
    S.update_quartiles(quart, mode, N, P, **kw)
    assert np.array_equal(quart.array, q)


def test_quantiles_arbitrary_ranks_ties_and_negative_values(S, O):
    # a table with a -150 % month drives values negative; a 2-entry table makes massive ties
    tbl = np.array([-150.0, 10.0, 3.0, -2.0, 0.5], dtype=np.float32)
    N, P = 50_000, 9
    ranks = [0, 1, 17, N // 7, N // 2, N - 2, N - 1, N // 2]
    q = S.quantiles("bootstrap", N, P, ranks=ranks, returns=tbl, seed=3)
    f = S.simulate("bootstrap", N, P, want_finals=True, returns=tbl, seed=3).finals
    assert f.min() < 0 < f.max()
    assert np.array_equal(q, np.sort(f)[ranks])
    assert np.array_equal(q, O.quantiles_of(O.simulate("bootstrap", N, P, want_finals=True, returns=tbl, seed=3).finals, ranks))
    tbl2 = np.array([1.0, -1.0], dtype=np.float32)
    q2 = S.quantiles("bootstrap", N, 4, ranks=ranks, returns=tbl2, seed=3)
    f2 = S.simulate("bootstrap", N, 4, want_finals=True, returns=tbl2, seed=3).finals
    assert np.unique(f2).size <= 16 and np.array_equal(q2, np.sort(f2)[ranks])
    with pytest.raises(S.SmmcError):
        S.quantiles("gaussian", 100, 5, ranks=[100])
    with pytest.raises(S.SmmcError):
        S.quantiles("gaussian", 100, 5, ranks=list(range(9)))


def test_quantiles_do_not_depend_on_sharding_or_offset(S, table):
    N, P = 100_000, 24
    whole = S.quantiles("gaussian", N, P, seed=5, stream_offset=1 << 33)
    # the pass protocol with two shards whose histograms are summed = what two ranks + all_reduce do
    st = S.SelectState(S.quartile_ranks(N))
    while not st.done:
        h = None
        for off, cnt in ((0, 60_160), (60_160, N - 60_160)):
            cfg, keep = S.make_config("gaussian", cnt, P, seed=5, stream_offset=(1 << 33) + off)
            part = st.run_pass(cfg)
            h = part if h is None else h + part
        st.advance(h)
    assert np.array_equal(st.values(), whole)


def test_select_pass_device_pointers_async(S, table):
    import torch
    from stock_market_monte_carlo_b200 import _native as Nn
    from stock_market_monte_carlo_b200.distributed import quantiles_sharded
    N, P = 70_000, 30
    got = quantiles_sharded("bootstrap", N, P, device=0, returns=table, seed=9)
    assert np.array_equal(got, S.quantiles("bootstrap", N, P, returns=table, seed=9))
    # first-pass histogram on the device: row 0 counts every path, the other rows are zeroed
    dev = torch.device("cuda", 0)
    hist = torch.full((Nn.SELECT_MAX_TARGETS * Nn.SELECT_BINS,), 7, dtype=torch.int64, device=dev)
    tdev = torch.from_numpy(table).to(dev)
    cfg, _ = S.make_config("bootstrap", N, P, seed=9, flags=Nn.FLAG_DEVICE_POINTERS,
                           stream=torch.cuda.current_stream(dev).cuda_stream, returns_ptr=tdev.data_ptr(), n_returns=table.size)
    S.SelectState([0, N - 1]).run_pass(cfg, hist_ptr=hist.data_ptr())
    h = hist.cpu().numpy().reshape(Nn.SELECT_MAX_TARGETS, Nn.SELECT_BINS)
    assert h[0].sum() == N and h[1:].sum() == 0


def test_quartiles_at_baseline_size(S, table):
    """C2-sized: 360 months x 100 M Gaussian paths; 3 replays instead of a 400 MB download and a host sort"""
    N, P = 100_000_000, 360
    q, ms = S.quantiles("gaussian", N, P, seed=1234, return_ms=True)
    assert np.all(np.diff(q) > 0) and ms > 0.0
    # median of a product of 360 factors 1.005 + N(0, 1/120): exp(360 * E[ln g]) ~ 5948.6; quartiles ~ exp(mu_ln -+ 0.6745 s_ln)
    z, w = np.polynomial.hermite_e.hermegauss(64)  # E over N(0,1) by Gauss-Hermite quadrature
    w = w / w.sum()
    lg = np.log(1.005 + z / 120)
    m1 = (w * lg).sum()
    mu_ln, s_ln = 360 * m1, np.sqrt(360 * (w * (lg - m1) ** 2).sum())
    want = 1000 * np.exp(mu_ln + s_ln * np.array([-0.674490, 0.0, 0.674490]))
    np.testing.assert_allclose(q[1:4], want, rtol=2e-3)
    # histogram cross-check: the count below each quartile value, from an independent run of the statistics kernel
    for k, frac in ((1, 0.25), (2, 0.5), (3, 0.75)):
        r = S.simulate("gaussian", N, P, seed=1234, target=float(q[k]))
        assert r.below_count <= S.quartile_ranks(N)[k] < r.below_count + 50  # ties at one float value are rare
        assert abs(r.below_count / N - frac) < 1e-6 + 1e-9


# ---- trajectories ---------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("mode", ALL_MODES)
@pytest.mark.parametrize("n,P,off", [(1, 1, 0), (300, 50, 0), (77, 361, 512), (1000, 360, 1 << 20)])
def test_trajectories_against_oracle_and_finals(S, O, table, mode, n, P, off):
    kw = mode_kwargs(mode, table)
    v, r = S.trajectories(mode, n, P, stream_offset=off, want_returns=True, **kw)
    ov, orr = O.trajectories(mode, n, P, stream_offset=off, want_returns=True, **kw)
    assert v.shape == (n, P + 1) and r.shape == (n, P)
    if mode in EXACT_MODES:
        assert np.array_equal(v, ov) and np.array_equal(r, orr)
    else:
        np.testing.assert_allclose(v, ov, rtol=GAUSS_RTOL)
        np.testing.assert_allclose(r, orr, rtol=2e-4, atol=2e-5)  # 100 (g - 1): cancellation amplifies the 1-ulp differences of g
    # same streams, same arithmetic as the statistics kernel: last column == finals, bit for bit, in every mode
    f = S.simulate(mode, n, P, stream_offset=off, want_finals=True, **kw).finals
    assert np.array_equal(v[:, -1].view(np.uint32), f.view(np.uint32))
    assert np.all(v[:, 0] == np.float32(1000.0))
    assert np.array_equal(S.trajectories(mode, n, P, stream_offset=off, **kw), v)


def test_trajectories_zero_periods_and_validation(S, table):
    v = S.trajectories("gaussian", 5, 0)
    assert v.shape == (5, 1) and np.all(v == 1000.0)
    assert S.trajectories("gaussian", 0, 10).shape == (0, 11)
    with pytest.raises(S.SmmcError):
        S.trajectories("bootstrap", 5, 10)  # no table
    with pytest.raises(S.SmmcError):
        S.trajectories("sobol_gaussian", 5, 1112)


def test_mc_simulations_keepdata_mirror(S, O, table1127):
    """reference call shape (src/simulations.cpp:139-202): mc_data[i] = P+1 values, final_values[i] = the last"""
    n_sim = S.AtomicLong(0)
    N, P = 1000, 120
    mc_data = [None] * N
    finals = S.FloatVector()
    S.mc_simulations_keepdata(n_sim, N, P, 1000.0, S.FloatVector(table1127), mc_data, finals)
    assert n_sim.load() == N and len(mc_data) == N and all(len(row) == P + 1 for row in mc_data)
    totals = S.FloatVector()
    S.mc_simulations_gpu(S.AtomicLong(0), N, P, 1000.0, S.FloatVector(table1127), totals, 1)
    assert np.array_equal(finals.array, totals.array)  # the same paths the drop-in entry point simulates
    assert np.array_equal(np.stack(mc_data), O.trajectories("bootstrap_compat", N, P, returns=table1127))


def test_trajectory_file_roundtrip_through_reference_format(S, table, tmp_path):
    """one path -> write_data_file -> the reference's pandas reader (python/plot_returns.py:40)"""
    import pandas as pd
    v, r = S.trajectories("bootstrap", 3, 24, returns=table, seed=1, want_returns=True)
    out = tmp_path / "historical_00002.csv"
    S.write_data_file(out, r[2], v[2])
    df = pd.read_csv(out, index_col=0, header=None).T[:-1]
    np.testing.assert_allclose(df["Values"].astype(float).to_numpy()[:25], v[2], rtol=1e-5)


# ---- warps per CTA ---------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("mode", ALL_MODES)
def test_results_do_not_depend_on_warps_per_cta(S, O, table, mode):
    kw = mode_kwargs(mode, table)
    N, P = 148 * 27 * 256 + 300, 12  # just over one wave of 27-warp CTAs
    base = S.simulate(mode, N, P, hist_bins=64, want_finals=True, want_block_partials=True, **kw)
    try:
        for w in (20, 27, 32):
            S.set_variant("warps", w)
            r = S.simulate(mode, N, P, hist_bins=64, want_finals=True, want_block_partials=True, **kw)
            assert np.array_equal(r.finals.view(np.uint32), base.finals.view(np.uint32))
            assert np.array_equal(r.block_means.view(np.uint32), base.block_means.view(np.uint32))
            assert np.array_equal(r.hist, base.hist) and r.below_count == base.below_count
            assert abs(r.mean - base.mean) <= 1e-12 * abs(base.mean) and abs(r.m2 - base.m2) <= 1e-9 * abs(base.m2)
    finally:
        S.set_variant()


# ---- the C++ side of the same rows -----------------------------------------------------------------------------------
CALLER = r'''
#include <cstdio>
#include <algorithm>
#include "stock_market_monte_carlo/simulations.h"
int main(int argc, char **argv) {
  std::vector<float> returns = smmc::read_historical_returns(argv[1]);
  if (returns.size() != 1127) return 2;
  std::atomic<long> n_sim{0};
  const long N = 4000; const int P = 60;
  std::vector<float> totals, quart, finals;
  mc_simulations_gpu(n_sim, N, P, 1000.f, returns, totals, 1);
  smmc::update_quartiles(quart, N, P, 1000.f, returns);
  std::sort(totals.begin(), totals.end());  // the reference's update_quartiles (examples/visualize_returns_gpu.cpp:108-109)
  const float want[5] = {totals.at(0), totals.at(N / 4), totals.at(N / 2), totals.at(N / 4 + N / 2), totals.back()};
  if (quart.size() != 5) return 3;
  for (int i = 0; i < 5; i++) if (quart[i] != want[i]) return 4;
  std::vector<std::vector<float>> mc_data;
  smmc::mc_simulations_keepdata(n_sim, 50, P, 1000.f, returns, mc_data, finals);
  if (mc_data.size() != 50 || mc_data[7].size() != (size_t)P + 1 || finals[7] != mc_data[7].back() || mc_data[7][0] != 1000.f) return 5;
  std::vector<float> r(returns.begin(), returns.begin() + P);
  smmc::write_data_file(argv[2], r, mc_data[7]);
  bool threw = false;
  try { smmc::read_historical_returns("/nonexistent.csv"); } catch (const std::runtime_error &) { threw = true; }
  return threw ? 0 : 6;
}
'''


def test_cxx_callers_of_the_next_rows(S, table1127, tmp_path):
    import os
    import subprocess
    from conftest import ROOT
    from stock_market_monte_carlo_b200 import _native
    libdir = os.path.dirname(_native.LIB_PATH)
    (tmp_path / "caller.cpp").write_text(CALLER)
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", str(tmp_path / "caller"), str(tmp_path / "caller.cpp"),
                    f"-I{ROOT}/include", f"-L{libdir}", "-lsmmc", f"-Wl,-rpath,{libdir}"], check=True)
    with open(tmp_path / "t.csv", "w") as f:
        f.write("Date,returns\n")
        for x in table1127:
            f.write(f"1900-01-01,{float(x)!r}\n")
    subprocess.run([str(tmp_path / "caller"), str(tmp_path / "t.csv"), str(tmp_path / "out.csv")], check=True)
    text = (tmp_path / "out.csv").read_text()
    assert text.startswith("Returns,,") and "\nValues,1000," in text
    v = S.trajectories("bootstrap_compat", 50, 60, returns=table1127)[7]
    got = np.array([float(t) for t in text.split("\nValues,")[1].rstrip(",").split(",")])
    np.testing.assert_allclose(got, v, rtol=1e-5)


def test_more_than_2_pow_32_paths_in_one_call(S):
    """64-bit block / path indexing: 5e9 paths x 1 month in ONE call (no per-path output); closed-form mean, every
    path counted once in the histogram, and the tail of the range equals the same paths run on their own"""
    N = 5_000_000_000
    r = S.simulate("gaussian", N, 1, seed=3, hist_bins=8, hist_range=(950.0, 1050.0))
    assert r.count == N and int(r.hist.sum()) == N
    assert abs(r.mean - 1005.0) < 5 * (1000 / 120) / np.sqrt(N) + 1e-4   # E = V0 (1 + mu), sd = V0 sigma
    assert abs(r.std - 1000 / 120) < 1e-3
    head = S.simulate("gaussian", N - 1_000_192, 1, seed=3, hist_bins=8, hist_range=(950.0, 1050.0))
    tail = S.simulate("gaussian", 1_000_192, 1, seed=3, stream_offset=N - 1_000_192, hist_bins=8, hist_range=(950.0, 1050.0))
    assert np.array_equal(head.hist + tail.hist, r.hist) and head.below_count + tail.below_count == r.below_count


def test_c5_size_ten_billion_paths_closed_form_moments(S):
    """BASELINE config 5 on one GPU: 360 months x 10^10 Gaussian paths + histogram in ONE call (1.8 s): moments against
    the closed forms E[V] = V0 (1+mu)^P, E[V^2] = V0^2 ((1+mu)^2 + sigma^2)^P (north_star: <= 1e-5 relative)"""
    N, P = 10_000_000_000, 360
    r = S.simulate("gaussian", N, P, seed=1234, hist_bins=100, hist_range=(0.0, 50_000.0))
    mu, sg = 0.005, 1.0 / 120.0
    mean = 1000.0 * (1 + mu) ** P
    var = 1000.0 ** 2 * ((1 + mu) ** 2 + sg ** 2) ** P - mean ** 2
    assert r.count == N and int(r.hist.sum()) == N
    se_rel = np.sqrt(var / N) / mean                       # 1.6e-6
    assert abs(r.mean - mean) / mean < 5 * se_rel + 2e-6     # + the fp32 compounding bias budget
    assert abs(r.std - np.sqrt(var)) / np.sqrt(var) < 1e-5   # north_star bound; this seed measures +2.8e-6 (the std's own standard error at 1e10 paths is 7.7e-6)


def test_fused_benchmark_program_prints_the_reference_programs_statistics(S, O, table1127, tmp_path):
    """examples/benchmark_mc_fused.cpp: the reference's benchmark CLI on the fused entry points -- same mean / std /
    count-below lines as examples/benchmark_mc_gpu.cpp computes on the host from N finals, plus exact quartiles"""
    import os
    import re
    import subprocess
    from conftest import ROOT
    from stock_market_monte_carlo_b200 import _native
    libdir = os.path.dirname(_native.LIB_PATH)
    exe = tmp_path / "benchmark_mc_fused"
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", str(exe), os.path.join(ROOT, "examples", "benchmark_mc_fused.cpp"),
                    f"-I{ROOT}/include", f"-L{libdir}", "-lsmmc", f"-Wl,-rpath,{libdir}"], check=True)
    (tmp_path / "data").mkdir()
    with open(tmp_path / "data" / "SP500_monthly_returns.csv", "w") as f:
        f.write("Date,returns\n")
        for x in table1127:
            f.write(f"1900-01-01,{float(x)!r}\n")
    N, P = 200_000, 360
    out = subprocess.run([str(exe), "1", str(P), str(N)], cwd=tmp_path, capture_output=True, text=True, check=True).stdout
    fin = O.simulate("bootstrap_compat", N, P, returns=table1127, want_finals=True).finals
    mean, std = map(float, re.search(r"mean: ([\d.]+) \| std: ([\d.]+)", out).groups())
    x = fin.astype(np.float64)
    assert abs(mean - x.mean()) <= 0.006 + 1e-6 * x.mean() and abs(std - x.std()) <= 0.006 + 1e-5 * x.std()
    below = int(re.search(r"count_below [\d.]+: (\d+)", out).group(1))
    assert below == int((fin < np.float32(1000.0)).sum())
    q = [float(v) for v in re.search(r"min: ([\d.]+) \| Q1: ([\d.]+) \| median: ([\d.]+) \| Q3: ([\d.]+) \| max: ([\d.]+)", out).groups()]
    want = np.sort(fin)[S.quartile_ranks(N)].astype(np.float64)
    np.testing.assert_allclose(q, want, rtol=0, atol=0.006 + 1e-7 * want.max())
    out = subprocess.run([str(exe), "1", "120", "100000", "gaussian", "7"], cwd=tmp_path, capture_output=True, text=True,
                         check=True).stdout
    g = S.simulate("gaussian", 100_000, 120, seed=7)
    assert abs(float(re.search(r"mean: ([\d.]+)", out).group(1)) - g.mean) < 0.006

@pytest.mark.parametrize("mode", ALL_MODES)
def test_first_select_pass_rides_along(S, table, mode):
    """smmc_run_select: statistics + quartiles in three replays -- the statistics are bit-identical to smmc_run's, the
    order statistics to smmc_quantiles' (and so to the sorted finals)"""
    kw = mode_kwargs(mode, table)
    N, P = 300_001, 45
    full = dict(hist_bins=64, hist_range=(0.0, 20_000.0), want_block_partials=True, want_finals=True, **kw)
    plain = S.simulate(mode, N, P, **full)
    r, q, extra_ms = S.simulate_with_quantiles(mode, N, P, **full)
    assert (r.count, r.below_count, r.mean, r.m2) == (plain.count, plain.below_count, plain.mean, plain.m2)
    assert np.array_equal(r.hist, plain.hist) and np.array_equal(r.finals, plain.finals)
    assert np.array_equal(r.block_means, plain.block_means) and np.array_equal(r.block_vars, plain.block_vars)
    assert int(r.select_hist[:2048].sum()) == N and not r.select_hist[2048:].any()
    assert np.array_equal(q, np.sort(plain.finals)[S.quartile_ranks(N)])
    assert np.array_equal(q, S.quantiles(mode, N, P, **kw))
    assert extra_ms > 0.0
    ranks = [5, 5, N - 2]
    _, q2, _ = S.simulate_with_quantiles(mode, N, P, ranks=ranks, **kw)
    assert np.array_equal(q2, np.sort(plain.finals)[ranks])
