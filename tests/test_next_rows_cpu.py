"""SURVEY 8(f) rows, CPU side: the host logic of the replay radix select (order statistics), the launch planner,
the oracle's trajectory mode, and the data-format edges (returns CSV in, trajectory file out) against the
reference's own functions (live from oracle/_ref when present, and through committed fixtures).  No device."""
import os

import numpy as np
import pytest

from conftest import ALL_MODES, EXACT_MODES, GOLDEN, golden, mode_kwargs


def ordered_keys(x):
    b = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    return np.where(b & np.uint32(0x80000000), ~b, b | np.uint32(0x80000000)).astype(np.uint32)


def emulate_pass(keys, st):
    """what smmc_select_pass computes on the device: per-target digit histograms under the current prefixes"""
    from stock_market_monte_carlo_b200 import _native as N
    bits = int(st.st.bits_done)
    w = 10 if bits >= 22 else 11
    digit = (keys >> np.uint32(32 - bits - w)) & np.uint32((1 << w) - 1)
    hist = np.zeros((N.SELECT_MAX_TARGETS, N.SELECT_BINS), dtype=np.uint64)
    if bits == 0:
        hist[0, :1 << w] = np.bincount(digit, minlength=1 << w)
    else:
        pre = keys >> np.uint32(32 - bits)
        for t in range(int(st.st.n_targets)):
            hist[t, :1 << w] = np.bincount(digit[pre == st.st.prefix[t]], minlength=1 << w)
    return hist


@pytest.mark.parametrize("case", ["lognormal", "mixed_sign", "ties", "single"])
def test_radix_select_protocol_finds_the_exact_order_statistics(S, O, case):
    rng = np.random.default_rng(5)
    x = {"lognormal": rng.lognormal(8.0, 0.6, 20_001),
         "mixed_sign": np.concatenate([rng.normal(0, 50, 5000), [0.0, -0.0, 1e-40, -1e-40, 3e38, -3e38]]),
         "ties": np.repeat(rng.lognormal(8.0, 0.6, 37), 300),
         "single": np.array([1234.5])}[case].astype(np.float32)
    n = x.size
    ranks = sorted(set(list(S.quartile_ranks(n)) + [min(n - 1, 7), n // 3, (2 * n) // 3]))[:8]
    st = S.SelectState(ranks)
    keys = ordered_keys(x)
    passes = 0
    while not st.done:
        st.advance(emulate_pass(keys, st))
        passes += 1
    assert passes == 3
    got = st.values()
    want = O.quantiles_of(x, ranks)  # the reference's way: full sort, index (update_quartiles)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(want, np.sort(x)[np.asarray(ranks)])


def test_radix_select_rejects_ranks_beyond_the_population(S):
    x = np.arange(10, dtype=np.float32)
    st = S.SelectState([10])
    with pytest.raises(S.SmmcError):
        st.advance(emulate_pass(ordered_keys(x), st))
    with pytest.raises(S.SmmcError):
        S.SelectState(list(range(9)))  # more than SMMC_SELECT_MAX_TARGETS
    st = S.SelectState([0])
    with pytest.raises(S.SmmcError):
        st.values()  # no pass run yet


def test_quartile_ranks_are_the_reference_indices(S, O):
    """examples/visualize_returns_gpu.cpp:84-86, :109"""
    for n in (1, 2, 3, 4, 5, 1000, 1001, 10 ** 8, 10 ** 10):
        q1, q2 = n // 4, n // 2
        assert S.quartile_ranks(n) == [0, q1, q2, q1 + q2, n - 1]
        assert list(O.quartile_ranks(n)) == S.quartile_ranks(n)


def test_launch_plan_fills_the_last_wave(S):
    """one warp per 256-path block: 5 M paths are 19 531 blocks; 148 x 32 warps would run 4.12 waves (82 %),
    148 x 27 warps run 4.89 of 5 (98 %); large launches keep 32 warps; results never depend on the choice"""
    grid, threads, copies, smem = S.plan_launch("sobol_gaussian", 5_000_000, 1000, hist_bins=100)
    # 1..48 waves: blocks are handed out by ticket, warps per CTA a multiple of 4 (28: 4.71 of 5 waves)
    assert grid == 148 and threads == 28 * 32 and copies == 0
    assert S.plan_launch("sobol_gaussian", 5_000_000, 1000, hist_bins=100, full=True)["ticket"] is True
    assert S.plan_launch("gaussian", 100_000_000, 360, full=True)["ticket"] is False   # 82 waves: static
    assert S.plan_launch("gaussian", 1000, 360, full=True)["ticket"] is False          # less than one wave: static
    try:
        S.set_variant("ticket", 0)
        assert S.plan_launch("sobol_gaussian", 5_000_000, 1000, hist_bins=100)[1] == 27 * 32  # static: 4.89 of 5 waves
        assert S.plan_launch("sobol_gaussian", 5_000_000, 1000, full=True)["ticket"] is False
    finally:
        S.set_variant()
    for mode in ("gaussian", "bootstrap_compat"):
        g, t, c, sm = S.plan_launch(mode, 100_000_000, 360, n_returns=1200 if "boot" in mode else 0, hist_bins=100)
        assert (g, t) == (148, 1024) and sm <= 232448
    assert S.plan_launch("gaussian", 0, 360)[1] == 1024 and S.plan_launch("gaussian", 1, 360)[1] % 32 == 0
    for n in (1, 255, 4736 * 256, 4737 * 256, 5_000_000, 12_345_678):
        g, t, c, sm = S.plan_launch("gaussian", n, 12)
        nb = (n + 255) // 256
        slots = g * t // 32
        waves = -(-nb // slots)
        full = -(-nb // (g * 32))
        assert 20 * 32 <= t <= 1024 and waves <= full + 1  # never more than one extra wave for a fuller last one


def test_launch_plan_table_replication_and_select_histograms(S):
    assert S.plan_launch("bootstrap", 10 ** 6, 360, n_returns=1200, hist_bins=100)[2] == 32
    assert S.plan_launch("bootstrap", 10 ** 6, 360, n_returns=1200, hist_bins=100, select_targets=8)[2] == 32
    assert S.plan_launch("bootstrap", 10 ** 6, 360, n_returns=1700, hist_bins=100, select_targets=8)[2] == 16
    assert S.plan_launch("bootstrap", 10 ** 6, 360, n_returns=40_000)[2] == 1
    with pytest.raises(S.SmmcError) as e:
        S.plan_launch("bootstrap", 10 ** 6, 360, n_returns=60_000)
    assert e.value.status == 4
    base = S.plan_launch("gaussian", 10 ** 6, 360)[3]
    assert S.plan_launch("gaussian", 10 ** 6, 360, select_targets=5)[3] >= base + 5 * 2048 * 4


# ---- trajectories (oracle) ----------------------------------------------------------------------------------------------

@pytest.mark.parametrize("mode", ALL_MODES)
def test_oracle_trajectories_end_in_the_finals(O, table, mode):
    kw = mode_kwargs(mode, table)
    n, P = 300, 50
    v, r = O.trajectories(mode, n, P, want_returns=True, **kw)
    f = O.simulate(mode, n, P, want_finals=True, **kw).finals
    assert v.shape == (n, P + 1) and r.shape == (n, P)
    assert np.all(v[:, 0] == np.float32(1000.0)) and np.array_equal(v[:, -1], f)
    if "bootstrap" in mode:
        assert np.all(np.isin(r, table))
        # values follow from the returns with the kernel's arithmetic: v' = fma(v, r * 0.01f, v)
        g = (r * np.float32(0.01)).astype(np.float32)
        want = (v[:, :-1].astype(np.float64) * g.astype(np.float64) + v[:, :-1].astype(np.float64)).astype(np.float32)
        assert np.array_equal(want, v[:, 1:])
    else:
        assert abs(r.mean() - 0.5) < 0.05 and abs(r.std() - 10.0 / 12.0) < 0.05


def test_oracle_trajectory_matches_reference_many_updates(O, table):
    """the reference's CPU trajectory (many_updates over update_fund, src/simulations.cpp:14-39) on the SAME drawn
    returns: three roundings per month there, one FFMA here -> agreement to a few ulp per step, not bit-exact"""
    if not O.ref_available("libref_cpu.so"):
        pytest.skip("oracle/_ref not built")
    v, r = O.trajectories("bootstrap", 20, 360, returns=table, seed=3, want_returns=True)
    for p in range(20):
        ref = O.ref_many_updates(1000.0, r[p])
        np.testing.assert_allclose(v[p], ref, rtol=360 * 2.4e-7)


# ---- data-format edges ----------------------------------------------------------------------------------------------

CSV_TEXT = "Date,returns,extra\n1950-01-01, 1.25 ,x\n1950-02-01,-3.5e0,y\r\n1950-03-01,0,z\n\n1950-04-01,+40,w\n"


def test_read_returns_csv(S, O, tmp_path):
    p = tmp_path / "SP500_monthly_returns.csv"
    p.write_text(CSV_TEXT)
    got = S.read_historical_returns(p)
    assert got.dtype == np.float32 and list(got) == [1.25, -3.5, 0.0, 40.0]
    t = S.synthetic_sp500_table(1200)
    q = tmp_path / "table.csv"
    with open(q, "w") as f:
        f.write("Date,returns\n")
        for i, v in enumerate(t):
            f.write(f"{1900 + i // 12}-{i % 12 + 1:02d}-01,{float(v)!r}\n")
    assert np.array_equal(S.read_historical_returns(q), t)
    if O.ref_available("libref_cpu.so"):  # the reference's own reader on the same files (blank line removed: it rejects them)
        assert np.array_equal(O.ref_read_historical_returns(q), t)
        p2 = tmp_path / "noblank.csv"
        p2.write_text(CSV_TEXT.replace("\n\n", "\n"))
        assert list(O.ref_read_historical_returns(p2)) == [1.25, -3.5, 0.0, 40.0]


def test_read_returns_csv_errors(S, tmp_path):
    for name, text in (("nocol.csv", "Date,ret\n1,2\n"), ("bad.csv", "returns\nabc\n"), ("cols.csv", "a,returns\n1\n"),
                       ("empty.csv", "")):
        p = tmp_path / name
        p.write_text(text)
        with pytest.raises(S.SmmcError) as e:
            S.read_historical_returns(p)
        assert e.value.status == 1
    with pytest.raises(S.SmmcError):
        S.read_historical_returns(tmp_path / "missing.csv")
    p = tmp_path / "header_only.csv"
    p.write_text("returns\n")
    assert S.read_historical_returns(p).size == 0


def _data_file_case():
    r = np.array([1.5, -2.25, 0.0, 1e-7, 123456.789, -40.0, 3.14159274], dtype=np.float32)
    v = np.concatenate([[1000.0], 1000.0 * np.cumprod(1 + r.astype(np.float64) / 100)]).astype(np.float32)
    return r, v


def test_write_data_file_matches_reference_bytes(S, O, tmp_path):
    r, v = _data_file_case()
    out = tmp_path / "traj.csv"
    S.write_data_file(out, r, v)
    mine = out.read_bytes()
    assert mine == open(os.path.join(GOLDEN, "ref_write_data_file.csv"), "rb").read()
    assert mine.startswith(b"Returns,,1.5,-2.25,0,1e-07,123457,-40,3.14159,\nValues,1000,1015,")
    if O.ref_available("libref_cpu.so"):
        cwd = os.getcwd()
        os.chdir(tmp_path)
        try:
            O.ref_write_data_file("ref.csv", r, v)
        finally:
            os.chdir(cwd)
        assert (tmp_path / "outputs" / "ref.csv").read_bytes() == mine
    # the reference's reader of this format (python/plot_returns.py:40) is pandas: index_col=0, header=None, .T[:-1]
    import pandas as pd
    df = pd.read_csv(out, index_col=0, header=None).T[:-1]
    assert np.allclose(df["Values"].astype(float).to_numpy()[: v.size], v, rtol=1e-5)


def test_write_data_file_empty_and_errors(S, tmp_path):
    out = tmp_path / "e.csv"
    S.write_data_file(out, [], [])
    assert out.read_bytes() == b"Returns,,\nValues,"
    with pytest.raises(S.SmmcError):
        S.write_data_file(tmp_path / "no_such_dir" / "x.csv", [1.0], [1.0])
