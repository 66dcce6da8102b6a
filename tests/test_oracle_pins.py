"""Pins the CPU oracle (oracle/smmc_oracle.c) before anything is checked against it.

Pin sources, strongest first:
  * the reference itself, run in the build container -> committed fixtures (tests/golden/make_golden.py) and,
    when oracle/_ref is present, live calls into the compiled reference;
  * the reference's own CUDA kernels run on a B200 -> tests/golden/ref_gpu_compat.npz (make_golden_gpu.py);
  * published known-answer vectors (Random123 Philox4x32-10);
  * closed-form moments (RNG-independent).
"""
import numpy as np
import pytest

from conftest import ALL_MODES, golden, mode_kwargs


# ---- Philox4x32-10: Random123 known-answer vectors (Random123 tests/kat_vectors, "philox4x32 10") -----------------
PHILOX_KAT = [
    ([0x00000000] * 4, [0x00000000] * 2, [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
    ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
    ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
     [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
]


@pytest.mark.parametrize("ctr,key,want", PHILOX_KAT)
def test_philox_known_answers(O, ctr, key, want):
    assert [int(x) for x in O.philox([ctr], key)[0]] == want


# Random123 kat_vectors, "philox4x32 7": the opt-in round count of ABI 3 (smmc_config.philox_rounds = 7)
PHILOX7_KAT = [
    ([0x00000000] * 4, [0x00000000] * 2, [0x5F6FB709, 0x0D893F64, 0x4F121F81, 0x4F730A48]),
    ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x5207DDC2, 0x45165E59, 0x4D8EE751, 0x8C52F662]),
    ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
     [0x4DFCCABA, 0x190A87F0, 0xC47362BA, 0xB6B5242A]),
]


@pytest.mark.parametrize("ctr,key,want", PHILOX7_KAT)
def test_philox7_known_answers(O, ctr, key, want):
    assert [int(x) for x in O.philox([ctr], key, rounds=7)[0]] == want


def test_philox_rounds_select_the_stream_in_the_oracle(O, table):
    """7 and 10 rounds are different streams of the same contract: draws differ, 0 means 10, and both reproduce the
    closed-form mean of the bootstrap (RNG-independent) within a few standard errors."""
    kw = dict(seed=11, want_finals=True, returns=table)
    a = O.simulate("bootstrap", 40_000, 120, philox_rounds=10, **kw)
    b = O.simulate("bootstrap", 40_000, 120, philox_rounds=7, **kw)
    c = O.simulate("bootstrap", 40_000, 120, philox_rounds=0, **kw)
    assert np.array_equal(a.finals, c.finals)
    assert not np.array_equal(a.finals, b.finals)
    m1 = 1.0 + float(np.mean(table.astype(np.float64) * 0.01))
    want = 1000.0 * m1 ** 120
    for r in (a, b):
        se = r.std / np.sqrt(40_000)
        assert abs(r.mean - want) < 5 * se
    d10 = O.draws("gaussian", 300, 16, seed=3, philox_rounds=10)
    d7 = O.draws("gaussian", 300, 16, seed=3, philox_rounds=7)
    assert d10.shape == d7.shape and not np.array_equal(d10, d7)


def test_philox_matches_numpy_generator(O):
    """numpy's Philox is Philox4x64, not 4x32, so it cannot serve; instead check the round structure against an
    independent pure-Python restatement of the published algorithm on random counters."""
    rng = np.random.default_rng(0)
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
    for _ in range(50):
        c = [int(x) for x in rng.integers(0, 2 ** 32, 4)]
        k = [int(x) for x in rng.integers(0, 2 ** 32, 2)]
        c0, c1, c2, c3 = c
        k0, k1 = k
        for _r in range(10):
            p0, p1 = M0 * c0, M1 * c2
            c0, c1, c2, c3 = (p1 >> 32) ^ c1 ^ k0, p1 & 0xFFFFFFFF, (p0 >> 32) ^ c3 ^ k1, p0 & 0xFFFFFFFF
            k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
        assert [int(x) for x in O.philox([c], k)[0]] == [c0, c1, c2, c3]


# ---- reference PRNG pieces (reference src/simulations.cu:62-75, :148) ----------------------------------------------
def _rand_pcg_py(s):
    word = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & 0xFFFFFFFF
    return ((word >> 22) ^ word) & 0xFFFFFFFF


def _xorshift_py(y):
    y ^= (y << 11) & 0xFFFFFFFF
    y ^= y >> 7
    y ^= y >> 12
    return y


def test_reference_prng_pieces(O):
    lib = O.load()
    rng = np.random.default_rng(1)
    for s in [0, 1, 2, 0xFFFFFFFF, 0x80000000] + [int(x) for x in rng.integers(0, 2 ** 32, 200)]:
        assert lib.orc_rand_pcg(s) == _rand_pcg_py(s)
        assert lib.orc_xorshift(s) == _xorshift_py(s)
    assert lib.orc_xorshift(0) == 0  # the reference stream's fixed point (SURVEY hard parts)


def test_compat_index_map_edge(O):
    """idx = (int)(n * (float(state) * 2^-32f)): states >= 2^32-128 round to 2^32 and give idx == n (one past the
    end; reference src/simulations.cu:148)."""
    lib = O.load()
    n = 1127
    assert lib.orc_compat_index(0, n) == 0
    assert lib.orc_compat_index(0xFFFFFFFF, n) == n
    assert lib.orc_compat_index(0xFFFFFF80, n) == n
    assert lib.orc_compat_index(0xFFFFFF7F, n) == n - 1
    st = np.random.default_rng(2).integers(0, 2 ** 32, 2000, dtype=np.uint64)
    want = (np.float32(n) * (st.astype(np.float32) * np.float32(2.3283064e-10))).astype(np.int32)
    got = np.array([lib.orc_compat_index(int(s), n) for s in st], dtype=np.int32)
    assert np.array_equal(got, want)


def test_compat_against_reference_kernels_on_b200(O, S):
    """The fixture holds the output of the reference's own kernels (src/simulations.cu:112-152, :185-247) compiled
    unmodified for sm_100a and run on a B200: finals and reduceBlock partials must match the oracle BIT FOR BIT."""
    g = golden("ref_gpu_compat.npz")
    tbl = S.synthetic_sp500_table(int(g["table_n"]), seed=int(g["table_seed"]))
    n = int(g["n"])
    o = O.simulate("bootstrap_compat", n, 360, returns=tbl, want_finals=True, want_block_partials=True)
    assert np.array_equal(o.finals, g["totals"])
    assert np.array_equal(o.block_means, g["means"])
    assert np.array_equal(o.block_vars, g["variances"])
    o1 = O.simulate("bootstrap_compat", n, 1, returns=tbl, want_finals=True)
    assert np.array_equal(o1.finals, g["totals_p1"])


# ---- Sobol: the reference's vendored Burkardt generator ------------------------------------------------------------
def test_sobol_points_against_reference_fixture(O):
    g = golden("sobol_i4_reference.npz")
    lib = O.load()
    dims, x30 = g["dims"], g["x30"]
    for j, d in enumerate(dims):
        got = np.array([lib.orc_sobol_point(i, int(d)) for i in range(x30.shape[0])], dtype=np.uint32)
        assert np.array_equal(got >> 2, x30[:, j]), f"dimension {d}"
        assert not np.any(got & 3)  # i4_sobol words are 30 bits wide (2^-30 scale, reference sobol.cpp:325)


def test_sobol_points_against_compiled_reference(O):
    if not O.ref_available("libref_sobol.so"):
        pytest.skip("oracle/_ref/libref_sobol.so not built")
    lib = O.load()
    for i in (0, 1, 2, 3, 4, 255, 256, 257, 4095, 65536, 1 << 20, 4999999 + 256):
        q = O.ref_i4_sobol(1111, i).astype(np.float64) * float(1 << 30)
        got = np.array([lib.orc_sobol_point(i, d) for d in range(1111)], dtype=np.uint32)
        assert np.array_equal((got >> 2).astype(np.float64), q)


def test_sobol_survey_probe_points(O):
    """SURVEY 8(a) A8, probed with the reference's demo: points 1..4 of dims 1..4."""
    lib = O.load()
    want = {1: (.5, .5, .5, .5), 2: (.75, .25, .75, .25), 3: (.25, .75, .25, .75), 4: (.375, .375, .625, .125)}
    for i, row in want.items():
        assert tuple(lib.orc_sobol_point(i, d) / 2.0 ** 32 for d in range(4)) == row


# ---- the reference's CPU arithmetic ------------------------------------------------------------------------------
def test_update_fund_fixture_vs_growth_factor():
    """update_fund = fund * (100 + r) / 100 with three roundings (reference src/simulations.cpp:14-16); the device
    compounds with one FFMA, total += total * (r * 0.01f) (src/simulations.cu:126, :149).  They agree to a few ulp
    per step, never exactly: this pins the size of that documented drift."""
    g = golden("ref_update_fund.npz")
    fund, ret, out = g["fund"], g["ret"], g["out"]
    three = ((fund * (np.float32(100.0) + ret)) / np.float32(100.0)).astype(np.float32)
    assert np.array_equal(three, out)
    one = (fund.astype(np.float64) * (1.0 + (ret * np.float32(0.01)).astype(np.float64))).astype(np.float32)
    assert np.max(np.abs(one - out) / out) < 4e-7


def _ks_against_quantiles(sample, qs, qv, n_ref):
    """two-sample KS statistic between `sample` and a reference sample summarised by quantiles (qs, qv)"""
    s = np.sort(np.asarray(sample, dtype=np.float64))
    cdf_s_at_qv = np.searchsorted(s, qv, side="right") / s.size
    d = np.max(np.abs(cdf_s_at_qv - qs))
    n, m = s.size, int(n_ref)
    crit = 1.95 * np.sqrt((n + m) / (n * m)) + 1.0 / (qs.size - 1)  # alpha = 1e-3, plus the quantile grid step
    return d, crit


@pytest.mark.parametrize("mode", ["bootstrap_compat", "bootstrap", "sobol_bootstrap"])
def test_bootstrap_distribution_vs_reference_cpu_v2(O, table, mode):
    """reference CPU v2 (mc_simulations, src/simulations.cpp:204-266) takes no seed: distributional agreement."""
    g = golden("ref_cpu_quantiles.npz")
    r = O.simulate(mode, 200_000, 360, want_finals=True, **mode_kwargs(mode, table))
    d, crit = _ks_against_quantiles(r.finals, g["qs"], g["boot_q"], g["n"])
    assert d < crit, (d, crit)
    se = float(g["boot_std"]) * np.sqrt(1 / 200_000 + 1 / float(g["n"]))
    tol = 1e-3 if mode == "sobol_bootstrap" else 5 * se  # QMC mean is far tighter than MC; compare to MC ref anyway
    assert abs(r.mean - float(g["boot_mean"])) < max(5 * se, tol)


@pytest.mark.parametrize("mode", ["gaussian", "sobol_gaussian"])
def test_gaussian_distribution_vs_reference_cpu(O, mode):
    """reference sample_returns_gaussian + many_updates (src/simulations.cpp:41-55, :24-39)."""
    g = golden("ref_cpu_quantiles.npz")
    r = O.simulate(mode, 200_000, 360, want_finals=True, seed=77)
    d, crit = _ks_against_quantiles(r.finals, g["qs"], g["gauss_q"], g["n"])
    assert d < crit, (d, crit)
    se = float(g["gauss_std"]) * np.sqrt(1 / 200_000 + 1 / float(g["n"]))
    assert abs(r.mean - float(g["gauss_mean"])) < 5 * se
    assert abs(r.std / float(g["gauss_std"]) - 1.0) < 0.02


def test_live_reference_cpu_when_present(O, table):
    if not O.ref_available("libref_cpu.so"):
        pytest.skip("oracle/_ref/libref_cpu.so not built")
    from scipy import stats
    ref, _ = O.ref_mc_simulations(50_000, 120, 1000.0, table)
    r = O.simulate("bootstrap", 50_000, 120, returns=table, seed=3, want_finals=True)
    assert stats.ks_2samp(ref, r.finals).pvalue > 1e-3
    refg, _ = O.ref_gaussian_finals(50_000, 120, 1000.0, 0.5, 10.0 / 12.0)
    rg = O.simulate("gaussian", 50_000, 120, seed=3, want_finals=True)
    assert stats.ks_2samp(refg, rg.finals).pvalue > 1e-3


# ---- closed-form moments --------------------------------------------------------------------------
def _closed_form_bootstrap(table, P, v0=1000.0):
    t = (table * np.float32(0.01)).astype(np.float64)
    m1 = t.mean()
    m2 = ((1.0 + t) ** 2).mean()
    mean = v0 * (1 + m1) ** P
    var = v0 ** 2 * m2 ** P - mean ** 2
    return mean, np.sqrt(var)


@pytest.mark.parametrize("mode", ["bootstrap_compat", "bootstrap"])
def test_closed_form_bootstrap(O, table, mode):
    N, P = 400_000, 60
    mean, std = _closed_form_bootstrap(table, P)
    r = O.simulate(mode, N, P, **mode_kwargs(mode, table))
    assert abs(r.mean - mean) < 5 * std / np.sqrt(N)
    assert abs(r.std / std - 1.0) < 0.02


def test_closed_form_gaussian(O):
    N, P = 400_000, 360
    mu, sigma = np.float32(0.5) * np.float32(0.01), np.float32(10.0 / 12.0) * np.float32(0.01)
    mean = 1000.0 * (1.0 + float(mu)) ** P
    std = np.sqrt(1000.0 ** 2 * ((1.0 + float(mu)) ** 2 + float(sigma) ** 2) ** P - mean ** 2)
    assert abs(mean - 6022.575) < 0.01 and abs(std - 953.4) < 0.5  # the SURVEY's worked example
    r = O.simulate("gaussian", N, P, seed=5)
    assert abs(r.mean - mean) < 5 * std / np.sqrt(N)
    assert abs(r.std / std - 1.0) < 0.01
    q = O.simulate("sobol_gaussian", 1 << 17, P, seed=0)  # QMC: far tighter than MC at the same N
    assert abs(q.mean - mean) < 1.0 * std / np.sqrt(1 << 17)


def test_box_muller_pairs_are_jointly_normal(O):
    """One 32-bit word -> (z0, z1): radius from w[19:0], angle from the hashed word.  Marginals,
    joint moments, a 2-D chi-square on quantile cells and a polar-bin chi-square with a non-dyadic number of angle
    bins (which a raw 12-bit angle grid fails)."""
    from scipy import stats
    n = 8_000_000
    w = np.random.default_rng(3).integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    z = O.gauss_pairs(w, 0.0, 100.0).astype(np.float64) - 1.0
    z0, z1 = z[:, 0], z[:, 1]
    se = 1.0 / np.sqrt(n)
    assert abs(z0.mean()) < 5 * se and abs(z1.mean()) < 5 * se
    assert abs(z0.var() - 1) < 5 * np.sqrt(2) * se and abs(z1.var() - 1) < 5 * np.sqrt(2) * se
    assert abs((z0 * z1).mean()) < 5 * se
    assert abs((z0 ** 4).mean() - 3) < 5 * np.sqrt(96) * se and abs((z0 ** 2 * z1 ** 2).mean() - 1) < 5 * np.sqrt(8) * se
    assert 5.0 < np.abs(z).max() < 5.45  # radius field is 20 bits: |z| <= sqrt(2 ln 2^21) = 5.40
    assert stats.kstest(z0[:2_000_000], "norm").pvalue > 1e-3 and stats.kstest(z1[:2_000_000], "norm").pvalue > 1e-3
    k = 32
    edges = stats.norm.ppf(np.linspace(0, 1, k + 1)[1:-1])
    cells = np.bincount(np.searchsorted(edges, z0) * k + np.searchsorted(edges, z1), minlength=k * k)
    chi = ((cells - n / k ** 2) ** 2 / (n / k ** 2)).sum()
    assert stats.chi2.sf(chi, k * k - 1) > 1e-4
    ang = np.floor((np.arctan2(z1, z0) + np.pi) / (2 * np.pi) * 1000).astype(np.int64).clip(0, 999)
    rad = np.searchsorted(np.sqrt(-2 * np.log(1 - np.linspace(0, 1, 11)[1:-1])), np.hypot(z0, z1))
    cells = np.bincount(ang * 10 + rad, minlength=10_000)
    chi = ((cells - n / 1e4) ** 2 / (n / 1e4)).sum()
    assert stats.chi2.sf(chi, 9999) > 1e-4


def test_sobol_gaussian_pairs(O):
    """SOBOL_GAUSSIAN = Box-Muller on dimension pairs: month 2k takes the radius from the word of
    dimension 2k, month 2k+1 the angle from dimension 2k+1; an odd last month takes the inverse normal CDF.  Checked here:
    the pair transform against a float64 Box-Muller of the same words, normality on random words, the month <-> word
    wiring through orc_draws / trajectories, and the randomised-QMC accuracy against the closed-form moments."""
    from scipy import stats
    from scipy.special import ndtri
    rng = np.random.default_rng(11)
    n = 2_000_000
    xr = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    xa = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    edge = np.array([0, 1, 2 ** 32 - 1, 2 ** 31, 2 ** 31 - 1, 12345], dtype=np.uint32)  # u -> 2^-33 ... 1, v = -pi ... pi
    xr[:6], xa[:6] = edge, edge[::-1]
    z = O.sobol_gauss_pairs(xr, xa, 0.0, 100.0).astype(np.float64) - 1.0
    u = (xr.astype(np.float64) + 0.5) / 2.0 ** 32
    v = 2.0 * np.pi * xa.astype(np.int32).astype(np.float64) / 2.0 ** 32
    rad = np.sqrt(-2.0 * np.log(u))
    assert np.all(np.isfinite(z))
    err = np.maximum(np.abs(z[:, 0] - rad * np.cos(v)), np.abs(z[:, 1] - rad * np.sin(v)))
    # fp32 evaluation; u is a float, i.e. 24 bits: next to u = 1 (radius -> 0) 1 - u is only resolved to 2^-25, which moves
    # a radius of r by 3e-8 / r (at most 2.5e-4, for the 1e-7 of the draws with r < 5e-4)
    assert np.max(err[u < 0.99]) < 3e-6 and np.max(err) < 2.5e-4
    for k in (0, 1):
        assert stats.kstest(z[:, k], "norm").pvalue > 1e-3
    assert abs((z[:, 0] * z[:, 1]).mean()) < 5 / np.sqrt(n)
    # wiring: returns of a trajectory = 100 (g - 1) of the pair built from that path's own words, months (2k, 2k+1);
    # the odd last month is the inverse CDF of its word
    P, mean, std = 7, 0.5, 10.0 / 12.0
    x = O.draws("sobol_gaussian", 300, P, seed=5).reshape(300, P)
    _, rets = O.trajectories("sobol_gaussian", 300, P, seed=5, gauss_mean=mean, gauss_std=std, want_returns=True)
    g = O.sobol_gauss_pairs(x[:, 0:6:2].ravel(), x[:, 1:6:2].ravel(), mean, std).reshape(300, 3, 2)
    want = ((g.reshape(300, 6) - np.float32(1.0)) * np.float32(100.0)).astype(np.float32)
    assert np.array_equal(rets[:, :6], want)
    z_last = (rets[:, 6].astype(np.float64) - mean) / std
    assert np.max(np.abs(z_last - ndtri((x[:, 6].astype(np.float64) + 0.5) / 2.0 ** 32))) < 2e-5
    # randomised QMC (4 digital shifts) against the closed form: mean to ~1e-5, far below the MC standard error
    Pm, N = 360, 1 << 17
    mu, sg = float(np.float32(0.5) * np.float32(0.01)), float(np.float32(10.0 / 12.0) * np.float32(0.01))
    exact = 1000.0 * (1.0 + mu) ** Pm
    mc_se = np.sqrt(1000.0 ** 2 * ((1.0 + mu) ** 2 + sg ** 2) ** Pm - exact ** 2) / np.sqrt(N)
    errs = [O.simulate("sobol_gaussian", N, Pm, seed=s).mean - exact for s in (1, 2, 3, 4)]
    assert np.sqrt(np.mean(np.square(errs))) < 0.2 * mc_se, (errs, mc_se)


def test_bootstrap_index_is_the_exact_floor(O, table):
    """idx = floor(f n / 2^bits) for the six fields of a call; every index < n; near-uniform over the table"""
    d = O.draws("bootstrap", 2000, 600, returns=table, seed=9)
    assert d.max() < table.size
    counts = np.bincount(d.ravel(), minlength=table.size)
    from scipy import stats
    chi = ((counts - d.size / table.size) ** 2 / (d.size / table.size)).sum()
    assert stats.chi2.sf(chi, table.size - 1) > 1e-4
    # restate the contract in Python for one path
    lib = O.load()
    w = O.philox([[0, 5, 0x424F4F54, 0]], [9, 0])[0]
    f = [int(x) & 0x1FFFFF for x in w] + [((int(w[0]) >> 21) << 11) | (int(w[1]) >> 21), ((int(w[2]) >> 21) << 11) | (int(w[3]) >> 21)]
    want = [(f[j] * table.size) >> (21 if j < 4 else 22) for j in range(6)]
    got = O.draws("bootstrap", 6, 6, returns=table, seed=9)[5]
    assert list(got) == want


# ---- contract freeze ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ALL_MODES)
def test_oracle_regression_fixture(O, table, mode):
    g = golden("oracle_regression.npz")
    seeds = {"bootstrap_compat": 0, "bootstrap": 1234, "gaussian": 1234, "sobol_bootstrap": 0, "sobol_gaussian": 99}
    kw = mode_kwargs(mode, table, seed=seeds[mode])
    r = O.simulate(mode, 600, 37, hist_bins=16, hist_range=(500.0, 2500.0), want_finals=True,
                   want_block_partials=True, stream_offset=512, **kw)
    d = O.draws(mode, 8, 37, returns=kw.get("returns"), seed=kw["seed"], stream_offset=512)
    assert np.array_equal(d, g[mode + "_draws"])
    if "gaussian" in mode:  # libm in the loop: allow an ulp-level drift across libm builds
        np.testing.assert_allclose(r.finals, g[mode + "_finals"], rtol=1e-6)
    else:
        assert np.array_equal(r.finals, g[mode + "_finals"])
        assert np.array_equal(r.block_means, g[mode + "_bmeans"])
        assert np.array_equal(r.block_vars, g[mode + "_bvars"])
        assert np.array_equal(r.hist, g[mode + "_hist"])
    np.testing.assert_allclose([r.count, r.below_count, r.mean, r.m2], g[mode + "_scalars"], rtol=1e-6)


# ---- properties of the build-defined pieces --------------------------------------------------------------------------
def test_histogram_bin_rule(O):
    lib = O.load()
    lo, hi, bins = 0.0, 50000.0, 100
    inv = lib.orc_hist_inv_width(lo, hi, bins)
    assert lib.orc_hist_bin(-1.0, lo, inv, bins) == 0
    assert lib.orc_hist_bin(float("nan"), lo, inv, bins) == 0
    assert lib.orc_hist_bin(0.0, lo, inv, bins) == 1
    assert lib.orc_hist_bin(499.99, lo, inv, bins) == 1
    assert lib.orc_hist_bin(500.0, lo, inv, bins) == 2
    assert lib.orc_hist_bin(49999.9, lo, inv, bins) == bins
    assert lib.orc_hist_bin(50000.0, lo, inv, bins) == bins + 1
    assert lib.orc_hist_bin(float("inf"), lo, inv, bins) == bins + 1


@pytest.mark.parametrize("mode", ALL_MODES)
def test_oracle_sharding_identity(O, table, mode):
    """path id = RNG offset: two half-runs reproduce one full run exactly (finals, histogram, counts)."""
    kw = mode_kwargs(mode, table)
    full = O.simulate(mode, 1024 + 100, 25, hist_bins=32, want_finals=True, **kw)
    a = O.simulate(mode, 512, 25, hist_bins=32, want_finals=True, **kw)
    b = O.simulate(mode, 512 + 100, 25, hist_bins=32, want_finals=True, stream_offset=512, **kw)
    assert np.array_equal(np.concatenate([a.finals, b.finals]), full.finals)
    assert np.array_equal(a.hist + b.hist, full.hist)
    assert a.below_count + b.below_count == full.below_count


def test_oracle_empty_and_degenerate(O, table):
    r = O.simulate("bootstrap", 0, 360, returns=table, hist_bins=8, want_finals=True)
    assert r.count == 0 and r.mean == 0.0 and r.m2 == 0.0 and int(r.hist.sum()) == 0
    r = O.simulate("gaussian", 300, 0, hist_bins=8, want_finals=True)
    assert np.all(r.finals == 1000.0) and r.m2 == 0.0 and r.below_count == 0
    one = np.array([2.0], dtype=np.float32)  # single-entry table: deterministic growth
    r = O.simulate("bootstrap", 10, 3, returns=one, want_finals=True)
    v = np.float32(1000.0)
    for _ in range(3):
        v = np.float32(np.float64(v) + np.float64(v) * np.float64(np.float32(2.0) * np.float32(0.01)))
    assert np.all(r.finals == v)
