"""Generates the committed golden fixtures in tests/golden/ by RUNNING THE REFERENCE ITSELF in the build container.

    python tests/golden/make_golden.py            (needs oracle/_ref: `make -C oracle ref`, i.e. /root/reference)

Fixtures written (all small; the reference tree does not exist on the GPU box, these travel instead):
  sobol_i4_reference.npz   integers behind the reference's vendored Burkardt generator i4_sobol
                           (reference include/sobol/sobol.cpp:239-13896): points 0..2047 of dimensions
                           0..15, 359, 999, 1000, 1110, as round(quasi * 2^30) (exact: quasi = lastq * 2^-30)
  ref_cpu_quantiles.npz    the reference's CPU paths, which take no seed (src/simulations.cpp:45, :245), summarised
                           as 2001 empirical quantiles + moments of 400 000 finals:
                             boot_*   mc_simulations (CPU v2 OpenMP bootstrap, src/simulations.cpp:204-266), P=360
                             gauss_*  sample_returns_gaussian + many_updates (:41-55, :24-39), P=360, mu=0.5 %, sd=10/12 %
  ref_update_fund.npz      update_fund (src/simulations.cpp:14-16) on a grid of (fund, return) pairs
  oracle_regression.npz    the oracle's own outputs for every mode at N=600, P=37 (freezes the stream contract:
                           a change of the RNG/counter layout, field split, bin rule or Sobol skip shows up here)
  ref_write_data_file.csv  the bytes the reference's write_data_file (src/helpers.cpp:23-39) writes for a fixed
                           (returns, values) pair (tests/test_next_rows_cpu.py::_data_file_case)
The GPU-side fixture (reference .cu kernels run on a B200) is made by make_golden_gpu.py.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import stock_market_monte_carlo_b200 as S  # noqa: E402
from oracle import binding as O  # noqa: E402

SOBOL_DIMS = list(range(16)) + [359, 999, 1000, 1110]
SOBOL_POINTS = 2048
N_REF = 400_000
QS = np.linspace(0.0, 1.0, 2001)


def sobol_fixture():
    x = np.zeros((SOBOL_POINTS, len(SOBOL_DIMS)), dtype=np.uint32)
    for i in range(SOBOL_POINTS):
        q = O.ref_i4_sobol(1111, i)
        v = q[SOBOL_DIMS].astype(np.float64) * float(1 << 30)
        assert np.all(v == np.round(v))
        x[i] = v.astype(np.uint32)
    np.savez_compressed(os.path.join(HERE, "sobol_i4_reference.npz"), dims=np.array(SOBOL_DIMS, dtype=np.int32), x30=x)


def cpu_quantiles():
    tbl = S.synthetic_sp500_table(1200)
    boot, threads = O.ref_mc_simulations(N_REF, 360, 1000.0, tbl)
    gauss, _ = O.ref_gaussian_finals(N_REF, 360, 1000.0, 0.5, 10.0 / 12.0)
    np.savez_compressed(
        os.path.join(HERE, "ref_cpu_quantiles.npz"), qs=QS, n=np.int64(N_REF),
        boot_q=np.quantile(boot.astype(np.float64), QS), boot_mean=boot.astype(np.float64).mean(),
        boot_std=boot.astype(np.float64).std(),
        gauss_q=np.quantile(gauss.astype(np.float64), QS), gauss_mean=gauss.astype(np.float64).mean(),
        gauss_std=gauss.astype(np.float64).std())


def update_fund_fixture():
    rng = np.random.default_rng(5)
    fund = rng.uniform(1.0, 1e6, 512).astype(np.float32)
    ret = rng.normal(0.5, 5.0, 512).astype(np.float32)
    out = np.array([O.ref_update_fund(f, r) for f, r in zip(fund, ret)], dtype=np.float32)
    np.savez_compressed(os.path.join(HERE, "ref_update_fund.npz"), fund=fund, ret=ret, out=out)


def oracle_regression():
    tbl = S.synthetic_sp500_table(1200)
    out = {}
    for mode, kw in (("bootstrap_compat", dict(returns=tbl)), ("bootstrap", dict(returns=tbl, seed=1234)),
                     ("gaussian", dict(seed=1234)), ("sobol_bootstrap", dict(returns=tbl, seed=0)),
                     ("sobol_gaussian", dict(seed=99))):
        r = O.simulate(mode, 600, 37, hist_bins=16, hist_range=(500.0, 2500.0), want_finals=True,
                       want_block_partials=True, stream_offset=512, **kw)
        out[mode + "_finals"] = r.finals
        out[mode + "_bmeans"] = r.block_means
        out[mode + "_bvars"] = r.block_vars
        out[mode + "_hist"] = r.hist
        out[mode + "_scalars"] = np.array([r.count, r.below_count, r.mean, r.m2], dtype=np.float64)
        out[mode + "_draws"] = O.draws(mode, 8, 37, returns=kw.get("returns"), seed=kw.get("seed", 0), stream_offset=512)
    np.savez_compressed(os.path.join(HERE, "oracle_regression.npz"), **out)


def write_data_file_fixture():
    import tempfile
    sys.path.insert(0, os.path.dirname(HERE))
    from test_next_rows_cpu import _data_file_case
    r, v = _data_file_case()
    tmp = tempfile.mkdtemp()
    cwd = os.getcwd()
    os.chdir(tmp)  # the reference writes ./outputs/<fname>
    try:
        O.ref_write_data_file("ref.csv", r, v)
    finally:
        os.chdir(cwd)
    data = open(os.path.join(tmp, "outputs", "ref.csv"), "rb").read()
    open(os.path.join(HERE, "ref_write_data_file.csv"), "wb").write(data)


if __name__ == "__main__":
    for name in ("libref_cpu.so", "libref_sobol.so"):
        if not O.ref_available(name):
            raise SystemExit(f"oracle/_ref/{name} missing: run `make -C oracle ref` (needs /root/reference)")
    sobol_fixture()
    cpu_quantiles()
    update_fund_fixture()
    oracle_regression()
    write_data_file_fixture()
    print("golden fixtures written to", HERE)
