"""Per-period withdrawals (ABI 2; the reference README's open "withdrawal strategies" item, README.md:130-133: fixed
amount, fixed percentage, varying percentage per period).  No reference implementation exists, so the pins are closed
forms (CPU, oracle) and oracle parity (GPU): value <- max(value * (1 - pct_m/100) - w, 0) after every period's return."""
import numpy as np
import pytest

from conftest import ALL_MODES, EXACT_MODES, mode_kwargs

CASES = {
    "fixed": dict(withdraw_fixed=3.5),
    "pct": dict(withdraw_pct=0.25),
    "schedule": None,  # filled per P
    "ruin": dict(withdraw_fixed=40.0, withdraw_pct=0.5),
}


def _case(name, P):
    if name == "schedule":
        return dict(withdraw_fixed=1.0, withdraw_pct_schedule=np.linspace(0.0, 0.6, P).astype(np.float32))
    return dict(CASES[name])


def test_oracle_withdrawal_closed_forms(O):
    N, P = 200_000, 120
    mu = 0.005
    g = 1 + mu
    r = O.simulate("gaussian", N, P, seed=5, withdraw_pct=0.25)
    want = 1000.0 * (g * (1 - 0.0025)) ** P                              # E[V] = V0 ((1+mu)(1-p))^P
    assert abs(r.mean - want) < 5 * r.std / np.sqrt(N) + 1e-4 * want
    w = 3.5
    r = O.simulate("gaussian", N, P, seed=5, withdraw_fixed=w, want_finals=True)
    want = 1000.0 * g ** P - w * (g ** P - 1) / (g - 1)                    # annuity formula (no path is ruined here)
    assert r.finals.min() > 0 and abs(r.mean - want) < 5 * r.std / np.sqrt(N) + 1e-4 * want
    sched = np.linspace(0.0, 0.6, P).astype(np.float32)
    r = O.simulate("gaussian", N, P, seed=5, withdraw_pct_schedule=sched)
    want = 1000.0 * np.prod(g * (1 - sched.astype(np.float64) / 100))
    assert abs(r.mean - want) < 5 * r.std / np.sqrt(N) + 1e-4 * want
    # ruin is absorbing: once at 0 a path stays at 0
    v = O.trajectories("gaussian", 50, P, seed=5, withdraw_fixed=40.0)
    assert np.all(v[:, -1] == 0.0) and np.all(np.diff((v == 0).astype(int), axis=1) >= 0)
    # no withdrawals requested = the ABI-1 results, bit for bit
    a = O.simulate("gaussian", 1000, P, seed=5, want_finals=True).finals
    b = O.simulate("gaussian", 1000, P, seed=5, want_finals=True, withdraw_fixed=0.0, withdraw_pct=0.0).finals
    assert np.array_equal(a, b)


def test_withdrawal_arguments(S):
    with pytest.raises(ValueError):
        S.make_config("gaussian", 10, 12, withdraw_pct_schedule=[0.1, 0.2])
    with pytest.raises(S.SmmcError) as e:
        S.simulate("gaussian", 10, 12, withdraw_fixed=float("nan"))
    assert e.value.status == 1
    base = S.plan_launch("gaussian", 10 ** 6, 360)[3]
    cfg, keep = S.make_config("gaussian", 10 ** 6, 360, withdraw_fixed=1.0)
    import ctypes as C
    from stock_market_monte_carlo_b200 import _native as N
    plan = N.SmmcPlan()
    N.check(S.load().smmc_plan_launch(C.byref(cfg), 0, 148, 232448, C.byref(plan)))
    assert base < plan.smem_bytes <= base + 360 * 4 + 128  # the schedule is staged in shared memory


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ALL_MODES)
@pytest.mark.parametrize("case", list(CASES))
def test_withdrawals_against_oracle(S, O, table, mode, case):
    from test_gpu_parity import GAUSS_RTOL
    kw = mode_kwargs(mode, table)
    N, P = 3000 + 37, 121
    kw.update(_case(case, P))
    g = S.simulate(mode, N, P, hist_bins=50, hist_range=(0.0, 5000.0), want_finals=True, want_block_partials=True, **kw)
    o = O.simulate(mode, N, P, hist_bins=50, hist_range=(0.0, 5000.0), want_finals=True, want_block_partials=True, **kw)
    if mode in EXACT_MODES:
        assert np.array_equal(g.finals, o.finals) and np.array_equal(g.hist, o.hist) and g.below_count == o.below_count
        assert np.array_equal(g.block_means, o.block_means) and np.array_equal(g.block_vars, o.block_vars)
    else:
        # a fixed withdrawal turns relative errors of the value into larger relative errors of a small remainder
        np.testing.assert_allclose(g.finals, o.finals, rtol=GAUSS_RTOL, atol=GAUSS_RTOL * 1000.0 * 3)
    if case == "ruin":
        assert (g.finals == 0).mean() > 0.9 and (o.finals == 0).mean() > 0.9
    # trajectories and order statistics run the same withdrawal variant
    t = S.trajectories(mode, 64, P, **kw)
    assert np.array_equal(t[:, -1].view(np.uint32), g.finals[:64].view(np.uint32))
    q = S.quantiles(mode, N, P, **kw)
    assert np.array_equal(q, np.sort(g.finals)[S.quartile_ranks(N)])


@pytest.mark.gpu
def test_withdrawals_do_not_touch_the_default_path(S, table):
    a = S.simulate("bootstrap", 200_000, 360, returns=table, seed=2, want_finals=True)
    b = S.simulate("bootstrap", 200_000, 360, returns=table, seed=2, want_finals=True, withdraw_fixed=0.0)
    assert np.array_equal(a.finals, b.finals)
    N = 20_000_000
    S.simulate("gaussian", 256 * 148 * 32, 360, seed=1, withdraw_pct=0.1)  # warm-up (module load)
    t0 = S.simulate("gaussian", N, 360, seed=1).device_ms
    t1 = S.simulate("gaussian", N, 360, seed=1, withdraw_pct=0.1, withdraw_fixed=1.0).device_ms
    assert t0 > 0.0 and t1 > 0.0  # the cost of the withdrawal variant is a bench figure, not a test condition
