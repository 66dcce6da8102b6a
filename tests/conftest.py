"""Test configuration.

Markers: `gpu` = needs a B200 (sm_100a) and calls the product through the C ABI (libsmmc.so); everything else
runs on CPU: the oracle against the golden fixtures / the compiled reference, host logic, ABI export checks.

The oracle (oracle/) is the CHECKER: only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs use it.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 GPU; runs the CUDA path through the C ABI")


def _gpu_count():
    try:
        import stock_market_monte_carlo_b200 as S
        return int(S.load().smmc_device_count())
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    if _gpu_count() > 0:
        return
    skip = pytest.mark.skip(reason="no sm_100a device visible (the product has no CPU fallback)")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def S():
    import stock_market_monte_carlo_b200 as pkg
    pkg.load()
    return pkg


@pytest.fixture(scope="session")
def O():
    from oracle import binding
    binding.load()
    return binding


@pytest.fixture(scope="session")
def table(S):
    return S.synthetic_sp500_table(1200)


@pytest.fixture(scope="session")
def table1127(S):
    return S.synthetic_sp500_table(1127, seed=7)


def golden(name):
    path = os.path.join(GOLDEN, name)
    if not os.path.exists(path):
        pytest.skip(f"golden fixture {name} not generated yet")
    return np.load(path)


# keyword arguments that make each mode runnable (bootstrap modes need the table)
def mode_kwargs(mode, table, seed=1234):
    kw = {}
    if "bootstrap" in mode:
        kw["returns"] = table
    kw["seed"] = 0 if mode == "bootstrap_compat" else seed
    return kw


ALL_MODES = ["bootstrap_compat", "bootstrap", "gaussian", "sobol_bootstrap", "sobol_gaussian"]
EXACT_MODES = ["bootstrap_compat", "bootstrap", "sobol_bootstrap"]  # integer/table paths: bit-exact finals
