"""Synthetic inputs (no network: the reference's S&P 500 CSV, python/get_data.py, needs Yahoo Finance) and the
Joe-Kuo direction numbers extracted from the reference's vendored Sobol source (tools/extract_sobol_directions.py)."""
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SOBOL_PATH = os.path.join(_HERE, "sobol_joe_kuo_1111x32.bin")


def synthetic_sp500_table(n=1200, seed=20221101):
    """`n` S&P-500-shaped monthly returns in PERCENT (the reference's units, python/get_data.py:59):
    r = clip(0.60 + 4.3 * t4 / sqrt(2), -30, +40) with t4 ~ Student-t(4)  (SURVEY.md 8(d))."""
    rng = np.random.default_rng(seed)
    t4 = rng.standard_t(4, size=n)
    return np.clip(0.60 + 4.3 * t4 / np.sqrt(2.0), -30.0, 40.0).astype(np.float32)


def sobol_directions():
    """uint32 [1111][32]: V[d][k] = m_{d,k} << (31 - k)."""
    return np.fromfile(SOBOL_PATH, dtype="<u4").reshape(1111, 32)
