"""Generates tests/golden/ref_gpu_compat.npz by running the REFERENCE'S OWN CUDA kernels (reference
src/simulations.cu compiled unmodified for sm_100a into oracle/_ref/libref_gpu.so) on a B200:

    gpurun -- 'python tests/golden/make_golden_gpu.py gpurun_out/ref_gpu_compat.npz'     then copy into tests/golden/

Contents: mc_simulations_gpu totals (N=3000, P=360) and mc_simulations_gpu_reduceBlock means / variances (N=3000:
11 full blocks + one partial block of 184, which exercises the divisor-256 quirk, src/simulations.cu:232,:241-242)
on the synthetic 1127-row table (the reference kernels hold the table in `__shared__ float[1127]`, :123, :193),
plus a P=1 run (one draw per path: pins the seed hash + first xorshift + index map per path id).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import stock_market_monte_carlo_b200 as S  # noqa: E402
from oracle import binding as O  # noqa: E402

if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "ref_gpu_compat.npz")
    tbl = S.synthetic_sp500_table(1127, seed=7)
    N = 3000
    totals = O.ref_mc_simulations_gpu(N, 360, 1000.0, tbl)
    totals_p1 = O.ref_mc_simulations_gpu(N, 1, 1000.0, tbl)
    means, variances = O.ref_mc_simulations_gpu_reduceBlock(N, 360, 1000.0, tbl)
    np.savez_compressed(out, table_n=np.int32(1127), table_seed=np.int32(7), n=np.int32(N), totals=totals,
                        totals_p1=totals_p1, means=means, variances=variances)
    print("wrote", out, totals[:4], means[:2], variances[:2])
