"""Times the reference's own benchmark_mc_gpu (unmodified source, relinked against libsmmc.so by `make -C oracle relink`)
at a given size: the wall clock the example itself prints around the drop-in mc_simulations_gpu call.
usage: python tests/run_relinked_example.py [N] [P]      (GPU box; checker-side driver, lives with the tests)"""
import os
import pathlib
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import stock_market_monte_carlo_b200 as S  # noqa: E402
from test_gpu_parity import _run_relinked  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
P = int(sys.argv[2]) if len(sys.argv) > 2 else 360
exe = os.path.join(ROOT, "oracle", "_ref", "benchmark_mc_gpu")
with tempfile.TemporaryDirectory() as d:
    for _ in range(3):
        out = _run_relinked(pathlib.Path(d), exe, [1, P, N], S.synthetic_sp500_table(1127))
        print([l for l in out.splitlines() if "simulation done" in l or "mean:" in l], flush=True)
