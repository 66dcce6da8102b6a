"""Same-box reference number: the reference's own kernels (src/simulations.cu unmodified, compiled for sm_100a into
oracle/_ref/libref_gpu.so) on config C3's shape, next to this build's BOOTSTRAP_COMPAT on the same inputs.
Checker-side timing driver (lives under tests/: only tests may call into oracle/) (run it under `ncu --metrics gpu__time_duration.sum` for kernel-only times).
usage: python tests/ref_gpu_time.py [N] [P]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import stock_market_monte_carlo_b200 as S  # noqa: E402
from oracle import binding as O  # noqa: E402  (checker-side tool: times the reference itself)

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
P = int(sys.argv[2]) if len(sys.argv) > 2 else 360
table = S.synthetic_sp500_table(1127)  # the reference's shared array holds 1127 entries (src/simulations.cu:123)
for name, fn in (("reference mc_simulations_gpu_reduceBlock (block partials)", lambda: O.ref_mc_simulations_gpu_reduceBlock(N, P, 1000.0, table)),
                 ("reference mc_simulations_gpu (all finals)", lambda: O.ref_mc_simulations_gpu(N, P, 1000.0, table)),
                 ("this build, BOOTSTRAP_COMPAT + block partials", lambda: S.simulate("bootstrap_compat", N, P, returns=table, want_block_partials=True)),
                 ("this build, BOOTSTRAP_COMPAT + all finals", lambda: S.simulate("bootstrap_compat", N, P, returns=table, want_finals=True))):
    fn()
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        r = fn()
        ts.append(time.perf_counter() - t0)
    t = min(ts)
    extra = f" device_ms {r.device_ms:.3f}" if hasattr(r, "device_ms") else ""
    print(f"{name:60s} wall {1e3 * t:9.2f} ms  {N * P / t:.4e} path-months/s end to end{extra}", flush=True)
