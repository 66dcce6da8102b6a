"""Development driver: where the host-buffer (e2e) step of C2 spends its time beyond the kernel."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stock_market_monte_carlo_b200 as S

N, P = 100_000_000, 360
for label, kw in (("stats only", dict()), ("stats + hist", dict(hist_bins=100, hist_range=(0.0, 50000.0))),
                  ("+ block partials (C2 e2e)", dict(hist_bins=100, hist_range=(0.0, 50000.0), want_block_partials=True))):
    for i in range(3):
        S.simulate("gaussian", N, P, seed=1234, stream_offset=i * N, **kw)
    t0 = time.perf_counter()
    K = 20
    kms = []
    for i in range(K):
        r = S.simulate("gaussian", N, P, seed=1234, stream_offset=(10 + i) * N, **kw)
        kms.append(r.device_ms)
    dt = (time.perf_counter() - t0) / K * 1e3
    print(f"{label:28s} wall {dt:7.3f} ms/step  kernel {np.mean(kms):7.3f} ms  overhead {dt - np.mean(kms):6.3f} ms", flush=True)
