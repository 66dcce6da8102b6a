"""Development driver: kernel time of the 5 M x 1000 Sobol launches with a warm and with a flushed L2 (bench.py flushes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stock_market_monte_carlo_b200 as S

dev = torch.device("cuda", 0)
tbl = S.synthetic_sp500_table(1200)
tbl_dev = torch.from_numpy(tbl).to(dev)
stats = torch.zeros(S.stats_len(100), dtype=torch.float64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
stream = torch.cuda.current_stream(dev).cuda_stream
for mode, rng in (("sobol_bootstrap", (0.0, 1.0e8)), ("sobol_gaussian", (0.0, 500000.0)), ("gaussian", (0.0, 50000.0))):
    N, P = (5_000_000, 1000) if mode.startswith("sobol") else (5_000_000, 360)
    kw = dict(returns_ptr=tbl_dev.data_ptr(), n_returns=tbl.size) if "bootstrap" in mode else {}
    for cold in (False, True, False, True):
        ms = []
        for i in range(13):
            S.simulate_device(mode, N, P, stats.data_ptr(), seed=1234, stream_offset=(i * N // 256 * 256) % (1 << 30), hist_bins=100,
                              hist_range=rng, stream=stream, **kw)
            if cold:
                flush.zero_()
            torch.cuda.synchronize()
            ms.append(S.kernel_times(0, 1)[0])
        print(f"{mode:16s} {'flushed L2' if cold else 'warm L2   '} kernel ms mean {np.mean(ms[3:]):.4f} min {np.min(ms[3:]):.4f}", flush=True)
