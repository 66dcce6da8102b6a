import time, ctypes as C, torch
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stock_market_monte_carlo_b200 as S
lib = S.load()
big = torch.full((1 << 28,), 1.5, dtype=torch.float32, device="cuda")
for name, f in (("noalloc", lambda: S.reduce_mean_device_noalloc(big.data_ptr(), big.numel())), ("alloc", lambda: S.reduce_mean_device(big.data_ptr(), big.numel())[0])):
    for _ in range(3): f()
    t = time.perf_counter()
    for _ in range(20): r = f()
    dt = (time.perf_counter() - t) / 20
    print(name, r, "%.3f ms/call wall  %.0f GB/s" % (dt * 1e3, big.numel() * 4 / dt / 1e9))
