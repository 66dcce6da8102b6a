"""Multi-rank correctness check on real GPUs (run under torchrun, one rank per GPU, NCCL):
  * simulate_sharded == the 1-GPU run of the same global config (histogram / counts bit-identical, moments 1e-12)
  * quantiles_sharded == smmc_quantiles on one GPU (exact)
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/multi_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stock_market_monte_carlo_b200 as S  # noqa: E402
from stock_market_monte_carlo_b200.distributed import quantiles_sharded, simulate_sharded  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
os.environ.setdefault("NCCL_DEBUG", "WARN")
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
tbl = S.synthetic_sp500_table(1200)
N, P = 3_000_000 + 77, 60
for mode, kw in (("gaussian", dict(seed=5)), ("bootstrap", dict(seed=5, returns=tbl))):
    dkw = dict(kw)
    tdev = None
    if "returns" in dkw:
        tdev = torch.from_numpy(dkw.pop("returns")).cuda()
        dkw.update(returns_ptr=tdev.data_ptr(), n_returns=tbl.size)
    r, _ = simulate_sharded(mode, N, P, device=local, hist_bins=50, hist_range=(0.0, 5000.0), **dkw)
    q = quantiles_sharded(mode, N, P, device=local, **kw)
    if rank == 0:
        one = S.simulate(mode, N, P, hist_bins=50, hist_range=(0.0, 5000.0), device=0, **kw)
        q1 = S.quantiles(mode, N, P, device=0, **kw)
        assert r.count == one.count == N and r.below_count == one.below_count and np.array_equal(r.hist, one.hist), mode
        assert abs(r.mean - one.mean) <= 1e-12 * abs(one.mean) and abs(r.m2 - one.m2) <= 1e-9 * abs(one.m2), mode
        assert np.array_equal(q, q1), (mode, q, q1)
        print(f"multi_check {mode} world={world}: sharded == single (hist, counts, quartiles {q}) OK", flush=True)
    dist.barrier()
dist.destroy_process_group()
if rank == 0:  # the one-host-thread driver (smmc_run_multi / smmc_quantiles with n_gpus) after NCCL has let go of the devices
    for mode, kw in (("gaussian", dict(seed=5)), ("bootstrap", dict(seed=5, returns=tbl))):
        one = S.simulate(mode, N, P, hist_bins=50, hist_range=(0.0, 5000.0), **kw)
        m = S.simulate(mode, N, P, hist_bins=50, hist_range=(0.0, 5000.0), n_gpus=world, **kw)
        assert np.array_equal(m.hist, one.hist) and m.below_count == one.below_count
        assert np.array_equal(S.quantiles(mode, N, P, n_gpus=world, **kw), S.quantiles(mode, N, P, **kw))
        print(f"multi_check {mode} n_gpus={world} (one host thread): smmc_run_multi / smmc_quantiles == single OK", flush=True)
