"""The N>1 host path on CPU: world_size 2 (and 3) over gloo.  Each rank takes `shard_range` of the global path-id
range, produces the allreduce-ready statistics vector of include/smmc.h for its shard, and ONE all_reduce(SUM)
merges them (stock_market_monte_carlo_b200/distributed.py).  On the GPU box the per-shard vector comes from the
CUDA kernel; here the oracle stands in for the device so that the partition / collective / finalize logic is
covered without a GPU (tests may use the oracle; the product never does)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _stats_vector(finals, hist, v0, bins):
    x = finals.astype(np.float64)
    head = [x.size, float((finals < np.float32(v0)).sum()), (x - v0).sum(), ((x - v0) ** 2).sum(), 0.0, 0.0]
    return np.concatenate([np.array(head), hist.astype(np.float64)])


def _worker(rank, world, port, mode, n_total, n_periods, bins, out_path):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import stock_market_monte_carlo_b200 as S
    from oracle import binding as O
    from stock_market_monte_carlo_b200.distributed import allreduce_stats, shard_range

    tbl = S.synthetic_sp500_table(1200)
    kw = dict(seed=11)
    if "bootstrap" in mode:
        kw["returns"] = tbl
    off, cnt = shard_range(n_total, rank, world)
    r = O.simulate(mode, cnt, n_periods, hist_bins=bins, want_finals=True, stream_offset=off, **kw)
    stats = torch.from_numpy(_stats_vector(r.finals, r.hist, 1000.0, bins))
    assert stats.numel() == S.stats_len(bins)
    allreduce_stats(stats)
    if rank == 0:
        np.save(out_path, stats.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("mode", ["gaussian", "bootstrap"])
def test_sharded_allreduce_equals_single_run(S, O, table, tmp_path, world, mode):
    n_total, n_periods, bins = 2000 + 77, 24, 20
    out = str(tmp_path / "stats.npy")
    mp.spawn(_worker, args=(world, _free_port(), mode, n_total, n_periods, bins, out), nprocs=world, join=True)
    merged = S.finalize_stats(np.load(out), bins, 1000.0)
    kw = dict(seed=11)
    if "bootstrap" in mode:
        kw["returns"] = table
    whole = O.simulate(mode, n_total, n_periods, hist_bins=bins, want_finals=True, **kw)
    assert merged.count == whole.count == n_total
    assert merged.below_count == whole.below_count
    assert np.array_equal(merged.hist, whole.hist)          # integer partials: bit-identical for every world size
    # same vector built from the unsharded run: the fp64 sums only differ by summation order
    single = S.finalize_stats(_stats_vector(whole.finals, whole.hist, 1000.0, bins), bins, 1000.0)
    assert abs(merged.mean - single.mean) <= 1e-12 * abs(single.mean)
    assert abs(merged.m2 - single.m2) <= 1e-9 * abs(single.m2)
    # the oracle's own moments come from fp32 per-block tree sums merged in fp64 (reference reduceBlock layout)
    assert abs(merged.mean - whole.mean) <= 1e-6 * abs(whole.mean)
    assert abs(merged.m2 - whole.m2) <= 1e-5 * abs(whole.m2)
