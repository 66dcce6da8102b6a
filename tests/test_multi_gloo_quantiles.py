"""The N>1 order-statistic protocol on CPU: world_size 2 and 3 over gloo.  Each rank builds the per-pass digit
histograms of ITS shard (here from the oracle's finals, standing in for the device replay), ONE all_reduce(SUM,
int64) per pass merges them, and every rank advances the same select state -- the host logic of
distributed.quantiles_sharded without a GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from test_multi_gloo import ROOT, _free_port


def _worker(rank, world, port, n_total, n_periods, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import stock_market_monte_carlo_b200 as S
    from oracle import binding as O
    from stock_market_monte_carlo_b200.distributed import shard_range
    from test_next_rows_cpu import emulate_pass, ordered_keys

    off, cnt = shard_range(n_total, rank, world)
    finals = O.simulate("gaussian", cnt, n_periods, seed=11, stream_offset=off, want_finals=True).finals
    keys = ordered_keys(finals)
    st = S.SelectState(S.quartile_ranks(n_total))
    while not st.done:
        h = torch.from_numpy(emulate_pass(keys, st).astype(np.int64).reshape(-1))
        dist.all_reduce(h, op=dist.ReduceOp.SUM)
        st.advance(h.numpy().astype(np.uint64))
    np.save(out_path + f".{rank}.npy", st.values())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_select_equals_sorting_the_whole_run(S, O, tmp_path, world):
    n_total, n_periods = 5000 + 77, 24
    out = str(tmp_path / "q")
    mp.spawn(_worker, args=(world, _free_port(), n_total, n_periods, out), nprocs=world, join=True)
    whole = O.simulate("gaussian", n_total, n_periods, seed=11, want_finals=True).finals
    want = O.quantiles_of(whole, O.quartile_ranks(n_total))
    for r in range(world):
        assert np.array_equal(np.load(out + f".{r}.npy"), want)
