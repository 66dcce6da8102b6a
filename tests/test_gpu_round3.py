"""Round-3 GPU tests: ordering of the cached per-device workspace between streams, and the host-side merge that
smmc_run_multi / smmc_quantiles fall back to when NCCL cannot be used (driven by the `host_merge` test hook)."""
import numpy as np
import pytest

from conftest import ALL_MODES, mode_kwargs

pytestmark = pytest.mark.gpu


def test_async_run_then_reduce_mean_on_another_stream(S):
    """smmc_reduce_mean_device writes the same cached work buffer (partials, ticket) as the simulation kernel: a
    SMMC_FLAG_NO_SYNC run still in flight on stream A must not be disturbed by a reduction issued on stream B."""
    import torch
    N, P = 20_000_000, 120  # ~1 ms of kernel: still running when the reduction is enqueued
    ref = S.simulate("gaussian", N, P, seed=9, hist_bins=64)
    vec = torch.full((1 << 22,), 2.5, dtype=torch.float32, device="cuda")
    a, b = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    for _ in range(5):
        stats = torch.zeros(S.stats_len(64), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        S.simulate_device("gaussian", N, P, stats.data_ptr(), seed=9, hist_bins=64, stream=a.cuda_stream)
        got, _ = S.reduce_mean_device(vec.data_ptr(), vec.numel(), stream=b.cuda_stream)
        assert got == 2.5
        a.synchronize()
        r = S.finalize_stats(stats.cpu().numpy(), 64, 1000.0)
        assert (r.count, r.below_count, r.mean, r.m2) == (ref.count, ref.below_count, ref.mean, ref.m2)
        assert np.array_equal(r.hist, ref.hist)


def test_multi_gpu_host_merge_equals_device_merge(S, table):
    """smmc_run_multi / smmc_quantiles without NCCL: per-shard vectors added on the host, same results"""
    if S.load().smmc_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    try:
        for mode in ALL_MODES:
            kw = mode_kwargs(mode, table)
            S.set_variant()
            dev = S.simulate(mode, 600_001, 24, hist_bins=64, want_block_partials=True, n_gpus=2, **kw)
            qdev = S.quantiles(mode, 600_001, 24, n_gpus=2, **kw)
            S.set_variant("host_merge", 1)
            host = S.simulate(mode, 600_001, 24, hist_bins=64, want_block_partials=True, n_gpus=2, **kw)
            qhost = S.quantiles(mode, 600_001, 24, n_gpus=2, **kw)
            assert (host.count, host.below_count) == (dev.count, dev.below_count) and np.array_equal(host.hist, dev.hist)
            assert np.array_equal(host.block_means, dev.block_means) and np.array_equal(host.block_vars, dev.block_vars)
            assert abs(host.mean - dev.mean) <= 1e-12 * abs(dev.mean) and abs(host.m2 - dev.m2) <= 1e-9 * dev.m2
            assert np.array_equal(qhost, qdev)
            stats_only = S.simulate(mode, 600_001, 24, n_gpus=2, **kw)  # statistics-only: the branch that needed NCCL
            assert stats_only.count == 600_001 and abs(stats_only.mean - dev.mean) <= 1e-12 * abs(dev.mean)
    finally:
        S.set_variant()


def test_run_rank_issues_one_allreduce_per_call_on_every_path(S, table):
    """smmc_run_rank on a one-rank communicator (runs on a 1-GPU box): the plain path, the device-pointer path and the
    split-launch path (block partials of >= 16 Mi paths into host memory: two kernels) each issue exactly one ncclAllReduce,
    and the split path's statistics equal the monolithic launch's"""
    import torch
    comm = S.Comm(S.Comm.unique_id(), 1, 0, 0)
    try:
        assert comm.allreduce_count() == 0
        comm.simulate("bootstrap", 100_003, 12, returns=table, seed=3, hist_bins=16, want_finals=True)
        assert comm.allreduce_count() == 1
        stats = torch.zeros(S.stats_len(16), dtype=torch.float64, device="cuda")
        comm.simulate_device("gaussian", 300_000, 36, stats.data_ptr(), stream=torch.cuda.current_stream().cuda_stream, seed=5, hist_bins=16)
        torch.cuda.synchronize()
        assert comm.allreduce_count() == 2
        N, P = (16 << 20) + 148 * 32 * 256 * 2 + 1000, 4
        split = comm.simulate("gaussian", N, P, hist_bins=32, want_block_partials=True)
        assert split.gpu_launches >= 2 and comm.allreduce_count() == 3
        try:
            S.set_variant("finals_chunks", 1)  # one monolithic launch
            mono = S.simulate("gaussian", N, P, hist_bins=32, want_block_partials=True)
        finally:
            S.set_variant()
        assert np.array_equal(split.block_means, mono.block_means) and np.array_equal(split.block_vars, mono.block_vars)
        assert np.array_equal(split.hist, mono.hist) and (split.count, split.below_count) == (mono.count, mono.below_count)
        assert abs(split.mean - mono.mean) <= 1e-12 * abs(mono.mean) and abs(split.m2 - mono.m2) <= 1e-9 * abs(mono.m2)
    finally:
        comm.destroy()
