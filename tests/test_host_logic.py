"""Host-side logic that needs no device: statistics-vector finalisation, shard planning, the reference-shaped
Python wrappers' argument handling."""
import numpy as np
import pytest


def test_stats_len_and_finalize_roundtrip(S):
    bins = 10
    L = S.stats_len(bins)
    assert L == 6 + bins + 2 and S.stats_len(0) == 8
    x = np.random.default_rng(0).lognormal(8.0, 0.5, 5000)
    v0 = 1000.0
    stats = np.zeros(L)
    stats[0] = x.size
    stats[1] = (x < v0).sum()
    stats[2] = (x - v0).sum()
    stats[3] = ((x - v0) ** 2).sum()
    stats[6:] = np.arange(bins + 2)
    r = S.finalize_stats(stats, bins, v0)
    assert r.count == x.size and r.below_count == int((x < v0).sum())
    assert abs(r.mean - x.mean()) < 1e-9 * x.mean()
    assert abs(r.m2 - ((x - x.mean()) ** 2).sum()) < 1e-9 * r.m2
    assert list(r.hist) == list(range(bins + 2))
    assert abs(r.std - x.std()) < 1e-9 * x.std() and r.p_below_target == r.below_count / r.count


def test_finalize_sums_of_shards_equal_whole(S):
    """the vector is sum-reducible: adding shard vectors then finalising == finalising the whole"""
    rng = np.random.default_rng(1)
    x = rng.lognormal(8.0, 0.5, 9000)
    v0 = 1000.0

    def vec(a):
        return np.array([a.size, (a < v0).sum(), (a - v0).sum(), ((a - v0) ** 2).sum(), 0, 0, 0, 0], dtype=np.float64)

    whole = S.finalize_stats(vec(x), 0, v0)
    parts = S.finalize_stats(vec(x[:2000]) + vec(x[2000:7000]) + vec(x[7000:]), 0, v0)
    assert whole.count == parts.count and whole.below_count == parts.below_count
    assert abs(whole.mean - parts.mean) < 1e-12 * whole.mean and abs(whole.m2 - parts.m2) < 1e-10 * whole.m2


def test_finalize_empty(S):
    r = S.finalize_stats(np.zeros(S.stats_len(4)), 4, 1000.0)
    assert r.count == 0 and r.mean == 0.0 and r.m2 == 0.0 and r.std == 0.0 and r.p_below_target == 0.0


@pytest.mark.parametrize("n", [0, 1, 255, 256, 257, 1000, 100_000_000, 10 ** 10])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_shard_ranges_partition_the_path_ids(S, n, world):
    from stock_market_monte_carlo_b200.distributed import shard_range
    nxt, total = 1024, 0
    sizes = []
    for r in range(world):
        off, cnt = shard_range(n, r, world, stream_offset=1024)
        assert off % 256 == 0 and cnt >= 0
        assert off == nxt or cnt == 0
        nxt = off + ((cnt + 255) // 256) * 256 if cnt else nxt
        total += cnt
        sizes.append(cnt)
    assert total == n
    assert max(sizes) - min(sizes) <= 256 + 255  # balanced to one logical block


@pytest.mark.parametrize("n", [0, 1, 255, 256, 257, 1000, 1_000_003, 100_000_000, 10 ** 10])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_rank_range_of_the_c_abi_partitions_the_run(S, n, world):
    """smmc_rank_range (the partition smmc_run_rank / smmc_run_multi use): whole 256-path blocks, contiguous, complete,
    balanced to one block, empty trailing ranks are zero-sized -- and it is the partition the Python plumbing uses."""
    from stock_market_monte_carlo_b200.distributed import shard_range
    nxt, sizes = 0, []
    for r in range(world):
        first, cnt = S.rank_range(n, world, r)
        assert cnt >= 0 and first % 256 == 0
        if cnt:
            assert first == nxt
            nxt = first + cnt
            if r + 1 < world and S.rank_range(n, world, r + 1)[1]:
                assert cnt % 256 == 0
        sizes.append(cnt)
        off, c2 = shard_range(n, r, world, stream_offset=0)
        assert c2 == cnt and (cnt == 0 or off == first)
    assert sum(sizes) == n and nxt == n
    assert max(sizes) - min(sizes) <= 256 + 255


def test_rank_range_rejects_bad_requests(S):
    for args in ((10, 2, 2), (10, 2, -1), (-5, 1, 0), (10, 0, 0)):
        with pytest.raises(S.SmmcError):
            S.rank_range(*args)


def test_shard_range_rejects_bad_requests(S):
    from stock_market_monte_carlo_b200.distributed import shard_range
    for args in ((10, 2, 2), (10, -1, 2), (-5, 0, 1), (10, 0, 0)):
        with pytest.raises(ValueError):
            shard_range(*args)


def test_reference_shaped_wrappers_argument_behaviour(S):
    """mc_simulations_gpu_reduceBlock throws for n_gpus != 1 like the reference (src/simulations.cu:690-694)"""
    n = S.AtomicLong(0)
    means, variances = S.FloatVector(), S.FloatVector()
    with pytest.raises(ValueError, match="not yet supported"):
        S.mc_simulations_gpu_reduceBlock(n, 1000, 12, 1000.0, S.FloatVector([1.0, 2.0]), means, variances, 2)
    assert n.load() == 0
    v = S.FloatVector([1, 2, 3])
    v.resize(5)
    assert v.size() == 5 and list(v.array) == [1, 2, 3, 0, 0]
    v.resize(2)
    assert list(v.array) == [1, 2]
    out = np.zeros(4, dtype=np.float32)
    S.vector_add(out, np.arange(4, dtype=np.float32), np.ones(4, dtype=np.float32), 4)  # host loop twin (gpu.cpp:7-15)
    assert list(out) == [1, 2, 3, 4]


def test_synthetic_table_is_deterministic_and_sp500_shaped(S):
    t = S.synthetic_sp500_table(1200)
    assert t.dtype == np.float32 and t.size == 1200 and np.array_equal(t, S.synthetic_sp500_table(1200))
    assert 0.2 < t.mean() < 1.2 and 3.0 < t.std() < 7.0 and t.min() >= -30 and t.max() <= 40


def test_sobol_table_embedded_in_library_matches_data_file(S):
    from stock_market_monte_carlo_b200.data import sobol_directions
    v = sobol_directions()
    assert v.shape == (1111, 32)
    assert np.all(v[0] == (1 << 31) >> np.arange(32))            # dimension 0 is the van der Corput sequence
    assert np.all(v[:, 0] == 1 << 31)                            # every first direction number is 1/2
    # column k has k+1 significant bits: V[d][k] = m << (31-k) with m odd
    for k in range(32):
        assert np.all((v[:, k] >> (31 - k)) & 1 == 1) and not np.any(v[:, k] & ((1 << (31 - k)) - 1))
