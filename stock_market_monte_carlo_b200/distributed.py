"""Multi-GPU host logic: one process per GPU, paths sharded by RNG offset, ONE sum-allreduce of the statistics vector.

The reference's multi-GPU launcher re-runs the same paths on every device and gathers all finals over PCIe
(reference src/simulations.cu:576-655, :621-622).  Here rank r simulates the contiguous, 256-aligned path-id
range `shard_range(N, r, G)`; the path id IS the RNG stream offset, so the union over ranks is exactly the
1-GPU run.  Each rank's kernel emits the allreduce-ready statistics vector of include/smmc.h
(count, below, S1, S2, histogram -- all sum-reducible, integer fields exact in fp64), and a single
all-reduce merges them: on GPUs `ncclAllReduce` enqueued by `smmc_run_rank` inside libsmmc.so (simulate_sharded and
quantiles_sharded go through `simulations.Comm`), in the gloo CPU tests of this host logic `torch.distributed.all_reduce`.

Nothing here touches per-path data, and nothing here computes paths: the simulation itself is `simulate_device`
(CUDA only).
"""
from . import _native as N


def shard_range(n_simulations, rank, world_size, stream_offset=0):
    """(first_path_id, n_paths) of `rank`: contiguous ranges in units of whole 256-path logical blocks, the
    remainder blocks going to the lowest ranks -- smmc_rank_range of the C ABI (the partition of smmc_run_multi and
    smmc_run_rank), plus the run's stream_offset."""
    n = int(n_simulations)
    if n < 0 or world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad shard request")
    from . import simulations as _S
    first, count = _S.rank_range(n, world_size, rank)
    return int(stream_offset) + first, count


_comms = {}  # (device, id(group)) -> simulations.Comm: the communicators behind simulate_sharded / quantiles_sharded on GPUs


def _uses_nccl(group=None):
    """True when this is a multi-rank job on GPUs: the collective then runs inside libsmmc.so (smmc_run_rank /
    smmc_quantiles_rank), never through torch.distributed -- the package has ONE GPU data plane."""
    import torch.distributed as dist
    return (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
            and "nccl" in str(dist.get_backend(group)).lower())


def _comm_for(device, group=None):
    key = (int(device), id(group))
    if key not in _comms:
        _comms[key] = init_comm(device, group)  # collective: every rank gets here on its first sharded call
    return _comms[key]


def allreduce_stats(stats, group=None):
    """In-place SUM all-reduce of a statistics vector through torch.distributed: the stand-in for the NCCL call of
    smmc_run_rank in the gloo CPU tests of the host logic (CPU float64 tensor).  GPU jobs do not come through here."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def _first(n_simulations, rank, world):
    from . import simulations as _S
    return _S.rank_range(int(n_simulations), world, rank)[0]


def simulate_sharded(mode, n_simulations, n_periods, *, device, stream=None, group=None, hist_bins=0, **kw):
    """Rank-local shard of a global run of `n_simulations` paths + the allreduce.  Must be called by every rank
    of `group`.  Returns (Result of the GLOBAL run, launches on this rank).  Needs torch + a CUDA device."""
    import torch
    import torch.distributed as dist

    from .simulations import finalize_stats, simulate_device, stats_len

    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    off, n_local = shard_range(n_simulations, rank, world, kw.pop("stream_offset", 0))
    dev = torch.device("cuda", device)
    stats = torch.empty(stats_len(hist_bins), dtype=torch.float64, device=dev)
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    with torch.cuda.stream(st):
        if _uses_nccl(group):  # kernel -> ncclAllReduce on the same stream, both enqueued by smmc_run_rank
            launches = _comm_for(device, group).simulate_device(mode, n_simulations, n_periods, stats.data_ptr(), stream_offset=off - _first(n_simulations, rank, world),
                                                                hist_bins=hist_bins, stream=st.cuda_stream, **kw)
        else:
            launches = simulate_device(mode, n_local, n_periods, stats.data_ptr(), stream_offset=off,
                                       hist_bins=hist_bins, device=device, stream=st.cuda_stream, **kw)
            allreduce_stats(stats, group)  # one rank, or a non-NCCL group
        host = stats.cpu()
    return finalize_stats(host.numpy(), hist_bins, kw.get("initial_capital", 1000.0)), launches


def quantiles_sharded(mode, n_simulations, n_periods, ranks=None, *, device, stream=None, group=None, **kw):
    """Exact order statistics of a GLOBAL run of `n_simulations` paths sharded over the ranks of `group`:
    every rank replays its own shard three times (smmc_select_pass, device pointers, asynchronous), and the
    per-pass digit histograms -- exact integer counters -- are summed with ONE all_reduce(SUM, int64) per pass
    (16 Ki counters = 128 KB; NCCL over NVLink).  Must be called by every rank.  Returns the values (float32)."""
    import torch
    import torch.distributed as dist

    from .simulations import SelectState, make_config, quartile_ranks

    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if _uses_nccl(group):  # smmc_quantiles_rank: replays + ncclAllReduce(ncclUint64) of the digit histograms inside the C ABI
        return _comm_for(device, group).quantiles(mode, n_simulations, n_periods, ranks=ranks, **kw)
    off, n_local = shard_range(n_simulations, rank, world, kw.pop("stream_offset", 0))
    dev = torch.device("cuda", device)
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    returns = kw.pop("returns", None)
    table_dev = None
    if returns is not None:
        import numpy as np
        table_dev = torch.from_numpy(np.ascontiguousarray(np.asarray(returns, dtype=np.float32))).to(dev)
    cfg, _ = make_config(mode, n_local, n_periods, stream_offset=off, device=device,
                         flags=N.FLAG_DEVICE_POINTERS, stream=st.cuda_stream,
                         returns_ptr=table_dev.data_ptr() if table_dev is not None else None,
                         n_returns=int(table_dev.numel()) if table_dev is not None else None, **kw)
    sel = SelectState(quartile_ranks(n_simulations) if ranks is None else ranks)
    hist = torch.zeros(N.SELECT_MAX_TARGETS * N.SELECT_BINS, dtype=torch.int64, device=dev)
    with torch.cuda.stream(st):
        while not sel.done:
            sel.run_pass(cfg, hist_ptr=hist.data_ptr())
            if world > 1:
                dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
            sel.advance(hist.cpu().numpy().astype("uint64"))
    return sel.values()


def init_comm(device, group=None):
    """A `simulations.Comm` for this rank of the torch.distributed job: rank 0 asks NCCL (inside libsmmc.so) for a unique id,
    torch.distributed carries the 128 bytes to the other ranks, every rank calls ncclCommInitRank through the C ABI.
    From then on the collective of the path is `smmc_run_rank` -- torch is only the launcher."""
    import torch.distributed as dist

    from .simulations import Comm

    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    box = [Comm.unique_id() if rank == 0 else None]
    if world > 1:
        dist.broadcast_object_list(box, src=0, group=group)
    return Comm(box[0], world, rank, device)
