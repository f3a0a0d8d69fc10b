"""Multi-GPU partitioning of the ORB front-end: frame pairs are independent, so rank r
of `world` takes one contiguous block of pairs and nothing crosses ranks on the data path.  torch.distributed
is used only around it: a barrier, the max-over-ranks of the device time, and the gather of per-pair results
to rank 0 (host side, small)."""


def shard_pairs(n_pairs, world, rank):
    """Contiguous block [lo, hi) of pair indices for `rank`; blocks differ in size by at most one pair."""
    assert world >= 1 and 0 <= rank < world and n_pairs >= 0
    base, extra = divmod(n_pairs, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_rows(n_queries, world, rank):
    """Query-row block [q0, q1) of ONE pair's Hamming matrix for `rank` (brute-force matching of BASELINE.json configs[4]:
    F2's descriptors go to every rank by H2D copy, each rank computes the top lists of its rows with orbx_match_rows, the
    row tables are gathered on the host and one rank runs orbx_match_resolve; no collective on the data path)."""
    return shard_pairs(n_queries, world, rank)


def shard_frames(n_pairs, world, rank):
    """Frame index range [2*lo, 2*hi) of the pairs of `rank` (pair k = frames 2k, 2k+1)."""
    lo, hi = shard_pairs(n_pairs, world, rank)
    return 2 * lo, 2 * hi


def max_over_ranks(value, device=None):
    """MAX all-reduce of a scalar (the timed region's device milliseconds); identity without a process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_to_rank0(obj):
    """Per-rank result lists -> list over ranks on rank 0 (None elsewhere).  Host objects, off the hot path."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [obj]
    out = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
    dist.gather_object(obj, out, dst=0)
    return out
