"""Sharding of independent planning queries over ranks (one process per GPU) -- SURVEY 8(e).

The path has no exchange step: every query owns its patch, grids and heuristic, so there is no
data-path collective.  torch.distributed is used only to agree on timings and counters
(barrier, max / sum all-reduce): NCCL on the GPU box, gloo in the CPU tests."""
import numpy as np

QUERIES_PER_MAP = 64  # config 5 draws 64 start/goal pairs per synthetic map


def map_seed(query_id, base=50000):
    return base + query_id // QUERIES_PER_MAP


def shard_queries(n_queries, world_size, rank):
    """Contiguous blocks of whole maps per rank (a map upload is reused by its 64 queries); the
    first `n_maps % world_size` ranks take one more map.  Returns the query ids of `rank`."""
    if not 0 <= rank < world_size:
        raise ValueError("rank outside world")
    n_maps = (n_queries + QUERIES_PER_MAP - 1) // QUERIES_PER_MAP
    per, extra = divmod(n_maps, world_size)
    m0 = rank * per + min(rank, extra)
    m1 = m0 + per + (1 if rank < extra else 0)
    return np.arange(min(m0 * QUERIES_PER_MAP, n_queries), min(m1 * QUERIES_PER_MAP, n_queries))


def reduce_timing(dist, device, local_ms, local_units):
    """max over ranks of the device time, sum over ranks of the units processed (whole-job
    throughput = units / max time).  `dist` = torch.distributed or None for a single process."""
    import torch
    t = torch.tensor([float(local_ms), float(local_units)], dtype=torch.float64, device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        tmax, tsum = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        return float(tmax[0]), float(tsum[1])
    return float(t[0]), float(t[1])
