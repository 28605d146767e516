"""Multi-GPU host logic: the path shards by utterance, one process per GPU, no data-path collective.

generate() of the reference (models/fatchord_version.py:159-266) handles one utterance per call; a serving job
with many utterances gives each rank a disjoint subset and every rank runs the single-GPU path unchanged.
The only cross-rank traffic is the timing reduction of the benchmark (max of the per-rank device times, sum of
the samples produced), which goes through torch.distributed (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def shard_utterances(n_utterances: int, world: int, rank: int) -> List[int]:
    """Indices of the utterances rank `rank` generates: contiguous blocks, sizes differ by at most one."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f'bad rank {rank} for world size {world}')
    if n_utterances < 0:
        raise ValueError('n_utterances < 0')
    base, extra = divmod(n_utterances, world)
    lo = rank * base + min(rank, extra)
    return list(range(lo, lo + base + (1 if rank < extra else 0)))


def whole_job_rate(samples_local: int, ms_local: Sequence[float], group=None, device='cpu') -> Tuple[float, List[float]]:
    """(samples per second of the whole job, per-leg max-over-ranks milliseconds).

    samples_local: samples this rank produced in every timed leg; ms_local: this rank's time of each leg
    (device time by CUDA events for the resident-input leg, wall time for the end-to-end leg).
    The job is as slow as its slowest rank: times reduce with MAX, samples with SUM.
    The rate is computed on the first leg.
    """
    import torch
    import torch.distributed as dist
    ms = torch.tensor(list(ms_local), dtype=torch.float64, device=device)
    n = torch.tensor([float(samples_local)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(n, op=dist.ReduceOp.SUM, group=group)
    ms_max = ms.tolist()
    return float(n.item()) / (ms_max[0] / 1000.0), ms_max
