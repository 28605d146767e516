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


# ---------------------------------------------------------------------------------------------------------------
# Fold-range sharding: ONE long utterance over several GPUs (fewer utterances than GPUs).
# Folds never communicate (`fatchord_version.py:194-196`: every fold starts from zero state), so rank g runs a
# contiguous range of folds; the only exchange is the final host gather of the per-fold samples, and the float64
# cross-fade / unfold / fade-out of `generate` (`:243-258`, `:342-405`) is done once, where the shards meet.

def shard_folds(n_folds: int, world: int, rank: int) -> Tuple[int, int]:
    """[lo, hi) of the folds rank `rank` runs: contiguous, sizes differ by at most one, empty when world > n_folds."""
    idx = shard_utterances(n_folds, world, rank)
    return (idx[0], idx[-1] + 1) if idx else (0, 0)


def combine_fold_samples(samples, wave_len: int, target: int, overlap: int, hop_length: int, mulaw_classes=None):
    """Per-fold samples (n_folds, target + 2 * overlap) float32 -> float64 waveform: the host tail of `generate`.

    The same arithmetic, in the same order, as the reference's numpy code (`fatchord_version.py:243-258` and
    `xfade_and_unfold` `:374-405`), so a sharded run equals the single-GPU result (which tests pin bit for bit against it).
    `mulaw_classes`: RAW mode with mu_law: n_classes of `decode_mu_law(output, n_classes, False)` (`:247-248`,
    `utils/dsp.py:98-103`), applied with the reference's own numpy expression.
    """
    import numpy as np
    y = np.asarray(samples, dtype=np.float32).astype(np.float64)
    n_folds, length = y.shape
    if length != target + 2 * overlap:
        raise ValueError(f'samples have {length} steps, expected target + 2 * overlap = {target + 2 * overlap}')
    if mulaw_classes is not None:
        mu = mulaw_classes - 1
        y = np.sign(y) / mu * ((1 + mu) ** np.abs(y) - 1)
    # xfade_and_unfold
    silence_len = overlap // 2
    fade_len = overlap - silence_len
    silence = np.zeros((silence_len), dtype=np.float64)
    linear = np.ones((silence_len), dtype=np.float64)
    t = np.linspace(-1, 1, fade_len, dtype=np.float64)
    fade_in = np.concatenate([silence, np.sqrt(0.5 * (1 + t))])
    fade_out = np.concatenate([linear, np.sqrt(0.5 * (1 - t))])
    if overlap > 0:
        y[:, :overlap] *= fade_in
        y[:, -overlap:] *= fade_out
    unfolded = np.zeros(n_folds * (target + overlap) + overlap, dtype=np.float64)
    for i in range(n_folds):
        start = i * (target + overlap)
        unfolded[start:start + target + 2 * overlap] += y[i]
    # tail of generate: truncate, fade out the last 20 hops
    out = unfolded[:wave_len]
    fade = np.linspace(1, 0, 20 * hop_length)
    out[-20 * hop_length:] *= fade
    return out


def generate_sharded(vocoder, mel, target: int, overlap: int, mu_law: bool = False, noise=None, seed: int = 0, group=None,
                     stream=None, run_shard=None):
    """`generate(batched=True)` of one utterance with its folds sharded over the ranks of `group`.

    Every rank calls this with the same arguments; rank 0 returns the float64 waveform, the others None.  `vocoder` is
    this rank's `binding.Vocoder`.  The per-rank work is the single-GPU path on a fold range; the one exchange is
    `gather` of (folds, steps) float32 samples to rank 0.  `run_shard(lo, hi) -> samples` replaces the GPU call in tests.
    """
    import numpy as np
    import torch
    import torch.distributed as dist
    dims = vocoder.dims
    mel = np.ascontiguousarray(mel, dtype=np.float32)
    T = mel.shape[-1]
    n_folds = vocoder.num_folds(T * dims.hop_length, target, overlap)
    steps = target + 2 * overlap
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    world = dist.get_world_size(group) if multi else 1
    rank = dist.get_rank(group) if multi else 0
    lo, hi = shard_folds(n_folds, world, rank)
    if hi > lo:
        if run_shard is not None:
            mine = np.asarray(run_shard(lo, hi), dtype=np.float32)
        else:
            vocoder.set_fold_range(lo, hi)
            try:
                vocoder.stage([mel.reshape(dims.feat_dims, T)], True, target, overlap, mu_law, noise=noise, stream=stream)
                vocoder.run_staged(seed=seed, stream=stream)
                mine = vocoder.fetch_samples(stream=stream)
            finally:
                vocoder.set_fold_range(0, 0)
    else:
        mine = np.zeros((0, steps), dtype=np.float32)
    if mine.shape != (hi - lo, steps):
        raise RuntimeError(f'shard returned {mine.shape}, expected {(hi - lo, steps)}')
    if multi:
        # gather through a fixed-size buffer (gloo / NCCL want equal shapes): every rank pads to the largest shard
        per = -(-n_folds // world)
        buf = torch.zeros((per, steps), dtype=torch.float32)
        buf[:hi - lo] = torch.from_numpy(mine)
        dev = torch.device('cuda', torch.cuda.current_device()) if dist.get_backend(group) == 'nccl' else torch.device('cpu')
        buf = buf.to(dev)
        parts = [torch.empty_like(buf) for _ in range(world)] if rank == 0 else None
        dist.gather(buf, parts, dst=0, group=group)
        if rank != 0:
            return None
        rows = []
        for r in range(world):
            a, b = shard_folds(n_folds, world, r)
            rows.append(parts[r][:b - a].cpu().numpy())
        samples = np.concatenate(rows, axis=0)
    else:
        samples = mine
    classes = dims.n_classes if (dims.mode == 'RAW' and mu_law) else None
    return combine_fold_samples(samples, (T - 1) * dims.hop_length, target, overlap, dims.hop_length, classes)
