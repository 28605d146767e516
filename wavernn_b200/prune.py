"""Pruning tool: the reference's gradual magnitude-pruning schedule, producing the 16x1 block masks of SURVEY.md row S.

The reference's only pruning code is a scratch notebook (`notebooks/Pruning - Scratchpad.ipynb`): `Pruner`
(:110-160) raises the sparsity of the GRU / Linear layers along a cubic schedule during training
(`update_sparsity`, :127-131: z = Z * (1 - (1 - (t - t0) / S) ** 3), clamped to [0, Z]), recomputes the masks every
`prune_every` steps after `t0` (`prune_or_not`, :142-143) and multiplies the weights by the masks on every step from
`t0` on (`apply_or_not`, :146-147); `PruneMask.mask_from_matrix` (:92-108) prunes ELEMENTWISE by magnitude, separately
per gate split.  This module keeps that driver (same names, same arguments, same schedule) and swaps the mask for the
block structure the sparse loop kernel runs on (csrc/loop_res.cu `sparse_pass`; C entry `wrnn_prune_16x1`): per GRU
matrix and per gate, blocks of 16 consecutive output rows x 1 input column, kept by L1 norm.  The elementwise mask of
the notebook is available too (`elementwise_mask`), for comparison and for the Linear layers.

Host-side training utility, numpy in and out (torch tensors are accepted and updated in place): it is not on the
generation path.  A model pruned with it loads through `load_weights` / `WaveRNN.load`, which detect the block
structure and pick the sparse kernel (tests/test_variants_gpu.py).
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, Optional

import numpy as np

GRU_KEYS = ('rnn1.weight_ih_l0', 'rnn1.weight_hh_l0', 'rnn2.weight_ih_l0', 'rnn2.weight_hh_l0')


def clamp(x, lo=0.0, hi=1.0):
    """`clamp` of the notebook (cell 2)."""
    return max(lo, min(hi, x))


def sparsity_at(t, start_prune, prune_steps, target_sparsity):
    """Sparsity the schedule asks for at training step t (`Pruner.update_sparsity`, notebook :127-131)."""
    z = target_sparsity * (1 - (1 - (t - start_prune) / prune_steps) ** 3)
    return clamp(z, 0, target_sparsity)


def block_mask_16x1(w: np.ndarray, n_gates: int, sparsity: float) -> np.ndarray:
    """Mask (same shape as w, float32 0/1) of SURVEY.md row S at the given sparsity: per gate (row split), blocks of 16
    consecutive output rows x 1 input column, the round((1 - sparsity) * n_blocks) blocks of largest L1 norm are kept
    (at least one; ties go to the lower block index).  Equals what `wrnn_prune_16x1(keep_fraction = 1 - sparsity)` keeps."""
    rows, K = w.shape
    if rows % n_gates or (rows // n_gates) % 16:
        raise ValueError(f'{rows} rows do not split into {n_gates} gates of 16-row blocks')
    H = rows // n_gates
    keep = 1.0 - float(sparsity)
    mask = np.zeros(w.shape, dtype=np.float32)
    for g in range(n_gates):
        blk = np.abs(np.asarray(w[g * H:(g + 1) * H], dtype=np.float64)).reshape(H // 16, 16, K)
        norm = blk.sum(axis=1).reshape(-1)                              # [block_row * K + col]
        n_keep = min(norm.size, max(1, int(math.floor(keep * norm.size + 0.5))))
        order = np.lexsort((np.arange(norm.size), -norm))
        m = np.zeros(norm.size, dtype=bool)
        m[order[:n_keep]] = True
        mask[g * H:(g + 1) * H] = np.repeat(m.reshape(H // 16, 1, K), 16, axis=1).reshape(H, K)
    return mask


def elementwise_mask(w: np.ndarray, n_splits: int, z: float) -> np.ndarray:
    """`PruneMask.mask_from_matrix` of the notebook (:92-108): per split, zero the k = int(n * z) smallest magnitudes
    (everything below the k-th smallest; ties at the threshold are kept, as `W_abs >= threshold` does)."""
    rows = w.shape[0]
    if rows % n_splits:
        raise ValueError('rows do not split evenly')
    H = rows // n_splits
    out = []
    for s in range(n_splits):
        part = np.abs(np.asarray(w[s * H:(s + 1) * H]))
        sorted_abs = np.sort(part.reshape(-1))
        k = int(part.shape[0] * part.shape[1] * z)
        threshold = sorted_abs[min(k, sorted_abs.size - 1)]
        out.append((part >= threshold).astype(np.float32))
    return np.concatenate(out)


def _as_numpy(x):
    return x.detach().cpu().numpy() if hasattr(x, 'detach') else np.asarray(x)


class Pruner:
    """The notebook's `Pruner` (:110-160) over a state dict, with 16x1 block masks for the GRU matrices.

    `state` maps parameter names to numpy arrays or torch tensors (a `state_dict()`); `keys` are the matrices to prune
    (default: the four GRU matrices of the fatchord model; `prune_rnn_input=False` leaves `weight_ih` alone as the
    notebook's flag does).  Call `prune(state, t)` once per training step, after the optimiser step.
    """

    def __init__(self, start_prune, prune_steps, target_sparsity, prune_rnn_input=True, prune_every=500,
                 keys: Optional[Iterable[str]] = None, n_gates=3, block=True):
        self.z = 0.0                    # sparsity at time t
        self.t_0 = start_prune
        self.S = prune_steps
        self.Z = target_sparsity
        self.prune_every = prune_every
        self.n_gates = n_gates
        self.block = block
        keys = tuple(keys) if keys is not None else GRU_KEYS
        self.keys = tuple(k for k in keys if prune_rnn_input or 'weight_ih' not in k)
        self.masks: Dict[str, np.ndarray] = {}
        self.num_pruned = 0
        self.total_params = 0

    def update_sparsity(self, t):
        t = int(_as_numpy(t).reshape(-1)[0]) if not isinstance(t, (int, float)) else t
        self.z = sparsity_at(t, self.t_0, self.S, self.Z)
        return t

    def prune_or_not(self, t):
        return t % self.prune_every == 0 and t > self.t_0

    def apply_or_not(self, t):
        return t >= self.t_0

    def _mask(self, w):
        w = _as_numpy(w)
        return block_mask_16x1(w, self.n_gates, self.z) if self.block else elementwise_mask(w, self.n_gates, self.z)

    def update_masks(self, state):
        for k in self.keys:
            self.masks[k] = self._mask(state[k])
        self.count_num_pruned()

    def apply_masks(self, state):
        for k, m in self.masks.items():
            w = state[k]
            if hasattr(w, 'detach'):                # torch parameter / tensor: in place, like `W *= M`
                import torch
                with torch.no_grad():
                    w.mul_(torch.as_tensor(m, dtype=w.dtype, device=w.device))
            else:
                w *= m.astype(w.dtype)

    def prune(self, state, t):
        t = self.update_sparsity(t)
        if self.prune_or_not(t):
            self.update_masks(state)
        if self.apply_or_not(t) and self.masks:
            self.apply_masks(state)
        return self.z

    def restart(self, state, t):
        """After a stop: recompute the masks for the sparsity of step t (notebook :150-154)."""
        self.update_sparsity(t)
        self.update_masks(state)

    def count_num_pruned(self):
        self.num_pruned = int(sum(int((1 - m).sum()) for m in self.masks.values()))
        self.total_params = int(sum(m.size for m in self.masks.values()))
