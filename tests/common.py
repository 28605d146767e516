"""Shared helpers for the tests: golden cases, synthetic inputs, oracle runs."""
import os

import numpy as np

from wavernn_b200 import synth
from wavernn_b200.config import VocoderDims

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')

TINY = dict(rnn_dims=64, fc_dims=64, bits=6, pad=2, upsample_factors=(2, 2, 3), feat_dims=16,
            compute_dims=32, res_out_dims=32, res_blocks=2, hop_length=12, sample_rate=22050)

# must match tests/golden/make_golden.py CASES
CASES = {
    'tiny_mol_unbatched': (VocoderDims(mode='MOL', **TINY), 'trained', 40, False, 100, 20, False),
    'tiny_mol_batched': (VocoderDims(mode='MOL', **TINY), 'trained', 61, True, 100, 20, False),
    'tiny_mol_batched_nopad': (VocoderDims(mode='MOL', **TINY), 'init', 52, True, 100, 4, False),
    'tiny_raw_batched_mulaw': (VocoderDims(mode='RAW', **TINY), 'trained', 61, True, 100, 20, True),
    'tiny_raw_unbatched': (VocoderDims(mode='RAW', **TINY), 'init', 40, False, 100, 20, False),
    'full_mol_unbatched': (VocoderDims(mode='MOL'), 'init', 22, False, 11000, 550, False),
    'full_mol_batched': (VocoderDims(mode='MOL'), 'trained', 24, True, 2000, 550, False),
    'full_raw_batched_mulaw': (VocoderDims(mode='RAW'), 'init', 24, True, 2000, 550, True),
}


def golden(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'))


def loop_extent(dims, T, batched, target, overlap):
    L = T * dims.hop_length
    if not batched:
        return 1, L
    from oracle.fatchord import num_folds_for
    return num_folds_for(L, target, overlap), target + 2 * overlap


def noise_for(dims, S, B, seed=4321):
    if dims.mode == 'MOL':
        return synth.synth_noise_mol(S, B, seed)
    return synth.synth_noise_exp(S, B, dims.n_classes, seed)


def case_inputs(name):
    dims, style, T, batched, target, overlap, mu_law = CASES[name]
    sd = synth.synth_state_dict(dims, seed=0, style=style)
    mel = synth.synth_mel(T, seed=1234, feat_dims=dims.feat_dims)
    B, S = loop_extent(dims, T, batched, target, overlap)
    return dims, sd, mel, noise_for(dims, S, B), (T, batched, target, overlap, mu_law, B, S)
