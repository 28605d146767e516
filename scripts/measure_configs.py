"""Rates of the SURVEY.md section 8(d) configurations through the public binding on one GPU (not bench lines:
bench.py times config 2; these are the numbers DESIGN.md section 6 quotes for the other configurations).

    python scripts/measure_configs.py            # prints one JSON line per configuration
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth            # noqa: E402
from wavernn_b200.config import VocoderDims         # noqa: E402

SR = 22050


def timed(fn, reps=2):
    fn()
    t0 = time.time()
    for _ in range(reps):
        out = fn()
    return (time.time() - t0) / reps, out


def fatchord(name, dims, T, batched, mu_law=False, n_utt=1, prune=None, reps=2):
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    if prune:
        v.prune_16x1(prune)
    mels = [synth.synth_mel(T, seed=1234 + u) for u in range(n_utt)]
    dt, out = timed(lambda: v.generate([m[0] for m in mels], batched, 11000, 550, mu_law, seed=7), reps)
    t = v.last_timing()
    n = sum(len(o) for o in out)
    print(json.dumps({'config': name, 'utterances': n_utt, 'folds': t['folds'], 'steps': t['steps'],
                      'kernel': v.loop_kernel_name(t['folds']), 'samples': n, 'seconds': round(dt, 4),
                      'samples_per_s': round(n / dt), 'x_real_time': round(n / dt / SR, 2),
                      'loop_us_per_step': round(1000 * t['ms_loop'] / t['steps'], 2)}), flush=True)


def deepmind(n_seq, seq_len=44100):
    v = binding.DeepmindVocoder(896, 256, device=0)
    v.load_weights(synth.synth_dm_state_dict(896, 256, seed=0))
    dt, out = timed(lambda: v.generate(seq_len, n_seq, seed=3), 1)
    print(json.dumps({'config': f'deepmind 896, {n_seq} sequence(s) x {seq_len}', 'seconds': round(dt, 3),
                      'samples_per_s': round(n_seq * seq_len / dt), 'us_per_step': round(1e6 * dt / seq_len, 2)}), flush=True)


if __name__ == '__main__':
    mol = VocoderDims(mode='MOL')
    raw = VocoderDims(mode='RAW', bits=9)
    fatchord('1: MoL unbatched T=161 (B=1, S=44275)', mol, 161, False)
    fatchord('2: MoL batched T=801 (B=20)', mol, 801, True)
    fatchord('2b: MoL batched T=800 (B=19, no tail pad)', mol, 800, True)
    fatchord('3: RAW 9-bit mu-law batched T=801 (B=20)', raw, 801, True, mu_law=True, reps=1)
    fatchord('4: config 2 with 16x1 block-pruned GRU (10 % kept), 16x1 block-sparse tables', mol, 801, True, prune=0.1)
    fatchord('5, one GPU of eight (32 of 256 utterances): MoL batched T=801, 640 folds in one launch', mol, 801, True, n_utt=32, reps=1)
    fatchord('5 shape, 8 utterances: 160 folds in one launch', mol, 801, True, n_utt=8, reps=1)
    deepmind(1)
    deepmind(8)
