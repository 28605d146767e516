"""Bring-up: 1 152 folds (nine tiles, CTAs with a GRU job and an fc job) against the oracle: which folds / steps differ."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import fatchord as ofc                 # noqa: E402
from wavernn_b200 import binding, synth            # noqa: E402
from wavernn_b200.config import VocoderDims         # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1047
dims = VocoderDims(mode='MOL')
sd = synth.synth_state_dict(dims, seed=0, style='trained')
v = binding.Vocoder(dims, device=0)
v.load_weights(sd)
W = ofc.to_torch(sd)
target, overlap = 200, 50
mel = synth.synth_mel(T, seed=1234)
with torch.no_grad():
    m, aux, wave_len = ofc.conditioning(W, dims, mel, True, target, overlap)
    B, S = int(m.shape[0]), int(m.shape[1])
for rep, seed in enumerate((4321, 99, 7)):
    noise = synth.synth_noise_mol(S, B, seed)
    with torch.no_grad():
        res = ofc.run_loop(W, dims, m, aux, noise)
    ref = res['samples'].numpy()
    v.set_loop_kernel(4)
    v.generate([mel[0]], True, target, overlap, False, noise=noise)
    got = v.fetch_samples()
    d = np.abs(got.astype(np.float64) - ref)
    bad = np.argwhere(d.max(axis=1) > 5e-3).reshape(-1)
    first = [int(np.argmax(d[b] > 5e-3)) for b in bad]
    print(f'run {rep}: folds {B}, bad folds {len(bad)}: tiles {sorted(set(int(b) // 128 for b in bad))}, '
          f'in-tile {sorted(set(int(b) % 128 for b in bad))[:40]}, first bad steps {sorted(set(first))[:20]}', flush=True)
