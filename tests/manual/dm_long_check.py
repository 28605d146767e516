"""One-off: the deepmind kernel against the oracle over a whole 2 s sequence (44 100 samples, injected noise), and eight
sequences x 8 000 samples: every coarse / fine value must be equal."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import deepmind as odm
from wavernn_b200 import binding, synth

sd = synth.synth_dm_state_dict(896, 256, seed=0)
W = {k: torch.as_tensor(v) for k, v in sd.items()}
v = binding.DeepmindVocoder(896, 256, device=0)
v.load_weights(sd)
for n_seq, T in ((1, 44100), (8, 8000)):
    g = torch.Generator().manual_seed(99 + n_seq)
    qc = torch.empty(T, n_seq, 256).exponential_(1.0, generator=g).numpy()
    qf = torch.empty(T, n_seq, 256).exponential_(1.0, generator=g).numpy()
    out, c, f = v.generate(T, n_seq, q_coarse=qc, q_fine=qf)
    t0 = time.time()
    o_out, o_c, o_f = odm.generate(W, T, qc, qf, n_seq=n_seq)
    o_c, o_f = np.asarray(o_c).reshape(n_seq, T), np.asarray(o_f).reshape(n_seq, T)
    bad = int((c != o_c).sum() + (f != o_f).sum())
    first = int(np.argmax((c != o_c).any(axis=0) | (f != o_f).any(axis=0))) if bad else -1
    print(f'{n_seq} x {T}: {bad} differing values (first at step {first}); oracle {time.time() - t0:.1f} s', flush=True)
