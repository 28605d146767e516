"""Bring-up: repeat short teacher-forced runs of the tensor-core loop kernel; stops at the first error (prints it).

  python scripts/mma_stress.py [n_seq] [hops] [repeats]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from wavernn_b200 import binding, synth      # noqa: E402
from wavernn_b200.config import VocoderDims  # noqa: E402

n_seq = int(sys.argv[1]) if len(sys.argv) > 1 else 64
hops = int(sys.argv[2]) if len(sys.argv) > 2 else 4
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
v.set_loop_kernel(4)
rng = np.random.default_rng(0)
S = hops * dims.hop_length
xs = (rng.random((n_seq, S), dtype=np.float32) * 2 - 1)
mels = rng.random((n_seq, dims.feat_dims, hops + 2 * dims.pad), dtype=np.float32)
ref = None
for i in range(reps):
    t0 = time.time()
    try:
        out = v.forward_logits(xs, mels)
    except Exception as e:      # noqa: BLE001
        print(f'run {i}: FAILED after {time.time() - t0:.2f} s: {e}', flush=True)
        sys.exit(1)
    if ref is None:
        ref = out
    print(f'run {i}: ok {time.time() - t0:.3f} s, max |diff to run 0| = {np.abs(out - ref).max():.3g}', flush=True)
