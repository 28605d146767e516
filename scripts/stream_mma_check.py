"""Chunked delivery with chunks large enough for the tensor-core kernel (64 folds per group): equal to generate() bit for bit."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth            # noqa: E402
from wavernn_b200.config import VocoderDims         # noqa: E402

dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
target, overlap = 500, 50
T = 160 * (target + overlap) // dims.hop_length
mel = synth.synth_mel(T, seed=6)
B = binding.num_folds(T * dims.hop_length, target, overlap)
noise = synth.synth_noise_mol(target + 2 * overlap, B, 5)
ref = v.generate([mel[0]], True, target, overlap, False, noise=noise)[0]
print('folds', B, 'kernel', v.loop_kernel_name(B))
got = []
wave = v.generate_stream(mel[0], target, overlap, False, on_chunk=lambda first, x, last: got.append((first, len(x), last)), chunk_folds=64, noise=noise)
print('chunks', got, 'launches', v.last_timing()['launches'])
print('equal to generate():', np.array_equal(wave, ref), float(np.abs(wave - ref).max()))
assert np.array_equal(wave, ref)
