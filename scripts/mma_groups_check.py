"""More folds than one tensor-core launch holds (1 152): 20 utterances x 65 folds = 1 300 folds run as two launches of 650;
every utterance must equal its own single-utterance call under the same injected noise."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth            # noqa: E402
from wavernn_b200.config import VocoderDims         # noqa: E402

dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
target, overlap, T, n_utt = 200, 50, 59, 20
mels = [synth.synth_mel(T, seed=100 + u)[0] for u in range(n_utt)]
per = binding.num_folds(T * dims.hop_length, target, overlap)
S = target + 2 * overlap
noise = synth.synth_noise_mol(S, per * n_utt, 99)
v.set_loop_kernel(4)
waves = v.generate(mels, True, target, overlap, False, noise=noise)
t = v.last_timing()
print('folds', t['folds'], 'launches', t['launches'], 'kernel', v.loop_kernel_name(t['folds']))
worst = 0.0
for u in range(n_utt):
    nz = tuple(np.ascontiguousarray(a[:, u * per:(u + 1) * per]) for a in noise)
    one = v.generate([mels[u]], True, target, overlap, False, noise=nz)[0]
    worst = max(worst, float(np.abs(one - waves[u]).max()))
print('max |batch - single| over', n_utt, 'utterances:', worst)
assert worst == 0.0
