"""Timeline of one step of the tensor-core loop kernel (WRNN_MMA_PROF=<step>; the library prints it to stderr).

  WRNN_MMA_PROF=100 WRNN_MMA_STEPS=200 [MMA_NT=96] [MMA_STYLE=trained] [MMA_T=801 MMA_TARGET=11000 MMA_OVERLAP=550] \
      python scripts/mma_prof.py [n_utt ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from wavernn_b200 import binding, synth      # noqa: E402
from wavernn_b200.config import VocoderDims  # noqa: E402

dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style=os.environ.get('MMA_STYLE', 'init')))
v.set_loop_kernel(4)
v.set_mma_fold_tile(int(os.environ.get('MMA_NT', '0')))
T = int(os.environ.get('MMA_T', '801'))
target, overlap = int(os.environ.get('MMA_TARGET', '11000')), int(os.environ.get('MMA_OVERLAP', '550'))
for n_utt in [int(a) for a in sys.argv[1:]] or [3, 32]:
    mels = [synth.synth_mel(T, seed=1234 + u)[0] for u in range(n_utt)]
    v.stage(mels, True, target, overlap, False)
    for rep in range(2):
        v.run_staged(seed=1)
        v.fetch()
        t = v.last_timing()
        steps = min(int(os.environ.get('WRNN_MMA_STEPS', t['steps'])), t['steps'])
        print(f'utterances {n_utt} folds {t["folds"]} steps run {steps} (run {rep}): {1000 * t["ms_loop"] / steps:.2f} us/step', flush=True)
