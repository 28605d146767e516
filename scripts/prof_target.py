"""One short call of one loop kernel, for ncu captures (keeps the ~40 replays of `ncu --set full` short).

  python scripts/prof_target.py mma     640 folds x 300 steps on the tensor-core kernel
  python scripts/prof_target.py res     one utterance, 20 folds x 2 200 steps on the register-resident kernel
  python scripts/prof_target.py taco    Tacotron decoder loop, 150 characters, 200 decoder steps
  python scripts/prof_target.py dm      deepmind kernel, 8 sequences x 2 000 samples
`mma`: 640 folds x 300 steps on the tensor-core kernel (T=582, target 200, overlap 50: the shape tests/test_mma_gpu.py checks).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from wavernn_b200 import binding, synth      # noqa: E402
from wavernn_b200.config import VocoderDims  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else 'mma'
if what == 'dm':        # deepmind coarse/fine kernel: 8 sequences x 2 000 samples
    v = binding.DeepmindVocoder(896, 256, device=0)
    v.load_weights(synth.synth_dm_state_dict(896, 256, seed=0))
    out, c, f = v.generate(2000, 8, seed=3)
    print(what, out.shape, v.last_timing())
    sys.exit(0)
if what == 'taco':      # the Tacotron decoder loop: 150 characters, r = 2, 200 decoder steps
    dec = binding.TacotronDecoder(device=0)
    dec.load_weights(synth.synth_taco_state_dict(seed=0))
    mel, attn = dec.decode(synth.synth_encoder_seq(150, seed=78), steps=400, r=2, stop_threshold=-1e30)
    print(what, mel.shape, attn.shape, dec.last_timing())
    sys.exit(0)
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
if what == 'mma':
    mel, target, overlap = synth.synth_mel(582, seed=1234), 200, 50
else:
    mel, target, overlap = synth.synth_mel(161, seed=1234), 1100, 550
w = v.generate([mel[0]], True, target, overlap, False, seed=1)[0]
t = v.last_timing()
print(what, v.loop_kernel_name(t['folds']), t, float(abs(w).max()))
