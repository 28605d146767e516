"""Prints the resident kernel's exchange calibration (L2 round trip in SM clocks, delay scale) and the step time with and
without it; run several times to see the spread of the measurement."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth            # noqa: E402
from wavernn_b200.config import VocoderDims         # noqa: E402

dims = VocoderDims(mode='MOL')
for rep in range(4):
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    mel = synth.synth_mel(161, seed=1)
    v.generate([mel[0]], True, 1100, 550, False, seed=1)
    v.generate([mel[0]], True, 1100, 550, False, seed=1)
    t = v.last_timing()
    print('calibration', v.exchange_calibration(), f"{1000 * t['ms_loop'] / t['steps']:.2f} us/step at {t['folds']} folds", flush=True)
