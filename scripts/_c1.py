import sys, os, time, json
sys.path.insert(0, '/root/repo')
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
dims = VocoderDims(mode='MOL')
sd = synth.synth_state_dict(dims, seed=0, style='init')
v = binding.Vocoder(dims, device=0); v.load_weights(sd)
for T, batched in ((161, False), (201, True), (401, True)):
    mel = synth.synth_mel(T, seed=1234)
    v.generate([mel[0]], batched, 11000, 550, False, seed=7)
    t0 = time.time(); out = v.generate([mel[0]], batched, 11000, 550, False, seed=7); dt = time.time() - t0
    t = v.last_timing()
    print(T, batched, 'folds', t['folds'], 'us/step', round(1000 * t['ms_loop'] / t['steps'], 2), flush=True)
