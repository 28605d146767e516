import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
for T, batched in ((161, False), (401, True), (201, True)):
    mel = synth.synth_mel(T, seed=1234)
    for i in range(2):
        out = v.generate([mel[0]], batched, 11000, 550, False, seed=7)
        t = v.last_timing()
    print(json.dumps({'T': T, 'folds': t['folds'], 'us_per_step': round(1000 * t['ms_loop'] / t['steps'], 3)}), flush=True)
