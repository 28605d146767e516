"""Loop time of one variant of config 2 on one GPU: `python scripts/time_variant.py raw|sp|mol` (RAW 9-bit mu-law,
16x1 block-sparse GRU, dense MoL).  scripts/time_config2.py times the dense MoL kernel with a fold count argument."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
which = sys.argv[1]
dims = VocoderDims(mode='RAW', bits=9) if which == 'raw' else VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
if which == 'sp': v.prune_16x1(0.1)
mel = synth.synth_mel(801, seed=1234)
for i in range(2):
    out = v.generate([mel[0]], True, 11000, 550, which == 'raw', seed=7)
    t = v.last_timing()
print(json.dumps({'which': which, 'kernel': v.loop_kernel_name(20), 'us_per_step': round(1000 * t['ms_loop'] / t['steps'], 3)}), flush=True)
