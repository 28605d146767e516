"""Per-site clocks of the decoder loop kernel (WRNN_TACO_PROF=1), L = 150, r = 2, 400 steps."""
import os
import sys

sys.path.insert(0, '.')
os.environ['WRNN_TACO_PROF'] = '1'
from wavernn_b200 import synth
from wavernn_b200.binding import TacotronDecoder

sd = synth.synth_taco_state_dict(seed=0)
dec = TacotronDecoder()
dec.load_weights(sd)
enc = synth.synth_encoder_seq(int(sys.argv[1]) if len(sys.argv) > 1 else 150, seed=78)
dec.decode(enc, steps=800, r=2, stop_threshold=-1e9)
mel, attn = dec.decode(enc, steps=800, r=2, stop_threshold=-1e9)
tm = dec.last_timing()
print(f'{attn.shape[0]} steps, {1e3 * tm["ms_loop"] / attn.shape[0]:.2f} us/step')
