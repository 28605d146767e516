"""Decoder-loop timing at the fixture sizes (GPU box): us per decoder step, error against the oracle."""
import sys
import time

import numpy as np

sys.path.insert(0, '.')
from oracle import tacotron as otc
from wavernn_b200 import synth
from wavernn_b200.binding import TacotronDecoder

sd = synth.synth_taco_state_dict(seed=0)
dec = TacotronDecoder()
dec.load_weights(sd)
for L, r, steps in ((37, 2, 60), (150, 2, 400), (150, 2, 1600), (64, 7, 140), (400, 2, 400)):
    enc = synth.synth_encoder_seq(L, seed=78)
    t0 = time.time()
    mel, attn = dec.decode(enc, steps=steps, r=r, stop_threshold=-1e9)
    t1 = time.time()
    mel, attn = dec.decode(enc, steps=steps, r=r, stop_threshold=-1e9)
    t2 = time.time()
    tm = dec.last_timing()
    line = f'L={L} r={r} steps={steps}: {attn.shape[0]} decoder steps, loop {tm["ms_loop"]:.3f} ms = {1e3 * tm["ms_loop"] / attn.shape[0]:.2f} us/step, call {1e3 * (t2 - t1):.2f} ms (first call {1e3 * (t1 - t0):.0f} ms)'
    if steps <= 400:
        o_mel, o_attn = otc.decode(sd, enc, steps=steps, r=r, stop_threshold=-1e9)
        line += f'  max|mel - oracle| {np.abs(mel - o_mel).max():.2e}  max|attn - oracle| {np.abs(attn - o_attn).max():.2e}'
    print(line, flush=True)
