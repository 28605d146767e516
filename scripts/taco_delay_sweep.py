"""Sweep of the first-look delay and the poll back-off of the decoder loop's tagged exchange (GPU box)."""
import os
import subprocess
import sys

code = '''
import sys
sys.path.insert(0, ".")
from wavernn_b200 import synth
from wavernn_b200.binding import TacotronDecoder
sd = synth.synth_taco_state_dict(seed=0)
dec = TacotronDecoder(); dec.load_weights(sd)
enc = synth.synth_encoder_seq(150, seed=78)
best = 1e9
for _ in range(3):
    mel, attn = dec.decode(enc, steps=800, r=2, stop_threshold=-1e9)
    best = min(best, 1e3 * dec.last_timing()["ms_loop"] / attn.shape[0])
print(f"{best:.2f}")
'''
for d1 in (0, 300, 600, 900, 1200):
    for d2 in (0, 100, 300):
        env = dict(os.environ, WRNN_TACO_D1=str(d1), WRNN_TACO_D2=str(d2))
        out = subprocess.run([sys.executable, '-c', code], env=env, capture_output=True, text=True, timeout=120)
        print(f'd_first {d1:5d} d_poll {d2:4d}: {out.stdout.strip()} us/step {out.stderr.strip()[-200:]}', flush=True)
