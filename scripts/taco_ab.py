"""A/B of two builds of the library on the decoder loop (GPU box): python scripts/taco_ab.py variants/a.so variants/b.so"""
import os
import subprocess
import sys

code = '''
import sys
sys.path.insert(0, ".")
from wavernn_b200 import binding, synth
binding.LIB_PATH = sys.argv[1]
sd = synth.synth_taco_state_dict(seed=0)
dec = binding.TacotronDecoder(); dec.load_weights(sd)
enc = synth.synth_encoder_seq(150, seed=78)
best = 1e9
for _ in range(4):
    mel, attn = dec.decode(enc, steps=1600, r=2, stop_threshold=-1e30)
    best = min(best, 1e3 * dec.last_timing()["ms_loop"] / attn.shape[0])
print(f"{best:.2f}")
'''
for rep in range(2):
    for lib in sys.argv[1:]:
        out = subprocess.run([sys.executable, '-c', code, os.path.abspath(lib)], capture_output=True, text=True, timeout=120)
        print(lib, out.stdout.strip(), 'us/step', out.stderr.strip()[-200:], flush=True)
