"""Repeat run of the Tacotron decoder loop kernel over random text lengths, reduction factors and frame budgets (GPU box):
every call is checked against the oracle, the handle is reused across calls (repacking on a change of r)."""
import sys
import time

import numpy as np

sys.path.insert(0, '.')
from oracle import tacotron as otc
from wavernn_b200 import synth
from wavernn_b200.binding import TacotronDecoder

n_calls = int(sys.argv[1]) if len(sys.argv) > 1 else 150
rs = np.random.RandomState(2024)
sd = synth.synth_taco_state_dict(seed=7)
dec = TacotronDecoder()
dec.load_weights(sd)
worst_mel = worst_attn = 0.0
skipped = 0
t0 = time.time()
for i in range(n_calls):
    L = int(rs.choice([1, 2, 3, 7, 19, 20, 21, 40, 41, 64, 150, 333, 480, 512])) if i % 3 == 0 else int(rs.randint(1, 513))
    r = int(rs.choice([1, 2, 2, 2, 3, 5, 7]))
    steps = int(rs.randint(1, 90))
    thr = float(rs.choice([-3.4, -3.4, 0.0, 1e9]))
    enc = synth.synth_encoder_seq(L, seed=1000 + i)
    try:
        mel, attn = dec.decode(enc, steps=steps, r=r, stop_threshold=thr)
    except ValueError as e:            # the longest texts at a large r exceed the shared memory: a clear error, the handle stays usable
        if 'shared memory' not in str(e):
            raise
        skipped += 1
        continue
    o_mel, o_attn = otc.decode(sd, enc, steps=steps, r=r, stop_threshold=thr)
    if mel.shape != o_mel.shape or attn.shape != o_attn.shape:
        raise SystemExit(f'call {i}: L={L} r={r} steps={steps} thr={thr}: shapes {mel.shape} {attn.shape} vs oracle {o_mel.shape} {o_attn.shape}')
    em, ea = float(np.abs(mel - o_mel).max()), float(np.abs(attn - o_attn).max())
    worst_mel, worst_attn = max(worst_mel, em), max(worst_attn, ea)
    if em > 2e-5 or ea > 5e-6:
        raise SystemExit(f'call {i}: L={L} r={r} steps={steps} thr={thr}: max|mel| {em:.2e} max|attn| {ea:.2e}')
print(f'{n_calls} calls ({skipped} refused for shared memory) ok in {time.time() - t0:.1f} s: worst max|mel - oracle| {worst_mel:.2e}, worst max|attn - oracle| {worst_attn:.2e}')
