"""Golden vectors of the Tacotron decoder loop from the REAL reference (build container only).

    python tests/golden/make_taco_golden.py

For every case: build the reference's `Tacotron` (oracle/ref_shim.build_tacotron), load the seeded decoder
parameters, replace the encoder's output by a seeded `enc_seq`, run the reference's own `generate`, run the oracle
restatement (oracle/tacotron.py) on the same inputs, assert mel and attention are BIT-EQUAL, and write the
reference's outputs to tests/golden/taco_<case>.npz.  Inputs are rebuilt from the seeds stored in each file.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_shim                 # noqa: E402
from oracle import tacotron as otc          # noqa: E402
from wavernn_b200 import synth              # noqa: E402

# name -> (L, r, steps, stop_threshold, weight seed, enc seed)
CASES = {
    'short_r2': (37, 2, 60, -3.4, 0, 77),
    'long_r2': (150, 2, 400, -3.4, 0, 78),
    'r7': (64, 7, 140, -3.4, 1, 79),
    'r1_one_char': (1, 1, 24, -3.4, 2, 80),
    'stop_rule': (50, 2, 200, 1e9, 0, 81),       # every frame is "silent": the loop ends at the first t > 10
}


def main():
    for name, (L, r, steps, thr, wseed, eseed) in CASES.items():
        sd = synth.synth_taco_state_dict(seed=wseed)
        enc = synth.synth_encoder_seq(L, seed=eseed)
        model = ref_shim.build_tacotron(sd, enc, r=r, stop_threshold=thr)
        mel, _linear, attn = model.generate([1 + (i % 60) for i in range(L)], steps=steps)
        o_mel, o_attn = otc.decode(sd, enc, steps=steps, r=r, stop_threshold=thr)
        for a, b, what in ((mel, o_mel, 'mel'), (attn, o_attn, 'attn')):
            if a.shape != b.shape or not np.array_equal(a, b):
                d = np.abs(a.astype(np.float64) - b).max() if a.shape == b.shape else (a.shape, b.shape)
                raise SystemExit(f'ORACLE != REFERENCE for {name} {what}: {d}')
        np.savez_compressed(os.path.join(HERE, f'taco_{name}.npz'), mel=mel, attn=attn, L=L, r=r, steps=steps,
                            stop_threshold=thr, weight_seed=wseed, enc_seed=eseed)
        print(f'{name}: mel {mel.shape} range [{mel.min():.3f}, {mel.max():.3f}]  attn {attn.shape} '
              f'peak {attn.max():.4f}  bit-equal to the oracle')


if __name__ == '__main__':
    main()
