"""Scratch: config 3 (RAW 9-bit) timing / in-kernel profile."""
import os, sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
dims = VocoderDims(mode='RAW')
sd = synth.synth_state_dict(dims, seed=0, style='init')
mel = synth.synth_mel(801, seed=1, feat_dims=dims.feat_dims)
v = binding.Vocoder(dims, device=0)
v.load_weights(sd)
print(v.loop_kernel_name(20))
for i in range(2):
    w = v.generate([mel[0]], True, 11000, 550, True, seed=3)[0]
    t = v.last_timing()
    print(t['folds'], t['steps'], 'loop_ms', round(t['ms_loop'], 2), 'us/step', round(1e3 * t['ms_loop'] / t['steps'], 2))
