"""Step time of the register-resident kernel (3) and the tensor-core kernel (4) by fold count: where kernel selection should switch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
for n_utt in (2, 3, 4, 6, 16, 48, 57):
    mels = [synth.synth_mel(801, seed=1234 + u)[0] for u in range(n_utt)]
    for k in (3, 4):
        v.set_loop_kernel(k)
        v.stage(mels, True, 11000, 550, False)
        for rep in range(2):
            v.run_staged(seed=1); v.fetch()
        t = v.last_timing()
        print(n_utt, 'utt', t['folds'], 'folds', v.loop_kernel_name(t['folds']), f"{1000*t['ms_loop']/t['steps']:.2f} us/step  cond {t['ms_cond']:.2f} ms", flush=True)
