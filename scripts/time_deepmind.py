import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth
for n_seq in (1, 8):
    v = binding.DeepmindVocoder(896, 256, device=0)
    v.load_weights(synth.synth_dm_state_dict(896, 256, seed=0))
    v.generate(2000, n_seq, seed=3)
    t0 = time.time(); out = v.generate(44100, n_seq, seed=3); dt = time.time() - t0
    print(json.dumps({'n_seq': n_seq, 'us_per_step': round(1e6 * dt / 44100, 2), 'samples_per_s': round(n_seq * 44100 / dt)}), flush=True)
