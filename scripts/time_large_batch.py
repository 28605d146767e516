import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
k = int(sys.argv[1]); n = int(sys.argv[2]); T = int(sys.argv[3]) if len(sys.argv) > 3 else 201
v.set_loop_kernel(k)
mels = [synth.synth_mel(T, seed=1234 + u)[0] for u in range(n)]
for i in range(2):
    t0 = time.time(); out = v.generate(mels, True, 11000, 550, False, seed=7); dt = time.time() - t0
    t = v.last_timing()
    print(json.dumps({'force': k, 'folds': t['folds'], 'steps': t['steps'], 'ms_loop': round(t['ms_loop'], 1), 'us_per_step_all_folds': round(1000 * t['ms_loop'] / t['steps'], 2),
                      'fold_steps_per_us': round(t['folds'] * t['steps'] / (1000 * t['ms_loop']), 3)}), flush=True)
