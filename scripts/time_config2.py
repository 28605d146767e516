import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=0)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
T = int(sys.argv[1]) if len(sys.argv) > 1 else 801
batched = not (len(sys.argv) > 2 and sys.argv[2] == 'unbatched')
mel = synth.synth_mel(T, seed=1234)
for i in range(3):
    t0 = time.time()
    out = v.generate([mel[0]], batched, 11000, 550, False, seed=7)
    dt = time.time() - t0
    t = v.last_timing()
    print(json.dumps({'kernel': v.loop_kernel_name(t['folds']), 'folds': t['folds'], 'us_per_step': round(1000 * t['ms_loop'] / t['steps'], 3),
                      'x_rt': round(len(out[0]) / dt / 22050, 2)}), flush=True)
