"""One long utterance, folds sharded over the GPUs of the box; run under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
      scripts/sharded_long_utterance.py [T_frames]

T = 6401 frames (80 s of audio) -> 160 folds x 12 100 steps: 20 folds per GPU at 8 GPUs.  Rank 0 prints one JSON line:
wall time of the sharded generate (barrier on both sides, max over ranks), samples/s, and the check against the
unsharded run of the same utterance on rank 0 with the same kernel and the same Philox seed.
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from wavernn_b200 import binding, synth           # noqa: E402
from wavernn_b200.config import VocoderDims       # noqa: E402
from wavernn_b200.shard import generate_sharded   # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 6401
world, rank, local = int(os.environ.get('WORLD_SIZE', '1')), int(os.environ.get('RANK', '0')), int(os.environ.get('LOCAL_RANK', '0'))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
dims = VocoderDims(mode='MOL')
v = binding.Vocoder(dims, device=local)
v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
v.set_loop_kernel(3)
mel = synth.synth_mel(T, seed=1234)[0]


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


generate_sharded(v, mel, 11000, 550, seed=1)       # warm-up (packing, workspaces, NCCL)
times = []
for i in range(3):
    barrier()
    t0 = time.time()
    wave = generate_sharded(v, mel, 11000, 550, seed=7)
    barrier()
    times.append(time.time() - t0)
if rank == 0:
    t0 = time.time()
    whole = v.generate([mel], True, 11000, 550, False, seed=7)[0]
    t_whole = time.time() - t0
    n = (T - 1) * dims.hop_length
    print(json.dumps({'test': 'fold-sharded long utterance', 'T_frames': T, 'audio_s': n / 22050, 'folds': binding.num_folds(T * 275, 11000, 550),
                      'n_gpus': world, 'kernel': 'fma-resident (forced for shards and whole)', 'seconds_sharded': min(times),
                      'samples_per_s_sharded': n / min(times), 'x_real_time': n / min(times) / 22050,
                      'seconds_one_gpu_same_kernel': t_whole, 'max_abs_diff_vs_one_gpu': float(np.abs(wave - whole).max()),
                      'shape_ok': wave.shape == (n,), 'finite': bool(np.all(np.isfinite(wave)))}))
if world > 1:
    dist.destroy_process_group()
