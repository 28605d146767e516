"""Bring-up / measurement of the tensor-core loop kernel (csrc/loop_mma.cu) on a B200, through the public binding.

  python tests/manual/mma_check.py [parity] [timing]

parity: teacher-forced logits and free-running samples of the forced "tcgen05-bf16x3" kernel against the CPU oracle
        (oracle/fatchord.py) at 2 / 64 / 130 / 640 folds, for several fold-tile widths;
timing: config-5 shapes (3 / 8 / 32 utterances of T = 801 -> 60 / 160 / 640 folds x 12 100 steps), us per step.
Prints one JSON object per measurement (kept under profiles/ by the caller).
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import fatchord as ofc          # noqa: E402  (checker only)
from wavernn_b200 import binding, synth     # noqa: E402
from wavernn_b200.config import VocoderDims  # noqa: E402


def emit(**kw):
    print(json.dumps(kw), flush=True)


def relerr(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(np.abs(b).max(), 1e-30))


def teacher_forced(v, dims, W, n_seq, S, style, nts):
    g = torch.Generator().manual_seed(99 + n_seq)
    xs = (torch.rand((n_seq, S), generator=g) * 2 - 1).numpy()
    mels = torch.rand((n_seq, dims.feat_dims, S // dims.hop_length + 2 * dims.pad), generator=g).numpy()
    with torch.no_grad():
        ms, auxs = [], []
        for b in range(n_seq):
            m, aux = ofc.upsample(W, dims, torch.as_tensor(mels[b:b + 1]))
            ms.append(m[0])
            auxs.append(aux[0])
        m, aux = torch.stack(ms), torch.stack(auxs)
        h1 = torch.zeros(n_seq, dims.rnn_dims)
        h2 = torch.zeros(n_seq, dims.rnn_dims)
        logits = []
        d = dims.aux_dims
        for i in range(S):
            a_t = [aux[:, i, d * j:d * (j + 1)] for j in range(4)]
            lg, h1, h2 = ofc.step_logits(W, dims, torch.as_tensor(xs[:, i:i + 1]), m[:, i, :], a_t, h1, h2)
            logits.append(lg)
        ref = torch.stack(logits).transpose(0, 1).numpy()
    for which, nt in [(0, 0)] + [(4, n) for n in nts]:
        v.set_loop_kernel(which)
        v.set_mma_fold_tile(nt)
        t0 = time.time()
        out = v.forward_logits(xs, mels)
        dt = time.time() - t0
        err_t = np.abs(out.astype(np.float64) - ref).max(axis=(0, 2)) / np.abs(ref).max()
        emit(test='teacher_forced_logits', style=style, n_seq=n_seq, S=S, kernel=v.loop_kernel_name(n_seq), fold_tile=nt,
             rel_err=relerr(out, ref), rel_err_first_100=float(err_t[:100].max()), rel_err_last_100=float(err_t[-100:].max()),
             finite=bool(np.all(np.isfinite(out))), seconds=round(dt, 3))
    v.set_loop_kernel(0)
    v.set_mma_fold_tile(0)


def free_running(v, dims, sd, W, T, target, overlap, style, nts):
    mel = synth.synth_mel(T, seed=1234)
    B = binding.num_folds(T * dims.hop_length, target, overlap)
    S = target + 2 * overlap
    noise = synth.synth_noise_mol(S, B, 4321)
    with torch.no_grad():
        m, aux, _ = ofc.conditioning(W, dims, mel, True, target, overlap)
        t0 = time.time()
        res = ofc.run_loop(W, dims, m, aux, noise)
        cpu_s = time.time() - t0
    ref, rpk = res['samples'].numpy(), res['picks'].numpy().T
    for which, nt in [(0, 0)] + [(4, n) for n in nts]:
        v.set_loop_kernel(which)
        v.set_mma_fold_tile(nt)
        v.stage([mel[0]], True, target, overlap, False, noise=noise)
        v.run_staged()
        v.fetch()
        got = v.fetch_samples()
        tm = v.last_timing()
        diff = np.abs(got.astype(np.float64) - ref)
        big = np.argwhere(diff.max(axis=0) > 1e-3)
        emit(test='free_running', style=style, folds=B, steps=S, kernel=v.loop_kernel_name(B), fold_tile=nt,
             max_abs_diff=float(diff.max()), mean_abs_diff=float(diff.mean()), frac_within_1e_4=float((diff <= 1e-4).mean()),
             first_step_diff_gt_1e_3=int(big[0, 0]) if len(big) else None, finite=bool(np.all(np.isfinite(got))),
             us_per_step=1000 * tm['ms_loop'] / S, oracle_cpu_s=round(cpu_s, 2))
    v.set_loop_kernel(0)
    v.set_mma_fold_tile(0)


def timing(v, dims, n_utt, nts, reps=1):
    mels = [synth.synth_mel(801, seed=1234 + u)[0] for u in range(n_utt)]
    for which, nt in [(4, n) for n in nts]:
        v.set_loop_kernel(which)
        v.set_mma_fold_tile(nt)
        v.stage(mels, True, 11000, 550, False)
        best = None
        for r in range(reps):
            v.run_staged(seed=r)
            waves = v.fetch()
            tm = v.last_timing()
            best = tm if best is None or tm['ms_loop'] < best['ms_loop'] else best
        ok = all(np.all(np.isfinite(w)) and np.abs(w).max() <= 1.5 for w in waves)
        emit(test='timing', utterances=n_utt, folds=best['folds'], steps=best['steps'], kernel=v.loop_kernel_name(best['folds']),
             fold_tile=nt, ms_loop=best['ms_loop'], ms_cond=best['ms_cond'], ms_tail=best['ms_tail'],
             us_per_step=1000 * best['ms_loop'] / best['steps'],
             fold_steps_per_s=best['folds'] * best['steps'] / (best['ms_loop'] * 1e-3),
             samples_per_s=n_utt * 220000 / ((best['ms_loop'] + best['ms_cond'] + best['ms_tail']) * 1e-3), sane=ok)
    v.set_loop_kernel(0)
    v.set_mma_fold_tile(0)


def main():
    what = sys.argv[1:] or ['parity', 'timing']
    dims = VocoderDims(mode='MOL')
    for style in (['trained', 'init'] if 'parity' in what else ['init']):
        sd = synth.synth_state_dict(dims, seed=0, style=style)
        W = ofc.to_torch(sd)
        v = binding.Vocoder(dims, device=0)
        v.load_weights(sd)
        if 'parity' in what:
            teacher_forced(v, dims, W, 2, 2 * dims.hop_length, style, [16])
            if style == 'trained':
                teacher_forced(v, dims, W, 64, 4 * dims.hop_length, style, [16, 64])
                teacher_forced(v, dims, W, 6, 12 * dims.hop_length, style, [16])        # 3 300 recurrent steps
            free_running(v, dims, sd, W, 59, 200, 50, style, [16, 48])                    # 65 folds x 300 steps
            if style == 'trained':
                free_running(v, dims, sd, W, 119, 200, 50, style, [0])                    # 131 folds
                free_running(v, dims, sd, W, 582, 200, 50, style, [0, 128])               # 640 folds
        if 'timing' in what and style == 'init':
            timing(v, dims, 3, [0])
            timing(v, dims, 8, [0, 32, 64])
            timing(v, dims, 32, [0, 96, 128], reps=2)
        v.close()


if __name__ == '__main__':
    main()
