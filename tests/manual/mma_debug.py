"""Bring-up: state of the tensor-core loop kernel after n steps (WRNN_MMA_STEPS=n) against the oracle, per phase.

  WRNN_MMA_STEPS=1 python tests/manual/mma_debug.py [folds-target]
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import fatchord as ofc          # noqa: E402
from wavernn_b200 import binding, synth     # noqa: E402
from wavernn_b200.config import VocoderDims  # noqa: E402


def main():
    n = int(os.environ.get('WRNN_MMA_STEPS', '1'))
    T = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='trained')
    W = ofc.to_torch(sd)
    mel = synth.synth_mel(T, seed=1234)
    target, overlap = 200, 50
    with torch.no_grad():
        m, aux, _ = ofc.conditioning(W, dims, mel, True, target, overlap)
    B, S = int(m.shape[0]), int(m.shape[1])
    noise = synth.synth_noise_mol(S, B, 4321)
    d = dims.aux_dims
    h1 = torch.zeros(B, 512)
    h2 = torch.zeros(B, 512)
    x = torch.zeros(B, 1)
    with torch.no_grad():
        for i in range(n):
            a_t = [aux[:, i, d * j:d * (j + 1)] for j in range(4)]
            xi = F.linear(torch.cat([x, m[:, i, :], a_t[0]], dim=1), W['I.weight'], W['I.bias'])
            h1 = torch.gru_cell(xi, h1, W['rnn1.weight_ih_l0'], W['rnn1.weight_hh_l0'], W['rnn1.bias_ih_l0'], W['rnn1.bias_hh_l0'])
            xx = xi + h1
            h2 = torch.gru_cell(torch.cat([xx, a_t[1]], dim=1), h2, W['rnn2.weight_ih_l0'], W['rnn2.weight_hh_l0'],
                                W['rnn2.bias_ih_l0'], W['rnn2.bias_hh_l0'])
            xx = xx + h2
            y1 = F.relu(F.linear(torch.cat([xx, a_t[2]], dim=1), W['fc1.weight'], W['fc1.bias']))
            y2 = F.relu(F.linear(torch.cat([y1, a_t[3]], dim=1), W['fc2.weight'], W['fc2.bias']))
            logits = F.linear(y2, W['fc3.weight'], W['fc3.bias'])
            s, k = ofc.mol_sample(logits, torch.as_tensor(noise[0][i]), torch.as_tensor(noise[1][i]))
            x = s.reshape(B, 1)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    v.set_loop_kernel(4)
    for nt in (16, 32):
        v.set_mma_fold_tile(nt)
        v.stage([mel[0]], True, target, overlap, False, noise=noise)
        v.run_staged()
        v.fetch()
        got = [v.debug_mma_state(i, B) for i in range(3)]
        smp = v.fetch_samples()
        print(f'--- steps={n} folds={B} fold_tile={nt} kernel={v.loop_kernel_name(B)}')
        for name, g, r in (('h1', got[0], h1.numpy()), ('h2', got[1], h2.numpy()), ('y2', got[2], y2.numpy())):
            e = np.abs(g - r)
            print(f'{name}: max|diff| {e.max():.3e}  max|ref| {np.abs(r).max():.3f}  per 32-unit tile: '
                  + ' '.join(f'{e[:, 32 * i:32 * i + 32].max():.1e}' for i in range(16)))
            print(f'    per fold: ' + ' '.join(f'{e[b].max():.1e}' for b in range(min(B, 20))))
            print(f'    per unit%8: ' + ' '.join(f'{e[:, j::8].max():.1e}' for j in range(8)))
        print('sample(step n-1): got', smp[:4, n - 1], 'ref', s.numpy()[:4])


if __name__ == '__main__':
    main()
