"""Golden vectors for wavernn_b200/prune.py from the reference's own pruning code.

Runs cells 1-4 of /root/reference/notebooks/Pruning - Scratchpad.ipynb (np_now, clamp, PruneMask, Pruner) unchanged, on a
seeded GRU + Linear pair, and records the schedule and the masks at a few steps.  Run in the build container (the
reference does not exist on the GPU box):  python tests/golden/make_prune_golden.py
"""
import json
import os

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
NB = '/root/reference/notebooks/Pruning - Scratchpad.ipynb'


def main():
    nb = json.load(open(NB))
    ns = {'torch': torch, 'nn': nn}
    for i in (1, 2, 3, 4):
        exec(''.join(nb['cells'][i]['source']), ns)
    torch.manual_seed(7)
    rnn = nn.GRU(24, 32)
    fc = nn.Linear(32, 32)
    layers = [rnn, fc]
    t0, S, Z, every = 10, 2000, 0.9375, 50
    pruner = ns['Pruner'](layers, t0, S, Z, prune_rnn_input=True, prune_every=every)
    w0 = {'rnn.weight_ih_l0': rnn.weight_ih_l0.detach().numpy().copy(), 'rnn.weight_hh_l0': rnn.weight_hh_l0.detach().numpy().copy(),
          'fc.weight': fc.weight.detach().numpy().copy()}
    steps = np.array([0, 5, 10, 11, 50, 100, 450, 1000, 1999, 2010, 2500, 5000])
    zs = []
    for t in steps:
        pruner.update_sparsity(torch.tensor([float(t)]))
        zs.append(pruner.z)
    out = {'steps': steps, 'z': np.array(zs, dtype=np.float64), 't0': t0, 'S': S, 'Z': Z, 'every': every}
    out.update({'w0_' + k: v for k, v in w0.items()})
    # masks of the notebook at z = 0.5 and z = 0.9 on the untouched weights
    for z in (0.5, 0.9):
        m = ns['PruneMask'](rnn, True)
        m.update_mask(rnn, z)
        out[f'mask_ih_{z}'] = m.mask[0].numpy()
        out[f'mask_hh_{z}'] = m.mask[1].numpy()
        mf = ns['PruneMask'](fc, True)
        mf.update_mask(fc, z)
        out[f'mask_fc_{z}'] = mf.mask[0].numpy()
    # the driver: 600 steps of prune() with a fake optimiser step in between; record when masks were updated and the
    # sparsity of the weights at the end
    torch.manual_seed(8)
    updated = []
    for t in range(600):
        with torch.no_grad():
            for p in list(rnn.parameters())[:2] + [fc.weight]:
                p.add_(0.01 * torch.randn_like(p))          # "training"
        before = [m.mask[0].clone() for m in pruner.masks]
        pruner.prune(layers, torch.tensor([float(t)]))
        if any(not torch.equal(a, m.mask[0]) for a, m in zip(before, pruner.masks)):
            updated.append(t)
    out['updated_steps'] = np.array(updated)
    out['final_z'] = pruner.z
    out['final_zero_frac_hh'] = float((rnn.weight_hh_l0 == 0).float().mean())
    out['num_pruned'] = pruner.num_pruned
    out['total_params'] = pruner.total_params
    np.savez_compressed(os.path.join(HERE, 'prune.npz'), **out)
    print('wrote prune.npz:', {k: (v.shape if hasattr(v, 'shape') else v) for k, v in out.items() if not k.startswith('w0_')})


if __name__ == '__main__':
    main()
