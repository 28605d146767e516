"""CPU: the pruning tool (wavernn_b200/prune.py) against the reference notebook's own code (tests/golden/prune.npz,
made by tests/golden/make_prune_golden.py from `notebooks/Pruning - Scratchpad.ipynb` cells 1-4) and against the numpy
statement of SURVEY.md row S that the GPU test holds `wrnn_prune_16x1` to."""
import numpy as np
import pytest

from wavernn_b200 import prune

from common import golden


def test_schedule_matches_the_notebook():
    g = golden('prune')
    z = [prune.sparsity_at(int(t), int(g['t0']), int(g['S']), float(g['Z'])) for t in g['steps']]
    # the notebook evaluates the expression on a float32 step counter (np_now(t)[0]); here t is an integer: 1e-7 apart
    assert np.abs(np.array(z, dtype=np.float64) - g['z']).max() <= 2e-7
    assert z[0] == 0 and z[-1] == float(g['Z'])                           # clamped on both sides
    assert all(b >= a for a, b in zip(z, z[1:]))


@pytest.mark.parametrize('z', [0.5, 0.9])
def test_elementwise_mask_matches_the_notebook(z):
    g = golden('prune')
    assert np.array_equal(prune.elementwise_mask(g['w0_rnn.weight_ih_l0'], 3, z), g[f'mask_ih_{z}'])
    assert np.array_equal(prune.elementwise_mask(g['w0_rnn.weight_hh_l0'], 3, z), g[f'mask_hh_{z}'])
    assert np.array_equal(prune.elementwise_mask(g['w0_fc.weight'], 1, z), g[f'mask_fc_{z}'])


def test_driver_updates_and_applies_like_the_notebook():
    """Same steps at which masks change (every `prune_every` after t0), same final sparsity, with the elementwise mask."""
    import torch
    g = golden('prune')
    state = {'rnn.weight_ih_l0': torch.tensor(g['w0_rnn.weight_ih_l0']), 'rnn.weight_hh_l0': torch.tensor(g['w0_rnn.weight_hh_l0'])}
    fc = {'fc.weight': torch.tensor(g['w0_fc.weight'])}
    p_rnn = prune.Pruner(int(g['t0']), int(g['S']), float(g['Z']), prune_every=int(g['every']), keys=list(state), block=False)
    p_fc = prune.Pruner(int(g['t0']), int(g['S']), float(g['Z']), prune_every=int(g['every']), keys=list(fc), n_gates=1, block=False)
    torch.manual_seed(8)
    updated = []
    for t in range(600):
        for w in (state['rnn.weight_ih_l0'], state['rnn.weight_hh_l0'], fc['fc.weight']):
            w.add_(0.01 * torch.randn_like(w))
        before = {k: m.copy() for k, m in p_rnn.masks.items()}
        p_rnn.prune(state, torch.tensor([float(t)]))
        p_fc.prune(fc, t)
        if before.keys() != p_rnn.masks.keys() or any(not np.array_equal(before[k], p_rnn.masks[k]) for k in before):
            updated.append(t)
    assert np.array_equal(np.array(updated), g['updated_steps'])
    assert abs(p_rnn.z - float(g['final_z'])) <= 2e-7
    assert float((state['rnn.weight_hh_l0'] == 0).float().mean()) == float(g['final_zero_frac_hh'])
    assert p_rnn.num_pruned + p_fc.num_pruned == int(g['num_pruned'])
    assert p_rnn.total_params + p_fc.total_params == int(g['total_params'])


def _mask_16x1(w, H, keep):          # the statement tests/test_variants_gpu.py holds wrnn_prune_16x1 to
    w = w.copy()
    K = w.shape[1]
    for g in range(3):
        blk = w[g * H:(g + 1) * H].reshape(H // 16, 16, K)
        norm = np.abs(blk.astype(np.float64)).sum(axis=1).reshape(-1)
        n_keep = min(norm.size, max(1, int(np.floor(keep * norm.size + 0.5))))
        order = np.lexsort((np.arange(norm.size), -norm))
        m = np.zeros(norm.size, dtype=bool)
        m[order[:n_keep]] = True
        blk *= m.reshape(H // 16, 1, K)
    return w


def test_block_mask_is_row_S_and_the_schedule_reaches_the_target():
    rng = np.random.default_rng(0)
    H, K = 64, 40
    w = rng.standard_normal((3 * H, K)).astype(np.float32)
    for keep in (0.1, 0.33, 1.0):
        m = prune.block_mask_16x1(w, 3, 1.0 - keep)
        assert np.array_equal(w * m, _mask_16x1(w, H, keep))
        per_block = m.reshape(3, H // 16, 16, K)
        assert np.all(per_block.min(axis=2) == per_block.max(axis=2))            # whole 16x1 blocks
    state = {k: rng.standard_normal((3 * H, K)).astype(np.float32) for k in prune.GRU_KEYS}
    p = prune.Pruner(start_prune=10, prune_steps=400, target_sparsity=0.9, prune_every=20)
    for t in range(500):
        p.prune(state, t)
    assert p.z == 0.9
    for k in prune.GRU_KEYS:
        kept = (np.abs(state[k]).reshape(3, H // 16, 16, K).sum(axis=2) > 0).mean()
        assert abs(kept - 0.1) < 0.01
    q = prune.Pruner(10, 400, 0.9, prune_rnn_input=False)
    assert all('weight_ih' not in k for k in q.keys) and len(q.keys) == 2
