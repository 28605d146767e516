"""CPU: the oracle reproduces the REFERENCE's outputs stored in tests/golden/*.npz bit-for-bit.

The fixtures were produced by tests/golden/make_golden.py running /root/reference itself in the
build container; this is what pins the oracle on machines where the reference is absent.
"""
import numpy as np
import pytest
import torch

from oracle import deepmind as odm
from oracle import fatchord as ofc
from wavernn_b200 import synth

from common import CASES, case_inputs, golden


@pytest.mark.parametrize('name', [n for n in CASES if n.startswith('tiny')] + ['full_mol_batched'])
def test_generate_matches_reference_fixture(name):
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    g = golden(name)
    W = ofc.to_torch(sd)
    wave = ofc.generate(W, dims, mel, batched, target, overlap, mu_law, noise)
    assert wave.dtype == np.float64 and wave.shape == ((T - 1) * dims.hop_length,)
    assert np.array_equal(wave, g['wave'])


@pytest.mark.parametrize('name', ['tiny_mol_batched', 'tiny_raw_unbatched', 'full_mol_unbatched'])
def test_upsample_and_forward_match_reference_fixture(name):
    dims, sd, mel, _, (T, *_r) = case_inputs(name)
    g = golden(name)
    W = ofc.to_torch(sd)
    with torch.no_grad():
        mp = ofc.pad_time(torch.as_tensor(mel).transpose(1, 2), dims.pad).transpose(1, 2)
        m, aux = ofc.upsample(W, dims, mp)
    idx = g['up_idx']
    assert np.array_equal(m[0, idx].numpy(), g['up_mels'])
    assert np.array_equal(aux[0, idx].numpy(), g['up_aux'])
    assert m.double().sum().item() == g['up_mels_sum']
    # teacher-forced logits: reference forward() uses nn.GRU, the oracle a GRUCell loop (F9)
    Sf = 4 * dims.hop_length
    gen = torch.Generator().manual_seed(int(g['fwd_seed']))
    xs = torch.rand((2, Sf), generator=gen) * 2 - 1
    mels_f = torch.rand((2, dims.feat_dims, Sf // dims.hop_length + 2 * dims.pad), generator=gen)
    lg = ofc.forward_logits(W, dims, xs.numpy(), mels_f.numpy()).numpy()[:, ::(int(g['fwd_stride']) if 'fwd_stride' in g.files else 1)]
    ref = g['fwd_logits']
    assert np.abs(lg - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max())


def test_components_match_reference_fixture():
    g = golden('components')
    x = torch.arange(1, 11, dtype=torch.float32).view(1, 10, 1)
    f = ofc.fold_with_overlap(x, 2, 1)           # docstring example, fatchord_version.py:312-316
    assert f[:, :, 0].tolist() == [[1, 2, 3, 4], [4, 5, 6, 7], [7, 8, 9, 10]]
    assert np.array_equal(f.numpy(), g['fold_doc'])
    gen = torch.Generator().manual_seed(7)
    for L, tgt, ov in [(1000, 100, 20), (1220, 100, 20), (620, 100, 20), (97, 10, 3), (4400, 1100, 55)]:
        x = torch.rand((1, L, 5), generator=gen)
        f = ofc.fold_with_overlap(x, tgt, ov)
        assert list(f.shape) == list(g[f'fold_{L}_{tgt}_{ov}_shape'])
        assert f.double().sum().item() == g[f'fold_{L}_{tgt}_{ov}_sum']
    rs = np.random.RandomState(5)
    for B, tgt, ov in [(3, 100, 20), (1, 50, 7), (20, 11000, 550), (4, 30, 1)]:
        y = rs.uniform(-1, 1, size=(B, tgt + 2 * ov))
        o = ofc.xfade_and_unfold(y.copy(), tgt, ov)
        r = g[f'unfold_{B}_{tgt}_{ov}']
        assert np.array_equal(o if o.size < 5000 else o[::37], r)
    for n in (512, 64):
        y = (2 * torch.arange(n).float() / (n - 1.) - 1.).numpy().astype(np.float64)
        assert np.array_equal(ofc.decode_mu_law(y, n, False), g[f'mulaw_{n}'])
    o, _ = ofc.mol_sample(torch.from_numpy(g['mol_logits']), torch.from_numpy(g['mol_u1']), torch.from_numpy(g['mol_u2']))
    assert np.array_equal(o.numpy(), g['mol_out'])


def test_deepmind_matches_reference_fixture():
    g = golden('deepmind')
    for name, hs, S in [('dm_small', 64, 300), ('dm_full', 896, 400)]:
        W = ofc.to_torch(synth.synth_dm_state_dict(hs, 256, seed=0))
        qc = synth.synth_noise_exp(S, 1, 256, seed=4321)
        qf = synth.synth_noise_exp(S, 1, 256, seed=4322)
        out, c, f = odm.generate(W, S, qc, qf)
        assert np.array_equal(out, g[name + '_output'])
        assert np.array_equal(c, g[name + '_coarse']) and np.array_equal(f, g[name + '_fine'])
        assert out.dtype == np.int64 and np.array_equal(out, c * 256 + f - 32768)
