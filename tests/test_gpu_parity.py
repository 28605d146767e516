"""GPU: the CUDA path (through the C ABI) against the oracle and the reference's golden fixtures.

Tolerances (north_star): fold / unfold / mu-law decode bit-exact; conditioning <= 1e-5 relative;
teacher-forced logits <= 1e-3 relative (fp32 accumulate); free-running waveforms under identical
injected noise: identical discrete decisions (mixture / class picks), waveform within 2e-4 absolute.
"""
import numpy as np
import pytest
import torch

from oracle import fatchord as ofc
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims

from common import CASES, TINY, case_inputs, golden

pytestmark = pytest.mark.gpu

_VOC = {}


def vocoder(dims, sd_key, sd):
    k = (dims, sd_key)
    if k not in _VOC:
        v = binding.Vocoder(dims, device=0)
        v.load_weights(sd)
        _VOC[k] = v
    return _VOC[k]


def relerr(a, b):
    return np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(np.abs(b).max(), 1e-30)


# ------------------------------------------------------------------ bit-exact components
@pytest.mark.parametrize('L,tgt,ov,F', [(10, 2, 1, 1), (1000, 100, 20, 5), (1220, 100, 20, 5), (620, 100, 20, 8),
                                         (97, 10, 3, 3), (4400, 1100, 55, 16), (220275, 11000, 550, 80),
                                         (220000, 11000, 550, 128), (31, 10, 1, 4)])
def test_fold_bit_exact(L, tgt, ov, F):
    dims = VocoderDims(mode='MOL', **TINY)
    v = vocoder(dims, 'init', synth.synth_state_dict(dims, 0, 'init'))
    x = torch.rand((1, L, F), generator=torch.Generator().manual_seed(L))
    ref = ofc.fold_with_overlap(x, tgt, ov).numpy()
    out = v.fold(x[0].numpy(), tgt, ov)
    assert out.shape == ref.shape and np.array_equal(out, ref)


def test_fold_docstring_and_golden():
    dims = VocoderDims(mode='MOL', **TINY)
    v = vocoder(dims, 'init', synth.synth_state_dict(dims, 0, 'init'))
    x = np.arange(1, 11, dtype=np.float32).reshape(10, 1)
    assert np.array_equal(v.fold(x, 2, 1), golden('components')['fold_doc'])
    with pytest.raises(ValueError):
        v.fold(x[:1], 2, 1)


@pytest.mark.parametrize('B,tgt,ov', [(3, 100, 20), (1, 50, 7), (20, 11000, 550), (4, 30, 1), (19, 11000, 550)])
def test_unfold_bit_exact(B, tgt, ov):
    dims = VocoderDims(mode='MOL', **TINY)
    v = vocoder(dims, 'init', synth.synth_state_dict(dims, 0, 'init'))
    y = np.random.RandomState(5 + B).uniform(-1, 1, size=(B, tgt + 2 * ov))
    ref = ofc.xfade_and_unfold(y.copy(), tgt, ov)
    out = v.unfold(y, tgt, ov)
    assert np.array_equal(out, ref)


def test_unfold_golden():
    dims = VocoderDims(mode='MOL', **TINY)
    v = vocoder(dims, 'init', synth.synth_state_dict(dims, 0, 'init'))
    g = golden('components')
    rs = np.random.RandomState(5)
    for B, tgt, ov in [(3, 100, 20), (1, 50, 7), (20, 11000, 550), (4, 30, 1)]:
        y = rs.uniform(-1, 1, size=(B, tgt + 2 * ov))
        o = v.unfold(y, tgt, ov)
        assert np.array_equal(o if o.size < 5000 else o[::37], g[f'unfold_{B}_{tgt}_{ov}'])


@pytest.mark.parametrize('n', [512, 64, 1024, 256])
def test_mulaw_decode_bit_exact(n):
    dims = VocoderDims(mode='MOL', **TINY)
    v = vocoder(dims, 'init', synth.synth_state_dict(dims, 0, 'init'))
    y32 = (2 * torch.arange(n).float() / (n - 1.) - 1.).numpy()
    ref = ofc.decode_mu_law(y32.astype(np.float64), n, False)
    out = v.decode_mu_law(np.tile(y32, 3), n)
    assert np.array_equal(out, np.tile(ref, 3))
    if n in (512, 64):
        assert np.array_equal(out[:n], golden('components')[f'mulaw_{n}'])


# ------------------------------------------------------------------ conditioning
@pytest.mark.parametrize('name', ['tiny_mol_batched', 'tiny_raw_unbatched', 'full_mol_unbatched', 'full_mol_batched'])
def test_upsample_vs_oracle_and_golden(name):
    dims, sd, mel, _, (T, *_r) = case_inputs(name)
    v = vocoder(dims, CASES[name][1], sd)
    m, aux = v.upsample(mel[0])
    W = ofc.to_torch(sd)
    with torch.no_grad():
        mp = ofc.pad_time(torch.as_tensor(mel).transpose(1, 2), dims.pad).transpose(1, 2)
        om, oaux = ofc.upsample(W, dims, mp)
    assert m.shape == tuple(om.shape[1:]) and aux.shape == tuple(oaux.shape[1:])
    assert relerr(m, om[0].numpy()) <= 1e-5
    assert relerr(aux, oaux[0].numpy()) <= 1e-5
    g = golden(name)
    assert relerr(m[g['up_idx']], g['up_mels']) <= 1e-5
    assert relerr(aux[g['up_idx']], g['up_aux']) <= 1e-5


# ------------------------------------------------------------------ teacher-forced logits
@pytest.mark.parametrize('name', ['tiny_mol_batched', 'tiny_raw_batched_mulaw', 'full_mol_batched', 'full_raw_batched_mulaw'])
def test_forward_logits_vs_reference_fixture(name):
    dims, sd, *_ = case_inputs(name)
    v = vocoder(dims, CASES[name][1], sd)
    g = golden(name)
    Sf = 4 * dims.hop_length
    gen = torch.Generator().manual_seed(int(g['fwd_seed']))
    xs = torch.rand((2, Sf), generator=gen) * 2 - 1
    mels_f = torch.rand((2, dims.feat_dims, Sf // dims.hop_length + 2 * dims.pad), generator=gen)
    out = v.forward_logits(xs.numpy(), mels_f.numpy())
    stride = int(g['fwd_stride']) if 'fwd_stride' in g.files else 1
    ref = g['fwd_logits']
    assert out[:, ::stride].shape == ref.shape
    assert relerr(out[:, ::stride], ref) <= 1e-3      # north_star tolerance; fp32 path is ~1e-6
    assert relerr(out[:, ::stride], ref) <= 2e-5


# ------------------------------------------------------------------ free-running generation
@pytest.mark.parametrize('name', list(CASES))
def test_generate_vs_reference_fixture(name):
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    v = vocoder(dims, CASES[name][1], sd)
    wave = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    ref = golden(name)['wave']
    assert wave.dtype == np.float64 and wave.shape == ref.shape
    t = v.last_timing()
    assert (t['folds'], t['steps']) == (B, S)
    if dims.mode == 'RAW':
        # discrete samples: every class pick must agree, so the float64 tail is bit-exact
        assert np.array_equal(wave, ref)
    else:
        assert np.abs(wave - ref).max() <= 2e-4


@pytest.mark.parametrize('name', ['tiny_mol_batched', 'tiny_raw_batched_mulaw', 'full_mol_batched'])
def test_generate_picks_vs_oracle(name):
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    v = vocoder(dims, CASES[name][1], sd)
    v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)
    samples = v.fetch_samples()
    W = ofc.to_torch(sd)
    with torch.no_grad():
        m, aux, _ = ofc.conditioning(W, dims, mel, batched, target, overlap)
        res = ofc.run_loop(W, dims, m, aux, noise)
    ref = res['samples'].numpy()
    assert samples.shape == ref.shape
    if dims.mode == 'RAW':
        assert np.array_equal(samples, ref)
    else:
        assert np.abs(samples - ref).max() <= 2e-4


def test_philox_generation_is_deterministic_and_bounded():
    dims, sd, mel, _, (T, batched, target, overlap, mu_law, B, S) = case_inputs('tiny_mol_batched')
    v = vocoder(dims, CASES['tiny_mol_batched'][1], sd)
    a = v.generate([mel[0]], batched, target, overlap, mu_law, seed=7)[0]
    b = v.generate([mel[0]], batched, target, overlap, mu_law, seed=7)[0]
    c = v.generate([mel[0]], batched, target, overlap, mu_law, seed=8)[0]
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert np.abs(a).max() <= 1.5 and np.all(np.isfinite(a))


def test_multi_utterance_batch_equals_single_calls():
    dims, sd, mel, _, (T, batched, target, overlap, mu_law, B, S) = case_inputs('tiny_mol_batched')
    v = vocoder(dims, CASES['tiny_mol_batched'][1], sd)
    mel2 = synth.synth_mel(45, seed=99, feat_dims=dims.feat_dims)
    B2 = binding.num_folds(45 * dims.hop_length, target, overlap)
    n1 = synth.synth_noise_mol(S, B, 1)
    n2 = synth.synth_noise_mol(S, B2, 2)
    w1 = v.generate([mel[0]], True, target, overlap, False, noise=n1)[0]
    w2 = v.generate([mel2[0]], True, target, overlap, False, noise=n2)[0]
    both = (np.concatenate([n1[0], n2[0]], axis=1), np.concatenate([n1[1], n2[1]], axis=1))
    o1, o2 = v.generate([mel[0], mel2[0]], True, target, overlap, False, noise=both)
    assert np.array_equal(o1, w1) and np.array_equal(o2, w2)


def test_errors():
    dims = VocoderDims(mode='MOL', **TINY)
    v = vocoder(dims, 'init', synth.synth_state_dict(dims, 0, 'init'))
    with pytest.raises(ValueError):        # T < 21: the reference fails in the fade-out broadcast
        v.generate([synth.synth_mel(10, feat_dims=16)[0]], False, 100, 20, False)
    with pytest.raises(ValueError):
        v.generate([np.zeros((3, 40), np.float32)], False, 100, 20, False)
    fresh = binding.Vocoder(dims, device=0)
    with pytest.raises(RuntimeError):      # weights not loaded
        fresh.generate([synth.synth_mel(40, feat_dims=16)[0]], False, 100, 20, False)
    with pytest.raises(RuntimeError):
        VocoderDims(mode='XYZ').n_classes
