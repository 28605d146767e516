"""Call compatibility (SURVEY.md section 4 viii): the reference's own `gen_wavernn.gen_from_file`
(`gen_wavernn.py:38-65`), un-modified, drives the drop-in `wavernn_b200.fatchord_version.WaveRNN`.

CPU variant: the C-ABI vocoder behind the module is replaced by a stub that answers with the oracle's waveform, so
what is exercised is the boundary itself: argument types the reference passes (a `(1, 80, T)` tensor, a `Path` save
target, `hp.mu_law`), the return type, the wav side effect, and `train()` restored at exit.  GPU variant: the same
call on the real library.  Needs the reference tree (`/root/reference` in the build container, or the copies
`oracle/stage_ref.py` staged at build time on the GPU box); skipped when neither is there.
"""
import contextlib
import io
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import fatchord as ofc
from oracle import ref_shim
from wavernn_b200 import synth, wavio
from wavernn_b200.config import VocoderDims
from wavernn_b200.fatchord_version import WaveRNN

needs_ref = pytest.mark.skipif(not ref_shim.available(), reason='reference tree not present / not staged')


def _model(dims, sd):
    with contextlib.redirect_stdout(io.StringIO()):
        model = WaveRNN(**dims.ctor_kwargs())
    model.load_state_dict({k: torch.from_numpy(np.array(v)) for k, v in sd.items()}, strict=True)
    return model


class _StubVocoder:
    """Stands in for binding.Vocoder on a machine without a GPU: same generate() / last_timing() surface."""

    def __init__(self, dims, sd):
        self.dims, self.W = dims, ofc.to_torch(sd)

    def generate(self, mels, batched, target, overlap, mu_law, noise=None, seed=0, stream=None):
        assert len(mels) == 1 and mels[0].dtype == np.float32 and mels[0].shape[0] == self.dims.feat_dims
        mel = torch.from_numpy(mels[0])[None]
        L = mel.shape[-1] * self.dims.hop_length
        B = ofc.num_folds_for(L, target, overlap) if batched else 1
        S = target + 2 * overlap if batched else L
        self._t = dict(ms_cond=0., ms_loop=1., ms_tail=0., steps=S, folds=B, launches=0)
        return [ofc.generate(self.W, self.dims, mel, batched, target, overlap, mu_law, synth.synth_noise_mol(S, B, seed % 1000))]

    def last_timing(self):
        return self._t


@needs_ref
def test_reference_gen_from_file_drives_the_drop_in_class(tmp_path, monkeypatch):
    gw = ref_shim.load_gen_wavernn()
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    model = _model(dims, sd)
    stub = _StubVocoder(dims, sd)
    monkeypatch.setattr(model, '_vocoder', lambda: stub)
    monkeypatch.setattr(torch.cuda, 'current_stream', lambda *a, **k: type('S', (), {'cuda_stream': None})())
    T = 24
    mel = synth.synth_mel(T, seed=3)[0]
    np.save(tmp_path / 'utt.npy', mel)
    model.train()
    with contextlib.redirect_stdout(io.StringIO()):
        gw.gen_from_file(model, Path(tmp_path / 'utt.npy'), Path(tmp_path), True, 2000, 550)
    assert model.training                                    # generate() sets eval() and restores train() (:170,262)
    k = model.get_step() // 1000
    wav = tmp_path / f'__utt__{k}k_steps_gen_batched_target2000_overlap550.wav'
    assert wav.is_file()
    data, sr = wavio.load_wav(wav)
    assert sr == dims.sample_rate and data.dtype == np.float32 and data.shape == ((T - 1) * dims.hop_length,)
    assert np.all(np.isfinite(data)) and np.abs(data).max() <= 1.5
    # the reference validates the .npy itself: wrong rank -> its own ValueError path (it formats `wav.shape`, a latent bug)
    np.save(tmp_path / 'bad.npy', np.zeros((3, 4, 5), np.float32))
    with pytest.raises((ValueError, UnboundLocalError, NameError)):
        gw.gen_from_file(model, Path(tmp_path / 'bad.npy'), Path(tmp_path), True, 2000, 550)


@needs_ref
@pytest.mark.gpu
def test_reference_gen_from_file_on_the_gpu_library(tmp_path):
    gw = ref_shim.load_gen_wavernn()
    dims = VocoderDims(mode='MOL')
    model = _model(dims, synth.synth_state_dict(dims, seed=0, style='init')).cuda()
    T = 30
    np.save(tmp_path / 'utt.npy', synth.synth_mel(T, seed=4)[0])
    with contextlib.redirect_stdout(io.StringIO()):
        gw.gen_from_file(model, Path(tmp_path / 'utt.npy'), Path(tmp_path), True, 1100, 550)
        gw.gen_from_file(model, Path(tmp_path / 'utt.npy'), Path(tmp_path), False, 1100, 550)
    k = model.get_step() // 1000
    for name in (f'__utt__{k}k_steps_gen_batched_target1100_overlap550.wav', f'__utt__{k}k_steps_gen_NOT_BATCHED.wav'):
        data, sr = wavio.load_wav(tmp_path / name)
        assert sr == dims.sample_rate and data.shape == ((T - 1) * dims.hop_length,) and np.all(np.isfinite(data))
    assert model.training
