"""Text -> Tacotron decoder loop -> WaveRNN, as `gen_tacotron.py:142-168` chains them, both loops on the library.

The reference's driver does, per sentence: `_, m, attention = tts_model.generate(x)`; `m = (m + 4) / 8`; clip to [0, 1];
`voc_model.generate(torch.tensor(m).unsqueeze(0), save_path, batched, target, overlap, mu_law)`.  Here the TTS model is
the stand-in object of tests/taco_standin.py with `wavernn_b200.tacotron.generate` bound as its method, the vocoder the
drop-in `wavernn_b200.fatchord_version.WaveRNN`; what is checked is the hand-over between the two paths (shapes, value
range, the (T - 1) * hop contract, reproducibility under a seed) and that a stop decided by the decoder kernel carries
through to the waveform length.
"""
import contextlib
import io
import types

import numpy as np
import pytest

from wavernn_b200 import synth
from wavernn_b200.config import VocoderDims

pytestmark = pytest.mark.gpu


def _models(stop_threshold=-3.4):
    import torch
    from taco_standin import build_stand_in
    import wavernn_b200.tacotron as bt
    from wavernn_b200.fatchord_version import WaveRNN
    tts = build_stand_in(synth.synth_taco_state_dict(seed=0), stop_threshold=stop_threshold).cuda()
    tts.generate = types.MethodType(bt.generate, tts)                  # INTEGRATION.md, Tacotron section
    dims = VocoderDims(mode='MOL')
    with contextlib.redirect_stdout(io.StringIO()):
        voc = WaveRNN(**dims.ctor_kwargs()).cuda()
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    voc.load_state_dict({k: torch.from_numpy(np.array(v)) for k, v in sd.items()}, strict=True)
    return tts, voc, dims


def _speak(tts, voc, x, steps, path, seed):
    import torch
    _, m, attention = tts.generate(x, steps=steps)                     # gen_tacotron.py:145
    m = (m + 4) / 8                                                    # :147-148
    np.clip(m, 0, 1, out=m)
    m = torch.tensor(m).unsqueeze(0)
    with contextlib.redirect_stdout(io.StringIO()):
        wav = voc.generate(m, path, True, 1100, 550, False, seed=seed)         # :168 (short target: a short utterance)
    return m, attention, wav


def test_text_to_waveform(tmp_path):
    tts, voc, dims = _models()
    x = [5 + (i * 11) % 60 for i in range(40)]
    m, attention, wav = _speak(tts, voc, x, 120, tmp_path / 'a.wav', seed=7)
    assert m.shape == (1, 80, 120) and attention.shape == (60, 40)
    assert float(m.min()) >= 0.0 and float(m.max()) <= 1.0
    assert wav.shape == ((120 - 1) * dims.hop_length,) and wav.dtype == np.float64
    assert np.all(np.isfinite(wav)) and np.abs(wav).max() <= 1.5          # samples clip at +-1, the cross-fade of two folds can reach sqrt(2)
    assert (tmp_path / 'a.wav').stat().st_size > 44
    # the same text and seed again: both loops are deterministic
    m2, _, wav2 = _speak(tts, voc, x, 120, tmp_path / 'b.wav', seed=7)
    assert np.array_equal(m.numpy(), m2.numpy()) and np.array_equal(wav, wav2)
    assert tts.training and voc.training                                # both generate() calls end with train()


def test_stop_rule_sets_the_waveform_length(tmp_path):
    """Every frame counts as silent (threshold +1e9): the decoder kernel stops at the first t > 10.  With r = 7 that is
    t = 14, three steps = 21 frames, and the vocoder produces (21 - 1) * hop samples.  With r = 2 it is seven steps = 14
    frames, fewer than the 20 hops `generate` fades out over (fatchord_version.py:256-258): the reference fails there with
    a broadcast ValueError, the library with a ValueError that says so."""
    tts, voc, dims = _models(stop_threshold=1e9)
    x = [3 + (i * 7) % 60 for i in range(25)]
    tts.r = 7
    m, attention, wav = _speak(tts, voc, x, 400, tmp_path / 'c.wav', seed=1)
    assert m.shape == (1, 80, 21) and attention.shape == (3, 25)
    assert wav.shape == (20 * dims.hop_length,)
    tts.r = 2
    with pytest.raises(ValueError, match='too short'):
        _speak(tts, voc, x, 400, tmp_path / 'd.wav', seed=1)
