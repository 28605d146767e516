"""melspectrogram (utils/dsp.py:72-75): the oracle's anchors on CPU, the CUDA path against the oracle on the GPU."""
import numpy as np
import pytest

from oracle import melspec as om


def _signals():
    rng = np.random.default_rng(3)
    sr = 22050
    t = np.arange(3 * sr + 137) / sr
    speechy = 0.3 * np.sin(2 * np.pi * 180 * t) * (1 + 0.5 * np.sin(2 * np.pi * 3 * t)) + 0.1 * np.sin(2 * np.pi * 2400 * t) \
        + 0.02 * rng.standard_normal(t.size)
    return {'noise': (0.2 * rng.standard_normal(20000)).astype(np.float32), 'speechy': speechy.astype(np.float32),
            'short': (0.5 * rng.standard_normal(1025 + 7)).astype(np.float32),
            'tone': (0.5 * np.sin(2 * np.pi * 1000 * t[:22050])).astype(np.float32)}


def test_oracle_anchors():
    """Shape as gen_wavernn.py:44-47 consumes it, Slaney normalisation of the basis, a tone in the band holding it."""
    sig = _signals()
    m = om.melspectrogram(sig['tone'])
    assert m.shape == (80, 1 + 22050 // 275) and m.min() >= 0 and m.max() <= 1
    B = om.mel_basis(22050, 2048, 80, 40.0)
    assert B.shape == (80, 1025) and B.min() >= 0
    # area normalisation: every triangle integrates to ~1 over frequency (bin width sr / n_fft)
    area = B.sum(axis=1) * 22050 / 2048
    assert np.all(np.abs(area[5:] - 1) < 0.2)
    centre = np.array([np.argmax(B[i]) for i in range(80)]) * 22050 / 2048
    band = int(np.argmin(np.abs(centre - 1000)))
    assert abs(int(np.argmax(m[:, 40])) - band) <= 1
    # the window: periodic Hann centred in the frame, frame count 1 + n // hop
    D = om.stft(sig['noise'], 2048, 275, 1100)
    assert D.shape == (1025, 1 + 20000 // 275)


@pytest.mark.gpu
@pytest.mark.parametrize('name', ['noise', 'speechy', 'short', 'tone'])
def test_melspectrogram_vs_oracle(name):
    from wavernn_b200 import dsp
    y = _signals()[name]
    got = dsp.melspectrogram(y)
    ref = om.melspectrogram(y)
    assert got.dtype == np.float32 and got.shape == ref.shape
    d = np.abs(got.astype(np.float64) - ref)
    # fp32 direct DFT against the float64 FFT: 1e-4 on the normalised scale (= 0.01 dB)
    assert d.max() <= 1e-4, (name, d.max(), np.unravel_index(d.argmax(), d.shape))


@pytest.mark.gpu
def test_melspectrogram_other_parameters_and_errors():
    from wavernn_b200 import dsp
    y = _signals()['noise']
    got = dsp.melspectrogram(y, n_fft=1024, hop_length=256, win_length=1024, num_mels=40, fmin=0.0, sample_rate=16000)
    ref = om.melspectrogram(y, sample_rate=16000, n_fft=1024, hop_length=256, win_length=1024, num_mels=40, fmin=0.0)
    assert got.shape == ref.shape == (40, 1 + 20000 // 256) and np.abs(got - ref).max() <= 1e-4
    back = dsp.melspectrogram(y)                        # tables rebuilt for the default parameters
    assert np.abs(back - om.melspectrogram(y)).max() <= 1e-4
    with pytest.raises(ValueError):
        dsp.melspectrogram(y[:1000])                    # reflect padding needs more than n_fft / 2 samples
    with pytest.raises(ValueError):
        dsp.melspectrogram(y, n_fft=1000)


@pytest.mark.gpu
def test_wav_in_wav_out_through_the_drop_in():
    """gen_wavernn.py:44-49 for a .wav input: load_wav -> melspectrogram -> generate, all on the library."""
    import os
    import tempfile
    from wavernn_b200 import dsp
    from wavernn_b200.config import VocoderDims
    from wavernn_b200.fatchord_version import WaveRNN
    dims = VocoderDims(mode='MOL')
    y = _signals()['speechy'][:22050]
    with tempfile.TemporaryDirectory() as td:
        src, dst = os.path.join(td, 'in.wav'), os.path.join(td, 'out.wav')
        dsp.save_wav(y, src, 22050)
        wav = dsp.load_wav(src)
        with pytest.raises(ValueError):
            dsp.load_wav(src, sample_rate=16000)
        mel = dsp.melspectrogram(wav)
        assert mel.shape == (80, 1 + len(wav) // 275)
        model = WaveRNN(rnn_dims=dims.rnn_dims, fc_dims=dims.fc_dims, bits=dims.bits, pad=dims.pad, upsample_factors=dims.upsample_factors,
                        feat_dims=dims.feat_dims, compute_dims=dims.compute_dims, res_out_dims=dims.res_out_dims, res_blocks=dims.res_blocks,
                        hop_length=dims.hop_length, sample_rate=dims.sample_rate, mode='MOL').cuda()
        import torch
        out = model.generate(torch.tensor(mel).unsqueeze(0), dst, True, 2000, 200, False)
        assert len(out) == (mel.shape[1] - 1) * 275 and os.path.getsize(dst) > 44
