"""ORACLE (test infrastructure, never the product path): `melspectrogram` of the reference, utils/dsp.py:72-75.

PARITY UNPINNED for this function: the reference delegates the two numeric steps to librosa 0.6.3 (requirements.txt),
which is not installed in this image, and holds no fixture for them.  What follows restates librosa 0.6.3's published
algorithm for the calls the reference makes, in float64:

  stft (utils/dsp.py:78-81 -> librosa.core.spectrum.stft): window = scipy.signal.get_window('hann', win_length,
      fftbins=True), zero-padded to n_fft in the centre (util.pad_center); y reflect-padded by n_fft // 2 on both sides
      (np.pad mode='reflect'); frames of n_fft every hop_length: 1 + len(y) // hop_length of them; FFT, bins 0 .. n_fft / 2.
  linear_to_mel (utils/dsp.py:42-44 -> librosa.feature.melspectrogram(S=...)): np.dot(filters.mel(sr, n_fft, n_mels,
      fmin), S) with htk=False (Slaney scale: linear below 1 kHz, log above, step log(6.4) / 27), fmax = sr / 2, norm=1
      (each triangle scaled by 2 / (f[i + 2] - f[i])).
  amp_to_db (:60-61), normalize (:52-53): the reference's own numpy, restated verbatim.

The anchors that ARE checked: frame count and shape as `gen_wavernn.py:44-47` consumes them, the basis rows summing to
the Slaney normalisation, a pure tone landing in the band that contains its frequency.
"""
import numpy as np


def hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)


def mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_basis(sr, n_fft, n_mels, fmin):
    fmax = sr / 2.0
    fftfreqs = np.linspace(0, sr / 2.0, 1 + n_fft // 2)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    weights = np.zeros((n_mels, 1 + n_fft // 2))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    return weights * enorm[:, np.newaxis]


def stft(y, n_fft, hop_length, win_length):
    y = np.asarray(y, dtype=np.float64)
    n = np.arange(win_length)
    window = 0.5 - 0.5 * np.cos(2 * np.pi * n / win_length)            # periodic Hann
    lpad = (n_fft - win_length) // 2
    window = np.pad(window, (lpad, n_fft - win_length - lpad))
    yp = np.pad(y, n_fft // 2, mode='reflect')
    n_frames = 1 + (len(yp) - n_fft) // hop_length
    frames = np.stack([yp[t * hop_length:t * hop_length + n_fft] for t in range(n_frames)], axis=1)
    return np.fft.rfft(window[:, None] * frames, axis=0)


def melspectrogram(y, sample_rate=22050, n_fft=2048, hop_length=275, win_length=1100, num_mels=80, fmin=40.0, min_level_db=-100.0):
    D = stft(y, n_fft, hop_length, win_length)
    S = np.dot(mel_basis(sample_rate, n_fft, num_mels, fmin), np.abs(D))
    S = 20 * np.log10(np.maximum(1e-5, S))
    return np.clip((S - min_level_db) / -min_level_db, 0, 1)
