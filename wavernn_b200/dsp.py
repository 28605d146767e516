"""`utils/dsp.py` of the reference, the functions either side of the generation path, on the GPU library.

  melspectrogram(y)      utils/dsp.py:72-75 (librosa stft + mel basis in the reference): `gen_wavernn.py:44-47` calls it
                         when `gen_from_file` is handed a .wav instead of a .npy mel.
  load_wav / save_wav    utils/dsp.py:18-23: wavernn_b200/wavio.py (dependency-free).
Parameters default to hparams.py:20-27.  There is no CPU fallback: the C-ABI library does the transform on `device`.
"""
from __future__ import annotations

import numpy as np

from . import binding
from .config import VocoderDims
from . import wavio
from .wavio import save_wav      # noqa: F401  (re-exported: the reference keeps it in the same module)

HP = dict(sample_rate=22050, n_fft=2048, hop_length=275, win_length=1100, num_mels=80, fmin=40.0, min_level_db=-100.0)
_HANDLE = {}


def _vocoder(device=0):
    if device not in _HANDLE:
        _HANDLE[device] = binding.Vocoder(VocoderDims(), device=device)
    return _HANDLE[device]


def melspectrogram(y, device=0, **hp):
    """Normalised mel spectrogram [num_mels][1 + len(y) // hop_length], float32 (the reference returns float64 and the
    model casts it to float32 on entry, `gen_wavernn.py:49`)."""
    p = dict(HP)
    p.update(hp)
    return _vocoder(device).melspectrogram(np.asarray(y, dtype=np.float32), **p)


def load_wav(path, sample_rate=HP['sample_rate']):
    """utils/dsp.py:18-19: the samples as float32 in [-1, 1].  The reference lets librosa resample to hp.sample_rate; this
    reader does not resample: a file at another rate is an error."""
    data, rate = wavio.load_wav(path)
    if rate != sample_rate:
        raise ValueError(f'{path}: sample rate {rate}, the model expects {sample_rate} (resample the file first)')
    return data
