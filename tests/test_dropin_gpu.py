"""GPU: the drop-in `WaveRNN` module keeps the reference's call contract."""
import os

import numpy as np
import pytest
import torch

from wavernn_b200 import synth, wavio
from wavernn_b200.fatchord_version import WaveRNN

from common import CASES, case_inputs, golden

pytestmark = pytest.mark.gpu


def test_generate_contract(tmp_path):
    name = 'tiny_mol_batched'
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    model = WaveRNN(**dims.ctor_kwargs()).cuda()
    ckpt = tmp_path / 'latest_weights.pyt'
    torch.save({k: torch.from_numpy(v) for k, v in sd.items()}, ckpt)
    model.load(ckpt)                                   # checkpoint layout unchanged
    assert model.get_step() == 800000
    path = tmp_path / 'out.wav'
    out = model.generate(torch.as_tensor(mel), path, batched, target, overlap, mu_law, noise=noise)
    assert model.training                              # train() restored at exit (:262)
    assert isinstance(out, np.ndarray) and out.dtype == np.float64 and out.shape == ((T - 1) * dims.hop_length,)
    assert np.abs(out - golden(name)['wave']).max() <= 2e-4
    data, rate = wavio.load_wav(path)
    assert rate == dims.sample_rate and np.array_equal(data, out.astype(np.float32))
    # weights changed in place -> repacked
    with torch.no_grad():
        model.fc3.bias.add_(0.5)
    out2 = model.generate(torch.as_tensor(mel), path, batched, target, overlap, mu_law, noise=noise)
    assert np.abs(out2 - out).max() > 1e-3


def test_cpu_module_raises():
    dims = CASES['tiny_mol_batched'][0]
    model = WaveRNN(**dims.ctor_kwargs())
    with pytest.raises(RuntimeError):
        model.generate(torch.zeros(1, dims.feat_dims, 40), os.devnull, False, 100, 20, False)
    assert model.training
