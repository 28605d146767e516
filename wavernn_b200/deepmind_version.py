"""Drop-in for `models/deepmind_version.py` (generation path): same constructor, parameter names
and `generate(seq_len) -> (output, coarse, fine)` contract (`deepmind_version.py:8-35,75-165`).
The shipped reference crashes at its progress print (`:159`); this one does not
print per step.  Arithmetic runs in libwavernn_b200.so; no CPU fallback.
"""
import math

import numpy as np
import torch
import torch.nn as nn

from . import binding


class _Affine(nn.Module):
    def __init__(self, n_in, n_out, bias=True):
        super().__init__()
        b = 1.0 / math.sqrt(n_in)
        self.weight = nn.Parameter(torch.empty(n_out, n_in).uniform_(-b, b))
        if bias:
            self.bias = nn.Parameter(torch.empty(n_out).uniform_(-b, b))


class WaveRNN(nn.Module):
    def __init__(self, hidden_size=896, quantisation=256):
        super().__init__()
        self.hidden_size = hidden_size
        self.split_size = hidden_size // 2
        self.quantisation = quantisation
        s = self.split_size
        self.R = _Affine(hidden_size, 3 * hidden_size, bias=False)
        self.O1, self.O2 = _Affine(s, s), _Affine(s, quantisation)
        self.O3, self.O4 = _Affine(s, s), _Affine(s, quantisation)
        self.I_coarse = _Affine(2, 3 * s, bias=False)
        self.I_fine = _Affine(3, 3 * s, bias=False)
        self.bias_u = nn.Parameter(torch.zeros(hidden_size))
        self.bias_r = nn.Parameter(torch.zeros(hidden_size))
        self.bias_e = nn.Parameter(torch.zeros(hidden_size))
        self._voc, self._voc_key = None, None
        self.num_params()

    def _vocoder(self):
        p0 = next(self.parameters())
        if p0.device.type != 'cuda':
            raise RuntimeError('wavernn_b200 has no CPU path: move the model to a CUDA device (model.cuda())')
        dev = p0.device.index if p0.device.index is not None else torch.cuda.current_device()
        sd = self.state_dict()
        key = (dev,) + tuple((v._version, v.data_ptr()) for v in sd.values())
        if self._voc is None or self._voc_key is None or self._voc_key[0] != dev:
            self._voc = binding.DeepmindVocoder(self.hidden_size, self.quantisation, device=dev)
            self._voc_key = None
        if key != self._voc_key:
            self._voc.load_weights({k: v.detach().cpu().numpy() for k, v in sd.items()})
            self._voc_key = key
        return self._voc

    def generate(self, seq_len, n_seq=1, noise=None, seed=None):
        """-> (output, coarse, fine) int64 numpy, shape (seq_len,) (or (n_seq, seq_len) when n_seq > 1)."""
        voc = self._vocoder()
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        qc, qf = noise if noise is not None else (None, None)
        out, c, f = voc.generate(seq_len, n_seq, qc, qf, seed)
        if n_seq == 1:
            return out[0], c[0], f[0]
        return out, c, f

    def get_initial_hidden(self, batch_size=1):
        return torch.zeros(batch_size, self.hidden_size, device=next(self.parameters()).device)

    def num_params(self, print_out=True):
        n = sum(int(np.prod(p.size())) for p in self.parameters() if p.requires_grad) / 1_000_000
        if print_out:
            print('Trainable Parameters: %.3f million' % n)
        return n
