"""Drop-in for `models/fatchord_version.py` of qingyundou/WaveRNN, generation path only.

Same constructor, same `state_dict` keys, same `generate(mels, save_path, batched, target,
overlap, mu_law)` contract (`/root/reference/models/fatchord_version.py:93-95,169-264`), so
`gen_wavernn.py` / `gen_tacotron.py` can swap the import and nothing else.  The module holds the
parameters (PyTorch is the checkpoint container); all arithmetic of the path runs in
libwavernn_b200.so through `binding.Vocoder`.  There is no CPU or eager-PyTorch fallback: on a
machine without the library or a CUDA device `generate` raises.

The parameter tree is built from one table (`config.state_dict_spec`) instead of nested layer
classes: the path never runs these tensors through torch ops, it only needs them named and
loadable.
"""
import math
import time
from pathlib import Path
from typing import Union

import numpy as np
import torch
import torch.nn as nn

from . import binding
from .config import VocoderDims, state_dict_spec
from .wavio import save_wav


def _container(root: nn.Module, path):
    m = root
    for name in path:
        if name not in m._modules:
            m.add_module(name, nn.Module())
        m = m._modules[name]
    return m


class WaveRNN(nn.Module):
    def __init__(self, rnn_dims, fc_dims, bits, pad, upsample_factors,
                 feat_dims, compute_dims, res_out_dims, res_blocks,
                 hop_length, sample_rate, mode='RAW'):
        super().__init__()
        self.dims = VocoderDims(rnn_dims=rnn_dims, fc_dims=fc_dims, bits=bits, pad=pad,
                                upsample_factors=tuple(upsample_factors), feat_dims=feat_dims,
                                compute_dims=compute_dims, res_out_dims=res_out_dims, res_blocks=res_blocks,
                                hop_length=hop_length, sample_rate=sample_rate, mode=mode)
        self.mode = mode
        self.pad = pad
        self.n_classes = self.dims.n_classes          # raises RuntimeError on an unknown mode
        self.rnn_dims = rnn_dims
        self.aux_dims = res_out_dims // 4
        self.hop_length = hop_length
        self.sample_rate = sample_rate
        self._keys = []
        for key, shape, role in state_dict_spec(self.dims):
            *path, leaf = key.split('.')
            holder = _container(self, path)
            if role in ('bn_m', 'bn_v', 'bn_n', 'step'):
                init = {'bn_m': torch.zeros(shape), 'bn_v': torch.ones(shape),
                        'bn_n': torch.tensor(0, dtype=torch.long), 'step': torch.zeros(1, dtype=torch.long)}[role]
                holder.register_buffer(leaf, init)
            else:
                holder.register_parameter(leaf, nn.Parameter(self._init(shape, role)))
            self._keys.append(key)
        self._voc = None
        self._voc_key = None
        self.num_params()

    @staticmethod
    def _init(shape, role):
        if role == 'bn_w':
            return torch.ones(shape)
        if role == 'bn_b':
            return torch.zeros(shape)
        if role == 'fir':
            return torch.full(shape, 1.0 / shape[-1])
        fan_in = shape[1] if role == 'gru' else (int(np.prod(shape[1:])) if len(shape) > 1 else shape[0])
        if role == 'gru_b':
            fan_in = shape[0] // 3
        if role == 'bias':
            fan_in = max(shape[0], 1)
        bound = 1.0 / math.sqrt(max(fan_in, 1))
        return torch.empty(shape).uniform_(-bound, bound)

    # ------------------------------------------------------------------ C-ABI handle
    def _vocoder(self):
        p0 = next(self.parameters())
        if p0.device.type != 'cuda':
            raise RuntimeError('wavernn_b200 has no CPU path: move the model to a CUDA device (model.cuda())')
        dev = p0.device.index if p0.device.index is not None else torch.cuda.current_device()
        sd = self.state_dict()
        key = (dev,) + tuple((sd[k]._version, sd[k].data_ptr()) for k in self._keys)
        if self._voc is None or self._voc_key is None or self._voc_key[0] != dev:
            if self._voc is not None:
                self._voc.close()
            self._voc = binding.Vocoder(self.dims, device=dev)
            self._voc_key = None
        if key != self._voc_key:           # first use, or load()/optimizer touched the weights: repack
            self._voc.load_weights({k: sd[k].detach().cpu().numpy() for k in self._keys})
            self._voc_key = key
        return self._voc

    # ------------------------------------------------------------------ the path
    def generate(self, mels, save_path: Union[str, Path], batched, target, overlap, mu_law,
                 noise=None, seed=None):
        """`WaveRNN.generate` (fatchord_version.py:169-264).  `noise`/`seed` are extensions for
        parity tests (pre-drawn noise in the reference's draw order) and reproducible sampling."""
        self.eval()
        start = time.time()
        try:
            voc = self._vocoder()
            mels = torch.as_tensor(mels)
            if mels.dim() != 3 or mels.size(0) != 1:
                raise ValueError(f'mels must have shape (1, {self.dims.feat_dims}, T), got {tuple(mels.shape)}')
            m = mels[0].detach().to(torch.float32).cpu().numpy()
            if seed is None:
                seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            stream = torch.cuda.current_stream(next(self.parameters()).device).cuda_stream
            output = voc.generate([m], batched, target, overlap, mu_law, noise=noise, seed=seed, stream=stream)[0]
            t = voc.last_timing()
            self.gen_display(t['steps'] - 1, t['steps'], t['folds'], start)
            save_wav(output, save_path, self.sample_rate)
        finally:
            self.train()
        return output

    def generate_many(self, mels_list, batched, target, overlap, mu_law, noise=None, seed=0):
        """Several utterances as ONE fold batch (the reference takes one utterance per call, F11)."""
        self.eval()
        try:
            voc = self._vocoder()
            ms = [torch.as_tensor(m).reshape(self.dims.feat_dims, -1).to(torch.float32).cpu().numpy() for m in mels_list]
            return voc.generate(ms, batched, target, overlap, mu_law, noise=noise, seed=seed)
        finally:
            self.train()

    def forward(self, x, mels):
        """Teacher-forced logits, inference only (`forward`, fatchord_version.py:131-167)."""
        self.step += 1
        voc = self._vocoder()
        out = voc.forward_logits(x.detach().cpu().numpy(), mels.detach().cpu().numpy())
        return torch.from_numpy(out).to(x.device)

    def gen_display(self, i, seq_len, b_size, start):
        gen_rate = (i + 1) / (time.time() - start) * b_size / 1000
        done = (i * 16) // max(seq_len, 1)
        pbar = ''.join('█' if j <= done else '░' for j in range(16))
        print(f'\r| {pbar} {i * b_size}/{seq_len * b_size} | Batch Size: {b_size} | Gen Rate: {gen_rate:.1f}kHz | ',
              end='')

    # ------------------------------------------------------------------ checkpoint API (unchanged layout)
    def get_step(self):
        return self.step.data.item()

    def log(self, path, msg):
        with open(path, 'a') as f:
            print(msg, file=f)

    def load(self, path: Union[str, Path]):
        device = next(self.parameters()).device
        self.load_state_dict(torch.load(path, map_location=device), strict=False)
        self._voc_key = None

    def save(self, path: Union[str, Path]):
        torch.save(self.state_dict(), path)

    def num_params(self, print_out=True):
        n = sum(int(np.prod(p.size())) for p in self.parameters() if p.requires_grad) / 1_000_000
        if print_out:
            print('Trainable Parameters: %.3fM' % n)
        return n
