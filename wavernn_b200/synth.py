"""Deterministic synthetic checkpoints and mels (no network, no pretrained blobs).

The pretrained `ljspeech.wavernn.mol.800k` checkpoint is unavailable on this box
(SURVEY.md F3), so tests, `smoke()` and `bench.py` use weights drawn here from a
numpy `RandomState`, keyed by checkpoint key, so the same state_dict can be rebuilt
bit-for-bit in this container (to drive the real reference) and on the GPU box.

style 'init'    : PyTorch-default-like uniform bounds (1/sqrt(fan_in)), BatchNorm
                  running stats re-drawn so eval-mode BN is not the identity.
style 'trained' : rescaled so GRU gates saturate, MoL log-scales reach the clamp
                  region and samples hit the +-1 clip.
"""
import zlib

import numpy as np

from .config import VocoderDims, state_dict_spec


def _rs(seed, key):
    return np.random.RandomState((seed * 1000003 + zlib.crc32(key.encode())) % (2 ** 32))


def synth_state_dict(dims: VocoderDims, seed: int = 0, style: str = 'init'):
    """-> dict key -> np.ndarray (float32, int64 for counters) in checkpoint layout."""
    gain = {'init': 1.0, 'trained': 2.5}[style]
    out = {}
    for key, shape, role in state_dict_spec(dims):
        r = _rs(seed, key)
        if role in ('conv', 'lin', 'gru'):
            fan_in = int(np.prod(shape[1:]))
            b = 1.0 / np.sqrt(fan_in)
            g = gain if role in ('gru', 'lin') else 1.0
            w = r.uniform(-b * g, b * g, size=shape)
        elif role in ('bias', 'gru_b'):
            w = r.uniform(-0.1, 0.1, size=shape) * gain
        elif role == 'bn_w':
            w = r.uniform(0.5, 1.5, size=shape)
        elif role == 'bn_b':
            w = r.normal(0.0, 0.1, size=shape)
        elif role == 'bn_m':
            w = r.normal(0.0, 0.1, size=shape)
        elif role == 'bn_v':
            w = r.uniform(0.5, 1.5, size=shape)
        elif role == 'fir':
            n = shape[-1]
            w = np.full(shape, 1.0 / n) + r.uniform(-0.2 / n, 0.2 / n, size=shape)
        elif role == 'bn_n':
            out[key] = np.array(1000, dtype=np.int64)
            continue
        elif role == 'step':
            out[key] = np.array([800000], dtype=np.int64)
            continue
        else:
            raise KeyError(role)
        out[key] = np.ascontiguousarray(w, dtype=np.float32)
    if style == 'trained' and dims.mode == 'MOL':
        # push the ten log-scales towards small values and spread the means
        b3 = out['fc3.bias']
        b3[20:30] = _rs(seed, 'fc3.ls').uniform(-9.0, -3.0, size=10).astype(np.float32)
        b3[10:20] = _rs(seed, 'fc3.mu').uniform(-0.9, 0.9, size=10).astype(np.float32)
        out['fc3.weight'][20:30] *= 4.0
    return out


def synth_mel(T: int, seed: int = 1234, feat_dims: int = 80):
    """`torch.rand(1, feat, T)` with a seeded CPU generator -> np.float32."""
    import torch
    g = torch.Generator().manual_seed(seed)
    return torch.rand((1, feat_dims, T), generator=g).numpy()


def synth_noise_mol(S: int, B: int, seed: int = 4321, nr_mix: int = 10):
    """Pre-drawn uniforms in the draw order of `utils/distribution.py:106,118`."""
    import torch
    g = torch.Generator().manual_seed(seed)
    u1 = torch.empty(S, B, nr_mix).uniform_(1e-5, 1.0 - 1e-5, generator=g)
    u2 = torch.empty(S, B).uniform_(1e-5, 1.0 - 1e-5, generator=g)
    return u1.numpy(), u2.numpy()


def synth_noise_exp(S: int, B: int, n_classes: int, seed: int = 4321):
    """Pre-drawn Exp(1) race variables for `Categorical.sample` == argmax(p/q)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    q = torch.empty(S, B, n_classes).exponential_(1.0, generator=g)
    return q.numpy()


def synth_dm_state_dict(hidden_size: int = 896, quantisation: int = 256, seed: int = 0):
    """deepmind_version checkpoint layout (`models/deepmind_version.py:15-31`)."""
    s = hidden_size // 2
    shapes = [('R.weight', (3 * hidden_size, hidden_size)),
              ('O1.weight', (s, s)), ('O1.bias', (s,)), ('O2.weight', (quantisation, s)), ('O2.bias', (quantisation,)),
              ('O3.weight', (s, s)), ('O3.bias', (s,)), ('O4.weight', (quantisation, s)), ('O4.bias', (quantisation,)),
              ('I_coarse.weight', (3 * s, 2)), ('I_fine.weight', (3 * s, 3)),
              ('bias_u', (hidden_size,)), ('bias_r', (hidden_size,)), ('bias_e', (hidden_size,))]
    out = {}
    for key, shape in shapes:
        r = _rs(seed, 'dm.' + key)
        if len(shape) == 2:
            b = 1.0 / np.sqrt(shape[1])
            # a larger gain on the output layers makes the softmaxes non-uniform
            g = 3.0 if key[0] == 'O' else 1.5
            w = r.uniform(-b * g, b * g, size=shape)
        else:
            w = r.uniform(-0.2, 0.2, size=shape)
        out[key] = np.ascontiguousarray(w, dtype=np.float32)
    return out


def taco_decoder_spec(decoder_dims: int = 256, lstm_dims: int = 512, n_mels: int = 80, max_r: int = 20):
    """Keys and shapes of the Tacotron decoder loop's parameters (`models/tacotron.py:171-227,300-303`)."""
    D, H, M = decoder_dims, lstm_dims, n_mels
    return [('encoder_proj.weight', (D, D)),
            ('decoder.prenet.fc1.weight', (256, M)), ('decoder.prenet.fc1.bias', (256,)),
            ('decoder.prenet.fc2.weight', (128, 256)), ('decoder.prenet.fc2.bias', (128,)),
            ('decoder.attn_net.conv.weight', (32, 2, 31)),
            ('decoder.attn_net.L.weight', (D, 32)), ('decoder.attn_net.L.bias', (D,)),
            ('decoder.attn_net.W.weight', (D, D)), ('decoder.attn_net.W.bias', (D,)),
            ('decoder.attn_net.v.weight', (1, D)),
            ('decoder.attn_rnn.weight_ih', (3 * D, D + D // 2)), ('decoder.attn_rnn.weight_hh', (3 * D, D)),
            ('decoder.attn_rnn.bias_ih', (3 * D,)), ('decoder.attn_rnn.bias_hh', (3 * D,)),
            ('decoder.rnn_input.weight', (H, 2 * D)), ('decoder.rnn_input.bias', (H,)),
            ('decoder.res_rnn1.weight_ih', (4 * H, H)), ('decoder.res_rnn1.weight_hh', (4 * H, H)),
            ('decoder.res_rnn1.bias_ih', (4 * H,)), ('decoder.res_rnn1.bias_hh', (4 * H,)),
            ('decoder.res_rnn2.weight_ih', (4 * H, H)), ('decoder.res_rnn2.weight_hh', (4 * H, H)),
            ('decoder.res_rnn2.bias_ih', (4 * H,)), ('decoder.res_rnn2.bias_hh', (4 * H,)),
            ('decoder.mel_proj.weight', (M * max_r, H))]


def synth_taco_state_dict(decoder_dims: int = 256, lstm_dims: int = 512, n_mels: int = 80, seed: int = 0,
                          gain: float = 2.0):
    """Seeded decoder-loop checkpoint: uniform +-gain/sqrt(fan_in) matrices (the attention energies, the gates and
    the mel frames move over a useful range at gain 2), small biases."""
    out = {}
    for key, shape in taco_decoder_spec(decoder_dims, lstm_dims, n_mels):
        r = _rs(seed, 'taco.' + key)
        if len(shape) >= 2:
            b = gain / np.sqrt(int(np.prod(shape[1:])))
            if key.endswith('attn_net.v.weight'):
                b *= 6.0            # energies of +-8: a peaked attention that moves, not a flat one
            w = r.uniform(-b, b, size=shape)
        else:
            w = r.uniform(-0.2, 0.2, size=shape)
        out[key] = np.ascontiguousarray(w, dtype=np.float32)
    # a trained LSA keeps most energies far below zero (smoothed attention = sigmoid / sum of sigmoids): lean the query
    # bias against v so that only a few positions per step come near sigmoid's upper half
    out['decoder.attn_net.W.bias'] -= (0.15 * np.sign(out['decoder.attn_net.v.weight'][0])).astype(np.float32)
    return out


def synth_encoder_seq(L: int, decoder_dims: int = 256, seed: int = 77):
    """Stand-in for the CBHG encoder's output (a bidirectional GRU's hidden states: values in (-1, 1))."""
    r = np.random.RandomState(seed)
    return np.ascontiguousarray(np.tanh(r.normal(0.0, 0.8, size=(L, decoder_dims))), dtype=np.float32)
