"""Vocoder dimensions for the generation path.

Field names follow the reference constructor
(`models/fatchord_version.py:93-95`) and `hparams.py:20-60`; the derived
quantities (`aux_dims`, `n_classes`, `indent`) follow `:100-110` and `:69`.
"""
from dataclasses import dataclass, field
from typing import Tuple


@dataclass(frozen=True)
class VocoderDims:
    rnn_dims: int = 512
    fc_dims: int = 512
    bits: int = 9
    pad: int = 2
    upsample_factors: Tuple[int, ...] = (5, 5, 11)
    feat_dims: int = 80
    compute_dims: int = 128
    res_out_dims: int = 128
    res_blocks: int = 10
    hop_length: int = 275
    sample_rate: int = 22050
    mode: str = 'MOL'

    @property
    def aux_dims(self) -> int:
        return self.res_out_dims // 4

    @property
    def n_classes(self) -> int:
        if self.mode == 'RAW':
            return 2 ** self.bits
        if self.mode == 'MOL':
            return 30
        raise RuntimeError("Unknown model mode value - ", self.mode)

    @property
    def total_scale(self) -> int:
        s = 1
        for f in self.upsample_factors:
            s *= f
        return s

    @property
    def indent(self) -> int:
        return self.pad * self.total_scale

    def ctor_kwargs(self) -> dict:
        return dict(rnn_dims=self.rnn_dims, fc_dims=self.fc_dims, bits=self.bits, pad=self.pad,
                    upsample_factors=tuple(self.upsample_factors), feat_dims=self.feat_dims,
                    compute_dims=self.compute_dims, res_out_dims=self.res_out_dims,
                    res_blocks=self.res_blocks, hop_length=self.hop_length,
                    sample_rate=self.sample_rate, mode=self.mode)


def state_dict_spec(d: VocoderDims):
    """Ordered (key, shape, role) list of the fatchord checkpoint layout.

    Mirrors the key list of SURVEY.md §8(a) (`fatchord_version.py:114-125`); the
    loader accepts exactly these keys and nothing is renamed.
    role: 'conv'/'lin'/'gru' weights, 'bias', 'bn_w','bn_b','bn_m','bn_v','bn_n', 'fir', 'step'.
    """
    C, F, R = d.compute_dims, d.feat_dims, d.res_out_dims
    k = 2 * d.pad + 1
    spec = [('upsample.resnet.conv_in.weight', (C, F, k), 'conv')]

    def bn(prefix):
        return [(prefix + '.weight', (C,), 'bn_w'), (prefix + '.bias', (C,), 'bn_b'),
                (prefix + '.running_mean', (C,), 'bn_m'), (prefix + '.running_var', (C,), 'bn_v'),
                (prefix + '.num_batches_tracked', (), 'bn_n')]

    spec += bn('upsample.resnet.batch_norm')
    for i in range(d.res_blocks):
        p = f'upsample.resnet.layers.{i}'
        spec += [(p + '.conv1.weight', (C, C, 1), 'conv'), (p + '.conv2.weight', (C, C, 1), 'conv')]
        spec += bn(p + '.batch_norm1') + bn(p + '.batch_norm2')
    spec += [('upsample.resnet.conv_out.weight', (R, C, 1), 'conv'),
             ('upsample.resnet.conv_out.bias', (R,), 'bias')]
    for j, s in enumerate(d.upsample_factors):
        spec.append((f'upsample.up_layers.{2 * j + 1}.weight', (1, 1, 1, 2 * s + 1), 'fir'))
    H, A, FC = d.rnn_dims, d.aux_dims, d.fc_dims
    spec += [('I.weight', (H, F + A + 1), 'lin'), ('I.bias', (H,), 'bias')]
    for name, inp in (('rnn1', H), ('rnn2', H + A)):
        spec += [(f'{name}.weight_ih_l0', (3 * H, inp), 'gru'), (f'{name}.weight_hh_l0', (3 * H, H), 'gru'),
                 (f'{name}.bias_ih_l0', (3 * H,), 'gru_b'), (f'{name}.bias_hh_l0', (3 * H,), 'gru_b')]
    spec += [('fc1.weight', (FC, H + A), 'lin'), ('fc1.bias', (FC,), 'bias'),
             ('fc2.weight', (FC, FC + A), 'lin'), ('fc2.bias', (FC,), 'bias'),
             ('fc3.weight', (d.n_classes, FC), 'lin'), ('fc3.bias', (d.n_classes,), 'bias'),
             ('step', (1,), 'step')]
    return spec
