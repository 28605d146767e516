"""A model object with the attribute names `wavernn_b200.tacotron.generate` needs (the reference's `Tacotron`): an
embedding as the encoder, the seeded decoder parameters, one convolution as the postnet.  Test helper."""
import torch
from torch import nn


class Bag(nn.Module):
    pass


def bag_from(sd, prefix):
    root = Bag()
    for k, v in sd.items():
        if not k.startswith(prefix):
            continue
        parts = k[len(prefix):].split('.')
        node = root
        for name in parts[:-1]:
            if not hasattr(node, name):
                setattr(node, name, Bag())
            node = getattr(node, name)
        node.register_parameter(parts[-1], nn.Parameter(torch.from_numpy(v.copy())))
    return root


class Postnet(nn.Module):
    def __init__(self, n_mels):
        super().__init__()
        self.conv = nn.Conv1d(n_mels, 64, 3, padding=1)

    def forward(self, x):
        return torch.tanh(self.conv(x)).transpose(1, 2)


class StandIn(nn.Module):
    def __init__(self, sd, n_mels=80, r=2, stop_threshold=-3.4):
        super().__init__()
        torch.manual_seed(11)
        self.encoder = nn.Embedding(70, 256)
        self.encoder_proj = nn.Linear(256, 256, bias=False)
        self.encoder_proj.weight.data.copy_(torch.from_numpy(sd['encoder_proj.weight']))
        self.decoder = bag_from(sd, 'decoder.')
        self.postnet = Postnet(n_mels)
        self.post_proj = nn.Linear(64, n_mels, bias=False)
        self.register_buffer('stop_threshold', torch.tensor(stop_threshold))
        self.r = r


def build_stand_in(sd, **kw):
    return StandIn(sd, **kw)
