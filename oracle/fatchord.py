"""Oracle: fatchord WaveRNN generation path on CPU.  TEST INFRASTRUCTURE ONLY (see __init__).

Weights `W` are a dict key -> torch.float32 CPU tensor in checkpoint layout
(`wavernn_b200.config.state_dict_spec`).  Every function cites the reference lines it follows.
Parity pinned by `tests/golden/make_golden.py` (bit-equal to the reference in the build
container) and `tests/test_oracle_golden.py` (fixtures, everywhere).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


def to_torch(state_dict_np):
    return {k: torch.from_numpy(np.array(v, copy=True)) for k, v in state_dict_np.items()}


# ----------------------------------------------------------------------------- conditioning
def pad_time(x, pad, side='both'):
    """`pad_tensor` (`fatchord_version.py:281-291`): zero-pad axis 1 of (b, t, c)."""
    b, t, c = x.shape
    total = t + 2 * pad if side == 'both' else t + pad
    out = torch.zeros(b, total, c, dtype=x.dtype)
    if side in ('before', 'both'):
        out[:, pad:pad + t, :] = x
    elif side == 'after':
        out[:, :t, :] = x
    return out


def _bn_eval(x, W, prefix):
    """BatchNorm1d in eval mode, eps 1e-5 (`fatchord_version.py:18-19,36`)."""
    return F.batch_norm(x, W[prefix + '.running_mean'], W[prefix + '.running_var'],
                        W[prefix + '.weight'], W[prefix + '.bias'], False, 0.1, 1e-5)


def mel_resnet(W, dims, m):
    """`MelResNet.forward` + `ResBlock.forward` (`fatchord_version.py:21-28,42-48`).
    m: (1, feat, T + 2*pad) -> (1, res_out, T)."""
    p = 'upsample.resnet.'
    x = F.conv1d(m, W[p + 'conv_in.weight'])
    x = F.relu(_bn_eval(x, W, p + 'batch_norm'))
    for i in range(dims.res_blocks):
        q = f'{p}layers.{i}.'
        r = x
        x = F.conv1d(x, W[q + 'conv1.weight'])
        x = F.relu(_bn_eval(x, W, q + 'batch_norm1'))
        x = F.conv1d(x, W[q + 'conv2.weight'])
        x = _bn_eval(x, W, q + 'batch_norm2')
        x = x + r
    return F.conv1d(x, W[p + 'conv_out.weight'], W[p + 'conv_out.bias'])


def upsample(W, dims, m):
    """`UpsampleNetwork.forward` (`fatchord_version.py:82-89`), `Stretch2d` (`:57-61`).
    m: (1, feat, T + 2*pad) -> mels (1, T*hop, feat), aux (1, T*hop, res_out)."""
    aux = mel_resnet(W, dims, m)
    aux = aux.repeat_interleave(dims.total_scale, dim=2)          # nearest repeat along time
    x = m.unsqueeze(1)                                            # (1,1,feat,T')
    for j, s in enumerate(dims.upsample_factors):
        x = x.repeat_interleave(s, dim=3)
        x = F.conv2d(x, W[f'upsample.up_layers.{2 * j + 1}.weight'], padding=(0, s))
    x = x.squeeze(1)[:, :, dims.indent:-dims.indent]
    return x.transpose(1, 2), aux.transpose(1, 2)


def fold_with_overlap(x, target, overlap):
    """`fold_with_overlap` (`fatchord_version.py:293-340`). x (1, L, F) -> (folds, target+2*overlap, F)."""
    _, total_len, features = x.shape
    num_folds = (total_len - overlap) // (target + overlap)
    extended_len = num_folds * (overlap + target) + overlap
    remaining = total_len - extended_len
    if remaining != 0:
        num_folds += 1
        x = pad_time(x, target + 2 * overlap - remaining, side='after')
    folded = torch.zeros(num_folds, target + 2 * overlap, features, dtype=x.dtype)
    for i in range(num_folds):
        s = i * (target + overlap)
        folded[i] = x[0, s:s + target + 2 * overlap, :]
    return folded


def num_folds_for(total_len, target, overlap):
    n = (total_len - overlap) // (target + overlap)
    if total_len - (n * (overlap + target) + overlap) != 0:
        n += 1
    return n


def xfade_and_unfold(y, target, overlap):
    """`xfade_and_unfold` (`fatchord_version.py:342-405`); y float64 (folds, target+2*overlap).
    Mutates y in place like the reference.  `target` is recomputed from y (`:375`)."""
    num_folds, length = y.shape
    target = length - 2 * overlap
    total_len = num_folds * (target + overlap) + overlap
    silence_len = overlap // 2
    fade_len = overlap - silence_len
    t = np.linspace(-1, 1, fade_len, dtype=np.float64)
    fade_in = np.concatenate([np.zeros(silence_len, dtype=np.float64), np.sqrt(0.5 * (1 + t))])
    fade_out = np.concatenate([np.ones(silence_len, dtype=np.float64), np.sqrt(0.5 * (1 - t))])
    y[:, :overlap] *= fade_in
    y[:, -overlap:] *= fade_out
    unfolded = np.zeros(total_len, dtype=np.float64)
    for i in range(num_folds):
        s = i * (target + overlap)
        unfolded[s:s + target + 2 * overlap] += y[i]
    return unfolded


def decode_mu_law(y, mu, from_labels=True):
    """`decode_mu_law` (`utils/dsp.py:98-103`) incl. `label_2_float` (`:8-9`)."""
    if from_labels:
        y = 2 * y / (2 ** math.log2(mu) - 1.) - 1.
    mu = mu - 1
    return np.sign(y) / mu * ((1 + mu) ** np.abs(y) - 1)


# ----------------------------------------------------------------------------- sampling
def mol_sample(logits, u1, u2):
    """`sample_from_discretized_mix_logistic` (`utils/distribution.py:87-123`) with the two
    uniform draws injected.  logits (B, 30), u1 (B, 10), u2 (B,) -> (sample (B,), k (B,))."""
    log_scale_min = float(np.log(1e-14))
    nr_mix = logits.shape[1] // 3
    y = logits.unsqueeze(0)                                       # (1, B, 30) == y.transpose(1,2)
    logit_probs = y[:, :, :nr_mix]
    temp = logit_probs - torch.log(-torch.log(u1.unsqueeze(0)))
    _, argmax = temp.max(dim=-1)
    one_hot = F.one_hot(argmax, nr_mix).float()
    means = torch.sum(y[:, :, nr_mix:2 * nr_mix] * one_hot, dim=-1)
    log_scales = torch.clamp(torch.sum(y[:, :, 2 * nr_mix:3 * nr_mix] * one_hot, dim=-1), min=log_scale_min)
    u = u2.unsqueeze(0)
    x = means + torch.exp(log_scales) * (torch.log(u) - torch.log(1. - u))
    x = torch.clamp(torch.clamp(x, min=-1.), max=1.)
    return x.view(-1), argmax.view(-1)


def raw_sample(logits, q, n_classes):
    """RAW branch (`fatchord_version.py:231-237`): softmax -> Categorical.sample -> affine.
    `Categorical(probs)` renormalises probs; ATen multinomial == argmax(probs / Exp(1)) (F8)."""
    p = F.softmax(logits, dim=1)
    p = p / p.sum(-1, keepdim=True)
    k = torch.argmax(p / q, dim=-1)
    return 2 * k.float() / (n_classes - 1.) - 1., k


# ----------------------------------------------------------------------------- per-step network
def step_logits(W, dims, x, m_t, a_t, h1, h2):
    """One iteration of the hot loop up to fc3 (`fatchord_version.py:203-223`).
    x (B,1); m_t (B,feat); a_t = 4 tensors (B,aux); h1,h2 (B,rnn) -> logits, h1, h2."""
    x = torch.cat([x, m_t, a_t[0]], dim=1)
    x = F.linear(x, W['I.weight'], W['I.bias'])
    h1 = torch.gru_cell(x, h1, W['rnn1.weight_ih_l0'], W['rnn1.weight_hh_l0'],
                        W['rnn1.bias_ih_l0'], W['rnn1.bias_hh_l0'])
    x = x + h1
    inp = torch.cat([x, a_t[1]], dim=1)
    h2 = torch.gru_cell(inp, h2, W['rnn2.weight_ih_l0'], W['rnn2.weight_hh_l0'],
                        W['rnn2.bias_ih_l0'], W['rnn2.bias_hh_l0'])
    x = x + h2
    x = F.relu(F.linear(torch.cat([x, a_t[2]], dim=1), W['fc1.weight'], W['fc1.bias']))
    x = F.relu(F.linear(torch.cat([x, a_t[3]], dim=1), W['fc2.weight'], W['fc2.bias']))
    return F.linear(x, W['fc3.weight'], W['fc3.bias']), h1, h2


def conditioning(W, dims, mels, batched, target, overlap):
    """`generate` lines `:183-190`: pad, upsample, optional fold. mels (1, feat, T) float32."""
    mels = torch.as_tensor(mels, dtype=torch.float32)
    wave_len = (mels.size(-1) - 1) * dims.hop_length
    m = pad_time(mels.transpose(1, 2), dims.pad, 'both')
    m, aux = upsample(W, dims, m.transpose(1, 2))
    if batched:
        m = fold_with_overlap(m, target, overlap)
        aux = fold_with_overlap(aux, target, overlap)
    return m.contiguous(), aux.contiguous(), wave_len


def run_loop(W, dims, m, aux, noise, max_steps=None, teacher=None, want_logits=False):
    """The hot loop (`fatchord_version.py:192-241`).
    noise: MOL -> (u1 (S,B,10), u2 (S,B)); RAW -> q (S,B,C).
    teacher: optional (B,S) float32 of forced previous samples (x fed at step i is teacher[:, i-1]).
    -> dict(samples (B,S) float32, [logits (S,B,C)], [picks (S,B) int64])."""
    B, S, _ = m.shape
    if max_steps is not None:
        S = min(S, max_steps)
    d = dims.aux_dims
    h1 = torch.zeros(B, dims.rnn_dims)
    h2 = torch.zeros(B, dims.rnn_dims)
    x = torch.zeros(B, 1)
    out, logit_list, picks = [], [], []
    for i in range(S):
        a_t = [aux[:, i, d * j:d * (j + 1)] for j in range(4)]
        logits, h1, h2 = step_logits(W, dims, x, m[:, i, :], a_t, h1, h2)
        if want_logits:
            logit_list.append(logits)
        if dims.mode == 'MOL':
            s, k = mol_sample(logits, torch.as_tensor(noise[0][i]), torch.as_tensor(noise[1][i]))
        elif dims.mode == 'RAW':
            s, k = raw_sample(logits, torch.as_tensor(noise[i]), dims.n_classes)
        else:
            raise RuntimeError("Unknown model mode value - ", dims.mode)
        out.append(s)
        picks.append(k)
        x = (teacher[:, i] if teacher is not None else s).reshape(B, 1)
    res = {'samples': torch.stack(out).transpose(0, 1).contiguous(),
           'picks': torch.stack(picks)}
    if want_logits:
        res['logits'] = torch.stack(logit_list)
    return res


def finish(dims, samples, wave_len, batched, target, overlap, mu_law):
    """Host tail of `generate` (`fatchord_version.py:243-258`); samples (B,S) float32 -> float64 wave."""
    output = samples.numpy().astype(np.float64)
    mu_law = mu_law if dims.mode == 'RAW' else False
    if mu_law:
        output = decode_mu_law(output, dims.n_classes, False)
    if batched:
        output = xfade_and_unfold(output, target, overlap)
    else:
        output = output[0]
    fade_out = np.linspace(1, 0, 20 * dims.hop_length)
    output = output[:wave_len]
    output[-20 * dims.hop_length:] *= fade_out
    return output


def generate(W, dims, mels, batched, target, overlap, mu_law, noise):
    """`WaveRNN.generate` (`fatchord_version.py:169-264`) with injected noise; returns float64 wave."""
    with torch.no_grad():
        m, aux, wave_len = conditioning(W, dims, mels, batched, target, overlap)
        res = run_loop(W, dims, m, aux, noise)
    return finish(dims, res['samples'], wave_len, batched, target, overlap, mu_law)


def forward_logits(W, dims, x, mels):
    """Teacher-forced `WaveRNN.forward` (`fatchord_version.py:131-167`) as a GRUCell loop (F9).
    x (B, S) float32 previous samples, mels (B, feat, S/hop + 2*pad) -> logits (B, S, C)."""
    with torch.no_grad():
        outs = []
        for b in range(x.shape[0]):
            m, aux = upsample(W, dims, torch.as_tensor(mels[b:b + 1]))
            S = x.shape[1]
            d = dims.aux_dims
            h1 = torch.zeros(1, dims.rnn_dims)
            h2 = torch.zeros(1, dims.rnn_dims)
            row = []
            for i in range(S):
                a_t = [aux[:, i, d * j:d * (j + 1)] for j in range(4)]
                lg, h1, h2 = step_logits(W, dims, torch.as_tensor(x[b:b + 1, i:i + 1]), m[:, i, :], a_t, h1, h2)
                row.append(lg[0])
            outs.append(torch.stack(row))
        return torch.stack(outs)
