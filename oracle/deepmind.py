"""Oracle: deepmind coarse/fine WaveRNN generation on CPU.  TEST INFRASTRUCTURE ONLY.

Follows `models/deepmind_version.py:75-165` (generate) and `utils/dsp.py:33-34`
(combine_signal); `Categorical.sample` is replaced by argmax(probs / q) with injected
q ~ Exp(1).  The reference supports one sequence; `N` independent sequences
are run as independent batch rows (each row sees exactly the batch-1 arithmetic of the
reference up to ATen's batched-GEMM summation order).
"""
import numpy as np
import torch
import torch.nn.functional as F


def combine_signal(coarse, fine):
    """`combine_signal` (`utils/dsp.py:33-34`)."""
    return coarse * 256 + fine - 2 ** 15


def _categorical(logits, q):
    p = F.softmax(logits, dim=1)
    p = p / p.sum(-1, keepdim=True)
    return torch.argmax(p / q, dim=-1)


def generate(W, seq_len, q_coarse, q_fine, n_seq=1, want_logits=False):
    """q_coarse, q_fine: (seq_len, n_seq, quantisation) Exp(1) draws.
    -> output, coarse, fine int64 arrays of shape (n_seq, seq_len) (squeezed when n_seq == 1)."""
    hidden_size = W['R.weight'].shape[1]
    split = hidden_size // 2
    with torch.no_grad():
        b_cu, b_fu = torch.split(W['bias_u'], split)
        b_cr, b_fr = torch.split(W['bias_r'], split)
        b_ce, b_fe = torch.split(W['bias_e'], split)
        c_out, f_out, lc, lf = [], [], [], []
        out_coarse = torch.zeros(n_seq, dtype=torch.long)
        out_fine = torch.zeros(n_seq, dtype=torch.long)
        hidden = torch.zeros(n_seq, hidden_size)
        for i in range(seq_len):
            hidden_coarse, hidden_fine = torch.split(hidden, split, dim=1)
            pc = out_coarse.unsqueeze(1).float() / 127.5 - 1.
            pf = out_fine.unsqueeze(1).float() / 127.5 - 1.
            prev_outputs = torch.cat([pc, pf], dim=1)
            I_cu, I_cr, I_ce = torch.split(F.linear(prev_outputs, W['I_coarse.weight']), split, dim=1)
            R_hidden = F.linear(hidden, W['R.weight'])
            R_cu, R_fu, R_cr, R_fr, R_ce, R_fe = torch.split(R_hidden, split, dim=1)
            u = torch.sigmoid(R_cu + I_cu + b_cu)
            r = torch.sigmoid(R_cr + I_cr + b_cr)
            e = torch.tanh(r * R_ce + I_ce + b_ce)
            hidden_coarse = u * hidden_coarse + (1. - u) * e
            logits_c = F.linear(F.relu(F.linear(hidden_coarse, W['O1.weight'], W['O1.bias'])),
                                W['O2.weight'], W['O2.bias'])
            out_coarse = _categorical(logits_c, torch.as_tensor(q_coarse[i]))
            c_out.append(out_coarse)
            coarse_pred = out_coarse.float() / 127.5 - 1.
            fine_input = torch.cat([prev_outputs, coarse_pred.unsqueeze(1)], dim=1)
            I_fu, I_fr, I_fe = torch.split(F.linear(fine_input, W['I_fine.weight']), split, dim=1)
            u = torch.sigmoid(R_fu + I_fu + b_fu)
            r = torch.sigmoid(R_fr + I_fr + b_fr)
            e = torch.tanh(r * R_fe + I_fe + b_fe)
            hidden_fine = u * hidden_fine + (1. - u) * e
            logits_f = F.linear(F.relu(F.linear(hidden_fine, W['O3.weight'], W['O3.bias'])),
                                W['O4.weight'], W['O4.bias'])
            out_fine = _categorical(logits_f, torch.as_tensor(q_fine[i]))
            f_out.append(out_fine)
            if want_logits:
                lc.append(logits_c)
                lf.append(logits_f)
            hidden = torch.cat([hidden_coarse, hidden_fine], dim=1)
        coarse = torch.stack(c_out).transpose(0, 1).numpy()
        fine = torch.stack(f_out).transpose(0, 1).numpy()
    output = combine_signal(coarse, fine)
    if n_seq == 1:
        output, coarse, fine = output[0], coarse[0], fine[0]
    if want_logits:
        return output, coarse, fine, torch.stack(lc), torch.stack(lf)
    return output, coarse, fine
