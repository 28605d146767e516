"""Oracle: the Tacotron decoder loop of `Tacotron.generate` on CPU.  TEST INFRASTRUCTURE ONLY.

Restates `models/tacotron.py:438-465` (encoder projection + the free-running decoder loop with its stop rule) and
the modules the loop calls: `Decoder.forward` (:227-287, eval mode: no zoneout, no dropout), `PreNet` (:137-147),
`LSA.forward` (:187-205, smoothed attention: sigmoid(u) / sum sigmoid(u), cumulative += attention) and the
reshape of `mel_proj`'s output (`mels.view(batch, n_mels, max_r)[:, :, :r]`, :280-281).  ATen `linear`, `conv1d`,
`gru_cell` and `lstm_cell` are called at the reference's call sites, so on CPU the outputs equal the reference's
bit for bit (checked by `tests/golden/make_taco_golden.py`, which writes `tests/golden/taco_*.npz`).

The boundary is the loop: the CBHG encoder before it and the CBHG postnet after it run once per utterance and are not
restated; `enc_seq` is the encoder's output (L, decoder_dims).
"""
import numpy as np
import torch
import torch.nn.functional as F

MAX_R = 20          # `Decoder.max_r` (:210)


def to_torch(W):
    return {k: torch.as_tensor(np.asarray(v)) for k, v in W.items()}


def decode(W, enc_seq, steps=2000, r=2, stop_threshold=-3.4):
    """-> (mel (n_mels, n_iter * r) float32, attn (n_iter, L) float32) as `Tacotron.generate` returns them
    (before the postnet).  W: state dict (numpy or torch) with the `decoder.*` and `encoder_proj.weight` keys."""
    W = to_torch(W)
    g = lambda k: W['decoder.' + k]
    with torch.no_grad():
        encoder_seq = torch.as_tensor(np.asarray(enc_seq), dtype=torch.float32).unsqueeze(0)      # (1, L, D)
        D = encoder_seq.shape[2]
        H = g('res_rnn1.weight_hh').shape[1]
        n_mels = g('mel_proj.weight').shape[0] // MAX_R
        attn_hidden = torch.zeros(1, D)
        rnn1_hidden = torch.zeros(1, H)
        rnn2_hidden = torch.zeros(1, H)
        rnn1_cell = torch.zeros(1, H)
        rnn2_cell = torch.zeros(1, H)
        go_frame = torch.zeros(1, n_mels)
        context_vec = torch.zeros(1, D)
        encoder_seq_proj = F.linear(encoder_seq, W['encoder_proj.weight'])                         # :440
        L = encoder_seq.shape[1]
        cumulative = torch.zeros(1, L)
        attention = torch.zeros(1, L)
        pad = (g('attn_net.conv.weight').shape[2] - 1) // 2
        mel_outputs, attn_scores = [], []
        for t in range(0, steps, r):                                                               # :446
            prenet_in = mel_outputs[-1][:, :, -1] if t > 0 else go_frame
            x = F.relu(F.linear(prenet_in, g('prenet.fc1.weight'), g('prenet.fc1.bias')))
            prenet_out = F.relu(F.linear(x, g('prenet.fc2.weight'), g('prenet.fc2.bias')))
            attn_rnn_in = torch.cat([context_vec, prenet_out], dim=-1)
            attn_hidden = torch.gru_cell(attn_rnn_in, attn_hidden, g('attn_rnn.weight_ih'), g('attn_rnn.weight_hh'),
                                         g('attn_rnn.bias_ih'), g('attn_rnn.bias_hh'))
            # LSA (:187-205)
            processed_query = F.linear(attn_hidden, g('attn_net.W.weight'), g('attn_net.W.bias')).unsqueeze(1)
            location = torch.cat([cumulative.unsqueeze(1), attention.unsqueeze(1)], dim=1)
            processed_loc = F.linear(F.conv1d(location, g('attn_net.conv.weight'), padding=pad).transpose(1, 2),
                                     g('attn_net.L.weight'), g('attn_net.L.bias'))
            u = F.linear(torch.tanh(processed_query + encoder_seq_proj + processed_loc), g('attn_net.v.weight'))
            u = u.squeeze(-1)
            scores = torch.sigmoid(u) / torch.sigmoid(u).sum(dim=1, keepdim=True)
            attention = scores
            cumulative = cumulative + attention
            scores = scores.unsqueeze(-1).transpose(1, 2)                                          # (1, 1, L)
            context_vec = (scores @ encoder_seq).squeeze(1)
            x = torch.cat([context_vec, attn_hidden], dim=1)
            x = F.linear(x, g('rnn_input.weight'), g('rnn_input.bias'))
            rnn1_hidden, rnn1_cell = torch.lstm_cell(x, (rnn1_hidden, rnn1_cell), g('res_rnn1.weight_ih'),
                                                     g('res_rnn1.weight_hh'), g('res_rnn1.bias_ih'), g('res_rnn1.bias_hh'))
            x = x + rnn1_hidden
            rnn2_hidden, rnn2_cell = torch.lstm_cell(x, (rnn2_hidden, rnn2_cell), g('res_rnn2.weight_ih'),
                                                     g('res_rnn2.weight_hh'), g('res_rnn2.bias_ih'), g('res_rnn2.bias_hh'))
            x = x + rnn2_hidden
            mels = F.linear(x, g('mel_proj.weight'))
            mel_frames = mels.view(1, n_mels, MAX_R)[:, :, :r]
            mel_outputs.append(mel_frames)
            attn_scores.append(scores)
            if (mel_frames < stop_threshold).all() and t > 10:                                     # :455
                break
        mel = torch.cat(mel_outputs, dim=2)[0].numpy()
        attn = torch.cat(attn_scores, 1)[0].numpy()
    return mel, attn
