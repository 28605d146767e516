"""Drop-in for `Tacotron.generate` (models/tacotron.py:420-480) with the decoder loop on the B200 kernel.

The reference's model object stays what it is (its class, its checkpoint, its CBHG encoder and postnet, which run once
per utterance through torch); only the sequential decoder loop (:438-465), the part with one launch-bound step per r
frames, is replaced by `taco_loop_kernel` (csrc/taco.cu) through the C ABI (`wrnn_taco_decode`).  Use:

    import types
    from wavernn_b200 import tacotron as b200_taco
    tts_model.generate = types.MethodType(b200_taco.generate, tts_model)      # gen_tacotron.py:105 onwards unchanged

or call `b200_taco.generate(tts_model, x, steps)` directly.  The model is duck-typed: it needs `.encoder`,
`.encoder_proj`, `.decoder` (for the state dict), `.postnet`, `.post_proj`, `.r`, `.stop_threshold`, `.state_dict()`.
There is no CPU fallback: without the CUDA library the call raises.
"""
import numpy as np

from .binding import TacotronDecoder

_ATTR = '_b200_taco_decoder'


def _decoder_for(model, device_index):
    """One `TacotronDecoder` per model object, re-loaded when the model's parameters change identity or version."""
    sd = model.state_dict()
    keys = [k for k in sd if k.startswith('decoder.') or k == 'encoder_proj.weight']
    stamp = tuple((k, sd[k].data_ptr(), sd[k]._version) for k in keys)
    cached = model.__dict__.get(_ATTR)
    if cached is not None and cached[0] == stamp and cached[2] == device_index:
        return cached[1]
    dims = dict(decoder_dims=int(sd['encoder_proj.weight'].shape[0]),
                lstm_dims=int(sd['decoder.res_rnn1.weight_hh'].shape[1]),
                n_mels=int(sd['decoder.mel_proj.weight'].shape[0]) // 20)
    dec = TacotronDecoder(device=device_index, **dims)
    dec.load_weights({k: sd[k] for k in keys})
    model.__dict__[_ATTR] = (stamp, dec, device_index)
    return dec


def generate(model, x, steps=2000):
    """`Tacotron.generate(x, steps)` -> (mel_outputs (n_mels, T), linear (fft_bins, T), attn_scores (n_iter, L)) as
    numpy arrays, like the reference (:467-478)."""
    import torch
    model.eval()
    device = next(model.parameters()).device
    with torch.no_grad():
        xt = torch.as_tensor(x, dtype=torch.long, device=device).unsqueeze(0)
        encoder_seq = model.encoder(xt)                                               # (1, L, decoder_dims), :446
        dec = _decoder_for(model, device.index if device.type == 'cuda' and device.index is not None else 0)
        thr = float(model.stop_threshold.item()) if hasattr(model.stop_threshold, 'item') else float(model.stop_threshold)
        mel, attn = dec.decode(encoder_seq[0].float().cpu().numpy(), steps=int(steps), r=int(model.r), stop_threshold=thr)
        mel_outputs = torch.as_tensor(mel, device=device).unsqueeze(0)                # (1, n_mels, T)
        postnet_out = model.postnet(mel_outputs)                                      # :467-468
        linear = model.post_proj(postnet_out)
        linear = linear.transpose(1, 2)[0].cpu().numpy()
    model.train()                                                                     # as the reference does (:478)
    return np.ascontiguousarray(mel), linear, np.ascontiguousarray(attn)
