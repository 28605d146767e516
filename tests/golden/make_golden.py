"""Generate golden vectors by running the REAL reference (build container only).

    python tests/golden/make_golden.py

For every case: run the reference's own code from /root/reference (import shims only, see
oracle/ref_shim.py), run the oracle restatement on the same inputs and injected noise, assert
the two are BIT-EQUAL on CPU, then write the reference's outputs to tests/golden/<case>.npz.
The fixtures are what pins the oracle on machines without /root/reference (the GPU box).
Inputs are not stored: they are rebuilt from `wavernn_b200.synth` seeds recorded in each file.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import deepmind as odm          # noqa: E402
from oracle import fatchord as ofc          # noqa: E402
from oracle import ref_shim                 # noqa: E402
from wavernn_b200 import synth              # noqa: E402
from wavernn_b200.config import VocoderDims  # noqa: E402

TINY = dict(rnn_dims=64, fc_dims=64, bits=6, pad=2, upsample_factors=(2, 2, 3), feat_dims=16,
            compute_dims=32, res_out_dims=32, res_blocks=2, hop_length=12, sample_rate=22050)

# name -> (dims, weight style, T, batched, target, overlap, mu_law)
CASES = {
    'tiny_mol_unbatched': (VocoderDims(mode='MOL', **TINY), 'trained', 40, False, 100, 20, False),
    'tiny_mol_batched': (VocoderDims(mode='MOL', **TINY), 'trained', 61, True, 100, 20, False),
    'tiny_mol_batched_nopad': (VocoderDims(mode='MOL', **TINY), 'init', 52, True, 100, 4, False),
    'tiny_raw_batched_mulaw': (VocoderDims(mode='RAW', **TINY), 'trained', 61, True, 100, 20, True),
    'tiny_raw_unbatched': (VocoderDims(mode='RAW', **TINY), 'init', 40, False, 100, 20, False),
    'full_mol_unbatched': (VocoderDims(mode='MOL'), 'init', 22, False, 11000, 550, False),
    'full_mol_batched': (VocoderDims(mode='MOL'), 'trained', 24, True, 2000, 550, False),
    'full_raw_batched_mulaw': (VocoderDims(mode='RAW'), 'init', 24, True, 2000, 550, True),
}


def same(a, b, what):
    a = a.detach().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().numpy() if torch.is_tensor(b) else np.asarray(b)
    if a.shape != b.shape or not np.array_equal(a, b):
        d = np.abs(a.astype(np.float64) - b.astype(np.float64)).max() if a.shape == b.shape else 'shape'
        raise SystemExit(f'ORACLE != REFERENCE for {what}: {d}')


def noise_for(dims, S, B, seed=4321):
    if dims.mode == 'MOL':
        return synth.synth_noise_mol(S, B, seed)
    return synth.synth_noise_exp(S, B, dims.n_classes, seed)


def run_case(name, dims, style, T, batched, target, overlap, mu_law):
    sd = synth.synth_state_dict(dims, seed=0, style=style)
    W = ofc.to_torch(sd)
    model = ref_shim.build_fatchord(dims, sd)
    mel = synth.synth_mel(T, seed=1234, feat_dims=dims.feat_dims)
    out = {'T': T, 'batched': batched, 'target': target, 'overlap': overlap, 'mu_law': mu_law,
           'style': style, 'mel_seed': 1234, 'weight_seed': 0, 'noise_seed': 4321}

    model.eval()
    with torch.no_grad():
        # conditioning: reference upsample on the padded mel (fatchord_version.py:185-186)
        mp = model.pad_tensor(torch.as_tensor(mel).transpose(1, 2), pad=dims.pad, side='both').transpose(1, 2)
        r_m, r_aux = model.upsample(mp)
        o_m, o_aux = ofc.upsample(W, dims, mp)
        same(r_m, o_m, name + ' upsample.mels')
        same(r_aux, o_aux, name + ' upsample.aux')
        L = r_m.shape[1]
        # keep fixtures small: a strided sample of time steps plus both ends
        idx = np.unique(np.concatenate([np.arange(0, L, max(1, L // 97)), np.arange(min(64, L)),
                                        np.arange(max(0, L - 64), L)]))
        out['up_idx'] = idx
        out['up_mels'] = r_m[0, idx].numpy()
        out['up_aux'] = r_aux[0, idx].numpy()
        out['up_mels_sum'] = r_m.double().sum().item()
        out['up_aux_sum'] = r_aux.double().sum().item()
        if batched:
            rf = model.fold_with_overlap(r_m, target, overlap)
            of = ofc.fold_with_overlap(o_m, target, overlap)
            same(rf, of, name + ' fold')
            out['fold_shape'] = np.array(rf.shape)
            B, S = rf.shape[0], rf.shape[1]
        else:
            B, S = 1, L

    noise = noise_for(dims, S, B)
    if dims.mode == 'MOL':
        uni = []
        for i in range(S):
            uni += [torch.from_numpy(noise[0][i]).unsqueeze(0), torch.from_numpy(noise[1][i]).unsqueeze(0)]
        inj = ref_shim.injected_noise(uniforms=uni)
    else:
        inj = ref_shim.injected_noise(exps=[torch.from_numpy(noise[i]) for i in range(S)])
    with inj:
        r_wave = model.generate(torch.as_tensor(mel), '/dev/null', batched, target, overlap, mu_law)
    o_wave = ofc.generate(W, dims, mel, batched, target, overlap, mu_law, noise)
    same(r_wave, o_wave, name + ' generate')
    assert r_wave.dtype == np.float64 and r_wave.shape == ((T - 1) * dims.hop_length,)
    out['wave'] = r_wave

    # teacher-forced logits through the reference's training forward (nn.GRU), F9
    Sf = 4 * dims.hop_length
    Tf = Sf // dims.hop_length
    g = torch.Generator().manual_seed(99)
    xs = torch.rand((2, Sf), generator=g) * 2 - 1
    mels_f = torch.rand((2, dims.feat_dims, Tf + 2 * dims.pad), generator=g)
    model.eval()                      # generate() leaves the module in train() mode (:262)
    with torch.no_grad():
        step_before = model.step.clone()
        r_log = model.forward(xs, mels_f)
        model.step.copy_(step_before)
    o_log = ofc.forward_logits(W, dims, xs.numpy(), mels_f.numpy())
    err = (r_log - o_log).abs().max().item()
    if err > 2e-6 * max(1.0, r_log.abs().max().item()):
        raise SystemExit(f'{name}: forward logits oracle vs reference {err}')
    stride = 1 if r_log.numel() < 200000 else 8    # keep fixtures small
    out['fwd_logits'] = r_log.numpy()[:, ::stride]
    out['fwd_stride'] = stride
    out['fwd_seed'] = 99
    out['fwd_gru_vs_cell_maxabs'] = err
    np.savez_compressed(os.path.join(HERE, name + '.npz'), **out)
    print(f'{name}: ok  B={B} S={S} wave={r_wave.shape} fwd_err={err:.2e}')


def run_components():
    fv, dv, dist, dsp = ref_shim.load()
    out = {}
    dims = VocoderDims()
    sd = synth.synth_state_dict(VocoderDims(mode='MOL', **TINY), 0, 'init')
    model = ref_shim.build_fatchord(VocoderDims(mode='MOL', **TINY), sd)
    # fold: docstring example (fatchord_version.py:312-316) and ragged cases
    x = torch.arange(1, 11, dtype=torch.float32).view(1, 10, 1)
    f = model.fold_with_overlap(x, 2, 1)
    same(f, ofc.fold_with_overlap(x, 2, 1), 'fold docstring')
    assert f[:, :, 0].tolist() == [[1, 2, 3, 4], [4, 5, 6, 7], [7, 8, 9, 10]]
    out['fold_doc'] = f.numpy()
    g = torch.Generator().manual_seed(7)
    for L, tgt, ov in [(1000, 100, 20), (1220, 100, 20), (620, 100, 20), (97, 10, 3), (4400, 1100, 55)]:
        x = torch.rand((1, L, 5), generator=g)
        f = model.fold_with_overlap(x, tgt, ov)
        same(f, ofc.fold_with_overlap(x, tgt, ov), f'fold {L}')
        out[f'fold_{L}_{tgt}_{ov}_shape'] = np.array(f.shape)
        out[f'fold_{L}_{tgt}_{ov}_sum'] = f.double().sum().item()
    # unfold
    rs = np.random.RandomState(5)
    for B, tgt, ov in [(3, 100, 20), (1, 50, 7), (20, 11000, 550), (4, 30, 1)]:
        y = rs.uniform(-1, 1, size=(B, tgt + 2 * ov))
        r = model.xfade_and_unfold(y.copy(), tgt, ov)
        o = ofc.xfade_and_unfold(y.copy(), tgt, ov)
        same(r, o, f'unfold {B}')
        out[f'unfold_{B}_{tgt}_{ov}'] = r if r.size < 5000 else r[::37]
    # mu-law decode over every class value of 9-bit and 6-bit RAW models
    for n in (512, 64):
        k = torch.arange(n)
        y = (2 * k.float() / (n - 1.) - 1.).numpy().astype(np.float64)
        r = dsp.decode_mu_law(y, n, False)
        same(r, ofc.decode_mu_law(y, n, False), f'mulaw {n}')
        out[f'mulaw_{n}'] = r
    lab = np.arange(512, dtype=np.float64)
    same(dsp.decode_mu_law(lab, 512, True), ofc.decode_mu_law(lab, 512, True), 'mulaw labels')
    # MoL sampler alone
    g = torch.Generator().manual_seed(11)
    logits = torch.randn((37, 30), generator=g) * 3
    u1 = torch.empty(37, 10).uniform_(1e-5, 1 - 1e-5, generator=g)
    u2 = torch.empty(37).uniform_(1e-5, 1 - 1e-5, generator=g)
    with ref_shim.injected_noise(uniforms=[u1.unsqueeze(0), u2.unsqueeze(0)]):
        r = dist.sample_from_discretized_mix_logistic(logits.unsqueeze(0).transpose(1, 2))
    o, _ = ofc.mol_sample(logits, u1, u2)
    same(r.view(-1), o, 'mol sampler')
    out['mol_logits'], out['mol_u1'], out['mol_u2'], out['mol_out'] = logits.numpy(), u1.numpy(), u2.numpy(), r.view(-1).numpy()
    # Categorical.sample == argmax(p/q) with the generator-drawn q (F8)
    p = torch.softmax(torch.randn((5, 64), generator=g), dim=1)
    torch.manual_seed(123)
    s_ref = torch.distributions.Categorical(p).sample()
    torch.manual_seed(123)
    q = torch.empty_like(p).exponential_(1)
    pn = p / p.sum(-1, keepdim=True)
    assert torch.equal(s_ref, torch.argmax(pn / q, dim=-1)), 'multinomial != exponential race'
    np.savez_compressed(os.path.join(HERE, 'components.npz'), **out)
    print('components: ok')


def run_deepmind():
    out = {}
    for name, hs, nseq, S in [('dm_small', 64, 1, 300), ('dm_full', 896, 1, 400)]:
        sd = synth.synth_dm_state_dict(hs, 256, seed=0)
        model = ref_shim.build_deepmind(sd, hs, 256)
        W = ofc.to_torch(sd)
        qc = synth.synth_noise_exp(S, 1, 256, seed=4321)
        qf = synth.synth_noise_exp(S, 1, 256, seed=4322)
        exps = []
        for i in range(S):
            exps += [torch.from_numpy(qc[i]), torch.from_numpy(qf[i])]
        with ref_shim.injected_noise(exps=exps):
            r_out, r_c, r_f = model.generate(S)
        o_out, o_c, o_f = odm.generate(W, S, qc, qf)
        same(r_out, o_out, name + ' output')
        same(r_c, o_c, name + ' coarse')
        same(r_f, o_f, name + ' fine')
        out[name + '_output'], out[name + '_coarse'], out[name + '_fine'] = r_out, r_c, r_f
        print(f'{name}: ok  distinct coarse={len(np.unique(r_c))}')
    np.savez_compressed(os.path.join(HERE, 'deepmind.npz'), **out)


if __name__ == '__main__':
    torch.set_num_threads(8)
    only = sys.argv[1:]
    if not only or 'components' in only:
        run_components()
    if not only or 'deepmind' in only:
        run_deepmind()
    for name, spec in CASES.items():
        if not only or name in only:
            run_case(name, *spec)
