"""GPU: parity at the BASELINE.json configurations AT FULL SIZE against the CPU oracle under identical injected noise.

The fixtures under tests/golden stop at T = 24 (3 folds x 3 100 steps); these tests run the oracle itself on the GPU
box's host cores (6-15 s per case) for the sizes the configs are quoted on:
  config 2  MoL, T = 801 -> 20 folds x 12 100 steps        config 2b  T = 800 -> 19 folds, no tail pad
  config 4  config 2 with 16x1 block-pruned GRUs (oracle: dense reference arithmetic on the masked weights)
  config 1  MoL, unbatched, T = 161 -> 1 fold x 44 275 steps
  config 3  RAW 9 bit + mu-law, 21 folds x 3 100 steps (float64 waveform bit-exact)
  deepmind  hidden 896, 8 sequences x 2 000 steps (int64 outputs bit-exact)
Each MoL case records, next to its assertion: max |wave - oracle|, the mixture-pick agreement and the first loop step
at which a fold's sample differs from the oracle's by more than 1e-6 (profiles/r02_parity_full_size.jsonl is a copy of
what a B200 run printed).  Marked `slow`: together about two minutes of oracle time.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import deepmind as odm
from oracle import fatchord as ofc
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims

pytestmark = [pytest.mark.gpu, pytest.mark.slow]

REPORT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'gpurun_out', 'parity_full_size.jsonl')

# measured on B200 (profiles/r02_parity_full_size.jsonl); the fp32 kernels differ from ATen's CPU kernels in summation
# order and in the 22-bit rounding of exchanged activations, never in a discrete decision at these sizes
MOL_WAVE_TOL = 2e-4


def report(**kw):
    line = json.dumps(kw)
    print('\nPARITY ' + line)
    try:
        os.makedirs(os.path.dirname(REPORT), exist_ok=True)
        with open(REPORT, 'a') as f:
            f.write(line + '\n')
    except OSError:
        pass


def mol_case(tag, dims, sd, T, batched, target, overlap, sd_oracle=None, kernel=0, prune=None):
    mel = synth.synth_mel(T, seed=1234)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    if prune:
        masked = v.prune_16x1(prune, want_masked=True)
        sd_oracle = dict(sd)
        sd_oracle.update(masked)
    v.set_loop_kernel(kernel)
    W = ofc.to_torch(sd_oracle if sd_oracle is not None else sd)
    with torch.no_grad():
        m, aux, wave_len = ofc.conditioning(W, dims, mel, batched, target, overlap)
        B, S = int(m.shape[0]), int(m.shape[1])
        noise = synth.synth_noise_mol(S, B, 4321)
        res = ofc.run_loop(W, dims, m, aux, noise)
    ref_wave = ofc.finish(dims, res['samples'], wave_len, batched, target, overlap, False)
    wave = v.generate([mel[0]], batched, target, overlap, False, noise=noise)[0]
    tm = v.last_timing()
    assert (tm['folds'], tm['steps']) == (B, S)
    got = v.fetch_samples().astype(np.float64)
    ref = res['samples'].numpy().astype(np.float64)
    d = np.abs(got - ref)
    over = np.argwhere(d.max(axis=0) > 1e-6)
    # mixture picks as the loop kernel recorded them are not fetched through the ABI; a flipped pick moves the sample
    # by the distance between two mixture means, far above 1e-3, so "all |diff| <= 1e-3" states pick agreement
    report(case=tag, kernel=v.loop_kernel_name(B), folds=B, steps=S, max_abs_wave_diff=float(np.abs(wave - ref_wave).max()),
           max_abs_sample_diff=float(d.max()), mean_abs_sample_diff=float(d.mean()),
           first_step_sample_diff_gt_1e_6=int(over[0, 0]) if len(over) else None,
           frac_fold_steps_within_1e_6=float((d <= 1e-6).mean()), frac_within_1e_3=float((d <= 1e-3).mean()),
           us_per_step=1000 * tm['ms_loop'] / S)
    assert wave.shape == ref_wave.shape == (wave_len,)
    return wave, ref_wave, d, v


def test_config2_full_size_vs_oracle():
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    wave, ref, d, v = mol_case('config2_T801', dims, sd, 801, True, 11000, 550)
    assert v.loop_kernel_name(20) == 'fma-resident'
    assert np.abs(wave - ref).max() <= MOL_WAVE_TOL and (d <= 1e-3).all()


def test_config2_trained_like_weights_full_size_vs_oracle():
    """Same size with the "trained-like" weights: saturating samples, log-scales at the clamp."""
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='trained')
    wave, ref, d, v = mol_case('config2_T801_trained', dims, sd, 801, True, 11000, 550)
    assert np.abs(wave - ref).max() <= MOL_WAVE_TOL and (d <= 1e-3).all()


def test_config2b_T800_no_tail_pad_vs_oracle():
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    wave, ref, d, v = mol_case('config2b_T800', dims, sd, 800, True, 11000, 550)
    assert v.last_timing()['folds'] == 19                       # (220000 - 550) / 11550 = 19 exactly: no padded fold
    assert np.abs(wave - ref).max() <= MOL_WAVE_TOL and (d <= 1e-3).all()


def test_config4_block_sparse_full_size_vs_dense_oracle_on_masked_weights():
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    wave, ref, d, v = mol_case('config4_T801_sparse16x1', dims, sd, 801, True, 11000, 550, prune=0.1)
    assert v.loop_kernel_name(20) == 'fma-resident-sparse16x1'
    assert np.abs(wave - ref).max() <= MOL_WAVE_TOL and (d <= 1e-3).all()


def test_config1_unbatched_T161_vs_oracle():
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    wave, ref, d, v = mol_case('config1_T161_unbatched', dims, sd, 161, False, 11000, 550)
    assert wave.shape == (44000,) and v.last_timing()['steps'] == 44275
    assert np.abs(wave - ref).max() <= MOL_WAVE_TOL and (d <= 1e-3).all()


def test_exact_order_fma_kernel_mol_against_the_reference_fixture():
    """north_star: "sampled waveforms match exactly when identical uniform noise tensors are injected".  Measured on
    the kernel that keeps the reference's order of operations (no composed matrices, fp32 FMA, accurate libdevice
    exp/log/tanh): equality is NOT bit-for-bit.  ATen's CPU GEMV / gru_cell sum in a different order (vectorised
    partial sums) and libm differs from libdevice by <= 1-2 ulp, so continuous MoL samples differ in the last bits;
    what holds exactly is every discrete decision (mixture pick), and RAW / deepmind outputs (discrete) are bit-equal.
    The deviation "exact" -> "<= 2e-4 absolute, identical picks" is recorded in DESIGN.md section 2."""
    from common import case_inputs, golden
    name = 'full_mol_batched'
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    v.set_loop_kernel(1)
    wave = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    ref = golden(name)['wave']
    d = np.abs(wave - ref)
    report(case='exact_order_fma_vs_reference_fixture', kernel=v.loop_kernel_name(B), bit_equal=bool(np.array_equal(wave, ref)),
           frac_bit_equal=float((wave == ref).mean()), max_abs_wave_diff=float(d.max()),
           first_differing_sample=int(np.argmax(d > 0)) if (d > 0).any() else None)
    assert d.max() <= MOL_WAVE_TOL


def test_config3_raw_21_folds_3100_steps_bit_exact():
    dims = VocoderDims(mode='RAW', bits=9)
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    T, target, overlap = 188, 2000, 550
    mel = synth.synth_mel(T, seed=1234)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    W = ofc.to_torch(sd)
    with torch.no_grad():
        m, aux, wave_len = ofc.conditioning(W, dims, mel, True, target, overlap)
        B, S = int(m.shape[0]), int(m.shape[1])
        assert (B, S) == (21, 3100)
        noise = synth.synth_noise_exp(S, B, dims.n_classes, 4321)
        res = ofc.run_loop(W, dims, m, aux, noise)
    ref = ofc.finish(dims, res['samples'], wave_len, True, target, overlap, True)
    wave = v.generate([mel[0]], True, target, overlap, True, noise=noise)[0]
    got = v.fetch_samples()
    same = float((got == res['samples'].numpy()).mean())
    report(case='config3_raw_21x3100', kernel=v.loop_kernel_name(B), class_pick_agreement=same,
           wave_bit_equal=bool(np.array_equal(wave, ref)))
    assert same == 1.0 and np.array_equal(wave, ref)


def test_deepmind_896_eight_sequences_2000_steps_bit_exact():
    hs, S, N = 896, 2000, 8
    sd = synth.synth_dm_state_dict(hs, 256, seed=0)
    qc = synth.synth_noise_exp(S, N, 256, seed=4321)
    qf = synth.synth_noise_exp(S, N, 256, seed=4322)
    v = binding.DeepmindVocoder(hs, 256, device=0)
    v.load_weights(sd)
    out, c, f = v.generate(S, N, qc, qf)
    o_out, o_c, o_f = odm.generate(ofc.to_torch(sd), S, qc, qf, n_seq=N)
    report(case='deepmind_896_8x2000', coarse_equal=bool(np.array_equal(c, o_c)), fine_equal=bool(np.array_equal(f, o_f)),
           us_per_step=1000 * v.last_timing()['ms_loop'] / S)
    assert np.array_equal(c, o_c) and np.array_equal(f, o_f) and np.array_equal(out, o_out)
